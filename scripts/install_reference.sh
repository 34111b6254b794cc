#!/bin/bash
# Installs the UNMODIFIED reference extensions (deepshwang/vegs submodules diff_gaussian_rasterization
# and simple-knn) into baseline/_ref, which is git-ignored but travels to the GPU box with gpurun.
# They are the live parity oracle of the GPU tests and bench.py's reference arm.  Run in the
# container that has /root/reference; the build writes into its source tree, so it runs on a copy.
set -eu
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
TMP=${TMPDIR:-/tmp}/vegs_refbuild
rm -rf "$TMP" && mkdir -p "$TMP"
cp -r /root/reference/submodules/diff_gaussian_rasterization "$TMP/dgr"
cp -r /root/reference/submodules/simple-knn "$TMP/sknn"
cd "$ROOT"
export TORCH_CUDA_ARCH_LIST=10.0a MAX_JOBS=${MAX_JOBS:-6}
python -m pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse --target baseline/_ref "$TMP/dgr"
# simple_knn.cu uses FLT_MAX without including <cfloat>: add the include on the command line, not in the source
NVCC_APPEND_FLAGS="-include cfloat" python -m pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse --target baseline/_ref "$TMP/sknn"
ls baseline/_ref
