#!/bin/bash
# Multi-GPU bench: bash scripts/gpu_scale.sh N   (run under gpurun --gpus N)
set -u
N=${1:-2}
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
nvidia-smi --query-gpu=index,name --format=csv > gpurun_out/gpus_$N.txt 2>&1
PORT=29531
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT \
  bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/scale_${N}.json 2> gpurun_out/scale_${N}.err
echo "ours N=$N rc=$?"
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((PORT+1)) \
  bench.py --impl reference --gpus $N --steps 3 --warmup 3 > gpurun_out/scale_${N}_ref.json 2> gpurun_out/scale_${N}_ref.err
echo "reference N=$N rc=$?"
tail -c 600 gpurun_out/scale_${N}.json; echo; tail -c 300 gpurun_out/scale_${N}_ref.json; echo
tail -5 gpurun_out/scale_${N}.err
