#!/bin/bash
# One GPU-box session: smoke, GPU tests in separate processes (a CUDA fault in one group must not
# hide the others), golden-vector generation from the reference extension, a short bench.
# Usage: gpurun --timeout 1500 -- 'bash scripts/gpu_ci.sh [stage ...]'
set -u
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
stages=${@:-smoke tests golden bench_small}
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
nproc >> gpurun_out/gpu.txt
for s in $stages; do
  case $s in
    smoke)
      timeout 180 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/summary.txt ;;
    tests)
      for grp in "dist2 and not noalloc" "mark_visible or edge or debug" "against_oracle or modes or deterministic or fused" "golden" "reference_live" "full_size" "noalloc"; do
        name=$(echo "$grp" | tr ' ' '_')
        timeout 240 python -m pytest tests -m gpu -q -k "$grp" -p no:cacheprovider > "gpurun_out/pytest_${name}.log" 2>&1
        echo "pytest[$grp] rc=$? $(tail -1 gpurun_out/pytest_${name}.log)" | tee -a gpurun_out/summary.txt
      done ;;
    golden)
      timeout 600 python tests/golden/make_golden.py gpurun_out/golden > gpurun_out/golden.log 2>&1; echo "golden rc=$?" | tee -a gpurun_out/summary.txt ;;
    bench_small)
      timeout 600 python bench.py --P 1000000 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/bench_small.json 2> gpurun_out/bench_small.err; echo "bench_small rc=$?" | tee -a gpurun_out/summary.txt
      timeout 600 python bench.py --impl reference --P 1000000 --steps 2 --warmup 1 > gpurun_out/bench_small_ref.json 2> gpurun_out/bench_small_ref.err; echo "bench_small_ref rc=$?" | tee -a gpurun_out/summary.txt ;;
    bench)
      timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" | tee -a gpurun_out/summary.txt
      timeout 900 python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench_ref rc=$?" | tee -a gpurun_out/summary.txt ;;
    configs)
      timeout 1200 python scripts/bench_configs.py > gpurun_out/configs.json 2> gpurun_out/configs.err; echo "configs rc=$?" | tee -a gpurun_out/summary.txt ;;
    sweep)
      # super-batch size of the compositing kernels (threads per CTA): rebuild and time each
      for sz in ${VEGS_SWEEP:-128}; do
        touch vegs_b200/csrc/composite_fwd.cu vegs_b200/csrc/composite_bwd.cu
        VEGS_NVCC_EXTRA="-DVEGS_FWD_SUPER=$sz" python vegs_b200/build.py > gpurun_out/sweep_build_$sz.log 2>&1
        timeout 600 python bench.py --steps 6 --warmup 3 --no-cpu-baseline > gpurun_out/sweep_$sz.json 2> gpurun_out/sweep_$sz.err
        echo "sweep $sz rc=$?" | tee -a gpurun_out/summary.txt
        timeout 600 python -m pytest tests -m gpu -q -k "against_oracle or golden" -p no:cacheprovider > gpurun_out/sweep_pytest_$sz.log 2>&1
        echo "sweep pytest $sz rc=$? $(tail -1 gpurun_out/sweep_pytest_$sz.log)" | tee -a gpurun_out/summary.txt
      done
      touch vegs_b200/csrc/composite_fwd.cu vegs_b200/csrc/composite_bwd.cu
      python vegs_b200/build.py > gpurun_out/sweep_build_default.log 2>&1 ;;
    tilestats)
      touch vegs_b200/csrc/composite_fwd.cu vegs_b200/csrc/composite_bwd.cu
      VEGS_TILE_DEBUG=1 python vegs_b200/build.py > gpurun_out/dbg_build.log 2>&1
      timeout 600 python scripts/tile_stats.py > gpurun_out/tile_stats.log 2>&1; echo "tilestats rc=$?" | tee -a gpurun_out/summary.txt
      touch vegs_b200/csrc/composite_fwd.cu vegs_b200/csrc/composite_bwd.cu
      python vegs_b200/build.py >> gpurun_out/dbg_build.log 2>&1 ;;
    ncu_pre)
      CMD="python bench.py --P 6000000 --views 1 --steps 1 --warmup 1 --no-cpu-baseline"
      timeout 600 $CMD > gpurun_out/ncu_plain.log 2>&1 &&
      timeout 900 ncu --set full --clock-control none --import-source on -k regex:'preprocess_(fwd|bwd)_kernel|emit_kernel|radix_scatter' -s 6 -c 8 -o gpurun_out/prof_pre $CMD > gpurun_out/ncu_pre.log 2>&1
      echo "ncu pre rc=$?" | tee -a gpurun_out/summary.txt ;;
    ncu)
      # launch list (cold-cache, serialised: compare shares) and one full capture of the top kernels;
      # the same command must have exited 0 without ncu first
      CMD="python bench.py --P 6000000 --views 2 --steps 1 --warmup 1 --no-cpu-baseline"
      timeout 600 $CMD > gpurun_out/ncu_plain.log 2>&1 &&
      timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
      echo "ncu launches rc=$?" | tee -a gpurun_out/summary.txt
      timeout 900 ncu --set full --clock-control none --import-source on -k regex:'composite_(fwd|bwd)_kernel' -s 2 -c 2 -o gpurun_out/prof_composite $CMD > gpurun_out/ncu_full.log 2>&1
      echo "ncu full rc=$?" | tee -a gpurun_out/summary.txt ;;
  esac
done
cat gpurun_out/summary.txt
