#!/bin/bash
# Round-2 GPU-box stages.  Usage: gpurun --timeout N -- 'bash scripts/gpu_r2.sh <tag> stage ...'
# Every stage runs in its own process (a CUDA fault in one must not hide the others) under its own
# timeout; everything lands under gpurun_out/<tag>_*.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
tag=$1; shift
S=gpurun_out/${tag}_summary.txt
: > $S
run_pytest() {  # name, -k expression, timeout
  timeout $3 python -m pytest tests -m gpu -q -x -k "$2" -p no:cacheprovider > "gpurun_out/${tag}_pytest_$1.log" 2>&1
  echo "pytest[$1] rc=$? $(tail -1 gpurun_out/${tag}_pytest_$1.log)" | tee -a $S
}
for s in "$@"; do
  case $s in
    smoke)
      timeout 240 python __graft_entry__.py smoke > gpurun_out/${tag}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $S ;;
    tests)
      run_pytest base "not full and not noalloc and not config and not callers and not giant" 600
      run_pytest noalloc "noalloc" 240 ;;
    tests_full)
      run_pytest config2 "config2" 600
      run_pytest config3 "config3" 900
      run_pytest config5 "config5" 900
      run_pytest config4 "config4" 300
      run_pytest callers "callers or giant" 600 ;;
    tests_all)   # what the driver runs
      timeout 2400 python -m pytest tests -m gpu -q -x -p no:cacheprovider > gpurun_out/${tag}_pytest_all.log 2>&1
      echo "pytest[all] rc=$? $(tail -1 gpurun_out/${tag}_pytest_all.log)" | tee -a $S ;;
    bench)
      timeout 900 python bench.py --steps ${STEPS:-10} --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?" | tee -a $S ;;
    bench_ref)
      timeout 900 python bench.py --impl reference --steps ${STEPS:-10} --warmup 3 > gpurun_out/${tag}_bench_ref.json 2> gpurun_out/${tag}_bench_ref.err; echo "bench_ref rc=$?" | tee -a $S
      # proof that the reference process never maps the repo's library
      grep -c libvegs_b200 gpurun_out/${tag}_bench_ref.err >> $S ;;
    bench_nobulk)   # A/B: the per-Gaussian kernels on their 16-byte-load paths instead of bulk copies
      VEGS_SH_BULK=0 timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-fused-leg > gpurun_out/${tag}_bench_nobulk.json 2> gpurun_out/${tag}_bench_nobulk.err; echo "bench_nobulk rc=$?" | tee -a $S ;;
    bench_nolookback)   # A/B: three-kernel radix passes instead of the look-back scatter
      VEGS_SORT_LOOKBACK=0 timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-fused-leg > gpurun_out/${tag}_bench_nolookback.json 2> gpurun_out/${tag}_bench_nolookback.err; echo "bench_nolookback rc=$?" | tee -a $S ;;
    tests_nobulk)
      VEGS_SH_BULK=0 timeout 600 python -m pytest tests -m gpu -q -x -k "against_oracle or golden or reference_live or forward_raw" -p no:cacheprovider > gpurun_out/${tag}_pytest_nobulk.log 2>&1
      echo "pytest[nobulk] rc=$? $(tail -1 gpurun_out/${tag}_pytest_nobulk.log)" | tee -a $S ;;
    multi)      # needs gpurun --gpus N (NGPU = N)
      N=${NGPU:-2}
      if [ "$N" = "2" ]; then
        timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q -x -p no:cacheprovider > gpurun_out/${tag}_pytest_multi.log 2>&1
        echo "pytest[multi] rc=$? $(tail -1 gpurun_out/${tag}_pytest_multi.log)" | tee -a $S
      fi
      timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps ${STEPS:-10} --warmup 3 > gpurun_out/${tag}_scale_$N.json 2> gpurun_out/${tag}_scale_$N.err
      echo "scale $N rc=$?" | tee -a $S
      if [ -n "${MULTI_OURS_ONLY:-}" ]; then
        if [ -n "${MULTI_BUSBW:-}" ]; then
          timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 scripts/nccl_busbw.py > gpurun_out/${tag}_nccl_busbw_$N.json 2> gpurun_out/${tag}_nccl_busbw_$N.err
          echo "nccl busbw $N rc=$? $(cat gpurun_out/${tag}_nccl_busbw_$N.json)" | tee -a $S
        fi
        cat $S; exit 0
      fi
      timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29518 bench.py --impl reference --gpus $N --steps 4 --warmup 3 > gpurun_out/${tag}_scale_${N}_ref.json 2> gpurun_out/${tag}_scale_${N}_ref.err
      echo "scale ref $N rc=$?" | tee -a $S
      timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 scripts/nccl_busbw.py > gpurun_out/${tag}_nccl_busbw_$N.json 2> gpurun_out/${tag}_nccl_busbw_$N.err
      echo "nccl busbw $N rc=$? $(cat gpurun_out/${tag}_nccl_busbw_$N.json)" | tee -a $S
      for algo in NVLS Ring Tree; do
        NCCL_ALGO=$algo timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29520 scripts/nccl_busbw.py > gpurun_out/${tag}_nccl_busbw_${N}_$algo.json 2> gpurun_out/${tag}_nccl_busbw_${N}_$algo.err
        echo "nccl busbw $N NCCL_ALGO=$algo rc=$? $(cat gpurun_out/${tag}_nccl_busbw_${N}_$algo.json)" | tee -a $S
      done ;;
    sweep)      # working-set size of the compositing kernels: VEGS_SWEEP="F:B F:B ..." (threads = candidates per set)
      for fb in ${VEGS_SWEEP:-512:256 256:512 512:512}; do
        F=${fb%%:*}; B=${fb##*:}
        touch vegs_b200/csrc/composite_fwd.cu vegs_b200/csrc/composite_bwd.cu
        VEGS_NVCC_EXTRA="-DVEGS_FWD_SUPER=$F -DVEGS_BWD_SUPER=$B" python vegs_b200/build.py > gpurun_out/${tag}_sweep_build_${F}_${B}.log 2>&1
        timeout 600 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --no-fused-leg > gpurun_out/${tag}_sweep_${F}_${B}.json 2> gpurun_out/${tag}_sweep_${F}_${B}.err
        echo "sweep $F:$B rc=$? $(python -c "import json;d=json.load(open('gpurun_out/${tag}_sweep_${F}_${B}.json'));print(round(d['value'],1), {k: round(v['ms_per_view'],3) for k,v in d['kernels'].items()})" 2>&1 | tail -1)" | tee -a $S
        timeout 900 python -m pytest tests -m gpu -q -x -k "against_oracle or golden or config3 or config5" -p no:cacheprovider > gpurun_out/${tag}_sweep_pytest_${F}_${B}.log 2>&1
        echo "sweep pytest $F:$B rc=$? $(tail -1 gpurun_out/${tag}_sweep_pytest_${F}_${B}.log)" | tee -a $S
      done
      touch vegs_b200/csrc/composite_fwd.cu vegs_b200/csrc/composite_bwd.cu
      python vegs_b200/build.py > gpurun_out/${tag}_sweep_build_default.log 2>&1 ;;
    tilestats)  # per-tile cycles of the compositing kernels (default build: clocks only)
      timeout 600 python scripts/tile_stats.py 6000000 gpurun_out/${tag}_tile_stats.npz > gpurun_out/${tag}_tile_stats.log 2>&1; echo "tilestats rc=$?" | tee -a $S ;;
    flags)      # A/B of build flags: VEGS_FLAG_SETS="name=-DA=1 -DB=2;name2=-DC=1" (each: rebuild, bench, parity tests)
      IFS=';' read -ra SETS <<< "${VEGS_FLAG_SETS:-bulklist=-DVEGS_FWD_BULK_LIST=1}"
      for fs in "${SETS[@]}"; do
        NAME=${fs%%=*}; FL=${fs#*=}
        touch vegs_b200/csrc/*.cu
        VEGS_NVCC_EXTRA="$FL" python vegs_b200/build.py > gpurun_out/${tag}_flags_build_${NAME}.log 2>&1
        timeout 600 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --no-fused-leg > gpurun_out/${tag}_flags_${NAME}.json 2> gpurun_out/${tag}_flags_${NAME}.err
        echo "flags $NAME [$FL] rc=$? $(python -c "import json;d=json.load(open('gpurun_out/${tag}_flags_${NAME}.json'));print(round(d['value'],1), {k: round(v['ms_per_view'],3) for k,v in d['kernels'].items()})" 2>&1 | tail -1)" | tee -a $S
        if [ -z "$VEGS_FLAGS_NOTEST" ]; then
          timeout 900 python -m pytest tests -m gpu -q -x -k "against_oracle or golden or config3 or config5" -p no:cacheprovider > gpurun_out/${tag}_flags_pytest_${NAME}.log 2>&1
          echo "flags pytest $NAME rc=$? $(tail -1 gpurun_out/${tag}_flags_pytest_${NAME}.log)" | tee -a $S
        fi
      done
      touch vegs_b200/csrc/*.cu
      python vegs_b200/build.py > gpurun_out/${tag}_flags_build_default.log 2>&1 ;;
    ncu_phases)  # per-phase stall / instruction shares of the two compositing kernels (source page cut at the barriers)
      CMD="python bench.py --P 6000000 --views 2 --steps 1 --warmup 3 --no-cpu-baseline --no-fused-leg"
      timeout 900 ncu --set full --clock-control none --import-source on -k regex:'composite_(fwd|bwd)_kernel' -s 6 -c 2 -f -o gpurun_out/${tag}_prof_composite $CMD > gpurun_out/${tag}_ncu_composite.log 2>&1
      echo "ncu composite rc=$?" | tee -a $S
      ncu -i gpurun_out/${tag}_prof_composite.ncu-rep --page source --csv --print-source sass > gpurun_out/${tag}_composite_source.csv 2> gpurun_out/${tag}_composite_source.err
      python scripts/ncu_phases.py gpurun_out/${tag}_composite_source.csv > gpurun_out/${tag}_ncu_phases.txt 2>&1
      python scripts/ncu_table.py gpurun_out/${tag}_prof_composite.ncu-rep gpurun_out/${tag}_ncu_composite.md "ncu --set full --clock-control none --import-source on, compositing kernels ($tag): \`$CMD\`" > /dev/null 2>&1
      rm -f gpurun_out/${tag}_composite_source.csv
      echo "ncu phases rc=$? $(wc -l < gpurun_out/${tag}_ncu_phases.txt) lines" | tee -a $S ;;
    bench_quick)
      timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench_quick rc=$?" | tee -a $S ;;
    configs)
      timeout 1200 python scripts/bench_configs.py > gpurun_out/${tag}_configs.json 2> gpurun_out/${tag}_configs.err; echo "configs rc=$?" | tee -a $S ;;
    ncu_launches)
      CMD="python bench.py --P 6000000 --views 2 --steps 1 --warmup 3 --no-cpu-baseline --no-fused-leg"
      timeout 600 $CMD > gpurun_out/${tag}_ncu_plain.log 2>&1 &&
      timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 400 --csv --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu_launches.log 2>&1
      echo "ncu launches rc=$?" | tee -a $S
      python scripts/ncu_launch_table.py gpurun_out/${tag}_launches.csv gpurun_out/${tag}_ncu_launches.md "ncu launch list ($tag): \`ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 400 $CMD\`" > /dev/null 2>&1 ;;
    ncu_launches_ref)
      CMD="python bench.py --impl reference --P 6000000 --views 2 --steps 1 --warmup 3"
      timeout 600 $CMD > gpurun_out/${tag}_ncu_plain_ref.log 2>&1 &&
      timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 200 -c 400 --csv --log-file gpurun_out/${tag}_launches_ref.csv $CMD > gpurun_out/${tag}_ncu_launches_ref.log 2>&1
      echo "ncu launches ref rc=$?" | tee -a $S
      python scripts/ncu_launch_table.py gpurun_out/${tag}_launches_ref.csv gpurun_out/${tag}_ncu_launches_ref.md "ncu launch list of the REFERENCE's CUDA extension ($tag): \`ncu --metrics gpu__time_duration.sum --clock-control none -s 200 -c 400 $CMD\`" > /dev/null 2>&1 ;;
    ncu_all)
      # one --set full capture of every kernel of the second pass of scripts/profile_all.py
      CMD="python scripts/profile_all.py"
      timeout 600 $CMD > gpurun_out/${tag}_profile_all.log 2>&1 &&
      timeout 1500 ncu --set full --clock-control none -k regex:'adam_kernel|bbox|build_boxes|composite|emit_kernel|gather_points|knn3|mark_visible|morton|normal_guidance|preprocess|radix|sh_grad_from_views|tile_' -s ${NCU_SKIP:-51} -c ${NCU_COUNT:-51} -f -o gpurun_out/${tag}_prof_all $CMD > gpurun_out/${tag}_ncu_all.log 2>&1
      echo "ncu all rc=$?" | tee -a $S
      python scripts/ncu_table.py gpurun_out/${tag}_prof_all.ncu-rep gpurun_out/${tag}_ncu_all.md "ncu --set full --clock-control none, every kernel of one pass over the 6 M scene ($tag): \`$CMD\`" > /dev/null 2>&1
      ls -la gpurun_out/${tag}_prof_all.ncu-rep >> $S
      # keep the merge under the 64 MiB limit
      if [ $(stat -c %s gpurun_out/${tag}_prof_all.ncu-rep 2>/dev/null || echo 0) -gt 40000000 ]; then rm -f gpurun_out/${tag}_prof_all.ncu-rep; fi ;;
  esac
done
cat $S
