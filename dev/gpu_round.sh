#!/bin/bash
# One GPU round trip: parity tests, smoke, bench (both arms), ncu launch list + full captures.
# Usage: gpurun --timeout 1500 -- 'bash dev/gpu_round.sh'
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
bash dev/gpu_check.sh; rc=$?
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "== smoke exit $?"; tail -n 3 gpurun_out/smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "== bench exit $?"
tail -c 6000 gpurun_out/bench.json; tail -n 5 gpurun_out/bench.err
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "== ref exit $?"
tail -c 600 gpurun_out/bench_ref.json
if [ "$1" != "nocu" ]; then
  export CP2_BENCH_CUDA_PROFILER=1
  STEP="python bench.py --steps 2 --warmup 3 --no-sweep --no-cpu-baseline"
  timeout 300 $STEP > gpurun_out/plain_step.log 2>&1 &&
  timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
      --log-file gpurun_out/launches_step.csv $STEP > gpurun_out/ncu_step.log 2>&1
  echo "== ncu launch list exit $?"
  timeout 300 $STEP > gpurun_out/plain_step2.log 2>&1 &&
  timeout 900 ncu --profile-from-start off --set full --clock-control none --import-source on \
      -k regex:'ema_kernel|global_nce_kernel|global_finalize|composite_kernel|prep_kernel|enqueue_kernel|dense_|top1' -c 20 \
      -o gpurun_out/prof_step -f $STEP > gpurun_out/ncu_step_full.log 2>&1
  echo "== ncu step full exit $?"
  unset CP2_BENCH_CUDA_PROFILER
  SW="python bench.py --sweep-only --sweep-points 3136 --sweep-dims 128 --sweep-iters 1"
  timeout 300 $SW > gpurun_out/plain_sweep.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"dense_(lse|grad)_kernel" -s 6 -c 2 \
      -o gpurun_out/prof_dense -f $SW > gpurun_out/ncu_dense.log 2>&1
  echo "== ncu dense exit $?"
fi
exit $rc
