#!/bin/bash
# tests of the small path, short step bench, ncu launch list of the step (durations of the library's kernels)
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider -x -k "small or loss or model" > gpurun_out/test_quick.log 2>&1
echo "== quick tests exit $?"; tail -n 4 gpurun_out/test_quick.log
export CP2_BENCH_CUDA_PROFILER=1
STEP="python bench.py --steps 2 --warmup 3 --no-sweep --no-cpu-baseline --no-eager-baseline"
timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
    -k regex:'ema_kernel|global_nce_kernel|composite_kernel|prep_kernel|enqueue_kernel|dense_|loss_combine|id_iou' \
    --log-file gpurun_out/launches_lib.csv $STEP > gpurun_out/ncu_lib.log 2>&1
echo "== ncu exit $?"
