#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_kernels_gpu.py -m gpu -q --tb=short -p no:cacheprovider -k top1 2>&1 | tail -n 3
export CP2_BENCH_CUDA_PROFILER=1
STEP="python bench.py --steps 2 --warmup 3 --no-sweep --no-cpu-baseline"
timeout 300 $STEP > gpurun_out/plain_step.log 2>&1 &&
timeout 600 ncu --profile-from-start off --set full --clock-control none --import-source on \
   -k regex:'top1_kernel' -c 4 -o gpurun_out/prof_top1 -f $STEP > gpurun_out/ncu_top1.log 2>&1
echo "== ncu top1 exit $?"
