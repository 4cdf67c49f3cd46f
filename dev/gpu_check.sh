#!/bin/bash
# Runs each GPU test file in its own process (a faulting kernel poisons the CUDA context) and
# keeps the logs under gpurun_out/.  Usage: gpurun --timeout 900 -- 'bash dev/gpu_check.sh'
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
rc=0
for f in tests/test_kernels_gpu.py tests/test_loss_gpu.py tests/test_mirror_gpu.py tests/test_weighted_gpu.py tests/test_densecl_gpu.py; do
  n=$(basename $f .py)
  timeout 600 python -m pytest $f -m gpu -q --tb=line -p no:cacheprovider > gpurun_out/$n.log 2>&1
  r=$?
  echo "== $f exit $r"; tail -n 40 gpurun_out/$n.log
  [ $r -ne 0 ] && rc=$r
done
exit $rc
