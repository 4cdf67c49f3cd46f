#!/bin/bash
# Mirror-variant parity tests + microbenchmark.  Usage: gpurun --timeout 900 -- 'bash dev/gpu_mirror.sh'
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_mirror_gpu.py -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/test_mirror_gpu.log 2>&1
rc=$?; echo "== mirror tests exit $rc"; tail -n 40 gpurun_out/test_mirror_gpu.log
timeout 300 python dev/bench_mirror.py > gpurun_out/mirror_bench.log 2>&1; echo "== mirror bench exit $?"; tail -n 8 gpurun_out/mirror_bench.log
exit $rc
