#!/bin/bash
# parity tests then the FMA-pipe share sweep of the dense kernels.  gpurun --timeout 1200 -- 'bash dev/gpu_dense.sh'
mkdir -p gpurun_out
bash dev/gpu_check.sh; rc=$?
timeout 300 python dev/tune_dense.py 3136 128 256 > gpurun_out/tune_3136_128.log 2>&1; echo "== tune exit $?"; cat gpurun_out/tune_3136_128.log | tail -12
timeout 300 python dev/tune_dense.py 3136 256 256 > gpurun_out/tune_3136_256.log 2>&1; echo "== tune exit $?"; cat gpurun_out/tune_3136_256.log | tail -12
timeout 300 python dev/tune_dense.py 196 128 256 > gpurun_out/tune_196_128.log 2>&1; echo "== tune exit $?"; cat gpurun_out/tune_196_128.log | tail -12
exit $rc
