#!/bin/bash
mkdir -p gpurun_out
bash dev/gpu_check.sh; rc=$?
timeout 300 python dev/tune_dense.py 3136 128 256 > gpurun_out/tune_3136_128.log 2>&1; echo "== tune exit $?"; tail -8 gpurun_out/tune_3136_128.log
timeout 300 python dev/tune_dense.py 3136 256 256 > gpurun_out/tune_3136_256.log 2>&1; echo "== tune exit $?"; tail -8 gpurun_out/tune_3136_256.log
timeout 300 python dev/tune_dense.py 196 128 256 > gpurun_out/tune_196_128.log 2>&1; echo "== tune exit $?"; tail -8 gpurun_out/tune_196_128.log
rm -f gpurun_out/trace.txt
for a in "196 128 256" "3136 128 64 0" "3136 128 64 2"; do timeout 200 python dev/trace_dense.py $a 2>&1 | tee -a gpurun_out/trace.txt; done
exit $rc
