#!/bin/bash
mkdir -p gpurun_out
for a in "196 128 256" "3136 128 64 0" "3136 128 64 2" "3136 256 64 0"; do
  timeout 200 python dev/trace_dense.py $a 2>&1 | tee -a gpurun_out/trace.txt
done
