#!/bin/bash
# ncu --set full of prep_kernel at B=256 for the map size $1 (default 3136)
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"prep_kernel" -s 6 -c 3 \
   -o gpurun_out/prof_prep -f python dev/time_dense.py ${1:-3136} > gpurun_out/ncu_prep.log 2>&1
echo "== ncu prep exit $?"
tail -3 gpurun_out/ncu_prep.log
