#!/bin/bash
# event timing + one ncu --set full capture of mirror_up_kernel at B=32, 512x512 labels
mkdir -p gpurun_out
python dev/time_mirror.py 32
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"mirror_up_kernel" -s 4 -c 2 \
   -o gpurun_out/prof_mirror -f python dev/time_mirror.py 32 > gpurun_out/ncu_mirror.log 2>&1
echo "== ncu mirror exit $?"
