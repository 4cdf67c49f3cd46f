#!/bin/bash
# event timing + ncu durations (and with "full" one --set full capture) of the fused mirror kernel at B=32 / B=8, 512x512 labels
mkdir -p gpurun_out
python dev/time_mirror.py 32; python dev/time_mirror.py 8
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv -k regex:"mirror_up" -s 8 -c 4 \
   python dev/time_mirror.py 32 2>/dev/null | grep mirror_up | awk -F'","' '{print "ncu B=32:", $NF}'
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv -k regex:"mirror_up" -s 8 -c 4 \
   python dev/time_mirror.py 8 2>/dev/null | grep mirror_up | awk -F'","' '{print "ncu B=8:", $NF}'
if [ "$1" == "full" ]; then
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:"mirror_up" -s 4 -c 2 \
     -o gpurun_out/prof_mirror -f python dev/time_mirror.py 32 > gpurun_out/ncu_mirror.log 2>&1
  echo "== ncu mirror exit $?"
fi
