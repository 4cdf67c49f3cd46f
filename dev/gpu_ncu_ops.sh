#!/bin/bash
# ncu durations (cold cache, isolated) of the kernels dev/prof_f2f3.py launches -> gpurun_out/ncu_ops.csv
mkdir -p gpurun_out
timeout 900 ncu --metrics gpu__time_duration.sum,sm__cycles_active.avg --clock-control none --csv --log-file gpurun_out/ncu_ops.csv \
   -k regex:"dense_wsmall|dense_small|dense_grad_kernel|densecl_match|global_nce|prep_kernel|weighted_stats" python dev/prof_f2f3.py 1 > gpurun_out/ncu_ops.log 2>&1
echo "== ncu ops exit $?"
if [ "$1" == "full" ]; then
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"densecl_match|dense_small_kernel|global_nce" -c 6 \
     -o gpurun_out/prof_ops -f python dev/prof_f2f3.py 1 > gpurun_out/ncu_ops_full.log 2>&1
  echo "== ncu ops full exit $?"
fi
