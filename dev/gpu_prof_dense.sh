#!/bin/bash
mkdir -p gpurun_out
./dev/mb/microbench.bin > gpurun_out/microbench.txt 2>&1; echo "== microbench exit $?"; cat gpurun_out/microbench.txt
for pf in 0 2; do
  export CP2_DENSE_PF_FWD=$pf CP2_DENSE_PF_BWD=$pf
  CMD="python dev/prof_dense.py 3136 128 64 ones"
  timeout 200 $CMD > gpurun_out/plain_prof_pf$pf.log 2>&1 &&
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:'dense_(lse|grad)_kernel' -s 2 -c 2 \
     -o gpurun_out/prof_dense_pf$pf -f $CMD > gpurun_out/ncu_dense_pf$pf.log 2>&1
  echo "== ncu pf=$pf exit $?"
done
