#!/bin/bash
# A/B of the dilated-convolution folding (short bench runs, no sweeps / baselines)
mkdir -p gpurun_out
for flag in ${FLAGS:-"--no-fold" "" "--fold-cover 0.6"}; do
  tag=$(echo "$flag" | tr -d ' -.')
  timeout 400 python bench.py --steps 10 --warmup 3 --no-sweep --no-cpu-baseline --no-eager-baseline $flag > gpurun_out/bench_fold_$tag.json 2> gpurun_out/bench_fold_$tag.err
  echo "== bench [$flag] exit $?"; grep -v "Warning\|warn\|run_backward\|^$" gpurun_out/bench_fold_$tag.err | tail -n 3 | cut -c1-300
  python - "gpurun_out/bench_fold_$tag.json" <<'P'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(round(d['value'],1), round(d['ms_per_step'],3), round(d['e2e']['value'],1), d['config'].get('encoder_cuda_graphs'))
P
done
