#!/bin/bash
mkdir -p gpurun_out
for cover in 0.15 0.5; do
  timeout 400 python bench.py --steps 10 --warmup 3 --no-sweep --no-cpu-baseline --no-eager-baseline --fold-cover $cover > gpurun_out/bench_fold_c$cover.json 2> gpurun_out/bench_fold_c$cover.err
  echo "== bench cover $cover exit $?"
  python - "gpurun_out/bench_fold_c$cover.json" <<'P'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(round(d['value'],1), round(d['ms_per_step'],3), round(d['e2e']['value'],1))
P
done
