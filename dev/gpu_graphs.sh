#!/bin/bash
# graph-captured encoders: the test, then the bench with and without graphs (short, no sweeps)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_loss_gpu.py -m gpu -q -x -p no:cacheprovider -k "graph or golden" 2>&1 | tail -15
for flag in "--no-graphs" ""; do
  timeout 600 python bench.py --no-sweep --no-cpu-baseline --no-eager-baseline $flag > gpurun_out/bench_graphs$flag.json 2> gpurun_out/bench_graphs$flag.err
  echo "== bench $flag exit $?"; tail -2 gpurun_out/bench_graphs$flag.err | cut -c1-300
  python - "gpurun_out/bench_graphs$flag.json" <<'P'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(round(d['value'],1), round(d['ms_per_step'],3), d['e2e']['value'], d['config'].get('encoder_cuda_graphs'), d['gpu_launches'])
P
done
