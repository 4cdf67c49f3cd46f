#!/bin/bash
# last check of the round: GPU tests, smoke, the default bench line (-> gpurun_out/bench.json)
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x -p no:cacheprovider 2>&1 | tail -3
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "== bench exit $?"
python - <<'P'
import json
d=json.loads(open("gpurun_out/bench.json").read().strip().splitlines()[-1])
print(round(d['value'],1), round(d['ms_per_step'],3), d['e2e'], d['config'].get('encoder_cuda_graphs'), d['gpu_launches'], d['clocks'])
print(d['gpu_eager_baseline'])
P
