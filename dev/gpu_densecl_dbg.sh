#!/bin/bash
mkdir -p gpurun_out
for i in 1 2 3; do python -c "import bench, json; print(json.dumps(bench.densecl_sweep(bench.load_peaks(), iters=10))[:160])"; done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_densecl.csv \
  python -c "import bench; bench.densecl_sweep(bench.load_peaks(), iters=1)" > gpurun_out/ncu_densecl.log 2>&1
echo "== ncu exit $?"
timeout 600 python -m pytest tests/test_mirror_gpu.py -m gpu -q -p no:cacheprovider 2>&1 | tail -3
python bench.py --workload mirror --steps 5 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('mirror', d['value'], d['ms_per_step'], json.dumps(d.get('kernels'))[:300])"
python -c "import bench, json; [print(json.dumps(r)) for r in bench.mirror_sweep(bench.load_peaks())]"
