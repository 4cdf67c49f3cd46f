#!/bin/bash
# f2 checks: the densecl GPU tests, then the densecl microbench alone.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_densecl_gpu.py -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/test_densecl_gpu.log 2>&1
echo "== densecl tests exit $?"; tail -n 60 gpurun_out/test_densecl_gpu.log
timeout 600 python - > gpurun_out/densecl_bench.json 2> gpurun_out/densecl_bench.err <<'PY'
import json, bench
print(json.dumps(bench.densecl_sweep(bench.load_peaks())))
PY
echo "== densecl bench exit $?"; cat gpurun_out/densecl_bench.json; tail -n 5 gpurun_out/densecl_bench.err
