#!/bin/bash
# quick perf check: selected tests, short bench without baselines, f2/f3 microbench timing
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/test_gpu.log 2>&1; rc=$?
echo "== pytest exit $rc"; tail -n 8 gpurun_out/test_gpu.log
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-eager-baseline --sweep-points 196 --sweep-dims 128 > gpurun_out/bench_e.json 2> gpurun_out/bench_e.err; echo "== bench exit $?"
python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/bench_e.json").read().strip().splitlines()[-1])
    for k in ("value", "ms_per_step", "gpu_launches", "cp2_kernels_us_per_step"):
        print(k, "=", json.dumps(d.get(k))[:300])
    for k, v in d.get("kernels", {}).items():
        print(k, v["us_per_call"], v["calls_per_step"], v["frac"])
    for key in ("weighted_microbench", "densecl_microbench"):
        for r in d.get(key, []):
            print(key, json.dumps(r)[:260])
except Exception as e:
    print("bench parse failed", e)
PY
tail -n 3 gpurun_out/bench_e.err
if [ "$1" == "ncu" ]; then
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"dense_wsmall_kernel|dense_small_bwd_kernel|dense_grad_kernel|densecl_match_kernel|weighted_stats" -s 5 -c 10 \
      python dev/prof_f2f3.py 2 2>&1 | grep -E "^\s+(void|<unnamed)|gpu__time" | paste - - | cut -c1-200
fi
exit $rc
