#!/bin/bash
# Round-2 GPU trip C: the new / changed kernels first (small dense path, K3 prologue), then everything, then a short
# bench + sweep at the small maps, then an ncu capture of the small-path kernels.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -q --tb=short -p no:cacheprovider -k "small or global" > gpurun_out/test_new.log 2>&1
echo "== new tests exit $?"; tail -n 25 gpurun_out/test_new.log
timeout 900 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/test_gpu.log 2>&1; rc=$?
echo "== pytest exit $rc"; tail -n 15 gpurun_out/test_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "== smoke exit $?"; tail -n 3 gpurun_out/smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-eager-baseline --sweep-points 196,784 --sweep-dims 128 > gpurun_out/bench_c.json 2> gpurun_out/bench_c.err; echo "== bench exit $?"
python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/bench_c.json").read().strip().splitlines()[-1])
    for k in ("value", "ms_per_step", "e2e", "gpu_launches", "cp2_kernels_us_per_step"):
        print(k, "=", json.dumps(d.get(k))[:600])
    for k, v in d.get("kernels", {}).items():
        print(k, v["us_per_call"], v["calls_per_step"], v["frac"])
    for r in d.get("dense_microbench", []):
        print(r["P"], r["C"], r["masks"], r["us"], r["frac"])
except Exception as e:
    print("bench parse failed", e)
PY
tail -n 5 gpurun_out/bench_c.err
if [ "$1" != "nocu" ]; then
  export CP2_BENCH_CUDA_PROFILER=1
  STEP="python bench.py --steps 2 --warmup 3 --no-sweep --no-cpu-baseline --no-eager-baseline"
  timeout 900 ncu --profile-from-start off --set full --clock-control none --import-source on \
      -k regex:'global_nce_kernel|dense_small|prep_kernel|loss_combine|id_iou|dense_top1' -c 14 \
      -o gpurun_out/prof_small -f $STEP > gpurun_out/ncu_small.log 2>&1
  echo "== ncu small exit $?"
  unset CP2_BENCH_CUDA_PROFILER
fi
exit $rc
