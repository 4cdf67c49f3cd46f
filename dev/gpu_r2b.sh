#!/bin/bash
# Round-2 GPU trip B: all GPU tests, smoke, both bench arms (no sweeps), ncu launch list + full capture of the step.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/test_gpu.log 2>&1; rc=$?
echo "== pytest exit $rc"; tail -n 40 gpurun_out/test_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "== smoke exit $?"; tail -n 3 gpurun_out/smoke.log
timeout 400 python bench.py --impl reference --steps 4 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "== ref exit $?"
tail -c 700 gpurun_out/bench_ref.json; tail -n 3 gpurun_out/bench_ref.err
timeout 900 python bench.py --steps 10 --warmup 3 --no-sweep > gpurun_out/bench_nosweep.json 2> gpurun_out/bench_nosweep.err; echo "== bench exit $?"
python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/bench_nosweep.json").read().strip().splitlines()[-1])
    for k in ("value", "ms_per_step", "e2e", "gpu_launches", "roofline", "cp2_kernels_us_per_step", "gpu_eager_baseline", "cpu_baseline"):
        print(k, "=", json.dumps(d.get(k))[:1200])
    for k, v in d.get("kernels", {}).items():
        print(k, v["us_per_call"], v["calls_per_step"], v["frac"])
except Exception as e:
    print("bench parse failed", e)
PY
tail -n 8 gpurun_out/bench_nosweep.err
if [ "$1" != "nocu" ]; then
  export CP2_BENCH_CUDA_PROFILER=1
  STEP="python bench.py --steps 2 --warmup 3 --no-sweep --no-cpu-baseline --no-eager-baseline"
  timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
      --log-file gpurun_out/launches_step.csv $STEP > gpurun_out/ncu_step.log 2>&1
  echo "== ncu launch list exit $?"
  timeout 900 ncu --profile-from-start off --set full --clock-control none --import-source on \
      -k regex:'ema_kernel|global_nce_kernel|composite_kernel|prep_kernel|enqueue_kernel|dense_|top1|loss_combine|id_iou|rescore' -c 24 \
      -o gpurun_out/prof_step -f $STEP > gpurun_out/ncu_step_full.log 2>&1
  echo "== ncu step full exit $?"
  unset CP2_BENCH_CUDA_PROFILER
fi
exit $rc
