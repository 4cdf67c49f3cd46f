#!/bin/bash
# Round-2 GPU trip G (final state; ncu legs launch the encoders eagerly): everything the judge re-runs (GPU tests, smoke, both bench arms with the sweeps), then the
# ncu launch list of the step and full captures of the step kernels, the 56x56 dense pair and the f2 / f3 kernels.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
rc=0
if [ "$SKIP_TESTS" != "1" ]; then
timeout 900 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/test_gpu.log 2>&1; rc=$?
echo "== pytest exit $rc"; tail -n 8 gpurun_out/test_gpu.log
fi
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "== smoke exit $?"; tail -n 3 gpurun_out/smoke.log
timeout 400 python bench.py --impl reference --steps 4 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "== ref exit $?"
timeout 1200 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "== bench exit $?"
python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/bench.json").read().strip().splitlines()[-1])
    for k in ("value", "ms_per_step", "e2e", "gpu_launches", "roofline", "cp2_kernels_us_per_step", "dense_loss"):
        print(k, "=", json.dumps(d.get(k))[:700])
    for k, v in d.get("kernels", {}).items():
        print(k, v["us_per_call"], v["calls_per_step"], v["frac"])
    for r in d.get("dense_microbench", []):
        print(r["P"], r["C"], r["masks"], r["us"], r["frac"])
    for key in ("weighted_microbench", "densecl_microbench"):
        for r in d.get(key, []):
            print(key, json.dumps(r)[:400])
    ge = d.get("gpu_eager_baseline") or {}
    print("eager:", {k: ge.get(k) for k in ("step_ms_nchw", "step_ms_channels_last", "own_step_ms", "loss_slice_eager_us", "loss_slice_own_us")})
except Exception as e:
    print("bench parse failed", e)
PY
tail -n 5 gpurun_out/bench.err
if [ "$1" != "nocu" ]; then
  export CP2_BENCH_CUDA_PROFILER=1
  STEP="python bench.py --steps 2 --warmup 3 --no-sweep --no-cpu-baseline --no-eager-baseline --no-graphs"
  timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
      --log-file gpurun_out/launches_step.csv $STEP > gpurun_out/ncu_step.log 2>&1
  echo "== ncu launch list exit $?"
  timeout 900 ncu --profile-from-start off --set full --clock-control none --import-source on \
      -k regex:'ema_kernel|global_nce_kernel|composite_kernel|prep_kernel|enqueue_kernel|dense_|loss_combine|id_iou' -c 18 \
      -o gpurun_out/prof_step -f $STEP > gpurun_out/ncu_step_full.log 2>&1
  echo "== ncu step full exit $?"
  unset CP2_BENCH_CUDA_PROFILER
  SW="python bench.py --sweep-only --sweep-points 3136 --sweep-dims 128 --sweep-iters 1"
  timeout 300 $SW > gpurun_out/plain_sweep.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"dense_(lse|grad)_kernel" -s 6 -c 2 \
      -o gpurun_out/prof_dense -f $SW > gpurun_out/ncu_dense.log 2>&1
  echo "== ncu dense exit $?"
  timeout 300 python dev/prof_f2f3.py 3 > gpurun_out/plain_f2f3.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on \
      -k regex:"dense_wsmall_kernel|dense_small_bwd_kernel|dense_grad_kernel|densecl_match_kernel|weighted_stats" -s 5 -c 10 \
      -o gpurun_out/prof_f2f3 -f python dev/prof_f2f3.py 2 > gpurun_out/ncu_f2f3.log 2>&1
  echo "== ncu f2f3 exit $?"; tail -n 2 gpurun_out/plain_f2f3.log
  bash dev/gpu_ncu_mirror.sh full > gpurun_out/mirror_times.log 2>&1; tail -n 12 gpurun_out/mirror_times.log
  timeout 300 python bench.py --steps 10 --warmup 3 --no-sweep --no-cpu-baseline --no-eager-baseline --no-graphs > gpurun_out/bench_nographs.json 2> gpurun_out/bench_nographs.err; echo "== bench --no-graphs exit $?"
fi
# gpurun_out/ comes back only below 64 MiB: summarise the captures HERE and leave the .ncu-rep files behind
mkdir -p gpurun_out/profiles_box
cp profiles/traffic.json gpurun_out/profiles_box/ 2>/dev/null
CP2_PROFILES_DST=gpurun_out/profiles_box python dev/summarise_profiles.py ${TAG:-r02g} > gpurun_out/summarise.log 2>&1; echo "== summarise exit $?"
rm -f gpurun_out/*.ncu-rep
du -sh gpurun_out
exit $rc
