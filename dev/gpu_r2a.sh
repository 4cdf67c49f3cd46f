#!/bin/bash
# Round-2 GPU trip A: parity tests (incl. the new headline-shape tests), smoke, bench (both arms), ncu of the
# all-ones 56x56 dense launch.   gpurun --timeout 1700 -- 'bash dev/gpu_r2a.sh'
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/test_gpu.log 2>&1; rc=$?
echo "== pytest exit $rc"; tail -n 25 gpurun_out/test_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "== smoke exit $?"; tail -n 3 gpurun_out/smoke.log
timeout 400 python bench.py --impl reference --steps 4 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "== ref exit $?"
tail -c 900 gpurun_out/bench_ref.json; tail -n 3 gpurun_out/bench_ref.err
timeout 1200 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "== bench exit $?"
python - <<'EOF'
import json
try:
    d = json.loads(open("gpurun_out/bench.json").read().strip().splitlines()[-1])
    for k in ("value", "ms_per_step", "e2e", "gpu_launches", "roofline", "cp2_kernels_us_per_step", "dense_loss", "gpu_eager_baseline", "cpu_baseline"):
        print(k, "=", json.dumps(d.get(k))[:900])
    for r in d.get("dense_microbench", []):
        print({k: r.get(k) for k in ("P", "C", "masks", "us", "frac", "frac_dense_algorithmic", "eager_fp32_us", "eager_bf16_bmm_us")})
    for k, v in d.get("kernels", {}).items():
        print(k, v["us_per_call"], v["calls_per_step"], v["frac"])
except Exception as e:
    print("bench parse failed", e)
EOF
tail -n 8 gpurun_out/bench.err
if [ "$1" != "nocu" ]; then
  SW="python bench.py --sweep-only --sweep-points 3136 --sweep-dims 128 --sweep-iters 1 --no-eager-baseline"
  timeout 300 $SW > gpurun_out/plain_sweep.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"dense_(lse|grad)_kernel" -s 2 -c 2 \
      -o gpurun_out/prof_dense_ones -f $SW > gpurun_out/ncu_dense_ones.log 2>&1
  echo "== ncu dense (all-ones) exit $?"
fi
exit $rc
