#!/bin/bash
# SM clock / power while the 56x56 dense pair runs back to back
nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,power.draw,power.limit,clocks_throttle_reasons.active --format=csv,noheader -lms 200 > gpurun_out/clk.log 2>&1 &
SMI=$!
python - <<'PY'
import torch, time, sys, os
sys.path.insert(0, os.getcwd())
from cp2_b200 import ops
P, C, B = 3136, 128, 256
dev = torch.device("cuda")
g = torch.Generator(device=dev).manual_seed(7)
q = torch.randn(B, C, P, device=dev, generator=g); k = torch.randn(B, C, P, device=dev, generator=g)
alpha = torch.full((B,), 1.0 / B, device=dev); v1 = torch.zeros(B, C, device=dev)
ma = torch.ones(B, P, device=dev); mb = torch.ones(B, P, device=dev)
qh, q_inv, _, _, _ = ops.prep_features(q, ma); kh, _, _, _, cnt = ops.prep_features(k, mb, compact=True)
torch.cuda.synchronize(); time.sleep(1.0)
for name, fn in (("lse", lambda: ops.dense_lse(qh, kh, 5.0, cnt)),):
    t0 = time.time(); n = 0
    while time.time() - t0 < 3.0:
        lse = fn(); n += 1
        if n % 50 == 0: torch.cuda.synchronize()
    torch.cuda.synchronize(); print(name, "avg us", (time.time() - t0) / n * 1e6, flush=True)
time.sleep(0.5)
t0 = time.time(); n = 0
while time.time() - t0 < 3.0:
    dq = ops.dense_grad(qh, kh, lse, ma, None, q_inv, alpha, v1, None, 5.0, key_count=cnt); n += 1
    if n % 50 == 0: torch.cuda.synchronize()
torch.cuda.synchronize(); print("grad avg us", (time.time() - t0) / n * 1e6, flush=True)
PY
kill $SMI
awk 'NR%3==0' gpurun_out/clk.log | head -40
