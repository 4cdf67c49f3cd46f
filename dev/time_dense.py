"""GPU: prep / cp2_dense_lse / cp2_dense_grad of the PRODUCT library at B=256 for the map sizes on argv (C = $C or 128)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cp2_b200 import ops
C, B = int(os.environ.get("C", "128")), int(os.environ.get("B", "256"))
dev = torch.device("cuda")
g = torch.Generator(device=dev).manual_seed(7)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def timed(fn, n=7):
    ts = []
    for i in range(n + 3):
        flush.zero_(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); out = fn(); e1.record(); torch.cuda.synchronize()
        if i >= 3: ts.append(e0.elapsed_time(e1) * 1e3)
    return sum(ts) / len(ts), out
for P in [int(a) for a in sys.argv[1:]] or [784]:
    q = torch.randn(B, C, P, device=dev, generator=g); k = torch.randn(B, C, P, device=dev, generator=g)
    alpha = torch.full((B,), 1.0 / B, device=dev); v1 = torch.zeros(B, C, device=dev)
    for masks in ("ones", "bern"):
        ma = torch.ones(B, P, device=dev) if masks == "ones" else (torch.rand(B, P, device=dev, generator=g) < 0.65).float()
        mb = torch.ones(B, P, device=dev) if masks == "ones" else (torch.rand(B, P, device=dev, generator=g) < 0.65).float()
        tp, _ = timed(lambda: ops.prep_features_pair(q, ma, k, mb))
        tq, _ = timed(lambda: ops.prep_features(q, ma))
        tk, _ = timed(lambda: ops.prep_features(k, mb, compact=True))
        qh, q_inv, _, _, _ = ops.prep_features(q, ma); kh, _, _, _, cnt = ops.prep_features(k, mb, compact=True)
        tl, lse = timed(lambda: ops.dense_lse(qh, kh, 5.0, cnt))
        tg, dq = timed(lambda: ops.dense_grad(qh, kh, lse, ma, None, q_inv, alpha, v1, None, 5.0, key_count=cnt))
        gb = 2 * (6.0 * B * C * P + 4.0 * B * P) / 1e3
        print(f"P={P} C={C} B={B} masks={masks}: prep pair {tp:7.1f} us ({gb/tp:6.0f} GB/s; q {tq:6.1f} k {tk:6.1f})  lse {tl:8.1f} us  "
              f"grad {tg:8.1f} us  sum {tl+tg:8.1f} us -> frac {4.0*B*P*P*C/(tl+tg)/1e6/1669.2:.4f}", flush=True)
