"""GPU: cp2_dense_lse / cp2_dense_grad of the PRODUCT library at the headline shape (B=256, P=3136, C=128)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cp2_b200 import ops
P, C, B = 3136, int(sys.argv[1]) if len(sys.argv) > 1 else 128, 256
dev = torch.device("cuda")
g = torch.Generator(device=dev).manual_seed(7)
q = torch.randn(B, C, P, device=dev, generator=g); k = torch.randn(B, C, P, device=dev, generator=g)
alpha = torch.full((B,), 1.0 / B, device=dev); v1 = torch.zeros(B, C, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def timed(fn, n=5):
    ts = []
    for i in range(n + 2):
        flush.zero_(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); out = fn(); e1.record(); torch.cuda.synchronize()
        if i >= 2: ts.append(e0.elapsed_time(e1) * 1e3)
    return sum(ts) / len(ts), out
for masks in ("ones", "bern"):
    ma = torch.ones(B, P, device=dev) if masks == "ones" else (torch.rand(B, P, device=dev, generator=g) < 0.65).float()
    mb = torch.ones(B, P, device=dev) if masks == "ones" else (torch.rand(B, P, device=dev, generator=g) < 0.65).float()
    qh, q_inv, _, _, _ = ops.prep_features(q, ma); kh, _, _, _, cnt = ops.prep_features(k, mb, compact=True)
    tl, lse = timed(lambda: ops.dense_lse(qh, kh, 5.0, cnt))
    tg, dq = timed(lambda: ops.dense_grad(qh, kh, lse, ma, None, q_inv, alpha, v1, None, 5.0, key_count=cnt))
    print(f"C={C} masks={masks}: lse {tl:8.1f} us  grad {tg:8.1f} us  sum {tl+tg:8.1f} us -> frac {4.0*B*P*P*C/(tl+tg)/1e6/1669.2:.4f}", flush=True)
