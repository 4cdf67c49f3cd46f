"""GPU: per-CTA wait-time accounting of cp2_dense_grad (cp2_dbg_dense_trace).  python dev/trace_dense.py P C B [pf]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cp2_b200 import _lib, ops  # noqa: E402

P, C, B = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
if len(sys.argv) > 4:
    os.environ["CP2_DENSE_PF_BWD"] = sys.argv[4]
lib = _lib.load()
dev = torch.device("cuda")
g = torch.Generator(device=dev).manual_seed(7)
q = torch.randn(B, C, P, device=dev, generator=g)
k = torch.randn(B, C, P, device=dev, generator=g)
ma = torch.ones(B, P, device=dev)
alpha = torch.full((B,), 1.0 / B, device=dev)
v1 = torch.zeros(B, C, device=dev)
qh, q_inv, _, _, _ = ops.prep_features(q, ma)
kh, _, _, _, cnt = ops.prep_features(k, ma, compact=True)
lse = ops.dense_lse(qh, kh, 5.0, cnt)
for _ in range(2):
    ops.dense_grad(qh, kh, lse, ma, None, q_inv, alpha, v1, None, 5.0, key_count=cnt)
torch.cuda.synchronize()
tr = torch.zeros(148 * 16, dtype=torch.int64, device=dev)
lib.cp2_dbg_dense_trace(tr.data_ptr())
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
ops.dense_grad(qh, kh, lse, ma, None, q_inv, alpha, v1, None, 5.0, key_count=cnt)
e1.record()
torch.cuda.synchronize()
lib.cp2_dbg_dense_trace(None)
t = tr.view(148, 16).double().cpu()
names = ["prod.q_empty", "prod.kempty", "mma.q_full", "mma.d_empty", "mma.kfull", "mma.p_full", "smx.kfull",
         "smx.s_full", "epi.d_full", "-", "life.prod", "life.smx", "life.epi"]
nxt = (P + 127) // 128
ty = 128 if C == 128 else 64
items = B * nxt / min(148, B * nxt)
tiles = items * ((P + ty - 1) // ty)
print(f"P={P} C={C} B={B}: {e0.elapsed_time(e1) * 1e3:.1f} us; items/CTA {items:.1f}, key tiles/CTA {tiles:.0f}")
for i, n in enumerate(names):
    if n != "-":
        print(f"  {n:14s} mean {t[:, i].mean():12.0f} cycles  max {t[:, i].max():12.0f}   per tile {t[:, i].mean() / tiles:8.1f}")
