"""One warm-up + one profiled pass of prep -> dense_lse -> dense_grad (ncu target).  python dev/prof_dense.py P C B masks"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cp2_b200 import ops  # noqa: E402

P, C, B = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
variant = sys.argv[4] if len(sys.argv) > 4 else "ones"
dev = torch.device("cuda")
g = torch.Generator(device=dev).manual_seed(7)
q = torch.randn(B, C, P, device=dev, generator=g)
k = torch.randn(B, C, P, device=dev, generator=g)
ma = torch.ones(B, P, device=dev) if variant == "ones" else (torch.rand(B, P, device=dev, generator=g) < 0.65).float()
mb = torch.ones(B, P, device=dev) if variant == "ones" else (torch.rand(B, P, device=dev, generator=g) < 0.65).float()
alpha = torch.full((B,), 1.0 / B, device=dev)
v1 = torch.zeros(B, C, device=dev)
for _ in range(2):
    qh, q_inv, _, _, _ = ops.prep_features(q, ma)
    kh, _, _, _, cnt = ops.prep_features(k, mb, compact=True)
    lse = ops.dense_lse(qh, kh, 5.0, cnt)
    dq = ops.dense_grad(qh, kh, lse, ma, None, q_inv, alpha, v1, None, 5.0, key_count=cnt)
torch.cuda.synchronize()
print("ok", float(lse.mean()), float(dq.abs().mean()))
