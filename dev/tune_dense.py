"""GPU: times cp2_dense_lse / cp2_dense_grad separately over the share of exponentials moved to the FMA
pipe (CP2_DENSE_PF_FWD / _BWD, read by the library at every call).  Usage: python dev/tune_dense.py [P] [C] [B]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cp2_b200 import ops  # noqa: E402

P = int(sys.argv[1]) if len(sys.argv) > 1 else 3136
C = int(sys.argv[2]) if len(sys.argv) > 2 else 128
B = int(sys.argv[3]) if len(sys.argv) > 3 else 256
dev = torch.device("cuda")
g = torch.Generator(device=dev).manual_seed(7)
q = torch.randn(B, C, P, device=dev, generator=g)
k = torch.randn(B, C, P, device=dev, generator=g)
alpha = torch.full((B,), 1.0 / B, device=dev)
v1 = torch.zeros(B, C, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timed(fn, n=5):
    ts = []
    for i in range(n + 2):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = fn()
        e1.record()
        torch.cuda.synchronize()
        if i >= 2:
            ts.append(e0.elapsed_time(e1) * 1e3)
    return sum(ts) / len(ts), out


for variant in ("ones", "bern"):
    ma = torch.ones(B, P, device=dev) if variant == "ones" else (torch.rand(B, P, device=dev, generator=g) < 0.65).float()
    mb = torch.ones(B, P, device=dev) if variant == "ones" else (torch.rand(B, P, device=dev, generator=g) < 0.65).float()
    qh, q_inv, _, _, _ = ops.prep_features(q, ma)
    kh, _, _, _, cnt = ops.prep_features(k, mb, compact=True)
    ref_lse = ref_dq = None
    for pf in (0, 1, 2, 3):
        os.environ["CP2_DENSE_PF_FWD"] = str(pf)
        os.environ["CP2_DENSE_PF_BWD"] = str(pf)
        t_l, lse = timed(lambda: ops.dense_lse(qh, kh, 5.0, cnt))
        t_g, dq = timed(lambda: ops.dense_grad(qh, kh, lse, ma, None, q_inv, alpha, v1, None, 5.0, key_count=cnt))
        if pf == 0:
            ref_lse, ref_dq = lse.clone(), dq.clone()
        el = float((lse - ref_lse).abs().max())
        eg = float((dq - ref_dq).norm() / ref_dq.norm())
        flop = 4.0 * B * P * P * C
        print(f"P={P} C={C} B={B} masks={variant} pf={pf}: lse {t_l:8.1f} us  grad {t_g:8.1f} us  sum {t_l + t_g:8.1f} us "
              f"-> {flop / (t_l + t_g) / 1e6:7.1f} TFLOP/s algorithmic | d_lse {el:.2e} d_dq {eg:.2e}", flush=True)
