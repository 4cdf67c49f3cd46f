"""GPU: times cp2_dense_grad of the DEBUG library (libcp2_b200_debug.so honours CP2_DENSE_BWD_VARIANT / CP2_DENSE_PF_BWD)
at the headline shape and checks the variants against each other.  Usage: python dev/tune_bwd.py [P] [B]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cp2_b200 import _lib, ops  # noqa: E402

P = int(sys.argv[1]) if len(sys.argv) > 1 else 3136
B = int(sys.argv[2]) if len(sys.argv) > 2 else 256
C = 128
dev = torch.device("cuda")
dbg = _lib.load_debug()
g = torch.Generator(device=dev).manual_seed(7)
q = torch.randn(B, C, P, device=dev, generator=g)
k = torch.randn(B, C, P, device=dev, generator=g)
alpha = torch.full((B,), 1.0 / B, device=dev)
v1 = torch.zeros(B, C, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
ptr = lambda t: None if t is None else t.data_ptr()
st = lambda: torch.cuda.current_stream().cuda_stream


def grad(qh, kh, lse, ma, q_inv, cnt):
    dq = torch.empty((B, C, P), device=dev, dtype=torch.float32)
    nl = torch.empty((B, dbg.cp2_dense_pad(P)), device=dev, dtype=torch.float32)
    rc = dbg.cp2_dense_grad(ptr(qh), ptr(kh), ptr(cnt), ptr(lse), ptr(ma), None, ptr(q_inv), ptr(alpha), ptr(v1), None, None,
                            ptr(dq), ptr(nl), 5.0, B, C, P, st())
    assert rc == 0, dbg.cp2_last_error()
    return dq


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


for masks in ("ones", "bern"):
    ma = torch.ones(B, P, device=dev) if masks == "ones" else (torch.rand(B, P, device=dev, generator=g) < 0.65).float()
    mb = torch.ones(B, P, device=dev) if masks == "ones" else (torch.rand(B, P, device=dev, generator=g) < 0.65).float()
    qh, q_inv, _, _, _ = ops.prep_features(q, ma)
    kh, _, _, _, cnt = ops.prep_features(k, mb, compact=True)
    lse = ops.dense_lse(qh, kh, 5.0, cnt)
    ref = None
    for nsw in (8, 4):
        for pf in (0, 1, 2, 3):
            os.environ["CP2_DENSE_NSW_FWD"] = str(nsw)
            os.environ["CP2_DENSE_PF_FWD"] = str(pf)
            lse_t = torch.empty((B, P), device=dev, dtype=torch.float32)

            def fwd():
                rc = dbg.cp2_dense_lse(ptr(qh), ptr(kh), ptr(cnt), ptr(lse_t), 5.0, B, C, P, st())
                assert rc == 0, dbg.cp2_last_error()
                return lse_t
            t, out = timed(fwd)
            print(f"P={P} B={B} masks={masks} fwd nsw={nsw} pf={pf}: lse {t:8.1f} us   max|d lse| {float((out - lse).abs().max()):.2e}", flush=True)
    for variant in (0,):
        for pf in (0, 1, 2, 3):
            os.environ["CP2_DENSE_BWD_VARIANT"] = str(variant)
            os.environ["CP2_DENSE_PF_BWD"] = str(pf)
            t, dq = timed(lambda: grad(qh, kh, lse, ma, q_inv, cnt))
            if ref is None:
                ref = dq.clone()
            err = float((dq - ref).norm() / ref.norm())
            print(f"P={P} B={B} masks={masks} variant={variant} pf={pf}: grad {t:8.1f} us   d_dq vs first {err:.2e}", flush=True)
