"""Microbenchmark of the mirror-variant loss (BASELINE configs[4] shapes): fused kernels vs the eager
torch formulation of networks/mirror_network.py:47-63 on the same GPU.  CUDA events, L2 flushed."""
import json
import sys
import os

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cp2_b200 import ops  # noqa: E402


def timeit(fn, iters=30, warm=5):
    flush = torch.empty(1 << 30, dtype=torch.uint8, device="cuda")   # ~330 us of memset hides the CPU launch path
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def eager(s_lo, t_lo, masks, lmbd, temp):
    s = F.interpolate(s_lo, size=masks.shape[-2:], mode="bilinear", align_corners=False)
    t = F.interpolate(t_lo, size=masks.shape[-2:], mode="bilinear", align_corners=False)
    cmp = F.cross_entropy(torch.softmax(s / temp, 1), torch.softmax(t / temp, 1))
    cls = F.cross_entropy(torch.cat([s, t]), torch.cat([masks, masks]))
    return cls + lmbd * cmp


def main():
    rows = []
    for B, Cls, h, H in [(8, 2, 32, 512), (8, 2, 64, 512), (32, 2, 32, 512), (32, 2, 64, 512)]:
        s_lo = torch.randn(B, Cls, h, h, device="cuda", requires_grad=True)
        t_lo = torch.randn(B, Cls, h, h, device="cuda", requires_grad=True)
        masks = torch.randint(0, Cls, (B, H, H), device="cuda")
        s_hi = F.interpolate(s_lo.detach(), size=(H, H), mode="bilinear", align_corners=False)
        t_hi = F.interpolate(t_lo.detach(), size=(H, H), mode="bilinear", align_corners=False)

        def ref():
            s_lo.grad = t_lo.grad = None
            eager(s_lo, t_lo, masks, 0.01, 2.0).backward()

        up = timeit(lambda: ops.mirror_loss_upsampled_raw(s_lo.detach(), t_lo.detach(), masks, 0.01, 2.0, True))
        up_fwd = timeit(lambda: ops.mirror_loss_upsampled_raw(s_lo.detach(), t_lo.detach(), masks, 0.01, 2.0, False))
        full = timeit(lambda: ops.mirror_loss_raw(s_hi, t_hi, masks, 0.01, 2.0, True))
        full_fwd = timeit(lambda: ops.mirror_loss_raw(s_hi, t_hi, masks, 0.01, 2.0, False))
        e = timeit(ref, iters=10, warm=3)
        px = B * H * H
        rows.append({"B": B, "Cls": Cls, "lo": h, "hi": H,
                     "up_fwd_bwd_us": round(up, 2), "up_fwd_us": round(up_fwd, 2),
                     "up_label_GBps": round(px * 8 / up / 1e3, 1),
                     "full_fwd_bwd_us": round(full, 2), "full_fwd_us": round(full_fwd, 2),
                     "full_GBps": round(px * (8 + 16 * Cls) / full / 1e3, 1),
                     "full_fwd_GBps": round(px * (8 + 8 * Cls) / full_fwd / 1e3, 1),
                     "eager_fwd_bwd_us": round(e, 1), "speedup_vs_eager": round(e / up, 1)})
        print(json.dumps(rows[-1]), flush=True)
    json.dump(rows, open("gpurun_out/mirror_bench.json", "w"), indent=1)


if __name__ == "__main__":
    main()
