"""K3 scaling probe: cp2_global_nce at several queue lengths (ncu target / event timing).  python dev/prof_global.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cp2_b200 import ops  # noqa: E402

dev = torch.device("cuda")
flush = torch.empty(1 << 30, dtype=torch.uint8, device=dev)
for B in (32, 256):
    for K in (4096, 16384, 65536, 262144, 1048576):
        q = torch.nn.functional.normalize(torch.randn(B, 128, device=dev), dim=1)
        queue = torch.nn.functional.normalize(torch.randn(128, K, device=dev), dim=0)
        ts = []
        for i in range(8):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); ops.global_nce(q, queue, 5.0); b.record()
            torch.cuda.synchronize()
            if i >= 3:
                ts.append(a.elapsed_time(b) * 1e3)
        t = sorted(ts)[len(ts) // 2]
        print(f"B={B} K={K}: {t:8.1f} us  -> {4 * 128 * K / t / 1e3:8.1f} GB/s", flush=True)
