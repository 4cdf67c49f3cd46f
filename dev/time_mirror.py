"""GPU: the fused mirror loss (cp2_mirror_loss_upsampled) at 512x512 labels, B from argv (default 32)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cp2_b200 import ops
B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
dev = torch.device("cuda")
g = torch.Generator(device=dev).manual_seed(3)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for lo in (32, 64):
    s = torch.randn(B, 2, lo, lo, device=dev, generator=g); t = torch.randn(B, 2, lo, lo, device=dev, generator=g)
    lab = torch.randint(0, 2, (B, 512, 512), device=dev, generator=g)
    ts = []
    for i in range(8):
        flush.zero_(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); ops.mirror_loss_upsampled_raw(s, t, lab, 0.01, 2.0, True); e1.record(); torch.cuda.synchronize()
        if i >= 3: ts.append(e0.elapsed_time(e1) * 1e3)
    print(f"B={B} logits {lo}x{lo}: {sum(ts)/len(ts):.1f} us", flush=True)
