"""GPU: one call of each round-2 kernel outside the step (f3 tensor-core path, f2 queue sums + backbone match) at the
bench shapes, for `ncu` captures.  Usage: python dev/prof_f2f3.py [iters]"""
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cp2_b200 import ops  # noqa: E402

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 9
dev = torch.device("cuda")
g = torch.Generator(device=dev).manual_seed(5)
B, C, hw, os_, K = 32, 128, 14, 16, 65536
P, H = hw * hw, hw * 16
q = torch.randn(B, C, P, device=dev, generator=g)
k = torch.randn(B, C, P, device=dev, generator=g)
ma = (torch.rand(B, P, device=dev, generator=g) < 0.65).float()
mb = (torch.rand(B, P, device=dev, generator=g) < 0.65).float()
pa = torch.arange(1, H * H + 1, device=dev).reshape(1, H, H).repeat(B, 1, 1)
pb = torch.roll(pa, shifts=(os_, os_), dims=(1, 2)).contiguous()
ra = (pa - 1) // (H * 64) * 4 + ((pa - 1) % H) // 64
rb = torch.roll(ra, shifts=(os_, os_), dims=(1, 2)).contiguous()
pw = ops.PairWeights(pa, pb, ra, rb, os_, 10.0, 2.0, 0.5, "FIXED", 2.0)
alpha = torch.full((B,), 1.0 / B, device=dev)
v1 = torch.zeros(B, C, device=dev)
qh, q_inv, sq, nsq, _ = ops.prep_features(q, ma)
kh, _, sk, nsk, _ = ops.prep_features(k, mb)
queue2 = F.normalize(torch.randn(C, K, device=dev, generator=g), dim=0)
shadow = ops.QueueShadow()
rows_bf = qh.view(B * P, C)
eq = F.normalize(torch.randn(B, 2048, P, device=dev, generator=g).relu(), dim=1)
ek = F.normalize(torch.randn(B, 2048, P, device=dev, generator=g).relu(), dim=1)
torch.cuda.synchronize()
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timed(name, fn, n=iters):
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
    ts.sort()
    print(f"{name:28s} {ts[len(ts) // 2]:9.1f} us (median of {n}, L2 flushed)", flush=True)
    return out


_, _, _, cnt, _, acc = timed("dense_weighted_small", lambda: ops.dense_weighted_small(qh, kh, ma, mb, pw, sq, nsq, sk, nsk, 5.0))
timed("dense_small_bwd", lambda: ops.dense_small_bwd(acc, qh, ma, q_inv, alpha, v1, None, key_count=cnt))
zp, gacc = timed("queue_sums_tc", lambda: ops.queue_sums_tc(rows_bf, queue2, 5.0, shadow))
idx = timed("densecl_match", lambda: ops.densecl_match(eq, ek))
(qh2, q_inv2, *_), (kh2, k_inv2, _, _, cnt2, _, _, _, perm2) = ops.prep_features_pair(q, ma, k, mb, want_perm=True)
timed("dense_small (no meter)", lambda: ops.dense_small(qh2, kh2, 5.0, cnt2))
timed("dense_small (acc_seg)", lambda: ops.dense_small(qh2, kh2, 5.0, cnt2, seg=(q, k, q_inv2, k_inv2, perm2, ma, mb)))
qpos = F.normalize(torch.randn(B, C, device=dev, generator=g), dim=1)
kpos = F.normalize(torch.randn(B, C, device=dev, generator=g), dim=1)
timed("global_nce (acc_ins)", lambda: ops.global_nce(qpos, queue2, 5.0, k_pos=kpos, shadow=shadow))
timed("prep_features_pair", lambda: ops.prep_features_pair(q, ma, k, mb))
# f3 at 28x28: the streaming tensor-core pair vs the SIMT pair
hw2, P2, H2 = 28, 784, 28 * 16
q2 = torch.randn(B, C, P2, device=dev, generator=g)
k2 = torch.randn(B, C, P2, device=dev, generator=g)
ma2 = (torch.rand(B, P2, device=dev, generator=g) < 0.65).float()
mb2 = (torch.rand(B, P2, device=dev, generator=g) < 0.65).float()
pa2 = torch.arange(1, H2 * H2 + 1, device=dev).reshape(1, H2, H2).repeat(B, 1, 1)
pb2 = torch.roll(pa2, shifts=(os_, os_), dims=(1, 2)).contiguous()
ra2 = (pa2 - 1) // (H2 * 64) * 4 + ((pa2 - 1) % H2) // 64
rb2 = torch.roll(ra2, shifts=(os_, os_), dims=(1, 2)).contiguous()
pw2 = ops.PairWeights(pa2, pb2, ra2, rb2, os_, 10.0, 2.0, 0.5, "FIXED", 2.0)
qh2b, q_inv2b, sq2, nsq2, _ = ops.prep_features(q2, ma2)
kh2b, _, sk2, nsk2, _ = ops.prep_features(k2, mb2)
lse_s, _, ws_s, _, _, ids_pad, b_pad = timed("weighted_stream fwd (28x28)", lambda: ops.dense_weighted_stream(qh2b, kh2b, ma2, mb2, pw2, sq2, nsq2, sk2, nsk2, 5.0))
timed("weighted_stream bwd (28x28)", lambda: ops.dense_weighted_stream_grad(qh2b, kh2b, ma2, mb2, pw2, lse_s, ws_s, ids_pad, b_pad, q_inv2b, alpha, v1, None, 5.0))
lse_m, _, ws_m, _, _ = timed("weighted SIMT fwd (28x28)", lambda: ops.dense_weighted_lse(qh2b, kh2b, ma2, mb2, pw2, sq2, nsq2, sk2, nsk2, 5.0))
timed("weighted SIMT bwd (28x28)", lambda: ops.dense_weighted_grad(qh2b, kh2b, ma2, mb2, pw2, lse_m, ws_m, q_inv2b, alpha, v1, None, 5.0))
print("ok", float(zp.sum()), int(idx.sum()))
