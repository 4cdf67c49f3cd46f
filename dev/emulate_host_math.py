"""Dev check (CPU): replaces the C-ABI kernels by torch restatements of their CONTRACTS
(include/cp2_b200.h) and verifies that the host algebra in ops.CP2LossFunction reproduces the
oracle's loss and autograd gradient.  Not a test of the kernels; not used by the product."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cp2_b200 import ops
from oracle import cp2_oracle as O

def prep(feat, mask):
    n = feat.norm(dim=1, keepdim=True).clamp_min(1e-12)
    fh = feat / n
    return fh.transpose(1, 2).contiguous(), (1.0 / n).squeeze(1), torch.einsum("ncx,nx->nc", fh, mask), torch.einsum("ncx,nx->nc", fh, 1 - mask)
def lse(qh, kh, it):
    return torch.logsumexp(torch.einsum("nxc,nyc->nxy", qh, kh) * it, dim=1)
def gnce(qp, queue, it):
    e = torch.exp((qp @ queue - 1.0) * it)
    return e.sum(1), e @ queue.T
def dgrad(qh, kh, l, ma, mb, inr, alpha, v1, v2, it):
    s = torch.einsum("nxc,nyc->nxy", qh, kh) * it
    p = torch.exp(s - l[:, None, :]) * mb[:, None, :]
    acc = torch.einsum("nxy,nyc->nxc", p, kh)
    g = alpha[:, None, None] * acc + ma[:, :, None] * v1[:, None, :]
    if v2 is not None: g = g + (1 - ma)[:, :, None] * v2[:, None, :]
    return ((g - qh * (qh * g).sum(-1, keepdim=True)) * inr[:, :, None]).transpose(1, 2)
ops.prep_features, ops.dense_lse, ops.global_nce, ops.dense_grad = prep, lse, gnce, dgrad
ops._f32c = lambda t: t.contiguous()
torch.set_default_dtype(torch.float64)
for bg in (False, True):
    for tl in (1.0, 0.2):
        g = torch.Generator().manual_seed(1)
        B, C, hw, K = 3, 32, 5, 64
        q = torch.randn(B, C, hw, hw, generator=g); k = torch.randn(B, C, hw, hw, generator=g)
        ma = (torch.rand(B, hw*hw, generator=g) < .6).double(); mb = (torch.rand(B, hw*hw, generator=g) < .6).double()
        queue = torch.nn.functional.normalize(torch.randn(C, K, generator=g), dim=0)
        qr = q.clone().requires_grad_(True)
        r = O.cp2_loss(qr, k, ma, mb, queue, 0.2, tl, 0.3, include_background=bg)
        (r["loss"] * 1.7).backward()
        qd = q.clone().requires_grad_(True)
        out = ops.cp2_loss(qd, k, ma, mb, queue, 0.2, tl, 0.3, bg)
        (out[0] * 1.7).backward()
        print(bg, tl, float(out[0] - r["loss"]), float(out[1] - r["loss_ins"]), float(out[2] - r["loss_dense"]),
              float((qd.grad - qr.grad).norm() / qr.grad.norm()))
