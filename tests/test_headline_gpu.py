"""GPU parity at the HEADLINE shapes, against the fp64 CPU oracle (not against another CUDA kernel):

  * BASELINE.json configs[3]'s largest point, 56x56 maps (P = 3136), C in {128, 256}: `ops.dense_lse`,
    `ops.dense_grad` and the full `ops.cp2_loss`, with all-ones masks and with Bernoulli masks + compacted keys
    (reference builder.py:1289-1292, 1431-1437 and their autograd);
  * BASELINE.json configs[1] itself: B=32, P=196, C=128, K=65536, tau_d in {1.0, 0.2}, include_background both
    ways (builder.py:1260-1268, 1279-1292, 1392-1439), and K3 alone at (32, 65536) with |d log z| <= 1e-4 * loss.

Tolerances (north_star): losses 1e-4 relative; gradients rel-L2 <= 2e-2 (bf16 operands, fp32 accumulate).
The fp64 reference of one 3136 x 3136 image is 79 MB; everything here finishes in seconds on the host.
"""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import cp2_oracle as O

pytestmark = pytest.mark.gpu
LOSS_RTOL = 1e-4
GRAD_REL_L2 = 2e-2


@pytest.fixture(scope="module")
def ops():
    from cp2_b200 import _lib, ops as _ops
    _lib.load()
    return _ops


def rel_l2(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def _unit_bf16(B, P, C, seed):
    g = torch.Generator().manual_seed(seed)
    return F.normalize(torch.randn(B, P, C, generator=g), dim=2).to(torch.bfloat16)


def _masks(B, P, kind, seed):
    if kind == "ones":
        return torch.ones(B, P), torch.ones(B, P)
    g = torch.Generator().manual_seed(seed)
    return (torch.rand(B, P, generator=g) < 0.65).float(), (torch.rand(B, P, generator=g) < 0.65).float()


def _lse64(qh, kh, inv_tau):
    s = torch.einsum("nxc,nyc->nxy", qh.double(), kh.double()) * inv_tau
    return torch.logsumexp(s, dim=1)                      # over the QUERY axis (LogSoftmax(dim=1))


def _grad64(qh, kh, lse, ma, mb, inv_norm, alpha, v1, inv_tau):
    """fp64 restatement of the cp2_dense_grad contract (include/cp2_b200.h)."""
    qh, kh = qh.double(), kh.double()
    s = torch.einsum("nxc,nyc->nxy", qh, kh) * inv_tau
    p = torch.exp(s - lse.double()[:, None, :]) * mb.double()[:, None, :]
    acc = torch.einsum("nxy,nyc->nxc", p, kh)
    g = alpha.double()[:, None, None] * acc + ma.double()[:, :, None] * v1.double()[:, None, :]
    d = (g - qh * (qh * g).sum(-1, keepdim=True)) * inv_norm.double()[:, :, None]
    return d.transpose(1, 2)


@pytest.mark.parametrize("C", [128, 256])
@pytest.mark.parametrize("masks", ["ones", "bernoulli"])
def test_dense_kernels_vs_fp64_oracle_56x56(ops, C, masks):
    B, P, tau = 1, 3136, 0.2
    g = torch.Generator().manual_seed(C + 7)
    qh = _unit_bf16(B, P, C, 101)
    feat_k = torch.randn(B, C, P, generator=g)
    ma, mb = _masks(B, P, masks, 8)
    inv_norm = torch.rand(B, P, generator=g) + 0.5
    alpha = torch.rand(B, generator=g) + 0.1
    v1 = torch.randn(B, C, generator=g) * 0.01
    # key operand exactly as the step builds it: normalised, bf16, foreground rows first (cp2_prep_features)
    kh_full, _, _, _, _ = ops.prep_features(feat_k.cuda(), mb.cuda())
    kh_c, _, _, _, cnt = ops.prep_features(feat_k.cuda(), mb.cuda(), compact=True)
    assert int(cnt[0]) == int(mb.sum())
    lse = ops.dense_lse(qh.cuda(), kh_c, 1.0 / tau, cnt)
    dq = ops.dense_grad(qh.cuda(), kh_c, lse, ma.cuda(), None, inv_norm.cuda(), alpha.cuda(), v1.cuda(), None,
                        1.0 / tau, key_count=cnt)
    torch.cuda.synchronize()
    kf = kh_full.float().cpu()
    lse_ref = _lse64(qh.float(), kf, 1.0 / tau)                                  # [B,P] in original key order
    order = torch.cat([torch.nonzero(mb[0]).flatten(), torch.nonzero(mb[0] == 0).flatten()])
    n_fg = int(mb.sum())
    got = lse[0, :n_fg].double().cpu()
    ref = lse_ref[0, order[:n_fg]]
    # a quarter of the exponentials is a degree-3 polynomial (max rel. error 7.6e-5): |d lse| < 5e-5 on lse ~ 8
    assert torch.allclose(got, ref, rtol=0, atol=5e-5), float((got - ref).abs().max())
    gref = _grad64(qh.float(), kf, lse_ref.float(), ma, mb, inv_norm, alpha, v1, 1.0 / tau)
    assert rel_l2(dq, gref) < 5e-3, rel_l2(dq, gref)


@pytest.mark.parametrize("C,masks,tl", [(128, "ones", 0.2), (128, "bernoulli", 0.2), (256, "bernoulli", 1.0)])
def test_cp2_loss_vs_fp64_oracle_56x56(ops, C, masks, tl):
    B, hw, K = 2, 56, 4096
    P = hw * hw
    g = torch.Generator().manual_seed(C + 3)
    q = torch.randn(B, C, hw, hw, generator=g)
    k = torch.randn(B, C, hw, hw, generator=g)
    ma, mb = _masks(B, P, masks, 9)
    queue = F.normalize(torch.randn(C, K, generator=g), dim=0)
    qr = q.double().requires_grad_(True)
    r = O.cp2_loss(qr, k.double(), ma.double(), mb.double(), queue.double(), 0.2, tl, 0.2)
    r["loss"].backward()
    qd = q.cuda().requires_grad_(True)
    loss, li, ld, k_pos, q_pos, st = ops.cp2_loss(qd, k.cuda(), ma.cuda(), mb.cuda(), queue.cuda(), 0.2, tl, 0.2, False)
    loss.backward()
    np.testing.assert_allclose(ld.item(), r["loss_dense"].item(), rtol=LOSS_RTOL)
    np.testing.assert_allclose(li.item(), r["loss_ins"].item(), rtol=LOSS_RTOL)
    np.testing.assert_allclose(loss.item(), r["loss"].item(), rtol=LOSS_RTOL)
    assert rel_l2(qd.grad, qr.grad) < GRAD_REL_L2, rel_l2(qd.grad, qr.grad)


@pytest.mark.parametrize("tl", [1.0, 0.2])
@pytest.mark.parametrize("bg", [False, True])
def test_cp2_loss_at_configs1_shape(ops, tl, bg):
    """BASELINE.json configs[1]: B=32 per GPU, 14x14 map, C=128, K=65536, tau_g=0.2."""
    B, C, hw, K = 32, 128, 14, 65536
    v = O.synth_views(B, 224, 224, 16, seed=77)
    ma = O.stride_subsample(O.composite(v["img_a"], v["bg0"])[1], 16).reshape(B, -1)
    mb = O.stride_subsample(O.composite(v["img_b"], v["bg1"])[1], 16).reshape(B, -1)
    g = torch.Generator().manual_seed(5)
    q = torch.randn(B, C, hw, hw, generator=g)
    k = q * 0.5 + torch.randn(B, C, hw, hw, generator=g)          # correlated views: l_pos well above the negatives
    queue = F.normalize(torch.randn(C, K, generator=g), dim=0)
    qr = q.double().requires_grad_(True)
    r = O.cp2_loss(qr, k.double(), ma.double(), mb.double(), queue.double(), 0.2, tl, 0.2, include_background=bg)
    r["loss"].backward()
    qd = q.cuda().requires_grad_(True)
    loss, li, ld, k_pos, q_pos, st = ops.cp2_loss(qd, k.cuda(), ma.cuda(), mb.cuda(), queue.cuda(), 0.2, tl, 0.2, bg)
    loss.backward()
    np.testing.assert_allclose(li.item(), r["loss_ins"].item(), rtol=LOSS_RTOL)
    np.testing.assert_allclose(ld.item(), r["loss_dense"].item(), rtol=LOSS_RTOL)
    np.testing.assert_allclose(loss.item(), r["loss"].item(), rtol=LOSS_RTOL)
    assert rel_l2(qd.grad, qr.grad) < GRAD_REL_L2, rel_l2(qd.grad, qr.grad)
    assert torch.allclose(k_pos.cpu().double(), r["k_pos"], rtol=1e-5, atol=1e-6)


def test_global_nce_log_z_at_configs1_shape(ops):
    """K3 alone at (B, K) = (32, 65536): |d log z| <= 1e-4 * loss for every row, i.e. the queue pass alone stays
    inside the north-star loss tolerance (the loss of a row is log(z + e_pos) + (1 - l_pos)/tau >= log z)."""
    B, C, K, tau = 32, 128, 65536, 0.2
    g = torch.Generator().manual_seed(B + K)
    qp = F.normalize(torch.randn(B, C, generator=g), dim=1)
    queue = F.normalize(torch.randn(C, K, generator=g), dim=0)
    e = torch.exp((qp.double() @ queue.double() - 1.0) / tau)
    z_ref, g_ref = e.sum(dim=1), e @ queue.double().T
    z, gg = ops.global_nce(qp.cuda(), queue.cuda(), 1.0 / tau)[:2]
    torch.cuda.synchronize()
    loss_rows = torch.log(z_ref)                          # lower bound of the row loss (l_pos <= 1, e_pos >= 0)
    assert bool((loss_rows > 1).all())
    dlog = (torch.log(z.double().cpu()) - torch.log(z_ref)).abs()
    assert bool((dlog <= 1e-4 * loss_rows).all()), float((dlog / loss_rows).max())
    assert rel_l2(gg, g_ref) < 5e-3, rel_l2(gg, g_ref)
