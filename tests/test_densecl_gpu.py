"""GPU parity of the forward_densecl loss path (reference builder.py:667-999):
(a) against golden vectors recorded from the reference's own `forward_densecl` (DENSECL and PROPOSED_V2),
(b) the two new C-ABI entries against the CPU oracle on seeded inputs, (c) the drop-in MODEL in these modes.

Tolerances: loss 1e-4 relative; gradients rel-L2 <= 2e-2 (cp2_global_nce reads the queue as bf16).
"""
import numpy as np
import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

from conftest import T
from oracle import cp2_oracle as O
from test_oracle_golden import DENSECL_TAGS, densecl_case

pytestmark = pytest.mark.gpu
LOSS_RTOL = 1e-4
GRAD_REL_L2 = 2e-2


def rel_l2(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


@pytest.fixture(scope="module")
def ops():
    from cp2_b200 import _lib, ops as _ops
    _lib.load()
    return _ops


@pytest.mark.parametrize("tag", DENSECL_TAGS)
def test_densecl_loss_golden(ops, tag):
    g, f = densecl_case(tag)
    ql = T(g["q_local_raw"]).cuda().requires_grad_(True)
    qg = T(g["q_global_raw"]).cuda().requires_grad_(True)
    cu = {k: v.detach().cuda() for k, v in f.items()}
    loss, lg, ll = ops.densecl_loss(cu["embd_q"], cu["embd_k"], F.normalize(ql.flatten(2), dim=1), cu["k_local"],
                                    F.normalize(qg, dim=1), cu["k_global"], T(g["queue0"]).cuda(),
                                    T(g["queue2_0"]).cuda(), 0.2, 0.2, 0.5, T(g["pix_q"]).cuda(), T(g["pix_k"]).cuda(),
                                    1, float(g["lmbd_coordinate"]))
    loss.backward()
    np.testing.assert_allclose(loss.item(), g["loss"], rtol=LOSS_RTOL)
    np.testing.assert_allclose(lg.item(), g["loss_global"], rtol=LOSS_RTOL)
    np.testing.assert_allclose(ll.item(), g["loss_local"], rtol=LOSS_RTOL)
    assert rel_l2(ql.grad, T(g["grad_q_local"])) < GRAD_REL_L2, rel_l2(ql.grad, T(g["grad_q_local"]))
    assert rel_l2(qg.grad, T(g["grad_q_global"])) < GRAD_REL_L2, rel_l2(qg.grad, T(g["grad_q_global"]))
    # both queues after the step (builder.py:975-976)
    queue, queue2 = T(g["queue0"]).cuda().clone(), T(g["queue2_0"]).cuda().clone()
    p1, p2 = torch.zeros(1, dtype=torch.int64).cuda(), torch.zeros(1, dtype=torch.int64).cuda()
    ops.enqueue(cu["k_global"], queue, p1)
    ops.enqueue(cu["k_pooled"], queue2, p2)
    assert int(p1) == int(g["queue_ptr1"][0]) and int(p2) == int(g["queue2_ptr1"][0])
    np.testing.assert_allclose(queue.cpu().numpy(), g["queue1"], rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(queue2.cpu().numpy(), g["queue2_1"], rtol=1e-5, atol=1e-7)


@pytest.mark.parametrize("N,K,tau", [(70, 1000, 0.2), (32, 4096, 0.07), (333, 2048, 0.2), (1300, 4096, 0.2),
                                     (1024, 1000, 0.1), (1500, 5000, 0.2), (2600, 8192, 0.2), (6272, 3000, 0.2)])
def test_rows_infonce_vs_oracle(ops, N, K, tau):
    gen = torch.Generator().manual_seed(N + K)
    C = 128
    rows = F.normalize(torch.randn(N, C, generator=gen), dim=1)
    keys = F.normalize(rows * 0.8 + torch.randn(N, C, generator=gen) * 0.06, dim=1)
    keys[::3] *= 1.7                                        # mixed positives need not be unit vectors
    queue = F.normalize(torch.randn(C, K, generator=gen), dim=0)
    rr = rows.clone().requires_grad_(True)
    ref = O.rows_infonce(rr, keys, queue, tau)
    (ref * 2.5).backward()
    rd = rows.cuda().requires_grad_(True)
    loss = ops.rows_infonce(rd, keys.cuda(), queue.cuda(), tau)
    (loss * 2.5).backward()
    np.testing.assert_allclose(loss.item(), ref.item(), rtol=LOSS_RTOL)
    # N >= ops.TC_ROWS_MIN: queue sums from cp2_queue_sums (the dense backward kernel, rectangular mode), partial
    # sums added by cp2_rows_nce; below: cp2_global_nce
    assert rel_l2(rd.grad, rr.grad) < GRAD_REL_L2, rel_l2(rd.grad, rr.grad)


@pytest.mark.parametrize("B,Cb,P", [(4, 2048, 196), (2, 256, 49), (3, 64, 128), (2, 512, 256), (1, 2048, 50),
                                    (2, 96, 131)])
def test_densecl_match_vs_oracle(ops, B, Cb, P):
    """cp2_densecl_match (tcgen05, bf16 hi+lo operands) against the fp64 arg-max of builder.py:817-821, on
    post-ReLU-like (non-negative) and signed features; planted winners; exact duplicates keep the first index."""
    gen = torch.Generator().manual_seed(Cb + P)
    for signed in (False, True):
        eq = torch.randn(B, Cb, P, generator=gen)
        ek = torch.randn(B, Cb, P, generator=gen)
        if not signed:
            eq, ek = eq.relu(), ek.relu()
        eq, ek = F.normalize(eq, dim=1), F.normalize(ek, dim=1)
        ek[:, :, 3] = eq[:, :, P - 1] * 1.5                  # planted: query P-1 matches key 3
        if P > 10:
            ek[:, :, 7] = ek[:, :, 5]                        # exact duplicate columns: the first one must win
        sim = torch.einsum("ncx,ncy->nxy", eq.double(), ek.double())
        want = sim.argmax(dim=2)
        got = ops.densecl_match(eq.cuda(), ek.cuda()).cpu()
        assert got.dtype == torch.int64 and got.shape == (B, P)
        assert (got[:, P - 1] == 3).all()
        assert not (got == 7).any() or P <= 10
        bad = got != want
        if bad.any():                                        # only ties below the stated 2^-16 resolution may differ
            top = sim.gather(2, want.unsqueeze(2)).squeeze(2)
            alt = sim.gather(2, got.unsqueeze(2)).squeeze(2)
            scale = eq.double().norm(dim=1) * ek.double().norm(dim=1).max(dim=1, keepdim=True).values
            assert ((top - alt)[bad] <= 2.0 ** -15 * scale[bad]).all(), float((top - alt)[bad].max())
            assert bad.float().mean() < 0.01


@pytest.mark.parametrize("B,C,hw,os_,lmbd", [(3, 128, 6, 2, 0.3), (2, 128, 14, 1, 0.5), (2, 64, 5, 4, 0.0)])
def test_positive_keys_vs_oracle(ops, B, C, hw, os_, lmbd):
    gen = torch.Generator().manual_seed(hw * 10 + B)
    P, H = hw * hw, hw * os_
    k_local = F.normalize(torch.randn(B, C, P, generator=gen), dim=1)
    idx = torch.randint(0, P, (B, P), generator=gen)
    pix_q = torch.arange(1, H * H + 1).reshape(1, H, H).repeat(B, 1, 1)
    pix_k = torch.roll(pix_q, shifts=(os_, 0), dims=(1, 2)).clone()
    pix_k[:, :, : H // 2] = pix_k[:, :, H // 2: 2 * (H // 2)].clone()      # duplicated ids: several matches per query
    pix_k[:, : H // 4] += 10 * H * H                                       # and a band without any
    sub = lambda t: t[:, os_ // 2::os_, os_ // 2::os_]
    want = O.densecl_positive_keys(k_local, idx, sub(pix_q), sub(pix_k), lmbd)
    corr = sub(pix_q).reshape(B, -1)[:, :, None] == sub(pix_k).reshape(B, -1)[:, None, :]
    assert lmbd == 0.0 or (corr.sum(-1).max() >= 2 and (corr.sum(-1) == 0).any())
    got = ops.positive_keys(k_local.cuda(), idx.cuda(), pix_q.cuda(), pix_k.cuda(), os_, lmbd)
    np.testing.assert_allclose(got.cpu().numpy(), want.reshape(B * P, C).numpy(), rtol=1e-5, atol=1e-6)
    plain = ops.positive_keys(k_local.cuda(), idx.cuda())
    np.testing.assert_allclose(plain.cpu().numpy(),
                               O.densecl_positive_keys(k_local, idx, sub(pix_q), sub(pix_k), 0.0).reshape(B * P, C).numpy(),
                               rtol=1e-6, atol=1e-7)


class _Enc(nn.Module):
    """2048-channel stride-16 backbone + a 128-d head for the output-stride probe."""

    class _BB(nn.Module):
        def __init__(self):
            super().__init__()
            self.proj = nn.Conv2d(3, 2048, 16, stride=16)

        def forward(self, x):
            y = self.proj(x)
            return (y, y, y, y)

    def __init__(self):
        super().__init__()
        self.backbone = _Enc._BB()
        self.head = nn.Conv2d(2048, 128, 1)

    def forward(self, x):
        return self.head(self.backbone(x)[3])


@pytest.mark.parametrize("mode,sym,lmbd_c,pred", [("DENSECL", False, 0.0, False), ("PROPOSED_V2", True, 0.3, True)])
def test_model_densecl_modes(monkeypatch, mode, sym, lmbd_c, pred):
    """MODEL.forward_densecl end to end: the loss equals the oracle's on the features the model itself
    extracted, both queues and pointers advance with the right keys, the EMA ran once per key pass."""
    from cp2_b200 import builder
    monkeypatch.setattr(torch.backends.cudnn, "allow_tf32", False)
    monkeypatch.setattr(torch.backends.cuda.matmul, "allow_tf32", False)
    torch.manual_seed(5)
    B, H, K = 4, 96, 512
    model = builder.MODEL(None, rank=1, dim=128, K=K, lmbd_cp2_dense_loss=0.5, instance_logits_temp=0.2,
                          dense_logits_temp=0.2, pretrain_type=mode, use_symmetrical_loss=sym, lmbd_coordinate=lmbd_c,
                          use_predictor=pred, pretrain_from_scratch=True, device=torch.device("cuda"),
                          encoder_factory=_Enc, output_stride=16).cuda()
    assert model.backbone_output_stride == 16 and isinstance(model.encoder_k.neck, builder.DenseCLNeck)
    with torch.no_grad():
        for p in model.encoder_q.parameters():
            p.add_(0.02 * torch.randn_like(p))
    rec = {"q": [], "k": []}
    qf, kf = model._query_features, model._key_features
    monkeypatch.setattr(model, "_query_features", lambda x: (rec["q"].append(qf(x)), rec["q"][-1])[1])
    monkeypatch.setattr(model, "_key_features", lambda x: (rec["k"].append(kf(x)), rec["k"][-1])[1])
    k_before = [p.detach().clone() for p in model.encoder_k.parameters()]
    queue0, queue2_0 = model.queue.clone(), model.queue2.clone()
    img_a, img_b = torch.rand(B, 3, H, H, device="cuda"), torch.rand(B, 3, H, H, device="cuda")
    ids_a = torch.arange(1, H * H + 1, device="cuda").reshape(1, H, H).repeat(B, 1, 1)
    ids_b = torch.roll(ids_a, shifts=(16, 32), dims=(1, 2)).contiguous()       # shared ids on the lattice
    loss = model(img_a=img_a, img_b=img_b, bg0=None, bg1=None, visualize=False, step=0, new_epoch=False,
                 pixel_ids_a=ids_a, pixel_ids_b=ids_b, region_ids_a=ids_a, region_ids_b=ids_b)
    loss.backward()
    assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in model.encoder_q.backbone.parameters())
    sub = lambda t: t[:, 8::16, 8::16].cpu()
    tot_g = tot_l = 0.0
    dirs = [(ids_a, ids_b)] + ([(ids_b, ids_a)] if sym else [])
    for (eq, ql, qg), (ek, kl, kg, kpool), (pq, pk) in zip(rec["q"], rec["k"], dirs):
        r = O.densecl_loss(eq.cpu(), ek.cpu(), ql.detach().cpu(), kl.cpu(), qg.detach().cpu(), kg.cpu(), sub(pq),
                           sub(pk), queue0.cpu(), queue2_0.cpu(), 0.2, 0.2, 0.5, lmbd_c)
        tot_g, tot_l = tot_g + r["loss_global"].item(), tot_l + r["loss_local"].item()
    np.testing.assert_allclose(float(model.loss_i.val), tot_g, rtol=LOSS_RTOL)
    np.testing.assert_allclose(float(model.loss_d.val), tot_l, rtol=LOSS_RTOL)
    np.testing.assert_allclose(loss.item(), 0.5 * tot_g + 0.5 * tot_l, rtol=LOSS_RTOL)
    # queues: step 0 with the symmetrical loss enqueues the second direction's keys (builder.py:963-968)
    kg, kpool = rec["k"][-1][2], rec["k"][-1][3]
    assert int(model.queue_ptr) == B and int(model.queue2_ptr) == B
    assert torch.allclose(model.queue[:, :B].t(), kg, atol=1e-6) and torch.equal(model.queue[:, B:], queue0[:, B:])
    assert torch.allclose(model.queue2[:, :B].t(), kpool, atol=1e-6)
    # one EMA per key pass (builder.py:724): k <- k*m^n + ... differs from before, neck included
    assert all(not torch.equal(a, p.detach()) for a, p in zip(k_before, model.encoder_k.parameters()))
    assert len(rec["k"]) == (2 if sym else 1)
