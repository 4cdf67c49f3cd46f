"""GPU parity of the PROPOSED-mode dense loss (pair weights from the pixel /
region id maps, NegativeType squash; reference builder.py:1204-1243, 1332-1392) through the C ABI:
(a) against golden vectors recorded from the reference's own `forward_cp2` in PretrainType.PROPOSED,
(b) against the CPU oracle on seeded inputs, dense part alone, with the id maps at full resolution so the
    in-kernel [OS//2::OS] sub-sample is exercised, (c) the drop-in MODEL in PROPOSED mode.

Tolerances: loss 1e-4 relative (north_star); gradients rel-L2 <= 2e-2 (bf16 operands, fp32 accumulate).
"""
import numpy as np
import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

from conftest import T, load_golden
from oracle import cp2_oracle as O
from test_oracle_golden import WEIGHTED_TAGS, weighted_case

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


@pytest.mark.parametrize("tag", WEIGHTED_TAGS)
def test_weighted_loss_golden(ops, tag):
    g, mask_a, mask_b, ids = weighted_case(tag)
    wp, wr, wn = (float(x) for x in g["weights"])
    pw = ops.PairWeights(ids["pixel_ids_a"].cuda(), ids["pixel_ids_b"].cuda(), ids["region_ids_a"].cuda(),
                         ids["region_ids_b"].cuda(), 1, wp, wr, wn, str(g["negative_type"]),
                         float(g["negative_scale"]))
    q = T(g["q"]).cuda().requires_grad_(True)
    loss, li, ld, _, _, _ = ops.cp2_loss(q, T(g["k"]).cuda(), mask_a.cuda(), mask_b.cuda(), T(g["queue0"]).cuda(),
                                         float(g["temp_global"]), float(g["temp_local"]), float(g["lmbd"]),
                                         bool(g["include_background"]), pw)
    loss.backward()
    np.testing.assert_allclose(loss.item(), g["loss"], rtol=LOSS_RTOL)
    np.testing.assert_allclose(li.item(), g["loss_ins"], rtol=LOSS_RTOL)
    np.testing.assert_allclose(ld.item(), g["loss_dense"], rtol=LOSS_RTOL)
    ref = T(g["grad_q"])
    assert rel_l2(q.grad, ref) < GRAD_REL_L2, rel_l2(q.grad, ref)
    if not bool(g["include_background"]):
        # background query pixels receive the dense loss's gradient only: the part this mode changes
        bgpix = (mask_a == 0).reshape(2, 1, 14, 14).expand_as(ref)
        assert bgpix.any()
        assert rel_l2(q.grad.cpu()[bgpix], ref[bgpix]) < GRAD_REL_L2, rel_l2(q.grad.cpu()[bgpix], ref[bgpix])


def _ids(B, hw, os_, gen):
    """Full-resolution id maps whose [os//2::os] lattice has pixel matches, region matches and unknowns."""
    H = hw * os_
    pa = torch.arange(1, H * H + 1).reshape(1, H, H).repeat(B, 1, 1)
    pb = torch.roll(pa, shifts=(os_, -os_), dims=(1, 2)).clone()
    pb[:, : H // 3] += H * H                                     # a band of view b that shares no pixel with view a
    blk = max(2 * os_, H // 3)
    yy, xx = torch.meshgrid(torch.arange(H), torch.arange(H), indexing="ij")
    ra = ((yy // blk) * 4 + xx // blk).repeat(B, 1, 1)              # block 0 = unknown region
    rb = torch.roll(ra, shifts=(os_, 0), dims=(1, 2)).clone()
    rb[torch.rand(B, H, H, generator=gen) < 0.1] = 0
    return pa, pb, ra, rb


@pytest.mark.parametrize("B,C,hw,os_,neg,tl,weights", [
    (3, 128, 14, 2, "FIXED", 0.2, (10.0, 2.0, 0.5)),
    (2, 128, 9, 4, "AVERAGE", 0.5, (3.0, 1.0, 0.0)),
    (2, 256, 7, 1, "NONE", 1.0, (10.0, 1.0, 1.0)),
    (2, 256, 12, 2, "AVERAGE", 0.2, (1.0, 4.0, 0.25)),
    (1, 128, 28, 2, "FIXED", 0.2, (5.0, 2.0, 1.0)),
    (3, 128, 14, 2, "MEDIAN", 0.2, (10.0, 2.0, 0.5)),
    (2, 256, 9, 1, "MEDIAN", 1.0, (2.0, 1.0, 1.0)),
])
def test_weighted_dense_vs_oracle(ops, B, C, hw, os_, neg, tl, weights):
    gen = torch.Generator().manual_seed(100 * hw + C)
    P = hw * hw
    q = torch.randn(B, C, P, generator=gen)
    k = q.roll(1, dims=2) * 0.7 + torch.randn(B, C, P, generator=gen) * 0.7      # structured similarities
    ma = (torch.rand(B, P, generator=gen) < 0.65).float()
    mb = (torch.rand(B, P, generator=gen) < 0.65).float()
    pa, pb, ra, rb = _ids(B, hw, os_, gen)
    sub = lambda t: O.stride_subsample(t, os_) if os_ > 1 else t
    w = O.cp2_corr_weights(sub(pa), sub(pb), sub(ra), sub(rb), ma, mb, *weights)
    assert len(torch.unique(w)) >= 2
    # oracle on the bf16-rounded unit vectors the kernels consume (straight-through for the gradient)
    qr = q.clone().requires_grad_(True)
    qh = F.normalize(qr, dim=1)
    qh = qh + (qh.bfloat16().float() - qh).detach()
    kh = F.normalize(k, dim=1).bfloat16().float()
    ref = O.dense_infonce_weighted(qh, kh, ma, mb, w, tl, neg, 2.0)
    ref.backward()

    qd, kd, mad, mbd = q.cuda(), k.cuda(), ma.cuda(), mb.cuda()
    qt, q_inv, sq, nsq, _ = ops.prep_features(qd, mad)
    kt, _, sk, nsk, _ = ops.prep_features(kd, mbd)
    pw = ops.PairWeights(pa.cuda(), pb.cuda(), ra.cuda(), rb.cuda(), os_, *weights, neg, 2.0)
    lse, dense_n, wstats, cnt, acc_seg = ops.dense_weighted_lse(qt, kt, mad, mbd, pw, sq, nsq, sk, nsk, 1.0 / tl)
    np.testing.assert_allclose(dense_n.mean().item(), ref.item(), rtol=LOSS_RTOL)
    assert torch.equal(cnt.cpu().long(), mb.sum(1).long())
    # acc_seg on the weighted logits (builder.py:1442-1448)
    z = O.negative_transform(torch.einsum("ncx,ncy->nxy", qh.detach(), kh), torch.einsum("nx,ny->nxy", ma, mb),
                             neg, 2.0) * w / tl
    idx = z.reshape(B, -1).argmax(dim=1)
    want = (ma[torch.arange(B), idx // P] * mb[torch.arange(B), idx % P]).mean() * 100.0
    np.testing.assert_allclose(acc_seg.item(), want.item(), rtol=1e-6)
    np.testing.assert_allclose(wstats[:, 1].cpu().numpy(), ma.sum(1).numpy())
    if neg in ("AVERAGE", "MEDIAN"):
        lab = torch.einsum("nx,ny->nxy", ma, mb)
        st = O.dense_loss_stats(torch.einsum("ncx,ncy->nxy", qh.detach(), kh), lab)["negative"]
        centre = st["average"] if neg == "AVERAGE" else st["quartiles"][1]
        np.testing.assert_allclose(wstats[:, 0].cpu().numpy(), centre.numpy(), rtol=2e-3, atol=2e-5)
    alpha = (1.0 / tl) / (B * mb.sum(1))
    v1 = torch.zeros(B, C)
    dq = ops.dense_weighted_grad(qt, kt, mad, mbd, pw, lse, wstats, q_inv, alpha.cuda(), v1.cuda(), None, 1.0 / tl)
    assert rel_l2(dq, qr.grad) < GRAD_REL_L2, rel_l2(dq, qr.grad)
    # gscale multiplies the whole gradient
    dq3 = ops.dense_weighted_grad(qt, kt, mad, mbd, pw, lse, wstats, q_inv, alpha.cuda(), v1.cuda(), None, 1.0 / tl,
                                  gscale=torch.tensor([3.0], device="cuda"))
    assert rel_l2(dq3, 3.0 * dq) < 1e-5
    # the tensor-core implementation for maps that fit TMEM: same contract, acc instead of a second P^2 pass
    if ops.dense_small_ok(C, P):
        lse2, dn2, ws2, cnt2, hit, acc = ops.dense_weighted_small(qt, kt, mad, mbd, pw, sq, nsq, sk, nsk, 1.0 / tl)
        np.testing.assert_allclose(dn2.mean().item(), ref.item(), rtol=LOSS_RTOL)
        np.testing.assert_allclose(dn2.cpu().numpy(), dense_n.cpu().numpy(), rtol=2e-5)
        np.testing.assert_allclose(hit.mean().item() * 100.0, want.item(), rtol=1e-6)
        assert torch.equal(cnt2, cnt) and torch.equal(ws2, wstats)
        assert torch.allclose(lse2, lse, rtol=0, atol=1e-4), float((lse2 - lse).abs().max())
        dq2 = ops.dense_small_bwd(acc, qt, mad, q_inv, alpha.cuda(), v1.cuda(), None)
        assert rel_l2(dq2, qr.grad) < GRAD_REL_L2, rel_l2(dq2, qr.grad)
        assert rel_l2(dq2, dq) < 6e-3, rel_l2(dq2, dq)          # g is rounded to bf16 for the tensor cores


@pytest.mark.parametrize("B,hw,os_,neg,tl,weights", [
    (2, 28, 2, "FIXED", 0.2, (5.0, 2.0, 1.0)),
    (3, 14, 2, "AVERAGE", 0.2, (10.0, 2.0, 0.5)),
    (2, 23, 1, "MEDIAN", 0.5, (3.0, 1.0, 0.25)),
    (1, 18, 2, "NONE", 1.0, (10.0, 1.0, 1.0)),
])
def test_weighted_stream_vs_oracle_and_simt(ops, B, hw, os_, neg, tl, weights):
    """The streaming tensor-core pair of the PROPOSED modes (cp2_dense_weighted_stream / _stream_grad) against the
    oracle and against the fp32 SIMT pair it replaces on maps that do not fit TMEM."""
    C = 128
    gen = torch.Generator().manual_seed(7 * hw + B)
    P = hw * hw
    q = torch.randn(B, C, P, generator=gen)
    k = q.roll(1, dims=2) * 0.7 + torch.randn(B, C, P, generator=gen) * 0.7
    ma = (torch.rand(B, P, generator=gen) < 0.65).float()
    mb = (torch.rand(B, P, generator=gen) < 0.65).float()
    pa, pb, ra, rb = _ids(B, hw, os_, gen)
    sub = lambda t: O.stride_subsample(t, os_) if os_ > 1 else t
    w = O.cp2_corr_weights(sub(pa), sub(pb), sub(ra), sub(rb), ma, mb, *weights)
    qr = q.clone().requires_grad_(True)
    qh = F.normalize(qr, dim=1)
    qh = qh + (qh.bfloat16().float() - qh).detach()
    kh = F.normalize(k, dim=1).bfloat16().float()
    ref = O.dense_infonce_weighted(qh, kh, ma, mb, w, tl, neg, 2.0)
    ref.backward()
    assert ops.dense_weighted_stream_ok(C, P, 1.0 / tl)
    qd, kd, mad, mbd = q.cuda(), k.cuda(), ma.cuda(), mb.cuda()
    qt, q_inv, sq, nsq, _ = ops.prep_features(qd, mad)
    kt, _, sk, nsk, _ = ops.prep_features(kd, mbd)
    pw = ops.PairWeights(pa.cuda(), pb.cuda(), ra.cuda(), rb.cuda(), os_, *weights, neg, 2.0)
    lse0, dn0, ws0, cnt0, acc0 = ops.dense_weighted_lse(qt, kt, mad, mbd, pw, sq, nsq, sk, nsk, 1.0 / tl)
    lse, dn, ws, cnt, acc_seg, ids_pad, b_pad = ops.dense_weighted_stream(qt, kt, mad, mbd, pw, sq, nsq, sk, nsk, 1.0 / tl)
    np.testing.assert_allclose(dn.mean().item(), ref.item(), rtol=LOSS_RTOL)
    np.testing.assert_allclose(dn.cpu().numpy(), dn0.cpu().numpy(), rtol=3e-5)
    assert torch.equal(cnt, cnt0) and torch.equal(ws, ws0)
    assert torch.allclose(lse, lse0, rtol=0, atol=1e-4), float((lse - lse0).abs().max())
    np.testing.assert_allclose(acc_seg.item(), acc0.item(), rtol=1e-6)
    alpha = (1.0 / tl) / (B * mb.sum(1))
    v1 = torch.zeros(B, C)
    dq0 = ops.dense_weighted_grad(qt, kt, mad, mbd, pw, lse0, ws0, q_inv, alpha.cuda(), v1.cuda(), None, 1.0 / tl)
    dq = ops.dense_weighted_stream_grad(qt, kt, mad, mbd, pw, lse, ws, ids_pad, b_pad, q_inv, alpha.cuda(), v1.cuda(), None,
                                        1.0 / tl, gscale=torch.tensor([1.0], device="cuda"))
    assert rel_l2(dq, qr.grad) < GRAD_REL_L2, rel_l2(dq, qr.grad)
    assert rel_l2(dq, dq0) < 6e-3, rel_l2(dq, dq0)                 # g is rounded to bf16 for the tensor cores


def test_weighted_all_ones_equals_cp2_path(ops):
    """w = 1 everywhere and no squash is the CP2 loss: the SIMT pair must agree with the tensor-core path."""
    gen = torch.Generator().manual_seed(3)
    B, C, hw, K = 4, 128, 14, 2048
    q = torch.randn(B, C, hw, hw, generator=gen)
    k = torch.randn(B, C, hw, hw, generator=gen)
    ma = (torch.rand(B, hw * hw, generator=gen) < 0.65).float().cuda()
    mb = (torch.rand(B, hw * hw, generator=gen) < 0.65).float().cuda()
    queue = F.normalize(torch.randn(C, K, generator=gen), dim=0).cuda()
    ids = torch.arange(1, hw * hw + 1).reshape(1, hw, hw).repeat(B, 1, 1).cuda()
    pw = ops.PairWeights(ids, ids, ids, ids, 1, 1.0, 1.0, 1.0, "HARD", 2.0)
    outs = []
    for pair in (None, pw):
        qd = q.cuda().requires_grad_(True)
        loss, li, ld, _, _, _ = ops.cp2_loss(qd, k.cuda(), ma, mb, queue, 0.2, 0.2, 0.5, True, pair)
        loss.backward()
        outs.append((loss.item(), ld.item(), qd.grad))
    np.testing.assert_allclose(outs[1][0], outs[0][0], rtol=1e-5)
    np.testing.assert_allclose(outs[1][1], outs[0][1], rtol=1e-5)
    assert rel_l2(outs[1][2], outs[0][2]) < 2e-3


def test_weighted_rejects_bad_arguments(ops):
    ids = torch.zeros(2, 14, 14, dtype=torch.int64, device="cuda")
    with pytest.raises(NotImplementedError):
        ops.PairWeights(ids, ids, ids, ids, 1, 1, 1, 1, "SOFTEST")
    with pytest.raises(TypeError):
        ops.PairWeights(ids.int(), ids, ids, ids, 1, 1, 1, 1)
    pw = ops.PairWeights(ids, ids, ids, ids, 1, 1, 1, 1)           # 14x14 ids, but P = 49 below
    q = torch.randn(2, 128, 49, device="cuda")
    m = torch.ones(2, 49, device="cuda")
    qt, _, sq, nsq, _ = ops.prep_features(q, m)
    with pytest.raises(RuntimeError, match="id maps"):
        ops.dense_weighted_lse(qt, qt, m, m, pw, sq, nsq, sq, nsq, 1.0)


class _Feat(nn.Module):
    def __init__(self, feat):
        super().__init__()
        self.feat = nn.Parameter(feat.clone())

    def forward(self, x):
        return self.feat[: x.shape[0]] * 1.0


@pytest.mark.parametrize("tag", ["fixed_t05", "avg_t1_bg", "median_t05"])
def test_model_proposed_mode_golden(tag, monkeypatch):
    """MODEL(pretrain_type=PROPOSED, mapping_type=PIXEL_REGION_ID) against the reference's own forward_cp2."""
    from cp2_b200 import builder
    g, mask_a, mask_b, ids = weighted_case(tag)
    wp, wr, wn = (float(x) for x in g["weights"])
    feats = [T(g["q"]), T(g["k"])]
    # m = 1: the EMA leaves the stored key features alone (the fixture's k is the key encoder's output)
    model = builder.MODEL(None, rank=1, dim=128, K=1024, m=1.0, pretrain_from_scratch=True,
                          include_background=bool(g["include_background"]), lmbd_cp2_dense_loss=float(g["lmbd"]),
                          lmbd_pixel_corr_weight=wp, lmbd_region_corr_weight=wr, lmbd_not_corr_weight=wn,
                          negative_type=str(g["negative_type"]), negative_scale=float(g["negative_scale"]),
                          pretrain_type="PROPOSED", mapping_type="PIXEL_REGION_ID",
                          dense_logits_temp=float(g["temp_local"]), device=torch.device("cuda"),
                          encoder_factory=lambda: _Feat(feats.pop(0)), output_stride=16).cuda()
    with torch.no_grad():
        model.encoder_k.feat.copy_(T(g["k"]).cuda())             # the constructor copied encoder_q's parameters
        model.queue.copy_(T(g["queue0"]).cuda())
    # the stand-in key encoder ignores its input, so the batch shuffle must not permute its output
    monkeypatch.setattr(model, "_batch_shuffle_ddp", lambda x: (x, None))
    monkeypatch.setattr(model, "_batch_unshuffle_ddp", lambda x, idx: x)
    # inputs whose composite reproduces the fixture's masks; ids at full resolution (lattice = the fixture's)
    bg0 = (~T(g["bg0_ch0_zero"])).float()[:, None].repeat(1, 3, 1, 1).cuda()
    bg1 = (~T(g["bg1_ch0_zero"])).float()[:, None].repeat(1, 3, 1, 1).cuda()
    full = {}
    for name, t in ids.items():
        f = torch.zeros(2, 224, 224, dtype=torch.int64)
        f[:, 8::16, 8::16] = t
        full[name] = f.cuda()
    img = torch.rand(2, 3, 224, 224, device="cuda")
    loss = model(img_a=img, img_b=img.clone(), bg0=bg0, bg1=bg1, **full)
    loss.backward()
    np.testing.assert_allclose(loss.item(), g["loss"], rtol=LOSS_RTOL)
    np.testing.assert_allclose(float(model.loss_d.val), g["loss_dense"], rtol=LOSS_RTOL)
    assert rel_l2(model.encoder_q.feat.grad, T(g["grad_q"])) < GRAD_REL_L2
    np.testing.assert_allclose(np.asarray(model.masked_correlation_ious, dtype=np.float32), g["iou_masked"], rtol=1e-6)
