"""GPU parity of the fused CP2 loss (ops.cp2_loss: prep + dense LSE + global NCE + one-kernel
backward) and of the drop-in MODEL against (a) golden vectors recorded from the unmodified
reference and (b) the CPU oracle on seeded inputs.

Tolerances: loss 1e-4 relative (north_star); gradient rel-L2 <= 2e-2 (bf16 operands on the
tensor cores, fp32 accumulate; measured ~3e-3).
"""
import numpy as np
import pytest
import torch
import torch.nn as nn

from conftest import T, load_golden
from oracle import cp2_oracle as O

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


@pytest.mark.parametrize("tag", ["t1", "t02", "t02_bg"])
def test_loss_golden_p196(ops, tag):
    g = load_golden(f"cp2_loss_p196_{tag}.npz")
    mask_a = O.stride_subsample(T(g["bg0_ch0_zero"]).float(), 16).reshape(2, -1)
    mask_b = O.stride_subsample(T(g["bg1_ch0_zero"]).float(), 16).reshape(2, -1)
    q = T(g["q"]).cuda().requires_grad_(True)
    loss, li, ld, k_pos, q_pos, st = ops.cp2_loss(
        q, T(g["k"]).cuda(), mask_a.cuda(), mask_b.cuda(), T(g["queue0"]).cuda(),
        float(g["temp_global"]), float(g["temp_local"]), float(g["lmbd"]), bool(g["include_background"]),
        track_accuracy=True)
    loss.backward()
    np.testing.assert_allclose(st["acc_ins"].item(), g["acc_ins"], atol=1e-3)   # from the loss's own queue pass
    np.testing.assert_allclose(st["acc_seg"].item(), g["acc_seg"], atol=1e-3)   # from the dense kernel itself (P=196)
    # the reference's own accuracy meters (builder.py:1441-1448) from the two arg-max kernels
    bg = bool(g["include_background"])
    acc_ins, _ = ops.global_top1(q_pos, k_pos, T(g["queue0"]).cuda(), st["q_neg"] if bg else None,
                                 st["k_neg"] if bg else None)
    acc_seg, _ = ops.dense_top1(q.detach(), T(g["k"]).cuda(), mask_a.cuda(), mask_b.cuda())
    np.testing.assert_allclose(acc_ins.item(), g["acc_ins"], atol=1e-3)
    np.testing.assert_allclose(acc_seg.item(), g["acc_seg"], atol=1e-3)
    np.testing.assert_allclose(loss.item(), g["loss"], rtol=LOSS_RTOL)
    np.testing.assert_allclose(li.item(), g["loss_ins"], rtol=LOSS_RTOL)
    np.testing.assert_allclose(ld.item(), g["loss_dense"], rtol=LOSS_RTOL)
    assert rel_l2(q.grad, T(g["grad_q"])) < GRAD_REL_L2, rel_l2(q.grad, T(g["grad_q"]))
    # enqueue of this step's keys
    queue = T(g["queue0"]).cuda().clone()
    ptr = torch.zeros(1, dtype=torch.int64).cuda()
    ops.enqueue(k_pos, queue, ptr)
    assert int(ptr.item()) == int(g["queue_ptr1"][0])
    np.testing.assert_allclose(queue[:, :8].cpu().numpy(), g["queue1_head"], rtol=1e-5, atol=1e-7)


@pytest.mark.parametrize("B,C,hw,K,tl,bg", [(4, 128, 14, 4096, 0.2, False), (2, 128, 28, 1024, 1.0, True),
                                             (3, 256, 7, 512, 0.2, False), (2, 128, 9, 256, 0.2, True)])
def test_loss_vs_oracle(ops, B, C, hw, K, tl, bg):
    g = torch.Generator().manual_seed(B * 7 + hw)
    q = torch.randn(B, C, hw, hw, generator=g)
    k = torch.randn(B, C, hw, hw, generator=g)
    ma = (torch.rand(B, hw * hw, generator=g) < 0.65).float()
    mb = (torch.rand(B, hw * hw, generator=g) < 0.65).float()
    queue = torch.nn.functional.normalize(torch.randn(C, K, generator=g), dim=0)
    qr = q.clone().requires_grad_(True)
    r = O.cp2_loss(qr, k, ma, mb, queue, 0.2, tl, 0.2, include_background=bg)
    (r["loss"] * 1.7).backward()
    qd = q.cuda().requires_grad_(True)
    loss, li, ld, k_pos, q_pos, st = ops.cp2_loss(qd, k.cuda(), ma.cuda(), mb.cuda(), queue.cuda(), 0.2, tl,
                                                  0.2, bg, track_accuracy=True)
    (loss * 1.7).backward()
    # the cross-image std meters (builder.py:1553-1563: q_pos.std(0).mean()) and l_pos come out of the same kernel
    np.testing.assert_allclose(st["std_q_pos"].item(), r["q_pos"].detach().std(0).mean().item(), rtol=1e-4)
    np.testing.assert_allclose(st["std_k_pos"].item(), r["k_pos"].std(0).mean().item(), rtol=1e-4)
    assert torch.allclose(st["l_pos"].cpu(), (r["q_pos"].detach() * r["k_pos"]).sum(1), rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(loss.item(), r["loss"].item(), rtol=LOSS_RTOL)
    np.testing.assert_allclose(li.item(), r["loss_ins"].item(), rtol=LOSS_RTOL)
    np.testing.assert_allclose(ld.item(), r["loss_dense"].item(), rtol=LOSS_RTOL)
    assert torch.allclose(k_pos.cpu(), r["k_pos"], rtol=1e-5, atol=1e-6)
    assert torch.allclose(q_pos.cpu(), r["q_pos"].detach(), rtol=1e-5, atol=1e-6)
    assert rel_l2(qd.grad, qr.grad) < GRAD_REL_L2, rel_l2(qd.grad, qr.grad)
    np.testing.assert_allclose(st["acc_ins"].item(), r["acc1"].item(), atol=1e-3)   # meter of the same queue pass


def test_dense_loss_linearity_full_size(ops):
    """Size-independent property at the BASELINE sweep's largest map (56x56): the backward is
    linear in the upstream gradient, and dq is orthogonal to q per pixel (normalize backward)."""
    B, C, hw = 2, 128, 56
    g = torch.Generator().manual_seed(2)
    q = torch.randn(B, C, hw, hw, generator=g).cuda()
    k = torch.randn(B, C, hw, hw, generator=g).cuda()
    ma = (torch.rand(B, hw * hw, generator=g) < 0.65).float().cuda()
    mb = (torch.rand(B, hw * hw, generator=g) < 0.65).float().cuda()
    queue = torch.nn.functional.normalize(torch.randn(C, 4096, generator=g), dim=0).cuda()
    grads = []
    for scale in (1.0, -2.5):
        qd = q.clone().requires_grad_(True)
        loss = ops.cp2_loss(qd, k, ma, mb, queue, 0.2, 0.2, 0.2, False)[0]
        (loss * scale).backward()
        grads.append(qd.grad)
    assert torch.isfinite(loss)
    assert rel_l2(grads[1], grads[0] * -2.5) < 1e-5
    dots = (grads[0] * q).sum(dim=1).abs().max() / grads[0].abs().max()
    assert dots < 2e-2, float(dots)                       # bf16 q-hat in the projection term


# ------------------------------------------------------------------------ MODEL
class _Patch(nn.Module):
    """Same stand-in as oracle/ref_harness.PatchEncoder (parameter names included) so the
    golden state dict loads."""

    def __init__(self, dim=128, os_=16):
        super().__init__()
        self.proj = nn.Conv2d(3, dim, kernel_size=os_, stride=os_)

    def forward(self, x):
        return self.proj(x)


def test_model_two_steps_against_reference_golden(monkeypatch):
    from cp2_b200.builder import MODEL
    g = load_golden("cp2_step_small.npz")
    # the golden run's stand-in conv ran in IEEE fp32 on the CPU; cuDNN defaults to TF32 convs
    monkeypatch.setattr(torch.backends.cudnn, "allow_tf32", False)
    B, H, C, K = int(g["B"]), int(g["H"]), int(g["C"]), int(g["K"])
    dev = torch.device("cuda:0")
    model = MODEL(None, rank=1, dim=C, K=K, dense_logits_temp=0.2, device=dev,
                  encoder_factory=lambda: _Patch(C, 16), output_stride=16, track_accuracy=True).to(dev)
    with torch.no_grad():
        model.queue.copy_(T(g["queue0"]))
        for name, p in model.encoder_q.named_parameters():
            p.copy_(T(g[f"q0.{name}"]))
        for name, p in model.encoder_k.named_parameters():
            p.copy_(T(g[f"k0.{name}"]))
    opt = torch.optim.SGD(model.encoder_q.parameters(), lr=0.05)
    ids = torch.arange(1, H * H + 1, dtype=torch.int64).reshape(1, H, H).repeat(B, 1, 1).to(dev)
    for s in (0, 1):
        opt.zero_grad()
        loss = model(img_a=T(g[f"s{s}.img_a"]).to(dev), img_b=T(g[f"s{s}.img_b"]).to(dev),
                     bg0=T(g[f"s{s}.bg0"]).to(dev), bg1=T(g[f"s{s}.bg1"]).to(dev), visualize=False, step=s,
                     new_epoch=False, pixel_ids_a=ids, pixel_ids_b=ids, region_ids_a=ids, region_ids_b=ids)
        loss.backward()
        np.testing.assert_allclose(loss.item(), g[f"s{s}.loss"], rtol=LOSS_RTOL)
        np.testing.assert_allclose(model.loss_i.val, g[f"s{s}.loss_ins"], rtol=LOSS_RTOL)
        np.testing.assert_allclose(model.loss_d.val, g[f"s{s}.loss_dense"], rtol=LOSS_RTOL)
        assert int(model.queue_ptr.item()) == int(g[f"s{s}.queue_ptr"][0])
        # opt-in accuracy meters (builder.py:1441-1448): values of the reference's own meters
        np.testing.assert_allclose(float(model.acc_ins.val), g[f"s{s}.acc_ins"], atol=1e-3)
        np.testing.assert_allclose(float(model.acc_seg.val), g[f"s{s}.acc_seg"], atol=1e-3)
        # queue contents: the keys come from a cuDNN conv here vs a CPU conv in the golden run
        np.testing.assert_allclose(model.queue.cpu().numpy(), g[f"s{s}.queue"], rtol=0, atol=2e-5)
        for name, p in model.encoder_k.named_parameters():
            if s == 0:      # EMA of identical inputs is bit-exact; step 1 inherits SGD/cuDNN rounding
                assert torch.equal(p.detach().cpu(), T(g[f"s{s}.k.{name}"])), name
            else:
                np.testing.assert_allclose(p.detach().cpu().numpy(), g[f"s{s}.k.{name}"], rtol=1e-4, atol=1e-6)
        for name, p in model.encoder_q.named_parameters():
            want = T(g[f"s{s}.gradw.{name}"])
            assert rel_l2(p.grad, want) < GRAD_REL_L2, (name, rel_l2(p.grad, want))
        opt.step()
    assert model.loss_o.count == 2 * B and abs(model.loss_o.val - float(g["s1.loss"])) < 1e-3
    # region-id IoUs recorded per step (builder.py:1252-1257): identical id maps in both views -> IoU 1,
    # one entry per image and step; reset_metrics restarts the meters but, as in the reference
    # (builder.py:1666-1688), not the IoU lists: those restart in the new-epoch branch (builder.py:1466,1486)
    assert len(model.correlation_ious) == 2 * B and len(model.masked_correlation_ious) == 2 * B
    assert all(v == 1.0 for v in model.correlation_ious)
    model.reset_metrics()
    assert model.loss_o.count == 0 and len(model.correlation_ious) == 2 * B
    logged = []
    model.rank, model._logger = 0, logged.append            # rank 0 with a logger: the new-epoch branch logs
    loss = model(img_a=T(g["s1.img_a"]).to(dev), img_b=T(g["s1.img_b"]).to(dev), bg0=T(g["s1.bg0"]).to(dev),
                 bg1=T(g["s1.bg1"]).to(dev), visualize=False, step=2, new_epoch=True, pixel_ids_a=ids,
                 pixel_ids_b=ids, region_ids_a=ids, region_ids_b=ids)
    assert torch.isfinite(loss) and model.epoch == 1
    assert model.correlation_ious == [] and model.masked_correlation_ious == []
    keys = set().union(*[set(d) for d in logged])
    assert {"feature_space_non_zero_iou", "dense-heatmaps", "train/loss_step", "train/+ive_scores_step",
            "step/instance_median_negative_scores"} <= keys, sorted(keys)
    assert model.last_heatmap_grid.dim() == 2
    out = model.on_train_epoch_end(3)
    assert out["train/loss"] == pytest.approx(loss.item(), rel=1e-5) and model.loss_o.count == 0


def test_loss_nan_hazard_matches_reference(ops):
    """a view without a foreground sample on the lattice gives `labels.sum() == 0`
    and a NaN dense loss in the reference (builder.py:1434-1437); the kernels do not hide it, and the other
    images of the batch are unaffected (their gradients stay finite and equal to the oracle's)."""
    g = torch.Generator().manual_seed(9)
    B, C, hw, K = 3, 128, 14, 512
    q = torch.randn(B, C, hw, hw, generator=g)
    k = torch.randn(B, C, hw, hw, generator=g)
    ma = (torch.rand(B, hw * hw, generator=g) < 0.6).float()
    mb = (torch.rand(B, hw * hw, generator=g) < 0.6).float()
    mb[1] = 0.0                                            # image 1: no foreground key pixel
    queue = torch.nn.functional.normalize(torch.randn(C, K, generator=g), dim=0)
    qr = q.clone().requires_grad_(True)
    r = O.cp2_loss(qr, k, ma, mb, queue, 0.2, 0.2, 0.2)
    assert torch.isnan(r["loss"]) and torch.isnan(r["loss_dense"]) and torch.isfinite(r["loss_ins"])
    qd = q.cuda().requires_grad_(True)
    loss, li, ld, _, _, _ = ops.cp2_loss(qd, k.cuda(), ma.cuda(), mb.cuda(), queue.cuda(), 0.2, 0.2, 0.2)
    assert torch.isnan(loss).item() and torch.isnan(ld).item()
    np.testing.assert_allclose(li.item(), r["loss_ins"].item(), rtol=LOSS_RTOL)


# ------------------------------------------------------------------------ CUDA-graph capture of the encoders
class _ConvBN(nn.Module):
    """Small segmentor with BatchNorm (running statistics are part of what the graphs must reproduce)."""

    def __init__(self, dim=128):
        super().__init__()
        self.stem = nn.Sequential(nn.Conv2d(3, 16, 4, stride=4, bias=False), nn.BatchNorm2d(16), nn.ReLU(inplace=True))
        self.body = nn.Sequential(nn.Conv2d(16, 32, 4, stride=4, bias=False), nn.BatchNorm2d(32), nn.ReLU(inplace=True))
        self.proj = nn.Conv2d(32, dim, 1)

    def forward(self, x):
        return self.proj(self.body(self.stem(x)))


def test_graph_encoders_matches_eager():
    """MODEL.graph_encoders: three training steps with both encoders replayed from CUDA graphs give the losses,
    parameters, BatchNorm statistics and queue of the same steps launched eagerly; capture itself leaves the model
    (BN running statistics included) untouched; eval mode and other shapes fall back to the eager path."""
    from cp2_b200.builder import MODEL
    dev = torch.device("cuda:0")
    B, H, C, K = 8, 224, 128, 256

    def make():
        torch.manual_seed(3)
        m = MODEL(None, rank=1, dim=C, K=K, dense_logits_temp=0.2, device=dev, pretrain_from_scratch=True,
                  encoder_factory=lambda: _ConvBN(C), output_stride=16).to(dev)
        return m, torch.optim.SGD(m.encoder_q.parameters(), lr=0.05, momentum=0.9)

    eager, opt_e = make()
    graphed, opt_g = make()
    staged, opt_s = make()
    before = {k: v.clone() for k, v in graphed.state_dict().items()}
    graphed.graph_encoders((B, 3, H, H))
    assert graphed.graphed == {"encoder_q": True, "encoder_k": True}
    # one graph pair per stage of the query encoder (what bench.py uses under DDP)
    staged.graph_encoders((B, 3, H, H), query_stages=[staged.encoder_q.stem, staged.encoder_q.body,
                                                       staged.encoder_q.proj])
    assert staged.graphed == {"encoder_q": "3 stages", "encoder_k": True}
    for k, v in staged.state_dict().items():
        assert torch.equal(v, before[k]), f"staged capture changed {k}"
    for k, v in graphed.state_dict().items():
        assert torch.equal(v, before[k]), f"capture changed {k}"
    g = torch.Generator().manual_seed(5)
    for s in range(3):
        img_a, img_b = torch.rand(B, 3, H, H, generator=g).to(dev), torch.rand(B, 3, H, H, generator=g).to(dev)
        bg0, bg1 = torch.rand(B, 3, H, H, generator=g), torch.rand(B, 3, H, H, generator=g)
        bg0[:, :, 30:200, 20:150] = 0.0
        bg1[:, :, 50:220, 60:224] = 0.0
        losses = []
        for m, opt in ((eager, opt_e), (graphed, opt_g), (staged, opt_s)):
            torch.manual_seed(100 + s)                      # the batch shuffle draws a CPU randperm
            opt.zero_grad(set_to_none=True)
            loss = m(img_a=img_a, img_b=img_b, bg0=bg0.to(dev), bg1=bg1.to(dev), visualize=False, step=s,
                     new_epoch=False)
            loss.backward()
            opt.step()
            losses.append(loss.item())
        np.testing.assert_allclose(losses[1], losses[0], rtol=1e-5)
        np.testing.assert_allclose(losses[2], losses[0], rtol=1e-5)
    se, sg = eager.state_dict(), graphed.state_dict()
    assert se.keys() == sg.keys()
    for other in (sg, staged.state_dict()):
        for k in se:
            a, b = se[k].float(), other[k].float()
            assert torch.allclose(a, b, rtol=1e-4, atol=1e-6), (k, float((a - b).abs().max()))
    assert int(sg["encoder_k.stem.1.num_batches_tracked"]) == 3 and int(sg["encoder_q.stem.1.num_batches_tracked"]) == 3
    # eval mode and another batch size run the eager path
    graphed.eval()
    with torch.no_grad():
        out = graphed.encoder_q(torch.rand(2, 3, H, H, device=dev))
        assert out.shape == (2, C, H // 16, H // 16)
        assert graphed._encode_keys(torch.rand(2, 3, H, H, device=dev)).shape == (2, C, H // 16, H // 16)
