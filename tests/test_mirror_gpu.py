"""GPU parity of the mirror-variant kernels (cp2_mirror_loss, cp2_mirror_loss_upsampled) through the
C ABI against golden vectors recorded from the unmodified reference (tests/golden/mirror.npz) and
against the CPU oracle (oracle.cp2_oracle.mirror_step) on seeded inputs.

Tolerances: loss within 1e-4 relative (north_star); gradients rel-L2 <= 1e-4 and max-abs error
<= 1e-4 of the largest gradient entry (fp32 arithmetic with MUFU ex2/lg2/rcp approximations)."""
import numpy as np
import pytest
import torch

from conftest import T, load_golden
from oracle import cp2_oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    from cp2_b200 import _lib, ops as _ops
    _lib.load()
    return _ops


def rel_l2(a, b):
    a, b = a.double().cpu(), torch.as_tensor(b).double().cpu()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def close_grad(a, b):
    b = torch.as_tensor(b)
    assert rel_l2(a, b) <= 1e-4, rel_l2(a, b)
    assert float((a.cpu() - b).abs().max()) <= 1e-4 * float(b.abs().max()) + 1e-12


@pytest.mark.parametrize("tag", ["a", "b", "c", "d", "e"])
def test_golden_upsampled(ops, tag):
    g = load_golden("mirror.npz")
    lmbd, temp = (float(v) for v in g[f"{tag}_cfg"])
    none = str(g[f"{tag}_variant"]) == "NONE"
    s_lo = T(g[f"{tag}_s_lo"]).cuda().requires_grad_(True)
    t_lo = None if none else T(g[f"{tag}_t_lo"]).cuda().requires_grad_(True)
    masks = T(g[f"{tag}_masks"]).cuda()
    loss, aux = ops.mirror_loss_upsampled(s_lo, t_lo, masks, lmbd, temp)
    (loss * 3.0).backward()                               # non-unit upstream gradient
    np.testing.assert_allclose(loss.item(), g[f"{tag}_loss"], rtol=1e-4)
    np.testing.assert_allclose(aux[0].item(), g[f"{tag}_class"], rtol=1e-4)
    assert int(aux[2].item()) == masks.numel()
    close_grad(s_lo.grad / 3.0, g[f"{tag}_ds_lo"])
    if not none:
        np.testing.assert_allclose(aux[1].item(), g[f"{tag}_compare"], rtol=1e-4)
        close_grad(t_lo.grad / 3.0, g[f"{tag}_dt_lo"])


@pytest.mark.parametrize("tag", ["a", "b", "c", "d", "e"])
def test_golden_full_resolution(ops, tag):
    g = load_golden("mirror.npz")
    lmbd, temp = (float(v) for v in g[f"{tag}_cfg"])
    none = str(g[f"{tag}_variant"]) == "NONE"
    masks = T(g[f"{tag}_masks"])
    size = masks.shape[-2:]
    s_hi = O.mirror_resize(T(g[f"{tag}_s_lo"]), size).cuda().requires_grad_(True)
    t_hi = None if none else O.mirror_resize(T(g[f"{tag}_t_lo"]), size).cuda().requires_grad_(True)
    loss, aux = ops.mirror_loss(s_hi, t_hi, masks.cuda(), lmbd, temp)
    (loss * 0.5).backward()
    np.testing.assert_allclose(loss.item(), g[f"{tag}_loss"], rtol=1e-4)
    close_grad(s_hi.grad * 2.0, g[f"{tag}_ds_hi"])
    if not none:
        close_grad(t_hi.grad * 2.0, g[f"{tag}_dt_hi"])
    # one-shot ABI mode: loss and gradients from a single launch, upstream gradient 1
    out, ds, dt = ops.mirror_loss_raw(s_hi.detach(), None if none else t_hi.detach(), masks.cuda(), lmbd, temp,
                                      want_grads=True)
    np.testing.assert_allclose(out[0].item(), g[f"{tag}_loss"], rtol=1e-4)
    close_grad(ds, g[f"{tag}_ds_hi"])


@pytest.mark.parametrize("B,Cls,h,w,H,W", [(8, 2, 32, 32, 512, 512),      # BASELINE configs[4] at OS16
                                            (2, 2, 64, 64, 512, 512),      # ... and the OS8 64x64 map
                                            (3, 4, 14, 14, 224, 224), (1, 8, 9, 5, 131, 77),
                                            (2, 2, 16, 16, 16, 16), (2, 3, 20, 12, 30, 50),
                                            # two-class difference kernel with 32 / 64 rows per CTA: ragged last
                                            # row tile and dead lanes in the last column tile; the B=32 bench shape
                                            (20, 2, 16, 16, 500, 390), (32, 2, 32, 32, 512, 512),
                                            (1, 2, 7, 5, 45, 131)])
def test_upsampled_vs_oracle(ops, B, Cls, h, w, H, W):
    gen = torch.Generator().manual_seed(B * 100 + Cls * 10 + h)
    s_lo = 2.5 * torch.randn(B, Cls, h, w, generator=gen)
    t_lo = 2.5 * torch.randn(B, Cls, h, w, generator=gen)
    masks = torch.randint(0, Cls, (B, H, W), generator=gen)
    lmbd, temp = 0.01, 2.0
    so, to = s_lo.clone().requires_grad_(True), t_lo.clone().requires_grad_(True)
    ref, ref_cls, ref_cmp = O.mirror_step(so, to, masks, lmbd, temp)
    ref.backward()
    sg, tg = s_lo.cuda().requires_grad_(True), t_lo.cuda().requires_grad_(True)
    loss, aux = ops.mirror_loss_upsampled(sg, tg, masks.cuda(), lmbd, temp)
    loss.backward()
    np.testing.assert_allclose(loss.item(), ref.item(), rtol=1e-4)
    np.testing.assert_allclose(aux[0].item(), ref_cls.item(), rtol=1e-4)
    np.testing.assert_allclose(aux[1].item(), ref_cmp.item(), rtol=1e-4)
    close_grad(sg.grad, so.grad)
    close_grad(tg.grad, to.grad)
    # the full-resolution kernel on the resized logits agrees as well
    s_hi = O.mirror_resize(s_lo, (H, W)).cuda()
    t_hi = O.mirror_resize(t_lo, (H, W)).cuda()
    out, _, _ = ops.mirror_loss_raw(s_hi, t_hi, masks.cuda(), lmbd, temp, want_grads=False)
    np.testing.assert_allclose(out[0].item(), ref.item(), rtol=1e-4)


def test_ignore_index_and_valid_count(ops):
    """nn.CrossEntropyLoss() ignores label -100 (mirror_network.py:30): the class mean runs over the
    valid pixels only; the compare term still runs over all of them."""
    gen = torch.Generator().manual_seed(5)
    B, Cls, h, w, H, W = 2, 3, 6, 6, 48, 40
    s_lo = torch.randn(B, Cls, h, w, generator=gen)
    t_lo = torch.randn(B, Cls, h, w, generator=gen)
    masks = torch.randint(0, Cls, (B, H, W), generator=gen)
    masks[torch.rand(B, H, W, generator=gen) < 0.3] = -100
    n_valid = int((masks >= 0).sum())
    so, to = s_lo.clone().requires_grad_(True), t_lo.clone().requires_grad_(True)
    ref, _, _ = O.mirror_step(so, to, masks, 0.2, 1.0)
    ref.backward()
    sg, tg = s_lo.cuda().requires_grad_(True), t_lo.cuda().requires_grad_(True)
    loss, aux = ops.mirror_loss_upsampled(sg, tg, masks.cuda(), 0.2, 1.0, n_valid=n_valid)
    loss.backward()
    assert int(aux[2].item()) == n_valid
    np.testing.assert_allclose(loss.item(), ref.item(), rtol=1e-4)
    close_grad(sg.grad, so.grad)
    close_grad(tg.grad, to.grad)
    s_hi = O.mirror_resize(s_lo, (H, W)).cuda().requires_grad_(True)
    t_hi = O.mirror_resize(t_lo, (H, W)).cuda().requires_grad_(True)
    loss2, aux2 = ops.mirror_loss(s_hi, t_hi, masks.cuda(), 0.2, 1.0, n_valid=n_valid)
    np.testing.assert_allclose(loss2.item(), ref.item(), rtol=1e-4)
    assert int(aux2[2].item()) == n_valid


def test_two_class_ignore_index_large_batch(ops):
    """The two-class kernel (logit differences, 64 rows per CTA at this size) with ignored labels,
    a strong compare weight and T != 1; and without a teacher."""
    gen = torch.Generator().manual_seed(9)
    B, Cls, h, w, H, W = 40, 2, 12, 10, 384, 300
    s_lo = 3.0 * torch.randn(B, Cls, h, w, generator=gen)
    t_lo = 3.0 * torch.randn(B, Cls, h, w, generator=gen)
    masks = torch.randint(0, Cls, (B, H, W), generator=gen)
    masks[torch.rand(B, H, W, generator=gen) < 0.25] = -100
    n_valid = int((masks >= 0).sum())
    so, to = s_lo.clone().requires_grad_(True), t_lo.clone().requires_grad_(True)
    ref, _, _ = O.mirror_step(so, to, masks, 0.7, 0.5)
    ref.backward()
    sg, tg = s_lo.cuda().requires_grad_(True), t_lo.cuda().requires_grad_(True)
    loss, aux = ops.mirror_loss_upsampled(sg, tg, masks.cuda(), 0.7, 0.5, n_valid=n_valid)
    loss.backward()
    assert int(aux[2].item()) == n_valid
    np.testing.assert_allclose(loss.item(), ref.item(), rtol=1e-4)
    close_grad(sg.grad, so.grad)
    close_grad(tg.grad, to.grad)
    # no teacher: class CE only (mirror_network.py:47-52 with one network)
    out, ds, dt = ops.mirror_loss_upsampled_raw(s_lo.cuda(), None, masks.cuda(), 0.7, 0.5, True, n_valid=n_valid)
    assert dt is None
    s2 = s_lo.clone().requires_grad_(True)
    ce = torch.nn.functional.cross_entropy(O.mirror_resize(s2, (H, W)), masks)
    ce.backward()
    np.testing.assert_allclose(out[0].item(), ce.item(), rtol=1e-4)
    close_grad(ds, s2.grad)


def test_mirror_module_shared_step(ops):
    """MirrorModule.shared_step end to end (segmentor -> fused resize+loss -> backward) against the
    oracle run on the same segmentor weights."""
    import copy
    import torch.nn as nn
    from cp2_b200.mirror_network import MirrorModule, MirrorVariant, Stage
    torch.manual_seed(0)
    make = lambda: nn.Sequential(nn.Conv2d(3, 8, 3, stride=4, padding=1), nn.ReLU(), nn.Conv2d(8, 2, 3, stride=4, padding=1))
    mod = MirrorModule(None, "RANDOM", 1e-3, 1e-4, 2, (3, 128, 128), 0.01, 2.0, MirrorVariant.OUTPUT,
                       segmentor_factory=make)
    ref_net = copy.deepcopy(mod.model)
    mod = mod.cuda()
    s_img, t_img = torch.rand(2, 3, 128, 128), torch.rand(2, 3, 128, 128)
    masks = (torch.rand(2, 128, 128) > 0.6).long()
    loss = mod.training_step((s_img.cuda(), t_img.cuda(), masks.cuda()), 0)
    loss.backward()
    ref, ref_cls, ref_cmp = O.mirror_step(ref_net(s_img), ref_net(t_img), masks, 0.01, 2.0)
    ref.backward()
    np.testing.assert_allclose(loss.item(), ref.item(), rtol=1e-4)
    np.testing.assert_allclose(mod.logged["train_compare_loss"].item(), ref_cmp.item(), rtol=1e-4)
    for p, q in zip(mod.model.parameters(), ref_net.parameters()):
        assert rel_l2(p.grad, q.grad) <= 1e-3
    assert "optimizer" in mod.configure_optimizers()
    assert mod.shared_step((s_img.cuda(), t_img.cuda(), masks.cuda()), Stage.VAL).ndim == 0
