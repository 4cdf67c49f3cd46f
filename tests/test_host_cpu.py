"""Host-side logic of cp2_b200.builder.MODEL that needs no GPU: constructor validation, the epoch-end
logging / meter restart (reference builder.py:1608-1688), the new-epoch heat-map arithmetic
(builder.py:1489-1532) and the log-only score statistics (tools/correlation_mapping.py:11-53)."""
import pytest
import torch
import torch.nn as nn

from oracle import cp2_oracle as O


class _Tiny(nn.Module):
    def __init__(self, dim=128):
        super().__init__()
        self.proj = nn.Conv2d(3, dim, 16, 16)

    def forward(self, x):
        return self.proj(x)


def _model(**kw):
    from cp2_b200.builder import MODEL
    kw.setdefault("encoder_factory", _Tiny)
    kw.setdefault("output_stride", 16)
    kw.setdefault("K", 256)
    return MODEL(None, kw.pop("rank", 0), **kw)


def test_constructor_validates_kernel_limits():
    with pytest.raises(ValueError, match="dim"):
        _model(dim=64)
    with pytest.raises(ValueError, match="multiple of 4"):
        _model(K=1001)
    with pytest.raises(ValueError, match="instance_logits_temp"):
        _model(instance_logits_temp=0.01)
    with pytest.raises(ValueError, match="dense_logits_temp"):
        _model(dense_logits_temp=0.05)
    m = _model(dim=256, encoder_factory=lambda: _Tiny(256))
    assert m.queue.shape == (256, 256)


def test_backbone_output_stride_kwarg():
    m = _model(backbone_output_stride=32)
    assert m.output_stride == 16 and m.backbone_output_stride == 32
    assert _model().backbone_output_stride == 16


def test_epoch_end_logs_and_restarts_meters():
    logged = []
    m = _model(logger=logged.append)
    for epoch in range(2):
        for step in range(3):
            m.loss_o.update(torch.tensor(1.0 + epoch), 4)
            m.loss_i.update(0.5 + epoch, 4)
            m.loss_d.update(0.25, 4)
            m.acc_ins.update(torch.tensor(50.0), 4)
            m.acc_seg.update(25.0, 4)
            m.cross_image_variance_source.update(0.1, 4)
            m.cross_image_variance_target.update(0.2, 4)
        out = m.on_train_epoch_end(step=3 * (epoch + 1))
        # the second epoch's averages are the second epoch's only: the meters restarted
        assert out["train/loss"] == pytest.approx(1.0 + epoch)
        assert out["train/loss_ins"] == pytest.approx(0.5 + epoch)
        assert out["train/acc_ins"] == pytest.approx(50.0) and out["train/acc_seg"] == pytest.approx(25.0)
        assert m.loss_o.count == 0 and m.acc_seg.count == 0
    assert len(logged) == 2 and set(logged[0]) == {
        "train/loss", "train/acc_ins", "train/loss_ins", "train/loss_dense", "train/acc_seg",
        "train/cross_image_variance_source", "train/cross_image_variance_target"}
    # ranks != 0 never log (builder.py:1609) but still restart
    quiet = _model(rank=1, logger=logged.append)
    quiet.loss_o.update(3.0, 2)
    quiet.on_train_epoch_end(1)
    assert len(logged) == 2 and quiet.loss_o.count == 0


def test_heatmap_grid_matches_reference_formula():
    torchvision = pytest.importorskip("torchvision")
    B, h, w = 3, 4, 5
    g = torch.Generator().manual_seed(0)
    logits = torch.randn(B, h * w, h * w, generator=g)
    ma = (torch.rand(B, h * w, generator=g) < 0.6).float()
    mb = (torch.rand(B, h * w, generator=g) < 0.6).float()
    ma[:, 0] = 1
    mb[:, 1] = 1
    m = _model()
    got = m.heatmap_grid(logits, ma, mb, (h, w))
    # builder.py:1493-1532 restated
    hm_a, hm_b = [], []
    for i in range(B):
        b_ = logits[i][ma[i].bool(), :]
        a_ = logits[i][:, mb[i].bool()]
        hm_b.append((b_.sum(0) / b_.shape[0]).reshape(h, w))
        hm_a.append((a_.sum(1) / a_.shape[1]).reshape(h, w))
    hm_a, hm_b = torch.stack(hm_a).unsqueeze(1), torch.stack(hm_b).unsqueeze(1)
    m_a, m_b = 2 * ma.reshape(B, 1, h, w) - 1, 2 * mb.reshape(B, 1, h, w) - 1
    imgs = torch.stack([m_a, hm_a, m_b, hm_b], dim=1).flatten(0, 1)
    grid = torchvision.utils.make_grid(imgs, nrow=4)[0]
    scale = 224 // h
    ref = grid.repeat_interleave(scale, 0).repeat_interleave(scale, 1)     # NEAREST_EXACT at an integer factor
    assert got.shape == ref.shape
    assert torch.allclose(got, ref, atol=1e-6)


def test_step_score_stats_match_oracle():
    B, C, hw = 2, 128, 3
    g = torch.Generator().manual_seed(1)
    q, k = torch.randn(B, C, hw, hw, generator=g), torch.randn(B, C, hw, hw, generator=g)
    ma = (torch.rand(B, hw * hw, generator=g) < 0.6).float()
    mb = (torch.rand(B, hw * hw, generator=g) < 0.6).float()
    ma[:, 0] = 1
    mb[:, 0] = 1
    m = _model()
    qh, q_pos, _ = O.normalize_and_pool(q.reshape(B, C, -1), ma)
    kh, k_pos, _ = O.normalize_and_pool(k.reshape(B, C, -1), mb)
    st = m._step_score_stats(q, k, ma, mb, q_pos, k_pos, None)
    ref = O.dense_loss_stats(O.dense_logits(qh, kh), torch.einsum("nx,ny->nxy", ma, mb))
    for name in ("positive", "negative"):
        assert torch.allclose(st[name]["average"], ref[name]["average"], atol=1e-6)
        for i in range(3):
            assert torch.allclose(st[name]["quartiles"][i], ref[name]["quartiles"][i], atol=1e-6)
    l_neg = q_pos @ m.queue
    assert torch.allclose(st["instance_neg_q"], torch.quantile(l_neg, torch.tensor([0.25, 0.5, 0.75]), dim=1))


def test_fold_dilated_convs_is_exact():
    """MODEL.fold_dilated_convs / encoder.fold_dilated_convs: a 3x3 convolution with padding == dilation applied tap
    by tap on the windows where a tap sees data equals the stock convolution (fp64: to rounding), output and all
    three gradients; parameters and state_dict keys are untouched; large maps keep the stock kernel."""
    import copy
    import torch.nn as nn
    from cp2_b200.encoder import ASPPContrastHead, fold_dilated_convs, folded_dilated_conv2d
    torch.manual_seed(0)
    for H, W, d, bias in [(14, 14, 18, False), (14, 14, 12, False), (9, 20, 7, True), (32, 32, 18, False), (7, 7, 12, True)]:
        conv = nn.Conv2d(8, 12, 3, padding=d, dilation=d, bias=bias).double()
        ref = copy.deepcopy(conv)
        assert fold_dilated_convs(conv, max_cover=0.5) == 1 and fold_dilated_convs(conv) == 0     # idempotent
        x = torch.randn(2, 8, H, W, dtype=torch.double, requires_grad=True)
        x2 = x.detach().clone().requires_grad_(True)
        assert folded_dilated_conv2d(x, conv.weight, d, 0.5) is not None                  # the folded path runs
        y, yr = conv(x), ref(x2)
        g = torch.randn_like(y)
        y.backward(g)
        yr.backward(g)
        assert torch.allclose(y, yr, rtol=0, atol=1e-12)
        assert torch.allclose(x.grad, x2.grad, rtol=0, atol=1e-12)
        assert torch.allclose(conv.weight.grad, ref.weight.grad, rtol=0, atol=1e-11)
        if bias:
            assert torch.allclose(conv.bias.grad, ref.bias.grad, rtol=0, atol=1e-11)
        assert list(conv.state_dict()) == list(ref.state_dict())
    # a map too large for the rewrite to pay (taps cover > 50 % of the products) runs the stock convolution
    assert folded_dilated_conv2d(torch.randn(1, 8, 64, 64), torch.randn(4, 8, 3, 3), 6, 0.5) is None
    # default threshold: the centre-tap-only case folds, the 2-pixel-border case stays on the stock kernel
    assert folded_dilated_conv2d(torch.randn(1, 8, 14, 14), torch.randn(4, 8, 3, 3), 18) is not None
    assert folded_dilated_conv2d(torch.randn(1, 8, 14, 14), torch.randn(4, 8, 3, 3), 12) is None
    # the stand-in head at the 14x14 map of 224x224 views: dilation 18 folds, 6 and 12 run the stock kernel
    head = ASPPContrastHead(in_channels=16, channels=8, dim=4).double()
    ref = copy.deepcopy(head)
    assert fold_dilated_convs(head) == 3                      # 3x3 branches with dilation 6, 12, 18
    x = torch.randn(2, 16, 14, 14, dtype=torch.double)
    head.eval(), ref.eval()
    assert torch.allclose(head(x), ref(x), rtol=0, atol=1e-12)


class _TinyAspp(nn.Module):
    """Patch embedding + two dilated 3x3 branches (dilation 18 folds at 14x14, dilation 2 never does)."""

    def __init__(self, dim=128):
        super().__init__()
        self.proj = nn.Conv2d(3, 8, 16, 16)
        self.far = nn.Conv2d(8, dim, 3, padding=18, dilation=18, bias=False)
        self.near = nn.Conv2d(8, dim, 3, padding=2, dilation=2)

    def forward(self, x):
        x = self.proj(x)
        return self.far(x) + self.near(x)


def test_model_fold_dilated_convs_patches_both_encoders():
    """MODEL.fold_dilated_convs: both encoders are patched, their outputs and the checkpoint keys stay the same,
    and gradients still reach the full 3x3 weight (zero on the taps that only ever read padding)."""
    torch.manual_seed(1)
    m = _model(encoder_factory=_TinyAspp)
    keys = list(m.state_dict())
    x = torch.randn(2, 3, 224, 224)
    with torch.no_grad():
        want_q, want_k = m.encoder_q(x), m.encoder_k(x)
    assert m.fold_dilated_convs() == 4 and m.fold_dilated_convs() == 0        # 2 eligible convs per encoder
    with torch.no_grad():
        assert torch.allclose(m.encoder_q(x), want_q, rtol=1e-5, atol=1e-6)
        assert torch.allclose(m.encoder_k(x), want_k, rtol=1e-5, atol=1e-6)
    assert list(m.state_dict()) == keys
    m.encoder_q(x).square().sum().backward()
    g = m.encoder_q.far.weight.grad
    assert g.shape == (128, 8, 3, 3) and float(g[:, :, 1, 1].abs().sum()) > 0
    off_centre = g.clone()
    off_centre[:, :, 1, 1] = 0
    assert float(off_centre.abs().max()) == 0.0             # those taps multiply zero padding everywhere
    # a larger map (448 x 448 views -> 28 x 28): the dilation-18 taps see data, the stock convolution runs
    with torch.no_grad():
        big = torch.randn(1, 3, 448, 448)
        ref = nn.functional.conv2d(m.encoder_q.proj(big), m.encoder_q.far.weight, padding=18, dilation=18)
        assert torch.allclose(m.encoder_q.far(m.encoder_q.proj(big)), ref, rtol=1e-5, atol=1e-6)
