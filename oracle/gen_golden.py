"""Generates tests/golden/*.npz by running the UNMODIFIED reference in this container.

TEST INFRASTRUCTURE ONLY.  Run once here (`python oracle/gen_golden.py`); the
resulting small fixtures are committed, because `/root/reference` does not exist on
the GPU box.  Every tensor stored below is an input to, or an output of, the
reference's own code:

  cp2_step_small.npz   two full `MODEL.forward_cp2` + backward steps (builder.py:1124)
                       with a one-conv encoder: B=4, 64x64 views, OS16 (P=16), C=128,
                       K=256 -> composite, losses, grads, EMA'd key weights, queue, ptr.
  cp2_loss_p196_*.npz  the loss path on given 14x14 feature maps (P=196, C=128,
                       K=1024) for tau_d in {1.0, 0.2} and include_background on/off.
  ema.npz              `_momentum_update_key_encoder` (builder.py:557) on odd-sized tensors.
  enqueue.npz          `_dequeue_and_enqueue` (builder.py:569) incl. wrap-around, K=100.
  corr.npz             `get_masked_correlation_map` / `calcuate_dense_loss_stats`
                       (tools/correlation_mapping.py) on the reference's own test cases.
"""
import os
import sys

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_harness as rh  # noqa: E402
from oracle.cp2_oracle import synth_views  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def _np(t):
    return t.detach().cpu().numpy().copy()


class _Capture:
    """Records encoder inputs/outputs and the unshuffle permutation without touching
    the reference's arithmetic."""

    def __init__(self, model):
        self.rec = {}
        model.encoder_q.register_forward_hook(self._q)
        model.encoder_k.register_forward_hook(self._k)
        orig = model._batch_shuffle_ddp

        def wrapped(x):
            y, idx = orig(x)
            self.rec["idx_unshuffle"] = idx.clone()
            return y, idx

        model._batch_shuffle_ddp = wrapped

    def _q(self, _m, inp, out):
        self.rec["img_a_comp"] = inp[0].detach().clone()
        out.retain_grad()
        self.rec["q"] = out

    def _k(self, _m, inp, out):
        self.rec["k_shuffled"] = out.detach().clone()


def gen_step_small():
    torch.manual_seed(0)
    B, H, C, K = 4, 64, 128, 256
    _, model = rh.make_reference_model(lambda: rh.PatchEncoder(C, 16), K=K, dim=C,
                                       dense_logits_temp=0.2)
    cap = _Capture(model)
    # make encoder_q differ from encoder_k so the EMA is not a no-op
    g = torch.Generator().manual_seed(5)
    with torch.no_grad():
        for p in model.encoder_q.parameters():
            p.add_(0.05 * torch.randn(p.shape, generator=g))
    out = {"queue0": _np(model.queue), "B": B, "H": H, "C": C, "K": K,
           "temp_local": 0.2, "temp_global": 0.2, "lmbd": 0.2, "m": 0.999}
    for name, p in model.encoder_q.named_parameters():
        out[f"q0.{name}"] = _np(p)
    for name, p in model.encoder_k.named_parameters():
        out[f"k0.{name}"] = _np(p)
    opt = torch.optim.SGD(model.encoder_q.parameters(), lr=0.05)
    for step in range(2):
        v = synth_views(B, H, H, 16, seed=100 + step)
        for k_, t in v.items():
            if "ids" not in k_:      # ids are arange(1, H*W+1) per image (synth_views)
                out[f"s{step}.{k_}"] = _np(t)
        opt.zero_grad()
        loss = model(visualize=False, step=step, new_epoch=False, **v)
        loss.backward()
        out[f"s{step}.loss"] = _np(loss)
        out[f"s{step}.loss_ins"] = np.float32(model.loss_i.val)
        out[f"s{step}.loss_dense"] = np.float32(model.loss_d.val)
        out[f"s{step}.acc_ins"] = np.float32(float(model.acc_ins.val))
        out[f"s{step}.acc_seg"] = np.float32(model.acc_seg.val)
        out[f"s{step}.img_a_comp"] = _np(cap.rec["img_a_comp"])
        out[f"s{step}.q"] = _np(cap.rec["q"])
        out[f"s{step}.grad_q"] = _np(cap.rec["q"].grad)
        out[f"s{step}.k"] = _np(cap.rec["k_shuffled"][cap.rec["idx_unshuffle"]])
        out[f"s{step}.queue"] = _np(model.queue)
        out[f"s{step}.queue_ptr"] = _np(model.queue_ptr)
        for name, p in model.encoder_k.named_parameters():
            out[f"s{step}.k.{name}"] = _np(p)
        for name, p in model.encoder_q.named_parameters():
            out[f"s{step}.gradw.{name}"] = _np(p.grad)
        opt.step()
    np.savez_compressed(os.path.join(OUT, "cp2_step_small.npz"), **out)
    print("cp2_step_small", out["s0.loss"], out["s1.loss"], out["s1.queue_ptr"])


class _FeatEncoder(nn.Module):
    """Returns a stored feature map whatever the input (identity 'encoder')."""

    class _BB(nn.Module):
        def forward(self, x):
            y = torch.zeros(x.shape[0], 1, 14, 14)
            return (y, y, y, y)

        def init_weights(self):
            pass

    def __init__(self, B, C, h):
        super().__init__()
        self.feat = nn.Parameter(torch.zeros(B, C, h, h))
        self.backbone = _FeatEncoder._BB()

    def forward(self, x):
        return self.feat[: x.shape[0]] * 1.0


def gen_loss_p196(tag, temp_local, include_background):
    torch.manual_seed(1)
    B, C, h, K = 2, 128, 14, 1024
    _, model = rh.make_reference_model(lambda: _FeatEncoder(B, C, h), K=K, dim=C,
                                       dense_logits_temp=temp_local,
                                       include_background=include_background)
    assert model.output_stride == 16
    cap = _Capture(model)
    g = torch.Generator().manual_seed(7)
    with torch.no_grad():
        model.encoder_q.feat.copy_(torch.randn(B, C, h, h, generator=g))
        model.encoder_k.feat.copy_(torch.randn(B, C, h, h, generator=g))
    v = synth_views(B, 224, 224, 16, seed=42)
    out = {"queue0": _np(model.queue), "temp_local": temp_local, "temp_global": 0.2,
           "lmbd": 0.2, "include_background": include_background,
           "bg0_ch0_zero": _np(v["bg0"][:, 0] == 0), "bg1_ch0_zero": _np(v["bg1"][:, 0] == 0)}
    loss = model(visualize=False, step=0, new_epoch=False, **v)
    loss.backward()
    out["q"] = _np(cap.rec["q"])
    out["k"] = _np(cap.rec["k_shuffled"][cap.rec["idx_unshuffle"]])
    out["grad_q"] = _np(model.encoder_q.feat.grad)
    out["loss"] = _np(loss)
    out["loss_ins"] = np.float32(model.loss_i.val)
    out["loss_dense"] = np.float32(model.loss_d.val)
    out["acc_seg"] = np.float32(model.acc_seg.val)
    out["acc_ins"] = np.float32(float(model.acc_ins.val))
    out["queue1_head"] = _np(model.queue[:, :8])   # only columns 0..B-1 changed
    out["queue_ptr1"] = _np(model.queue_ptr)
    out["iou"] = np.asarray(model.correlation_ious, dtype=np.float32)
    out["iou_masked"] = np.asarray(model.masked_correlation_ious, dtype=np.float32)
    np.savez_compressed(os.path.join(OUT, f"cp2_loss_p196_{tag}.npz"), **out)
    print("cp2_loss_p196", tag, out["loss"], out["loss_ins"], out["loss_dense"])


class _Holder(nn.Module):
    def __init__(self, shapes, seed):
        super().__init__()
        g = torch.Generator().manual_seed(seed)
        self.ps = nn.ParameterList([nn.Parameter(torch.randn(s, generator=g)) for s in shapes])
        self.backbone = _FeatEncoder._BB()
        self.h = nn.Parameter(torch.zeros(1, 1, 14, 14))

    def forward(self, x):
        return self.h.expand(x.shape[0], 1, 14, 14)


def gen_ema():
    shapes = [(64, 3, 7, 7), (64,), (1,), (257, 33), (5, 1, 3), (100003,), (128, 512, 1, 1), (19,)]
    seeds = iter([11, 12])
    _, model = rh.make_reference_model(lambda: _Holder(shapes, next(seeds)), K=16, dim=8)
    g = torch.Generator().manual_seed(13)
    with torch.no_grad():
        for p in model.encoder_k.ps:
            p.copy_(torch.randn(p.shape, generator=g) * 3.0)
        # specials: zeros, denormals, negative zero, large
        model.encoder_q.ps[1][:4] = torch.tensor([0.0, -0.0, 1e-40, 3e38])
        model.encoder_k.ps[1][:4] = torch.tensor([-0.0, 0.0, -1e-41, 3e38])
    out = {"m": 0.999, "n": len(shapes)}
    for i, (pq, pk) in enumerate(zip(model.encoder_q.ps, model.encoder_k.ps)):
        out[f"q{i}"], out[f"k{i}"] = _np(pq), _np(pk)
    model._momentum_update_key_encoder()
    for i, pk in enumerate(model.encoder_k.ps):
        out[f"k1_{i}"] = _np(pk)
    model.momentum = 0.99
    model._momentum_update_key_encoder()
    for i, pk in enumerate(model.encoder_k.ps):
        out[f"k2_{i}"] = _np(pk)
    np.savez_compressed(os.path.join(OUT, "ema.npz"), **out)
    print("ema ok")


def gen_enqueue():
    _, model = rh.make_reference_model(lambda: _Holder([(2,)], 1), K=100, dim=16)
    g = torch.Generator().manual_seed(21)
    out = {"queue0": _np(model.queue)}
    batches = [32, 32, 32, 32, 7, 100, 1, 29]   # wraps at step 3, exact fill, single row
    out["batches"] = np.asarray(batches)
    for i, n in enumerate(batches):
        keys = torch.randn(n, 16, generator=g)
        out[f"keys{i}"] = _np(keys)
        model._dequeue_and_enqueue(keys)
        out[f"queue{i + 1}"] = _np(model.queue)
        out[f"ptr{i + 1}"] = _np(model.queue_ptr)
    np.savez_compressed(os.path.join(OUT, "enqueue.npz"), **out)
    print("enqueue ok", out[f"ptr{len(batches)}"])


def gen_corr():
    rh.import_reference()
    from tools.correlation_mapping import (calcuate_dense_loss_stats,
                                           get_masked_correlation_map)
    out = {}
    # reference test case "uniqueIds" (tests/test_correlation_mapping.py:27-76), fixed shuffle
    bs, hh, ww = 4, 10, 10
    rng = np.random.RandomState(3)
    base = torch.arange(1, bs * hh * ww + 1)[torch.from_numpy(rng.permutation(bs * hh * ww))]
    base = base.reshape(bs, hh, ww)
    map_a, map_b = base[:, :5, :5], base[:, 1:6, 2:7]
    mask_a = torch.zeros(bs, 5, 5); mask_a[:, 2:4, 1:3] = 1
    mask_b = torch.zeros(bs, 5, 5); mask_b[:, 1:3, 0:2] = 1
    r = get_masked_correlation_map(map_a=map_a, map_b=map_b, mask_a=mask_a, mask_b=mask_b)
    out.update(u_map_a=_np(map_a), u_map_b=_np(map_b), u_mask_a=_np(mask_a), u_mask_b=_np(mask_b),
               u_iou=_np(r["iou"]), u_iou_masked=_np(r["iou_masked"]), u_corr=_np(r["corr_map"]))
    # reference test case "sharedIds" (tests/test_correlation_mapping.py:81-129)
    base = torch.Tensor([[[1, 2, 2, 3, 4, 5], [6, 2, 2, 3, 3, 3],
                          [7, 8, 9, 10, 11, 12], [13, 8, 8, 8, 14, 15]]])
    map_a, map_b = base[:, 0:3, 1:4], base[:, 0:3, 2:5]
    mask_a = torch.Tensor([[[1, 1, 1], [1, 1, 1], [0, 0, 0]]])
    mask_b = torch.Tensor([[[1, 0, 0], [1, 0, 0], [1, 0, 0]]])
    r = get_masked_correlation_map(map_a=map_a, map_b=map_b, mask_a=mask_a, mask_b=mask_b)
    out.update(s_map_a=_np(map_a), s_map_b=_np(map_b), s_mask_a=_np(mask_a), s_mask_b=_np(mask_b),
               s_iou=_np(r["iou"]), s_iou_masked=_np(r["iou_masked"]), s_corr=_np(r["corr_map"]))
    # dense stats
    g = torch.Generator().manual_seed(9)
    logits = torch.randn(3, 12, 12, generator=g)
    labels = (torch.rand(3, 12, generator=g) > 0.4).float()[:, :, None] * \
             (torch.rand(3, 12, generator=g) > 0.4).float()[:, None, :]
    st = calcuate_dense_loss_stats(logits, labels)
    out.update(d_logits=_np(logits), d_labels=_np(labels))
    for side in ("positive", "negative"):
        out[f"d_{side}_avg"] = _np(st[side]["average"])
        out[f"d_{side}_q"] = _np(torch.stack(list(st[side]["quartiles"])))
    np.savez_compressed(os.path.join(OUT, "corr.npz"), **out)
    print("corr ok", out["u_iou"][0], out["u_iou_masked"][0], out["s_iou"], out["s_iou_masked"])


class _MirrorSelf:
    """The attributes `MirrorModule.shared_step` (networks/mirror_network.py:40-86) reads.  The
    reference's own `__init__` cannot run (TypeError at HEAD, SURVEY.md header), so its UNMODIFIED
    `shared_step` and `SegmentationModule.forward` are called unbound on this holder."""

    def __init__(self, ref_seg, ref_mirror, variant, lo_by_image, image_shape, lmbd, temp):
        self._seg = ref_seg
        self.mirror_variant = variant
        self.image_shape = image_shape
        self.lmbd_compare_loss = lmbd
        self.softmax_temp = temp
        self.class_loss = nn.CrossEntropyLoss()          # mirror_network.py:30-32
        self.compare_loss = nn.CrossEntropyLoss()
        self.softmax = nn.Softmax(dim=1)
        self.logged = {}
        self._lo = lo_by_image
        self.model = lambda img: self._lo[id(img)]       # "segmentor": returns the prepared low-res logits
        self.train_metrics = type("M", (), {"update": lambda *a, **k: None, "items": lambda self_: []})()

    def forward(self, images):                            # segment_network.py:219-231, unmodified
        return self._seg.SegmentationModule.forward(self, images)

    def log(self, name, value, **_k):
        self.logged[name] = value

    def log_dict(self, *_a, **_k):
        pass


def gen_mirror():
    rh.import_reference()
    import types
    for name, attrs in (("albumentations", {}),):
        if name not in sys.modules:
            m = types.ModuleType(name)
            m.__dict__.update(attrs)
            sys.modules[name] = m
    sys.modules["lightning"].LightningDataModule = object
    import networks.mirror_network as ref_mirror
    import networks.segment_network as ref_seg
    out = {}
    cases = [  # tag, B, Cls, h, w, H, W, lmbd, T, variant
        ("a", 2, 2, 4, 4, 64, 64, 0.01, 2.0, "OUTPUT"),        # the reference's defaults (mirror_pretrain.py:50-51), x16
        ("b", 3, 2, 5, 7, 37, 45, 0.5, 0.7, "OUTPUT"),         # ragged, non-integer scale
        ("c", 2, 5, 6, 6, 24, 24, 0.1, 1.5, "OUTPUT"),         # multi-class
        ("d", 2, 3, 8, 8, 8, 8, 0.3, 2.0, "OUTPUT"),           # no resize
        ("e", 2, 2, 4, 4, 64, 64, 0.01, 2.0, "NONE"),          # MirrorVariant.NONE
    ]
    g = torch.Generator().manual_seed(77)
    for tag, B, Cls, h, w, H, W, lmbd, temp, variant in cases:
        s_lo = (3.0 * torch.randn(B, Cls, h, w, generator=g)).requires_grad_()
        t_lo = (3.0 * torch.randn(B, Cls, h, w, generator=g)).requires_grad_()
        masks = torch.randint(0, Cls, (B, H, W), generator=g)
        s_img, t_img = torch.zeros(1), torch.zeros(1)
        holder = _MirrorSelf(ref_seg, ref_mirror, ref_mirror.MirrorVariant[variant],
                             {id(s_img): s_lo, id(t_img): t_lo}, (3, H, W), lmbd, temp)
        batch = (s_img, t_img, masks) if variant == "OUTPUT" else (s_img, masks)
        loss = ref_mirror.MirrorModule.shared_step(holder, batch, ref_seg.Stage.TRAIN)
        loss.backward()
        out.update({f"{tag}_s_lo": _np(s_lo), f"{tag}_masks": _np(masks), f"{tag}_loss": _np(loss),
                    f"{tag}_class": _np(holder.logged["train_class_loss"]),
                    f"{tag}_ds_lo": _np(s_lo.grad),
                    f"{tag}_cfg": np.asarray([lmbd, temp], dtype=np.float64),
                    f"{tag}_variant": np.asarray(variant)})
        if variant == "OUTPUT":
            out.update({f"{tag}_t_lo": _np(t_lo), f"{tag}_dt_lo": _np(t_lo.grad),
                        f"{tag}_compare": _np(holder.logged["train_compare_loss"])})
        # full-resolution logits and their gradients, for the non-fused kernel
        s_hi = F.interpolate(s_lo.detach(), size=(H, W), mode="bilinear", align_corners=False).requires_grad_()
        t_hi = F.interpolate(t_lo.detach(), size=(H, W), mode="bilinear", align_corners=False).requires_grad_()
        holder2 = _MirrorSelf(ref_seg, ref_mirror, ref_mirror.MirrorVariant[variant], {}, (3, H, W), lmbd, temp)
        holder2.forward = lambda images, _m={id(s_img): s_hi, id(t_img): t_hi}: (_m[id(images)], None)
        loss2 = ref_mirror.MirrorModule.shared_step(holder2, batch, ref_seg.Stage.TRAIN)
        loss2.backward()
        assert abs(float(loss2) - float(loss)) < 1e-6
        out[f"{tag}_ds_hi"] = _np(s_hi.grad)
        if variant == "OUTPUT":
            out[f"{tag}_dt_hi"] = _np(t_hi.grad)
        print("mirror", tag, float(loss))
    out["tags"] = np.asarray([c[0] for c in cases])
    np.savez_compressed(os.path.join(OUT, "mirror.npz"), **out)


def proposed_ids(batch, height, width, seed):
    """Pixel / region id maps for the PROPOSED-mode fixtures: view b is view a shifted by a multiple
    of the output stride (so pixel ids meet on the [OS/2::OS] lattice), out-of-view pixels get fresh
    ids; regions are 64x64 blocks, a few of them "unknown" (id 0, builder.py:1221-1232)."""
    g = torch.Generator().manual_seed(seed)
    pa = torch.arange(1, height * width + 1, dtype=torch.int64).reshape(1, height, width).repeat(batch, 1, 1)
    pb = torch.zeros_like(pa)
    ra = torch.zeros_like(pa)
    rb = torch.zeros_like(pa)
    yy, xx = torch.meshgrid(torch.arange(height), torch.arange(width), indexing="ij")
    for i in range(batch):
        dy, dx = (int(v) * 16 for v in torch.randint(-3, 4, (2,), generator=g))
        sy, sx = yy + dy, xx + dx
        inside = (sy >= 0) & (sy < height) & (sx >= 0) & (sx < width)
        fresh = height * width + 1 + yy * width + xx
        pb[i] = torch.where(inside, pa[i][sy.clamp(0, height - 1), sx.clamp(0, width - 1)], fresh)
        blocks = (yy // 64) * ((width + 63) // 64) + xx // 64          # block 0 is "unknown"
        drop = torch.randint(1, int(blocks.max()) + 1, (2,), generator=g)
        ra[i] = torch.where((blocks == drop[0]), torch.zeros_like(blocks), blocks)
        rb_full = torch.where((blocks == drop[1]), torch.zeros_like(blocks), blocks)
        rb[i] = torch.where(inside, rb_full[sy.clamp(0, height - 1), sx.clamp(0, width - 1)], torch.zeros_like(blocks))
    return {"pixel_ids_a": pa, "pixel_ids_b": pb, "region_ids_a": ra, "region_ids_b": rb}


def gen_weighted(tag, negative_type, temp_local, weights=(10.0, 2.0, 0.5), include_background=False):
    """`forward_cp2` in PretrainType.PROPOSED / MappingType.PIXEL_REGION_ID (builder.py:1204-1243,
    1332-1392) on given 14x14 feature maps -> tests/golden/cp2_weighted_<tag>.npz."""
    torch.manual_seed(1)
    B, C, h, K = 2, 128, 14, 1024
    builder, model = rh.make_reference_model(
        lambda: _FeatEncoder(B, C, h), K=K, dim=C, dense_logits_temp=temp_local,
        include_background=include_background, pretrain_type=rh.import_reference().PretrainType.PROPOSED,
        mapping_type=rh.import_reference().MappingType.PIXEL_REGION_ID,
        negative_type=rh.import_reference().NegativeType[negative_type], negative_scale=2.0,
        lmbd_pixel_corr_weight=weights[0], lmbd_region_corr_weight=weights[1], lmbd_not_corr_weight=weights[2])
    cap = _Capture(model)
    g = torch.Generator().manual_seed(11)
    with torch.no_grad():
        model.encoder_q.feat.copy_(torch.randn(B, C, h, h, generator=g))
        model.encoder_k.feat.copy_(torch.randn(B, C, h, h, generator=g))
    v = synth_views(B, 224, 224, 16, seed=43)
    v.update(proposed_ids(B, 224, 224, seed=5))
    sub = lambda t: t[:, 8::16, 8::16]
    out = {"queue0": _np(model.queue), "temp_local": temp_local, "temp_global": 0.2, "lmbd": 0.2,
           "include_background": include_background, "negative_type": negative_type, "negative_scale": 2.0,
           "weights": np.asarray(weights, dtype=np.float32),
           "bg0_ch0_zero": _np(v["bg0"][:, 0] == 0), "bg1_ch0_zero": _np(v["bg1"][:, 0] == 0)}
    for name in ("pixel_ids_a", "pixel_ids_b", "region_ids_a", "region_ids_b"):
        out[name + "_ds"] = _np(sub(v[name]))          # the [OS/2::OS] lattice is all the path reads
    loss = model(visualize=False, step=0, new_epoch=False, **v)
    loss.backward()
    out["q"] = _np(cap.rec["q"])
    out["k"] = _np(cap.rec["k_shuffled"][cap.rec["idx_unshuffle"]])
    out["grad_q"] = _np(model.encoder_q.feat.grad)
    out["loss"] = _np(loss)
    out["loss_ins"] = np.float32(model.loss_i.val)
    out["loss_dense"] = np.float32(model.loss_d.val)
    out["iou"] = np.asarray(model.correlation_ious, dtype=np.float32)
    out["iou_masked"] = np.asarray(model.masked_correlation_ious, dtype=np.float32)
    np.savez_compressed(os.path.join(OUT, f"cp2_weighted_{tag}.npz"), **out)
    print("cp2_weighted", tag, out["loss"], out["loss_ins"], out["loss_dense"])


class _DenseCLEncoder(nn.Module):
    """Stand-in segmentor for the DenseCL modes: `.backbone(img)[3]` is a 2048-channel stride-16 map
    (DenseCLNeck's in_channels is hard-coded, builder.py:436-441), the module itself returns a 128-d map for
    the output-stride probe (builder.py:392-402)."""

    class _BB(nn.Module):
        def __init__(self):
            super().__init__()
            self.proj = nn.Conv2d(3, 2048, 16, stride=16)

        def forward(self, x):
            y = self.proj(x)
            return (y, y, y, y)

        def init_weights(self):
            pass

    def __init__(self):
        super().__init__()
        self.backbone = _DenseCLEncoder._BB()
        self.head = nn.Conv2d(2048, 128, 1)

    def forward(self, x):
        return self.head(self.backbone(x)[3])


def gen_densecl(tag, pretrain_type, lmbd_coordinate, use_predictor=False):
    """`forward_densecl` (builder.py:667-999) + backward on 96x96 views (S=6, P=36), K=512 ->
    tests/golden/densecl_<tag>.npz: the unit features the loss consumes, loss values, gradients w.r.t. the
    neck outputs, both queues after the step."""
    torch.manual_seed(3)
    B, H, K = 3, 96, 512
    ref = rh.import_reference()
    _, model = rh.make_reference_model(
        _DenseCLEncoder, K=K, dim=128, pretrain_type=ref.PretrainType[pretrain_type], lmbd_cp2_dense_loss=0.5,
        instance_logits_temp=0.2, dense_logits_temp=0.2, lmbd_coordinate=lmbd_coordinate,
        use_predictor=use_predictor)
    rec = {}

    def neck_hook(side):
        def hook(_m, _inp, out):
            for name, t in out.items():
                if side == "q" and t.requires_grad:
                    t.retain_grad()
                rec[f"{side}.{name}"] = t
        return hook

    def bb_hook(side):
        def hook(_m, _inp, out):
            rec[f"{side}.embd"] = out[3].detach().clone()
        return hook

    model.encoder_q.neck.register_forward_hook(neck_hook("q"))
    model.encoder_k.neck.register_forward_hook(neck_hook("k"))
    model.encoder_q.backbone.register_forward_hook(bb_hook("q"))
    model.encoder_k.backbone.register_forward_hook(bb_hook("k"))
    orig = model._batch_shuffle_ddp
    model._batch_shuffle_ddp = lambda x: (lambda y, idx: (rec.__setitem__("idx_unshuffle", idx.clone()), (y, idx))[1])(*orig(x))
    g = torch.Generator().manual_seed(21)
    with torch.no_grad():                                  # make the two encoders differ
        for p in model.encoder_q.parameters():
            p.add_(0.02 * torch.randn(p.shape, generator=g))
    v = synth_views(B, H, H, 16, seed=77)
    # The reference raises IndexError at builder.py:861 (`.max(dim=2)` on a 2-d tensor) as soon as ONE pixel id
    # is shared between the two lattices, so its own code can only be recorded on views without overlap.
    v["pixel_ids_b"] = v["pixel_ids_a"] + H * H
    out = {"queue0": _np(model.queue), "queue2_0": _np(model.queue2), "lmbd_coordinate": lmbd_coordinate,
           "use_predictor": use_predictor, "pretrain_type": pretrain_type,
           "pix_q": _np(v["pixel_ids_a"][:, 8::16, 8::16]), "pix_k": _np(v["pixel_ids_b"][:, 8::16, 8::16])}
    loss = model(visualize=False, step=0, new_epoch=False, **v)
    loss.backward()
    un = rec["idx_unshuffle"]
    ql = rec["q.x_local_pred"] if use_predictor else rec["q.x_local_proj"]
    qg = rec["q.x_global_pred"] if use_predictor else rec["q.x_global_proj"]
    out.update({"q_local_raw": _np(ql), "q_global_raw": _np(qg), "grad_q_local": _np(ql.grad),
                "grad_q_global": _np(qg.grad), "embd_q": _np(rec["q.embd"]), "embd_k": _np(rec["k.embd"][un]),
                "k_local_raw": _np(rec["k.x_local_proj"][un]), "k_global_raw": _np(rec["k.x_global_proj"][un]),
                "k_pooled_raw": _np(rec["k.x_avgpool_local_proj"][un]),
                "loss": _np(loss), "loss_global": np.float32(model.loss_i.val),
                "loss_local": np.float32(model.loss_d.val), "queue1": _np(model.queue), "queue2_1": _np(model.queue2),
                "queue_ptr1": _np(model.queue_ptr), "queue2_ptr1": _np(model.queue2_ptr)})
    np.savez_compressed(os.path.join(OUT, f"densecl_{tag}.npz"), **out)
    print("densecl", tag, out["loss"], out["loss_global"], out["loss_local"])


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    if len(sys.argv) > 1 and sys.argv[1] == "densecl":
        gen_densecl("plain", "DENSECL", 0.0)
        gen_densecl("v2_pred", "PROPOSED_V2", 0.3, use_predictor=True)
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "weighted":
        gen_weighted("none_t1", "NONE", 1.0)
        gen_weighted("fixed_t05", "FIXED", 0.5)
        gen_weighted("avg_t1_bg", "AVERAGE", 1.0, include_background=True)
        gen_weighted("hard_t1", "HARD", 1.0, weights=(3.0, 1.0, 0.0))
        gen_weighted("median_t05", "MEDIAN", 0.5, weights=(5.0, 2.0, 1.0))
        sys.exit(0)
    gen_step_small()
    gen_loss_p196("t1", 1.0, False)
    gen_loss_p196("t02", 0.2, False)
    gen_loss_p196("t02_bg", 0.2, True)
    gen_ema()
    gen_enqueue()
    gen_corr()
    gen_mirror()
    gen_weighted("none_t1", "NONE", 1.0)
    gen_weighted("fixed_t05", "FIXED", 0.5)
    gen_weighted("avg_t1_bg", "AVERAGE", 1.0, include_background=True)
    gen_weighted("hard_t1", "HARD", 1.0, weights=(3.0, 1.0, 0.0))
    gen_weighted("median_t05", "MEDIAN", 0.5, weights=(5.0, 2.0, 1.0))
    gen_densecl("plain", "DENSECL", 0.0)
    gen_densecl("v2_pred", "PROPOSED_V2", 0.3, use_predictor=True)
