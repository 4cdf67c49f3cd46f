"""TEST INFRASTRUCTURE ONLY — never imported by the product path (cp2_b200/).

Imports the *unmodified* reference (`/root/reference/builder.py`, read-only) in this
dev container so that `oracle/gen_golden.py` can record golden vectors from the
reference's own code.  The reference cannot travel to the GPU box, so nothing that
runs there (`-m gpu` tests, `smoke()`, `bench.py`) may import this module.

Recipe: the reference imports several non-arithmetic
packages at module level (`builder.py:7-27`, `networks/segment_network.py:3-9`,
`tools/correlation_mapping.py:6`) that are absent from this image; they are replaced
by empty stub modules.  `mmseg.models.build_segmentor` is replaced by a factory for
a caller-supplied encoder so that the *loss path* (`builder.py:1124-1606`) runs on
known feature maps.  All arithmetic on the path is the reference's own.
"""
import os
import sys
import types

import torch
import torch.nn as nn
import torch.nn.functional as F

REFERENCE_ROOT = os.environ.get("CP2_REFERENCE_ROOT", "/root/reference")


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "builder.py"))


class PatchEncoder(nn.Module):
    """Tiny stand-in encoder: one strided conv `3 -> dim`, output stride `os_`.

    It plays the role of `build_segmentor(cfg.model)` (`builder.py:366-371`): callable
    on images, has `.backbone(x)` returning a 4-tuple (probed at `builder.py:398-401`)
    and `.backbone.init_weights()`.
    """

    class _Backbone(nn.Module):
        def __init__(self, os_):
            super().__init__()
            self.os_ = os_

        def forward(self, x):
            y = F.avg_pool2d(x, self.os_)
            return (y, y, y, y)

        def init_weights(self):
            pass

    def __init__(self, dim=128, os_=16):
        super().__init__()
        self.backbone = PatchEncoder._Backbone(os_)
        self.proj = nn.Conv2d(3, dim, kernel_size=os_, stride=os_)

    def forward(self, x):
        return self.proj(x)


_encoder_factory = [lambda: PatchEncoder()]


def _install_stubs():
    def mod(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
        return m

    if "matplotlib" not in sys.modules or not hasattr(sys.modules["matplotlib"], "__file__"):
        mpl = mod("matplotlib")
        mpl.cm = mod("matplotlib.cm")
        mpl.pyplot = mod("matplotlib.pyplot")
    mod("segmentation_models_pytorch")
    mm = mod("mmseg")
    mm.models = mod("mmseg.models", build_segmentor=lambda *a, **k: _encoder_factory[0]())
    mm.models.decode_heads = mod("mmseg.models.decode_heads", FCNHead=object)
    mm.models.utils = mod(
        "mmseg.models.utils",
        resize=lambda input, size, mode, align_corners: F.interpolate(
            input, size=size, mode=mode, align_corners=align_corners
        ),
    )

    class _NoMetric:
        def __init__(self, *a, **k):
            pass

        def update(self, *a, **k):
            pass

    tm = mod(
        "torchmetrics",
        Accuracy=None, Dice=None, F1Score=None, JaccardIndex=None,
        MetricCollection=None, Precision=None, Recall=None,
    )
    tm.aggregation = mod("torchmetrics.aggregation", MeanMetric=_NoMetric)
    mod("lightning", LightningModule=object)
    if "wandb" not in sys.modules:
        try:
            import wandb  # noqa: F401
        except Exception:
            mod("wandb", log=lambda *a, **k: None, define_metric=lambda *a, **k: None,
                Image=lambda *a, **k: None)


def import_reference():
    """Returns the reference's `builder` module (unmodified source)."""
    if not reference_available():
        raise RuntimeError(f"reference not found under {REFERENCE_ROOT}")
    _install_stubs()
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import builder  # the reference's builder.py

    return builder


def ensure_process_group():
    import torch.distributed as dist

    if not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29577")
        dist.init_process_group("gloo", rank=0, world_size=1)


class _Cfg:
    model = None

    def get(self, *_a, **_k):
        return None


def make_reference_model(encoder_factory, **kwargs):
    """Builds the reference `MODEL` (`builder.py:277`) on CPU with `encoder_factory()`
    standing in for both segmentors.  rank=1 disables every wandb call (all gated on
    rank==0)."""
    builder = import_reference()
    ensure_process_group()
    # the reference hard-codes `.cuda()` (builder.py:175,618,1420); identity on CPU
    if not torch.cuda.is_available():
        torch.Tensor.cuda = lambda self, *a, **k: self
    _encoder_factory[0] = encoder_factory
    kwargs.setdefault("pretrain_from_scratch", True)
    kwargs.setdefault("device", torch.device("cpu"))
    model = builder.MODEL(_Cfg(), rank=1, **kwargs)
    return builder, model
