"""TEST / BASELINE INFRASTRUCTURE ONLY — never imported by the product path (cp2_b200/).

Imports the *unmodified* reference `builder.py`: from the read-only source tree
(`/root/reference`) in this dev container, where `oracle/gen_golden.py` records golden
vectors from the reference's own code, or — on the GPU box, where that tree does not
exist — from the sourceless byte-compiled modules `oracle/build_ref.py` puts under
`oracle/_ref/` (git-ignored build artefact, the Python analogue of a compiled
`oracle/_ref/*.so`).  Users: `oracle/gen_golden.py`, `bench.py --impl reference` (the
reference's own CPU path) and bench.py's `gpu_eager_baseline` leg (the reference's eager
PyTorch path on the same B200).  The `-m gpu` tests and `smoke()` use the committed golden
vectors and `oracle/cp2_oracle.py` only.

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

_HERE = os.path.dirname(os.path.abspath(__file__))


def _find_reference():
    """(root, kind): the source tree if present, else the byte-compiled modules, else (None, None)."""
    env = os.environ.get("CP2_REFERENCE_ROOT")
    for root in ([env] if env else []) + ["/root/reference"]:
        if os.path.isfile(os.path.join(root, "builder.py")):
            return root, "source"
    pyc = os.path.join(_HERE, "_ref")
    if os.path.isfile(os.path.join(pyc, "builder.pycode")):
        return pyc, "pyc"
    return None, None


class _ByteCodeFinder:
    """Meta-path finder for the sourceless byte-code modules of oracle/_ref (`<name>.pycode` =
    a .pyc under another extension, packages as `<name>/__init__.pycode`, namespace dirs allowed)."""

    def __init__(self, root):
        self.root = root

    def find_spec(self, fullname, path=None, target=None):
        import importlib.machinery
        import importlib.util
        rel = os.path.join(self.root, *fullname.split("."))
        if os.path.isfile(rel + ".pycode"):
            return importlib.util.spec_from_loader(fullname, self, origin=rel + ".pycode")
        if os.path.isfile(os.path.join(rel, "__init__.pycode")):
            return importlib.util.spec_from_loader(fullname, self, origin=os.path.join(rel, "__init__.pycode"),
                                                   is_package=True)
        if os.path.isdir(rel) and any(f.endswith(".pycode") for f in os.listdir(rel)):
            spec = importlib.machinery.ModuleSpec(fullname, self, is_package=True)      # namespace-like (tools/)
            spec.origin = None
            return spec
        return None

    def create_module(self, spec):
        return None

    def exec_module(self, module):
        import marshal
        origin = module.__spec__.origin
        if module.__spec__.submodule_search_locations is not None:
            module.__path__ = [os.path.join(self.root, *module.__name__.split("."))]
        if origin is None:
            return
        with open(origin, "rb") as f:
            data = f.read()
        module.__file__ = origin
        exec(marshal.loads(data[16:]), module.__dict__)                                # 16-byte pyc header


REFERENCE_ROOT, REFERENCE_KIND = _find_reference()


def reference_available() -> bool:
    return REFERENCE_ROOT is not None


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
        raise RuntimeError("reference not found: neither /root/reference nor oracle/_ref/builder.pycode "
                           "(run `python oracle/build_ref.py` where the reference tree exists)")
    _install_stubs()
    if REFERENCE_KIND == "pyc":
        if not any(isinstance(f, _ByteCodeFinder) for f in sys.meta_path):
            sys.meta_path.insert(0, _ByteCodeFinder(REFERENCE_ROOT))
    elif REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import builder  # the reference's builder.py

    return builder


def ensure_process_group(backend="gloo"):
    """The reference calls torch.distributed unconditionally (builder.py:612-649, 1710-1722): a
    world-size-1 group (gloo for the CPU arm, nccl when the model runs on the GPU) through a private file
    store (independent of MASTER_* / torchrun's env)."""
    import tempfile

    import torch.distributed as dist

    if not dist.is_initialized():
        store = dist.FileStore(tempfile.NamedTemporaryFile(prefix="cp2_ref_pg_", delete=False).name, 1)
        dist.init_process_group(backend, store=store, rank=0, world_size=1)


class _Cfg:
    model = None

    def get(self, *_a, **_k):
        return None


def make_reference_model(encoder_factory, cpu=True, **kwargs):
    """Builds the reference `MODEL` (`builder.py:277`) with `encoder_factory()` standing in for both
    segmentors.  rank=1 disables every wandb call (all gated on rank==0).  cpu=True: the model is meant
    to run on the host (the caller keeps it there), so the reference's hard-coded `.cuda()` calls
    (builder.py:175,618,1420) are made the identity FOR THIS PROCESS; cpu=False (GPU eager baseline):
    nothing is patched and the caller moves the model to the device."""
    builder = import_reference()
    ensure_process_group("gloo" if cpu else "nccl")
    if cpu:
        torch.Tensor.cuda = lambda self, *a, **k: self
    _encoder_factory[0] = encoder_factory
    kwargs.setdefault("pretrain_from_scratch", True)
    kwargs.setdefault("device", torch.device("cpu"))
    model = builder.MODEL(_Cfg(), rank=1, **kwargs)
    return builder, model
