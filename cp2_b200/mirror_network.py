"""Drop-in for the reference's `networks/mirror_network.py` (MirrorModule) on the B200 path.

The reference module is a Lightning module around an mmseg segmentor; neither package is
installable here, so this is a plain `torch.nn.Module` with the same constructor arguments,
the same method names (`forward`, `shared_step`, `training_step`, `validation_step`,
`test_step`, `configure_optimizers`) and the same loss arithmetic, with the whole of
`shared_step`'s loss (mirror_network.py:47-63) and the bilinear `resize` of
`SegmentationModule.forward` (segment_network.py:219-231) running as ONE kernel of
libcp2_b200.so (`cp2_mirror_loss_upsampled`): the [B,Cls,H,W] logits are never written to HBM.

Reference quirk kept on record (SURVEY.md header, Appendix A.8): `MirrorModule.__init__` at the
reference's HEAD does not pass `use_backbone_only` to `SegmentationModule.__init__`
(mirror_network.py:21-28 vs segment_network.py:49-58) and therefore raises `TypeError`; here the
argument exists with the default the fine-tuning script uses (`False`).
"""
from enum import Enum

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import ops
from .builder import PretrainType

BACKGROUND_CLASS = 0                 # segment_network.py:11


class Stage(Enum):                   # segment_network.py:41-45
    TRAIN = 0
    VAL = 1
    TEST = 2
    PSEUDOTEST = 3


class MirrorVariant(Enum):           # datasets/pretrain_dataset.py:187-189
    NONE = 0
    OUTPUT = 1


def _default_segmentor(model_config, num_classes):
    try:
        from mmseg.models import build_segmentor  # type: ignore
        return build_segmentor(model_config.model)
    except ImportError:
        from .encoder import DeepLabV3Encoder
        return DeepLabV3Encoder(dim=num_classes)


class MirrorModule(nn.Module):
    """networks/mirror_network.py:8-86 (+ the parts of SegmentationModule it inherits).

    Extra keyword-only arguments: `segmentor_factory` (callable returning the segmentor, replaces
    `build_segmentor(model_config.model)`), `use_backbone_only` (see module docstring)."""

    def __init__(self, model_config, pretrain_type, learning_rate, weight_decay, num_classes, image_shape,
                 lmbd_compare_loss, softmax_temp, mirror_variant, *, use_backbone_only=False,
                 segmentor_factory=None):
        super().__init__()
        pretrain_type = PretrainType[getattr(pretrain_type, "name", pretrain_type)]
        if pretrain_type not in (PretrainType.RANDOM, PretrainType.NONE):
            raise NotImplementedError(
                f"{pretrain_type = }: checkpoint surgery (segment_network.py:63-160) is fine-tuning code, out of scope")
        self.model = segmentor_factory() if segmentor_factory is not None \
            else _default_segmentor(model_config, num_classes)
        if pretrain_type == PretrainType.NONE and segmentor_factory is None and hasattr(self.model, "backbone") \
                and hasattr(self.model.backbone, "init_weights"):
            self.model.backbone.init_weights()                   # segment_network.py:64-67 (ImageNet init)
        self.learning_rate = learning_rate
        self.weight_decay = weight_decay
        self.num_classes = num_classes
        self.image_shape = tuple(image_shape)                     # (C, H, W)
        self.use_backbone_only = use_backbone_only
        self.lmbd_compare_loss = lmbd_compare_loss
        self.softmax_temp = softmax_temp
        mirror_variant = MirrorVariant[getattr(mirror_variant, "name", mirror_variant)]
        self.mirror_variant = mirror_variant
        self.logged = {}                                          # stand-in for Lightning's self.log

    # segment_network.py:219-231 — kept for inference / evaluation callers; shared_step does not use it
    def forward(self, images):
        logits = self.model(images)
        logits = F.interpolate(logits, size=self.image_shape[1:], mode="bilinear", align_corners=False)
        return logits, logits.argmax(dim=1)

    def log(self, name, value, **_kw):
        self.logged[name] = value.detach() if torch.is_tensor(value) else value

    def shared_step(self, batch, stage):
        """mirror_network.py:40-63.  Returns the 0-d loss (autograd to the segmentor)."""
        size = self.image_shape[1:]
        if self.mirror_variant == MirrorVariant.OUTPUT:
            s_img, t_img, masks = batch
            assert tuple(masks.shape[-2:]) == tuple(size), f"{masks.shape = } vs {self.image_shape = }"
            s_lo = self.model(s_img)
            t_lo = self.model(t_img)
            loss, aux = ops.mirror_loss_upsampled(s_lo, t_lo, masks, self.lmbd_compare_loss, self.softmax_temp)
        elif self.mirror_variant == MirrorVariant.NONE:
            img, masks = batch
            assert tuple(masks.shape[-2:]) == tuple(size), f"{masks.shape = } vs {self.image_shape = }"
            loss, aux = ops.mirror_loss_upsampled(self.model(img), None, masks, self.lmbd_compare_loss,
                                                  self.softmax_temp)
        else:
            raise NotImplementedError(f"{self.mirror_variant = }")
        name = stage.name.lower()
        self.log(f"{name}_loss", loss)
        self.log(f"{name}_compare_loss", aux[1])
        self.log(f"{name}_class_loss", aux[0])
        return loss

    def training_step(self, batch, batch_idx):                   # segment_network.py:283-284
        return self.shared_step(batch, Stage.TRAIN)

    def validation_step(self, batch, batch_idx, dataloader_idx=0):
        return self.shared_step(batch, Stage.VAL if dataloader_idx == 0 else Stage.PSEUDOTEST)

    def test_step(self, batch, batch_idx):
        return self.shared_step(batch, Stage.TEST)

    def configure_optimizers(self):                               # segment_network.py:294-310
        optimizer = torch.optim.Adam(self.parameters(), lr=self.learning_rate, weight_decay=self.weight_decay)
        return {"optimizer": optimizer}
