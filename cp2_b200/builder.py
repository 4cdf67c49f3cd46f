"""Host-side mirror of the reference's `builder.py` for the CP2 pre-training loss path.

`MODEL` keeps the reference constructor, keyword `forward(...) -> 0-d loss`, buffers
(`queue`, `queue_ptr`, `queue2`, `queue2_ptr`), sub-modules (`encoder_q`, `encoder_k`),
meters and method names (builder.py:277-304, 476-482, 543-553, 557-649, 1124-1606), so it
drops into the reference's `main.py` loop (main.py:390-461, 616-644) and its checkpoints
interchange.  Everything between the encoder outputs and the scalar loss, plus the
compositing, the EMA and the enqueue, runs in the sm_100a kernels of libcp2_b200.so
(include/cp2_b200.h); torch supplies device memory, streams, autograd plumbing and NCCL.

Scope: `forward_cp2`, i.e. PretrainType.CP2 (MappingType.CP2 / NegativeType.NONE)
and PretrainType.PROPOSED (pair weights from the id maps, every NegativeType; row f3), and `forward_densecl`
(DENSECL / PROPOSED_V2; row f2).  MOCO / BYOL / PIXPRO raise NotImplementedError here.

Deliberate, documented deviations from the reference (none changes the loss):
  * no `int(queue_ptr)` / `.item()` / `.cpu()` host syncs on the step: the queue pointer
    lives on the device and the meters accumulate on the device, syncing only when read;
  * no `queue.clone()` (builder.py:1396): the global kernel produces the backward's
    queue-weighted sum in the forward pass, so enqueue may overwrite the queue before
    `loss.backward()`;
  * the correlation IoUs (builder.py:1252-1257) come from one kernel and stay on the device until read;
  * the per-step / per-epoch wandb logs of rank 0 (builder.py:1450-1550, 1565-1604, 1608-1664) go through
    `MODEL.logger` (default: `wandb.log` when a wandb run is active on rank 0, else nothing).  The score
    statistics that only feed those logs (dense score means / quartiles, builder.py:1298; queue-logit
    quartiles, builder.py:1401-1403) and the new-epoch heat-maps (builder.py:1489-1549) are computed by a
    plain-torch branch off the hot path (`MODEL.dense_logits_for_visualisation`, `_step_score_stats`), taken
    only when somebody logs: rank 0 with an active logger, as in the reference.
"""
import time
from enum import Enum

import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F

from . import ops


# ---------------------------------------------------------------------------- enums
class BackboneType(Enum):          # builder.py:30-38
    DEEPLABV3 = 0
    UNET_ENCODER_ONLY = 1
    UNET_TRUNCATED = 2


class MappingType(Enum):           # builder.py:41-49
    CP2 = 0
    PIXEL_ID = 1
    REGION_ID = 2
    PIXEL_REGION_ID = 3


class NegativeType(Enum):          # builder.py:140-147
    NONE = 0
    FIXED = 1
    AVERAGE = 2
    MEDIAN = 3
    HARD = 4


class PretrainType(Enum):          # networks/segment_network.py:14-38
    RANDOM = 0
    NONE = 1
    CP2 = 2
    MIRROR = 3
    BYOL = 4
    MOCO = 5
    PROPOSED = 6
    PIXPRO = 7
    DENSECL_IMGNET = 8
    DINO_IMGNET = 9
    BARLOWTWINS_IMGNET = 10
    VICEREGL_IMGNET = 11
    MOCO_IMGNET = 12
    PIXPRO_IMGNET = 13
    BYOL_IMGNET = 14
    CP2_IMGNET = 15
    MOSREP_IMGNET = 16
    CLOVE_IMGNET = 17
    DENSECL = 18
    PROPOSED_V2 = 19


# --------------------------------------------------------------------------- meters
class AverageMeter(object):
    """builder.py:51-73, with one change: `update` accepts 0-d device tensors and keeps the
    running sum on the device, so a training step never waits for the GPU.  `val`, `avg`,
    `sum` read back (and synchronise) only when somebody looks at them."""

    def __init__(self, name, fmt=":f"):
        self.name = name
        self.fmt = fmt
        self.reset()

    def reset(self):
        self._val = 0
        self._sum = 0
        self.count = 0

    def update(self, val, n=1):
        if torch.is_tensor(val):
            val = val.detach()
        self._val = val
        self._sum = self._sum + val * n
        self.count += n

    @staticmethod
    def _f(x):
        return float(x) if torch.is_tensor(x) else x

    @property
    def val(self):
        return self._f(self._val)

    @property
    def sum(self):
        return self._f(self._sum)

    @property
    def avg(self):
        return self._f(self._sum) / self.count if self.count else 0

    def __str__(self):
        fmtstr = "{name} {val" + self.fmt + "} ({avg" + self.fmt + "})"
        return fmtstr.format(name=self.name, val=self.val, avg=self.avg)


# ---------------------------------------------------------------------- collectives
def _world():
    return dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1


def _rank():
    return dist.get_rank() if dist.is_available() and dist.is_initialized() else 0


@torch.no_grad()
def concat_all_gather(tensor):
    """builder.py:1710-1722: rank-ordered concatenation along dim 0, no gradient.  One
    `all_gather_into_tensor` (ncclAllGather over NVLink on the B200 box, gloo on CPU tests)
    straight into the contiguous output: no per-rank temporaries, no `cat`."""
    world = _world()
    if world == 1:
        return tensor
    tensor = tensor.contiguous()
    out = torch.empty((world * tensor.shape[0],) + tuple(tensor.shape[1:]), dtype=tensor.dtype,
                      device=tensor.device)
    dist.all_gather_into_tensor(out, tensor)
    return out


def _default_encoder(cfg, dim):
    """`build_segmentor(cfg.model)` (builder.py:365-371) when mmseg is importable, else the
    torchvision ResNet-50 (OS16) + ASPP + contrast_conv stand-in of cp2_b200.encoder."""
    try:
        from mmseg.models import build_segmentor  # type: ignore
        return build_segmentor(cfg.model, train_cfg=cfg.get("train_cfg"), test_cfg=cfg.get("test_cfg"))
    except ImportError:
        from .encoder import DeepLabV3Encoder
        return DeepLabV3Encoder(dim=dim)


# ---------------------------------------------------------------------------- MODEL
class DenseCLNeck(nn.Module):
    """builder.py:179-276: the DenseCL neck (global fc-relu-fc, dense 1x1 conv-relu-conv, predictors on top),
    same parameter names and initialisation.  Plain torch layers: cuDNN / cuBLAS, outside the loss path."""

    def __init__(self, in_channels, hid_channels, out_channels, num_grid=None):
        super().__init__()
        self.avgpool_global = nn.AdaptiveAvgPool2d((1, 1))
        self.global_projector = nn.Sequential(nn.Linear(in_channels, hid_channels), nn.ReLU(inplace=True),
                                              nn.Linear(hid_channels, out_channels))
        self.global_predictor = nn.Sequential(nn.Linear(out_channels, hid_channels), nn.ReLU(inplace=True),
                                              nn.Linear(hid_channels, out_channels))
        self.with_pool = num_grid is not None
        if self.with_pool:
            self.pool = nn.AdaptiveAvgPool2d((num_grid, num_grid))
        self.local_projector = nn.Sequential(nn.Conv2d(in_channels, hid_channels, 1), nn.ReLU(inplace=True),
                                             nn.Conv2d(hid_channels, out_channels, 1))
        self.local_predictor = nn.Sequential(nn.Conv2d(out_channels, hid_channels, 1), nn.ReLU(inplace=True),
                                             nn.Conv2d(hid_channels, out_channels, 1))
        self.avgpool_local = nn.AdaptiveAvgPool2d((1, 1))
        for m in self.modules():                                # builder.py:217-238
            if isinstance(m, (nn.Conv2d, nn.Linear)):
                nn.init.xavier_normal_(m.weight)
                if m.bias is not None:
                    nn.init.constant_(m.bias, 0)

    def forward(self, x):                                       # builder.py:240-275
        x_global_proj = self.global_projector(self.avgpool_global(x).flatten(1))
        x_global_pred = self.global_predictor(x_global_proj)
        if self.with_pool:
            x = self.pool(x)
        x_local_proj = self.local_projector(x)
        x_local_pred = self.local_predictor(x_local_proj)
        return {"x_global_proj": x_global_proj, "x_local_proj": x_local_proj, "x_global_pred": x_global_pred,
                "x_local_pred": x_local_pred, "x_avgpool_local_pred": self.avgpool_local(x_local_pred).flatten(1),
                "x_avgpool_local_proj": self.avgpool_local(x_local_proj).flatten(1)}


class MODEL(nn.Module):
    """Drop-in for the reference's `builder.MODEL` (builder.py:277-1706), CP2 mode.

    Supported range (validated in __init__): dim in {128, 256}, K % 4 == 0, both temperatures >= 1/16.

    Extra keyword-only arguments (not in the reference):
      encoder_factory : callable returning one segmentor; replaces `build_segmentor(cfg.model)`
                        (mmseg/mmcv are not installable here).  Called twice (q, k).
      track_accuracy  : evaluate the loss-neutral `acc_ins` / `acc_seg` meters (builder.py:1441-1448) each
                        step, as the reference does.  They are the top-1 over the 65537 global logits —
                        taken inside the loss's own queue pass (cp2_global_nce: candidates within the bf16
                        error bound of l_pos are re-scored exactly in fp32) — and the arg-max over all P x P
                        dense logits; neither tensor is materialised and nothing syncs
                        the host.  On by default; False skips them and the meters read NaN.
      output_stride   : skip the reference's CPU probe forward (builder.py:392-397) and use
                        this value.  When None the probe runs exactly as in the reference,
                        including its side effect on encoder_q's BN running statistics.
      backbone_output_stride : stride of `encoder.backbone(x)[3]` (builder.py:398-401), used by
                        forward_densecl for the id lattice.  None: probed as in the reference when the
                        encoder has a `.backbone` (also when `output_stride` is given), else = output_stride.
      logger          : callable(dict) receiving what the reference hands to `wandb.log` on rank 0
                        (builder.py:1450-1604, 1608-1664).  None: `wandb.log` if a wandb run is active on
                        rank 0, otherwise nothing is logged and none of the log-only statistics is computed.
      overlap_exchange: world > 1 only — issue the BN-shuffle image exchange (builder.py:609-630) on a side
                        stream right after the compositing, so that it runs under `encoder_q` instead of
                        between the two encoders.  The permutation, RNG draws and results are unchanged.
    """

    def __init__(self, cfg, rank, dim=128, K=65536, m=0.999, instance_logits_temp=0.2,
                 pretrain_from_scratch=False, include_background=False, lmbd_cp2_dense_loss=0.2,
                 lmbd_pixel_corr_weight=1, lmbd_region_corr_weight=1, lmbd_not_corr_weight=1,
                 negative_type=NegativeType.NONE, negative_scale=2, pretrain_type=PretrainType.CP2,
                 backbone_type=BackboneType.DEEPLABV3, mapping_type=MappingType.CP2,
                 dense_logits_temp=1, unet_truncated_dec_blocks=2, use_predictor=False,
                 use_avgpool_global=False, use_symmetrical_loss=False, lmbd_coordinate=0,
                 device=None, *, encoder_factory=None, output_stride=None, backbone_output_stride=None,
                 track_accuracy=True, logger=None, overlap_exchange=True):
        super().__init__()
        # limits of the kernels, checked here instead of inside the first forward (include/cp2_b200.h)
        if dim not in ops.SUPPORTED_DIMS:
            raise ValueError(f"cp2_b200: dim={dim} unsupported; the tensor-core kernels take {ops.SUPPORTED_DIMS}")
        if K <= 0 or K % 4 != 0:
            raise ValueError(f"cp2_b200: K={K} must be a positive multiple of 4 (16-byte queue rows for TMA)")
        for name, t in (("instance_logits_temp", instance_logits_temp), ("dense_logits_temp", dense_logits_temp)):
            if not (t >= 1.0 / ops.MAX_INV_TAU):
                raise ValueError(f"cp2_b200: {name}={t} must be >= {1.0 / ops.MAX_INV_TAU} (the kernels use a "
                                 f"fixed soft-max shift, valid for 1/tau <= {ops.MAX_INV_TAU:g})")
        self.queue_len = K
        self.momentum = m
        self.temp_global = instance_logits_temp
        self.dim = dim
        self.include_background = include_background
        self.lmbd_dense_loss = lmbd_cp2_dense_loss
        self.device = device
        self.rank = rank
        self.epoch = 0
        self.temp_local = dense_logits_temp
        self.use_predictor = use_predictor
        self.use_avgpool_global = use_avgpool_global
        self.use_symmetrical_loss = use_symmetrical_loss
        assert 0 <= lmbd_coordinate <= 1, f"{lmbd_coordinate = }"
        self.lmbd_coordinate = lmbd_coordinate

        # builder.py:325-347 — the same validation, by enum NAME so that the reference's
        # own enum members (main.py:136-140) are accepted as well
        mapping_type = MappingType[getattr(mapping_type, "name", mapping_type)]
        pretrain_type = PretrainType[getattr(pretrain_type, "name", pretrain_type)]
        negative_type = NegativeType[getattr(negative_type, "name", negative_type)]
        backbone_type = BackboneType[getattr(backbone_type, "name", backbone_type)]
        if mapping_type == MappingType.CP2:
            assert lmbd_pixel_corr_weight == 1
            assert lmbd_region_corr_weight == 1
            assert lmbd_not_corr_weight == 1
        elif mapping_type == MappingType.PIXEL_ID:
            assert lmbd_region_corr_weight == 1
            assert lmbd_pixel_corr_weight > 1
        elif mapping_type == MappingType.REGION_ID:
            assert lmbd_pixel_corr_weight == 1
            assert lmbd_region_corr_weight > 1
        self.mapping_type = mapping_type
        self.lmbd_pixel_corr_weight = lmbd_pixel_corr_weight
        self.lmbd_region_corr_weight = lmbd_region_corr_weight
        self.lmbd_not_corr_weight = lmbd_not_corr_weight
        self.pretrain_type = pretrain_type
        self.negative_type = negative_type
        self.negative_scale = negative_scale
        self.backbone_type = backbone_type
        self._densecl = pretrain_type in (PretrainType.DENSECL, PretrainType.PROPOSED_V2)
        if pretrain_type not in (PretrainType.CP2, PretrainType.PROPOSED) and not self._densecl:
            raise NotImplementedError(
                f"cp2_b200 implements forward_cp2 (PretrainType.CP2 / PROPOSED) and forward_densecl "
                f"(DENSECL / PROPOSED_V2) only ({pretrain_type = })")
        if self._densecl:                                       # builder.py:443-460
            assert self.momentum == 0.999, f"{self.momentum = }"
            assert self.lmbd_dense_loss == 0.5, f"{self.lmbd_dense_loss = }"
            assert self.temp_global == 0.2, f"{self.temp_global = }"
            assert self.temp_local == 0.2, f"{self.temp_local = }"
        if pretrain_type == PretrainType.DENSECL:
            assert not self.use_predictor and not self.use_avgpool_global
            assert not self.use_symmetrical_loss and self.lmbd_coordinate == 0
        if pretrain_type == PretrainType.CP2:
            assert negative_type == NegativeType.NONE   # builder.py:431-433
            assert mapping_type == MappingType.CP2
        # the PROPOSED modes that change the loss; all-ones weights with no
        # negative squash are the CP2 loss and take the tensor-core path
        self._weighted = not (lmbd_pixel_corr_weight == 1 and lmbd_region_corr_weight == 1
                              and lmbd_not_corr_weight == 1
                              and negative_type in (NegativeType.NONE, NegativeType.HARD))
        if backbone_type != BackboneType.DEEPLABV3 and encoder_factory is None:
            raise NotImplementedError(f"{backbone_type = } needs an encoder_factory (smp is not installed)")

        make = encoder_factory if encoder_factory is not None else (lambda: _default_encoder(cfg, dim))
        self.encoder_q = make()
        self.encoder_k = make()
        if encoder_factory is None and not pretrain_from_scratch:
            for enc in (self.encoder_q, self.encoder_k):       # builder.py:375-378
                if hasattr(enc, "backbone") and hasattr(enc.backbone, "init_weights"):
                    enc.backbone.init_weights()

        if output_stride is None:                              # builder.py:392-402
            test_sample = torch.rand(2, 3, 224, 224)
            self.output_stride = int(test_sample.shape[2] / self.encoder_q(test_sample).shape[2])
        else:
            self.output_stride = int(output_stride)
        if backbone_output_stride is not None:
            self.backbone_output_stride = int(backbone_output_stride)
        elif hasattr(self.encoder_q, "backbone") and (output_stride is None or self._densecl):
            # builder.py:398-401.  With `output_stride` given the probe is only run for the modes that use the
            # value (forward_densecl), on the backbone alone and without touching BN statistics.
            test_sample = torch.rand(2, 3, 224, 224)
            if output_stride is None:
                shp = self.encoder_q.backbone(test_sample)[3].shape
            else:
                was_training = self.encoder_q.training
                self.encoder_q.eval()
                with torch.no_grad():
                    shp = self.encoder_q.backbone(test_sample)[3].shape
                self.encoder_q.train(was_training)
            self.backbone_output_stride = int(test_sample.shape[2] / shp[2])
        else:
            self.backbone_output_stride = self.output_stride

        if self._densecl:                                       # builder.py:436-441, 452-457
            if not hasattr(self.encoder_q, "backbone"):
                raise NotImplementedError("the DenseCL modes need an encoder with a `.backbone` (4 stage outputs)")
            self.encoder_q.neck = DenseCLNeck(in_channels=2048, hid_channels=2048, out_channels=self.dim)
            self.encoder_k.neck = DenseCLNeck(in_channels=2048, hid_channels=2048, out_channels=self.dim)

        # builder.py:464-469: parameters only — BN buffers are not copied
        for param_q, param_k in zip(self.encoder_q.parameters(), self.encoder_k.parameters()):
            param_k.data.copy_(param_q.data)
            param_k.requires_grad = False

        # builder.py:476-482 (same RNG draws, same order)
        self.register_buffer("queue", torch.randn(self.dim, self.queue_len))
        self.queue = F.normalize(self.queue, dim=0)
        self.register_buffer("queue_ptr", torch.zeros(1, dtype=torch.long))
        self.register_buffer("queue2", torch.randn(self.dim, self.queue_len))
        self.queue2 = F.normalize(self.queue2, dim=0)
        self.register_buffer("queue2_ptr", torch.zeros(1, dtype=torch.long))

        self.loss_o = AverageMeter("Loss_overall", ":.4f")      # builder.py:543-553
        self.loss_i = AverageMeter("Loss_ins", ":.4f")
        self.loss_d = AverageMeter("Loss_den", ":.4f")
        self.acc_ins = AverageMeter("Acc_ins", ":6.2f")
        self.acc_seg = AverageMeter("Acc_seg", ":6.2f")
        self.cross_image_variance_source = AverageMeter("Cross_Image_Variance_Source", ":6.2f")
        self.cross_image_variance_target = AverageMeter("Cross_Image_Variance_Target", ":6.2f")
        self._iou_host, self._iou_dev = [[], []], [[], []]      # (all pixels, masked): lists + pending device chunks
        self._iou_pending = [[], []]                            # (pinned host tensor, copy-done event) pairs
        self._ema_plan = None
        self._shadow, self._shadow2 = ops.QueueShadow(), ops.QueueShadow()   # bf16 operands of cp2_global_nce
        self.track_accuracy = bool(track_accuracy)
        self._logger = logger
        self.overlap_exchange = bool(overlap_exchange)
        self._side_stream = None
        self._cpu_group = None
        self._key_graph = None                 # (CUDAGraph, static_in, static_out) of graph_encoders()
        self.graphed = {"encoder_q": False, "encoder_k": False}
        self.collective_times = None        # bench.py: dict name -> [seconds | (event, event)] when not None

    @torch.no_grad()
    def _dense_accuracy(self, q, k, mask_a, mask_b, stats):
        """builder.py:1442-1448: arg-max over all P x P dense logits (no [B,P,P] tensor, no host sync).  Maps that
        fit TMEM: found by cp2_dense_small while S is on chip (stats["acc_seg"]); larger maps: cp2_dense_top1.
        In the PROPOSED modes the arg-max runs on the weighted logits inside cp2_dense_weighted_lse.  acc_ins (builder.py:1441) comes out of the loss's own queue pass (cp2_global_nce)."""
        if stats.get("acc_seg") is not None:                   # weighted modes / maps that fit TMEM: already there
            return stats["acc_seg"]
        return ops.dense_top1(q, k, mask_a, mask_b)[0]

    # ----------------------------------------------------------- a4 (metrics)
    # builder.py:1252-1257 extends two python lists with per-image IoUs every step, through a .cpu() sync.
    # Here the per-step [B] results stay on the device; every _IOU_FLUSH steps they are copied to pinned host
    # memory asynchronously (no sync, bounded device memory) and the lists are materialised when somebody looks
    # at them (the reference does so once per epoch, builder.py:1455-1486).
    _IOU_FLUSH = 64

    def _iou_flush_async(self, which):
        chunks, self._iou_dev[which] = self._iou_dev[which], []
        if not chunks:
            return
        dev_t = torch.cat(chunks)
        host_t = torch.empty(dev_t.shape, dtype=dev_t.dtype, pin_memory=True)
        host_t.copy_(dev_t, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        self._iou_pending[which].append((host_t, ev, dev_t))     # dev_t kept alive until the copy is done

    def _iou_list(self, which):
        self._iou_flush_async(which)
        for host_t, ev, _dev_t in self._iou_pending[which]:
            ev.synchronize()
            self._iou_host[which].extend(host_t.numpy().tolist())
        self._iou_pending[which] = []
        return self._iou_host[which]

    def _iou_set(self, which, v):
        self._iou_dev[which], self._iou_pending[which], self._iou_host[which] = [], [], list(v)

    correlation_ious = property(lambda self: self._iou_list(0), lambda self, v: self._iou_set(0, v))
    masked_correlation_ious = property(lambda self: self._iou_list(1), lambda self, v: self._iou_set(1, v))

    # ------------------------------------------------------------------ K4
    @torch.no_grad()
    def _momentum_update_key_encoder(self):
        """builder.py:557-567 as ONE kernel launch over every parameter (bit-exact)."""
        pq = [p.data for p in self.encoder_q.parameters()]
        pk = [p.data for p in self.encoder_k.parameters()]
        if self._ema_plan is None:
            self._ema_plan = ops.EmaPlan(pq, pk)
        else:
            self._ema_plan.rebind(pq, pk)
        self._ema_plan.apply(self.momentum)

    # ------------------------------------------------------------------ K5
    @torch.no_grad()
    def _dequeue_and_enqueue(self, keys):
        """builder.py:569-587: all-gather, then transposed write at the device-side pointer
        with wrap-around.  No host sync."""
        ops.enqueue(self._timed("nccl_all_gather_keys", lambda: concat_all_gather(keys)), self.queue, self.queue_ptr,
                    self._shadow)

    @torch.no_grad()
    def _dequeue_and_enqueue2(self, keys):
        """builder.py:589-607."""
        ops.enqueue(concat_all_gather(keys), self.queue2, self.queue2_ptr, self._shadow2)

    # ----------------------------------------------------------------- a11 / f1
    @staticmethod
    def _exchange_rows(x, perm_host):
        """out_global[i] = in_global[perm[i]] for a batch sharded in rank order, moving ONLY the rows each
        rank needs (one all_to_all_single = grouped ncclSend/ncclRecv over NVLink) instead of the
        reference's all-gather of the whole global batch (builder.py:612,635: W x the data).
        `perm_host` is the full permutation on the HOST, identical on every rank."""
        W, me, B = _world(), _rank(), x.shape[0]
        need = perm_host.view(W, B)                              # need[r] = global rows rank r must end up with
        owner = torch.div(need, B, rounding_mode="floor")
        # what I send to rank r: my local rows among need[r], in the order they appear in need[r]
        send_rows = [need[r][owner[r] == me] - me * B for r in range(W)]
        in_splits = [int(t.numel()) for t in send_rows]
        # what I receive: grouped by source rank, each group in need[me] order (stable sort by owner)
        order = torch.sort(owner[me], stable=True).indices       # recv[i] belongs at position order[i]
        out_splits = torch.bincount(owner[me], minlength=W).tolist()
        send = x.index_select(0, torch.cat(send_rows).to(x.device, non_blocking=True)).contiguous()
        recv = torch.empty_like(x)
        dist.all_to_all_single(recv, send, output_split_sizes=out_splits, input_split_sizes=in_splits)
        inv = torch.empty_like(order)
        inv[order] = torch.arange(B)
        return recv.index_select(0, inv.to(x.device, non_blocking=True))

    def _timed(self, name, fn):
        """Runs one collective; with `collective_times` set (bench.py) brackets it with CUDA events on the
        issuing stream (device-side duration incl. the NCCL stream hand-over)."""
        if self.collective_times is None:
            return fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = fn()
        e1.record()
        self.collective_times.setdefault(name, []).append((e0, e1))
        return out

    def _broadcast_perm_host(self, idx_host):
        """Rank 0's CPU permutation to every rank WITHOUT a device->host read: the 8*B*W bytes travel over a
        gloo side group on the host (the device-side copy is broadcast separately by NCCL).  Falls back to the
        default group when that already is a CPU backend."""
        if self._cpu_group is None:
            backend = dist.get_backend()
            self._cpu_group = dist.group.WORLD if backend == "gloo" else dist.new_group(backend="gloo")
        t0 = time.perf_counter()
        dist.broadcast(idx_host, src=0, group=self._cpu_group)
        if self.collective_times is not None:
            self.collective_times.setdefault("gloo_broadcast_perm_host_s", []).append(time.perf_counter() - t0)
        return idx_host

    @torch.no_grad()
    def _batch_shuffle_ddp(self, x):
        """builder.py:609-630.  Same RNG contract: every rank draws a CPU `randperm`, rank 0's wins.
        Returns (x_shuffled_for_this_rank, idx_unshuffle) like the reference; the rows travel by all-to-all
        (SURVEY.md section 8 row f1).  No rank reads the permutation back from the device: rank 0's draw is
        broadcast on the host (gloo) and uploaded by every rank."""
        batch_size_this = x.shape[0]
        W = _world()
        batch_size_all = batch_size_this * W
        idx_host = torch.randperm(batch_size_all)                # CPU generator, advanced on every rank
        if W > 1:
            idx_host = self._broadcast_perm_host(idx_host)
        idx_shuffle = idx_host.to(x.device, non_blocking=True)
        unshuffle_host = torch.argsort(idx_host)
        idx_unshuffle = unshuffle_host.to(x.device, non_blocking=True)
        self._unshuffle_host = (idx_unshuffle, unshuffle_host)
        if W == 1:
            return x[idx_shuffle], idx_unshuffle
        return self._timed("nccl_all_to_all_shuffle", lambda: self._exchange_rows(x, idx_host)), idx_unshuffle

    @torch.no_grad()
    def _batch_unshuffle_ddp(self, x, idx_unshuffle):
        """builder.py:632-649."""
        if _world() == 1:
            return x[idx_unshuffle]
        cached = getattr(self, "_unshuffle_host", None)
        if cached is not None and cached[0] is idx_unshuffle:
            perm_host = cached[1]                                # no second device->host read
        else:
            perm_host = idx_unshuffle.cpu()
        return self._timed("nccl_all_to_all_unshuffle", lambda: self._exchange_rows(x, perm_host))

    def fold_dilated_convs(self, max_cover=0.15):
        """Exact rewrite of the encoders' strongly dilated 3x3 convolutions (the ASPP branches) for small feature
        maps: every tap is applied as a 1x1 convolution on the window where it reads data instead of padding
        (`cp2_b200.encoder.fold_dilated_convs`; at the 14x14 map of 224x224 views the dilation-18 branch is a 1x1
        convolution; the default `max_cover` keeps the dilation-12 branch, which touches 2-pixel borders, on cuDNN's
        kernel because that measured faster).  Same parameters, same sums; maps that are too
        large for it to pay run the stock convolution.  Call before `graph_encoders`.  -> convolutions patched."""
        from .encoder import fold_dilated_convs
        return fold_dilated_convs(self.encoder_q, max_cover) + fold_dilated_convs(self.encoder_k, max_cover)

    def graph_encoders(self, image_shape, *, query=True, key=True, warmup=3, query_stages=None):
        """Capture the encoders for ONE input shape in CUDA graphs (no counterpart in the reference, which launches
        every convolution / BN / ReLU kernel from Python: ~1100 launches per step, the host-bound part of the step).

        * `encoder_q`: forward and backward, through `torch.cuda.make_graphed_callables` (its forward becomes one
          graph launch, its backward another; parameters keep their storage, so the EMA kernel, the optimizer and
          DDP see the same tensors).  Call BEFORE wrapping the model in DistributedDataParallel.  The whole backward
          is then ONE autograd node: DDP's gradient hooks fire at its end, so the all-reduce no longer overlaps
          the backward.  `query_stages` (a list of sub-modules that `encoder_q.forward` calls one after the other,
          each on the previous one's output tensor) captures every stage on its own instead: autograd sees one node
          per stage and DDP's buckets start reducing while the earlier stages are still in their backward.
        * `encoder_k`: the no-grad forward of `forward_cp2` in one `torch.cuda.CUDAGraph` (static input / output).
        Only used while the module is in training mode and the input has `image_shape`; any other call runs the
        eager path.  The graphs hold the parameters' and buffers' addresses: load_state_dict / optimizer steps /
        the EMA kernel (all in place) are fine, `.to(...)`, `.half()` or re-creating parameters afterwards are not.  The warm-up forwards run on zeros: BatchNorm running statistics of both encoders are
        restored afterwards, so capturing does not change the model."""
        dev = self.queue.device
        assert dev.type == "cuda", "graph_encoders needs the model on a CUDA device"
        assert self.training, "capture in training mode (the graphs replay train-mode BatchNorm)"
        image_shape = tuple(int(v) for v in image_shape)

        def with_buffers_restored(module, fn):
            saved = [b.detach().clone() for b in module.buffers()]
            try:
                return fn()
            finally:
                with torch.no_grad():
                    for b, v in zip(module.buffers(), saved):
                        b.copy_(v)

        if query and not self.graphed["encoder_q"]:
            sample = torch.zeros(image_shape, device=dev, dtype=torch.float32)
            if query_stages:
                stages, samples, x = tuple(query_stages), [], sample

                def capture_stages():
                    nonlocal x
                    with torch.no_grad():                        # shapes of the tensors handed from stage to stage
                        for i, st in enumerate(stages):
                            samples.append((torch.zeros_like(x).requires_grad_(i > 0),))
                            x = st(x)
                    torch.cuda.make_graphed_callables(stages, tuple(samples), num_warmup_iters=warmup,
                                                      allow_unused_input=True)
                with_buffers_restored(self.encoder_q, capture_stages)
                self.graphed["encoder_q"] = f"{len(stages)} stages"
            else:
                with_buffers_restored(self.encoder_q, lambda: torch.cuda.make_graphed_callables(
                    self.encoder_q, (sample,), num_warmup_iters=warmup, allow_unused_input=True))
                self.graphed["encoder_q"] = True
        if key and self._key_graph is None:
            static_in = torch.zeros(image_shape, device=dev, dtype=torch.float32)

            def capture():
                side = torch.cuda.Stream(device=dev)
                side.wait_stream(torch.cuda.current_stream(dev))
                with torch.cuda.stream(side), torch.no_grad():
                    for _ in range(max(1, warmup)):             # cuDNN algorithm search happens outside the capture
                        self.encoder_k(static_in)
                torch.cuda.current_stream(dev).wait_stream(side)
                graph = torch.cuda.CUDAGraph()
                with torch.no_grad(), torch.cuda.graph(graph):
                    static_out = self.encoder_k(static_in)
                return graph, static_out
            graph, static_out = with_buffers_restored(self.encoder_k, capture)
            self._key_graph = (graph, static_in, static_out)
            self.graphed["encoder_k"] = True
        return self

    def _encode_keys(self, img_b):
        """builder.py:1275 `k = self.encoder_k(img_b)` (inside no_grad)."""
        kg = self._key_graph
        if kg is None or not self.training or tuple(img_b.shape) != tuple(kg[1].shape) or not img_b.is_cuda:
            return self.encoder_k(img_b)
        graph, static_in, static_out = kg
        static_in.copy_(img_b)
        graph.replay()
        return static_out                          # read by this step's loss before the next replay overwrites it

    def forward(self, **kwargs):                               # builder.py:651-665
        if self._densecl:
            return self.forward_densecl(**kwargs)
        return self.forward_cp2(**kwargs)

    # ------------------------------------------------------- forward_densecl (f2)
    def _query_features(self, img):
        """builder.py:700-720."""
        embd = self.encoder_q.backbone(img)[3]
        out = self.encoder_q.neck(embd)
        tag = "pred" if self.use_predictor else "proj"
        q_local, q_global = out[f"x_local_{tag}"], out[f"x_global_{tag}"]
        if self.use_avgpool_global:
            q_global = out[f"x_avgpool_local_{tag}"]
        return (F.normalize(embd.detach().flatten(2), dim=1), F.normalize(q_local.flatten(2), dim=1),
                F.normalize(q_global, dim=1))

    @torch.no_grad()
    def _key_features(self, img):
        """builder.py:722-760: EMA, shuffle, key encoder, normalise, unshuffle."""
        self._momentum_update_key_encoder()
        img, idx_unshuffle = self._batch_shuffle_ddp(img)
        embd = self.encoder_k.backbone(img)[3]
        out = self.encoder_k.neck(embd)
        pooled = out["x_avgpool_local_proj"]
        k_global = pooled if self.use_avgpool_global else out["x_global_proj"]
        feats = [F.normalize(embd.flatten(2), dim=1), F.normalize(out["x_local_proj"].flatten(2), dim=1),
                 F.normalize(k_global, dim=1), F.normalize(pooled, dim=1)]
        return [self._batch_unshuffle_ddp(t, idx_unshuffle) for t in feats]

    def forward_densecl(self, img_a, img_b, bg0=None, bg1=None, visualize=False, step=0, new_epoch=False,
                        pixel_ids_a=None, pixel_ids_b=None, region_ids_a=None, region_ids_b=None):
        """builder.py:667-999 (DENSECL / PROPOSED_V2).  The images are used as they come (no compositing in this
        mode); the losses run through `ops.densecl_loss`: no [N,1+K] logits, no [B,S^2,S^2] local similarity
        matrix, no gather.  The reference's per-step wandb statistics (quartiles of the negatives, matching
        rate) are logging only and are not computed; its IndexError at builder.py:861 on views that share a
        pixel id is not reproduced."""
        n = len(img_a)
        bos = self.backbone_output_stride

        def one_direction(x_q, x_k, ids_q, ids_k):
            embd_q, q_local, q_global = self._query_features(x_q)
            embd_k, k_local, k_global, k_pooled = self._key_features(x_k)
            mix = self.lmbd_coordinate != 0 and ids_q is not None and ids_k is not None
            loss, loss_g, loss_l = ops.densecl_loss(
                embd_q, embd_k, q_local, k_local, q_global, k_global, self.queue, self.queue2, self.temp_global,
                self.temp_local, self.lmbd_dense_loss, ids_q if mix else None, ids_k if mix else None, bos,
                self.lmbd_coordinate, shadow=self._shadow, shadow2=self._shadow2)
            return loss_g, loss_l, k_global, k_pooled

        loss_global, loss_local, k_glob, k_loc = one_direction(img_a, img_b, pixel_ids_a, pixel_ids_b)
        if self.use_symmetrical_loss:                          # builder.py:942-968
            lg2, ll2, k_glob2, k_loc2 = one_direction(img_b, img_a, pixel_ids_b, pixel_ids_a)
            loss_global, loss_local = loss_global + lg2, loss_local + ll2
            if step % 2 == 0:
                k_glob, k_loc = k_glob2, k_loc2
        loss = (1 - self.lmbd_dense_loss) * loss_global + self.lmbd_dense_loss * loss_local
        self._dequeue_and_enqueue(k_glob)                      # builder.py:975-976
        self._dequeue_and_enqueue2(k_loc)
        with torch.no_grad():
            self.loss_o.update(loss, n)
            self.loss_i.update(loss_global, n)
            self.loss_d.update(loss_local, n)
        return loss

    # ---------------------------------------------------------- forward_cp2
    def _shuffled_keys_input(self, img_b):
        """builder.py:1274 `_batch_shuffle_ddp(img_b)`.  World > 1 with `overlap_exchange`: the exchange depends
        only on the compositing kernel, so it is issued on a side stream and joined just before `encoder_k`
        (it then runs under `encoder_q`).  -> (img_b_shuffled, idx_unshuffle, join) with join() to be called
        on the compute stream before the result is used."""
        if _world() == 1 or not self.overlap_exchange or not img_b.is_cuda:
            out, idx = self._batch_shuffle_ddp(img_b)
            return out, idx, (lambda: None)
        cur = torch.cuda.current_stream()
        if self._side_stream is None:
            self._side_stream = torch.cuda.Stream(device=img_b.device)
        side = self._side_stream
        side.wait_stream(cur)                                  # the composited img_b is ready
        with torch.cuda.stream(side):
            out, idx = self._batch_shuffle_ddp(img_b)
        img_b.record_stream(side)

        def join():
            cur.wait_stream(side)
            out.record_stream(cur)
            idx.record_stream(cur)
        return out, idx, join

    def forward_cp2(self, img_a, img_b, bg0, bg1, visualize=False, step=0, new_epoch=False,
                    pixel_ids_a=None, pixel_ids_b=None, region_ids_a=None, region_ids_b=None):
        """builder.py:1124-1606.

        MappingType.CP2 / NegativeType.NONE: the pixel / region id maps only feed loss-neutral bookkeeping
        (corr_weights is all ones, builder.py:329-332, 1238-1243): the region-id IoUs the reference records
        per step (builder.py:1252-1257) come from `cp2_id_iou`; the pixel-id maps are not needed.
        PretrainType.PROPOSED with other weights or a NegativeType: the four id maps define the pair
        weights and are read by the dense-loss kernels themselves (`cp2_dense_weighted_*`).
        """
        os_ = self.output_stride
        keep = bool(visualize) and self.rank == 0
        # K1: builder.py:1146-1159 (+ the mask part of the [OS/2::OS] subsample), both views in one launch
        (img_a, mask_a, full_a), (img_b, mask_b, full_b) = ops.composite_pair(img_a, bg0, img_b, bg1, os_,
                                                                            want_full_mask=keep)
        if keep:
            self.last_visualization = {"img_a": img_a, "img_b": img_b, "mask_a": full_a, "mask_b": full_b}
        n = img_a.shape[0]
        hidden_image_size = tuple(mask_a.shape[1:])
        with torch.no_grad():                                  # builder.py:1271-1277, exchange started early
            img_b, idx_unshuffle, join_exchange = self._shuffled_keys_input(img_b)
        if region_ids_a is not None and region_ids_b is not None:
            # a2 + a4: builder.py:1211-1217, 1252-1257 (the region-id IoUs are the ones the reference keeps)
            iou, iou_m = ops.id_iou(region_ids_a, region_ids_b, mask_a, mask_b, os_)
            self._iou_dev[0].append(iou)
            self._iou_dev[1].append(iou_m)
            if len(self._iou_dev[0]) >= self._IOU_FLUSH:
                self._iou_flush_async(0)
                self._iou_flush_async(1)
        mask_a = mask_a.reshape(n, -1)
        mask_b = mask_b.reshape(n, -1)

        q = self.encoder_q(img_a)                              # builder.py:1260 (cuDNN, out of scope)
        with torch.no_grad():                                  # builder.py:1271-1277
            self._momentum_update_key_encoder()
            join_exchange()
            k = self._encode_keys(img_b)
            k = self._batch_unshuffle_ddp(k, idx_unshuffle)

        # K2a/K2/K3 + autograd: builder.py:1261-1268, 1279-1292, 1392-1397, 1414-1439
        pair_weights = None
        if self._weighted:                                     # builder.py:1204-1243, 1332-1392
            pair_weights = ops.PairWeights(pixel_ids_a, pixel_ids_b, region_ids_a, region_ids_b, os_,
                                           self.lmbd_pixel_corr_weight, self.lmbd_region_corr_weight,
                                           self.lmbd_not_corr_weight, self.negative_type, self.negative_scale)
        loss, loss_ins, loss_dense, k_pos, q_pos, stats = ops.cp2_loss(
            q, k, mask_a, mask_b, self.queue, self.temp_global, self.temp_local, self.lmbd_dense_loss,
            self.include_background, pair_weights, queue_shadow=self._shadow, track_accuracy=self.track_accuracy)
        if self.track_accuracy:                                # acc_ins came with the queue pass (before the enqueue)
            stats["acc_seg"] = self._dense_accuracy(q, k, mask_a, mask_b, stats)
        log = self._active_logger()
        score_stats = None
        if log is not None:                                    # log-only statistics: queue BEFORE the enqueue
            score_stats = self._step_score_stats(q, k, mask_a, mask_b, q_pos, k_pos, pair_weights)
        self._dequeue_and_enqueue(k_pos)                       # builder.py:1426

        if new_epoch:
            self._new_epoch_branch(log, q, k, mask_a, mask_b, hidden_image_size, pair_weights)
        with torch.no_grad():                                  # builder.py:1553-1563, no syncs
            self.loss_o.update(loss, n)
            self.loss_i.update(loss_ins, n)
            self.loss_d.update(loss_dense, n)
            self.acc_ins.update(stats["acc_ins"], n)
            self.acc_seg.update(stats["acc_seg"], n)
            self.cross_image_variance_source.update(stats["std_q_pos"], n)    # q_pos.std(0).mean()
            self.cross_image_variance_target.update(stats["std_k_pos"], n)
        if log is not None:
            self._log_step(log, score_stats)
        return loss

    # ------------------------------------------------ a12: logging / visualisation (off the hot path)
    def _active_logger(self):
        """The callable that receives the reference's rank-0 `wandb.log` dicts, or None (nothing is logged and
        none of the log-only statistics is computed)."""
        if self.rank != 0:
            return None
        if self._logger is not None:
            return self._logger
        try:
            import wandb
        except Exception:
            return None
        return wandb.log if getattr(wandb, "run", None) is not None else None

    @torch.no_grad()
    def dense_logits_for_visualisation(self, q, k, pair_weights=None, mask_a=None, mask_b=None):
        """The materialised `logits_dense` [B,P,P] of builder.py:1289-1292 (/ tau_d at :1431 is NOT applied, as at
        the point the reference slices it, :1493-1510), in plain torch.  Slow branch: only the new-epoch
        heat-maps and the log-only score statistics use it.  PROPOSED modes: times the pair weights after the
        NegativeType squash (builder.py:1332-1392) via the same torch formulation the tests use."""
        B, C = q.shape[0], q.shape[1]
        qh = F.normalize(q.detach().float().reshape(B, C, -1), dim=1)
        kh = F.normalize(k.detach().float().reshape(B, C, -1), dim=1)
        logits = torch.einsum("ncx,ncy->nxy", qh, kh)
        if pair_weights is None or mask_a is None:
            return logits
        # PROPOSED modes (slow, visualisation only): squash of the non-label pairs, then the pair weights
        labels = torch.einsum("nx,ny->nxy", mask_a, mask_b)
        w_pix, w_reg, w_not, code, scale = pair_weights.scalars
        if code != 0:
            neg = labels == 0
            if code == 1:                                      # FIXED
                centre = logits.new_zeros(B, 1, 1)
            else:
                x = torch.where(neg, logits, torch.full_like(logits, float("nan"))).reshape(B, -1)
                centre = (torch.nanmean(x, dim=1) if code == 2 else torch.nanquantile(x, 0.5, dim=1)).reshape(B, 1, 1)
            logits = torch.where(neg, 2.0 / (1.0 + torch.exp((logits - centre) * -scale)) - 1.0, logits)
        o, st = pair_weights.os // 2, pair_weights.os
        pa, pb, ra, rb = (t[:, o::st, o::st].reshape(B, -1) for t in pair_weights.ids)
        pix = pa[:, :, None] == pb[:, None, :]
        reg = (ra[:, :, None] == rb[:, None, :]) & ((ra[:, :, None] * rb[:, None, :]) != 0)
        w = w_reg * reg.to(logits.dtype)
        w = torch.where(pix, torch.full_like(w, w_pix), w)
        w = w + (w == 0) * w_not
        return logits * w

    @torch.no_grad()
    def _step_score_stats(self, q, k, mask_a, mask_b, q_pos, k_pos, pair_weights):
        """Log-only statistics of builder.py:1294-1303 (`calcuate_dense_loss_stats`,
        tools/correlation_mapping.py:11-53) and :1399-1412 (queue-logit quartiles), plain torch."""
        logits = self.dense_logits_for_visualisation(q, k)          # raw cosines: builder.py:1298 runs before the squash
        labels = torch.einsum("nx,ny->nxy", mask_a, mask_b)
        B = logits.shape[0]
        qs = torch.tensor([0.25, 0.5, 0.75], device=logits.device)
        out = {}
        for name, sel in (("positive", labels == 1), ("negative", labels == 0)):
            x = torch.where(sel, logits, torch.full_like(logits, float("nan"))).reshape(B, -1)
            out[name] = {"average": torch.nanmean(x, dim=1), "quartiles": torch.nanquantile(x, qs, dim=1)}
        l_neg = q_pos.detach() @ self.queue
        out["instance_pos"] = (q_pos.detach() * k_pos.detach()).sum(1)
        out["instance_neg_avg"] = l_neg.mean(dim=1)
        out["instance_neg_q"] = torch.quantile(l_neg, qs, dim=1)
        return out

    def _log_step(self, log, st):
        """builder.py:1565-1604 (rank 0).  Reads the meters' `.val` -> one device sync per logged step, as in
        the reference."""
        log({"train/loss_step": self.loss_o.val})
        log({"train/acc_ins_step": self.acc_ins.val})
        d = {"train/loss_ins_step": self.loss_i.val, "train/loss_dense_step": self.loss_d.val,
             "train/acc_seg_step": self.acc_seg.val,
             "train/cross_image_variance_source_step": self.cross_image_variance_source.val,
             "train/cross_image_variance_target_step": self.cross_image_variance_target.val}
        if st is not None:
            pos, neg = st["positive"], st["negative"]
            d.update({
                "train/+ive_scores_step": float(pos["average"].mean()),
                "train/-ive_scores_step": float(neg["average"].mean()),
                "step/dense_per_sample_average_positive_scores": float(pos["average"].mean()),
                "step/dense_per_sample_lower_positive_scores": float(pos["quartiles"][0].mean()),
                "step/dense_per_sample_median_positive_scores": float(pos["quartiles"][1].mean()),
                "step/dense_per_sample_upper_positive_scores": float(pos["quartiles"][2].mean()),
                "step/dense_per_sample_average_negative_scores": float(neg["average"].mean()),
                "step/dense_per_sample_lower_negative_scores": float(neg["quartiles"][0].mean()),
                "step/dense_per_sample_median_negative_scores": float(neg["quartiles"][1].mean()),
                "step/dense_per_sample_upper_negative_scores": float(neg["quartiles"][2].mean()),
                "step/instance_average_positive_scores": float(st["instance_pos"].mean()),
                "step/instance_average_negative_scores": float(st["instance_neg_avg"].mean()),
                "step/instance_lower_negative_scores": float(st["instance_neg_q"][0].mean()),
                "step/instance_median_negative_scores": float(st["instance_neg_q"][1].mean()),
                "step/instance_upper_negative_scores": float(st["instance_neg_q"][2].mean()),
            })
        log(d)

    @torch.no_grad()
    def heatmap_grid(self, logits_dense, mask_a, mask_b, hidden_image_size):
        """builder.py:1489-1532: per image the mean similarity of every key pixel to the foreground queries
        (hm_b) and of every query pixel to the foreground keys (hm_a), laid out as [m_a, hm_a, m_b, hm_b] rows of
        a 4-column grid, nearest-upscaled by 224 // h.  -> [H_grid, W_grid] float tensor (first channel)."""
        B = logits_dense.shape[0]
        a, b = mask_a.reshape(B, -1), mask_b.reshape(B, -1)
        hm_b = torch.einsum("nxy,nx->ny", logits_dense, a) / a.sum(1, keepdim=True)
        hm_a = torch.einsum("nxy,ny->nx", logits_dense, b) / b.sum(1, keepdim=True)
        h, w = hidden_image_size
        shape = (B, 1, h, w)
        m_a, m_b = 2 * a.reshape(shape) - 1, 2 * b.reshape(shape) - 1
        imgs = torch.stack([m_a, hm_a.reshape(shape), m_b, hm_b.reshape(shape)], dim=1).flatten(0, 1)
        # torchvision.utils.make_grid(imgs, nrow=4): 2-pixel zero padding between / around the cells
        nrow, pad = 4, 2
        ncell = imgs.shape[0]
        rows = (ncell + nrow - 1) // nrow
        grid = imgs.new_zeros(rows * (h + pad) + pad, nrow * (w + pad) + pad)
        for i in range(ncell):
            r, c = divmod(i, nrow)
            grid[pad + r * (h + pad): pad + r * (h + pad) + h, pad + c * (w + pad): pad + c * (w + pad) + w] = imgs[i, 0]
        scale = max(1, 224 // h)
        return grid.repeat_interleave(scale, dim=0).repeat_interleave(scale, dim=1)

    def _new_epoch_branch(self, log, q, k, mask_a, mask_b, hidden_image_size, pair_weights):
        """builder.py:1450-1550.  Rank 0 with a logger: IoU statistics + heat-maps, then the lists restart.
        Every rank: the IoU lists restart (the reference leaks them on ranks != 0)."""
        if log is not None:
            import numpy as np
            for which, tag in ((0, "feature_space"), (1, "feature_space_masked")):
                vals = list(self._iou_list(which))
                key = "iou" if which == 0 else "masked_iou"
                nz = int(np.count_nonzero(vals))
                log({f"feature_space_non_zero_{key}": nz})
                log({f"feature_space_non_zero_{key}_ratio": nz / len(vals) if vals else float("nan")})
                try:                                            # the histogram figure needs matplotlib + wandb.Image
                    import matplotlib.pyplot as plt
                    import wandb
                    plt.figure(figsize=(10, 4))
                    plt.hist(vals, bins="auto")
                    plt.title("Histogram of IoU values" if which == 0 else "Histogram of Masked IoU values")
                    plt.xlabel("IoU")
                    plt.ylabel("Frequency")
                    log({f"feature_space_{key}": wandb.Image(plt)})
                except Exception:
                    log({f"feature_space_{key}_histogram": np.histogram(vals, bins=10, range=(0, 1))[0].tolist()
                         if vals else []})
            grid = self.heatmap_grid(self.dense_logits_for_visualisation(q, k, pair_weights, mask_a, mask_b), mask_a, mask_b,
                                     hidden_image_size)
            self.last_heatmap_grid = grid
            g = grid.detach().cpu().numpy()
            try:                                                # viridis colour map as builder.py:1536-1547
                import matplotlib
                import matplotlib.cm as cm
                import wandb
                norm = matplotlib.colors.Normalize(vmin=g.min(), vmax=g.max(), clip=True)
                rgb = cm.ScalarMappable(norm=norm, cmap="viridis").to_rgba(g)[:, :, :3]
                log({"dense-heatmaps": wandb.Image(rgb)})
            except Exception:
                log({"dense-heatmaps": g})
        if self.rank == 0:
            self.epoch += 1
        self.correlation_ious = []
        self.masked_correlation_ious = []

    def on_train_epoch_end(self, step):
        """builder.py:1608-1664: rank 0 logs the epoch averages, every rank restarts the meters.  Returns the
        logged dict (the reference returns None; main.py:656 ignores the value)."""
        out = {"train/loss": self.loss_o.avg}
        if self.pretrain_type in (PretrainType.CP2, PretrainType.PROPOSED):
            out["train/acc_ins"] = self.acc_ins.avg
        if self.pretrain_type in (PretrainType.DENSECL, PretrainType.PROPOSED_V2, PretrainType.CP2):
            out["train/loss_ins"] = self.loss_i.avg
            out["train/loss_dense"] = self.loss_d.avg
        if self.pretrain_type == PretrainType.CP2:
            out["train/acc_seg"] = self.acc_seg.avg
        if self.pretrain_type in (PretrainType.PROPOSED_V2, PretrainType.CP2):
            out["train/cross_image_variance_source"] = self.cross_image_variance_source.avg
            out["train/cross_image_variance_target"] = self.cross_image_variance_target.avg
        log = self._active_logger()
        if log is not None:
            log(out)
        self.reset_metrics()
        return out

    def reset_metrics(self):                                   # builder.py:1666-1688 (the IoU lists are not touched)
        for mtr in (self.loss_o, self.loss_i, self.loss_d, self.acc_ins, self.acc_seg,
                    self.cross_image_variance_source, self.cross_image_variance_target):
            mtr.reset()

    def _accuracy(self, output, target, topk=(1,)):            # builder.py:1690-1706
        with torch.no_grad():
            maxk = max(topk)
            batch_size = target.size(0)
            _, pred = output.topk(maxk, 1, True, True)
            correct = pred.t().eq(target.view(1, -1))
            return [correct[:k_].reshape(-1).float().sum(0, keepdim=True).mul_(100.0 / batch_size)
                    for k_ in topk]
