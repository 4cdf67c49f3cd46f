"""CPU oracle for the CP2 pre-training loss hot path.

TEST INFRASTRUCTURE ONLY.  This file restates, op by op, what the reference
(`kimathikaai/CP2`, `/root/reference`) computes on the path named by
BASELINE.json.north_star, in plain torch on the CPU (the reference itself is plain
torch, so torch's fp32/fp64 CPU ops *are* the reference arithmetic).  Only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s CPU-baseline / `--impl reference` legs
may import it; `cp2_b200/` never does.

Parity status: PINNED.  `oracle/gen_golden.py` runs the unmodified reference
(`builder.MODEL.forward_cp2`, `_momentum_update_key_encoder`,
`_dequeue_and_enqueue`, `tools.correlation_mapping.*`) in the dev container and
stores its outputs under `tests/golden/`; `tests/test_oracle_golden.py` checks every
function below against those vectors and against the reference's own unit-test
constants (`tests/test_correlation_mapping.py:73-74,126-127`,
`tests/test_contrastive_metrics.py:53-55`).

All `file:line` citations are relative to `/root/reference/`.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F


# --------------------------------------------------------------------------- a1
def composite(img: torch.Tensor, bg: torch.Tensor):
    """Copy-paste compositing of one view.  builder.py:1146-1152.

    The fg mask is 1 where channel 0 of the background is exactly 0; the image is
    multiplied by the mask and the background is *added* (not selected).
    Returns (composited image [B,3,H,W], full-resolution mask [B,H,W] f32).
    """
    mask = (bg[:, 0] == 0).float()
    out = img * mask.unsqueeze(1) + bg
    return out, mask


# --------------------------------------------------------------------------- a2
def stride_subsample(x: torch.Tensor, output_stride: int) -> torch.Tensor:
    """`x[:, OS//2::OS, OS//2::OS]` on a [B,H,W] map.  builder.py:1155-1186."""
    o = output_stride // 2
    return x[:, o::output_stride, o::output_stride]


# --------------------------------------------------------------------------- a5
def normalize_and_pool(feat: torch.Tensor, mask: torch.Tensor):
    """Per-pixel L2 normalisation and masked pooling.  builder.py:1261-1268 (query)
    and 1279-1285 (key).

    feat [B,C,P] (already flattened), mask [B,P] in {0,1}.
    Returns (f_hat [B,C,P], pos [B,C], neg [B,C]) where pos/neg are the normalised
    sums of f_hat over foreground / background pixels.
    """
    f_hat = F.normalize(feat, dim=1)
    pos = F.normalize(torch.einsum("ncx,nx->nc", f_hat, mask), dim=1)
    inv = (~mask.bool()).to(feat.dtype)
    neg = F.normalize(torch.einsum("ncx,nx->nc", f_hat, inv), dim=1)
    return f_hat, pos, neg


# --------------------------------------------------------------------------- a6
def dense_logits(q_hat: torch.Tensor, k_hat: torch.Tensor) -> torch.Tensor:
    """Pixel x pixel cosine similarities [B,P,P].  builder.py:1289."""
    return torch.einsum("ncx,ncy->nxy", q_hat, k_hat)


def dense_infonce(q_hat, k_hat, mask_a, mask_b, temp_local: float):
    """Dense pixel InfoNCE.  builder.py:1289-1292, 1392 (CP2 weights are all ones),
    1431-1437.

    The log-softmax runs over dim=1 of [B,Pq,Pk], i.e. over the QUERY pixels for each
    key pixel; labels are the outer product mask_a (x) mask_b.
    """
    logits = dense_logits(q_hat, k_hat) / temp_local
    labels = torch.einsum("nx,ny->nxy", mask_a, mask_b)
    neg_log = -torch.log_softmax(logits, dim=1)
    n = logits.shape[0]
    num = (neg_log.reshape(n, -1) * labels.reshape(n, -1)).sum(dim=1)
    den = labels.reshape(n, -1).sum(dim=1)
    return (num / den).mean()


def dense_accuracy(q_hat, k_hat, mask_a, mask_b) -> torch.Tensor:
    """`acc_seg`: is the arg-max logit of each image a fg-fg pair?  builder.py:1442-1448.
    (Temperature does not move an arg-max, so it is omitted.)"""
    logits = dense_logits(q_hat, k_hat)
    n = logits.shape[0]
    labels = torch.einsum("nx,ny->nxy", mask_a, mask_b).reshape(n, -1)
    idx = logits.reshape(n, -1).argmax(dim=1)
    return labels[torch.arange(n), idx].float().mean() * 100.0


# --------------------------------------------------------------------------- a7
def global_logits(q_pos, k_pos, queue, q_neg=None, k_neg=None, include_background=False):
    """MoCo logits before the temperature.  builder.py:1395-1397, 1414-1418."""
    l_pos = (q_pos * k_pos).sum(dim=1, keepdim=True)
    l_neg = q_pos @ queue
    parts = [l_pos, l_neg]
    if include_background:
        parts.append((q_pos * q_neg).sum(dim=1, keepdim=True))
        parts.append((q_pos * k_neg).sum(dim=1, keepdim=True))
    return torch.cat(parts, dim=1)


def global_infonce(q_pos, k_pos, queue, temp_global: float, q_neg=None, k_neg=None,
                   include_background=False):
    """Instance-level InfoNCE against the negative queue.  builder.py:1395-1428.
    Cross-entropy with the positive at index 0, mean over the batch."""
    logits = global_logits(q_pos, k_pos, queue, q_neg, k_neg, include_background) / temp_global
    target = torch.zeros(logits.shape[0], dtype=torch.long)
    return F.cross_entropy(logits, target), logits


def topk_accuracy(logits: torch.Tensor, topk=(1, 5)):
    """builder.py:1690-1706 with the positive at column 0."""
    maxk = max(topk)
    pred = logits.topk(maxk, dim=1).indices
    hit = pred == 0
    return [hit[:, :k].any(dim=1).float().sum() * (100.0 / logits.shape[0]) for k in topk]


# ------------------------------------------------------------------ a5+a6+a7+a12
def cp2_loss(q, k, mask_a, mask_b, queue, temp_global=0.2, temp_local=1.0,
             lmbd_dense=0.2, include_background=False, weights=None, negative_type="NONE",
             negative_scale=2.0):
    """Everything `forward_cp2` does between the two encoder outputs and the scalar
    loss, for MappingType.CP2 / NegativeType.NONE.  builder.py:1260-1268, 1279-1292,
    1392-1397, 1414-1439.

    q [B,C,h,w] (requires grad if a gradient is wanted), k [B,C,h,w] (treated as
    detached, builder.py:1271), masks [B,P].  `queue` is the queue BEFORE this step's
    enqueue (builder.py:1396 precedes :1426).
    """
    b, c = q.shape[0], q.shape[1]
    q = q.reshape(b, c, -1)
    k = k.detach().reshape(b, c, -1)
    q_hat, q_pos, q_neg = normalize_and_pool(q, mask_a)
    with torch.no_grad():
        k_hat, k_pos, k_neg = normalize_and_pool(k, mask_b)
    if weights is None and negative_type == "NONE":
        loss_d = dense_infonce(q_hat, k_hat, mask_a, mask_b, temp_local)
    else:                                  # PROPOSED modes
        w = torch.ones(()) if weights is None else weights
        loss_d = dense_infonce_weighted(q_hat, k_hat, mask_a, mask_b, w, temp_local, negative_type,
                                        negative_scale)
    loss_i, logits = global_infonce(q_pos, k_pos, queue.detach(), temp_global, q_neg, k_neg,
                                    include_background)
    loss = loss_i + loss_d * lmbd_dense
    with torch.no_grad():
        acc1, acc5 = topk_accuracy(logits)
        acc_d = dense_accuracy(q_hat, k_hat, mask_a, mask_b)
    return {
        "loss": loss, "loss_ins": loss_i, "loss_dense": loss_d,
        "k_pos": k_pos, "q_pos": q_pos, "acc1": acc1, "acc5": acc5, "acc_dense": acc_d,
        "cross_image_variance_source": q_pos.detach().std(0).mean(),
        "cross_image_variance_target": k_pos.std(0).mean(),
    }


# ------------------------------------------------------------------ f2 (forward_densecl)
def rows_infonce(rows, positives, queue, temp: float):
    """`ContrastiveHead` (builder.py:150-176) for rows whose positive logit is rows . positives:
    CE(cat([pos, rows @ queue]) / temp, 0), mean over the rows."""
    pos = (rows * positives).sum(dim=1, keepdim=True)
    logits = torch.cat([pos, rows @ queue], dim=1) / temp
    return F.cross_entropy(logits, torch.zeros(rows.shape[0], dtype=torch.long))


def densecl_match(embd_q, embd_k):
    """builder.py:817-821: for every query pixel the key pixel with the most similar BACKBONE feature.
    embd_* [B,Cb,P] unit columns -> idx [B,P]."""
    return torch.einsum("ncx,ncy->nxy", embd_q, embd_k).max(dim=2)[1]


def densecl_positive_keys(k_local, idx, pix_q, pix_k, lmbd_coordinate: float):
    """The (detached) vector each query pixel is pulled towards, so that pos_local = q_local_x . kp_x:
    the matched key pixel (builder.py:824-835), mixed with the sum of the key pixels sharing the query
    pixel's id wherever there is one (builder.py:839-855).  k_local [B,C,P] -> kp [B,P,C]."""
    b, c, p = k_local.shape
    kp = torch.gather(k_local, 2, idx[:, None, :].expand(b, c, p)).permute(0, 2, 1)
    corr = (pix_q.reshape(b, -1)[:, :, None] == pix_k.reshape(b, -1)[:, None, :])
    overlap = corr.sum(-1) > 0
    coord = torch.einsum("nxy,ncy->nxc", corr.to(k_local.dtype), k_local)
    return torch.where(overlap[:, :, None], kp * (1 - lmbd_coordinate) + coord * lmbd_coordinate, kp)


def densecl_loss(embd_q, embd_k, q_local, k_local, q_global, k_global, pix_q, pix_k, queue, queue2,
                 temp_global=0.2, temp_local=0.2, lmbd_dense=0.5, lmbd_coordinate=0.0):
    """One direction of `forward_densecl` between the feature extractors and the scalar loss
    (builder.py:762-999): global InfoNCE on the pooled projections, local InfoNCE of every query pixel
    against queue2 with the matched / coordinate-mixed key pixel as its positive.  All inputs are the unit
    vectors `get_query_features` / `get_key_features` return (builder.py:700-760); keys detached."""
    b, c, p = q_local.shape
    loss_g = rows_infonce(q_global, k_global.detach(), queue.detach(), temp_global)
    idx = densecl_match(embd_q.detach(), embd_k.detach())
    kp = densecl_positive_keys(k_local.detach(), idx, pix_q, pix_k, lmbd_coordinate)
    rows = q_local.permute(0, 2, 1).reshape(-1, c)
    loss_l = rows_infonce(rows, kp.reshape(-1, c), queue2.detach(), temp_local)
    return {"loss": (1 - lmbd_dense) * loss_g + lmbd_dense * loss_l, "loss_global": loss_g,
            "loss_local": loss_l, "idx": idx}


# --------------------------------------------------------------------------- a8
def ema_update(params_q, params_k, m: float):
    """Momentum update of the key encoder, parameters only.  builder.py:557-567.

    Three separately rounded fp32 ops per element: k*m, q*(1-m) with (1-m) evaluated
    in Python double and rounded to fp32 by torch's scalar handling, then the add.
    Returns new tensors (the reference rebinds `param_k.data`).
    """
    return [pk * m + pq * (1.0 - m) for pq, pk in zip(params_q, params_k)]


# ---------------------------------------------------------------------- a9 / a10
def concat_all_gather(per_rank: list) -> torch.Tensor:
    """What every rank holds after `concat_all_gather`: the rank-ordered concatenation
    along dim 0.  builder.py:1710-1722."""
    return torch.cat(list(per_rank), dim=0)


def dequeue_and_enqueue(queue: torch.Tensor, ptr: int, keys_all: torch.Tensor):
    """Writes the gathered keys [n,C] as columns of queue [C,K] starting at `ptr`, with
    wrap-around, and advances the pointer.  builder.py:569-587.  In place on `queue`;
    returns the new pointer.  (As in the reference, n <= K is assumed.)"""
    c, kq = queue.shape
    n = keys_all.shape[0]
    if ptr + n > kq:
        head = kq - ptr
        queue[:, ptr:kq] = keys_all[:head].T
        queue[:, 0:n - head] = keys_all[head:].T
    else:
        queue[:, ptr:ptr + n] = keys_all.T
    return (ptr + n) % kq


# -------------------------------------------------------------------------- a11
def batch_shuffle(x_all: torch.Tensor, idx_shuffle: torch.Tensor, rank: int, world: int):
    """Rows rank `rank` feeds to its key encoder, and the inverse permutation.
    builder.py:609-630."""
    idx_unshuffle = torch.argsort(idx_shuffle)
    return x_all[idx_shuffle.view(world, -1)[rank]], idx_unshuffle


def batch_unshuffle(x_all: torch.Tensor, idx_unshuffle: torch.Tensor, rank: int, world: int):
    """builder.py:632-649."""
    return x_all[idx_unshuffle.view(world, -1)[rank]]


# --------------------------------------------------------------------------- a4
def masked_iou(map_a, map_b, mask_a, mask_b) -> torch.Tensor:
    """IoU of the id sets of two flattened id maps restricted to their masks.
    tools/correlation_mapping.py:103-138.  Ids are shifted by +1 so that 0 can mean
    "masked out"; union = number of distinct surviving ids, intersection = number of
    ids that survive more than once."""
    b = map_a.shape[0]
    out = torch.zeros(b)
    for i in range(b):
        ids = torch.cat([(map_a[i] + 1) * mask_a[i], (map_b[i] + 1) * mask_b[i]])
        ids = ids[ids != 0]
        u, cnt = torch.unique(ids, return_counts=True)
        out[i] = float((cnt > 1).sum()) / float(len(u))
    return out


def correlation_map(map_a, map_b) -> torch.Tensor:
    """[B,P,P] bool: ids equal.  tools/correlation_mapping.py:141-161."""
    b = map_a.shape[0]
    return map_a.reshape(b, -1)[:, :, None] == map_b.reshape(b, -1)[:, None, :]


def masked_correlation(map_a, map_b, mask_a, mask_b):
    """tools/correlation_mapping.py:192-247 (the fields forward_cp2 consumes)."""
    b = map_a.shape[0]
    corr = correlation_map(map_a, map_b)
    ones = torch.ones(b, map_a[0].numel())
    fa, fb = map_a.reshape(b, -1), map_b.reshape(b, -1)
    return {
        "corr_map": corr,
        "iou": masked_iou(fa, fb, ones, ones),
        "iou_masked": masked_iou(fa, fb, mask_a.reshape(b, -1), mask_b.reshape(b, -1)),
    }


def cp2_corr_weights(pixel_ids_a, pixel_ids_b, region_ids_a, region_ids_b, mask_a, mask_b,
                     w_pixel=1, w_region=1, w_not=1):
    """builder.py:1204-1243: region weight where region ids match (and both are known,
    i.e. non-zero), overridden by the pixel weight where pixel ids match, and the
    "not" weight wherever the result is still zero.  All ones in CP2 mode.  The dtype follows
    the reference's `w_region * bool_tensor` (int64 for python ints, f32 for floats as main.py:75-77
    passes them)."""
    b = mask_a.shape[0]
    pix = correlation_map(pixel_ids_a, pixel_ids_b)
    reg = correlation_map(region_ids_a, region_ids_b)
    known = (region_ids_a.reshape(b, -1)[:, :, None] * region_ids_b.reshape(b, -1)[:, None, :]) != 0
    w = w_region * (reg & known)
    w[pix] = w_pixel
    w = w + (w == 0) * w_not
    return w


def negative_transform(logits, labels, negative_type: str, negative_scale: float):
    """builder.py:1332-1386 on a copy: the non-label entries are squashed to
    2*sigmoid(scale*(s - centre)) - 1 with centre 0 (FIXED) or the per-sample mean of the non-label
    scores (AVERAGE) or their per-sample median, `torch.nanquantile(q=0.5)` (MEDIAN), both detached.  HARD
    leaves the logits untouched (its `*= 1.5` lands on a temporary)."""
    if negative_type in ("NONE", "HARD"):
        return logits
    neg = ~labels.bool()
    if negative_type == "FIXED":
        centre = torch.zeros(logits.shape[0], 1, 1)
    elif negative_type == "AVERAGE":
        centre = dense_loss_stats(logits, labels)["negative"]["average"].detach().reshape(-1, 1, 1)
    elif negative_type == "MEDIAN":
        centre = dense_loss_stats(logits, labels)["negative"]["quartiles"][1].detach().reshape(-1, 1, 1)
    else:
        raise NotImplementedError(negative_type)
    squashed = 2.0 / (1.0 + torch.exp((logits - centre) * -negative_scale)) - 1.0
    return torch.where(neg, squashed, logits)


def dense_infonce_weighted(q_hat, k_hat, mask_a, mask_b, weights, temp_local: float,
                           negative_type: str = "NONE", negative_scale: float = 2.0):
    """The dense loss of `forward_cp2` for the PROPOSED modes: builder.py:1289-1292, 1332-1392
    (negative transform, then `* corr_weights`), 1431-1437."""
    logits = dense_logits(q_hat, k_hat)
    labels = torch.einsum("nx,ny->nxy", mask_a, mask_b)
    logits = negative_transform(logits, labels, negative_type, negative_scale) * weights / temp_local
    neg_log = -torch.log_softmax(logits, dim=1)
    n = logits.shape[0]
    num = (neg_log.reshape(n, -1) * labels.reshape(n, -1)).sum(dim=1)
    return (num / labels.reshape(n, -1).sum(dim=1)).mean()


def dense_loss_stats(logits, labels):
    """Per-sample mean and quartiles of positive / negative dense scores.
    tools/correlation_mapping.py:11-53 (torch.nanquantile, linear interpolation)."""
    q = torch.tensor([0.25, 0.5, 0.75])

    def stats(s):
        return {"average": s.nanmean((1, 2)),
                "quartiles": tuple(torch.nanquantile(s.flatten(1), q=q, dim=1))}

    pos = logits.detach().clone()
    pos[~labels.bool()] = float("nan")
    neg = logits.detach().clone()
    neg[labels.bool()] = float("nan")
    return {"positive": stats(pos), "negative": stats(neg)}


# -------------------------------------------------------------------------- a13
def mirror_loss(s_logits, t_logits, masks, lmbd_compare: float, softmax_temp: float):
    """MirrorModule.shared_step, OUTPUT variant.  networks/mirror_network.py:40-63.

    compare = CE(input=softmax(s/T), target=softmax(t/T)) — i.e. the *input* to the
    cross-entropy is already a probability map (double softmax) and the target is a
    soft distribution over the class dim; class = CE(cat(s,t), cat(masks,masks)).
    """
    ps = torch.softmax(s_logits / softmax_temp, dim=1)
    pt = torch.softmax(t_logits / softmax_temp, dim=1)
    compare = F.cross_entropy(ps, pt)
    cls = F.cross_entropy(torch.cat([s_logits, t_logits]), torch.cat([masks, masks]))
    return cls + lmbd_compare * compare, cls, compare


def mirror_loss_none(logits, masks):
    """MirrorVariant.NONE (networks/mirror_network.py:53-55,62-63): class CE only."""
    cls = F.cross_entropy(logits, masks)
    return cls, cls, torch.zeros(())


def mirror_resize(logits_lo, size):
    """SegmentationModule.forward's `resize` (networks/segment_network.py:222-228)."""
    return F.interpolate(logits_lo, size=tuple(size), mode="bilinear", align_corners=False)


def mirror_step(s_lo, t_lo, masks, lmbd_compare: float, softmax_temp: float):
    """shared_step from the segmentor's low-resolution logits: resize then the loss
    (networks/mirror_network.py:40-63).  t_lo None = MirrorVariant.NONE."""
    size = masks.shape[-2:]
    if t_lo is None:
        return mirror_loss_none(mirror_resize(s_lo, size), masks)
    return mirror_loss(mirror_resize(s_lo, size), mirror_resize(t_lo, size), masks, lmbd_compare, softmax_temp)


# --------------------------------------------------------------- synthetic inputs
def synth_views(batch, height, width, output_stride, seed, fg_lo=0.5, fg_hi=0.8):
    """Deterministic synthetic inputs of the shape `forward_cp2` takes:
    images ~U[0,1); backgrounds ~U[0,1) with one rectangle per image set to exactly 0
    on all channels (mirrors RandomErasing at main.py:220-225), area fraction
    ~U(fg_lo,fg_hi), aspect ~logU(0.8,1.25).  The rectangle always covers at least one
    point of the [OS/2::OS] lattice so the dense loss is finite (builder.py:1436).
    """
    g = torch.Generator().manual_seed(seed)
    out = {}
    for name in ("img_a", "img_b", "bg0", "bg1"):
        out[name] = torch.rand(batch, 3, height, width, generator=g)
    for name in ("bg0", "bg1"):
        for i in range(batch):
            frac = fg_lo + (fg_hi - fg_lo) * torch.rand((), generator=g).item()
            log_ar = torch.empty(()).uniform_(-0.2231, 0.2231, generator=g).item()
            ar = float(torch.tensor(log_ar).exp())
            rh = int(min(height, max(output_stride, round((frac * height * width * ar) ** 0.5))))
            rw = int(min(width, max(output_stride, round((frac * height * width / ar) ** 0.5))))
            top = int(torch.randint(0, height - rh + 1, (1,), generator=g))
            left = int(torch.randint(0, width - rw + 1, (1,), generator=g))
            out[name][i, :, top:top + rh, left:left + rw] = 0.0
    ids = torch.arange(1, height * width + 1, dtype=torch.int64).reshape(1, height, width)
    out["pixel_ids_a"] = ids.repeat(batch, 1, 1)
    out["pixel_ids_b"] = ids.repeat(batch, 1, 1)
    out["region_ids_a"] = out["pixel_ids_a"].clone()
    out["region_ids_b"] = out["pixel_ids_b"].clone()
    return out


# ------------------------------------------------- whole pre-training step (CPU port)
class CP2StepPort:
    """CPU restatement of ONE reference pre-training iteration: `train()` body
    main.py:616-644 around `MODEL.forward_cp2` (builder.py:1124-1439) in the order the
    reference runs it (Appendix A.3 of SURVEY.md): composite both views, encoder_q,
    EMA, encoder_k, losses against the queue BEFORE enqueue, enqueue, backward, SGD.

    Used only as bench.py's CPU baseline / `--impl reference` arm (kind "port": the
    reference is pure Python and cannot travel to the GPU box).  World size 1, so the
    batch shuffle (builder.py:609-649) is the identity up to a permutation that BN
    batch statistics do not see.  `encoder_factory()` builds one segmentor (the
    out-of-scope backbone; plain torch on the CPU, as in the reference).
    """

    def __init__(self, encoder_factory, dim=128, K=65536, m=0.999, temp_global=0.2, temp_local=1.0,
                 lmbd_dense=0.2, output_stride=16, lr=0.001, seed=0):
        torch.manual_seed(seed)
        self.encoder_q = encoder_factory()
        self.encoder_k = encoder_factory()
        for pq, pk in zip(self.encoder_q.parameters(), self.encoder_k.parameters()):
            pk.data.copy_(pq.data)                      # builder.py:464-469
            pk.requires_grad = False
        self.queue = F.normalize(torch.randn(dim, K), dim=0)   # builder.py:476-477
        self.ptr = 0
        self.correlation_ious, self.masked_correlation_ious = [], []
        self.m, self.tg, self.tl, self.lmbd, self.os = m, temp_global, temp_local, lmbd_dense, output_stride
        self.opt = torch.optim.SGD(self.encoder_q.parameters(), lr, momentum=0.9, weight_decay=1e-4)  # main.py:467-471

    def step(self, img_a, img_b, bg0, bg1, pixel_ids_a=None, pixel_ids_b=None, region_ids_a=None,
             region_ids_b=None):
        img_a, mask_a = composite(img_a, bg0)
        img_b, mask_b = composite(img_b, bg1)
        n = img_a.shape[0]
        mask_a = stride_subsample(mask_a, self.os)
        mask_b = stride_subsample(mask_b, self.os)
        if region_ids_a is not None:
            # builder.py:1165-1186, 1203-1217, 1252-1257: both correlation maps and the region-id IoUs
            pa, pb = stride_subsample(pixel_ids_a, self.os), stride_subsample(pixel_ids_b, self.os)
            ra, rb = stride_subsample(region_ids_a, self.os), stride_subsample(region_ids_b, self.os)
            masked_correlation(pa, pb, mask_a, mask_b)
            reg = masked_correlation(ra, rb, mask_a, mask_b)
            self.correlation_ious.extend(reg["iou"].tolist())
            self.masked_correlation_ious.extend(reg["iou_masked"].tolist())
        mask_a = mask_a.reshape(n, -1)
        mask_b = mask_b.reshape(n, -1)
        q = self.encoder_q(img_a)
        with torch.no_grad():
            pq = [p.data for p in self.encoder_q.parameters()]
            pk = [p.data for p in self.encoder_k.parameters()]
            for p, new in zip(self.encoder_k.parameters(), ema_update(pq, pk, self.m)):
                p.data = new
            k = self.encoder_k(img_b)
        r = cp2_loss(q, k, mask_a, mask_b, self.queue.clone(), self.tg, self.tl, self.lmbd)  # clone: builder.py:1396
        self.ptr = dequeue_and_enqueue(self.queue, self.ptr, r["k_pos"])
        self.opt.zero_grad()
        r["loss"].backward()
        self.opt.step()
        return float(r["loss"].detach())
