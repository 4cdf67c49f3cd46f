"""Host-side wrappers: torch tensors in, C-ABI calls on the current CUDA stream.

torch is used for device memory, streams and autograd plumbing only; every op on the
hot path is a kernel of libcp2_b200.so.  All wrappers raise if their inputs are not
CUDA tensors — there is no CPU path.
"""
import math

import torch

from . import _lib

_EPS = 1e-12   # F.normalize eps
SUPPORTED_DIMS = (128, 256)     # feature dims of the tcgen05 kernels (cp2_dense_*, cp2_global_nce)
MAX_INV_TAU = 16.0              # fixed soft-max shift: |logit| <= 1/tau <= 16 keeps exp2 in range

# ------------------------------------------------------------ launch accounting
# Every C-ABI call goes through `_call`: it counts the kernels launched (bench.py's
# `gpu_launches`) and, when a KernelTimer is active, brackets the call with CUDA events on
# the launching stream (bench.py's per-kernel roofline figures).
LAUNCHES = 0
_timer = None


class KernelTimer:
    """with KernelTimer() as t: ...   then t.summary() -> {name: (n_calls, mean_us)}."""

    def __init__(self):
        self.records = {}

    def __enter__(self):
        global _timer
        self._prev, _timer = _timer, self
        return self

    def __exit__(self, *exc):
        global _timer
        _timer = self._prev

    def summary(self):
        torch.cuda.synchronize()
        out = {}
        for name, evs in self.records.items():
            ms = [a.elapsed_time(b) for a, b in evs]
            out[name] = (len(ms), 1e3 * sum(ms) / max(1, len(ms)))
        return out


def _call(name, n_kernels, fn, *args):
    global LAUNCHES
    LAUNCHES += n_kernels
    if _timer is None:
        _lib.check(fn(*args), name)
        return
    e0 = torch.cuda.Event(enable_timing=True)
    e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    _lib.check(fn(*args), name)
    e1.record()
    _timer.records.setdefault(name, []).append((e0, e1))


def _ptr(t):
    return None if t is None else t.data_ptr()


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _need_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError("cp2_b200 ops need CUDA tensors (B200, sm_100a); there is no CPU fallback")


def _f32c(t):
    if t.dtype != torch.float32:
        t = t.float()
    return t if t.is_contiguous() else t.contiguous()


# Persistent zero-initialised scratch (done counters / fixed-point accumulators): the kernels that use one
# leave it all-zero again, so it is allocated and cleared once per (device, stream, tag, size).
_scratch = {}


def _persistent_zero(tag, nbytes, device):
    key = (device.index, torch.cuda.current_stream(device).cuda_stream, tag, int(nbytes))
    t = _scratch.get(key)
    if t is None:
        t = torch.zeros((int(nbytes) + 7) // 8, device=device, dtype=torch.int64)
        _scratch[key] = t
    return t


# ------------------------------------------------------------------ K1
def composite(img, bg, output_stride, want_full_mask=False, out=None):
    """builder.py:1146-1159.  Returns (composited [B,3,H,W], mask_ds [B,hs,ws], mask_full|None)."""
    _need_cuda(img, bg)
    img, bg = _f32c(img), _f32c(bg)
    B, ch, H, W = img.shape
    assert ch == 3 and bg.shape == img.shape, f"{img.shape = }, {bg.shape = }"
    o = output_stride // 2
    hs, ws = len(range(o, H, output_stride)), len(range(o, W, output_stride))
    out = torch.empty_like(img) if out is None else out
    mask_ds = torch.empty((B, hs, ws), device=img.device, dtype=torch.float32)
    mask_full = torch.empty((B, H, W), device=img.device, dtype=torch.float32) if want_full_mask else None
    lib = _lib.load()
    _call("cp2_composite", 1, lib.cp2_composite, _ptr(img), _ptr(bg), _ptr(out), _ptr(mask_full), _ptr(mask_ds),
                                 B, H, W, output_stride, _stream())
    return out, mask_ds, mask_full


def composite_pair(img_a, bg_a, img_b, bg_b, output_stride, want_full_mask=False):
    """Both views of builder.py:1146-1159 in ONE launch -> ((out_a, mask_ds_a, full_a), (out_b, mask_ds_b, full_b))."""
    _need_cuda(img_a, bg_a, img_b, bg_b)
    img_a, bg_a, img_b, bg_b = _f32c(img_a), _f32c(bg_a), _f32c(img_b), _f32c(bg_b)
    B, ch, H, W = img_a.shape
    assert ch == 3 and bg_a.shape == img_a.shape and img_b.shape == img_a.shape and bg_b.shape == img_a.shape
    o = output_stride // 2
    hs, ws = len(range(o, H, output_stride)), len(range(o, W, output_stride))
    dev = img_a.device
    outs = [torch.empty_like(img_a), torch.empty_like(img_b)]
    mds = [torch.empty((B, hs, ws), device=dev, dtype=torch.float32) for _ in range(2)]
    full = [torch.empty((B, H, W), device=dev, dtype=torch.float32) if want_full_mask else None for _ in range(2)]
    _call("cp2_composite2", 1, _lib.load().cp2_composite2, _ptr(img_a), _ptr(bg_a), _ptr(outs[0]), _ptr(full[0]),
          _ptr(mds[0]), _ptr(img_b), _ptr(bg_b), _ptr(outs[1]), _ptr(full[1]), _ptr(mds[1]), B, H, W, output_stride,
          _stream())
    return (outs[0], mds[0], full[0]), (outs[1], mds[1], full[1])


# ----------------------------------------------------------------- K2a
def prep_features(feat, mask, compact=False, unit=False, perm=None):
    """builder.py:1261-1268.  feat [B,C,P] f32, mask [B,P] -> (fhat_t bf16 [B,P,C], inv_norm [B,P],
    pos_sum [B,C], neg_sum [B,C], key_count).  compact=True (key side): fhat_t holds the foreground
    rows first (pixel order) then the background rows, and key_count [B] int32 is the number of foreground
    rows; else None.  perm: optional [B,P] int32 OUT, the pixel stored in each row (compact only).
    unit=True appends (pos_unit [B,C], neg_unit [B,C], norms [B,2]): the outer F.normalize of the pooled
    vectors (builder.py:1264,1268) and the raw norms its backward needs."""
    _need_cuda(feat, mask)
    feat, mask = _f32c(feat), _f32c(mask)
    B, C, P = feat.shape
    lib = _lib.load()
    dev = feat.device
    fhat_t = torch.empty((B, P, C), device=dev, dtype=torch.bfloat16)
    inv_norm = torch.empty((B, P), device=dev, dtype=torch.float32)
    pos = torch.empty((B, C), device=dev, dtype=torch.float32)
    neg = torch.empty((B, C), device=dev, dtype=torch.float32)
    scratch = torch.empty((B, lib.cp2_prep_tiles(P), 2, C), device=dev, dtype=torch.float32)
    count = torch.empty((B,), device=dev, dtype=torch.int32) if compact else None
    pu = torch.empty((B, C), device=dev, dtype=torch.float32) if unit else None
    nu = torch.empty((B, C), device=dev, dtype=torch.float32) if unit else None
    norms = torch.empty((B, 2), device=dev, dtype=torch.float32) if unit else None
    _call("cp2_prep_features", 1, lib.cp2_prep_features, _ptr(feat), _ptr(mask), _ptr(fhat_t), _ptr(inv_norm),
          _ptr(pos), _ptr(neg), _ptr(scratch), _ptr(count), _ptr(perm), _ptr(pu), _ptr(nu), _ptr(norms),
          _ptr(_persistent_zero("prep", 4 * B, dev)), B, C, P, _stream())
    if unit:
        return fhat_t, inv_norm, pos, neg, count, pu, nu, norms
    return fhat_t, inv_norm, pos, neg, count


def prep_features_pair(feat_q, mask_q, feat_k, mask_k, want_perm=False):
    """Query side and (compacted) key side of one step in ONE launch.  -> two tuples shaped like
    prep_features(..., unit=True): (fhat_t, inv_norm, pos_sum, neg_sum, key_count|None, pos_unit, neg_unit, norms);
    want_perm appends perm_k [B,P] int32 (pixel of each compacted key row) to the key tuple."""
    _need_cuda(feat_q, mask_q, feat_k, mask_k)
    feat_q, mask_q, feat_k, mask_k = _f32c(feat_q), _f32c(mask_q), _f32c(feat_k), _f32c(mask_k)
    B, C, P = feat_q.shape
    assert feat_k.shape == feat_q.shape
    lib, dev = _lib.load(), feat_q.device
    f32 = dict(device=dev, dtype=torch.float32)
    tiles = lib.cp2_prep_tiles(P)

    def side():
        return [torch.empty((B, P, C), device=dev, dtype=torch.bfloat16), torch.empty((B, P), **f32),
                torch.empty((B, C), **f32), torch.empty((B, C), **f32), torch.empty((B, tiles, 2, C), **f32),
                torch.empty((B, C), **f32), torch.empty((B, C), **f32), torch.empty((B, 2), **f32)]
    q, k = side(), side()
    count = torch.empty((B,), device=dev, dtype=torch.int32)
    perm = torch.empty((B, P), device=dev, dtype=torch.int32) if want_perm else None
    _call("cp2_prep_features2", 1, lib.cp2_prep_features2,
          _ptr(feat_q), _ptr(mask_q), *[_ptr(t) for t in q],
          _ptr(feat_k), _ptr(mask_k), *[_ptr(t) for t in k[:5]], _ptr(count), _ptr(perm), *[_ptr(t) for t in k[5:]],
          _ptr(_persistent_zero("prep2", 8 * B, dev)), B, C, P, _stream())
    kt = (k[0], k[1], k[2], k[3], count, k[5], k[6], k[7])
    return ((q[0], q[1], q[2], q[3], None, q[5], q[6], q[7]), kt + (perm,) if want_perm else kt)


# ------------------------------------------------------------------ K2
def dense_lse(qhat_t, khat_t, inv_tau, key_count=None):
    """lse[n,y] = log sum_x exp(qhat[n,x].khat[n,y] * inv_tau).  builder.py:1289,1431-1432.
    key_count [B] int32 (compacted keys): rows >= key_count[n] are skipped / meaningless."""
    _need_cuda(qhat_t, khat_t)
    B, P, C = qhat_t.shape
    lse = torch.empty((B, P), device=qhat_t.device, dtype=torch.float32)
    _call("cp2_dense_lse", 1, _lib.load().cp2_dense_lse, _ptr(qhat_t), _ptr(khat_t), _ptr(key_count), _ptr(lse),
          float(inv_tau), B, C, P, _stream())
    return lse


def dense_grad(qhat_t, khat_t, lse, mask_a, mask_b, inv_norm, alpha, v1, v2, inv_tau, key_count=None, gscale=None):
    """Fused backward of the dense + global loss w.r.t. the query features -> dq [B,C,P].
    mask_b may be None when the keys are compacted (key_count given); gscale: None or a 1-element f32
    device tensor, the upstream gradient multiplied into alpha / v1 / v2 inside the kernel."""
    B, P, C = qhat_t.shape
    lib = _lib.load()
    dq = torch.empty((B, C, P), device=qhat_t.device, dtype=torch.float32)
    nl = torch.empty((B, lib.cp2_dense_pad(P)), device=qhat_t.device, dtype=torch.float32)
    _call("cp2_dense_grad", 2, lib.cp2_dense_grad, _ptr(qhat_t), _ptr(khat_t), _ptr(key_count), _ptr(lse),
          _ptr(mask_a), _ptr(mask_b), _ptr(inv_norm), _ptr(alpha), _ptr(v1), _ptr(v2), _ptr(gscale), _ptr(dq),
          _ptr(nl), float(inv_tau), B, C, P, _stream())
    return dq


def dense_small_ok(C, P):
    """True when the single-kernel path for maps that fit TMEM takes this shape (cp2_dense_small_ok)."""
    return bool(_lib.load().cp2_dense_small_ok(int(C), int(P)))


_SM_COUNT = {}


def dense_small_preferred(B, C, P, device=None):
    """The single-kernel path runs one CTA per SM (all of TMEM, both operands of the image in shared memory): it
    wins while the batch fits one wave (B * ceil(P/128) <= SM count; measured at P=196: 27 us vs 30 us at B=32,
    65 us vs 52 us at B=256), beyond that the streaming pair (2 CTAs per SM in the forward) has the throughput."""
    if not dense_small_ok(C, P):
        return False
    dev = None if device is None else torch.device(device).index
    if dev is None:
        dev = torch.cuda.current_device()
    if dev not in _SM_COUNT:
        _SM_COUNT[dev] = torch.cuda.get_device_properties(dev).multi_processor_count
    return B * ((P + 127) // 128) <= _SM_COUNT[dev]


def dense_small(qhat_t, khat_t, inv_tau, key_count=None, mask_b=None, seg=None):
    """cp2_dense_small: column LSE and the P^2-sized part of the gradient in ONE launch (P <= 256).
    -> (lse [B,P], acc [B,C,P]); acc[n,:,x] = sum_y on_y exp(S[x,y]/tau - lse[y]) khat[y,:].
    seg = (q_f32 [B,C,P], k_f32 [B,C,P], inv_norm_q [B,P], inv_norm_k [B,P], perm_k [B,P] int32 | None,
           mask_a [B,P], mask_b [B,P] | None) turns the acc_seg meter on (builder.py:1442-1448): a third result
    hit [B] = label of the image's arg-max pixel pair, decided in fp32."""
    _need_cuda(qhat_t, khat_t)
    B, P, C = qhat_t.shape
    dev = qhat_t.device
    lse = torch.empty((B, P), device=dev, dtype=torch.float32)
    acc = torch.empty((B, C, P), device=dev, dtype=torch.float32)
    segp, hit = [None] * 7, None
    if seg is not None:
        segp = list(seg)
        for t in segp[:4] + segp[5:]:
            assert t is None or (t.dtype == torch.float32 and t.is_contiguous())
        hit = torch.empty((B,), device=dev, dtype=torch.float32)
    _call("cp2_dense_small", 1, _lib.load().cp2_dense_small, _ptr(qhat_t), _ptr(khat_t), _ptr(key_count), _ptr(mask_b),
          _ptr(lse), _ptr(acc), *[_ptr(t) for t in segp], _ptr(hit), float(inv_tau), B, C, P, _stream())
    return (lse, acc) if seg is None else (lse, acc, hit)


def dense_small_bwd(acc, qhat_t, mask_a, inv_norm, alpha, v1, v2, key_count=None, gscale=None):
    """cp2_dense_small_bwd: rank-1 terms, upstream gradient (device scalar) and the F.normalize backward applied to
    cp2_dense_small's acc -> dq [B,C,P] (what cp2_dense_grad's epilogue does)."""
    B, P, C = qhat_t.shape
    dq = torch.empty((B, C, P), device=qhat_t.device, dtype=torch.float32)
    _call("cp2_dense_small_bwd", 1, _lib.load().cp2_dense_small_bwd, _ptr(acc), _ptr(qhat_t), _ptr(key_count),
          _ptr(mask_a), _ptr(inv_norm), _ptr(alpha), _ptr(v1), _ptr(v2), _ptr(gscale), _ptr(dq), B, C, P, _stream())
    return dq


def dense_fwd_bwd(qhat_t, khat_t, key_count, mask_a, inv_norm, alpha, v1, v2, inv_tau, gscale=None):
    """Column LSE + gradient w.r.t. the query features of the dense loss in one call -> (lse [B,P], dq [B,C,P]).
    The configs[3] microbench unit: cp2_dense_small + cp2_dense_small_bwd for maps that fit TMEM, else
    cp2_dense_lse + cp2_dense_grad."""
    B, P, C = qhat_t.shape
    if dense_small_preferred(B, C, P, qhat_t.device):
        lse, acc = dense_small(qhat_t, khat_t, inv_tau, key_count)
        return lse, dense_small_bwd(acc, qhat_t, mask_a, inv_norm, alpha, v1, v2, key_count=key_count, gscale=gscale)
    lse = dense_lse(qhat_t, khat_t, inv_tau, key_count)
    dq = dense_grad(qhat_t, khat_t, lse, mask_a, None, inv_norm, alpha, v1, v2, inv_tau, key_count=key_count,
                    gscale=gscale)
    return lse, dq


# ------------------------------------------------------------------ K3
class QueueShadow:
    """The bf16 entry-major [K,C] shadow of one fp32 [C,K] queue buffer that cp2_global_nce streams.
    cp2_enqueue rewrites the rows it enqueues, so in steady state the shadow never needs a conversion pass;
    it is rebuilt (cp2_queue_shadow, one transpose kernel) when the fp32 buffer was replaced or written by
    anything else (torch in-place ops bump `queue._version`: load_state_dict, copy_, normalisation at init)."""

    def __init__(self):
        self.bf = None           # [K,C] view of the storage below
        self._store = None       # [rows >= K, C]; rows past K stay zero (block padding of queue_sums_tc)
        self._key = None

    def _sig(self, queue):
        return (queue.data_ptr(), queue._version, tuple(queue.shape))

    def get(self, queue, min_rows=0):
        _need_cuda(queue)
        assert queue.dtype == torch.float32 and queue.is_contiguous() and queue.dim() == 2
        C, K = queue.shape
        if self.bf is None or self._key != self._sig(queue) or self._store.shape[0] < min_rows:
            rows = max(K, int(min_rows), 0 if self._store is None else self._store.shape[0])
            if self._store is None or self._store.shape != (rows, C) or self._store.device != queue.device:
                self._store = torch.zeros((rows, C), device=queue.device, dtype=torch.bfloat16)
            self.bf = self._store[:K]
            _call("cp2_queue_shadow", 1, _lib.load().cp2_queue_shadow, _ptr(queue), _ptr(self.bf), C, K, _stream())
            self._key = self._sig(queue)
        return self.bf

    def padded(self, queue, rows):
        """[rows, C] view (rows >= K): the shadow followed by zero rows."""
        self.get(queue, min_rows=rows)
        return self._store[:rows]


def global_nce(q_pos, queue, inv_tau, k_pos=None, q_neg=None, k_neg=None, shadow=None):
    """One launch, one pass over the queue: z[b] = sum_k exp((l-1)/tau), g[b,:] = sum_k exp((l-1)/tau) queue[:,k]
    and, with k_pos, the top-1 meter (hit [B], acc [1]).  -> (z, g, hit|None, acc|None).
    shadow: a QueueShadow kept across steps (MODEL does); None converts the queue for this call."""
    _need_cuda(q_pos, queue)
    q_pos = _f32c(q_pos)
    B, C = q_pos.shape
    assert queue.dtype == torch.float32 and queue.is_contiguous() and queue.shape[0] == C
    K = queue.shape[1]
    lib = _lib.load()
    dev = q_pos.device
    qbf = (shadow if shadow is not None else QueueShadow()).get(queue)
    z = torch.empty((B,), device=dev, dtype=torch.float32)
    g = torch.empty((B, C), device=dev, dtype=torch.float32)
    hit = acc = None
    if k_pos is not None:
        k_pos = _f32c(k_pos)
        hit = torch.empty((B,), device=dev, dtype=torch.float32)
        acc = torch.empty((1,), device=dev, dtype=torch.float32)
    if q_neg is not None:
        q_neg, k_neg = _f32c(q_neg), _f32c(k_neg)
    scratch = _persistent_zero("global", lib.cp2_global_scratch_bytes(B, C), dev)
    _call("cp2_global_nce", 1, lib.cp2_global_nce, _ptr(q_pos), _ptr(qbf), _ptr(queue), _ptr(k_pos), _ptr(q_neg),
          _ptr(k_neg), _ptr(z), _ptr(g), _ptr(hit), _ptr(acc), _ptr(scratch), float(inv_tau), B, C, K, _stream())
    return z, g, hit, acc


# ------------------------------------------------------------------ K4
class EmaPlan:
    """Device-side pointer / chunk tables for cp2_ema (built once per model; rebuilt if any
    parameter storage moves).  builder.py:557-567."""

    def __init__(self, params_q, params_k):
        self._key = None
        self.rebind(params_q, params_k)
        self._build()

    def rebind(self, params_q, params_k):
        """Points the plan at (possibly re-allocated) parameter tensors; the device tables are
        rebuilt lazily by `apply` only if an address or size actually changed."""
        self.params_q = list(params_q)
        self.params_k = list(params_k)
        assert len(self.params_q) == len(self.params_k)

    def _signature(self):
        return tuple((pq.data_ptr(), pk.data_ptr(), pk.numel(), pk.stride())
                     for pq, pk in zip(self.params_q, self.params_k))

    def _build(self):
        _need_cuda(*self.params_q, *self.params_k)
        for pq, pk in zip(self.params_q, self.params_k):
            if pq.dtype != torch.float32 or pk.dtype != torch.float32:
                raise RuntimeError("cp2_ema: fp32 parameters only")
            # elementwise over the storage: any dense layout works (contiguous or channels_last)
            # as long as both sides use the same one
            dense = pk.is_contiguous() or (pk.dim() == 4 and pk.is_contiguous(memory_format=torch.channels_last))
            if not dense or pq.shape != pk.shape or pq.stride() != pk.stride():
                raise RuntimeError("cp2_ema: parameters must be dense, shape- and stride-matched")
        dev = self.params_k[0].device
        self.n_params = sum(p.numel() for p in self.params_k)
        ct, ci = [], []
        for t, pk in enumerate(self.params_k):
            n = (pk.numel() + _lib.EMA_CHUNK - 1) // _lib.EMA_CHUNK
            ct.extend([t] * n)
            ci.extend(range(n))
        i64 = dict(device=dev, dtype=torch.int64)
        self.q_ptrs = torch.tensor([p.data_ptr() for p in self.params_q], **i64)
        self.k_ptrs = torch.tensor([p.data_ptr() for p in self.params_k], **i64)
        self.numel = torch.tensor([p.numel() for p in self.params_k], **i64)
        self.chunk_tensor = torch.tensor(ct, device=dev, dtype=torch.int32)
        self.chunk_index = torch.tensor(ci, device=dev, dtype=torch.int32)
        self.n_chunks = len(ct)
        self._key = self._signature()

    @torch.no_grad()
    def apply(self, m):
        if self._signature() != self._key:
            self._build()
        # the reference's scalars: m and (1.0 - m) evaluated in double, rounded to fp32 by the op
        _call("cp2_ema", 1, _lib.load().cp2_ema, _ptr(self.q_ptrs), _ptr(self.k_ptrs), _ptr(self.numel),
                                      _ptr(self.chunk_tensor), _ptr(self.chunk_index), self.n_chunks,
                                      float(m), float(1.0 - m), _stream())


# ------------------------------------------------------------------ K5
@torch.no_grad()
def enqueue(keys_all, queue, queue_ptr, shadow=None):
    """builder.py:574-587 on device, one launch: queue[:, (ptr+i) % K] = keys_all[i]; ptr = (ptr+n) % K.
    shadow: the queue's QueueShadow; its rows (ptr+i) % K are rewritten too (it stays valid)."""
    _need_cuda(keys_all, queue, queue_ptr)
    keys_all = _f32c(keys_all)
    n, C = keys_all.shape
    assert queue.dtype == torch.float32 and queue.is_contiguous() and queue.shape[0] == C
    assert queue_ptr.dtype == torch.int64 and queue_ptr.numel() == 1
    qbf = shadow.get(queue) if shadow is not None else None      # (re)built first if stale, then kept in step
    _call("cp2_enqueue", 1, _lib.load().cp2_enqueue, _ptr(keys_all), _ptr(queue), _ptr(qbf), _ptr(queue_ptr),
          _ptr(_persistent_zero("enqueue", 8, queue.device)), n, C, queue.shape[1], _stream())


# ------------------------------------------------- fused loss (a5 + a6 + a7 + a12)
def loss_combine(q_pos, k_pos, q_neg, k_neg, sq, sk, norms_q, z, G, lse, key_count, mask_a, temp_global,
                 temp_local, lmbd_dense, dense_n=None, hit_seg=None):
    """cp2_loss_combine: the [B,C]/[B]-sized remainder of the loss (1 launch).
    -> out [6] = (loss, loss_ins, loss_dense, std meter of q_pos, std meter of k_pos, acc_seg | NaN), l_pos [B],
       alpha [B], v1 [B,C], v2 [B,C]|None (inputs of cp2_dense_grad for upstream gradient 1).
    dense_n [B] (PROPOSED modes): per-image dense losses from dense_weighted_lse; sq, sk, lse are then unused."""
    B, C = q_pos.shape
    P = mask_a.shape[1]
    dev = q_pos.device
    out = torch.empty((6,), device=dev, dtype=torch.float32)
    l_pos = torch.empty((B,), device=dev, dtype=torch.float32)
    alpha = torch.empty((B,), device=dev, dtype=torch.float32)
    v1 = torch.empty((B, C), device=dev, dtype=torch.float32)
    v2 = torch.empty((B, C), device=dev, dtype=torch.float32) if q_neg is not None else None
    scratch = torch.empty((B, 2), device=dev, dtype=torch.float32)
    _call("cp2_loss_combine", 1, _lib.load().cp2_loss_combine, _ptr(q_pos), _ptr(k_pos), _ptr(q_neg), _ptr(k_neg),
          _ptr(sq), _ptr(sk), _ptr(norms_q), _ptr(z), _ptr(G), _ptr(lse), _ptr(key_count), _ptr(mask_a), _ptr(dense_n),
          _ptr(hit_seg), _ptr(out), _ptr(l_pos), _ptr(alpha), _ptr(v1), _ptr(v2), _ptr(scratch),
          _ptr(_persistent_zero("combine", 8, dev)), 1.0 / float(temp_global),
          1.0 / float(temp_local), float(lmbd_dense), B, C, P, _stream())
    return out, l_pos, alpha, v1, v2


class CP2LossFunction(torch.autograd.Function):
    """Dense + global CP2 loss from the two encoder outputs (builder.py:1260-1268, 1279-1292,
    1392-1397, 1414-1439), MappingType.CP2 / NegativeType.NONE.

    forward(q [B,C,h,w], k [B,C,h,w], mask_a [B,P], mask_b [B,P], queue [C,K], ...) ->
        (loss, loss_ins, loss_dense, k_pos, q_pos, l_pos, meters [2], q_neg, k_neg, acc_ins [1])
    Only `loss` carries a gradient, and only to q (keys and queue are detached in the
    reference, builder.py:1271,1396).  Forward = cp2_prep_features2, cp2_global_nce, cp2_dense_lse,
    cp2_loss_combine (4 launches); backward = cp2_dense_grad alone, reading the upstream gradient from device memory.
    Maps that fit TMEM (P <= 256): cp2_dense_small replaces cp2_dense_lse and already holds the P^2 part of the
    gradient, the backward is the elementwise cp2_dense_small_bwd.
    No ATen arithmetic in between: torch only allocates the buffers.
    """

    @staticmethod
    def forward(ctx, q, k, mask_a, mask_b, queue, temp_global, temp_local, lmbd_dense, include_background,
                queue_shadow, track_accuracy):
        B, C = q.shape[0], q.shape[1]
        qf = _f32c(q.detach()).reshape(B, C, -1)
        kf = _f32c(k.detach()).reshape(B, C, -1)
        mask_a, mask_b = _f32c(mask_a), _f32c(mask_b)
        tg, tl = float(temp_global), float(temp_local)
        ctx.small = dense_small_preferred(B, C, qf.shape[2], qf.device)
        seg_on = bool(track_accuracy) and ctx.small
        # one launch for both sides; key side compacted: fg keys first, bg keys after
        (qh, q_inv, sq, _, _, q_pos, q_neg, norms_q), (kh, k_inv, sk, _, kcnt, k_pos, k_neg, _, *perm) = \
            prep_features_pair(qf, mask_a, kf, mask_b, want_perm=seg_on)
        bg = bool(include_background)
        # one pass over the queue: fwd + bwd sums (+ the acc_ins meter)
        z, G, _, acc_ins = global_nce(q_pos, queue, 1.0 / tg, k_pos if track_accuracy else None,
                                      q_neg if (bg and track_accuracy) else None,
                                      k_neg if (bg and track_accuracy) else None, shadow=queue_shadow)
        hit_seg = None
        if ctx.small:                                          # map fits TMEM: LSE and the P^2 part of dq in one launch
            if seg_on:                                         # ... and the acc_seg meter (builder.py:1442-1448)
                lse, acc, hit_seg = dense_small(qh, kh, 1.0 / tl, kcnt,
                                                seg=(qf, kf, q_inv, k_inv, perm[0], mask_a, mask_b))
            else:
                lse, acc = dense_small(qh, kh, 1.0 / tl, kcnt)
        else:
            lse, acc = dense_lse(qh, kh, 1.0 / tl, kcnt), None  # column LSE on the tensor cores
        out, l_pos, alpha, v1, v2 = loss_combine(q_pos, k_pos, q_neg if bg else None, k_neg if bg else None,
                                                 sq, sk, norms_q, z, G, lse, kcnt, mask_a, tg, tl, lmbd_dense,
                                                 hit_seg=hit_seg)
        ctx.save_for_backward(qh, acc if ctx.small else kh, lse, mask_a, kcnt, q_inv, alpha, v1, v2)
        ctx.inv_tl = 1.0 / tl
        ctx.q_shape = q.shape
        loss, loss_ins, loss_dense, meters = out[0], out[1], out[2], out[3:]     # meters = (std q, std k, acc_seg | NaN)
        if acc_ins is None:
            acc_ins = torch.full((1,), float("nan"), device=q.device)
        ctx.mark_non_differentiable(loss_ins, loss_dense, k_pos, q_pos, l_pos, meters, q_neg, k_neg, acc_ins)
        return loss, loss_ins, loss_dense, k_pos, q_pos, l_pos, meters, q_neg, k_neg, acc_ins

    @staticmethod
    def backward(ctx, g_loss, *_unused):
        qh, kh, lse, mask_a, kcnt, q_inv, alpha, v1, v2 = ctx.saved_tensors
        g = g_loss.detach().reshape(1)
        if g.dtype != torch.float32 or not g.is_contiguous():
            g = g.float().contiguous()
        if ctx.small:                                          # kh holds acc: nothing is recomputed
            dq = dense_small_bwd(kh, qh, mask_a, q_inv, alpha, v1, v2, key_count=kcnt, gscale=g)
        else:
            dq = dense_grad(qh, kh, lse, mask_a, None, q_inv, alpha, v1, v2, ctx.inv_tl, key_count=kcnt, gscale=g)
        return (dq.reshape(ctx.q_shape),) + (None,) * 10


# ------------------------------------------------- f3: PROPOSED-mode dense loss
NEGATIVE_CODES = {"NONE": 0, "FIXED": 1, "AVERAGE": 2, "MEDIAN": 3, "HARD": 0}   # HARD: a no-op in the reference (Appendix A.8)


class PairWeights:
    """The id maps and scalars that define w[x,y] and the negative squash (builder.py:1204-1243, 1332-1386).
    ids: four int64 maps [B,H,W] (full resolution, sub-sampled inside the kernels at `output_stride`)."""

    def __init__(self, pixel_ids_a, pixel_ids_b, region_ids_a, region_ids_b, output_stride, w_pixel, w_region,
                 w_not, negative_type="NONE", negative_scale=2.0):
        name = getattr(negative_type, "name", negative_type)
        if name not in NEGATIVE_CODES:
            raise NotImplementedError(f"negative_type {name}")    # the reference's own fallthrough, builder.py:1386
        self.ids = []
        for t in (pixel_ids_a, pixel_ids_b, region_ids_a, region_ids_b):
            _need_cuda(t)
            if t.dtype != torch.int64 or t.dim() != 3:
                raise TypeError("id maps must be int64 [B,H,W]")
            self.ids.append(t.contiguous())
        self.h, self.w = self.ids[0].shape[1:]
        self.os = int(output_stride)
        self.scalars = (float(w_pixel), float(w_region), float(w_not), NEGATIVE_CODES[name], float(negative_scale))

    def _args(self):
        return [_ptr(t) for t in self.ids]

    def _tail(self, inv_tau, B, C, P):
        wp, wr, wn, code, scale = self.scalars
        return (wp, wr, wn, code, scale, float(inv_tau), B, C, P, self.h, self.w, self.os, _stream())


def dense_weighted_lse(qhat_t, khat_t, mask_a, mask_b, pw, pos_q, neg_q, pos_k, neg_k, inv_tau):
    """cp2_dense_weighted_lse -> (lse [B,P], dense_n [B], wstats [B,3], key_count [B] int32, acc_seg [1])."""
    B, P, C = qhat_t.shape
    lib, dev = _lib.load(), qhat_t.device
    lse = torch.empty((B, P), device=dev, dtype=torch.float32)
    dense_n = torch.empty((B,), device=dev, dtype=torch.float32)
    wstats = torch.empty((B, 3), device=dev, dtype=torch.float32)
    acc_seg = torch.empty((1,), device=dev, dtype=torch.float32)
    count = torch.empty((B,), device=dev, dtype=torch.int32)
    scratch = torch.empty((B, lib.cp2_dense_weighted_tiles(P)), device=dev, dtype=torch.float32)
    _call("cp2_dense_weighted_lse", 3, lib.cp2_dense_weighted_lse, _ptr(qhat_t), _ptr(khat_t), _ptr(mask_a),
          _ptr(mask_b), *pw._args(), _ptr(pos_q), _ptr(neg_q), _ptr(pos_k), _ptr(neg_k), _ptr(lse), _ptr(dense_n),
          _ptr(acc_seg), _ptr(wstats), _ptr(count), _ptr(scratch), *pw._tail(inv_tau, B, C, P))
    return lse, dense_n, wstats, count, acc_seg


def dense_weighted_small(qhat_t, khat_t, mask_a, mask_b, pw, pos_q, neg_q, pos_k, neg_k, inv_tau):
    """cp2_dense_weighted_small (maps that fit TMEM) -> (lse [B,P], dense_n [B], wstats [B,3], key_count [B] int32,
    hit [B], acc [B,C,P]): the PROPOSED-mode dense loss on the tensor cores; acc feeds dense_small_bwd."""
    B, P, C = qhat_t.shape
    lib, dev = _lib.load(), qhat_t.device
    f32 = dict(device=dev, dtype=torch.float32)
    lse, dense_n, wstats = torch.empty((B, P), **f32), torch.empty((B,), **f32), torch.empty((B, 3), **f32)
    hit, acc = torch.empty((B,), **f32), torch.empty((B, C, P), **f32)
    count = torch.empty((B,), device=dev, dtype=torch.int32)
    scratch = torch.empty((B, lib.cp2_dense_weighted_tiles(P)), **f32)
    _call("cp2_dense_weighted_small", 2, lib.cp2_dense_weighted_small, _ptr(qhat_t), _ptr(khat_t), _ptr(mask_a),
          _ptr(mask_b), *pw._args(), _ptr(pos_q), _ptr(neg_q), _ptr(pos_k), _ptr(neg_k), _ptr(lse), _ptr(dense_n),
          _ptr(hit), _ptr(wstats), _ptr(count), _ptr(scratch), _ptr(acc), *pw._tail(inv_tau, B, C, P))
    return lse, dense_n, wstats, count, hit, acc


def dense_weighted_stream_ok(C, P, inv_tau):
    """Shapes the streaming tensor-core pair of the PROPOSED modes takes (any map size; used when the single-kernel
    path does not apply: maps above 256 pixels, batches beyond one wave)."""
    return C == 128 and P >= 1 and 0.0 < inv_tau <= MAX_INV_TAU


def dense_weighted_stream(qhat_t, khat_t, mask_a, mask_b, pw, pos_q, neg_q, pos_k, neg_k, inv_tau):
    """cp2_dense_weighted_stream -> (lse [B,P], dense_n [B], wstats [B,3], key_count [B] int32, acc_seg [1],
    ids_pad, b_pad): cp2_dense_weighted_lse's results on the tensor cores + what the streaming backward needs."""
    B, P, C = qhat_t.shape
    lib, dev = _lib.load(), qhat_t.device
    f32 = dict(device=dev, dtype=torch.float32)
    lse, dense_n, wstats = torch.empty((B, P), **f32), torch.empty((B,), **f32), torch.empty((B, 3), **f32)
    acc_seg = torch.empty((1,), **f32)
    count = torch.empty((B,), device=dev, dtype=torch.int32)
    scratch = torch.empty((B, lib.cp2_dense_weighted_tiles(P)), **f32)
    ids_pad = torch.empty((lib.cp2_dense_weighted_stream_ids_bytes(B, P) // 8,), device=dev, dtype=torch.int64)
    b_pad = torch.empty((B, lib.cp2_dense_pad(P)), **f32)
    _call("cp2_dense_weighted_stream", 4, lib.cp2_dense_weighted_stream, _ptr(qhat_t), _ptr(khat_t), _ptr(mask_a),
          _ptr(mask_b), *pw._args(), _ptr(pos_q), _ptr(neg_q), _ptr(pos_k), _ptr(neg_k), _ptr(lse), _ptr(dense_n),
          _ptr(acc_seg), _ptr(wstats), _ptr(count), _ptr(scratch), _ptr(ids_pad), _ptr(b_pad),
          *pw._tail(inv_tau, B, C, P))
    return lse, dense_n, wstats, count, acc_seg, ids_pad, b_pad


def dense_weighted_stream_grad(qhat_t, khat_t, mask_a, mask_b, pw, lse, wstats, ids_pad, b_pad, inv_norm, alpha, v1, v2,
                               inv_tau, gscale=None):
    """cp2_dense_weighted_stream_grad -> dq [B,C,P] (the weighted backward on the dense backward kernel)."""
    B, P, C = qhat_t.shape
    lib = _lib.load()
    dq = torch.empty((B, C, P), device=qhat_t.device, dtype=torch.float32)
    nl = torch.empty((B, lib.cp2_dense_pad(P)), device=qhat_t.device, dtype=torch.float32)
    wp, wr, wn, code, scale = pw.scalars
    _call("cp2_dense_weighted_stream_grad", 2, lib.cp2_dense_weighted_stream_grad, _ptr(qhat_t), _ptr(khat_t),
          _ptr(mask_a), _ptr(mask_b), _ptr(lse), _ptr(ids_pad), _ptr(b_pad), _ptr(wstats), _ptr(inv_norm), _ptr(alpha),
          _ptr(v1), _ptr(v2), _ptr(gscale), _ptr(dq), _ptr(nl), wp, wr, wn, code, scale, float(inv_tau), B, C, P,
          _stream())
    return dq


def dense_weighted_grad(qhat_t, khat_t, mask_a, mask_b, pw, lse, wstats, inv_norm, alpha, v1, v2, inv_tau, gscale=None):
    """cp2_dense_weighted_grad -> dq [B,C,P]."""
    B, P, C = qhat_t.shape
    dq = torch.empty((B, C, P), device=qhat_t.device, dtype=torch.float32)
    _call("cp2_dense_weighted_grad", 1, _lib.load().cp2_dense_weighted_grad, _ptr(qhat_t), _ptr(khat_t), _ptr(mask_a),
          _ptr(mask_b), *pw._args(), _ptr(lse), _ptr(wstats), _ptr(inv_norm), _ptr(alpha), _ptr(v1), _ptr(v2),
          _ptr(gscale), _ptr(dq), *pw._tail(inv_tau, B, C, P))
    return dq


class CP2WeightedLossFunction(torch.autograd.Function):
    """CP2LossFunction for the PROPOSED modes: same outputs, the dense loss with
    pair weights and the negative squash.  Forward = cp2_prep_features x 2 (keys not compacted),
    cp2_global_nce, cp2_dense_weighted_lse, cp2_loss_combine; backward = cp2_dense_weighted_grad.  Maps that fit
    TMEM: cp2_dense_weighted_small (tcgen05) in the forward, cp2_dense_small_bwd in the backward."""

    @staticmethod
    def forward(ctx, q, k, mask_a, mask_b, queue, temp_global, temp_local, lmbd_dense, include_background, pw,
                queue_shadow, track_accuracy):
        B, C = q.shape[0], q.shape[1]
        qf = _f32c(q.detach()).reshape(B, C, -1)
        kf = _f32c(k.detach()).reshape(B, C, -1)
        mask_a, mask_b = _f32c(mask_a), _f32c(mask_b)
        tg, tl = float(temp_global), float(temp_local)
        qh, q_inv, sq, nsq, _, q_pos, q_neg, norms_q = prep_features(qf, mask_a, unit=True)
        kh, _, sk, nsk, _, k_pos, k_neg, _ = prep_features(kf, mask_b, unit=True)
        bg = bool(include_background)
        z, G, _, acc_ins = global_nce(q_pos, queue, 1.0 / tg, k_pos if track_accuracy else None,
                                      q_neg if (bg and track_accuracy) else None,
                                      k_neg if (bg and track_accuracy) else None, shadow=queue_shadow)
        ctx.small = dense_small_preferred(B, C, qh.shape[1], qh.device)
        ctx.stream_extra = None
        if ctx.small:                    # map fits TMEM: the tensor-core kernel, which also leaves the P^2 part of dq
            lse, dense_n, wstats, kcnt, hit, acc = dense_weighted_small(qh, kh, mask_a, mask_b, pw, sq, nsq, sk, nsk,
                                                                        1.0 / tl)
            acc_seg = None
        elif dense_weighted_stream_ok(C, qh.shape[1], 1.0 / tl):      # larger maps: the streaming tensor-core pair
            lse, dense_n, wstats, kcnt, acc_seg, ids_pad, b_pad = dense_weighted_stream(qh, kh, mask_a, mask_b, pw, sq,
                                                                                        nsq, sk, nsk, 1.0 / tl)
            hit = acc = None
            ctx.stream_extra = (ids_pad, b_pad)
        else:
            lse, dense_n, wstats, kcnt, acc_seg = dense_weighted_lse(qh, kh, mask_a, mask_b, pw, sq, nsq, sk, nsk,
                                                                     1.0 / tl)
            hit = acc = None
        out, l_pos, alpha, v1, v2 = loss_combine(q_pos, k_pos, q_neg if bg else None, k_neg if bg else None,
                                                 None, None, norms_q, z, G, None, kcnt, mask_a, tg, tl, lmbd_dense,
                                                 dense_n=dense_n, hit_seg=hit)
        if ctx.small:
            acc_seg = out[5:6]
        ctx.save_for_backward(qh, acc if ctx.small else kh, lse, mask_a, mask_b, wstats, q_inv, alpha, v1, v2, kcnt)
        ctx.inv_tl, ctx.pw, ctx.q_shape = 1.0 / tl, pw, q.shape
        loss, loss_ins, loss_dense, meters = out[0], out[1], out[2], out[3:]
        if acc_ins is None:
            acc_ins = torch.full((1,), float("nan"), device=q.device)
        ctx.mark_non_differentiable(loss_ins, loss_dense, k_pos, q_pos, l_pos, meters, q_neg, k_neg, acc_ins, acc_seg)
        return loss, loss_ins, loss_dense, k_pos, q_pos, l_pos, meters, q_neg, k_neg, acc_ins, acc_seg

    @staticmethod
    def backward(ctx, g_loss, *_unused):
        qh, kh, lse, mask_a, mask_b, wstats, q_inv, alpha, v1, v2, kcnt = ctx.saved_tensors
        g = g_loss.detach().reshape(1)
        if g.dtype != torch.float32 or not g.is_contiguous():
            g = g.float().contiguous()
        if ctx.small:                    # kh holds acc
            dq = dense_small_bwd(kh, qh, mask_a, q_inv, alpha, v1, v2, key_count=kcnt, gscale=g)
        elif ctx.stream_extra is not None:
            dq = dense_weighted_stream_grad(qh, kh, mask_a, mask_b, ctx.pw, lse, wstats, *ctx.stream_extra, q_inv, alpha,
                                            v1, v2, ctx.inv_tl, gscale=g)
        else:
            dq = dense_weighted_grad(qh, kh, mask_a, mask_b, ctx.pw, lse, wstats, q_inv, alpha, v1, v2, ctx.inv_tl,
                                     gscale=g)
        return (dq.reshape(ctx.q_shape),) + (None,) * 11


def cp2_loss(q, k, mask_a, mask_b, queue, temp_global=0.2, temp_local=1.0, lmbd_dense=0.2,
             include_background=False, pair_weights=None, queue_shadow=None, track_accuracy=False):
    """-> (loss, loss_ins, loss_dense, k_pos, q_pos, stats); only `loss` has a gradient (to q).
    pair_weights: None (MappingType.CP2 / NegativeType.NONE, the tensor-core path) or a PairWeights.
    queue_shadow: the queue's QueueShadow (None: the queue is converted to bf16 for this call).
    track_accuracy: stats["acc_ins"] from the same queue pass (builder.py:1441); stats["acc_seg"] from the dense
    kernel itself when the single-kernel path is taken (map fits TMEM, batch fits one wave: dense_small_preferred)
    or the weights are on, else None (ops.dense_top1 computes it)."""
    acc_seg = None
    if pair_weights is None:
        res = CP2LossFunction.apply(q, k, mask_a, mask_b, queue, temp_global, temp_local, lmbd_dense,
                                    include_background, queue_shadow, bool(track_accuracy))
        if track_accuracy and dense_small_preferred(q.shape[0], q.shape[1], q[0, 0].numel(), q.device):
            acc_seg = res[6][2]                      # came out of cp2_dense_small / cp2_loss_combine
    else:                    # the weighted forward finds the arg-max of its own logits on the way
        *res, acc_seg = CP2WeightedLossFunction.apply(q, k, mask_a, mask_b, queue, temp_global, temp_local,
                                                      lmbd_dense, include_background, pair_weights, queue_shadow,
                                                      bool(track_accuracy))
        acc_seg = acc_seg[0]
    loss, loss_ins, loss_dense, k_pos, q_pos, l_pos, meters, q_neg, k_neg, acc_ins = res
    stats = {"l_pos": l_pos, "acc_ins": acc_ins[0], "acc_seg": acc_seg, "std_q_pos": meters[0],
             "std_k_pos": meters[1], "q_neg": q_neg, "k_neg": k_neg}
    return loss, loss_ins, loss_dense, k_pos, q_pos, stats


# ------------------------------------------------- f2: forward_densecl losses
def rows_nce(rows, pos_keys, z, G, inv_tau, want_grad=True):
    """cp2_rows_nce -> (loss [1], terms [N], dq [N,C] | None).  z [N] / G [N,C], or partial sums
    z [nz, stride >= N] / G [ng, N, C] from queue_sums_tc (added inside the kernel, fixed order)."""
    N, C = rows.shape
    dev = rows.device
    loss = torch.empty((1,), device=dev, dtype=torch.float32)
    terms = torch.empty((N,), device=dev, dtype=torch.float32)
    dq = torch.empty((N, C), device=dev, dtype=torch.float32) if want_grad else None
    nz, zs = (z.shape[0], z.stride(0)) if z.dim() == 2 else (1, 0)
    ng, gs = (G.shape[0], G.stride(0)) if G.dim() == 3 else (1, 0)
    _call("cp2_rows_nce", 2, _lib.load().cp2_rows_nce, _ptr(rows), _ptr(pos_keys), _ptr(z), _ptr(G), _ptr(loss),
          _ptr(terms), _ptr(dq), float(inv_tau), N, C, nz, zs, ng, gs, _stream())
    return loss, terms, dq


TC_ROWS_MIN = 1024      # from this many rows on, the queue sums run on the dense loss's tensor-core kernel
_NL_CACHE = {}


def _queue_nl(K, inv_tau, device):
    """nl operand of cp2_queue_sums: -log2(e)/tau per real queue entry, -80 for the padding up to a multiple of
    128.  Built once per (K, tau, device)."""
    key = (int(K), float(inv_tau), str(device))
    if key not in _NL_CACHE:
        kp = (K + 127) // 128 * 128
        nl = torch.full((kp,), -float(inv_tau) * 1.4426950408889634, device=device, dtype=torch.float32)
        nl[K:] = -80.0
        _NL_CACHE[key] = nl
    return _NL_CACHE[key]


@torch.no_grad()
def queue_sums_tc(rows_bf, queue, inv_tau, shadow=None):
    """Partial sums of z [N] = sum_k exp((l_k-1)/tau) and G [N,C] = sum_k exp((l_k-1)/tau) u_k (cp2_global_nce's
    definitions) for MANY rows in ONE launch of the dense loss's backward kernel in its rectangular mode
    (cp2_queue_sums): queries = rows_bf [N,C] bf16 unit rows, keys = the queue's bf16 shadow.
    -> (zp [2*split, pad128(N)], gacc [split, N, C]) for rows_nce; no ATen arithmetic."""
    N, C = rows_bf.shape
    K = queue.shape[1]
    dev = rows_bf.device
    lib = _lib.load()
    qbf = (shadow if shadow is not None else QueueShadow()).get(queue)
    split = lib.cp2_queue_sums_split(N, K)
    zp = torch.empty((2 * split, lib.cp2_dense_pad(N)), device=dev, dtype=torch.float32)
    gacc = torch.empty((split, N, C), device=dev, dtype=torch.float32)
    _call("cp2_queue_sums", 1, lib.cp2_queue_sums, _ptr(rows_bf), _ptr(qbf), _ptr(_queue_nl(K, inv_tau, dev)), _ptr(zp),
          _ptr(gacc), float(inv_tau), N, C, K, _stream())
    return zp, gacc


class RowsInfoNCEFunction(torch.autograd.Function):
    """`ContrastiveHead(pos, neg, temperature)` (builder.py:150-176) with pos = rows . pos_keys and
    neg = rows @ queue, never materialising the [N,1+K] logits: the queue sums come from cp2_global_nce (few
    rows: one pass over the queue per 32 rows) or from the dense loss's tensor-core backward kernel in its
    rectangular mode (`queue_sums_tc` / cp2_queue_sums, N >= TC_ROWS_MIN, one launch), then cp2_rows_nce.
    Gradient to `rows` only (keys and queue are detached in the reference, builder.py:726, 768, 872)."""

    @staticmethod
    def forward(ctx, rows, pos_keys, queue, temp, shadow, rows_bf=None):
        rows_c, pos_c = _f32c(rows.detach()), _f32c(pos_keys.detach())
        if rows_c.shape[0] >= TC_ROWS_MIN and rows_c.shape[1] == 128:
            if rows_bf is None:
                rows_bf = rows_c.to(torch.bfloat16)
            z, G = queue_sums_tc(rows_bf, queue, 1.0 / float(temp), shadow)
        else:
            z, G = global_nce(rows_c, queue, 1.0 / float(temp), shadow=shadow)[:2]
        loss, _, dq = rows_nce(rows_c, pos_c, z, G, 1.0 / float(temp), want_grad=True)
        ctx.save_for_backward(dq)
        return loss[0]

    @staticmethod
    def backward(ctx, g):
        (dq,) = ctx.saved_tensors
        return dq * g, None, None, None, None, None


def rows_infonce(rows, pos_keys, queue, temp, shadow=None, rows_bf=None):
    """rows, pos_keys [N,C] (rows unit vectors), queue [C,K] -> 0-d loss with a gradient to rows.
    rows_bf: the rows as a bf16 [N,C] tensor if the caller already has it (cp2_prep_features' operand)."""
    return RowsInfoNCEFunction.apply(rows, pos_keys, queue, temp, shadow, rows_bf)


@torch.no_grad()
def positive_keys(k_local, idx, pix_q=None, pix_k=None, output_stride=1, lmbd_coordinate=0.0):
    """builder.py:824-855 as vectors: k_local [B,C,P] unit columns, idx [B,P] -> kp [B*P,C]."""
    _need_cuda(k_local, idx)
    k_local = _f32c(k_local)
    B, C, P = k_local.shape
    idx = idx.contiguous()
    assert idx.dtype == torch.int64 and tuple(idx.shape) == (B, P)
    kp = torch.empty((B * P, C), device=k_local.device, dtype=torch.float32)
    h = w = 1
    if pix_q is not None:
        assert pix_q.dtype == torch.int64 and pix_k.dtype == torch.int64 and pix_q.shape == pix_k.shape
        pix_q, pix_k = pix_q.contiguous(), pix_k.contiguous()
        h, w = pix_q.shape[1:]
    _call("cp2_positive_keys", 1, _lib.load().cp2_positive_keys, _ptr(k_local), _ptr(idx), _ptr(pix_q), _ptr(pix_k),
          _ptr(kp), float(lmbd_coordinate), B, C, P, h, w, int(output_stride), _stream())
    return kp


@torch.no_grad()
def densecl_match(embd_q, embd_k):
    """builder.py:817-821: idx[n,x] = argmax_y embd_q[n,:,x] . embd_k[n,:,y] on the 2048-d backbone features.
    cp2_densecl_match: one tcgen05 kernel (fp32 operands split into bf16 hi + lo on chip, the [B,S^2,S^2]
    similarities stay in TMEM).  Maps above 256 pixels or channel counts off the 16-grid: torch.bmm + argmax."""
    _need_cuda(embd_q, embd_k)
    B, Cb, P = embd_q.shape
    if P > 256 or Cb % 16 != 0:
        return torch.bmm(embd_q.transpose(1, 2), embd_k).argmax(dim=2)
    embd_q, embd_k = _f32c(embd_q), _f32c(embd_k)
    lib = _lib.load()
    idx = torch.empty((B, P), device=embd_q.device, dtype=torch.int64)
    scratch = _persistent_zero("match", lib.cp2_densecl_match_scratch_bytes(B, P), embd_q.device)
    _call("cp2_densecl_match", 1, lib.cp2_densecl_match, _ptr(embd_q), _ptr(embd_k), _ptr(idx), _ptr(scratch), B, Cb, P,
          _stream())
    return idx


def densecl_loss(embd_q, embd_k, q_local, k_local, q_global, k_global, queue, queue2, temp_global=0.2,
                 temp_local=0.2, lmbd_dense=0.5, pix_q=None, pix_k=None, output_stride=1, lmbd_coordinate=0.0,
                 shadow=None, shadow2=None):
    """One direction of forward_densecl (builder.py:762-931) on the unit features get_query_features /
    get_key_features return: -> (loss, loss_global, loss_local), gradients to q_global and q_local."""
    B, C, P = q_local.shape
    loss_g = rows_infonce(q_global, k_global, queue, temp_global, shadow)
    idx = densecl_match(embd_q.detach(), embd_k.detach())
    kp = positive_keys(k_local.detach(), idx, pix_q, pix_k, output_stride, lmbd_coordinate)
    rows = q_local.permute(0, 2, 1).reshape(B * P, C)
    rows_bf = None
    if B * P >= TC_ROWS_MIN and C == 128:       # bf16 [B,P,C] operand of the tensor-core queue sums (mask unused)
        rows_bf = prep_features(q_local.detach(), _persistent_zero("densecl_mask", 4 * B * P, q_local.device)
                                .view(torch.float32)[: B * P].view(B, P))[0].view(B * P, C)
    loss_l = rows_infonce(rows, kp, queue2, temp_local, shadow2, rows_bf)
    return (1 - lmbd_dense) * loss_g + lmbd_dense * loss_l, loss_g, loss_l


# ------------------------------------------------------------- f4: accuracy meters
@torch.no_grad()
def dense_top1(q, k, mask_a, mask_b):
    """builder.py:1442-1448 -> (acc_seg 0-d, hit [B]).  q, k [B,C,P] or [B,C,h,w] f32 encoder outputs."""
    _need_cuda(q, k, mask_a, mask_b)
    B, C = q.shape[0], q.shape[1]
    q, k = _f32c(q.detach()).reshape(B, C, -1), _f32c(k.detach()).reshape(B, C, -1)
    mask_a, mask_b = _f32c(mask_a), _f32c(mask_b)
    P = q.shape[2]
    lib, dev = _lib.load(), q.device
    hit = torch.empty((B,), device=dev, dtype=torch.float32)
    acc = torch.empty((1,), device=dev, dtype=torch.float32)
    scratch = torch.empty((lib.cp2_dense_top1_scratch_floats(B, P),), device=dev, dtype=torch.float32)
    _call("cp2_dense_top1", 2, lib.cp2_dense_top1, _ptr(q), _ptr(k), _ptr(mask_a), _ptr(mask_b), _ptr(hit), _ptr(acc),
          _ptr(scratch), B, C, P, _stream())
    return acc[0], hit


@torch.no_grad()
def global_top1(q_pos, k_pos, queue, q_neg=None, k_neg=None, shadow=None):
    """builder.py:1441 (top-1 of `_accuracy`) -> (acc_ins 0-d, hit [B]): the meter part of cp2_global_nce on its
    own (the training step gets it from the loss's queue pass).  Call BEFORE the enqueue."""
    _, _, hit, acc = global_nce(q_pos, queue, 5.0, k_pos, q_neg, k_neg, shadow=shadow)
    return acc[0], hit


# ------------------------------------------------------------- a2 + a4
@torch.no_grad()
def id_iou(ids_a, ids_b, mask_a_ds, mask_b_ds, output_stride):
    """tools/correlation_mapping.py:103-138 on the `[OS/2::OS]` subsample of two full-resolution id maps
    (builder.py:1165-1186, 1203-1217).  ids [B,H,W] int64, masks [B,hs,ws] f32 -> (iou [B], iou_masked [B])."""
    _need_cuda(ids_a, ids_b, mask_a_ds, mask_b_ds)
    ids_a, ids_b = _labels_i64(ids_a), _labels_i64(ids_b)
    mask_a_ds, mask_b_ds = _f32c(mask_a_ds), _f32c(mask_b_ds)
    B, H, W = ids_a.shape
    assert ids_b.shape == ids_a.shape and mask_a_ds.shape[0] == B and mask_a_ds.shape == mask_b_ds.shape
    iou = torch.empty((B,), device=ids_a.device, dtype=torch.float32)
    iou_m = torch.empty((B,), device=ids_a.device, dtype=torch.float32)
    _call("cp2_id_iou", 1, _lib.load().cp2_id_iou, _ptr(ids_a), _ptr(ids_b), _ptr(mask_a_ds), _ptr(mask_b_ds),
          _ptr(iou), _ptr(iou_m), B, H, W, int(output_stride), _stream())
    return iou, iou_m


# ------------------------------------------------ mirror variant (a13 / f5)
def _labels_i64(labels):
    if labels.dtype != torch.int64:
        labels = labels.long()
    return labels if labels.is_contiguous() else labels.contiguous()


def mirror_loss_raw(s_logits, t_logits, labels, lmbd_compare, softmax_temp, want_grads, gscale=None, n_valid=None):
    """One launch of cp2_mirror_loss on full-resolution logits [B,Cls,*spatial] (t_logits may be None:
    MirrorVariant.NONE).  -> (out [4] = loss, class, compare, n_valid_counted; ds|None; dt|None)."""
    _need_cuda(s_logits, t_logits, labels)
    s = _f32c(s_logits)
    t = None if t_logits is None else _f32c(t_logits)
    labels = _labels_i64(labels)
    B, Cls = s.shape[0], s.shape[1]
    HW = s[0, 0].numel()
    assert labels.numel() == B * HW and (t is None or t.shape == s.shape), f"{s.shape = } {labels.shape = }"
    lib = _lib.load()
    dev = s.device
    partial = torch.empty((lib.cp2_mirror_partials(B, HW), 3), device=dev, dtype=torch.float32)
    out = torch.empty((4,), device=dev, dtype=torch.float32)
    ds = torch.empty_like(s) if want_grads else None
    dt = torch.empty_like(s) if (want_grads and t is not None) else None
    _call("cp2_mirror_loss", 2, lib.cp2_mirror_loss, _ptr(s), _ptr(t), _ptr(labels), _ptr(partial), _ptr(out),
          _ptr(ds), _ptr(dt), _ptr(gscale), float(lmbd_compare), 1.0 / float(softmax_temp),
          int(n_valid) if n_valid else 0, B, Cls, HW, _stream())
    return out, ds, dt


def mirror_loss_upsampled_raw(s_lo, t_lo, labels, lmbd_compare, softmax_temp, want_grads, n_valid=None):
    """One launch of cp2_mirror_loss_upsampled: decoder-resolution logits [B,Cls,h,w] against labels
    [B,H,W]; the bilinear resize (segment_network.py:219-231) happens on chip.
    -> (out [4]; ds_lo|None; dt_lo|None) for upstream gradient 1."""
    _need_cuda(s_lo, t_lo, labels)
    s = _f32c(s_lo)
    t = None if t_lo is None else _f32c(t_lo)
    labels = _labels_i64(labels)
    B, Cls, h, w = s.shape
    _, H, W = labels.shape
    assert labels.shape[0] == B and (t is None or t.shape == s.shape), f"{s.shape = } {labels.shape = }"
    lib = _lib.load()
    dev = s.device
    partial = torch.empty((lib.cp2_mirror_up_partials(B, H, W), 3), device=dev, dtype=torch.float32)
    out = torch.empty((4,), device=dev, dtype=torch.float32)
    ds = dt = None
    if want_grads and t is not None:
        both = torch.empty((2,) + tuple(s.shape), device=dev, dtype=torch.float32)   # adjacent: ONE memset
        ds, dt = both[0], both[1]
    elif want_grads:
        ds = torch.empty_like(s)
    n_launch = 2 + (1 if ds is not None else 0)                                    # + memset of the grads
    _call("cp2_mirror_loss_upsampled", n_launch, lib.cp2_mirror_loss_upsampled, _ptr(s), _ptr(t), _ptr(labels),
          _ptr(partial), _ptr(out), _ptr(ds), _ptr(dt), float(lmbd_compare), 1.0 / float(softmax_temp),
          int(n_valid) if n_valid else 0, B, Cls, h, w, H, W, _stream())
    return out, ds, dt


class MirrorLossFunction(torch.autograd.Function):
    """loss = class + lmbd * compare on full-resolution logits (mirror_network.py:47-63).
    Forward = one loss-only pass; backward = one pass that recomputes the per-pixel terms and writes
    both gradients already multiplied by the upstream gradient (read from device memory)."""

    @staticmethod
    def forward(ctx, s_logits, t_logits, labels, lmbd_compare, softmax_temp, n_valid):
        out, _, _ = mirror_loss_raw(s_logits.detach(), None if t_logits is None else t_logits.detach(), labels,
                                    lmbd_compare, softmax_temp, want_grads=False, n_valid=n_valid)
        ctx.save_for_backward(s_logits, t_logits, labels)
        ctx.cfg = (lmbd_compare, softmax_temp, n_valid)
        aux = out[1:]
        ctx.mark_non_differentiable(aux)
        return out[0], aux

    @staticmethod
    def backward(ctx, g_loss, _g_aux):
        s_logits, t_logits, labels = ctx.saved_tensors
        lmbd_compare, softmax_temp, n_valid = ctx.cfg
        g = g_loss.detach().reshape(1).float().contiguous()
        _, ds, dt = mirror_loss_raw(s_logits.detach(), None if t_logits is None else t_logits.detach(), labels,
                                    lmbd_compare, softmax_temp, want_grads=True, gscale=g, n_valid=n_valid)
        return (ds.reshape(s_logits.shape), None if dt is None else dt.reshape(t_logits.shape),
                None, None, None, None)


class MirrorLossUpsampledFunction(torch.autograd.Function):
    """The same loss taken directly from the decoder's low-resolution logits: forward computes the
    loss AND both low-resolution gradients in one pass over the labels; backward only scales the two
    [B,Cls,h,w] gradients by the upstream gradient."""

    @staticmethod
    def forward(ctx, s_lo, t_lo, labels, lmbd_compare, softmax_temp, n_valid):
        need = s_lo.requires_grad or (t_lo is not None and t_lo.requires_grad)
        out, ds, dt = mirror_loss_upsampled_raw(s_lo.detach(), None if t_lo is None else t_lo.detach(), labels,
                                                lmbd_compare, softmax_temp, want_grads=need, n_valid=n_valid)
        ctx.save_for_backward(ds, dt)
        aux = out[1:]
        ctx.mark_non_differentiable(aux)
        return out[0], aux

    @staticmethod
    def backward(ctx, g_loss, _g_aux):
        ds, dt = ctx.saved_tensors
        return (ds * g_loss, None if dt is None else dt * g_loss, None, None, None, None)


def mirror_loss(s_logits, t_logits, labels, lmbd_compare=0.01, softmax_temp=2.0, n_valid=None):
    """-> (loss with autograd to s/t logits, aux [3] = class, compare, valid pixel count)."""
    return MirrorLossFunction.apply(s_logits, t_logits, labels, lmbd_compare, softmax_temp, n_valid)


def mirror_loss_upsampled(s_lo, t_lo, labels, lmbd_compare=0.01, softmax_temp=2.0, n_valid=None):
    """-> (loss with autograd to the low-resolution logits, aux [3])."""
    return MirrorLossUpsampledFunction.apply(s_lo, t_lo, labels, lmbd_compare, softmax_temp, n_valid)
