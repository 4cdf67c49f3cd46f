/* cp2_b200 — C ABI of the B200-native CP2 pre-training loss hot path.
 *
 * One shared library (cp2_b200/csrc/libcp2_b200.so, sm_100a only) exports exactly the
 * entry points below.  Each replaces a span of the reference's Python
 * (kimathikaai/CP2; citations are file:line under /root/reference) that today runs as
 * a string of ATen/cuBLAS/NCCL library calls.
 *
 * Conventions (all entry points)
 *   - plain pointers and sizes only; every pointer is a DEVICE pointer unless its
 *     name ends in _host; the caller (torch) owns every buffer, nothing is
 *     allocated, freed or retained across calls;
 *   - all work is enqueued on `stream` (a cudaStream_t passed as void*); no
 *     entry point synchronises the device or the stream;
 *   - returns 0 on success, a negative CP2_E_* code otherwise and never throws;
 *     cp2_last_error() then holds a thread-local human-readable message;
 *   - re-entrant across streams and host threads; no global mutable device state.
 */
#ifndef CP2_B200_H_
#define CP2_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CP2_ABI_VERSION 10

#define CP2_OK 0
#define CP2_E_INVALID (-1)      /* bad shape / null pointer / unsupported dim          */
#define CP2_E_CUDA (-2)         /* a CUDA runtime/driver call failed (see last_error)  */
#define CP2_E_UNSUPPORTED (-3)  /* device is not sm_100 or driver lacks TMA encode     */

int cp2_abi_version(void);
const char *cp2_last_error(void);

/* K1 — copy-paste compositing of one view.
 * Replaces builder.py:1146-1152 (mask from channel 0 of the background, img*mask+bg)
 * and the mask part of the stride subsample builder.py:1155-1159 / 1173-1177.
 *   img, bg, out : [B,3,H,W] f32 NCHW (out may alias img)
 *   mask_full    : [B,H,W] f32 or NULL (only the visualisation path needs it)
 *   mask_ds      : [B,hs,ws] f32, hs = len(range(os/2,H,os)), ws likewise
 * Bit-exact with the reference. */
int cp2_composite(const float *img, const float *bg, float *out, float *mask_full,
                  float *mask_ds, int B, int H, int W, int os, void *stream);
/* Both views of one step in one launch (same contract per view as cp2_composite). */
int cp2_composite2(const float *img_a, const float *bg_a, float *out_a, float *mask_full_a, float *mask_ds_a,
                   const float *img_b, const float *bg_b, float *out_b, float *mask_full_b, float *mask_ds_b,
                   int B, int H, int W, int os, void *stream);

/* K2a — per-pixel L2 normalisation + masked pooling + bf16 transpose (+ key compaction).
 * Replaces builder.py:1261-1268 (query) / 1279-1285 (key): F.normalize(dim=1) and the
 * two `ncx,nx->nc` einsums (the outer F.normalize of the pooled vectors is left to
 * the caller: it is [B,C]).
 *   feat      : [B,C,P] f32 (encoder output, NCHW flattened)
 *   mask      : [B,P] f32 in {0,1}
 *   fhat_t    : [B,P,C] bf16, pixel-major unit vectors (operand layout of K2)
 *   inv_norm  : [B,P] f32, 1/max(||feat_x||, 1e-12)
 *   pos_sum   : [B,C] f32, sum_x mask_x * fhat_x      (fp32 values, not the bf16 copy)
 *   neg_sum   : [B,C] f32, sum_x (1-mask_x) * fhat_x
 *   scratch   : [B, cp2_prep_tiles(P), 2, C] f32 work space (per-tile partial sums)
 *   key_count : NULL, or [B] int32 OUT.  When given, fhat_t is written COMPACTED (a stable partition): the
 *               unit vectors of the foreground pixels (mask != 0) first, in pixel order, then those of the
 *               background pixels, in pixel order; key_count[n] = number of foreground pixels.  Only the
 *               key side is compacted: background keys have label 0 for every query
 *               (builder.py:1291-1292,1435), so they enter neither the loss nor the gradient and K2 skips
 *               the rows past key_count (they are still there for the arg-max meter, builder.py:1442).
 *               inv_norm / pos_sum / neg_sum are unaffected (pixel order).
 *   perm      : NULL, or [B,P] int32 OUT (needs key_count): perm[n,row] = the pixel stored in row `row`.
 *   pos_unit, neg_unit, norms : all NULL, or OUT [B,C], [B,C], [B,2]: the outer F.normalize of the pooled
 *               vectors (builder.py:1264,1268,1282,1285), x / max(||x||, 1e-12), and the raw norms
 *               (||pos_sum||, ||neg_sum||) the normalisation's backward needs.
 *   scratch_counters : B uint32, PERSISTENT (zero before the first call, left zero): done counters — the
 *               last tile CTA of an image reduces the image's per-tile partial sums in a fixed order
 *               (one launch, deterministic).
 * C must be a multiple of 64 and <= 256. */
int cp2_prep_tiles(int P);
int cp2_prep_features(const float *feat, const float *mask, void *fhat_t, float *inv_norm,
                      float *pos_sum, float *neg_sum, float *scratch, int32_t *key_count, int32_t *perm,
                      float *pos_unit, float *neg_unit, float *norms, void *scratch_counters,
                      int B, int C, int P, void *stream);
/* cp2_prep_features without caller-owned work space, for callers that cannot allocate (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc).  It has NO counterpart in the reference.
 * Same inputs, outputs and results as cp2_prep_features; the
 * `scratch` / `scratch_counters` arguments are gone: both live in one file-scope `static __device__` array of
 * CP2_PREP_STATIC_BYTES (8 MiB) that every GPU which loads the library holds until the process exits.  The call
 * allocates nothing (no cudaMalloc / cudaFree, no allocating library routine): one cudaMemsetAsync of the done
 * counters and one kernel launch on `stream`.
 * Limits (CP2_E_INVALID with a message beyond them): B <= CP2_PREP_STATIC_MAX_B (4096) and
 *   B * cp2_prep_tiles(P) * 2 * C * 4 <= CP2_PREP_STATIC_BYTES - 4 * CP2_PREP_STATIC_MAX_B   (8 372 224 bytes;
 *   configs[1], B=32, P=196, C=128: 229 376; B=256, P=784, C=128: 6 553 600; B=256, P=3136 does not fit).
 * Every call on a device shares the array, which keeps the previous call's contents (the entry clears what must
 * start at zero): calls that OVERLAP on the GPU — different streams without an ordering between them — race on
 * it; calls on one stream, or ordered by events, are safe. */
#define CP2_PREP_STATIC_BYTES (8u << 20)
#define CP2_PREP_STATIC_MAX_B 4096
int cp2_prep_features_static(const float *feat, const float *mask, void *fhat_t, float *inv_norm,
                             float *pos_sum, float *neg_sum, int32_t *key_count, int32_t *perm,
                             float *pos_unit, float *neg_unit, float *norms, int B, int C, int P, void *stream);
/* Query side and key side (key side compacted through key_count) of one step in ONE launch; per side the
 * contract of cp2_prep_features (perm_k may be NULL).  scratch_counters: 2*B uint32. */
int cp2_prep_features2(const float *feat_q, const float *mask_q, void *qhat_t, float *inv_norm_q,
                       float *pos_sum_q, float *neg_sum_q, float *scratch_q, float *pos_unit_q,
                       float *neg_unit_q, float *norms_q, const float *feat_k, const float *mask_k,
                       void *khat_t, float *inv_norm_k, float *pos_sum_k, float *neg_sum_k,
                       float *scratch_k, int32_t *key_count, int32_t *perm_k, float *pos_unit_k,
                       float *neg_unit_k, float *norms_k, void *scratch_counters, int B, int C, int P,
                       void *stream);

/* K2 forward — dense pixel InfoNCE, column log-sum-exp.
 * Replaces builder.py:1289 (einsum ncx,ncy->nxy), :1431 (/T) and the LogSoftmax(dim=1)
 * of :1432 without materialising [B,P,P]:
 *   lse[n,y] = log sum_x exp( qhat[n,x,:].khat[n,y,:] * inv_tau )      ([B,P] f32)
 * tcgen05 (bf16 operands, fp32 accumulate in TMEM), operands fetched by TMA.
 *   key_count : NULL (all P key rows) or [B] int32: only rows y < key_count[n] are meaningful
 *               (compacted keys); key tiles entirely past it are skipped and their lse set to 0.
 * C in {128, 256}; 0 < inv_tau <= 16. */
int cp2_dense_lse(const void *qhat_t, const void *khat_t, const int32_t *key_count, float *lse,
                  float inv_tau, int B, int C, int P, void *stream);

/* K2 backward — fused gradient of the dense + global loss w.r.t. the query features.
 * Replaces autograd of builder.py:1261-1268, 1289-1292, 1431-1437:
 *   acc[x,:]  = sum_y on_y * exp(S[x,y]*inv_tau - lse[y]) * khat[y,:]   (tcgen05)
 *               on_y = (mask_b == NULL || mask_b[y] != 0) && (key_count == NULL || y < key_count[n])
 *   dqhat_x   = alpha[n]*acc[x,:] + mask_a[x]*v1[n,:] + (1-mask_a[x])*v2[n,:]
 *   dq[n,:,x] = (dqhat_x - qhat_x (qhat_x . dqhat_x)) * inv_norm[n,x]        ([B,C,P] f32)
 * alpha [B], v1 [B,C], v2 [B,C] (v2 may be NULL = zeros) carry every rank-1 term (cp2_loss_combine
 * writes them for a unit upstream gradient); gscale is NULL or a device scalar, the upstream
 * gradient d L / d loss, multiplied into all three on load (no host read, no scaling pass).
 *   scratch_nl : [B, cp2_dense_pad(P)] f32 work space (cp2_dense_pad: P rounded up to whole key tiles, a multiple of 384). */
int cp2_dense_pad(int P);
int cp2_dense_grad(const void *qhat_t, const void *khat_t, const int32_t *key_count, const float *lse,
                   const float *mask_a, const float *mask_b, const float *inv_norm, const float *alpha,
                   const float *v1, const float *v2, const float *gscale, float *dq, float *scratch_nl,
                   float inv_tau, int B, int C, int P, void *stream);

/* K2-small — the same loss for maps that fit one SM's tensor memory (P <= 256 at C = 128, P <= 128 at
 * C = 256; 14x14 = 196 is the map configs[1] trains at): ONE launch gives the column log-sum-exp of the
 * forward and the P^2-sized part of the gradient, so the backward recomputes nothing.
 * Replaces builder.py:1289-1292, 1431-1437 and their autograd like cp2_dense_lse + cp2_dense_grad:
 *   lse[n,y]   = log sum_x exp(S[x,y]*inv_tau)                               ([B,P] f32; 0 for y >= key_count[n])
 *   acc[n,:,x] = sum_y on_y * exp(S[x,y]*inv_tau - lse[y]) * khat[y,:]        ([B,C,P] f32)
 * cp2_dense_small_bwd then applies what cp2_dense_grad's epilogue applies (rank-1 terms, upstream
 * gradient from device memory, F.normalize backward):
 *   dq[n,:,x] = N_x( g * (alpha[n]*acc[n,:,x] + mask_a[x]*v1[n,:] + (1-mask_a[x])*v2[n,:]) )
 * The acc_seg meter of builder.py:1442-1448 (label of the arg-max over all pixel pairs) comes out of the same
 * launch when seg_hit is given: the largest similarity is found while S is in TMEM, every pair within 2^-7 of it
 * (the bound of the operands' bf16 rounding) is re-scored exactly in fp32 (at most 32 such pairs per image; an image
 * with more — e.g. collapsed features — is decided among the first 32 collected):
 *   seg_q, seg_k : [B,C,P] f32 encoder outputs; seg_inv_q, seg_inv_k : [B,P] (cp2_prep_features inv_norm);
 *   seg_perm_k   : [B,P] int32 pixel of each key row (cp2_prep_features perm) or NULL (keys in pixel order);
 *   seg_mask_a, seg_mask_b : [B,P] f32 masks in pixel order (seg_mask_b NULL: label = row < key_count);
 *   seg_hit      : [B] f32 OUT = mask_a[x*] * mask_b[y*]   (NULL: meter off, the other seg_* are ignored).
 * cp2_dense_small_ok(C, P) = 1 when the shape is taken by this path. */
int cp2_dense_small_ok(int C, int P);
int cp2_dense_small(const void *qhat_t, const void *khat_t, const int32_t *key_count, const float *mask_b,
                    float *lse, float *acc, const float *seg_q, const float *seg_k, const float *seg_inv_q,
                    const float *seg_inv_k, const int32_t *seg_perm_k, const float *seg_mask_a,
                    const float *seg_mask_b, float *seg_hit, float inv_tau, int B, int C, int P, void *stream);
int cp2_dense_small_bwd(const float *acc, const void *qhat_t, const int32_t *key_count, const float *mask_a,
                        const float *inv_norm, const float *alpha, const float *v1, const float *v2,
                        const float *gscale, float *dq, int B, int C, int P, void *stream);

/* K3 — global MoCo InfoNCE against the negative queue: forward sums, backward sums and the top-1
 * accuracy meter in one launch and one pass over the queue.  Replaces builder.py:1396
 * (q_pos @ queue.clone()), the queue part of F.cross_entropy at :1428 and its autograd, and
 * :1441 + _accuracy :1690-1706 (top-1 of logits_moco):
 *   z[b]   = sum_k exp((l[b,k] - 1) * inv_tau),      l = q_pos @ queue
 *   g[b,:] = sum_k exp((l[b,k] - 1) * inv_tau) * queue[:,k]
 *   hit[b] = l_pos[b] > every l[b,k] (and > the two background logits when q_neg/k_neg are given),
 *            l_pos = q_pos[b].k_pos[b];   acc[0] = 100 * mean(hit)
 * (fixed shift 1: |l| <= 1 for unit vectors, so no running max is needed.)
 *   q_pos    : [B,C] f32 unit rows (used as a bf16 hi+lo pair)
 *   queue_bf : [K,C] bf16, the entry-major SHADOW of the queue (cp2_queue_shadow / cp2_enqueue keep it in
 *              step with the fp32 buffer); the only queue bytes the pass streams (TMA, 128-byte swizzle)
 *   queue    : [C,K] f32 (the checkpointed buffer) — read only for the few entries whose bf16 logit lies
 *              within 3e-3 of l_pos, which are re-scored exactly in fp32; may be NULL when k_pos is NULL
 *   k_pos    : [B,C] f32 or NULL (no accuracy meter; hit / acc untouched)
 *   q_neg, k_neg : [B,C] f32 or both NULL (include_background, builder.py:1414-1418)
 *   z : [B] f32; g : [B,C] f32; hit : [B] f32; acc : [1] f32
 *   scratch  : cp2_global_scratch_bytes(B,C) bytes, 8-byte aligned, PERSISTENT: all zero before the first
 *              call and left all zero by every call (fixed-point accumulators of the deterministic
 *              cross-CTA reduction + the done counter); one scratch per concurrently running stream.
 * C in {128, 256}; any K > 0; 0 < inv_tau <= 16; queue_bf 16-byte aligned. */
int64_t cp2_global_scratch_bytes(int B, int C);
int cp2_queue_shadow(const float *queue, void *queue_bf, int C, int K, void *stream);
int cp2_global_nce(const float *q_pos, const void *queue_bf, const float *queue, const float *k_pos,
                   const float *q_neg, const float *k_neg, float *z, float *g, float *hit, float *acc,
                   void *scratch, float inv_tau, int B, int C, int K, void *stream);

/* a5/a7/a12 glue — the [B,C] / [B]-sized remainder of the CP2 loss in two launches.
 * Replaces builder.py:1395 (l_pos), :1414-1418 (background logits), :1428 (cross entropy at the
 * positive), :1431-1437 (rank-1 positive term and masked mean of the dense loss), :1439 (loss
 * combination), the cross-image std meters (:1260-1285 stats) and, for the backward, the global loss's
 * gradient pushed through the outer F.normalize of :1264,1268.
 *   q_pos,k_pos,(q_neg,k_neg) : [B,C] unit pooled vectors (cp2_prep_features pos_unit / neg_unit);
 *                               q_neg = k_neg = v2 = NULL when include_background is off
 *   sq, sk   : [B,C] pooled sums before normalisation (pos_sum of the query / key side)
 *   norms_q  : [B,2] (||sq||, ||neg_sum_q||) from cp2_prep_features
 *   z, G     : [B], [B,C] from cp2_global_nce ; lse : [B,P] from cp2_dense_lse (compacted key order)
 *   key_count: [B] int32 foreground keys per image ; mask_a : [B,P] f32
 *   dense_n  : NULL, or [B] per-image dense losses from cp2_dense_weighted_lse (PROPOSED modes): then
 *              sq, sk and lse are not read and v1 carries the global loss's gradient only
 *   hit_seg  : NULL or [B] f32 (cp2_dense_small's seg_hit): out[5] = 100 * mean(hit_seg), the acc_seg meter
 *   out      : [6] f32 = {loss, loss_ins, loss_dense, mean_c std_b(q_pos), mean_c std_b(k_pos), acc_seg | NaN}
 *   l_pos [B]; alpha [B], v1 [B,C], v2 [B,C] : inputs of cp2_dense_grad for upstream gradient 1
 *   scratch  : [B,2] f32 */
int cp2_loss_combine(const float *q_pos, const float *k_pos, const float *q_neg, const float *k_neg,
                     const float *sq, const float *sk, const float *norms_q, const float *z, const float *G,
                     const float *lse, const int32_t *key_count, const float *mask_a, const float *dense_n,
                     const float *hit_seg, float *out, float *l_pos, float *alpha, float *v1, float *v2,
                     float *scratch, void *scratch_counter, float inv_tau_global,
                     float inv_tau_dense, float lmbd_dense, int B, int C, int P, void *stream);

/* K4 — momentum (EMA) update of the key encoder, all parameters in one launch.
 * Replaces builder.py:557-567 (3 ATen kernels x ~183 tensors).
 *   k = fadd_rn(fmul_rn(k, m), fmul_rn(q, one_minus_m)), bit-exact with the reference.
 *   q_ptrs, k_ptrs : device arrays [n_tensors] of tensor base pointers
 *   chunk_tensor, chunk_index : device arrays [n_chunks] (int32): chunk i covers
 *       elements [chunk_index[i]*CP2_EMA_CHUNK, ...) of tensor chunk_tensor[i]
 *   numel : device array [n_tensors] (int64) */
#define CP2_EMA_CHUNK 4096
int cp2_ema(const float *const *q_ptrs, float *const *k_ptrs, const int64_t *numel,
            const int32_t *chunk_tensor, const int32_t *chunk_index, int n_chunks, float m,
            float one_minus_m, void *stream);

/* K5 — queue enqueue of the all-gathered keys with wrap-around, device-side pointer, ONE launch.
 * Replaces builder.py:574-587 (int(queue_ptr) D2H sync, 1-2 transposed slice writes,
 * pointer update).  The all-gather itself (builder.py:1710-1722) is issued by the
 * caller as one ncclAllGather into keys_all on the same stream.
 *   keys_all : [n,C] f32 ; queue : [C,K] f32 ; queue_ptr : [1] int64 (read, then advanced)
 *   queue_bf : NULL, or the [K,C] bf16 shadow cp2_global_nce streams: rows (ptr+i) mod K are rewritten too
 *   scratch_counter : 4 bytes, PERSISTENT: zero before the first call, left zero by every call (done
 *              counter of "the last CTA to have read the pointer advances it")
 * queue and queue_ptr bit-exact with the reference (n <= K). */
int cp2_enqueue(const float *keys_all, float *queue, void *queue_bf, int64_t *queue_ptr, void *scratch_counter,
                int n, int C, int K, void *stream);

/* a2 + a4 — sub-sampled id maps and their (masked) correlation IoU.
 * Replaces builder.py:1165-1186 (the `[OS/2::OS]` subsample of the id maps) and
 * tools/correlation_mapping.py:103-138 get_masked_iou (a python loop of torch.unique per image
 * followed by .cpu(), builder.py:1252-1257), for one pair of id maps:
 *   values   = (id + 1) of both sub-sampled maps, 0 where the sub-sampled mask is 0
 *   iou[n]   = #distinct values occurring more than once / #distinct values         (all pixels)
 *   iou_masked[n] likewise over the unmasked pixels only; 0/0 -> NaN (the reference raises)
 *   ids_a, ids_b : [B,H,W] int64 (full resolution) ; mask_a_ds, mask_b_ds : [B,hs,ws] f32 of K1
 *   iou, iou_masked : [B] f32.  Bit-exact with the reference. */
int cp2_id_iou(const int64_t *ids_a, const int64_t *ids_b, const float *mask_a_ds,
               const float *mask_b_ds, float *iou, float *iou_masked, int B, int H, int W, int os,
               void *stream);

/* a13/f5 — MirrorModule loss (networks/mirror_network.py:40-63): class CE on both logit maps +
 * the "mirror" output-consistency term, forward and backward in one pass.
 *   class   = CE(cat(s,t), cat(labels,labels))               nn.CrossEntropyLoss(), ignore_index -100
 *   compare = CE(input = softmax(s/T), target = softmax(t/T)) soft-target CE on a probability map
 *   loss    = class + lmbd_compare * compare                  (mirror_network.py:62-63)
 * Replaces the ~10 ATen passes of mirror_network.py:47-63 and their autograd.
 *   s_logits, t_logits : [B,Cls,HW] f32.  t_logits == NULL selects MirrorVariant.NONE
 *                        (mirror_network.py:53-55): class CE on s only, compare = 0.
 *   labels  : [B,HW] int64; -100 (and any value outside [0,Cls)) is ignored by the class term
 *   partial : [cp2_mirror_partials(B,HW), 3] f32 work space (per-CTA sums, reduced in fixed order)
 *   out     : [4] f32 = {loss, class, compare, number of valid labelled pixels counted}
 *   ds, dt  : NULL (loss only) or [B,Cls,HW] f32 OUT: d loss/d s_logits, d loss/d t_logits,
 *             multiplied by *gscale when gscale != NULL (the upstream gradient, a device scalar)
 *   n_valid : number of labelled pixels that are not ignored, <= 0 meaning "all B*HW" (the mean of
 *             the class term divides by it; out[3] reports the counted value for verification)
 * 2 <= Cls <= 8; inv_temp = 1/T > 0. */
int cp2_mirror_partials(int B, int64_t HW);
int cp2_mirror_loss(const float *s_logits, const float *t_logits, const int64_t *labels,
                    float *partial, float *out, float *ds, float *dt, const float *gscale,
                    float lmbd_compare, float inv_temp, int64_t n_valid, int B, int Cls, int64_t HW,
                    void *stream);

/* a13/f5 — the same loss with the bilinear `resize` of SegmentationModule.forward
 * (networks/segment_network.py:219-231; F.interpolate(mode="bilinear", align_corners=False))
 * fused in: the kernel reads the decoder's low-resolution logits, interpolates them on chip
 * and accumulates the upsample's backward on chip, so the [B,Cls,H,W] logits and their gradients
 * never exist in HBM (algorithmic traffic: the labels).
 *   s_lo, t_lo   : [B,Cls,h,w] f32 (t_lo == NULL: MirrorVariant.NONE); labels : [B,H,W] int64
 *   partial      : [cp2_mirror_up_partials(B,H,W), 3] f32 work space; out : [4] f32 as above
 *   ds_lo, dt_lo : NULL or [B,Cls,h,w] f32 OUT (zeroed here, then accumulated with fp32 atomics:
 *                  last-bit run-to-run differences, as ATen's own upsample backward); when dt_lo directly
 *                  follows ds_lo in memory both are cleared by one memset
 * Two classes (Cls == 2, the polyp masks of mirror_pretrain.py) run a specialised kernel on the logit
 * DIFFERENCES s1 - s0, t1 - t0 (every term of the loss depends on them alone); other class counts the generic one.
 * Gradients are for upstream gradient 1 (they are tiny: scale them on the caller's side). */
int cp2_mirror_up_partials(int B, int H, int W);
int cp2_mirror_loss_upsampled(const float *s_lo, const float *t_lo, const int64_t *labels,
                              float *partial, float *out, float *ds_lo, float *dt_lo,
                              float lmbd_compare, float inv_temp, int64_t n_valid, int B, int Cls,
                              int h, int w, int H, int W, void *stream);

/* f3 — dense loss of forward_cp2 in the PROPOSED modes: per-pair correspondence weights from the id maps
 * and the NegativeType squash, applied on the fly.  Replaces builder.py:1204-1243 (two
 * get_masked_correlation_map calls, the known-region einsum, corr_weights), :1332-1392 (negative
 * transform, `* corr_weights`) and :1431-1437 with their autograd — nothing [B,P,P]-sized is stored.
 *   z[x,y]   = w[x,y] * t(S[x,y]) * inv_tau,  S = qhat_x . khat_y
 *   w[x,y]   = w_pix if pixel ids agree, else w_reg if region ids agree and are non-zero, else 0;
 *              then w_not wherever that is 0
 *   t(s)     = s on label pairs (mask_a[x] != 0 and mask_b[y] != 0) or neg_type 0;
 *              2/(1+exp(-neg_scale*(s - centre))) - 1 elsewhere, centre = 0 (neg_type 1, FIXED) or the
 *              image's mean (neg_type 2, AVERAGE) or median (neg_type 3, MEDIAN: torch.nanquantile 0.5, by a
 *              4-pass radix select that recomputes S; 8 more launches) of the non-label S.
 *              NegativeType.HARD is a no-op in the reference: pass 0.
 *   lse[n,y] = log sum_x exp(z[x,y]);  dense_n[n] = sum_xy a_x b_y (lse_y - z[x,y]) / (sum a * sum b)
 *   qhat_t, khat_t : [B,P,C] bf16 from cp2_prep_features, keys NOT compacted; masks [B,P] f32 (binary)
 *   *_ids_*  : [B,ids_h,ids_w] int64 id maps; pixel p = (r,c) of the P = hs*ws lattice reads element
 *              (r*os + os/2, c*os + os/2), the reference's [OS//2::OS] subsample (builder.py:1165-1186);
 *              pass os = 1 for maps that are already subsampled
 *   pos_*, neg_* : [B,C] pooled sums of cp2_prep_features (query / key side), for the AVERAGE centre
 *   acc_seg  : NULL or [1] f32 out: 100 * mean_n [arg-max pair of z[n] is a label pair] (builder.py:1442-1448
 *              on the weighted logits; bf16 operands, so near-ties may fall differently from the reference)
 *   wstats   : [B,3] f32 out = {centre, sum mask_a, sum mask_b}; key_count [B] int32 out
 *   scratch  : [B, cp2_dense_weighted_tiles(P)] f32.   C in {128, 256}. */
int cp2_dense_weighted_tiles(int P);
int cp2_dense_weighted_lse(const void *qhat_t, const void *khat_t, const float *mask_a, const float *mask_b,
                           const int64_t *pixel_ids_a, const int64_t *pixel_ids_b, const int64_t *region_ids_a,
                           const int64_t *region_ids_b, const float *pos_q, const float *neg_q,
                           const float *pos_k, const float *neg_k, float *lse, float *dense_n, float *acc_seg,
                           float *wstats, int32_t *key_count, float *scratch, float w_pix, float w_reg, float w_not,
                           int neg_type, float neg_scale, float inv_tau, int B, int C, int P, int ids_h,
                           int ids_w, int os, void *stream);

/* f3 backward — d loss / d q for the PROPOSED modes, one launch (recomputes S):
 *   acc[x,:]  = sum_y b_y (exp(z[x,y] - lse[y]) - a_x / sum a) * w[x,y] * t'(S[x,y]) * khat[y,:]
 *   dqhat_x   = alpha[n]*acc[x,:] + mask_a[x]*v1[n,:] + (1-mask_a[x])*v2[n,:]
 *   dq[n,:,x] = (dqhat_x - qhat_x (qhat_x . dqhat_x)) * inv_norm[n,x]         ([B,C,P] f32)
 * alpha, v1, v2 from cp2_loss_combine called with dense_n; gscale as in cp2_dense_grad. */
int cp2_dense_weighted_grad(const void *qhat_t, const void *khat_t, const float *mask_a, const float *mask_b,
                            const int64_t *pixel_ids_a, const int64_t *pixel_ids_b, const int64_t *region_ids_a,
                            const int64_t *region_ids_b, const float *lse, const float *wstats,
                            const float *inv_norm, const float *alpha, const float *v1, const float *v2,
                            const float *gscale, float *dq, float w_pix, float w_reg, float w_not, int neg_type,
                            float neg_scale, float inv_tau, int B, int C, int P, int ids_h, int ids_w, int os,
                            void *stream);

/* f3 on the tensor cores — cp2_dense_weighted_lse for maps that fit TMEM (cp2_dense_small_ok(C, P)): the
 * statistics pre-pass, then ONE tcgen05 launch gives lse, the per-image loss dense_n, the label of the arg-max
 * pair hit [B] (acc_seg = 100 * mean(hit); may be NULL) and
 *   acc[n,:,x] = sum_y b_y (exp(z[x,y] - lse_y) - a_x/A) w[x,y] t'(S[x,y]) khat[y,:]            ([B,C,P] f32)
 * so that the backward is cp2_dense_small_bwd (alpha * acc + rank-1 terms, normalize backward) instead of
 * cp2_dense_weighted_grad.  Other arguments as cp2_dense_weighted_lse. */
int cp2_dense_weighted_small(const void *qhat_t, const void *khat_t, const float *mask_a, const float *mask_b,
                             const int64_t *pixel_ids_a, const int64_t *pixel_ids_b, const int64_t *region_ids_a,
                             const int64_t *region_ids_b, const float *pos_q, const float *neg_q, const float *pos_k,
                             const float *neg_k, float *lse, float *dense_n, float *hit, float *wstats,
                             int32_t *key_count, float *scratch, float *acc, float w_pix, float w_reg, float w_not,
                             int neg_type, float neg_scale, float inv_tau, int B, int C, int P, int ids_h, int ids_w,
                             int os, void *stream);

/* f3 on the tensor cores for maps that do not fit TMEM (P > 256, C = 128; e.g. 28x28): the streaming tcgen05
 * pair.  cp2_dense_weighted_stream = cp2_dense_weighted_lse (same outputs: lse, dense_n, acc_seg, wstats,
 * key_count) and additionally leaves the ids gathered at the lattice and the padded key mask for the backward:
 *   ids_pad : cp2_dense_weighted_stream_ids_bytes(B,P) bytes = [4][B][cp2_dense_pad(P)] int64 (pid_a, pid_b, rid_a, rid_b)
 *   b_pad   : [B][cp2_dense_pad(P)] f32
 * cp2_dense_weighted_stream_grad = cp2_dense_weighted_grad on the tensor cores (dense backward kernel, weighted
 * soft-max in place in TMEM); scratch_nl [B, cp2_dense_pad(P)] f32.  0 < inv_tau <= 16. */
int64_t cp2_dense_weighted_stream_ids_bytes(int B, int P);
int cp2_dense_weighted_stream(const void *qhat_t, const void *khat_t, const float *mask_a, const float *mask_b,
                              const int64_t *pixel_ids_a, const int64_t *pixel_ids_b, const int64_t *region_ids_a,
                              const int64_t *region_ids_b, const float *pos_q, const float *neg_q, const float *pos_k,
                              const float *neg_k, float *lse, float *dense_n, float *acc_seg, float *wstats,
                              int32_t *key_count, float *scratch, void *ids_pad, float *b_pad, float w_pix,
                              float w_reg, float w_not, int neg_type, float neg_scale, float inv_tau, int B, int C,
                              int P, int ids_h, int ids_w, int os, void *stream);
int cp2_dense_weighted_stream_grad(const void *qhat_t, const void *khat_t, const float *mask_a, const float *mask_b,
                                   const float *lse, const void *ids_pad, const float *b_pad, const float *wstats,
                                   const float *inv_norm, const float *alpha, const float *v1, const float *v2,
                                   const float *gscale, float *dq, float *scratch_nl, float w_pix, float w_reg,
                                   float w_not, int neg_type, float neg_scale, float inv_tau, int B, int C, int P,
                                   void *stream);

/* f4 — acc_seg without the [B,P,P] logits.  Replaces builder.py:1442-1448:
 *   hit[n] = mask_a[n,x*] * mask_b[n,y*] at (x*,y*) = argmax_{x,y} qhat_x . khat_y (first occurrence, flat
 *   index x*P+y as the reference's reshape); acc[0] = 100 * mean_n hit[n].
 * q, k: the encoder outputs [B,C,P] f32 (normalised inside, fp32 CUDA-core dot products so that near-ties
 * fall as in the reference's fp32 einsum).  scratch: cp2_dense_top1_scratch_floats(B,P) f32. */
int64_t cp2_dense_top1_scratch_floats(int B, int P);
int cp2_dense_top1(const float *q, const float *k, const float *mask_a, const float *mask_b, float *hit,
                   float *acc, float *scratch, int B, int C, int P, void *stream);

/* f2 — ContrastiveHead (builder.py:150-176) for N rows against a queue, given cp2_global_nce's z [N], G [N,C]
 * for those rows: the global loss of forward_densecl (builder.py:762-775; N = B) and its local loss
 * (builder.py:866-873, 906-908; N = B*S^2, rows = q_local pixels, queue2).
 *   l_r = rows[r].pos_keys[r];  terms[r] = log(z[r] + exp((l_r-1)/tau)) + (1-l_r)/tau;  loss[0] = mean_r terms
 *   dq[r,:] = d loss / d rows[r,:] = ((G[r,:] + exp((l_r-1)/tau) pos_keys[r,:]) / Z_r - pos_keys[r,:]) / (N tau)
 * rows, pos_keys [N,C] f32 (rows unit vectors); terms [N] scratch/out; dq may be NULL (no gradient wanted).
 * z / G may be given as nz / ng partial sums, zstride / gstride floats apart (cp2_queue_sums); 1, 0 otherwise. */
int cp2_rows_nce(const float *rows, const float *pos_keys, const float *z, const float *G, float *loss,
                 float *terms, float *dq, float inv_tau, int N, int C, int nz, int64_t zstride, int ng,
                 int64_t gstride, void *stream);

/* f2 — the queue sums z, G of cp2_global_nce for MANY rows (the local InfoNCE of forward_densecl has
 * N = B*S^2 = 6272 rows against the 65536-entry queue2: builder.py:866-873) on the tensor cores: the dense
 * loss's backward kernel in its rectangular mode (queries = the rows, keys = the queue entries, cut into
 * cp2_queue_sums_split(N,K) contiguous ranges so that ranges x 128-row tiles fill the SMs).  PARTIAL sums come
 * back; cp2_rows_nce adds them in a fixed order (nz = 2*split, zstride = cp2_dense_pad(N); ng = split,
 * gstride = N*C):
 *   zp   [2*split, cp2_dense_pad(N)] f32:  sum_i zp[i,r]   = sum_k exp((rows[r].u_k - 1)/tau)
 *   gacc [split, N, C] f32:                sum_s gacc[s,r] = sum_k exp((rows[r].u_k - 1)/tau) u_k
 *   rows_bf  [N,C] bf16 unit rows (cp2_prep_features' operand layout); queue_bf [K,C] bf16 (the queue's shadow);
 *   nl [ceil128(K)] f32 = -log2(e)/tau for k < K, -80 for the padding (built once per (K, tau)).
 * C = 128; 0 < inv_tau <= 16. */
int cp2_queue_sums_split(int N, int K);
int cp2_queue_sums(const void *rows_bf, const void *queue_bf, const float *nl, float *zp, float *gacc,
                   float inv_tau, int N, int C, int K, void *stream);

/* f2 — the positive of every query pixel (builder.py:824-855), as a vector: kp[n*P+x,:] = k_local[n,:,idx[n,x]],
 * and where key pixels share query pixel x's id: (1-lmbd) * that + lmbd * sum_{y: pix_k[y]==pix_q[x]} k_local[n,:,y]
 * so that pos_local = q_local_x . kp_x without the [B,S^2,S^2] local similarity matrix.
 *   k_local [B,C,P] f32 unit columns; idx [B,P] int64 (the backbone-feature match, builder.py:817-821);
 *   pix_q, pix_k: [B,ids_h,ids_w] int64 id maps read at the [os//2::os] lattice (os = 1: already sub-sampled),
 *   or both NULL (no coordinate mixing); kp [B*P,C] f32 out.
 * Note: the reference itself raises IndexError at builder.py:861 whenever one id is shared; this implements the
 * formula of :839-855. */
int cp2_positive_keys(const float *k_local, const int64_t *idx, const int64_t *pix_q, const int64_t *pix_k,
                      float *kp, float lmbd_coordinate, int B, int C, int P, int ids_h, int ids_w, int os,
                      void *stream);

/* f2 — the backbone match of forward_densecl (builder.py:817-821) in one tcgen05 kernel:
 *   idx[n,x] = argmax_y sum_c embd_q[n,c,x] * embd_k[n,c,y]        (first largest y)
 * embd_q, embd_k : [B,Cb,P] f32 (the backbone feature maps, pixel axis contiguous), 16-byte aligned;
 * idx : [B,P] int64.  Each fp32 value enters as two bf16 terms (hi + lo; three products per pair), i.e. the
 * similarity to ~2^-16 relative; the [B,P,P] matrix exists only in TMEM.
 *   scratch : cp2_densecl_match_scratch_bytes(B,P) bytes, 8-byte aligned, PERSISTENT: all zero before the first call
 *             and left all zero by every call (per-pixel 64-bit arg-max keys combined across key tiles + per-image
 *             done counters).
 * P <= 256, Cb a multiple of 16. */
int64_t cp2_densecl_match_scratch_bytes(int B, int P);
int cp2_densecl_match(const float *embd_q, const float *embd_k, int64_t *idx, void *scratch, int B, int Cb, int P,
                      void *stream);

#ifdef __cplusplus
}
#endif
#endif /* CP2_B200_H_ */
