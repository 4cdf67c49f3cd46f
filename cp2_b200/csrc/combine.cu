// a5/a7/a12 glue — everything of the CP2 loss that lives on [B,C] / [B] tensors, in ONE launch (the CTA of
// the last image to finish sums the batch; done counter in a persistent, self-resetting scratch word):
// the global InfoNCE's positive (and optional background) logits and its cross-entropy
// (reference builder.py:1395, 1414-1418, 1428), the rank-1 positive term and the masked mean of the
// dense loss (builder.py:1431-1437), the loss combination (builder.py:1439), and the per-image
// coefficient vectors the one-kernel backward (cp2_dense_grad) consumes: the global loss's gradient
// pushed through the outer F.normalize of the pooled query vectors (builder.py:1264,1268), plus the
// dense loss's rank-1 term.  The [B,P,P] work is done by cp2_dense_lse and the queue pass by
// cp2_global_nce; this kernel only combines their [B,P] / [B,C] results, replacing ~40 tiny ATen
// launches of the eager formulation.
#include "common.cuh"

namespace {

constexpr float kEps = 1e-12f;     // F.normalize eps

template <int N>
__device__ __forceinline__ void block_sum(float (&v)[N], float *red /* [N][4] */) {
#pragma unroll
  for (int k = 0; k < N; ++k)
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v[k] += __shfl_down_sync(0xffffffffu, v[k], off);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  __syncthreads();                                   // red may still be read from a previous round
  if (lane == 0)
#pragma unroll
    for (int k = 0; k < N; ++k) red[k * 4 + wid] = v[k];
  __syncthreads();
#pragma unroll
  for (int k = 0; k < N; ++k) v[k] = red[k * 4] + red[k * 4 + 1] + red[k * 4 + 2] + red[k * 4 + 3];
}

// out = {loss, loss_ins, loss_dense, mean_c std_b(q_pos), mean_c std_b(k_pos), acc_seg}; fixed summation order.
// 128 threads.  acc_seg = 100 * mean(hit_seg) (builder.py:1442-1448) or NaN without hit_seg.
__device__ void loss_finalize(const float *terms, const float *__restrict__ q_pos, const float *__restrict__ k_pos,
                              const float *hit_seg, float *__restrict__ out, float lmbd, int B, int C,
                              float *red /* >= 8 floats */) {
  float sd[2] = {0.f, 0.f};
  for (int c = threadIdx.x; c < C; c += 128) {
#pragma unroll
    for (int w = 0; w < 2; ++w) {
      const float *src = w == 0 ? q_pos : k_pos;
      float m = 0.f;
      for (int b = 0; b < B; ++b) m += src[(long long)b * C + c];
      m /= (float)B;
      float var = 0.f;
      for (int b = 0; b < B; ++b) { const float d = src[(long long)b * C + c] - m; var = fmaf(d, d, var); }
      sd[w] += sqrtf(var / (float)(B - 1));                             // torch.std: unbiased; B = 1 -> NaN
    }
  }
#pragma unroll
  for (int w = 0; w < 2; ++w)
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) sd[w] += __shfl_down_sync(0xffffffffu, sd[w], off);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) { red[wid] = sd[0]; red[4 + wid] = sd[1]; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float a = 0.f, d = 0.f, s0 = 0.f, s1 = 0.f;
    for (int b = 0; b < B; ++b) { a += __ldcg(terms + 2 * b); d += __ldcg(terms + 2 * b + 1); }
    for (int w = 0; w < 4; ++w) { s0 += red[w]; s1 += red[4 + w]; }
    a /= (float)B;
    d /= (float)B;
    out[0] = a + d * lmbd;
    out[1] = a;
    out[2] = d;
    out[3] = s0 / (float)C;
    out[4] = s1 / (float)C;
    float hs = __int_as_float(0x7fc00000);
    if (hit_seg) {
      hs = 0.f;
      for (int b = 0; b < B; ++b) hs += __ldcg(hit_seg + b);
      hs *= 100.f / (float)B;
    }
    out[5] = hs;
  }
}

// One CTA (128 threads) per image.
__global__ void __launch_bounds__(128)
loss_combine_kernel(const float *__restrict__ q_pos, const float *__restrict__ k_pos,
                    const float *__restrict__ q_neg, const float *__restrict__ k_neg,
                    const float *__restrict__ sq, const float *__restrict__ sk,
                    const float *__restrict__ norms_q, const float *__restrict__ z,
                    const float *__restrict__ G, const float *__restrict__ lse,
                    const int *__restrict__ key_count, const float *__restrict__ mask_a,
                    const float *__restrict__ dense_n, const float *hit_seg, float *__restrict__ terms,
                    float *__restrict__ l_pos_out, float *__restrict__ alpha,
                    float *__restrict__ v1, float *__restrict__ v2, float *__restrict__ out,
                    unsigned int *__restrict__ counter, float inv_tg, float inv_tl, float lmbd,
                    int B, int C, int P) {
  __shared__ float red[6 * 4];
  __shared__ int is_last;
  const int b = blockIdx.x, tid = threadIdx.x;
  const bool bg = q_neg != nullptr;
  const long long row = (long long)b * C;
  const int cnt = min(max(key_count[b], 0), P);
  // every operand of this image is read ONCE, up front (C <= 256: two channels per thread), so the kernel is one
  // memory round trip + two block reductions instead of three dependent passes over global memory
  constexpr int NC = 2;
  float qp[NC], kp[NC], qn[NC], kn[NC], Gv[NC], sqv[NC], skv[NC];
#pragma unroll
  for (int i = 0; i < NC; ++i) {
    const int c = tid + 128 * i;
    const bool v = c < C;
    qp[i] = v ? q_pos[row + c] : 0.f;
    kp[i] = v ? k_pos[row + c] : 0.f;
    Gv[i] = v ? G[row + c] : 0.f;
    qn[i] = (v && bg) ? q_neg[row + c] : 0.f;
    kn[i] = (v && bg) ? k_neg[row + c] : 0.f;
    sqv[i] = (v && !dense_n) ? sq[row + c] : 0.f;
    skv[i] = (v && !dense_n) ? sk[row + c] : 0.f;
  }
  const float zb = z[b], nq = norms_q[2 * b], nnq = norms_q[2 * b + 1];
  // ---- round 1: the dot products and the [P]-sized sums
  float r1[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};     // l_pos, q_pos.q_neg, q_pos.k_neg, sq.sk, A, sum lse
#pragma unroll
  for (int i = 0; i < NC; ++i) {
    r1[0] = fmaf(qp[i], kp[i], r1[0]);
    r1[1] = fmaf(qp[i], qn[i], r1[1]);
    r1[2] = fmaf(qp[i], kn[i], r1[2]);
    r1[3] = fmaf(sqv[i], skv[i], r1[3]);
  }
  for (int x = tid; x < P; x += 128) {
    r1[4] += mask_a[(long long)b * P + x];
    if (!dense_n && x < cnt) r1[5] += lse[(long long)b * P + x];   // compacted keys: rows < cnt are foreground
  }
  block_sum<6>(r1, red);
  const float l_pos = r1[0], A = r1[4], Bm = (float)cnt;
  // shift 1 as in cp2_global_nce: |l| <= 1 for unit vectors
  const float e_pos = expf((l_pos - 1.f) * inv_tg);
  float Z = zb + e_pos, e_nq = 0.f, e_nk = 0.f;
  if (bg) {
    e_nq = expf((r1[1] - 1.f) * inv_tg);
    e_nk = expf((r1[2] - 1.f) * inv_tg);
    Z += e_nq + e_nk;
  }
  const float rZ = 1.f / Z, invB = 1.f / (float)B;
  const float dense_scale = inv_tl / (A * Bm);       // 1 / (tau_d * A * Bm); 0/0 -> NaN as the reference
  // ---- round 2: projections of the outer-normalize backward
  float dqpos[NC], r2[2] = {0.f, 0.f};
#pragma unroll
  for (int i = 0; i < NC; ++i) {
    float num = fmaf(e_pos, kp[i], Gv[i]);
    if (bg) num = fmaf(e_nq, qn[i], fmaf(e_nk, kn[i], num));
    dqpos[i] = (num * rZ - kp[i]) * invB * inv_tg;
    r2[0] = fmaf(qp[i], dqpos[i], r2[0]);
    if (bg) r2[1] = fmaf(qn[i], e_nq * rZ * qp[i] * invB * inv_tg, r2[1]);
  }
  block_sum<2>(r2, red);
  const float dq_d = 1.f / fmaxf(nq, kEps), dn_d = 1.f / fmaxf(nnq, kEps);
#pragma unroll
  for (int i = 0; i < NC; ++i) {
    const int c = tid + 128 * i;
    if (c < C) {
      const float proj = nq > kEps ? dqpos[i] - qp[i] * r2[0] : dqpos[i];
      // dense_n given (PROPOSED modes): the dense loss has no rank-1 term, cp2_dense_weighted_grad does it all
      v1[row + c] = dense_n ? proj * dq_d : proj * dq_d - lmbd * invB * skv[i] * dense_scale;
      if (bg) {
        const float dqneg = e_nq * rZ * qp[i] * invB * inv_tg;
        v2[row + c] = (nnq > kEps ? dqneg - qn[i] * r2[1] : dqneg) * dn_d;
      }
    }
  }
  if (tid == 0) {
    terms[2 * b] = logf(Z) + (1.f - l_pos) * inv_tg;                 // -log softmax at the positive
    terms[2 * b + 1] = dense_n ? dense_n[b] : r1[5] / Bm - r1[3] * dense_scale;   // loss_dense of this image
    l_pos_out[b] = l_pos;
    alpha[b] = lmbd * invB * inv_tl / Bm;
    __threadfence();                      // this image's terms are visible before the done counter moves
    const int last = atomicAdd(counter, 1u) == (unsigned)B - 1u;
    if (last) *counter = 0u;
    is_last = last;
  }
  __syncthreads();
  if (is_last) {                          // the last image's CTA sums the batch in a fixed order
    __threadfence();
    loss_finalize(terms, q_pos, k_pos, hit_seg, out, lmbd, B, C, red);
  }
}

}  // namespace

extern "C" int cp2_loss_combine(const float *q_pos, const float *k_pos, const float *q_neg, const float *k_neg,
                                const float *sq, const float *sk, const float *norms_q, const float *z,
                                const float *G, const float *lse, const int32_t *key_count,
                                const float *mask_a, const float *dense_n, const float *hit_seg, float *out,
                                float *l_pos, float *alpha, float *v1,
                                float *v2, float *scratch, void *scratch_counter, float inv_tau_global,
                                float inv_tau_dense, float lmbd_dense, int B, int C, int P, void *stream) {
  CP2_REQUIRE(B > 0 && C > 0 && C <= 256 && P > 0, "cp2_loss_combine: bad shape B=%d C=%d P=%d (C <= 256)", B, C, P);
  CP2_REQUIRE(q_pos && k_pos && norms_q && z && G && key_count && mask_a && out && l_pos && alpha && v1 && scratch && scratch_counter,
              "cp2_loss_combine: null pointer");
  CP2_REQUIRE(dense_n || (sq && sk && lse), "cp2_loss_combine: sq, sk and lse are needed unless dense_n is given");
  CP2_REQUIRE((q_neg == nullptr) == (k_neg == nullptr) && (q_neg == nullptr) == (v2 == nullptr),
              "cp2_loss_combine: q_neg, k_neg and v2 go together (include_background)");
  CP2_REQUIRE(inv_tau_global > 0.f && inv_tau_dense > 0.f, "cp2_loss_combine: bad temperature");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  loss_combine_kernel<<<B, 128, 0, st>>>(q_pos, k_pos, q_neg, k_neg, sq, sk, norms_q, z, G, lse, key_count, mask_a,
                                         dense_n, hit_seg, scratch, l_pos, alpha, v1, v2, out,
                                         static_cast<unsigned int *>(scratch_counter), inv_tau_global, inv_tau_dense,
                                         lmbd_dense, B, C, P);
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}
