// f3 — the dense loss of forward_cp2 in the PROPOSED modes: per-pair
// correspondence weights from the pixel / region id maps (reference builder.py:1204-1243) and the
// NegativeType squash of the non-label similarities (builder.py:1332-1386), applied on the fly to
// S = qhat^T khat so that neither the [B,P,P] id-equality maps, nor corr_weights, nor the logits are
// ever materialised.  The modes are only run at the 14x14 / 28x28 feature maps the reference can hold,
// where the tensor-core kernel of the CP2 mode is launch-bound anyway, so this is an fp32 SIMT kernel
// pair on the same bf16 operands: forward = column log-sum-exp + the label-weighted positive sum,
// backward = recomputation + the second contraction + the per-pixel normalize backward, one launch.
#include "common.cuh"
#include "weighted_common.cuh"

using namespace weighted;

namespace {

constexpr int WT = 32;                  // tile edge (pixels)
constexpr int WTHREADS = 256;
constexpr int SEL_INTS = 4 + 512;         // MEDIAN selection state per image

struct Best { float v; int i; };
__device__ __forceinline__ void take(Best &a, float v, int i) {
  if (v > a.v || (v == a.v && i < a.i)) { a.v = v; a.i = i; }     // first occurrence on ties
}

// 8 bf16 -> 8 floats at dst
__device__ __forceinline__ void unpack8(uint4 v, float *dst) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    dst[2 * j] = __uint_as_float(w[j] << 16);
    dst[2 * j + 1] = __uint_as_float(w[j] & 0xffff0000u);
  }
}

// rows [row0, row0+32) of one image's [P,C] bf16 matrix -> fp32 smem tile with row stride `ld`
template <int C>
__device__ __forceinline__ void load_rows(const __nv_bfloat16 *src, int row0, int P, float *dst, int ld) {
  for (int i = threadIdx.x; i < WT * (C / 8); i += WTHREADS) {
    const int row = i / (C / 8), c8 = i - row * (C / 8);
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (row0 + row < P) v = *reinterpret_cast<const uint4 *>(src + (long long)(row0 + row) * C + c8 * 8);
    float f[8];
    unpack8(v, f);
#pragma unroll
    for (int j = 0; j < 8; ++j) dst[row * ld + c8 * 8 + j] = f[j];
  }
}

// wstats[b] = {centre, A = sum mask_a, Bm = sum mask_b}; key_count[b] = number of foreground keys.
// centre (NegativeType.AVERAGE, builder.py:1299,1342-1357) = mean of S over the non-label pairs, from
// pooled sums: sum_all S = (sum_x qhat_x).(sum_y khat_y), sum_label S = pos_q.pos_k (binary masks).
__global__ void __launch_bounds__(128)
weighted_stats_kernel(const float *__restrict__ mask_a, const float *__restrict__ mask_b,
                      const float *__restrict__ pos_q, const float *__restrict__ neg_q,
                      const float *__restrict__ pos_k, const float *__restrict__ neg_k,
                      float *__restrict__ wstats, int32_t *__restrict__ key_count, int *__restrict__ sel,
                      int C, int P) {
  __shared__ float red[4][4];
  const int b = blockIdx.x, tid = threadIdx.x;
  float v[4] = {0.f, 0.f, 0.f, 0.f};                 // A, Bm, sum_all, sum_label
  for (int x = tid; x < P; x += 128) {
    v[0] += mask_a[(long long)b * P + x];
    v[1] += mask_b[(long long)b * P + x];
  }
  for (int c = tid; c < C; c += 128) {
    const long long i = (long long)b * C + c;
    v[2] = fmaf(pos_q[i] + neg_q[i], pos_k[i] + neg_k[i], v[2]);
    v[3] = fmaf(pos_q[i], pos_k[i], v[3]);
  }
#pragma unroll
  for (int k = 0; k < 4; ++k)
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v[k] += __shfl_down_sync(0xffffffffu, v[k], off);
  if ((tid & 31) == 0)
#pragma unroll
    for (int k = 0; k < 4; ++k) red[k][tid >> 5] = v[k];
  __syncthreads();
  if (tid == 0) {
    float s[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) s[k] = red[k][0] + red[k][1] + red[k][2] + red[k][3];
    wstats[3 * b] = (s[2] - s[3]) / ((float)P * (float)P - s[0] * s[1]);
    wstats[3 * b + 1] = s[0];
    wstats[3 * b + 2] = s[1];
    key_count[b] = (int)(s[1] + 0.5f);
    if (sel) {                           // NegativeType.MEDIAN: ranks of the two middle order statistics
      const long long n = (long long)P * P - (long long)(s[0] + 0.5f) * (long long)(s[1] + 0.5f);
      int *st = sel + (long long)b * SEL_INTS;
      st[0] = st[1] = 0;                 // key prefixes
      st[2] = (int)((n - 1) / 2);        // floor(0.5 (n-1))
      st[3] = (int)(n / 2);              // ceil (0.5 (n-1))
    }
  }
  if (sel)
    for (int i = tid; i < 512; i += 128) sel[(long long)b * SEL_INTS + 4 + i] = 0;
}

// ---- NegativeType.MEDIAN (builder.py:1358-1374): the per-image median of the non-label similarities, i.e.
// torch.nanquantile(q=0.5) = the mean of the two middle order statistics, found by a 4-pass radix select
// (8 key bits per pass) that recomputes S instead of storing it.  sel[b] = {prefix0, prefix1, rank0, rank1,
// hist0[256], hist1[256]}.
__device__ __forceinline__ uint32_t sortable(float s) {
  const uint32_t u = __float_as_uint(s);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float unsortable(uint32_t k) {
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

template <int C>
__global__ void __launch_bounds__(WTHREADS)
weighted_select_kernel(const __nv_bfloat16 *__restrict__ qhat_t, const __nv_bfloat16 *__restrict__ khat_t,
                       const float *__restrict__ mask_a, const float *__restrict__ mask_b, int *__restrict__ sel,
                       int P, int pass) {
  extern __shared__ __align__(16) float sm[];
  float *ks = sm;
  float *qs = sm + WT * (C + 4);
  __shared__ float a_s[WT];
  __shared__ int hist[2][256];
  const int b = blockIdx.y, y0 = blockIdx.x * WT, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const __nv_bfloat16 *qb = qhat_t + (long long)b * P * C, *kb = khat_t + (long long)b * P * C;
  int *st = sel + (long long)b * SEL_INTS;
  for (int i = threadIdx.x; i < 512; i += WTHREADS) (&hist[0][0])[i] = 0;
  load_rows<C>(kb, y0, P, ks, C + 1);
  const int y = y0 + lane;
  const bool vy = y < P;
  const float by = vy ? mask_b[(long long)b * P + y] : 0.f;
  const uint32_t pre0 = (uint32_t)st[0], pre1 = (uint32_t)st[1];
  const int shift = 24 - 8 * pass;
  for (int x0 = 0; x0 < P; x0 += WT) {
    __syncthreads();
    load_rows<C>(qb, x0, P, qs, C);
    if (threadIdx.x < WT) a_s[threadIdx.x] = x0 + threadIdx.x < P ? mask_a[(long long)b * P + x0 + threadIdx.x] : 0.f;
    __syncthreads();
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    const float *kr = ks + lane * (C + 1);
#pragma unroll 4
    for (int c = 0; c < C; c += 4) {
      const float k0 = kr[c], k1 = kr[c + 1], k2 = kr[c + 2], k3 = kr[c + 3];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float4 q4 = *reinterpret_cast<const float4 *>(qs + (warp * 4 + i) * C + c);
        acc[i] = fmaf(q4.x, k0, fmaf(q4.y, k1, fmaf(q4.z, k2, fmaf(q4.w, k3, acc[i]))));
      }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int xl = warp * 4 + i;
      if (vy && x0 + xl < P && !(a_s[xl] != 0.f && by != 0.f)) {       // non-label pairs only
        const uint32_t key = sortable(acc[i]);
        const uint32_t hi = pass == 0 ? 0u : key >> (shift + 8);
        if (hi == pre0) atomicAdd(&hist[0][(key >> shift) & 255u], 1);
        if (hi == pre1) atomicAdd(&hist[1][(key >> shift) & 255u], 1);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 512; i += WTHREADS) {
    const int v = (&hist[0][0])[i];
    if (v) atomicAdd(st + 4 + i, v);
  }
}

// Picks the bin holding each target rank, extends the prefixes, clears the histograms; after the last pass
// writes the centre = v0 + frac (v1 - v0) into wstats[b][0].
__global__ void __launch_bounds__(32)
weighted_pick_kernel(int *__restrict__ sel, float *__restrict__ wstats, int pass) {
  int *st = sel + (long long)blockIdx.x * SEL_INTS;
  const int lane = threadIdx.x;
  if (lane < 2) {
    const int *h = st + 4 + 256 * lane;
    int rank = st[2 + lane], cum = 0, bin = 255;
    for (int i = 0; i < 256; ++i) {
      const int c = h[i];
      if (rank < cum + c) { bin = i; break; }
      cum += c;
    }
    st[lane] = (int)(((uint32_t)st[lane] << 8) | (uint32_t)bin);
    st[2 + lane] = rank - cum;
  }
  __syncwarp();
  for (int i = lane; i < 512; i += 32) st[4 + i] = 0;
  __syncwarp();
  if (pass == 3 && lane == 0) {
    const float v0 = unsortable((uint32_t)st[0]), v1 = unsortable((uint32_t)st[1]);
    wstats[3 * blockIdx.x] = v0 + 0.5f * (v1 - v0);     // identical ranks (odd count) give v0
  }
}

// Forward: CTA = 32 key pixels y of one image, all query pixels x in chunks of 32.  Thread (warp, lane):
// key y0+lane, query rows 4*warp..4*warp+3 of the chunk.  lse[y] = log sum_x exp(z[x,y]) with
// z = w*t(S)*inv_tau; partial[b,tile] = sum_y b_y (A lse_y - sum_x a_x z[x,y]); best_v/best_i[b,tile] = the
// tile's largest z and its flat index x*P+y (acc_seg, builder.py:1442-1448, on the weighted logits).
template <int C>
__global__ void __launch_bounds__(WTHREADS)
weighted_lse_kernel(const __nv_bfloat16 *__restrict__ qhat_t, const __nv_bfloat16 *__restrict__ khat_t,
                    const float *__restrict__ mask_a, const float *__restrict__ mask_b,
                    const long long *__restrict__ pid_a, const long long *__restrict__ pid_b,
                    const long long *__restrict__ rid_a, const long long *__restrict__ rid_b,
                    const float *__restrict__ wstats, float *__restrict__ lse, float *__restrict__ partial,
                    float *__restrict__ best_v, int *__restrict__ best_i, WParams p) {
  extern __shared__ __align__(16) float sm[];
  float *ks = sm;                        // [32][C+1]
  float *qs = sm + WT * (C + 4);         // [32][C], 16-byte aligned rows
  __shared__ long long pa_s[WT], ra_s[WT];
  __shared__ float a_s[WT];
  __shared__ float red[3][WTHREADS / 32][WT];
  const int b = blockIdx.y, y0 = blockIdx.x * WT, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int P = p.P;
  const __nv_bfloat16 *qb = qhat_t + (long long)b * P * C, *kb = khat_t + (long long)b * P * C;
  load_rows<C>(kb, y0, P, ks, C + 1);
  const int y = y0 + lane;
  const bool vy = y < P;
  const float by = vy ? mask_b[(long long)b * P + y] : 0.f;
  const long long pb = vy ? pid_b[id_index(p, b, y)] : 0, rb = vy ? rid_b[id_index(p, b, y)] : 0;
  const float centre = p.neg_type >= 2 ? wstats[3 * b] : 0.f, A = wstats[3 * b + 1];
  float m = -INFINITY, l = 0.f, pos = 0.f;
  Best best{-INFINITY, 0x7fffffff};
  for (int x0 = 0; x0 < P; x0 += WT) {
    __syncthreads();
    load_rows<C>(qb, x0, P, qs, C);
    if (threadIdx.x < WT) {
      const int x = x0 + threadIdx.x;
      a_s[threadIdx.x] = x < P ? mask_a[(long long)b * P + x] : 0.f;
      pa_s[threadIdx.x] = x < P ? pid_a[id_index(p, b, x)] : 0;
      ra_s[threadIdx.x] = x < P ? rid_a[id_index(p, b, x)] : 0;
    }
    __syncthreads();
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    const float *kr = ks + lane * (C + 1);
#pragma unroll 4
    for (int c = 0; c < C; c += 4) {
      const float k0 = kr[c], k1 = kr[c + 1], k2 = kr[c + 2], k3 = kr[c + 3];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float4 q4 = *reinterpret_cast<const float4 *>(qs + (warp * 4 + i) * C + c);
        acc[i] = fmaf(q4.x, k0, fmaf(q4.y, k1, fmaf(q4.z, k2, fmaf(q4.w, k3, acc[i]))));
      }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int xl = warp * 4 + i;
      if (vy && x0 + xl < P) {
        const float ax = a_s[xl];
        float t, dt;
        squash(acc[i], ax != 0.f && by != 0.f, centre, p, t, dt);
        const float z = pair_weight(pa_s[xl], pb, ra_s[xl], rb, p) * t * p.inv_tau;
        const float mn = fmaxf(m, z);
        l = l * expf(m - mn) + expf(z - mn);
        m = mn;
        pos = fmaf(ax, z, pos);
        take(best, z, (x0 + xl) * P + y);
      }
    }
  }
  red[0][warp][lane] = m;
  red[1][warp][lane] = l;
  red[2][warp][lane] = pos;
  __syncthreads();
  if (warp == 0) {
    float M = -INFINITY, L = 0.f, Ps = 0.f;
#pragma unroll
    for (int w = 0; w < WTHREADS / 32; ++w) M = fmaxf(M, red[0][w][lane]);
#pragma unroll
    for (int w = 0; w < WTHREADS / 32; ++w) {
      const float mw = red[0][w][lane];
      if (mw > -INFINITY) L += red[1][w][lane] * expf(mw - M);
      Ps += red[2][w][lane];
    }
    const float ls = M + logf(L);
    if (vy) lse[(long long)b * P + y] = ls;
    float contrib = (vy && by != 0.f) ? by * (A * ls - Ps) : 0.f;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) contrib += __shfl_down_sync(0xffffffffu, contrib, off);
    if (lane == 0) partial[(long long)b * gridDim.x + blockIdx.x] = contrib;
  }
  // arg-max of the tile
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    const float v = __shfl_down_sync(0xffffffffu, best.v, off);
    const int i = __shfl_down_sync(0xffffffffu, best.i, off);
    take(best, v, i);
  }
  __syncthreads();                       // red[] is free again
  if (lane == 0) { red[0][warp][0] = best.v; red[1][warp][0] = __int_as_float(best.i); }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < WTHREADS / 32; ++w) take(best, red[0][w][0], __float_as_int(red[1][w][0]));
    best_v[(long long)b * gridDim.x + blockIdx.x] = best.v;
    best_i[(long long)b * gridDim.x + blockIdx.x] = best.i;
  }
}

// dense_n[b] = sum_tiles partial / (A * Bm): builder.py:1434-1437 for one image (0/0 -> NaN as the reference);
// acc_seg[0] = 100 * mean_b mask_a[x*] mask_b[y*] at the image's arg-max pair.  One CTA.
__global__ void __launch_bounds__(WTHREADS)
weighted_finalize_kernel(const float *__restrict__ partial, const float *__restrict__ best_v,
                         const int *__restrict__ best_i, const float *__restrict__ wstats,
                         const float *__restrict__ mask_a, const float *__restrict__ mask_b,
                         float *__restrict__ dense_n, float *__restrict__ acc_seg, int B, int P, int tiles) {
  __shared__ float red[WTHREADS / 32];
  float hits = 0.f;
  for (int b = threadIdx.x; b < B; b += WTHREADS) {
    float s = 0.f;
    Best best{-INFINITY, 0x7fffffff};
    for (int t = 0; t < tiles; ++t) {
      s += partial[(long long)b * tiles + t];
      take(best, best_v[(long long)b * tiles + t], best_i[(long long)b * tiles + t]);
    }
    dense_n[b] = s / (wstats[3 * b + 1] * wstats[3 * b + 2]);
    if (best.i != 0x7fffffff) {
      const int x = best.i / P, y = best.i - x * P;
      hits += mask_a[(long long)b * P + x] * mask_b[(long long)b * P + y];
    }
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) hits += __shfl_down_sync(0xffffffffu, hits, off);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = hits;
  __syncthreads();
  if (threadIdx.x == 0 && acc_seg) {
    float t = 0.f;
    for (int w = 0; w < WTHREADS / 32; ++w) t += red[w];
    acc_seg[0] = t * (100.f / (float)B);
  }
}

// Backward: CTA = 32 query pixels x of one image, all key pixels y in chunks of 32.
//   phase 1 (lane = x, warp -> 4 keys): g[x,y] = b_y (exp(z - lse_y) - a_x/A) * w * dt/ds
//   phase 2 (warp -> 4 rows x, lane -> channels lane+32j): acc[x,:] += g[x,y] khat[y,:]
//   epilogue: dqhat = alpha*acc + a_x v1 + (1-a_x) v2, then the per-pixel normalize backward, like
//   cp2_dense_grad; dq leaves through an smem transpose so the [B,C,P] stores run along P.
template <int C>
__global__ void __launch_bounds__(WTHREADS)
weighted_grad_kernel(const __nv_bfloat16 *__restrict__ qhat_t, const __nv_bfloat16 *__restrict__ khat_t,
                     const float *__restrict__ mask_a, const float *__restrict__ mask_b,
                     const long long *__restrict__ pid_a, const long long *__restrict__ pid_b,
                     const long long *__restrict__ rid_a, const long long *__restrict__ rid_b,
                     const float *__restrict__ lse, const float *__restrict__ wstats,
                     const float *__restrict__ inv_norm, const float *__restrict__ alpha,
                     const float *__restrict__ v1, const float *__restrict__ v2,
                     const float *__restrict__ gscale, float *__restrict__ dq, WParams p) {
  constexpr int NJ = C / 32;
  extern __shared__ __align__(16) float sm[];
  float *qs = sm;                        // [32][C+1]   this CTA's query rows
  float *kc = sm + WT * (C + 4);         // [32][C]     key chunk; reused as the [C][33] output tile
  float *gs = kc + (WT + 1) * C;         // [32][33]
  __shared__ long long pb_s[WT], rb_s[WT];
  __shared__ float b_s[WT], lse_s[WT], a_s[WT], inr_s[WT];
  const int b = blockIdx.y, x0 = blockIdx.x * WT, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int P = p.P;
  const __nv_bfloat16 *qb = qhat_t + (long long)b * P * C, *kb = khat_t + (long long)b * P * C;
  load_rows<C>(qb, x0, P, qs, C + 1);
  const int x = x0 + lane;
  const bool vx = x < P;
  const float ax = vx ? mask_a[(long long)b * P + x] : 0.f;
  const long long pa = vx ? pid_a[id_index(p, b, x)] : 0, ra = vx ? rid_a[id_index(p, b, x)] : 0;
  if (warp == 0) {
    a_s[lane] = ax;
    inr_s[lane] = vx ? inv_norm[(long long)b * P + x] : 0.f;
  }
  const float centre = p.neg_type >= 2 ? wstats[3 * b] : 0.f, a_over_A = ax / wstats[3 * b + 1];
  float acc[4][NJ];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < NJ; ++j) acc[i][j] = 0.f;
  for (int y0 = 0; y0 < P; y0 += WT) {
    __syncthreads();
    load_rows<C>(kb, y0, P, kc, C);
    if (threadIdx.x < WT) {
      const int y = y0 + threadIdx.x;
      const bool vy = y < P;
      b_s[threadIdx.x] = vy ? mask_b[(long long)b * P + y] : 0.f;
      lse_s[threadIdx.x] = vy ? lse[(long long)b * P + y] : 0.f;
      pb_s[threadIdx.x] = vy ? pid_b[id_index(p, b, y)] : 0;
      rb_s[threadIdx.x] = vy ? rid_b[id_index(p, b, y)] : 0;
    }
    __syncthreads();
    float s[4] = {0.f, 0.f, 0.f, 0.f};
    const float *qr = qs + lane * (C + 1);
#pragma unroll 4
    for (int c = 0; c < C; c += 4) {
      const float q0 = qr[c], q1 = qr[c + 1], q2 = qr[c + 2], q3 = qr[c + 3];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float4 k4 = *reinterpret_cast<const float4 *>(kc + (warp * 4 + i) * C + c);
        s[i] = fmaf(q0, k4.x, fmaf(q1, k4.y, fmaf(q2, k4.z, fmaf(q3, k4.w, s[i]))));
      }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int yl = warp * 4 + i;
      const float by = b_s[yl];
      float g = 0.f;
      if (vx && by != 0.f) {             // b_y = 0 (or y >= P): no loss term on this column
        float t, dt;
        squash(s[i], ax != 0.f, centre, p, t, dt);
        const float w = pair_weight(pa, pb_s[yl], ra, rb_s[yl], p);
        g = by * (expf(w * t * p.inv_tau - lse_s[yl]) - a_over_A) * w * dt;
      }
      gs[lane * (WT + 1) + yl] = g;
    }
    __syncthreads();
#pragma unroll 4
    for (int yl = 0; yl < WT; ++yl) {
      float kv[NJ];
#pragma unroll
      for (int j = 0; j < NJ; ++j) kv[j] = kc[yl * C + lane + 32 * j];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float g = gs[(warp * 4 + i) * (WT + 1) + yl];
#pragma unroll
        for (int j = 0; j < NJ; ++j) acc[i][j] = fmaf(g, kv[j], acc[i][j]);
      }
    }
  }
  __syncthreads();                       // kc is free: becomes the output tile [C][33]
  const float gsc = gscale ? gscale[0] : 1.f;
  const float al = alpha[b] * gsc;
  float va[NJ], vb[NJ];
#pragma unroll
  for (int j = 0; j < NJ; ++j) {
    const float u2 = v2 ? v2[(long long)b * C + lane + 32 * j] * gsc : 0.f;
    va[j] = v1[(long long)b * C + lane + 32 * j] * gsc - u2;
    vb[j] = u2;
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int xl = warp * 4 + i;
    const float a = a_s[xl];
    float gq[NJ], dot = 0.f;
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      gq[j] = fmaf(al, acc[i][j], fmaf(a, va[j], vb[j]));
      dot = fmaf(qs[xl * (C + 1) + lane + 32 * j], gq[j], dot);
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, off);
#pragma unroll
    for (int j = 0; j < NJ; ++j)
      kc[(lane + 32 * j) * (WT + 1) + xl] = (gq[j] - qs[xl * (C + 1) + lane + 32 * j] * dot) * inr_s[xl];
  }
  __syncthreads();
  for (int i = threadIdx.x; i < C * WT; i += WTHREADS) {
    const int c = i >> 5, xl = i & 31;
    if (x0 + xl < P) dq[((long long)b * C + c) * P + x0 + xl] = kc[c * (WT + 1) + xl];
  }
}

size_t lse_smem(int C) { return (size_t)(WT * (C + 4) + WT * C) * sizeof(float); }
size_t grad_smem(int C) { return (size_t)(WT * (C + 4) + (WT + 1) * C + WT * (WT + 1)) * sizeof(float); }

}  // namespace

// Pre-pass shared by the SIMT and the tensor-core forward: wstats / key_count, and for NegativeType.MEDIAN the
// 4-pass radix select of the centre.  scratch: [B, cp2_dense_weighted_tiles(P)] floats.
int weighted::prepass(const __nv_bfloat16 *q, const __nv_bfloat16 *k, const float *mask_a, const float *mask_b,
                      const float *pos_q, const float *neg_q, const float *pos_k, const float *neg_k, float *wstats,
                      int32_t *key_count, float *scratch, int neg_type, int B, int C, int P, cudaStream_t st) {
  const int tiles = cp2_cdiv(P, WT);
  int *sel = neg_type == 3 ? reinterpret_cast<int *>(scratch + 3 * (size_t)B * tiles) : nullptr;
  weighted_stats_kernel<<<B, 128, 0, st>>>(mask_a, mask_b, pos_q, neg_q, pos_k, neg_k, wstats, key_count, sel, C, P);
  CP2_LAUNCH_CHECK();
  const dim3 grid(tiles, B);
  if (sel) {
    if (C == 128) CP2_CUDA(cudaFuncSetAttribute(weighted_select_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lse_smem(128)));
    else CP2_CUDA(cudaFuncSetAttribute(weighted_select_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lse_smem(256)));
    for (int pass = 0; pass < 4; ++pass) {
      if (C == 128) weighted_select_kernel<128><<<grid, WTHREADS, lse_smem(128), st>>>(q, k, mask_a, mask_b, sel, P, pass);
      else weighted_select_kernel<256><<<grid, WTHREADS, lse_smem(256), st>>>(q, k, mask_a, mask_b, sel, P, pass);
      CP2_LAUNCH_CHECK();
      weighted_pick_kernel<<<B, 32, 0, st>>>(sel, wstats, pass);
      CP2_LAUNCH_CHECK();
    }
  }
  return CP2_OK;
}

int weighted::finalize(const float *partial, const float *best_v, const int *best_i, const float *wstats,
                       const float *mask_a, const float *mask_b, float *dense_n, float *acc_seg, int B, int P, int tiles,
                       cudaStream_t st) {
  weighted_finalize_kernel<<<1, WTHREADS, 0, st>>>(partial, best_v, best_i, wstats, mask_a, mask_b, dense_n, acc_seg, B, P,
                                                   tiles);
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}

extern "C" int cp2_dense_weighted_tiles(int P) { return 3 * cp2_cdiv(P, WT) + SEL_INTS; }

extern "C" int cp2_dense_weighted_lse(const void *qhat_t, const void *khat_t, const float *mask_a, const float *mask_b,
                                      const int64_t *pixel_ids_a, const int64_t *pixel_ids_b,
                                      const int64_t *region_ids_a, const int64_t *region_ids_b, const float *pos_q,
                                      const float *neg_q, const float *pos_k, const float *neg_k, float *lse,
                                      float *dense_n, float *acc_seg, float *wstats, int32_t *key_count,
                                      float *scratch,
                                      float w_pix, float w_reg, float w_not, int neg_type, float neg_scale,
                                      float inv_tau, int B, int C, int P, int ids_h, int ids_w, int os, void *stream) {
  CP2_REQUIRE(B > 0 && P > 0 && (C == 128 || C == 256), "cp2_dense_weighted_lse: bad shape B=%d C=%d P=%d", B, C, P);
  CP2_REQUIRE(qhat_t && khat_t && mask_a && mask_b && pixel_ids_a && pixel_ids_b && region_ids_a && region_ids_b &&
              pos_q && neg_q && pos_k && neg_k && lse && dense_n && wstats && key_count && scratch,
              "cp2_dense_weighted_lse: null pointer");
  WParams p;
  if (int rc = make_params(p, w_pix, w_reg, w_not, neg_type, neg_scale, inv_tau, P, ids_h, ids_w, os)) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int tiles = cp2_cdiv(P, WT);
  const auto *pa = reinterpret_cast<const long long *>(pixel_ids_a), *pb = reinterpret_cast<const long long *>(pixel_ids_b);
  const auto *ra = reinterpret_cast<const long long *>(region_ids_a), *rb = reinterpret_cast<const long long *>(region_ids_b);
  const auto *q = static_cast<const __nv_bfloat16 *>(qhat_t), *k = static_cast<const __nv_bfloat16 *>(khat_t);
  if (int rc = weighted::prepass(q, k, mask_a, mask_b, pos_q, neg_q, pos_k, neg_k, wstats, key_count, scratch, neg_type, B, C, P, st))
    return rc;
  const dim3 grid(tiles, B);
  float *bv = scratch + (size_t)B * tiles;
  int *bi = reinterpret_cast<int *>(scratch + 2 * (size_t)B * tiles);
  if (C == 128) {
    CP2_CUDA(cudaFuncSetAttribute(weighted_lse_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lse_smem(128)));
    weighted_lse_kernel<128><<<grid, WTHREADS, lse_smem(128), st>>>(q, k, mask_a, mask_b, pa, pb, ra, rb, wstats, lse, scratch, bv, bi, p);
  } else {
    CP2_CUDA(cudaFuncSetAttribute(weighted_lse_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lse_smem(256)));
    weighted_lse_kernel<256><<<grid, WTHREADS, lse_smem(256), st>>>(q, k, mask_a, mask_b, pa, pb, ra, rb, wstats, lse, scratch, bv, bi, p);
  }
  CP2_LAUNCH_CHECK();
  weighted_finalize_kernel<<<1, WTHREADS, 0, st>>>(scratch, bv, bi, wstats, mask_a, mask_b, dense_n, acc_seg, B, P,
                                                   tiles);
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}

extern "C" int cp2_dense_weighted_grad(const void *qhat_t, const void *khat_t, const float *mask_a, const float *mask_b,
                                       const int64_t *pixel_ids_a, const int64_t *pixel_ids_b,
                                       const int64_t *region_ids_a, const int64_t *region_ids_b, const float *lse,
                                       const float *wstats, const float *inv_norm, const float *alpha, const float *v1,
                                       const float *v2, const float *gscale, float *dq, float w_pix, float w_reg,
                                       float w_not, int neg_type, float neg_scale, float inv_tau, int B, int C, int P,
                                       int ids_h, int ids_w, int os, void *stream) {
  CP2_REQUIRE(B > 0 && P > 0 && (C == 128 || C == 256), "cp2_dense_weighted_grad: bad shape B=%d C=%d P=%d", B, C, P);
  CP2_REQUIRE(qhat_t && khat_t && mask_a && mask_b && pixel_ids_a && pixel_ids_b && region_ids_a && region_ids_b &&
              lse && wstats && inv_norm && alpha && v1 && dq, "cp2_dense_weighted_grad: null pointer");
  WParams p;
  if (int rc = make_params(p, w_pix, w_reg, w_not, neg_type, neg_scale, inv_tau, P, ids_h, ids_w, os)) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const auto *pa = reinterpret_cast<const long long *>(pixel_ids_a), *pb = reinterpret_cast<const long long *>(pixel_ids_b);
  const auto *ra = reinterpret_cast<const long long *>(region_ids_a), *rb = reinterpret_cast<const long long *>(region_ids_b);
  const auto *q = static_cast<const __nv_bfloat16 *>(qhat_t), *k = static_cast<const __nv_bfloat16 *>(khat_t);
  const dim3 grid(cp2_cdiv(P, WT), B);
  if (C == 128) {
    CP2_CUDA(cudaFuncSetAttribute(weighted_grad_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)grad_smem(128)));
    weighted_grad_kernel<128><<<grid, WTHREADS, grad_smem(128), st>>>(q, k, mask_a, mask_b, pa, pb, ra, rb, lse, wstats,
                                                                      inv_norm, alpha, v1, v2, gscale, dq, p);
  } else {
    CP2_CUDA(cudaFuncSetAttribute(weighted_grad_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)grad_smem(256)));
    weighted_grad_kernel<256><<<grid, WTHREADS, grad_smem(256), st>>>(q, k, mask_a, mask_b, pa, pb, ra, rb, lse, wstats,
                                                                      inv_norm, alpha, v1, v2, gscale, dq, p);
  }
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}
