// f2 — the two loss-side pieces of forward_densecl that are not already
// covered by cp2_global_nce:
//   cp2_rows_nce        the ContrastiveHead (reference builder.py:150-176) for N rows whose negatives are a
//                       queue: takes z / G of cp2_global_nce, adds the positive logit, produces the mean
//                       cross-entropy and the gradient w.r.t. every row.  N = B for the global loss
//                       (builder.py:762-775), N = B*S^2 for the local one (builder.py:866-873, 906-908).
//   cp2_positive_keys   the vector every query pixel is pulled towards: the key pixel matched through the
//                       backbone features (builder.py:824-835), mixed with the key pixels that share the
//                       query pixel's id (builder.py:839-855), so that pos_local = q_local_x . kp_x and the
//                       [B,S^2,S^2] local similarity matrix, the gather and the masked sum never exist.
#include "common.cuh"

namespace {

constexpr int RT = 256;                 // threads: 8 rows per CTA, one warp per row

// terms[r] = -log softmax_0([pos, negs] / tau);  dq[r,:] = d mean(terms) / d rows[r,:]
__global__ void __launch_bounds__(RT)
rows_nce_kernel(const float *__restrict__ rows, const float *__restrict__ pos_keys, const float *__restrict__ z,
                const float *__restrict__ G, float *__restrict__ terms, float *__restrict__ dq, float inv_tau,
                int N, int C, int nz, long long zstride, int ng, long long gstride) {
  // z and G may arrive as partial sums (cp2_queue_sums): z[r] = sum_i z[i*zstride + r], G[r,:] likewise (fixed order)
  const int lane = threadIdx.x & 31, r = blockIdx.x * (RT / 32) + (threadIdx.x >> 5);
  if (r >= N) return;
  const float *q = rows + (long long)r * C, *kp = pos_keys + (long long)r * C, *g = G + (long long)r * C;
  float zr = 0.f;
  for (int i = 0; i < nz; ++i) zr += z[(long long)i * zstride + r];
  float lp = 0.f;
  for (int c = lane; c < C; c += 32) lp = fmaf(q[c], kp[c], lp);
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) lp += __shfl_xor_sync(0xffffffffu, lp, off);
  // shift 1 as in cp2_global_nce: z = sum_k exp((l_k - 1) / tau)
  const float e_pos = expf((lp - 1.f) * inv_tau);
  const float Z = zr + e_pos, rZ = 1.f / Z, scale = inv_tau / (float)N;
  if (lane == 0) terms[r] = logf(Z) + (1.f - lp) * inv_tau;
  if (dq)
    for (int c = lane; c < C; c += 32) {
      float gc = 0.f;
      for (int i = 0; i < ng; ++i) gc += g[(long long)i * gstride + c];
      dq[(long long)r * C + c] = (fmaf(e_pos, kp[c], gc) * rZ - kp[c]) * scale;
    }
}

// loss[0] = mean(terms), fixed summation order.  One CTA.
__global__ void __launch_bounds__(RT)
rows_mean_kernel(const float *__restrict__ terms, float *__restrict__ loss, int N) {
  __shared__ float red[RT];
  float s = 0.f;
  const int per = (N + RT - 1) / RT, lo = threadIdx.x * per;
  for (int i = lo; i < min(lo + per, N); ++i) s += terms[i];
  red[threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int i = 0; i < RT; ++i) t += red[i];
    loss[0] = t / (float)N;
  }
}

struct Lattice { int ws, ids_w, os; long long ids_hw; };
__device__ __forceinline__ long long lat_index(const Lattice &l, int b, int pix) {
  const int r = pix / l.ws, c = pix - r * l.ws;
  return (long long)b * l.ids_hw + (long long)(r * l.os + l.os / 2) * l.ids_w + c * l.os + l.os / 2;
}

// One warp per query pixel x of image b; lanes over channels.  kp[b*P+x, :] = k_local[b,:,idx[b,x]], mixed
// with lmbd * sum_{y: pix_k[y] == pix_q[x]} k_local[b,:,y] where such y exist.
__global__ void __launch_bounds__(RT)
positive_keys_kernel(const float *__restrict__ k_local, const long long *__restrict__ idx,
                     const long long *__restrict__ pix_q, const long long *__restrict__ pix_k,
                     float *__restrict__ kp, float lmbd, int C, int P, Lattice lat) {
  extern __shared__ long long pk_s[];   // [P] key-side pixel ids of this image
  const int b = blockIdx.y, lane = threadIdx.x & 31, x = blockIdx.x * (RT / 32) + (threadIdx.x >> 5);
  for (int y = threadIdx.x; y < P; y += RT) pk_s[y] = pix_k ? pix_k[lat_index(lat, b, y)] : 0;
  __syncthreads();
  if (x >= P) return;
  const float *kb = k_local + (long long)b * C * P;
  const long long m = idx[(long long)b * P + x];
  const long long id = pix_q ? pix_q[lat_index(lat, b, x)] : 0;
  float *dst = kp + ((long long)b * P + x) * C;
  for (int c0 = 0; c0 < C; c0 += 32) {
    const int c = c0 + lane;
    if (c >= C) break;
    const float km = kb[(long long)c * P + m];
    float coord = 0.f;
    bool any = false;
    if (pix_q && lmbd != 0.f)
      for (int y = 0; y < P; ++y)
        if (pk_s[y] == id) { coord += kb[(long long)c * P + y]; any = true; }
    dst[c] = any ? km * (1.f - lmbd) + coord * lmbd : km;
  }
}

}  // namespace

extern "C" int cp2_rows_nce(const float *rows, const float *pos_keys, const float *z, const float *G, float *loss,
                            float *terms, float *dq, float inv_tau, int N, int C, int nz, int64_t zstride, int ng,
                            int64_t gstride, void *stream) {
  CP2_REQUIRE(N > 0 && C > 0, "cp2_rows_nce: bad shape N=%d C=%d", N, C);
  CP2_REQUIRE(rows && pos_keys && z && G && loss && terms, "cp2_rows_nce: null pointer");
  CP2_REQUIRE(inv_tau > 0.f, "cp2_rows_nce: bad temperature");
  CP2_REQUIRE(nz >= 1 && ng >= 1, "cp2_rows_nce: bad partial counts nz=%d ng=%d", nz, ng);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  rows_nce_kernel<<<cp2_cdiv(N, RT / 32), RT, 0, st>>>(rows, pos_keys, z, G, terms, dq, inv_tau, N, C, nz,
                                                      (long long)zstride, ng, (long long)gstride);
  CP2_LAUNCH_CHECK();
  rows_mean_kernel<<<1, RT, 0, st>>>(terms, loss, N);
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}

extern "C" int cp2_positive_keys(const float *k_local, const int64_t *idx, const int64_t *pix_q,
                                 const int64_t *pix_k, float *kp, float lmbd_coordinate, int B, int C, int P,
                                 int ids_h, int ids_w, int os, void *stream) {
  CP2_REQUIRE(B > 0 && C > 0 && P > 0, "cp2_positive_keys: bad shape B=%d C=%d P=%d", B, C, P);
  CP2_REQUIRE(k_local && idx && kp, "cp2_positive_keys: null pointer");
  CP2_REQUIRE((pix_q == nullptr) == (pix_k == nullptr), "cp2_positive_keys: pix_q and pix_k go together");
  Lattice lat{1, 1, 1, 1};
  if (pix_q) {
    CP2_REQUIRE(os >= 1 && ids_h >= 1 && ids_w >= 1, "cp2_positive_keys: bad id map %dx%d stride %d", ids_h, ids_w, os);
    const int hs = (ids_h - os / 2 + os - 1) / os, ws = (ids_w - os / 2 + os - 1) / os;
    CP2_REQUIRE(hs * ws == P, "cp2_positive_keys: id maps %dx%d at stride %d give %dx%d pixels, P=%d", ids_h, ids_w,
                os, hs, ws, P);
    lat = Lattice{ws, ids_w, os, (long long)ids_h * ids_w};
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t smem = (size_t)P * sizeof(long long);
  CP2_REQUIRE(smem <= 200 * 1024, "cp2_positive_keys: P=%d too large", P);
  CP2_CUDA(cudaFuncSetAttribute(positive_keys_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  positive_keys_kernel<<<dim3(cp2_cdiv(P, RT / 32), B), RT, smem, st>>>(
      k_local, reinterpret_cast<const long long *>(idx), reinterpret_cast<const long long *>(pix_q),
      reinterpret_cast<const long long *>(pix_k), kp, lmbd_coordinate, C, P, lat);
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}
