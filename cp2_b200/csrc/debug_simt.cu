// TEST-ONLY CUDA-core reference kernels (include/cp2_b200_debug.h).  Not on the
// product path; see the header.
#include "common.cuh"
#include "cp2_b200_debug.h"

namespace {

template <int C>
__device__ __forceinline__ void load_rows(__nv_bfloat16 (*dst)[C + 2], const __nv_bfloat16 *__restrict__ src,
                                          int row0, int P) {
  // 32 rows x C bf16, zero-filled past P
  for (int i = threadIdx.x; i < 32 * (C / 2); i += blockDim.x) {
    const int r = i / (C / 2), cp = i % (C / 2);
    __nv_bfloat162 v = __floats2bfloat162_rn(0.f, 0.f);
    if (row0 + r < P) v = reinterpret_cast<const __nv_bfloat162 *>(src + (long long)(row0 + r) * C)[cp];
    reinterpret_cast<__nv_bfloat162 *>(&dst[r][0])[cp] = v;
  }
}

template <int C>
__device__ __forceinline__ float dot_rows(const __nv_bfloat16 *a, const __nv_bfloat16 *b) {
  float acc = 0.f;
#pragma unroll 8
  for (int c = 0; c < C; c += 2) {
    const float2 x = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162 *>(a + c));
    const float2 y = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162 *>(b + c));
    acc = fmaf(x.x, y.x, acc);
    acc = fmaf(x.y, y.y, acc);
  }
  return acc;
}

template <int C>
__global__ void __launch_bounds__(256)
dense_lse_simt(const __nv_bfloat16 *__restrict__ q, const __nv_bfloat16 *__restrict__ k,
               float *__restrict__ lse, float inv_tau, int P) {
  __shared__ __align__(16) __nv_bfloat16 Ks[32][C + 2];
  __shared__ __align__(16) __nv_bfloat16 Qs[32][C + 2];
  __shared__ float red[8][32];
  const int b = blockIdx.y, y0 = blockIdx.x * 32;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  q += (long long)b * P * C;
  k += (long long)b * P * C;
  load_rows<C>(Ks, k, y0, P);
  float colsum = 0.f;
  for (int x0 = 0; x0 < P; x0 += 32) {
    __syncthreads();
    load_rows<C>(Qs, q, x0, P);
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int x = warp * 4 + i;
      if (x0 + x < P) colsum += expf(dot_rows<C>(&Qs[x][0], &Ks[lane][0]) * inv_tau);
    }
  }
  red[warp][lane] = colsum;
  __syncthreads();
  if (warp == 0 && y0 + lane < P) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += red[w][lane];
    lse[(long long)b * P + y0 + lane] = logf(s);
  }
}

template <int C>
__global__ void __launch_bounds__(256)
dense_grad_simt(const __nv_bfloat16 *__restrict__ q, const __nv_bfloat16 *__restrict__ k,
                const float *__restrict__ lse, const float *__restrict__ mask_a,
                const float *__restrict__ mask_b, const float *__restrict__ inv_norm,
                const float *__restrict__ alpha, const float *__restrict__ v1,
                const float *__restrict__ v2, float *__restrict__ dq, float inv_tau, int P) {
  constexpr int CG = C / 8;
  __shared__ __align__(16) __nv_bfloat16 Ks[32][C + 2];
  __shared__ __align__(16) __nv_bfloat16 Qs[32][C + 2];
  __shared__ float Ps[32][33];
  __shared__ float lse_s[32], mb_s[32];
  const int b = blockIdx.y, x0 = blockIdx.x * 32;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int xr = threadIdx.x >> 3, cg = threadIdx.x & 7;
  q += (long long)b * P * C;
  k += (long long)b * P * C;
  load_rows<C>(Qs, q, x0, P);
  float acc[CG];
#pragma unroll
  for (int j = 0; j < CG; ++j) acc[j] = 0.f;
  for (int y0 = 0; y0 < P; y0 += 32) {
    __syncthreads();
    load_rows<C>(Ks, k, y0, P);
    if (threadIdx.x < 32) {
      const int y = y0 + threadIdx.x;
      lse_s[threadIdx.x] = y < P ? lse[(long long)b * P + y] : 0.f;
      mb_s[threadIdx.x] = y < P ? mask_b[(long long)b * P + y] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int x = warp * 4 + i;
      float p = 0.f;
      if (x0 + x < P && y0 + lane < P && mb_s[lane] != 0.f)
        p = mb_s[lane] * expf(dot_rows<C>(&Qs[x][0], &Ks[lane][0]) * inv_tau - lse_s[lane]);
      Ps[x][lane] = p;
    }
    __syncthreads();
    for (int y = 0; y < 32; ++y) {
      const float p = Ps[xr][y];
#pragma unroll
      for (int j = 0; j < CG; ++j) acc[j] = fmaf(p, __bfloat162float(Ks[y][cg * CG + j]), acc[j]);
    }
  }
  const int x = x0 + xr;
  const bool valid = x < P;
  const float a = valid ? mask_a[(long long)b * P + x] : 0.f;
  const float al = alpha[b];
  float dot = 0.f;
#pragma unroll
  for (int j = 0; j < CG; ++j) {
    const int c = cg * CG + j;
    const float w2 = v2 ? v2[(long long)b * C + c] : 0.f;
    acc[j] = al * acc[j] + a * v1[(long long)b * C + c] + (1.f - a) * w2;
    dot = fmaf(__bfloat162float(Qs[xr][c]), acc[j], dot);
  }
  dot += __shfl_xor_sync(0xffffffffu, dot, 1);
  dot += __shfl_xor_sync(0xffffffffu, dot, 2);
  dot += __shfl_xor_sync(0xffffffffu, dot, 4);
  if (valid) {
    const float inr = inv_norm[(long long)b * P + x];
#pragma unroll
    for (int j = 0; j < CG; ++j) {
      const int c = cg * CG + j;
      dq[((long long)b * C + c) * P + x] = (acc[j] - __bfloat162float(Qs[xr][c]) * dot) * inr;
    }
  }
}

// queue tile [C][129] + e[32][128] in dynamic shared memory
__global__ void __launch_bounds__(128)
global_nce_simt(const float *__restrict__ q_pos, const float *__restrict__ queue, float *__restrict__ z,
                float *__restrict__ g, float inv_tau, int B, int C, int K) {
  extern __shared__ float sm[];
  float *tile = sm;                 // [C][129]
  float *es = sm + C * 129;         // [32][128]
  const int t = threadIdx.x, k0 = blockIdx.x * 128;
  for (int c = 0; c < C; ++c) tile[c * 129 + t] = (k0 + t < K) ? queue[(long long)c * K + k0 + t] : 0.f;
  __syncthreads();
  for (int b0 = 0; b0 < B; b0 += 32) {
    for (int bb = 0; bb < 32 && b0 + bb < B; ++bb) {
      const float *qp = q_pos + (long long)(b0 + bb) * C;
      float l = 0.f;
      for (int c = 0; c < C; ++c) l = fmaf(qp[c], tile[c * 129 + t], l);
      float e = (k0 + t < K) ? expf((l - 1.0f) * inv_tau) : 0.f;
      es[bb * 128 + t] = e;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) e += __shfl_xor_sync(0xffffffffu, e, o);
      if ((t & 31) == 0) atomicAdd(&z[b0 + bb], e);
    }
    __syncthreads();
    for (int c = t; c < C; c += 128) {
      for (int bb = 0; bb < 32 && b0 + bb < B; ++bb) {
        float s = 0.f;
        for (int kk = 0; kk < 128; ++kk) s = fmaf(es[bb * 128 + kk], tile[c * 129 + kk], s);
        atomicAdd(&g[(long long)(b0 + bb) * C + c], s);
      }
    }
    __syncthreads();
  }
}

}  // namespace

extern "C" int cp2_dbg_dense_lse_simt(const void *qhat_t, const void *khat_t, float *lse, float inv_tau,
                                      int B, int C, int P, void *stream) {
  CP2_REQUIRE(qhat_t && khat_t && lse, "cp2_dbg_dense_lse_simt: null pointer");
  CP2_REQUIRE(C == 128 || C == 256, "cp2_dbg_dense_lse_simt: C=%d", C);
  if (B == 0 || P == 0) return CP2_OK;
  dim3 grid(cp2_cdiv(P, 32), B);
  auto *q = static_cast<const __nv_bfloat16 *>(qhat_t);
  auto *k = static_cast<const __nv_bfloat16 *>(khat_t);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (C == 128) dense_lse_simt<128><<<grid, 256, 0, st>>>(q, k, lse, inv_tau, P);
  else dense_lse_simt<256><<<grid, 256, 0, st>>>(q, k, lse, inv_tau, P);
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}

extern "C" int cp2_dbg_dense_grad_simt(const void *qhat_t, const void *khat_t, const float *lse,
                                       const float *mask_a, const float *mask_b, const float *inv_norm,
                                       const float *alpha, const float *v1, const float *v2, float *dq,
                                       float inv_tau, int B, int C, int P, void *stream) {
  CP2_REQUIRE(qhat_t && khat_t && lse && mask_a && mask_b && inv_norm && alpha && v1 && dq,
              "cp2_dbg_dense_grad_simt: null pointer");
  CP2_REQUIRE(C == 128 || C == 256, "cp2_dbg_dense_grad_simt: C=%d", C);
  if (B == 0 || P == 0) return CP2_OK;
  dim3 grid(cp2_cdiv(P, 32), B);
  auto *q = static_cast<const __nv_bfloat16 *>(qhat_t);
  auto *k = static_cast<const __nv_bfloat16 *>(khat_t);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (C == 128)
    dense_grad_simt<128><<<grid, 256, 0, st>>>(q, k, lse, mask_a, mask_b, inv_norm, alpha, v1, v2, dq, inv_tau, P);
  else
    dense_grad_simt<256><<<grid, 256, 0, st>>>(q, k, lse, mask_a, mask_b, inv_norm, alpha, v1, v2, dq, inv_tau, P);
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}

extern "C" int cp2_dbg_global_nce_simt(const float *q_pos, const float *queue, float *z, float *g,
                                       float inv_tau, int B, int C, int K, void *stream) {
  CP2_REQUIRE(q_pos && queue && z && g, "cp2_dbg_global_nce_simt: null pointer");
  CP2_REQUIRE(C > 0 && C <= 256 && K > 0 && B >= 0, "cp2_dbg_global_nce_simt: bad shape");
  if (B == 0) return CP2_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CP2_CUDA(cudaMemsetAsync(z, 0, sizeof(float) * B, st));
  CP2_CUDA(cudaMemsetAsync(g, 0, sizeof(float) * (size_t)B * C, st));
  const size_t smem = sizeof(float) * ((size_t)C * 129 + 32 * 128);
  CP2_CUDA(cudaFuncSetAttribute(global_nce_simt, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  global_nce_simt<<<cp2_cdiv(K, 128), 128, smem, st>>>(q_pos, queue, z, g, inv_tau, B, C, K);
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}
