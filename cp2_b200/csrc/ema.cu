// K4 — momentum (EMA) update of the key encoder (reference builder.py:557-567).
// One launch over every parameter tensor.  HBM-bound: 12 bytes per parameter
// (read k, read q, write k).  Work unit = one chunk of CP2_EMA_CHUNK elements of one
// tensor; the host precomputes the (tensor, chunk) table once per model.
// Arithmetic is the reference's three separately rounded fp32 ops (no FMA).
#include "common.cuh"

namespace {

__device__ __forceinline__ float ema1(float k, float q, float m, float om) {
  return __fadd_rn(__fmul_rn(k, m), __fmul_rn(q, om));
}

__global__ void __launch_bounds__(256)
ema_kernel(const float *const *__restrict__ q_ptrs, float *const *__restrict__ k_ptrs,
           const int64_t *__restrict__ numel, const int32_t *__restrict__ chunk_tensor,
           const int32_t *__restrict__ chunk_index, int n_chunks, float m, float om) {
  for (int ch = blockIdx.x; ch < n_chunks; ch += gridDim.x) {
    const int t = chunk_tensor[ch];
    const long long start = (long long)chunk_index[ch] * CP2_EMA_CHUNK;
    const long long n = numel[t];
    const long long len = (n - start < CP2_EMA_CHUNK) ? (n - start) : CP2_EMA_CHUNK;
    const float *__restrict__ q = q_ptrs[t] + start;
    float *__restrict__ k = k_ptrs[t] + start;
    const bool vec = ((reinterpret_cast<uintptr_t>(q) | reinterpret_cast<uintptr_t>(k)) & 15u) == 0;
    if (vec) {
      const int nv = (int)(len >> 2);
      // CP2_EMA_CHUNK/4 = 1024 float4 per chunk = 4 per thread: issue all loads first
      float4 kv[4], qv[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int i = threadIdx.x + j * 256;
        if (i < nv) {
          kv[j] = reinterpret_cast<const float4 *>(k)[i];
          qv[j] = __ldcs(reinterpret_cast<const float4 *>(q) + i);
        }
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int i = threadIdx.x + j * 256;
        if (i < nv) {
          float4 r;
          r.x = ema1(kv[j].x, qv[j].x, m, om);
          r.y = ema1(kv[j].y, qv[j].y, m, om);
          r.z = ema1(kv[j].z, qv[j].z, m, om);
          r.w = ema1(kv[j].w, qv[j].w, m, om);
          reinterpret_cast<float4 *>(k)[i] = r;
        }
      }
      for (int i = (nv << 2) + threadIdx.x; i < len; i += 256) k[i] = ema1(k[i], q[i], m, om);
    } else {
      for (int i = threadIdx.x; i < len; i += 256) k[i] = ema1(k[i], q[i], m, om);
    }
  }
}

}  // namespace

extern "C" int cp2_ema(const float *const *q_ptrs, float *const *k_ptrs, const int64_t *numel,
                       const int32_t *chunk_tensor, const int32_t *chunk_index, int n_chunks,
                       float m, float one_minus_m, void *stream) {
  CP2_REQUIRE(n_chunks >= 0, "cp2_ema: n_chunks=%d", n_chunks);
  if (n_chunks == 0) return CP2_OK;
  CP2_REQUIRE(q_ptrs && k_ptrs && numel && chunk_tensor && chunk_index, "cp2_ema: null table");
  const int cap = cp2_num_sms() * 8;          // 8 resident CTAs of 256 threads per SM
  const int blocks = n_chunks < cap ? n_chunks : cap;
  ema_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(q_ptrs, k_ptrs, numel, chunk_tensor,
                                                                     chunk_index, n_chunks, m, one_minus_m);
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}
