// a2 + a4 — stride subsample of the id maps and the (masked) correlation IoU of two id maps
// (reference builder.py:1165-1186 `x[:, OS//2::OS, OS//2::OS]`, tools/correlation_mapping.py:103-138
// get_masked_iou, called through get_masked_correlation_map at builder.py:1203-1217, 1252-1257).
//
// Per image the reference concatenates both (sub-sampled) id maps, adds 1, zeroes the masked entries and
// runs torch.unique(return_counts=True) on the host side of a python loop:
//   union        = number of distinct non-zero values
//   intersection = number of distinct non-zero values that occur more than once
// Integer work, bit-exact.  One CTA per image builds an open-addressing hash set of the 2P (id+1) values in
// shared memory (64-bit keys claimed by atomicCAS, linear probing, load <= 0.5) with two occurrence counters per
// slot (all pixels / pixels under the masks): O(P) work per image instead of the (2P)^2 all-pairs scan of
// round 1, so P = 3136 costs what P = 196 did.
#include "common.cuh"

namespace {

__device__ __forceinline__ unsigned int hash64(unsigned long long v) {      // splitmix64 finaliser
  v ^= v >> 30; v *= 0xbf58476d1ce4e5b9ULL;
  v ^= v >> 27; v *= 0x94d049bb133111ebULL;
  v ^= v >> 31;
  return (unsigned int)v;
}

__global__ void __launch_bounds__(512)
id_iou_kernel(const long long *__restrict__ ids_a, const long long *__restrict__ ids_b,
              const float *__restrict__ mask_a, const float *__restrict__ mask_b, float *__restrict__ iou,
              float *__restrict__ iou_masked, int H, int W, int os, int hs, int ws, int slots) {
  extern __shared__ unsigned long long keys[];               // [slots] id + 1, 0 = empty
  unsigned int *cnt = reinterpret_cast<unsigned int *>(keys + slots);   // [slots] low 16: all, high 16: masked
  const int P = hs * ws, n = 2 * P, b = blockIdx.x, o = os / 2;
  for (int i = threadIdx.x; i < slots; i += blockDim.x) { keys[i] = 0ULL; cnt[i] = 0u; }
  __syncthreads();
  const unsigned int smask = (unsigned int)slots - 1u;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int p = i < P ? i : i - P;
    const int y = o + (p / ws) * os, x = o + (p % ws) * os;
    const long long *src = i < P ? ids_a : ids_b;
    const float *msk = i < P ? mask_a : mask_b;
    const unsigned long long v = (unsigned long long)(src[((long long)b * H + y) * W + x] + 1);
    if (v == 0ULL) continue;                                  // the reference drops the value 0 (id == -1)
    const unsigned int m = msk[(long long)b * P + p] != 0.f ? 1u : 0u;
    unsigned int s = hash64(v) & smask;
    while (true) {
      const unsigned long long prev = atomicCAS(&keys[s], 0ULL, v);
      if (prev == 0ULL || prev == v) break;
      s = (s + 1u) & smask;
    }
    atomicAdd(&cnt[s], 1u + (m << 16));
  }
  __syncthreads();
  int uni = 0, inter = 0, uni_m = 0, inter_m = 0;
  for (int i = threadIdx.x; i < slots; i += blockDim.x) {
    const unsigned int c = cnt[i], ca = c & 0xffffu, cm = c >> 16;
    uni += ca > 0u;
    inter += ca > 1u;
    uni_m += cm > 0u;
    inter_m += cm > 1u;
  }
  __shared__ int red[4][16];
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    uni += __shfl_down_sync(0xffffffffu, uni, off);
    inter += __shfl_down_sync(0xffffffffu, inter, off);
    uni_m += __shfl_down_sync(0xffffffffu, uni_m, off);
    inter_m += __shfl_down_sync(0xffffffffu, inter_m, off);
  }
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) { red[0][wid] = uni; red[1][wid] = inter; red[2][wid] = uni_m; red[3][wid] = inter_m; }
  __syncthreads();
  if (threadIdx.x == 0) {
    int t[4] = {0, 0, 0, 0};
    for (int k = 0; k < 4; ++k)
      for (int w = 0; w < 16; ++w) t[k] += red[k][w];
    // python: intersection / union in double, stored into a float32 tensor; union == 0 raises there -> NaN here
    iou[b] = t[0] ? (float)((double)t[1] / (double)t[0]) : __int_as_float(0x7fc00000);
    iou_masked[b] = t[2] ? (float)((double)t[3] / (double)t[2]) : __int_as_float(0x7fc00000);
  }
}

}  // namespace

extern "C" int cp2_id_iou(const int64_t *ids_a, const int64_t *ids_b, const float *mask_a_ds,
                          const float *mask_b_ds, float *iou, float *iou_masked, int B, int H, int W, int os,
                          void *stream) {
  CP2_REQUIRE(B >= 0 && H > 0 && W > 0 && os > 0, "cp2_id_iou: bad shape B=%d H=%d W=%d os=%d", B, H, W, os);
  if (B == 0) return CP2_OK;
  CP2_REQUIRE(ids_a && ids_b && mask_a_ds && mask_b_ds && iou && iou_masked, "cp2_id_iou: null pointer");
  const int o = os / 2;
  const int hs = H > o ? (H - o + os - 1) / os : 0;
  const int ws = W > o ? (W - o + os - 1) / os : 0;
  CP2_REQUIRE(hs > 0 && ws > 0, "cp2_id_iou: empty sub-sampled map (H=%d W=%d os=%d)", H, W, os);
  int slots = 1024;
  while (slots < 4 * hs * ws && (size_t)slots * 2 * 12 <= 200 * 1024) slots <<= 1;   // 2P values at load <= 0.5 ...
  const size_t smem = (size_t)slots * 12;
  CP2_REQUIRE(2.0 * hs * ws <= 0.85 * slots && 2 * hs * ws < 65536,                  // ... <= 0.85 at the largest maps
              "cp2_id_iou: %d x %d sub-sampled ids do not fit the shared-memory hash set", hs, ws);
  if (smem > 48 * 1024)
    CP2_CUDA(cudaFuncSetAttribute(id_iou_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  id_iou_kernel<<<B, 512, smem, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const long long *>(ids_a), reinterpret_cast<const long long *>(ids_b), mask_a_ds,
      mask_b_ds, iou, iou_masked, H, W, os, hs, ws, slots);
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}
