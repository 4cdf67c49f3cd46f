// a2 + a4 — stride subsample of the id maps and the (masked) correlation IoU of two id maps
// (reference builder.py:1165-1186 `x[:, OS//2::OS, OS//2::OS]`, tools/correlation_mapping.py:103-138
// get_masked_iou, called through get_masked_correlation_map at builder.py:1203-1217, 1252-1257).
//
// Per image the reference concatenates both (sub-sampled) id maps, adds 1, zeroes the masked entries and
// runs torch.unique(return_counts=True) on the host side of a python loop:
//   union        = number of distinct non-zero values
//   intersection = number of distinct non-zero values that occur more than once
// Integer work, bit-exact.  One CTA per image; the 2P values sit in shared memory and every thread
// classifies its elements ("first occurrence?", "occurs again?") by a linear scan — P = 196 at the
// reference's 224x224 / OS16, where the whole problem is 392^2 compares per image.
#include "common.cuh"

namespace {

__global__ void __launch_bounds__(512)
id_iou_kernel(const long long *__restrict__ ids_a, const long long *__restrict__ ids_b,
              const float *__restrict__ mask_a, const float *__restrict__ mask_b, float *__restrict__ iou,
              float *__restrict__ iou_masked, int H, int W, int os, int hs, int ws) {
  extern __shared__ long long sv[];                 // [2P] id + 1
  const int P = hs * ws, n = 2 * P, b = blockIdx.x, o = os / 2;
  unsigned char *sm = reinterpret_cast<unsigned char *>(sv + n);     // [2P] mask != 0
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int p = i < P ? i : i - P;
    const int y = o + (p / ws) * os, x = o + (p % ws) * os;
    const long long *src = i < P ? ids_a : ids_b;
    const float *msk = i < P ? mask_a : mask_b;
    sv[i] = src[((long long)b * H + y) * W + x] + 1;
    sm[i] = msk[(long long)b * P + p] != 0.f;
  }
  __syncthreads();
  int uni = 0, inter = 0, uni_m = 0, inter_m = 0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const long long v = sv[i];
    const bool mi = sm[i];
    // every thread walks the same j: broadcast shared-memory reads, unrolled so 8 are in flight
    int before = 0, after = 0, before_m = 0, after_m = 0;       // equal values at j < i / j > i
#pragma unroll 8
    for (int j = 0; j < n; ++j) {
      const int eq = sv[j] == v, m = sm[j];
      const int lo = j < i, hi = j > i;
      before += eq & lo;
      after += eq & hi;
      before_m += eq & lo & m;
      after_m += eq & hi & m;
    }
    const bool first = before == 0, dup = (before + after) > 0;
    const bool first_m = before_m == 0, dup_m = (before_m + after_m) > 0;
    if (v != 0) {                                   // the reference drops the value 0 (id == -1)
      uni += first;
      inter += first && dup;
      if (mi) { uni_m += first_m; inter_m += first_m && dup_m; }
    }
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
  const size_t smem = (size_t)2 * hs * ws * 9 + 16;
  CP2_REQUIRE(smem <= 200 * 1024, "cp2_id_iou: %d x %d sub-sampled ids do not fit shared memory", hs, ws);
  if (smem > 48 * 1024)
    CP2_CUDA(cudaFuncSetAttribute(id_iou_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  id_iou_kernel<<<B, 512, smem, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const long long *>(ids_a), reinterpret_cast<const long long *>(ids_b), mask_a_ds,
      mask_b_ds, iou, iou_masked, H, W, os, hs, ws);
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}
