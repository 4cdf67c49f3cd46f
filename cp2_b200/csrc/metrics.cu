// f4 — the dense accuracy meter forward_cp2 logs every step, without materialising the logits:
//   acc_seg (reference builder.py:1442-1448): arg-max of the [P*P] dense logits of each image, hit if
//           that pair is a label pair (mask_a[x] * mask_b[y]).
// An arg-max decision, taken in fp32 on CUDA cores straight from the fp32 inputs (bf16 operands could flip
// near-ties).  Positive scalings (1/tau, all-ones weights) do not move an arg-max and are skipped.
// (acc_ins, builder.py:1441, comes out of cp2_global_nce's single queue pass.)
#include "common.cuh"

namespace {

constexpr int MT = 256;                 // threads per CTA
constexpr int YT = 64;                  // key pixels per CTA (lane = keys y0+lane and y0+32+lane)
constexpr int XT = 64;                  // query pixels per chunk (warp w owns rows 8w..8w+7)
constexpr float kEps = 1e-12f;          // F.normalize eps

struct Best { float v; int i; };
__device__ __forceinline__ void take(Best &a, float v, int i) {
  if (v > a.v || (v == a.v && i < a.i)) { a.v = v; a.i = i; }     // first occurrence on ties
}

// rows [c][W] <- feat[b, c, p0 .. p0+W) (zero past P); feat is [B,C,P] f32.  128-bit loads when the rows are
// 16-byte aligned (P % 4 == 0), all of a thread's loads issued before the first store.
template <int W>
__device__ __forceinline__ void load_cols(const float *feat_b, int C, int P, int p0, float *dst) {
  if ((P & 3) == 0) {
    constexpr int W4 = W / 4;
    const int n4 = C * W4;
    for (int i0 = threadIdx.x; i0 < n4; i0 += 8 * MT) {
      float4 v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int i = i0 + u * MT, c = i / W4, j = (i - c * W4) * 4;
        v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (i < n4 && p0 + j < P) v[u] = *reinterpret_cast<const float4 *>(feat_b + (long long)c * P + p0 + j);
      }
#pragma unroll
      for (int u = 0; u < 8; ++u)
        if (i0 + u * MT < n4) *reinterpret_cast<float4 *>(dst + 4 * (i0 + u * MT)) = v[u];
    }
    return;
  }
  for (int i = threadIdx.x; i < C * W; i += MT) {
    const int c = i / W, j = i - c * W;
    dst[i] = p0 + j < P ? feat_b[(long long)c * P + p0 + j] : 0.f;
  }
}

// One CTA = 64 key pixels x 64 query pixels of one image (grid: key tiles, images, query chunks):
// best (value, flat index x*P+y) of the tile.
__global__ void __launch_bounds__(MT)
dense_top1_kernel(const float *__restrict__ q, const float *__restrict__ k, float *__restrict__ part_v,
                  int *__restrict__ part_i, int C, int P) {
  extern __shared__ __align__(16) float sm[];
  float *ks = sm;                        // [C][32]
  float *qs = sm + C * YT;               // [C][64]
  __shared__ float invk_s[YT], invq_s[XT];
  __shared__ float red_v[MT / 32];
  __shared__ int red_i[MT / 32];
  const int b = blockIdx.y, y0 = blockIdx.x * YT, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const float *qb = q + (long long)b * C * P, *kb = k + (long long)b * C * P;
  load_cols<YT>(kb, C, P, y0, ks);
  __syncthreads();
  if (threadIdx.x < YT) {
    float s = 0.f;
    for (int c = 0; c < C; ++c) s = fmaf(ks[c * YT + threadIdx.x], ks[c * YT + threadIdx.x], s);
    invk_s[threadIdx.x] = 1.f / fmaxf(sqrtf(s), kEps);
  }
  const int y = y0 + lane;
  Best best{-INFINITY, 0x7fffffff};
  {
    const int x0 = blockIdx.z * XT;
    load_cols<XT>(qb, C, P, x0, qs);
    __syncthreads();
    if (threadIdx.x < XT) {
      float s = 0.f;
      for (int c = 0; c < C; ++c) s = fmaf(qs[c * XT + threadIdx.x], qs[c * XT + threadIdx.x], s);
      invq_s[threadIdx.x] = 1.f / fmaxf(sqrtf(s), kEps);
    }
    // two keys per thread: every 128-bit broadcast read of the query tile feeds 8 FMAs instead of 4
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, bcc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll 4
    for (int c = 0; c < C; ++c) {
      const float kv = ks[c * YT + lane], kw = ks[c * YT + 32 + lane];
      const float4 a = *reinterpret_cast<const float4 *>(qs + c * XT + warp * 8);
      const float4 d = *reinterpret_cast<const float4 *>(qs + c * XT + warp * 8 + 4);
      acc[0] = fmaf(a.x, kv, acc[0]); acc[1] = fmaf(a.y, kv, acc[1]);
      acc[2] = fmaf(a.z, kv, acc[2]); acc[3] = fmaf(a.w, kv, acc[3]);
      acc[4] = fmaf(d.x, kv, acc[4]); acc[5] = fmaf(d.y, kv, acc[5]);
      acc[6] = fmaf(d.z, kv, acc[6]); acc[7] = fmaf(d.w, kv, acc[7]);
      bcc[0] = fmaf(a.x, kw, bcc[0]); bcc[1] = fmaf(a.y, kw, bcc[1]);
      bcc[2] = fmaf(a.z, kw, bcc[2]); bcc[3] = fmaf(a.w, kw, bcc[3]);
      bcc[4] = fmaf(d.x, kw, bcc[4]); bcc[5] = fmaf(d.y, kw, bcc[5]);
      bcc[6] = fmaf(d.z, kw, bcc[6]); bcc[7] = fmaf(d.w, kw, bcc[7]);
    }
    __syncthreads();                     // invq_s written
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int x = x0 + warp * 8 + i;
      if (x < P) {
        const float iq = invq_s[warp * 8 + i];
        if (y < P) take(best, acc[i] * iq * invk_s[lane], x * P + y);
        if (y + 32 < P) take(best, bcc[i] * iq * invk_s[32 + lane], x * P + y + 32);
      }
    }
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    const float v = __shfl_down_sync(0xffffffffu, best.v, off);
    const int i = __shfl_down_sync(0xffffffffu, best.i, off);
    take(best, v, i);
  }
  if (lane == 0) { red_v[warp] = best.v; red_i[warp] = best.i; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < MT / 32; ++w) take(best, red_v[w], red_i[w]);
    const long long slot = ((long long)b * gridDim.x + blockIdx.x) * gridDim.z + blockIdx.z;
    part_v[slot] = best.v;
    part_i[slot] = best.i;
  }
}

// hit[b] = mask_a[b, x*] * mask_b[b, y*] at the image's arg-max pair; acc = 100 * mean(hit).  One CTA.
__global__ void __launch_bounds__(MT)
dense_top1_finalize_kernel(const float *__restrict__ part_v, const int *__restrict__ part_i,
                           const float *__restrict__ mask_a, const float *__restrict__ mask_b,
                           float *__restrict__ hit, float *__restrict__ acc, int B, int P, int tiles) {
  __shared__ float red[MT / 32];
  float s = 0.f;
  for (int b = threadIdx.x; b < B; b += MT) {
    Best best{-INFINITY, 0x7fffffff};
    for (int t = 0; t < tiles; ++t) take(best, part_v[(long long)b * tiles + t], part_i[(long long)b * tiles + t]);
    float h = 0.f;
    if (best.i != 0x7fffffff) {
      const int x = best.i / P, y = best.i - x * P;
      h = mask_a[(long long)b * P + x] * mask_b[(long long)b * P + y];
    }
    hit[b] = h;
    s += h;
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) s += __shfl_down_sync(0xffffffffu, s, off);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < MT / 32; ++w) t += red[w];
    acc[0] = t * (100.f / (float)B);
  }
}

}  // namespace

extern "C" int64_t cp2_dense_top1_scratch_floats(int B, int P) {
  return 2LL * B * cp2_cdiv(P, YT) * cp2_cdiv(P, XT);
}

extern "C" int cp2_dense_top1(const float *q, const float *k, const float *mask_a, const float *mask_b, float *hit,
                              float *acc, float *scratch, int B, int C, int P, void *stream) {
  CP2_REQUIRE(B > 0 && P > 0 && C > 0 && C <= 384 && (long long)P * P < 0x7fffffffLL,
              "cp2_dense_top1: bad shape B=%d C=%d P=%d", B, C, P);
  CP2_REQUIRE(q && k && mask_a && mask_b && hit && acc && scratch, "cp2_dense_top1: null pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int ytiles = cp2_cdiv(P, YT), xchunks = cp2_cdiv(P, XT), tiles = ytiles * xchunks;
  CP2_REQUIRE(xchunks <= 65535 && B <= 65535, "cp2_dense_top1: B=%d or P=%d too large for the grid", B, P);
  float *part_v = scratch;
  int *part_i = reinterpret_cast<int *>(scratch + (size_t)B * tiles);
  const size_t smem = (size_t)C * (YT + XT) * sizeof(float);
  CP2_CUDA(cudaFuncSetAttribute(dense_top1_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dense_top1_kernel<<<dim3(ytiles, B, xchunks), MT, smem, st>>>(q, k, part_v, part_i, C, P);
  CP2_LAUNCH_CHECK();
  dense_top1_finalize_kernel<<<1, MT, 0, st>>>(part_v, part_i, mask_a, mask_b, hit, acc, B, P, tiles);
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}
