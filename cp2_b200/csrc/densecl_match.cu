// f2 — the backbone match of forward_densecl (reference builder.py:817-821):
//     idx[n,x] = argmax_y sum_c embd_q[n,c,x] * embd_k[n,c,y]        (c over the 2048 backbone channels)
// as ONE kernel on the tensor cores; the [B,S^2,S^2] similarity matrix lives only in TMEM.
//
// The operands are the fp32 backbone feature maps as cuDNN leaves them, [B,Cb,P] with the pixel axis contiguous.
// CTA = (image, 128 query pixels).  Per 32-channel slab, 8 converter warps read the fp32 values (coalesced along
// P), split each into two bf16 terms (hi = RN(v), lo = RN(v - hi)) and write them as MN-major, 128-byte-swizzled
// tcgen05 operand tiles (3 slabs in flight); the MMA warp accumulates  S += q_hi k_hi + q_hi k_lo + q_lo k_hi  in
// TMEM (fp32), i.e. the similarity to ~2^-16 relative — ties closer than that are below the noise of the
// reference's own fp32 summation order on 2048 channels.  After the last slab four warps scan their TMEM lane
// (one query pixel each) for the first largest column.  P <= 256.
#include "common.cuh"
#include "umma.cuh"

using namespace umma;

namespace {

constexpr int MT_THREADS = 288;          // warp 0: MMA issuer + TMEM owner; warps 1-8: converters (1-4 also arg-max)
constexpr int KS = 32;                   // channels per slab (two K = 16 MMA slices)
constexpr int NS = 3;                    // slabs in flight
constexpr int A_BYTES = 2 * KS * 128;    // one bf16 term of the query slab: 2 chunks of 64 pixels x 32 channel rows
constexpr int B_BYTES = 4 * KS * 128;    // one bf16 term of the key slab: 4 chunks of 64 pixels
constexpr int STAGE = 2 * A_BYTES + 2 * B_BYTES;
constexpr size_t MT_SMEM = 1024 + (size_t)NS * STAGE + 2 * NS * 8 + 8 + 16;

__device__ __forceinline__ uint8_t *align1024m(uint8_t *p) {
  return reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(p) + 1023) & ~uintptr_t(1023));
}

// 8 consecutive pixels of channel row `c` (0..31) starting at pixel 8*unit: byte offset inside an MN-major
// operand term whose 64-pixel chunks are KS*128 bytes apart (LBO) and whose 8-row groups are 1024 bytes apart (SBO)
__device__ __forceinline__ uint32_t mn_off(int c, int unit) {
  return (uint32_t)((unit >> 3) * (KS * 128) + (c >> 3) * 1024 + (c & 7) * 128 + (((unit & 7) ^ (c & 7)) << 4));
}

// 8 fp32 -> hi / lo bf16 x 8 (two 16-byte vectors)
__device__ __forceinline__ void split8(const float (&v)[8], uint4 &hi, uint4 &lo) {
  uint32_t h[4], l[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const __nv_bfloat16 h0 = __float2bfloat16_rn(v[2 * i]), h1 = __float2bfloat16_rn(v[2 * i + 1]);
    h[i] = pack_bf16x2(__bfloat162float(h0), __bfloat162float(h1));
    l[i] = pack_bf16x2(v[2 * i] - __bfloat162float(h0), v[2 * i + 1] - __bfloat162float(h1));
  }
  hi = make_uint4(h[0], h[1], h[2], h[3]);
  lo = make_uint4(l[0], l[1], l[2], l[3]);
}

// 8 consecutive pixels p0..p0+7 of one channel row (zeros past P); vec: rows are 16-byte aligned (P % 4 == 0)
__device__ __forceinline__ void load8(const float *__restrict__ row, int p0, int P, bool vec, float (&v)[8]) {
  if (vec && p0 + 8 <= P) {
    const float4 a = __ldg(reinterpret_cast<const float4 *>(row + p0));
    const float4 b = __ldg(reinterpret_cast<const float4 *>(row + p0 + 4));
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = (p0 + i < P) ? __ldg(row + p0 + i) : 0.f;
  }
}

__global__ void __launch_bounds__(MT_THREADS, 1)
densecl_match_kernel(const float *__restrict__ eq, const float *__restrict__ ek, long long *__restrict__ idx,
                     int Cb, int P, int ncols) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = align1024m(smem_raw);
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + NS * STAGE);
  uint64_t *full = bars, *empty = bars + NS, *done = bars + 2 * NS;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(done + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.y, x0 = blockIdx.x * 128;
  const int nslab = Cb / KS;
  if (threadIdx.x == 0) {
    for (int s = 0; s < NS; ++s) { mbar_init(&full[s], 8); mbar_init(&empty[s], 1); }
    mbar_init(done, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc(tmem_slot, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------------------------ MMA issuer
    const uint32_t idesc = make_idesc(128, ncols, FMT_BF16, FMT_BF16, MAJOR_MN, MAJOR_MN);
    for (int i = 0; i < nslab; ++i) {
      const int s = i % NS, ph = (i / NS) & 1;
      mbar_wait(&full[s], ph);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t base = smem_u32(smem + s * STAGE);
        const uint32_t aH = base, aL = base + A_BYTES, bH = base + 2 * A_BYTES, bL = bH + B_BYTES;
#pragma unroll
        for (int kk = 0; kk < KS / 16; ++kk) {            // 16 channel rows (2 KB) per slice
          const uint64_t dAh = make_sdesc(aH + kk * 2048, KS * 128, 1024), dAl = make_sdesc(aL + kk * 2048, KS * 128, 1024);
          const uint64_t dBh = make_sdesc(bH + kk * 2048, KS * 128, 1024), dBl = make_sdesc(bL + kk * 2048, KS * 128, 1024);
          mma_f16_ss(tmem, dAh, dBh, idesc, (i > 0 || kk > 0) ? 1u : 0u);
          mma_f16_ss(tmem, dAh, dBl, idesc, 1u);
          mma_f16_ss(tmem, dAl, dBh, idesc, 1u);
        }
        mma_commit(&empty[s]);
        if (i == nslab - 1) mma_commit(done);
      }
      __syncwarp();
    }
  } else {
    // ------------------------------------------------------------------ converters
    const int t = threadIdx.x - 32;                      // 0..255
    const bool vec = (P & 3) == 0;
    const float *qb = eq + (long long)b * Cb * P, *kb = ek + (long long)b * Cb * P;
    // thread -> (channel row c, 8-pixel unit): query slab 32 x 16 units (2 per thread), key slab 32 x 32 (4)
    const int c = t >> 3;                                // 0..31
    const int u0 = t & 7;                                // units u0 + 8*j
    // register double buffer: the loads of slab i+1 are in flight while slab i is split and stored
    float vqa[2][8], vka[4][8], vqb[2][8], vkb[4][8];
    auto fetch = [&](int i, float (&vq)[2][8], float (&vk)[4][8]) {
      const float *qrow = qb + (long long)(i * KS + c) * P, *krow = kb + (long long)(i * KS + c) * P;
#pragma unroll
      for (int j = 0; j < 2; ++j) load8(qrow, x0 + (u0 + 8 * j) * 8, P, vec, vq[j]);
#pragma unroll
      for (int j = 0; j < 4; ++j) load8(krow, (u0 + 8 * j) * 8, P, vec, vk[j]);
    };
    auto emit = [&](int i, const float (&vq)[2][8], const float (&vk)[4][8]) {
      const int s = i % NS, ph = (i / NS) & 1;
      mbar_wait(&empty[s], ph ^ 1);                      // the MMAs that read this stage have completed
      uint8_t *st = smem + s * STAGE;
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        uint4 hi, lo;
        split8(vq[j], hi, lo);
        const uint32_t off = mn_off(c, u0 + 8 * j);
        *reinterpret_cast<uint4 *>(st + off) = hi;
        *reinterpret_cast<uint4 *>(st + A_BYTES + off) = lo;
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        uint4 hi, lo;
        split8(vk[j], hi, lo);
        const uint32_t off = mn_off(c, u0 + 8 * j);
        *reinterpret_cast<uint4 *>(st + 2 * A_BYTES + off) = hi;
        *reinterpret_cast<uint4 *>(st + 2 * A_BYTES + B_BYTES + off) = lo;
      }
      fence_proxy_async_smem();                          // generic-proxy writes -> visible to the tensor core
      __syncwarp();
      if (lane == 0) mbar_arrive(&full[s]);
    };
    fetch(0, vqa, vka);
    for (int i = 0; i < nslab; i += 2) {
      if (i + 1 < nslab) fetch(i + 1, vqb, vkb);
      emit(i, vqa, vka);
      if (i + 1 < nslab) {
        if (i + 2 < nslab) fetch(i + 2, vqa, vka);
        emit(i + 1, vqb, vkb);
      }
    }
    // ------------------------------------------------------------------ arg-max (warps 1-4, one TMEM lane each)
    if (warp <= 4) {
      mbar_wait(done, 0);
      tc_fence_after();
      const int quad = warp & 3, x = x0 + quad * 32 + lane;
      const uint32_t lane_base = tmem + (uint32_t(quad * 32) << 16);
      float best = -INFINITY;
      int bi = 0;
      for (int ch = 0; ch * 32 < ncols; ++ch) {
        uint32_t r[32];
        tmem_ld32(lane_base + ch * 32, r);
        tmem_wait_ld();
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const int y = ch * 32 + j;
          const float v = __uint_as_float(r[j]);
          if (y < P && v > best) { best = v; bi = y; }   // strict: the first largest column wins
        }
      }
      if (x < P) idx[(long long)b * P + x] = bi;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}

}  // namespace

extern "C" int cp2_densecl_match(const float *embd_q, const float *embd_k, int64_t *idx, int B, int Cb, int P,
                                 void *stream) {
  CP2_REQUIRE(B > 0 && B <= 65535 && P > 0 && P <= 256, "cp2_densecl_match: bad shape B=%d P=%d (P <= 256)", B, P);
  CP2_REQUIRE(Cb > 0 && Cb % KS == 0, "cp2_densecl_match: Cb=%d must be a positive multiple of %d", Cb, KS);
  CP2_REQUIRE(embd_q && embd_k && idx, "cp2_densecl_match: null pointer");
  CP2_REQUIRE((reinterpret_cast<uintptr_t>(embd_q) & 15u) == 0 && (reinterpret_cast<uintptr_t>(embd_k) & 15u) == 0,
              "cp2_densecl_match: embd_q / embd_k must be 16-byte aligned");
  const int ncols = ((P + 31) / 32) * 32;              // MMA N: every 32-column chunk the arg-max reads is written
  CP2_SMEM_ONCE(densecl_match_kernel, MT_SMEM);
  densecl_match_kernel<<<dim3(cp2_cdiv(P, 128), B), MT_THREADS, MT_SMEM, static_cast<cudaStream_t>(stream)>>>(
      embd_q, embd_k, reinterpret_cast<long long *>(idx), Cb, P, ncols);
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}
