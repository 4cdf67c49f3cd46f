// f2 — the backbone match of forward_densecl (reference builder.py:817-821):
//     idx[n,x] = argmax_y sum_c embd_q[n,c,x] * embd_k[n,c,y]        (c over the 2048 backbone channels)
// as ONE kernel on the tensor cores; the [B,S^2,S^2] similarity matrix lives only in TMEM.
//
// The operands are the fp32 backbone feature maps as cuDNN leaves them, [B,Cb,P] with the pixel axis contiguous.
// CTA = (image, 128 query pixels).  Per 16-channel slab, 8 converter warps bring the fp32 values into a shared-
// memory ring with cp.async (4 slabs in flight per SM: the kernel is a stream over 2 x 1.6 MB per image), split
// each into two bf16 terms (hi = RN(v), lo = RN(v - hi)) and write them as MN-major, 128-byte-swizzled tcgen05
// operand tiles (3 stages); the MMA warp accumulates  S += q_hi k_hi + q_hi k_lo + q_lo k_hi  in
// TMEM (fp32), i.e. the similarity to ~2^-16 relative — ties closer than that are below the noise of the
// reference's own fp32 summation order on 2048 channels.  After the last slab four warps scan their TMEM lane
// (one query pixel each) for the first largest column.  P <= 256.
#include "common.cuh"
#include "umma.cuh"

using namespace umma;

namespace {

constexpr int MT_THREADS = 288;          // warp 0: MMA issuer + TMEM owner; warps 1-8: converters (1-4 also arg-max)
constexpr int KS = 16;                   // channels per slab (one K = 16 MMA slice)
constexpr int NS = 3;                    // bf16 operand stages
constexpr int ND = 4;                    // fp32 slabs in flight (cp.async ring)
constexpr int A_BYTES = 2 * KS * 128;    // one bf16 term of the query slab: 2 chunks of 64 pixels x 16 channel rows
constexpr int B_BYTES = 4 * KS * 128;    // one bf16 term of the key slab: 4 chunks of 64 pixels
constexpr int STAGE = 2 * A_BYTES + 2 * B_BYTES;
constexpr int RING = 6 * 256 * 16;       // one fp32 slab: 6 16-byte pieces per converter thread, piece-major
constexpr size_t MT_SMEM = 1024 + (size_t)NS * STAGE + (size_t)ND * RING + 2 * NS * 8 + 8 + 16;

__device__ __forceinline__ uint8_t *align1024m(uint8_t *p) {
  return reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(p) + 1023) & ~uintptr_t(1023));
}

// 8 consecutive pixels of channel row `c` (0..15) starting at pixel 8*unit: byte offset inside an MN-major
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

// One unit (8 consecutive pixels p0..p0+7 of a channel row) into two 16-byte ring pieces, asynchronously in every
// case (a synchronous load on the map edge would put one memory round trip per slab on the critical path):
// 16-byte cp.async when the rows are 16-byte aligned (P % 4 == 0: a piece is then entirely inside or outside the
// map), 4-byte cp.async per pixel otherwise; pieces / pixels past P are zero-filled (src-size 0 reads nothing).
__device__ __forceinline__ void stage8(const float *__restrict__ row, int p0, int P, bool vec, uint8_t *d0, uint8_t *d1) {
  if (vec) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int p = p0 + 4 * h;
      const uint32_t n = p < P ? 16u : 0u;
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(h ? d1 : d0)),
                   "l"(row + (p < P ? p : 0)), "r"(n) : "memory");
    }
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int p = p0 + i;
      const uint32_t n = p < P ? 4u : 0u;
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(smem_u32((i < 4 ? d0 : d1) + 4 * (i & 3))),
                   "l"(row + (p < P ? p : 0)), "r"(n) : "memory");
    }
  }
}

__global__ void __launch_bounds__(MT_THREADS, 1)
densecl_match_kernel(const float *__restrict__ eq, const float *__restrict__ ek, long long *__restrict__ idx,
                     int Cb, int P, int ncols) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = align1024m(smem_raw);
  uint8_t *ring = smem + NS * STAGE;
  uint64_t *bars = reinterpret_cast<uint64_t *>(ring + ND * RING);
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
        const uint64_t dAh = make_sdesc(aH, KS * 128, 1024), dAl = make_sdesc(aL, KS * 128, 1024);
        const uint64_t dBh = make_sdesc(bH, KS * 128, 1024), dBl = make_sdesc(bL, KS * 128, 1024);
        mma_f16_ss(tmem, dAh, dBh, idesc, i > 0 ? 1u : 0u);
        mma_f16_ss(tmem, dAh, dBl, idesc, 1u);
        mma_f16_ss(tmem, dAl, dBh, idesc, 1u);
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
    // thread -> channel row c, one 8-pixel unit of the query slab (16 x 16 units) and two of the key slab (16 x 32);
    // a thread converts exactly what it copied, so the fp32 ring needs no barrier, only cp.async.wait_group
    const int c = t >> 4;                                // 0..15
    const int ua = t & 15, ub0 = t & 15, ub1 = (t & 15) + 16;
    auto piece = [&](int slot, int pc) { return ring + slot * RING + (pc * 256 + t) * 16; };
    auto fetch = [&](int i) {
      const int slot = i % ND;
      const float *qrow = qb + (long long)(i * KS + c) * P, *krow = kb + (long long)(i * KS + c) * P;
      stage8(qrow, x0 + ua * 8, P, vec, piece(slot, 0), piece(slot, 1));
      stage8(krow, ub0 * 8, P, vec, piece(slot, 2), piece(slot, 3));
      stage8(krow, ub1 * 8, P, vec, piece(slot, 4), piece(slot, 5));
    };
    auto take8 = [&](int slot, int pc, float (&v)[8]) {
      const float4 a = *reinterpret_cast<const float4 *>(piece(slot, pc));
      const float4 d = *reinterpret_cast<const float4 *>(piece(slot, pc + 1));
      v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = d.x; v[5] = d.y; v[6] = d.z; v[7] = d.w;
    };
    for (int i = 0; i < ND - 1; ++i) {                   // prologue: ND-1 slabs in flight
      if (i < nslab) fetch(i);
      asm volatile("cp.async.commit_group;" ::: "memory");
    }
    for (int i = 0; i < nslab; ++i) {
      if (i + ND - 1 < nslab) fetch(i + ND - 1);         // its ring slot was converted in iteration i-1
      asm volatile("cp.async.commit_group;" ::: "memory");
      asm volatile("cp.async.wait_group %0;" ::"n"(ND - 1) : "memory");   // slab i has landed (this thread's pieces)
      const int slot = i % ND, s = i % NS, ph = (i / NS) & 1;
      float v[8];
      uint4 hq, lq, hk0, lk0, hk1, lk1;
      take8(slot, 0, v); split8(v, hq, lq);
      take8(slot, 2, v); split8(v, hk0, lk0);
      take8(slot, 4, v); split8(v, hk1, lk1);
      mbar_wait(&empty[s], ph ^ 1);                      // the MMAs that read this stage have completed
      uint8_t *st = smem + s * STAGE;
      const uint32_t oa = mn_off(c, ua), o0 = mn_off(c, ub0), o1 = mn_off(c, ub1);
      *reinterpret_cast<uint4 *>(st + oa) = hq;
      *reinterpret_cast<uint4 *>(st + A_BYTES + oa) = lq;
      *reinterpret_cast<uint4 *>(st + 2 * A_BYTES + o0) = hk0;
      *reinterpret_cast<uint4 *>(st + 2 * A_BYTES + B_BYTES + o0) = lk0;
      *reinterpret_cast<uint4 *>(st + 2 * A_BYTES + o1) = hk1;
      *reinterpret_cast<uint4 *>(st + 2 * A_BYTES + B_BYTES + o1) = lk1;
      fence_proxy_async_smem();                          // generic-proxy writes -> visible to the tensor core
      __syncwarp();
      if (lane == 0) mbar_arrive(&full[s]);
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
