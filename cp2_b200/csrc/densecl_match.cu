// f2 — the backbone match of forward_densecl (reference builder.py:817-821):
//     idx[n,x] = argmax_y sum_c embd_q[n,c,x] * embd_k[n,c,y]        (c over the 2048 backbone channels)
// as ONE kernel on the tensor cores; the [B,S^2,S^2] similarity matrix lives only in TMEM.
//
// The operands are the fp32 backbone feature maps as cuDNN leaves them, [B,Cb,P] with the pixel axis contiguous.
// The kernel is a stream over them (2 x 1.6 MB per image at configs[1]), so it is laid out for bytes in flight:
// CTA = (image, 128 query pixels, 128 key pixels) — 4 CTAs per 14x14 image, 128 at B = 32 — and per 16-channel
// slab a group of 2 converter warps (4 groups, each owning every 4th slab so that their per-slab fixed costs
// overlap) brings the fp32 values into a shared-memory ring with cp.async (warp-wide, coalesced 512-byte row
// copies), splits each into two bf16 terms (hi = RN(v), lo = RN(v - hi)) and writes them as MN-major,
// 128-byte-swizzled tcgen05 operand tiles (6 stages);
// the MMA warp accumulates
// S += q_hi k_hi + q_hi k_lo + q_lo k_hi in TMEM (fp32), i.e. the similarity to ~2^-16 relative — ties closer than
// that are below the noise of the reference's own fp32 summation order on 2048 channels.  After the last slab four
// warps scan their TMEM lane (one query pixel each) for the first largest column of the tile and publish
// (value, column) with one 64-bit atomicMax per query pixel (order-preserving key, smaller column wins ties:
// deterministic); the last CTA of an image decodes the winners into idx and re-zeroes the scratch.  P <= 256.
#include "common.cuh"
#include "umma.cuh"

using namespace umma;

namespace {

constexpr int MT_THREADS = 288;          // warp 0: MMA issuer + TMEM owner; warps 1-8: converters (1-4 also arg-max)
// Tunables (template parameters of the kernel): KS channels per slab (K = 16 MMA slices: KS / 16), NG converter
// groups (256 / NG threads each; group g owns the slabs i = g (mod NG), so the per-slab fixed costs of the groups
// overlap), ND fp32 slabs in the cp.async ring (ND / NG per group, all but one in flight), NS bf16 operand stages.
template <int KS, int NG, int ND, int NS>
struct MatchCfg {
  static constexpr int TG = 256 / NG;                   // threads per converter group
  static constexpr int RPT = KS / (TG / 32);            // channel rows per warp and operand: a warp copies whole 512-byte rows
  static constexpr int T_BYTES = 2 * KS * 128;          // one bf16 term of a 128-pixel slab: 2 chunks of 64 pixels x KS channel rows
  static constexpr int STAGE = 4 * T_BYTES;             // q_hi, q_lo, k_hi, k_lo
  static constexpr int RING = 2 * KS * 512;             // one fp32 slab (query + key rows), piece-major per owning group
  static constexpr size_t SMEM = 1024 + (size_t)NS * STAGE + (size_t)ND * RING + 2 * NS * 8 + 8 + 32;
  static_assert(ND % NG == 0 && KS % 16 == 0 && KS % (TG / 32) == 0, "bad match configuration");
};

__device__ __forceinline__ uint8_t *align1024m(uint8_t *p) {
  return reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(p) + 1023) & ~uintptr_t(1023));
}

// 8 consecutive pixels of channel row `c` (0..KS-1) starting at pixel 8*unit: byte offset inside an MN-major
// operand term whose 64-pixel chunks are KS*128 bytes apart (LBO) and whose 8-row groups are 1024 bytes apart (SBO)
template <int KS>
__device__ __forceinline__ uint32_t mn_off(int c, int unit) {
  return (uint32_t)((unit >> 3) * (KS * 128) + (c >> 3) * 1024 + (c & 7) * 128 + (((unit & 7) ^ (c & 7)) << 4));
}

// 4 fp32 -> hi / lo bf16 x 4 (two 8-byte vectors).  hi pairs come from one packed conversion (cvt.rn.bf16x2.f32),
// their fp32 values back by shift / mask: no quarter-rate scalar F2F on the path.
__device__ __forceinline__ void split4(const float4 v, uint2 &hi, uint2 &lo) {
  hi.x = pack_bf16x2(v.x, v.y);
  hi.y = pack_bf16x2(v.z, v.w);
  lo.x = pack_bf16x2(v.x - __uint_as_float(hi.x << 16), v.y - __uint_as_float(hi.x & 0xffff0000u));
  lo.y = pack_bf16x2(v.z - __uint_as_float(hi.y << 16), v.w - __uint_as_float(hi.y & 0xffff0000u));
}

// Four consecutive pixels p0..p0+3 of a channel row into one 16-byte ring piece, asynchronously in every case (a
// synchronous load on the map edge would put one memory round trip per slab on the critical path): one 16-byte
// cp.async when the rows are 16-byte aligned (P % 4 == 0: a piece is then entirely inside or outside the map),
// 4-byte cp.async per pixel otherwise; pixels past P are zero-filled (src-size 0 reads nothing).  The 32 lanes of a
// warp take consecutive pieces: 512 contiguous bytes of one row per instruction.
__device__ __forceinline__ void stage4(const float *__restrict__ row, int p0, int P, bool vec, uint8_t *d) {
  if (vec) {
    const uint32_t n = p0 < P ? 16u : 0u;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(d)), "l"(row + (p0 < P ? p0 : 0)), "r"(n)
                 : "memory");
  } else {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int p = p0 + i;
      const uint32_t n = p < P ? 4u : 0u;
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(smem_u32(d + 4 * i)), "l"(row + (p < P ? p : 0)),
                   "r"(n) : "memory");
    }
  }
}

// order-preserving 32-bit key of a float (larger float <-> larger unsigned)
__device__ __forceinline__ uint32_t orderable(float v) {
  const uint32_t u = __float_as_uint(v);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

template <int KS, int NG, int ND, int NS>
__global__ void __launch_bounds__(MT_THREADS, 1)
densecl_match_kernel(const float *__restrict__ eq, const float *__restrict__ ek, long long *__restrict__ idx,
                     unsigned long long *__restrict__ best, unsigned int *__restrict__ counters, int Cb, int P) {
  using Cfg = MatchCfg<KS, NG, ND, NS>;
  constexpr int TG = Cfg::TG, RPT = Cfg::RPT, T_BYTES = Cfg::T_BYTES, STAGE = Cfg::STAGE, RING = Cfg::RING;
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = align1024m(smem_raw);
  uint8_t *ring = smem + NS * STAGE;
  uint64_t *bars = reinterpret_cast<uint64_t *>(ring + ND * RING);
  uint64_t *full = bars, *empty = bars + NS, *done = bars + 2 * NS;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(done + 1);
  int *is_last = reinterpret_cast<int *>(tmem_slot + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.z, x0 = blockIdx.x * 128, y0 = blockIdx.y * 128;
  const int nslab = Cb / KS;
  const int ncols = min(128, ((P - y0 + 31) / 32) * 32);   // MMA N: every 32-column chunk the arg-max reads is written
  if (threadIdx.x == 0) {
    for (int s = 0; s < NS; ++s) { mbar_init(&full[s], TG / 32); mbar_init(&empty[s], 1); }
    mbar_init(done, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc(tmem_slot, 128);
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
#pragma unroll
        for (int kk = 0; kk < KS / 16; ++kk) {             // 16 channel rows (2 KB) per slice
          const uint32_t o = base + kk * 2048;
          const uint64_t dAh = make_sdesc(o, KS * 128, 1024), dAl = make_sdesc(o + T_BYTES, KS * 128, 1024);
          const uint64_t dBh = make_sdesc(o + 2 * T_BYTES, KS * 128, 1024), dBl = make_sdesc(o + 3 * T_BYTES, KS * 128, 1024);
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
    // group g = t / TG converts the slabs i = g, g + NG, ...; inside the group warp wg takes the channel rows
    // wg + (TG/32) j (j < RPT) of both 16 x 128 slabs, lane l the pixels 4l..4l+3 of the row (coalesced 512-byte
    // copies).  A thread converts exactly what it copied, so the fp32 ring needs no barrier, only
    // cp.async.wait_group.
    constexpr int DG = ND / NG;                          // ring slots per group
    const int grp = t / TG, tg = t % TG, wg = tg >> 5;
    auto piece = [&](int slot, int pc) { return ring + slot * RING + (pc * TG + tg) * 16; };
    auto fetch = [&](int i) {
      const int slot = i % ND;
#pragma unroll
      for (int j = 0; j < RPT; ++j) {
        const long long ro = (long long)(i * KS + wg + (TG / 32) * j) * P;
        stage4(qb + ro, x0 + 4 * lane, P, vec, piece(slot, j));
        stage4(kb + ro, y0 + 4 * lane, P, vec, piece(slot, RPT + j));
      }
    };
    for (int k = 0; k < DG - 1; ++k) {                   // prologue: DG-1 of the group's slabs in flight
      if (grp + k * NG < nslab) fetch(grp + k * NG);
      asm volatile("cp.async.commit_group;" ::: "memory");
    }
    for (int i = grp; i < nslab; i += NG) {
      if (i + (DG - 1) * NG < nslab) fetch(i + (DG - 1) * NG);   // its ring slot was converted one round ago
      asm volatile("cp.async.commit_group;" ::: "memory");
      asm volatile("cp.async.wait_group %0;" ::"n"(DG - 1) : "memory");   // slab i has landed (this thread's pieces)
      const int slot = i % ND, s = i % NS, ph = (i / NS) & 1;
      uint2 hq[RPT], lq[RPT], hk[RPT], lk[RPT];
#pragma unroll
      for (int j = 0; j < RPT; ++j) {
        split4(*reinterpret_cast<const float4 *>(piece(slot, j)), hq[j], lq[j]);
        split4(*reinterpret_cast<const float4 *>(piece(slot, RPT + j)), hk[j], lk[j]);
      }
      mbar_wait(&empty[s], ph ^ 1);                      // the MMAs that read this stage have completed
#pragma unroll
      for (int j = 0; j < RPT; ++j) {
        // 8-pixel unit lane/2 of channel row c, its lower / upper four pixels
        uint8_t *st = smem + s * STAGE + mn_off<KS>(wg + (TG / 32) * j, lane >> 1) + (lane & 1) * 8;
        *reinterpret_cast<uint2 *>(st) = hq[j];
        *reinterpret_cast<uint2 *>(st + T_BYTES) = lq[j];
        *reinterpret_cast<uint2 *>(st + 2 * T_BYTES) = hk[j];
        *reinterpret_cast<uint2 *>(st + 3 * T_BYTES) = lk[j];
      }
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
      float bv = -INFINITY;
      int bi = 0;
      for (int ch = 0; ch * 32 < ncols; ++ch) {
        uint32_t r[32];
        tmem_ld32(lane_base + ch * 32, r);
        tmem_wait_ld();
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const int y = y0 + ch * 32 + j;
          const float v = __uint_as_float(r[j]);
          if (y < P && v > bv) { bv = v; bi = y; }       // strict: the first largest column of the tile wins
        }
      }
      if (x < P)                                         // across key tiles: larger value, then smaller column
        atomicMax(best + (long long)b * P + x, ((unsigned long long)orderable(bv) << 32) | (0xFFFFFFFFu - (uint32_t)bi));
    }
  }
  tc_fence_before();
  __threadfence();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 128);
  // the last CTA of the image decodes the winners and leaves the scratch zeroed for the next call
  if (threadIdx.x == 0) {
    const unsigned int total = gridDim.x * gridDim.y;
    const int last = atomicAdd(counters + b, 1u) == total - 1u;
    if (last) counters[b] = 0u;
    *is_last = last;
  }
  __syncthreads();
  if (*is_last) {
    __threadfence();
    for (int x = threadIdx.x; x < P; x += MT_THREADS) {
      const unsigned long long key = __ldcg(best + (long long)b * P + x);
      best[(long long)b * P + x] = 0ull;
      idx[(long long)b * P + x] = (long long)(0xFFFFFFFFu - (uint32_t)(key & 0xFFFFFFFFull));
    }
  }
}

}  // namespace

extern "C" int64_t cp2_densecl_match_scratch_bytes(int B, int P) { return (int64_t)B * P * 8 + (int64_t)B * 4 + 16; }

extern "C" int cp2_densecl_match(const float *embd_q, const float *embd_k, int64_t *idx, void *scratch, int B, int Cb,
                                 int P, void *stream) {
  CP2_REQUIRE(B > 0 && B <= 65535 && P > 0 && P <= 256, "cp2_densecl_match: bad shape B=%d P=%d (P <= 256)", B, P);
  CP2_REQUIRE(Cb > 0 && Cb % 16 == 0, "cp2_densecl_match: Cb=%d must be a positive multiple of 16", Cb);
  CP2_REQUIRE(embd_q && embd_k && idx && scratch, "cp2_densecl_match: null pointer");
  CP2_REQUIRE((reinterpret_cast<uintptr_t>(embd_q) & 15u) == 0 && (reinterpret_cast<uintptr_t>(embd_k) & 15u) == 0 &&
                  (reinterpret_cast<uintptr_t>(scratch) & 7u) == 0,
              "cp2_densecl_match: embd_q / embd_k must be 16-byte, scratch 8-byte aligned");
  auto *best = static_cast<unsigned long long *>(scratch);
  auto *counters = reinterpret_cast<unsigned int *>(best + (size_t)B * P);
  const int nt = cp2_cdiv(P, 128);
  const dim3 grid(nt, nt, B);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
#define CP2_MATCH(KSV, NGV, NDV, NSV)                                                                              \
  do {                                                                                                             \
    CP2_SMEM_ONCE((densecl_match_kernel<KSV, NGV, NDV, NSV>), (MatchCfg<KSV, NGV, NDV, NSV>::SMEM));              \
    densecl_match_kernel<KSV, NGV, NDV, NSV><<<grid, MT_THREADS, MatchCfg<KSV, NGV, NDV, NSV>::SMEM, st>>>(       \
        embd_q, embd_k, reinterpret_cast<long long *>(idx), best, counters, Cb, P);                                \
  } while (0)
  // measured on B200 at B=32, P=196, Cb=2048 (ncu, isolated): (KS, NG, ND, NS) = (16,4,8,6) 60 us, (32,2,4,3) 62 us,
  // (32,4,4,3) 64 us, (16,2,8,4) 67 us, (16,1,8,4) 75 us
  CP2_MATCH(16, 4, 8, 6);
#undef CP2_MATCH
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}
