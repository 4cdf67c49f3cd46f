// K3 — global MoCo InfoNCE against the negative queue: forward sums, backward sums and the top-1
// accuracy meter in ONE launch and ONE pass over the queue (reference builder.py:1396, 1420-1428 and
// autograd, 1441 + _accuracy 1690-1706; removes queue.clone()).  HBM-bound.
//
// Operand: a bf16 SHADOW of the queue, entry-major [K, C] (one 2*C-byte row per queue entry), kept in
// step with the fp32 [C,K] buffer (the checkpoint contract) by cp2_enqueue / cp2_queue_shadow.  A tile of
// 128 entries is C/64 TMA boxes with the 128-byte swizzle; that ONE image is the K-major A operand of the
// logits GEMM and the MN-major A operand of the gradient GEMM (no converter warps, half the bytes of the
// fp32 queue, exactly the bf16 rounding the round-1 kernel applied on chip).
//
// Per tile:   GEMM1  D1[k,b]  = sum_c tile[k,c] * (q_hi[b,c] + q_lo[b,c])            (warp 1, tcgen05)
//             epi    e = exp2((D1 - 1) * scale);  z[b] += e;  E[b,k] = bf16(e) -> smem; top-1 test  (warps 2-5)
//             GEMM2  D2[c,b] += sum_k tile[k,c] * E[b,k]                              (warp 1)
// q_pos is split into two bf16 terms (hi + lo) so only the queue's own bf16 rounding enters the logits.
// Every CTA requests ALL its stages up front (up to 128 KB in flight per SM).
//
// acc_ins without a second pass: a logit within MARGIN of l_pos (the bound of the bf16 rounding of a unit
// queue entry against a unit query) is re-scored exactly in fp32 from the fp32 queue; one above l_pos + MARGIN
// is a certain miss.  Rows whose l_pos clears every logit by the margin never touch the fp32 queue.
//
// Cross-CTA reduction: fixed-point (2^-32) 64-bit integer atomics — associative, hence deterministic —
// into a persistent scratch that the last CTA converts, publishes (z, g, hit, acc) and zeroes again.
#include "common.cuh"
#include "umma.cuh"

using namespace umma;

namespace {

constexpr int KT = 128;                 // queue entries per tile (MMA M of GEMM1, K of GEMM2)
constexpr int ROWS = 32;                // batch rows per CTA chunk
constexpr int NTHREADS = 192;           // warp 0 TMA, warp 1 MMA + TMEM owner, warps 2-5 epilogue
constexpr float LOG2E = 1.4426950408889634f;
constexpr float MARGIN = 3.0e-3f;       // > 2^-9 (bf16 RN of unit entries) + fp32 accumulation slack
constexpr float FIX = 4294967296.0f;    // 2^32
constexpr float UNFIX = 2.3283064365386963e-10f;

template <int C, int NB>
struct GCfg {
  static constexpr int NBOX = C / 64;
  static constexpr int TBOX = KT * 128;                 // one tile box: 128 entries x 64 channels (128 B rows)
  static constexpr int TILE = NBOX * TBOX;              // 32 KB (C=128) / 64 KB (C=256)
  static constexpr int NST = (C == 128) ? 4 : 2;
  static constexpr int QBOX = NB * 128;                 // one q / E box: NB rows x 128 B
  static constexpr int QOP = NBOX * QBOX;
  static constexpr int EOP = 2 * QBOX;                  // E[b][128 k]: two 64-column boxes
  static constexpr int NH = C / 128;                    // M=128 halves of the gradient GEMM
  static constexpr int TCOLS_USED = 2 * NB + NH * NB;
  static constexpr uint32_t TMEM_COLS = TCOLS_USED <= 32 ? 32 : TCOLS_USED <= 64 ? 64 : 128;
  static constexpr int NFLOATS = 4 * NB + 4 * NB;       // zs[4][NB], lpos/lhi/llo/miss [NB]
  static constexpr int NBARS = 2 * NST + 2 + 2 + 2 + 2 + 1;
  static constexpr size_t SMEM = 1024 + (size_t)NST * TILE + 2 * QOP + 2 * EOP + NFLOATS * 4 + NBARS * 8 + 32;
};

__device__ __forceinline__ uint8_t *align1024(uint8_t *p) {
  return reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(p) + 1023) & ~uintptr_t(1023));
}

// byte offset of element (row, col) of a K-major bf16 operand stored as 128-byte-swizzled boxes of 64 columns
__device__ __forceinline__ uint32_t sw128_off(int row, int col, int box_bytes) {
  return (uint32_t)((col >> 6) * box_bytes + row * 128 + ((((col & 63) >> 3) ^ (row & 7)) << 4) + (col & 7) * 2);
}

struct GlobalScratch {                  // views into the caller's persistent scratch (all zero between calls)
  long long *acc_z;                     // [chunks*NB]
  long long *acc_g;                     // [chunks*NB*C]
  int *miss;                            // [chunks*NB]
  float *lpos;                          // [chunks*NB] exact l_pos (written by the first CTA of a chunk; no zeroing needed)
  float *lbg;                           // [chunks*NB] max of the two background logits (or -3e38)
  unsigned int *counter;                // [1]
};

template <int C, int NB>
__global__ void __launch_bounds__(NTHREADS, 1)
global_nce_kernel(const __grid_constant__ CUtensorMap tm_q, const float *__restrict__ q_pos,
                  const float *__restrict__ queue_f32, const float *__restrict__ k_pos,
                  const float *__restrict__ q_neg, const float *__restrict__ k_neg, float *__restrict__ z,
                  float *__restrict__ g, float *__restrict__ hit, float *__restrict__ acc, GlobalScratch sc,
                  float scale_log2, int B, int K) {
  using Cfg = GCfg<C, NB>;
  constexpr int NST = Cfg::NST;
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = align1024(smem_raw);
  uint8_t *sT = smem;                               // NST tile stages
  uint8_t *sQh = sT + NST * Cfg::TILE;              // q_pos hi
  uint8_t *sQl = sQh + Cfg::QOP;                    // q_pos lo
  uint8_t *sE = sQl + Cfg::QOP;                     // 2 E buffers
  float *zs = reinterpret_cast<float *>(sE + 2 * Cfg::EOP);       // [4][NB]
  float *lpos = zs + 4 * NB, *lhi = lpos + NB, *llo = lhi + NB;
  int *miss_s = reinterpret_cast<int *>(llo + NB);                // [NB]
  uint64_t *bars = reinterpret_cast<uint64_t *>(miss_s + NB);
  uint64_t *tfull = bars, *tempty = tfull + NST, *d1_full = tempty + NST, *d1_empty = d1_full + 2,
           *e_full = d1_empty + 2, *e_empty = e_full + 2, *d2_full = e_empty + 2;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(d2_full + 1);
  int *last_flag = reinterpret_cast<int *>(tmem_slot + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int chunk = blockIdx.y, b0 = chunk * ROWS;
  const int cta = blockIdx.x, nctas = gridDim.x;
  const bool want_acc = k_pos != nullptr;
  // balanced contiguous range of 128-entry tiles
  const int tiles_all = (K + KT - 1) / KT;
  const int t0 = (int)((long long)cta * tiles_all / nctas);
  const int t1 = (int)((long long)(cta + 1) * tiles_all / nctas);
  const int my_tiles = t1 - t0;

  if (threadIdx.x == 0) {
    for (int s = 0; s < NST; ++s) { mbar_init(&tfull[s], 1); mbar_init(&tempty[s], 1); }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&d1_full[i], 1); mbar_init(&d1_empty[i], 4);
      mbar_init(&e_full[i], 4); mbar_init(&e_empty[i], 1);
    }
    mbar_init(d2_full, 1);
    fence_barrier_init();
    prefetch_tmap(&tm_q);
    // every stage is requested right here: the DRAM round trip overlaps the rest of the prologue
    for (int i = 0; i < min(NST, my_tiles); ++i) {
      mbar_expect_tx(&tfull[i], Cfg::TILE);
      for (int kb = 0; kb < Cfg::NBOX; ++kb)
        tma_load_2d(sT + i * Cfg::TILE + kb * Cfg::TBOX, &tm_q, &tfull[i], kb * 64, (t0 + i) * KT);
    }
  }
  if (warp == 1) tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
  // q_pos -> bf16 hi / lo operands (K-major, rows past B are zero).  Every global load of the prologue is
  // issued before the first use: one memory round trip instead of one per loop iteration.
  {
    constexpr int NPAIR = NB * (C / 2), NU = (NPAIR + NTHREADS - 1) / NTHREADS;
    float2 v[NU];
#pragma unroll
    for (int u = 0; u < NU; ++u) {
      const int idx = threadIdx.x + u * NTHREADS, row = idx / (C / 2), c = (idx % (C / 2)) * 2;
      v[u] = make_float2(0.f, 0.f);
      if (idx < NPAIR && b0 + row < B) v[u] = __ldg(reinterpret_cast<const float2 *>(q_pos + (long long)(b0 + row) * C + c));
    }
    // exact fp32 positive logits and the two decision thresholds of the top-1 test: item = (row, 32-channel
    // segment), the segments of a row sit in adjacent lanes
    constexpr int SEG = C / 32, NITEM = NB * SEG, NIT = (NITEM + NTHREADS - 1) / NTHREADS;
    float lp[NIT], lq[NIT], lk[NIT];
#pragma unroll
    for (int it = 0; it < NIT; ++it) {
      const int item = threadIdx.x + it * NTHREADS, r = item / SEG, seg = item % SEG;
      lp[it] = lq[it] = lk[it] = 0.f;
      if (want_acc && item < NITEM && b0 + r < B) {
        const long long ro = (long long)(b0 + r) * C + seg * 32;
        float4 a[8], bq[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          a[j] = __ldg(reinterpret_cast<const float4 *>(q_pos + ro) + j);
          bq[j] = __ldg(reinterpret_cast<const float4 *>(k_pos + ro) + j);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j)
          lp[it] = fmaf(a[j].x, bq[j].x, fmaf(a[j].y, bq[j].y, fmaf(a[j].z, bq[j].z, fmaf(a[j].w, bq[j].w, lp[it]))));
        if (q_neg) {
          float4 cq[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            bq[j] = __ldg(reinterpret_cast<const float4 *>(q_neg + ro) + j);
            cq[j] = __ldg(reinterpret_cast<const float4 *>(k_neg + ro) + j);
          }
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            lq[it] = fmaf(a[j].x, bq[j].x, fmaf(a[j].y, bq[j].y, fmaf(a[j].z, bq[j].z, fmaf(a[j].w, bq[j].w, lq[it]))));
            lk[it] = fmaf(a[j].x, cq[j].x, fmaf(a[j].y, cq[j].y, fmaf(a[j].z, cq[j].z, fmaf(a[j].w, cq[j].w, lk[it]))));
          }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < NU; ++u) {
      const int idx = threadIdx.x + u * NTHREADS, row = idx / (C / 2), c = (idx % (C / 2)) * 2;
      if (idx < NPAIR) {
        const __nv_bfloat16 h0 = __float2bfloat16_rn(v[u].x), h1 = __float2bfloat16_rn(v[u].y);
        const float l0 = v[u].x - __bfloat162float(h0), l1 = v[u].y - __bfloat162float(h1);
        const uint32_t off = sw128_off(row, c, Cfg::QBOX);
        *reinterpret_cast<uint32_t *>(sQh + off) = pack_bf16x2(__bfloat162float(h0), __bfloat162float(h1));
        *reinterpret_cast<uint32_t *>(sQl + off) = pack_bf16x2(l0, l1);
      }
    }
#pragma unroll
    for (int it = 0; it < NIT; ++it) {
      const int item = threadIdx.x + it * NTHREADS, r = item / SEG, seg = item % SEG;
#pragma unroll
      for (int o = SEG / 2; o > 0; o >>= 1) {
        lp[it] += __shfl_xor_sync(0xffffffffu, lp[it], o);
        lq[it] += __shfl_xor_sync(0xffffffffu, lq[it], o);
        lk[it] += __shfl_xor_sync(0xffffffffu, lk[it], o);
      }
      if (item < NITEM && seg == 0) {
        const bool on = want_acc && b0 + r < B;
        lpos[r] = lp[it];
        lhi[r] = on ? lp[it] + MARGIN : 3.0e38f;       // logit >= lhi: certainly >= l_pos
        llo[r] = on ? lp[it] - MARGIN : 3.0e38f;       // logit <  llo: certainly <  l_pos
        miss_s[r] = 0;
        if (on && cta == 0) {                          // published once per chunk for the last CTA's finalisation
          sc.lpos[chunk * NB + r] = lp[it];
          sc.lbg[chunk * NB + r] = q_neg ? fmaxf(lq[it], lk[it]) : -3.0e38f;
        }
      }
    }
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tmem_D2 = tmem + 2 * NB;

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer (tiles past the first NST)
    if (lane == 0) {
      for (int i = NST; i < my_tiles; ++i) {
        const int s = i % NST, ph = (i / NST) & 1;
        mbar_wait(&tempty[s], ph ^ 1);
        mbar_expect_tx(&tfull[s], Cfg::TILE);
        for (int kb = 0; kb < Cfg::NBOX; ++kb)
          tma_load_2d(sT + s * Cfg::TILE + kb * Cfg::TBOX, &tm_q, &tfull[s], kb * 64, (t0 + i) * KT);
      }
    }
  } else if (warp == 1) {
    // -------------------------------------------------------------- MMA issuer
    if (lane == 0 && my_tiles > 0) {
      constexpr uint32_t idesc1 = make_idesc(KT, NB, FMT_BF16, FMT_BF16, MAJOR_K, MAJOR_K);
      constexpr uint32_t idesc2 = make_idesc(128, NB, FMT_BF16, FMT_BF16, MAJOR_MN, MAJOR_K);
      const uint32_t aQh = smem_u32(sQh), aQl = smem_u32(sQl);
      auto gemm1 = [&](int i) {
        const int s = i % NST, ph = (i / NST) & 1, buf = i & 1, tph = (i >> 1) & 1;
        const uint32_t aT = smem_u32(sT + s * Cfg::TILE);
        mbar_wait(&tfull[s], ph);
        mbar_wait(&d1_empty[buf], tph ^ 1);
        tc_fence_after();
#pragma unroll
        for (int kk = 0; kk < C / 16; ++kk)       // K = channel, 16 per slice (32 B inside a 128-B row)
          mma_f16_ss(tmem + buf * NB, make_sdesc(aT + (kk >> 2) * Cfg::TBOX + (kk & 3) * 32, 16, 1024),
                     make_sdesc(aQh + (kk >> 2) * Cfg::QBOX + (kk & 3) * 32, 16, 1024), idesc1, kk > 0);
#pragma unroll
        for (int kk = 0; kk < C / 16; ++kk)
          mma_f16_ss(tmem + buf * NB, make_sdesc(aT + (kk >> 2) * Cfg::TBOX + (kk & 3) * 32, 16, 1024),
                     make_sdesc(aQl + (kk >> 2) * Cfg::QBOX + (kk & 3) * 32, 16, 1024), idesc1, 1u);
        mma_commit(&d1_full[buf]);
      };
      gemm1(0);
      for (int i = 0; i < my_tiles; ++i) {
        if (i + 1 < my_tiles) gemm1(i + 1);
        const int s = i % NST, eb = i & 1;
        const uint32_t aT = smem_u32(sT + s * Cfg::TILE), aE = smem_u32(sE + eb * Cfg::EOP);
        mbar_wait(&e_full[eb], (i >> 1) & 1);
        tc_fence_after();
        // K = queue entry, 16 tile rows (2 KB) per slice; A read MN-major: 64-channel chunks TBOX apart
#pragma unroll
        for (int h = 0; h < Cfg::NH; ++h)
#pragma unroll
          for (int kk = 0; kk < KT / 16; ++kk)
            mma_f16_ss(tmem_D2 + h * NB, make_sdesc(aT + h * 2 * Cfg::TBOX + kk * 2048, Cfg::TBOX, 1024),
                       make_sdesc(aE + (kk >> 2) * Cfg::QBOX + (kk & 3) * 32, 16, 1024), idesc2,
                       (i > 0 || kk > 0) ? 1u : 0u);
        mma_commit(&tempty[s]);
        mma_commit(&e_empty[eb]);
      }
      mma_commit(d2_full);
    }
  } else {
    // ---------------------------------------------------- epilogue warps 2-5 (TMEM quadrant warp&3)
    const int quad = warp & 3;
    const int row = quad * 32 + lane;               // tile row = queue entry within the tile
    const uint32_t lane_off = uint32_t(quad * 32) << 16;
    float zacc[NB], thi[NB], tlo[NB];                // thresholds in registers: no loads / branches per element
#pragma unroll
    for (int j = 0; j < NB; ++j) { zacc[j] = 0.f; thi[j] = lhi[j]; tlo[j] = llo[j]; }
    for (int i = 0; i < my_tiles; ++i) {
      const int buf = i & 1, tph = (i >> 1) & 1, eb = i & 1;
      const int kg = (t0 + i) * KT + row;
      const bool kvalid = kg < K;
      uint8_t *sEb = sE + eb * Cfg::EOP;
      mbar_wait(&d1_full[buf], tph);
      tc_fence_after();
      mbar_wait(&e_empty[eb], tph ^ 1);           // GEMM2 of tile i-2 has consumed this E buffer
      uint32_t certain = 0u, band = 0u;
#pragma unroll
      for (int ch = 0; ch < NB / 16; ++ch) {
        uint32_t r[16];
        tmem_ld16(tmem + lane_off + buf * NB + ch * 16, r);
        tmem_wait_ld();
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const int bb = ch * 16 + j;
          const float l = __uint_as_float(r[j]);
          float e = ex2_approx(fmaf(l, scale_log2, -scale_log2));
          e = kvalid ? e : 0.f;
          zacc[bb] += e;
          *reinterpret_cast<__nv_bfloat16 *>(sEb + sw128_off(bb, row, Cfg::QBOX)) = __float2bfloat16_rn(e);
          certain |= (l >= thi[bb]) ? (1u << bb) : 0u;
          band |= (l >= tlo[bb]) ? (1u << bb) : 0u;
        }
      }
      if (!kvalid) certain = band = 0u;
      band &= ~certain;
      fence_proxy_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) { mbar_arrive(&e_full[eb]); mbar_arrive(&d1_empty[buf]); }
      // ---- top-1 bookkeeping (off the MMA's critical path: the barriers above are already signalled)
      while (certain) {
        const int bb = __ffs(certain) - 1;
        certain &= certain - 1;
        miss_s[bb] = 1;
      }
      __syncwarp();
      while (band) {
        const int bb = __ffs(band) - 1;
        band &= band - 1;
        if (*reinterpret_cast<volatile int *>(&miss_s[bb])) continue;
        const float *qr = q_pos + (long long)(b0 + bb) * C;
        const float *qc = queue_f32 + kg;
        // 32 queue loads (one sector each, strided by K) in flight per round trip; q_pos as 128-bit broadcasts
        float exact = 0.f;
#pragma unroll 1
        for (int c0 = 0; c0 < C; c0 += 32) {
          float qv[32];
#pragma unroll
          for (int c = 0; c < 32; ++c) qv[c] = __ldg(qc + (long long)(c0 + c) * K);
#pragma unroll
          for (int c = 0; c < 32; c += 4) {
            const float4 q4 = __ldg(reinterpret_cast<const float4 *>(qr + c0 + c));
            exact = fmaf(q4.x, qv[c], fmaf(q4.y, qv[c + 1], fmaf(q4.z, qv[c + 2], fmaf(q4.w, qv[c + 3], exact))));
          }
        }
        if (exact >= lpos[bb]) miss_s[bb] = 1;
      }
    }
    // ---- this CTA's sums -> fixed-point atomics
#pragma unroll
    for (int j = 0; j < NB; ++j) {
      float v = zacc[j];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0) zs[quad * NB + j] = v;
    }
    asm volatile("bar.sync 1, 128;" ::: "memory");
    const int st = threadIdx.x - 64;
    if (my_tiles > 0) {
      if (st < NB) {
        const float zt = zs[st] + zs[NB + st] + zs[2 * NB + st] + zs[3 * NB + st];
        atomicAdd(reinterpret_cast<unsigned long long *>(sc.acc_z + chunk * NB + st),
                  (unsigned long long)__float2ll_rn(zt * FIX));
        if (want_acc && miss_s[st]) atomicOr(sc.miss + chunk * NB + st, 1);
      }
      mbar_wait(d2_full, 0);
      tc_fence_after();
#pragma unroll
      for (int h = 0; h < Cfg::NH; ++h)
#pragma unroll
        for (int ch = 0; ch < NB / 16; ++ch) {
          uint32_t r[16];
          tmem_ld16(tmem_D2 + lane_off + h * NB + ch * 16, r);
          tmem_wait_ld();
#pragma unroll
          for (int j = 0; j < 16; ++j)
            atomicAdd(reinterpret_cast<unsigned long long *>(
                          sc.acc_g + ((long long)(chunk * NB + ch * 16 + j)) * C + h * 128 + row),
                      (unsigned long long)__float2ll_rn(__uint_as_float(r[j]) * FIX));
        }
    }
    // ---- last CTA of the grid: publish and re-zero the scratch
    __threadfence();
    asm volatile("bar.sync 1, 128;" ::: "memory");
    if (st == 0) {
      const unsigned int total = gridDim.x * gridDim.y;
      *last_flag = (atomicAdd(sc.counter, 1u) == total - 1u) ? 1 : 0;
    }
    asm volatile("bar.sync 1, 128;" ::: "memory");
    if (*last_flag) {
      __threadfence();
      const int chunks = gridDim.y;
      const int total = chunks * NB * C;
      // batches of 16 independent L2 loads per thread: a handful of round trips instead of one per element
      for (int base = 0; base < total; base += 128 * 16) {
        long long v[16];
#pragma unroll
        for (int u = 0; u < 16; ++u) {
          const int idx = base + u * 128 + st;
          v[u] = idx < total ? __ldcg(sc.acc_g + idx) : 0;
        }
#pragma unroll
        for (int u = 0; u < 16; ++u) {
          const int idx = base + u * 128 + st;
          if (idx < total) {
            const int rr = idx / C, c = idx - rr * C;
            const int b = (rr / NB) * ROWS + (rr % NB);
            sc.acc_g[idx] = 0;
            if (b < B) g[(long long)b * C + c] = __ll2float_rn(v[u]) * UNFIX;
          }
        }
      }
      float hits = 0.f;
      for (int rr = st; rr < chunks * NB; rr += 128) {
        const int b = (rr / NB) * ROWS + (rr % NB);
        const long long v = __ldcg(sc.acc_z + rr);
        const int ms = __ldcg(sc.miss + rr);
        const float lp = __ldcg(sc.lpos + rr), lb = __ldcg(sc.lbg + rr);
        sc.acc_z[rr] = 0;
        sc.miss[rr] = 0;
        if (b < B) {
          z[b] = __ll2float_rn(v) * UNFIX;
          if (want_acc) {
            const float h = (!ms && lp > lb) ? 1.f : 0.f;
            hit[b] = h;
            hits += h;
          }
        }
      }
      if (want_acc) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) hits += __shfl_xor_sync(0xffffffffu, hits, o);
        if (lane == 0) zs[quad] = hits;
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (st == 0) acc[0] = (zs[0] + zs[1] + zs[2] + zs[3]) * (100.f / (float)B);
      }
      if (st == 0) *sc.counter = 0u;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, Cfg::TMEM_COLS);
}

// queue [C,K] f32 -> shadow [K,C] bf16 (RN), 32 entries x 64 channels per CTA through shared memory
__global__ void __launch_bounds__(256)
queue_shadow_kernel(const float *__restrict__ queue, __nv_bfloat16 *__restrict__ shadow, int C, int K) {
  __shared__ float tile[64][33];
  const int k0 = blockIdx.x * 32, c0 = blockIdx.y * 64;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
#pragma unroll
  for (int r = ty; r < 64; r += 8) {
    const int c = c0 + r, k = k0 + tx;
    tile[r][tx] = (c < C && k < K) ? queue[(long long)c * K + k] : 0.f;
  }
  __syncthreads();
#pragma unroll
  for (int r = ty; r < 32; r += 8) {
    const int k = k0 + r, c = c0 + 2 * tx;
    if (k < K && c + 1 < C)
      *reinterpret_cast<__nv_bfloat162 *>(shadow + (long long)k * C + c) =
          __floats2bfloat162_rn(tile[2 * tx][r], tile[2 * tx + 1][r]);
  }
}

int pick_nb(int B) { return B <= 16 ? 16 : 32; }

int pick_nctas(int K) {
  const int tiles = cp2_cdiv(K, KT);
  return tiles < cp2_num_sms() ? tiles : cp2_num_sms();
}

int make_shadow_tmap(CUtensorMap *tm, const void *base, int C, uint64_t K) {
  EncodeTiledFn enc = get_encode_tiled();
  if (!enc) { cp2_set_error("cuTensorMapEncodeTiled unavailable in this driver"); return CP2_E_UNSUPPORTED; }
  const cuuint64_t dims[2] = {(cuuint64_t)C, K};
  const cuuint64_t strides[1] = {(cuuint64_t)C * 2};
  const cuuint32_t box[2] = {64, KT};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void *>(base), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                         CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { cp2_set_error("cuTensorMapEncodeTiled(queue shadow) failed: CUresult %d", (int)r); return CP2_E_CUDA; }
  return CP2_OK;
}

GlobalScratch carve_scratch(void *scratch, int chunks, int NB, int C) {
  GlobalScratch sc;
  sc.acc_z = static_cast<long long *>(scratch);
  sc.acc_g = sc.acc_z + (size_t)chunks * NB;
  sc.miss = reinterpret_cast<int *>(sc.acc_g + (size_t)chunks * NB * C);
  sc.lpos = reinterpret_cast<float *>(sc.miss + (size_t)chunks * NB);
  sc.lbg = sc.lpos + (size_t)chunks * NB;
  sc.counter = reinterpret_cast<unsigned int *>(sc.lbg + (size_t)chunks * NB);
  return sc;
}

template <int C, int NB>
int launch_global(const CUtensorMap &tq, const float *q_pos, const float *queue, const float *k_pos,
                  const float *q_neg, const float *k_neg, float *z, float *g, float *hit, float *acc,
                  void *scratch, float scale_log2, int B, int K, int nctas, int chunks, cudaStream_t st) {
  using Cfg = GCfg<C, NB>;
  auto kern = global_nce_kernel<C, NB>;
  CP2_SMEM_ONCE(kern, Cfg::SMEM);
  kern<<<dim3(nctas, chunks), NTHREADS, Cfg::SMEM, st>>>(tq, q_pos, queue, k_pos, q_neg, k_neg, z, g, hit, acc,
                                                         carve_scratch(scratch, chunks, NB, C), scale_log2, B, K);
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}

}  // namespace

extern "C" int64_t cp2_global_scratch_bytes(int B, int C) {
  if (B <= 0) return 16;
  const int64_t rows = (int64_t)cp2_cdiv(B, ROWS) * pick_nb(B);
  return rows * 8 + rows * C * 8 + rows * 12 + 16;
}

extern "C" int cp2_queue_shadow(const float *queue, void *queue_bf, int C, int K, void *stream) {
  CP2_REQUIRE(queue && queue_bf, "cp2_queue_shadow: null pointer");
  CP2_REQUIRE(C > 0 && C % 64 == 0 && K > 0, "cp2_queue_shadow: bad shape C=%d K=%d", C, K);
  queue_shadow_kernel<<<dim3(cp2_cdiv(K, 32), C / 64), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      queue, static_cast<__nv_bfloat16 *>(queue_bf), C, K);
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}

extern "C" int cp2_global_nce(const float *q_pos, const void *queue_bf, const float *queue, const float *k_pos,
                              const float *q_neg, const float *k_neg, float *z, float *g, float *hit, float *acc,
                              void *scratch, float inv_tau, int B, int C, int K, void *stream) {
  CP2_REQUIRE(C == 128 || C == 256, "cp2_global_nce: C=%d unsupported (128 or 256)", C);
  CP2_REQUIRE(B >= 0 && K > 0, "cp2_global_nce: bad shape B=%d K=%d", B, K);
  CP2_REQUIRE(inv_tau > 0.f && inv_tau <= 16.f, "cp2_global_nce: 1/tau=%g outside (0, 16]", (double)inv_tau);
  if (B == 0) return CP2_OK;
  CP2_REQUIRE(q_pos && queue_bf && z && g && scratch, "cp2_global_nce: null pointer");
  CP2_REQUIRE((k_pos == nullptr) || (queue && hit && acc), "cp2_global_nce: the top-1 meter needs queue, hit and acc");
  CP2_REQUIRE((q_neg == nullptr) == (k_neg == nullptr), "cp2_global_nce: q_neg and k_neg go together");
  CP2_REQUIRE((reinterpret_cast<uintptr_t>(queue_bf) & 15u) == 0 && (reinterpret_cast<uintptr_t>(q_pos) & 15u) == 0 &&
                  (reinterpret_cast<uintptr_t>(k_pos) & 15u) == 0 && (reinterpret_cast<uintptr_t>(q_neg) & 15u) == 0 &&
                  (reinterpret_cast<uintptr_t>(k_neg) & 15u) == 0 && (reinterpret_cast<uintptr_t>(scratch) & 7u) == 0,
              "cp2_global_nce: queue_bf, q_pos, k_pos, q_neg, k_neg must be 16-byte, scratch 8-byte aligned");
  const int NB = pick_nb(B), chunks = cp2_cdiv(B, ROWS);
  CP2_REQUIRE(chunks <= 65535, "cp2_global_nce: B=%d too large", B);
  const int nctas = pick_nctas(K);
  CUtensorMap tq;
  int rc;
  if ((rc = make_shadow_tmap(&tq, queue_bf, C, (uint64_t)K)) != CP2_OK) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const float sl2 = inv_tau * LOG2E;
#define CP2_GLB(CC, NBV) launch_global<CC, NBV>(tq, q_pos, queue, k_pos, q_neg, k_neg, z, g, hit, acc, scratch, sl2, \
                                                B, K, nctas, chunks, st)
  if (C == 128) return NB == 16 ? CP2_GLB(128, 16) : CP2_GLB(128, 32);
  return NB == 16 ? CP2_GLB(256, 16) : CP2_GLB(256, 32);
#undef CP2_GLB
}
