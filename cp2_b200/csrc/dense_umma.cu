// K2 — dense pixel InfoNCE on the 5th-gen tensor cores (reference builder.py:1289-1292,
// 1431-1437 and their autograd).  The [B,P,P] logit matrix never exists in HBM.
//
// Operands: qhat_t / khat_t [B,P,C] bf16 (unit pixel vectors, written by cp2_prep_features),
// fetched by TMA as (rows x 64-channel) boxes with the 128-byte swizzle; rows past P are
// zero-filled by the TMA unit.  With `key_count` (compacted keys: foreground rows first, zero rows
// after) only ceil(count/tile) key tiles are visited by either pass: background key pixels carry
// no loss term (labels = mask_a (x) mask_b) and no gradient (P~ = P . mask_b).
//
// Forward  (cp2_dense_lse):  CTA = (image n, 128 key pixels y), 2 CTAs per SM.  S^T tile =
//   Khat_y . Qhat_x^T (M = y -> TMEM lanes, N = x -> TMEM columns), so the log-sum-exp over the
//   QUERY axis (the reference's LogSoftmax(dim=1)) is a per-lane running sum: no cross-lane
//   traffic and, because |S| <= 1/tau for unit vectors, no running max.
// Backward (cp2_dense_grad): persistent CTA (one per SM) over work items (image n, 128 query
//   pixels x).  Per key tile: S = Qhat_x . Khat_y^T (double-buffered in TMEM) -> 8 softmax warps
//   turn it into P~ = exp2(S*c - lse_y*log2e) as bf16 *in place in TMEM* -> dQhat += P~ . Khat_y
//   with A read from TMEM and B the same shared-memory key tile addressed MN-major.  While the
//   softmax warps work on tile j the tensor pipe computes S(j+1) and dQ(j-1).  4 epilogue warps add
//   the rank-1 terms, apply the F.normalize backward and write dq [B,C,P] coalesced along P while
//   the next item is already running (dQhat accumulator double-buffered when C = 128).
//
// The exponentials are the co-bottleneck (16 MUFU ex2 / clk / SM vs 8192 MMA flop / clk / SM at
// C = 128): PF of every 4 column pairs are evaluated on the FMA pipe instead (Cody-Waite +
// minimax polynomial, packed FFMA2; umma.cuh), the rest on the MUFU.
#include <cstdlib>

#include "common.cuh"
#include "cp2_b200_debug.h"
#include "umma.cuh"
#include "dense_common.cuh"
#include "weighted_common.cuh"

using namespace umma;
using namespace dense;

namespace {

// ------------------------------------------------------------------ forward
// warp 0 TMA, warp 1 MMA + TMEM owner, warps 2..2+NSW-1 softmax.  NSW = 4: one warp per TMEM lane quadrant
// takes all 128 query columns of a tile; NSW = 8: two warps per quadrant take 64 columns each.  The
// FMUL2 -> MUFU -> FADD2 stream of one warp issues in order, so the MUFU pipe only saturates with >= 4
// such warps per SM sub-partition (dev/mb/microbench.cu: 12.5 ex2/clk/SM with 2 warps, 15.1 with 4).
template <int C, int NST, int PF, int NSW>
__global__ void __launch_bounds__(64 + 32 * NSW, 2)
dense_lse_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                 const int *__restrict__ key_count, float *__restrict__ lse, float scale_log2, int P) {
  constexpr int BOX = TILE * 128;
  constexpr int OPER = (C / 64) * BOX;
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = align1024(smem_raw);
  uint8_t *sK = smem;
  uint8_t *sQ = smem + OPER;
  uint64_t *bars = reinterpret_cast<uint64_t *>(sQ + NST * OPER);
  uint64_t *bar_k = bars, *full = bars + 1, *empty = full + NST, *tfull = empty + NST, *tempty = tfull + 2;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tempty + 2);
  float *half_sum = reinterpret_cast<float *>(tmem_slot + 4);      // [TILE] partial sums of the upper column half
  constexpr int CW = TILE / (NSW / 4);                             // query columns per softmax warp

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.y, y0 = blockIdx.x * TILE;
  const int nx = (P + TILE - 1) / TILE;
  const int cnt = key_count ? min(max(key_count[b], 0), P) : P;
  if (y0 >= cnt) {                       // whole tile is background: no loss term reads it
    const int y = y0 + threadIdx.x;
    if (threadIdx.x < TILE && y < P) lse[(long long)b * P + y] = 0.f;
    return;
  }

  if (threadIdx.x == 0) {
    mbar_init(bar_k, 1);
    for (int s = 0; s < NST; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], NSW); }
    fence_barrier_init();
    prefetch_tmap(&tm_q);
    prefetch_tmap(&tm_k);
  }
  if (warp == 1) tmem_alloc(tmem_slot, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(bar_k, OPER);
      for (int kb = 0; kb < C / 64; ++kb) tma_load_3d(sK + kb * BOX, &tm_k, bar_k, kb * 64, y0, b);
      for (int i = 0; i < nx; ++i) {
        const int s = i % NST, ph = (i / NST) & 1;
        mbar_wait(&empty[s], ph ^ 1);
        mbar_expect_tx(&full[s], OPER);
        for (int kb = 0; kb < C / 64; ++kb)
          tma_load_3d(sQ + s * OPER + kb * BOX, &tm_q, &full[s], kb * 64, i * TILE, b);
      }
    }
  } else if (warp == 1) {
    // whole warp converged (loop state stays warp-uniform), one elected lane issues
    constexpr uint32_t idesc = make_idesc(TILE, TILE, FMT_BF16, FMT_BF16, MAJOR_K, MAJOR_K);
    const uint64_t dK = make_sdesc(smem_u32(sK), 16, 1024);
    const uint64_t dQ0 = make_sdesc(smem_u32(sQ), 16, 1024);
    mbar_wait(bar_k, 0);
    for (int i = 0; i < nx; ++i) {
      const int s = i % NST, ph = (i / NST) & 1, buf = i & 1, tph = (i >> 1) & 1;
      mbar_wait(&full[s], ph);
      mbar_wait(&tempty[buf], tph ^ 1);
      tc_fence_after();
      if (elect_one()) {
        const uint64_t dQ = sdesc_advance(dQ0, s * OPER);
        const uint32_t tacc = tmem + buf * TILE;
#pragma unroll
        for (int kk = 0; kk < C / 16; ++kk) {
          const uint32_t off = (kk >> 2) * BOX + (kk & 3) * 32;
          mma_f16_ss(tacc, sdesc_advance(dK, off), sdesc_advance(dQ, off), idesc, kk > 0);
        }
        mma_commit(&empty[s]);
        mma_commit(&tfull[buf]);
      }
      __syncwarp();
    }
  } else {
    const int quad = warp & 3, half = (warp - 2) >> 2, c0 = half * CW;
    const int y = y0 + quad * 32 + lane;
    const uint32_t lane_base = tmem + (uint32_t(quad * 32) << 16) + c0;
    const f2 scale2 = f2_pack(scale_log2, scale_log2);
    f2 acc_m = f2_pack(0.f, 0.f), acc_p = f2_pack(0.f, 0.f), acc_m2 = acc_m, acc_p2 = acc_m;
    float tail = 0.f;
    for (int i = 0; i < nx; ++i) {
      const int buf = i & 1, tph = (i >> 1) & 1;
      mbar_wait(&tfull[buf], tph);
      tc_fence_after();
      const int nvalid = min(CW, P - i * TILE - c0);     // valid columns of this warp's slice (may be <= 0)
      // software pipeline: the TMEM load of chunk ch+1 is in flight while chunk ch is exponentiated
      uint32_t r[2][32];
      tmem_ld32(lane_base + buf * TILE, r[0]);
#pragma unroll
      for (int ch = 0; ch < CW / 32; ++ch) {
        tmem_wait_ld();
        if (ch + 1 < CW / 32 && (ch + 1) * 32 < nvalid) tmem_ld32(lane_base + buf * TILE + (ch + 1) * 32, r[(ch + 1) & 1]);
        if (ch * 32 < nvalid) {                    // warp-uniform
          const uint32_t(&rc)[32] = r[ch & 1];
          if ((ch + 1) * 32 <= nvalid) {
            softmax_chunk_fwd<PF>(rc, scale2, acc_m, acc_m2, acc_p, acc_p2);
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (ch * 32 + j < nvalid) tail += ex2_approx(__uint_as_float(rc[j]) * scale_log2);
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[buf]);
    }
    float s0, s1;
    f2_unpack(f2_add(f2_add(acc_m, acc_m2), f2_add(acc_p, acc_p2)), s0, s1);
    float total = s0 + s1 + tail;
    if constexpr (NSW == 8) {
      if (half == 1) half_sum[quad * 32 + lane] = total;
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (half == 0) total += half_sum[quad * 32 + lane];
    }
    if (half == 0 && y < P) lse[(long long)b * P + y] = log2f(total) * LN2;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 256);
}

// ----------------------------------------------------------------- backward
// nl[b, y] = -lse[b,y]*log2(e) for key columns that carry gradient, NL_MASKED otherwise; rows padded
// to a multiple of 128 so every key tile's slice is one aligned bulk copy.
__global__ void dense_nl_kernel(const float *__restrict__ lse, const float *__restrict__ mask_b,
                                const int *__restrict__ key_count, float *__restrict__ nl, int P, int P_pad) {
  const int b = blockIdx.y;
  const int y = blockIdx.x * blockDim.x + threadIdx.x;
  if (y >= P_pad) return;
  float v = NL_MASKED;
  if (y < P) {
    bool on = true;
    if (mask_b) on = mask_b[(long long)b * P + y] != 0.f;
    if (key_count) on = on && (y < key_count[b]);
    if (on) v = -lse[(long long)b * P + y] * LOG2E;
  }
  nl[(long long)b * P_pad + y] = v;
}

// warp 0 TMA, warp 1 MMA + TMEM owner, then 4*NSG softmax warps (NSG column groups x 4 TMEM lane quadrants), then 4
// epilogue warps.  NSG = 2: 448 threads (TY = 128: two 64-column groups); NSG = 3: 576 threads (TY = 192).
constexpr int grad_threads(int nsg) { return 64 + 128 * nsg + 128; }
constexpr int GRAD_THREADS = grad_threads(2);

template <int C, int TY, int NST, int NQ, int ND>
struct GradCfg {
  static constexpr int QBOX = TILE * 128;              // query box: 128 rows x 128 B
  static constexpr int KBOX = TY * 128;                // key box:   TY rows x 128 B
  static constexpr int OPER_Q = (C / 64) * QBOX;
  static constexpr int OPER_K = (C / 64) * KBOX;
  static constexpr int NBARS = 2 * NQ + 2 * NST + 4 + 2 * ND;
  static constexpr size_t SMEM = 1024 + (size_t)NQ * OPER_Q + (size_t)NST * OPER_K + (size_t)NST * TY * 4 +
                                 2 * C * 4 + NBARS * 8 + 16;
  static constexpr size_t W_EXTRA = (size_t)NST * TY * 20;   // MODE 2: per-stage b (f32), pixel ids, region ids (int64) of the key tile
  static_assert(2 * TY + ND * C <= 512, "TMEM budget");
};

// MODE 2 (f3, PROPOSED modes on maps that do not fit TMEM): what the weighted soft-max needs besides the operands
struct GradW {
  const long long *ids;        // [4][B][P_pad] pid_a, pid_b, rid_a, rid_b gathered at the lattice (zero past P)
  const float *b_pad;          // [B][P_pad] key-side mask (zero past P)
  const float *wstats;         // [B][3] centre, A, Bm
  weighted::WParams p;
};

// MODE 0: the dense loss's backward (items = (image, 128 query pixels), keys = the image's key pixels).
// MODE 1: queue sums of an InfoNCE with MANY rows (f2, builder.py:866-873): one "image", queries = the N rows,
//         keys = the K queue entries cut into B = nsplit contiguous ranges of `tps` key tiles; items = (range, 128
//         rows).  nl holds -log2(e)/tau for real entries (NL_MASKED past K), so P~ = exp((l-1)/tau); the softmax
//         warps also keep fp32 row sums of P~ (z), and the epilogue writes the raw accumulator:
//           dq[(s*P + r)*C + c] = sum_{k in range s} P~[r,k] u_k[c],   zp[(2s+half)*P_pad + r] = partial z.
// MODE 2: MODE 0 with the PROPOSED-mode pair weights and NegativeType squash (builder.py:1204-1243, 1332-1386):
//         the softmax warps write g[x,y] = b_y (exp(z - lse_y) - a_x/A) w t'(S) instead of P~; the key tile's mask and
//         ids travel with the tile (bulk copies into the stage), the query pixel's own sit in registers.
template <int C, int TY, int NST, int NQ, int ND, int PF, int MODE = 0, int NSG = 2>
__global__ void __launch_bounds__(grad_threads(NSG), 1)
dense_grad_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                  const __nv_bfloat16 *__restrict__ qhat, const float *__restrict__ nl,
                  const int *__restrict__ key_count, const float *__restrict__ mask_a,
                  const float *__restrict__ inv_norm, const float *__restrict__ alpha,
                  const float *__restrict__ v1, const float *__restrict__ v2,
                  const float *__restrict__ gscale, float *__restrict__ dq,
                  float scale_log2, int B, int P, int P_pad, long long *__restrict__ trace,
                  float *__restrict__ zp, int tps, int ktiles, const GradW gw) {
  using Cfg = GradCfg<C, TY, NST, NQ, ND>;
  // optional per-CTA wait-time accounting (test-only; cp2_dbg_dense_trace): slot k of trace[blockIdx.x*16+k]
  long long tr_a = 0, tr_b = 0, tr_c = 0, tr_d = 0;
  const long long tr_start = trace ? clock64() : 0;
  auto twait = [&](uint64_t *bar, uint32_t parity, long long &acc) {
    if (trace) {
      const long long t = clock64();
      mbar_wait(bar, parity);
      acc += clock64() - t;
    } else {
      mbar_wait(bar, parity);
    }
  };
  constexpr int W = TY / NSG;                            // S columns per softmax warp
  static_assert(W % 32 == 0, "softmax column groups are whole 32-column chunks");
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = align1024(smem_raw);
  uint8_t *sQ = smem;
  uint8_t *sK = sQ + NQ * Cfg::OPER_Q;
  float *nlbuf = reinterpret_cast<float *>(sK + NST * Cfg::OPER_K);      // [NST][TY]
  float *sva = nlbuf + NST * TY, *svb = sva + C;                          // [C] each (epilogue)
  long long *pbbuf = reinterpret_cast<long long *>(svb + C);              // MODE 2: [NST][TY] each
  long long *rbbuf = pbbuf + (MODE == 2 ? NST * TY : 0);
  float *bbuf = reinterpret_cast<float *>(rbbuf + (MODE == 2 ? NST * TY : 0));
  uint64_t *bars = reinterpret_cast<uint64_t *>(bbuf + (MODE == 2 ? NST * TY : 0));
  uint64_t *q_full = bars, *q_empty = q_full + NQ, *kfull = q_empty + NQ, *kempty = kfull + NST,
           *s_full = kempty + NST, *p_full = s_full + 2, *d_full = p_full + 2, *d_empty = d_full + ND;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(d_empty + ND);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nxt = (P + TILE - 1) / TILE;
  const int items = B * nxt;

  if (threadIdx.x == 0) {
    for (int i = 0; i < NQ; ++i) { mbar_init(&q_full[i], 1); mbar_init(&q_empty[i], 1); }
    for (int s = 0; s < NST; ++s) { mbar_init(&kfull[s], 1); mbar_init(&kempty[s], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&s_full[i], 1); mbar_init(&p_full[i], 4 * NSG); }
    for (int i = 0; i < ND; ++i) { mbar_init(&d_full[i], 1); mbar_init(&d_empty[i], 4); }
    fence_barrier_init();
    prefetch_tmap(&tm_q);
    prefetch_tmap(&tm_k);
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tmem_D = tmem + 2 * TY;

  auto key_tiles = [&](int b) {
    if constexpr (MODE == 1) return max(0, min(tps, ktiles - b * tps));
    const int cnt = key_count ? min(max(key_count[b], 0), P) : P;
    return (cnt + TY - 1) / TY;
  };
  // MODE 1: the operands are single "images" (batch coordinate 0), the key tiles of range b start at b * tps
  auto img = [&](int b) { return MODE == 1 ? 0 : b; };
  auto tile0 = [&](int b) { return MODE == 1 ? b * tps : 0; };

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int g = 0, it = 0;
      for (int w = blockIdx.x; w < items; w += gridDim.x, ++it) {
        const int b = w / nxt, x0 = (w % nxt) * TILE, ny = key_tiles(b);
        const int qb = it % NQ;
        twait(&q_empty[qb], ((it / NQ) & 1) ^ 1, tr_a);
        mbar_expect_tx(&q_full[qb], Cfg::OPER_Q);
        for (int kb = 0; kb < C / 64; ++kb)
          tma_load_3d(sQ + qb * Cfg::OPER_Q + kb * Cfg::QBOX, &tm_q, &q_full[qb], kb * 64, x0, img(b));
        for (int j = 0; j < ny; ++j, ++g) {
          const int s = g % NST;
          twait(&kempty[s], ((g / NST) & 1) ^ 1, tr_b);
          mbar_expect_tx(&kfull[s], Cfg::OPER_K + TY * 4 + (MODE == 2 ? TY * 20 : 0));
          for (int kb = 0; kb < C / 64; ++kb)
            tma_load_3d(sK + s * Cfg::OPER_K + kb * Cfg::KBOX, &tm_k, &kfull[s], kb * 64, (tile0(b) + j) * TY, img(b));
          bulk_load_1d(nlbuf + s * TY, nl + (MODE == 1 ? 0 : (long long)b * P_pad) + (long long)(tile0(b) + j) * TY,
                       TY * 4, &kfull[s]);
          if constexpr (MODE == 2) {
            const long long o = (long long)b * P_pad + j * TY, plane = (long long)B * P_pad;
            bulk_load_1d(bbuf + s * TY, gw.b_pad + o, TY * 4, &kfull[s]);
            bulk_load_1d(pbbuf + s * TY, gw.ids + plane + o, TY * 8, &kfull[s]);
            bulk_load_1d(rbbuf + s * TY, gw.ids + 3 * plane + o, TY * 8, &kfull[s]);
          }
        }
      }
    }
  } else if (warp == 1) {
    // -------------------------------------------------------------- MMA issuer
    // The whole warp runs the loop (state stays warp-uniform -> descriptors live in uniform registers and
    // one MMA costs ~3 instructions to issue); an elected lane issues tcgen05.mma / commit.
    constexpr uint32_t idesc_s = make_idesc(TILE, TY, FMT_BF16, FMT_BF16, MAJOR_K, MAJOR_K);
    constexpr uint32_t idesc_d = make_idesc(TILE, C, FMT_BF16, FMT_BF16, MAJOR_K, MAJOR_MN);
    const uint64_t dQ0 = make_sdesc(smem_u32(sQ), 16, 1024);
    const uint64_t dK0 = make_sdesc(smem_u32(sK), 16, 1024);               // K-major view (S GEMM)
    const uint64_t dKmn0 = make_sdesc(smem_u32(sK), Cfg::KBOX, 1024);      // MN-major view (dQ GEMM)
    int g = 0, it = 0;
    for (int w = blockIdx.x; w < items; w += gridDim.x, ++it) {
      const int b = w / nxt, ny = key_tiles(b);
      const int qb = it % NQ, db = it % ND;
      const uint64_t dQ = sdesc_advance(dQ0, qb * Cfg::OPER_Q);
      const uint32_t tD = tmem_D + db * C;
      twait(&q_full[qb], (it / NQ) & 1, tr_a);
      twait(&d_empty[db], ((it / ND) & 1) ^ 1, tr_b);
      auto issue_s = [&](int gi) {                       // S(gi) = Qhat_x . Khat_y^T -> S buffer gi&1
        const int s = gi % NST;
        twait(&kfull[s], (gi / NST) & 1, tr_c);
        tc_fence_after();
        if (elect_one()) {
          const uint64_t dK = sdesc_advance(dK0, s * Cfg::OPER_K);
          const uint32_t tS = tmem + (gi & 1) * TY;
#pragma unroll
          for (int kk = 0; kk < C / 16; ++kk)
            mma_f16_ss(tS, sdesc_advance(dQ, (kk >> 2) * Cfg::QBOX + (kk & 3) * 32),
                       sdesc_advance(dK, (kk >> 2) * Cfg::KBOX + (kk & 3) * 32), idesc_s, kk > 0);
          mma_commit(&s_full[gi & 1]);
        }
        __syncwarp();
      };
      auto release_q = [&]() {                           // all S MMAs of the item issued: Qhat tile is free
        if (elect_one()) mma_commit(&q_empty[qb]);
        __syncwarp();
      };
      if (ny > 0) issue_s(g);
      if (ny <= 1) release_q();
      for (int j = 0; j < ny; ++j, ++g) {
        if (j + 1 < ny) {
          issue_s(g + 1);
          if (j + 2 == ny) release_q();
        }
        const int s = g % NST;
        twait(&p_full[g & 1], (g >> 1) & 1, tr_d);
        tc_fence_after();
        // dQhat += P~ . Khat_y : A = bf16 pairs in TMEM (8 columns per K=16 slice; each softmax warp
        // packed its W columns at the start of its own region), B = the key tile read MN-major:
        // 16 key rows (2 KB) per slice, 64-channel chunks KBOX apart.
        if (elect_one()) {
          const uint64_t dKmn = sdesc_advance(dKmn0, s * Cfg::OPER_K);
          const uint32_t tS = tmem + (g & 1) * TY;
#pragma unroll
          for (int kk = 0; kk < TY / 16; ++kk) {
            const int half = (kk * 16) / W, local = (kk * 16) % W;
            mma_f16_ts(tD, tS + half * W + local / 2, sdesc_advance(dKmn, kk * 2048), idesc_d,
                       (j > 0 || kk > 0) ? 1u : 0u);
          }
          mma_commit(&kempty[s]);
        }
        __syncwarp();
      }
      if (elect_one()) mma_commit(&d_full[db]);
      __syncwarp();
    }
  } else if (warp < 2 + 4 * NSG) {
    // ------------------------------------------------ softmax warps: S -> P~ (bf16, in place)
    const int quad = warp & 3, half = (warp - 2) >> 2;
    const uint32_t lane_off = uint32_t(quad * 32) << 16;
    const f2 scale2 = f2_pack(scale_log2, scale_log2);
    int g = 0;
    for (int w = blockIdx.x; w < items; w += gridDim.x) {
      const int ny = key_tiles(w / nxt);
      f2 zsum = f2_pack(0.f, 0.f);
      // MODE 2: this lane's query pixel
      float w_ax = 0.f, w_aoa = 0.f, w_centre = 0.f;
      long long w_pa = 0, w_ra = 0;
      bool w_vx = false;
      if constexpr (MODE == 2) {
        const int wb = w / nxt, x = (w % nxt) * TILE + quad * 32 + lane;
        w_vx = x < P;
        const long long o = (long long)wb * P_pad + (w_vx ? x : 0);
        w_ax = w_vx ? mask_a[(long long)wb * P + x] : 0.f;
        w_pa = w_vx ? gw.ids[o] : 0;
        w_ra = w_vx ? gw.ids[2ll * B * P_pad + o] : 0;
        w_aoa = w_ax / gw.wstats[3 * wb + 1];
        w_centre = gw.p.neg_type >= 2 ? gw.wstats[3 * wb] : 0.f;
      }
      for (int j = 0; j < ny; ++j, ++g) {
        const int s = g % NST;
        const uint32_t tS = tmem + (g & 1) * TY + lane_off + half * W;
        twait(&kfull[s], (g / NST) & 1, tr_a);           // nl slice of this tile landed (bulk copy)
        twait(&s_full[g & 1], (g >> 1) & 1, tr_b);
        tc_fence_after();
        const uint32_t nl_saddr = smem_u32(nlbuf + s * TY + half * W);
        uint32_t rb[2][32];
        tmem_ld32(tS, rb[0]);
#pragma unroll
        for (int ch = 0; ch < W / 32; ++ch) {
          uint32_t pk[16];
          tmem_wait_ld();
          if (ch + 1 < W / 32) tmem_ld32(tS + (ch + 1) * 32, rb[(ch + 1) & 1]);
          const uint32_t(&r)[32] = rb[ch & 1];
          if constexpr (MODE == 2) {
            const int c0 = s * TY + half * W + ch * 32;          // first key of the chunk inside the stage buffers
#pragma unroll
            for (int jj = 0; jj < 32; jj += 2) {
              float g2[2];
#pragma unroll
              for (int e = 0; e < 2; ++e) {
                const float by = bbuf[c0 + jj + e];              // 0 past P
                float gv = 0.f;
                if (by != 0.f && w_vx) {
                  float z2, wd;
                  weighted::pair_eval(__uint_as_float(r[jj + e]), w_ax != 0.f, w_pa, pbbuf[c0 + jj + e], w_ra,
                                      rbbuf[c0 + jj + e], w_centre, gw.p, scale_log2, z2, wd);
                  gv = by * (ex2_approx(z2 + nlbuf[c0 + jj + e]) - w_aoa) * wd;
                }
                g2[e] = gv;
              }
              pk[jj >> 1] = pack_bf16x2(g2[0], g2[1]);
            }
          } else {
            softmax_chunk_bwd<PF, MODE == 1>(r, nl_saddr + ch * 128, scale2, pk, &zsum);
          }
          tmem_st16(tS + ch * 16, pk);
        }
        tmem_wait_st();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&p_full[g & 1]);
      }
      if constexpr (MODE == 1) {
        float z0, z1;
        f2_unpack(zsum, z0, z1);
        const int x = (w % nxt) * TILE + quad * 32 + lane;
        static_assert(MODE == 0 || NSG == 2, "queue-sums mode: two column groups");
        zp[(long long)(2 * (w / nxt) + half) * P_pad + x] = z0 + z1;
      }
    }
  } else {
    // ---- epilogue warps: rank-1 terms, F.normalize backward, coalesced store along P
    const int quad = warp & 3;
    const int row = quad * 32 + lane;
    const uint32_t lane_off = uint32_t(quad * 32) << 16;
    const float gup = gscale ? __ldg(gscale) : 1.f;          // upstream gradient (device scalar)
    int it = 0;
    for (int w = blockIdx.x; w < items; w += gridDim.x, ++it) {
      const int b = w / nxt, x0 = (w % nxt) * TILE, ny = key_tiles(b);
      const int db = it % ND;
      const int x = x0 + row;
      const bool valid = x < P;
      if constexpr (MODE == 1) {                     // queue sums: the raw accumulator row, 512 B per thread
        twait(&d_full[db], (it / ND) & 1, tr_a);
        tc_fence_after();
        const uint32_t tDq = tmem_D + db * C + lane_off;
        float4 *dst = reinterpret_cast<float4 *>(dq + ((long long)b * P + (valid ? x : 0)) * C);
#pragma unroll 1
        for (int ch = 0; ch < C / 32; ++ch) {
          uint32_t r[32];
          if (ny > 0) {
            tmem_ld32(tDq + ch * 32, r);
            tmem_wait_ld();
          } else {
#pragma unroll
            for (int i = 0; i < 32; ++i) r[i] = 0u;
          }
          if (valid) {
#pragma unroll
            for (int u = 0; u < 8; ++u)
              dst[ch * 8 + u] = make_float4(__uint_as_float(r[4 * u]), __uint_as_float(r[4 * u + 1]),
                                            __uint_as_float(r[4 * u + 2]), __uint_as_float(r[4 * u + 3]));
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&d_empty[db]);
        continue;
      }
      const float a = valid ? mask_a[(long long)b * P + x] : 0.f;
      const float inr = valid ? inv_norm[(long long)b * P + x] : 0.f;
      const float al = (ny > 0) ? alpha[b] * gup : 0.f;
      const uint4 *qrow = reinterpret_cast<const uint4 *>(qhat + ((long long)b * P + (valid ? x : 0)) * C);
      // rank-1 vectors of this image -> shared memory: sva = v1 - v2, svb = v2 (read back as broadcasts)
      const int et = threadIdx.x - (64 + 128 * NSG);           // 0..127 among the epilogue threads
      asm volatile("bar.sync 1, 128;" ::: "memory");           // everyone is done with the previous item's copy
      for (int c = et; c < C; c += 128) {
        const float b2 = v2 ? v2[(long long)b * C + c] * gup : 0.f;
        sva[c] = v1[(long long)b * C + c] * gup - b2;
        svb[c] = b2;
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      // software pipeline over the 2 * C/32 (pass, 32-channel chunk) steps: the Qhat row chunk of step s+1 is
      // loaded while step s computes; the first one is in flight before the accumulator is ready
      constexpr int NCH = C / 32, NSTEP = 2 * NCH;
      uint4 qb[2][4];
#pragma unroll
      for (int u = 0; u < 4; ++u) qb[0][u] = __ldg(qrow + u);
      twait(&d_full[db], (it / ND) & 1, tr_a);
      tc_fence_after();
      const uint32_t tD = tmem_D + db * C + lane_off;
      f2 dot2 = f2_pack(0.f, 0.f), ndot2 = dot2;
      const f2 al2 = f2_pack(al, al), a2 = f2_pack(a, a), inr2 = f2_pack(inr, inr);
      float *dq_row = dq + (long long)b * C * P + x;
#pragma unroll 1
      for (int st2 = 0; st2 < NSTEP; st2 += 2) {
#pragma unroll
      for (int sub = 0; sub < 2; ++sub) {
        const int st = st2 + sub;
        const int ch = st % NCH, pass = st / NCH;
        if (st + 1 < NSTEP) {
#pragma unroll
          for (int u = 0; u < 4; ++u) qb[(sub + 1) & 1][u] = __ldg(qrow + ((st + 1) % NCH) * 4 + u);
        }
        uint32_t r[32];
        if (ny > 0) {
          tmem_ld32(tD + ch * 32, r);
          tmem_wait_ld();
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) r[i] = 0u;
        }
        // packed fp32x2 arithmetic on channel pairs (c, c+1): the bf16 pair of Qhat is one word, the TMEM
        // pair two adjacent registers, the rank-1 vectors come as float4 broadcasts
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const uint4 qv = qb[sub][u];
          const uint32_t qw[4] = {qv.x, qv.y, qv.z, qv.w};
#pragma unroll
          for (int e4 = 0; e4 < 2; ++e4) {
            const float4 va = *reinterpret_cast<const float4 *>(sva + ch * 32 + u * 8 + e4 * 4);
            const float4 vb = *reinterpret_cast<const float4 *>(svb + ch * 32 + u * 8 + e4 * 4);
#pragma unroll
            for (int h2 = 0; h2 < 2; ++h2) {
              const int e = e4 * 4 + h2 * 2, c = ch * 32 + u * 8 + e;          // channels c, c+1
              const uint32_t w = qw[e >> 1];
              const f2 qh2 = f2_pack_u(w << 16, w & 0xffff0000u);
              const f2 va2 = h2 == 0 ? f2_pack(va.x, va.y) : f2_pack(va.z, va.w);
              const f2 vb2 = h2 == 0 ? f2_pack(vb.x, vb.y) : f2_pack(vb.z, vb.w);
              const f2 gq2 = f2_fma(al2, f2_pack_u(r[u * 8 + e], r[u * 8 + e + 1]), f2_fma(a2, va2, vb2));
              if (pass == 0) {
                dot2 = f2_fma(qh2, gq2, dot2);
              } else {
                float o0, o1;
                f2_unpack(f2_mul(f2_fma(qh2, ndot2, gq2), inr2), o0, o1);
                if (valid) {
                  float *dst = dq_row + (long long)c * P;
                  dst[0] = o0;
                  dst[P] = o1;
                }
              }
            }
          }
        }
        if (pass == 0 && ch == NCH - 1) {          // end of the first pass: dot = sum of the two lanes of dot2
          float d0, d1;
          f2_unpack(dot2, d0, d1);
          const float dsum = d0 + d1;
          ndot2 = f2_pack(-dsum, -dsum);
        }
      }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&d_empty[db]);
    }
  }
  if (trace && lane == 0 && (warp <= 2 || warp == 10)) {
    const int slot = warp == 0 ? 0 : warp == 1 ? 2 : warp == 2 ? 6 : 8;
    long long *t = trace + blockIdx.x * 16 + slot;
    t[0] = tr_a; t[1] = tr_b;
    if (warp == 1) { t[2] = tr_c; t[3] = tr_d; }
    if (warp == 0) trace[blockIdx.x * 16 + 10] = clock64() - tr_start;
    if (warp == 2 || warp == 10) trace[blockIdx.x * 16 + (warp == 2 ? 11 : 12)] = clock64() - tr_start;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

// ---------------------------------------------------------------------- host

// Tuning knobs and the wait-time trace exist only in the DEBUG build of this file (libcp2_b200_debug.so,
// -DCP2_DEBUG_BUILD, loaded by tests / dev scripts); the product library has neither getenv on the launch path
// nor a trace entry point.
#ifdef CP2_DEBUG_BUILD
long long *g_dense_trace = nullptr;    // set by cp2_dbg_dense_trace, read by the next cp2_dense_grad
// share of the exponentials evaluated on the FMA pipe, in quarters (0..3)
int poly_quarters(const char *env, int dflt) {
  const char *e = getenv(env);
  if (!e || !*e) return dflt;
  const int v = atoi(e);
  return v < 0 ? 0 : (v > 3 ? 3 : v);
}
bool fwd_nsw8() {
  const char *e = getenv("CP2_DENSE_NSW_FWD");
  return !(e && atoi(e) == 4);
}
int bwd_variant(int dflt) {
  const char *e = getenv("CP2_DENSE_BWD_VARIANT");
  return (e && *e) ? atoi(e) : dflt;
}
#else
constexpr int bwd_variant(int dflt) { return dflt; }
constexpr long long *g_dense_trace = nullptr;
constexpr int poly_quarters(const char *, int dflt) { return dflt; }
constexpr bool fwd_nsw8() { return true; }
#endif

template <int C>
constexpr size_t lse_smem(int nst) { return 1024 + (size_t)(C / 64) * TILE * 128 * (1 + nst) + 256 + TILE * 4; }

template <int C, int PF, int NSW>
int launch_lse(const CUtensorMap &tq, const CUtensorMap &tk, const int *key_count, float *lse, float scale_log2,
               int B, int P, cudaStream_t st) {
  constexpr int NST = 2;
  const size_t smem = lse_smem<C>(NST);
  CP2_SMEM_ONCE((dense_lse_kernel<C, NST, PF, NSW>), smem);
  dim3 grid(cp2_cdiv(P, TILE), B);
  dense_lse_kernel<C, NST, PF, NSW><<<grid, 64 + 32 * NSW, smem, st>>>(tq, tk, key_count, lse, scale_log2, P);
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}

template <int C, int TY, int NST, int NQ, int ND, int PF, int NSG = 2>
int launch_grad(const CUtensorMap &tq, const CUtensorMap &tk, const void *qhat, const float *nl,
                const int *key_count, const float *mask_a, const float *inv_norm, const float *alpha,
                const float *v1, const float *v2, const float *gscale, float *dq, float scale_log2, int B, int P,
                int P_pad, cudaStream_t st) {
  using Cfg = GradCfg<C, TY, NST, NQ, ND>;
  auto kern = dense_grad_kernel<C, TY, NST, NQ, ND, PF, 0, NSG>;
  CP2_SMEM_ONCE(kern, Cfg::SMEM);
  const long long items = (long long)B * cp2_cdiv(P, TILE);
  const int grid = (int)(items < cp2_num_sms() ? items : cp2_num_sms());
  kern<<<grid, grad_threads(NSG), Cfg::SMEM, st>>>(tq, tk, static_cast<const __nv_bfloat16 *>(qhat), nl, key_count,
                                              mask_a, inv_norm, alpha, v1, v2, gscale, dq, scale_log2, B, P, P_pad,
                                              g_dense_trace, nullptr, 0, 0, GradW{});
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}

// f2: queue sums on the backward kernel (MODE 1).  C = 128.
int launch_queue_sums(const CUtensorMap &tq, const CUtensorMap &tk, const float *nl, float *zp, float *gacc,
                      float scale_log2, int nsplit, int N, int N_pad, int tps, int ktiles, cudaStream_t st) {
  using Cfg = GradCfg<128, 128, 5, 1, 2>;
  auto kern = dense_grad_kernel<128, 128, 5, 1, 2, 1, 1>;
  CP2_SMEM_ONCE(kern, Cfg::SMEM);
  const long long items = (long long)nsplit * cp2_cdiv(N, TILE);
  const int grid = (int)(items < cp2_num_sms() ? items : cp2_num_sms());
  kern<<<grid, GRAD_THREADS, Cfg::SMEM, st>>>(tq, tk, nullptr, nl, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr,
                                              nullptr, gacc, scale_log2, nsplit, N, N_pad, nullptr, zp, tps, ktiles,
                                              GradW{});
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}


// f3 backward on the streaming kernel (MODE 2).  C = 128.
int launch_grad_weighted(const CUtensorMap &tq, const CUtensorMap &tk, const void *qhat, const float *nl,
                         const float *mask_a, const float *inv_norm, const float *alpha, const float *v1,
                         const float *v2, const float *gscale, float *dq, float scale_log2, int B, int P, int P_pad,
                         const GradW &gw, cudaStream_t st) {
  using Cfg = GradCfg<128, 128, 5, 1, 2>;
  auto kern = dense_grad_kernel<128, 128, 5, 1, 2, 0, 2>;
  const size_t smem = Cfg::SMEM + Cfg::W_EXTRA;
  CP2_SMEM_ONCE(kern, smem);
  const long long items = (long long)B * cp2_cdiv(P, TILE);
  const int grid = (int)(items < cp2_num_sms() ? items : cp2_num_sms());
  kern<<<grid, GRAD_THREADS, smem, st>>>(tq, tk, static_cast<const __nv_bfloat16 *>(qhat), nl, nullptr, mask_a, inv_norm,
                                         alpha, v1, v2, gscale, dq, scale_log2, B, P, P_pad, nullptr, nullptr, 0, 0, gw);
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}

}  // namespace

extern "C" int cp2_dense_weighted_stream_grad(const void *qhat_t, const void *khat_t, const float *mask_a,
                                              const float *mask_b, const float *lse, const void *ids_pad,
                                              const float *b_pad, const float *wstats, const float *inv_norm,
                                              const float *alpha, const float *v1, const float *v2,
                                              const float *gscale, float *dq, float *scratch_nl, float w_pix,
                                              float w_reg, float w_not, int neg_type, float neg_scale, float inv_tau,
                                              int B, int C, int P, void *stream) {
  CP2_REQUIRE(C == 128, "cp2_dense_weighted_stream_grad: C=%d unsupported (128)", C);
  CP2_REQUIRE(B > 0 && P > 0 && B <= 65535, "cp2_dense_weighted_stream_grad: bad shape B=%d P=%d", B, P);
  CP2_REQUIRE(inv_tau > 0.f && inv_tau <= 16.f, "cp2_dense_weighted_stream_grad: 1/tau=%g outside (0, 16]", (double)inv_tau);
  CP2_REQUIRE(neg_type >= 0 && neg_type <= 3, "cp2_dense_weighted_stream_grad: negative type %d", neg_type);
  CP2_REQUIRE(qhat_t && khat_t && mask_a && mask_b && lse && ids_pad && b_pad && wstats && inv_norm && alpha && v1 && dq &&
              scratch_nl, "cp2_dense_weighted_stream_grad: null pointer");
  CUtensorMap tq, tk;
  int rc;
  if ((rc = make_feature_tmap(&tq, qhat_t, B, C, P, TILE)) != CP2_OK) return rc;
  if ((rc = make_feature_tmap(&tk, khat_t, B, C, P, 128)) != CP2_OK) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int P_pad = cp2_dense_pad(P);
  dense_nl_kernel<<<dim3(cp2_cdiv(P_pad, 256), B), 256, 0, st>>>(lse, mask_b, nullptr, scratch_nl, P, P_pad);
  CP2_LAUNCH_CHECK();
  GradW gw;
  gw.ids = static_cast<const long long *>(ids_pad);
  gw.b_pad = b_pad;
  gw.wstats = wstats;
  gw.p = weighted::WParams{w_pix, w_reg, w_not, neg_scale, inv_tau, neg_type, P, 0, 0, 0, 0};
  return launch_grad_weighted(tq, tk, qhat_t, scratch_nl, mask_a, inv_norm, alpha, v1, v2, gscale, dq, inv_tau * LOG2E, B,
                              P, P_pad, gw, st);
}


// ranges the queue is cut into so that (ranges x row tiles) fills the SMs
extern "C" int cp2_queue_sums_split(int N, int K) {
  if (N <= 0 || K <= 0) return 1;
  const int nxt = cp2_cdiv(N, TILE), ktiles = cp2_cdiv(K, 128);
  int s = cp2_num_sms() / nxt;
  s = s < 1 ? 1 : (s > 8 ? 8 : s);
  return s < ktiles ? s : ktiles;
}

extern "C" int cp2_queue_sums(const void *rows_bf, const void *queue_bf, const float *nl, float *zp, float *gacc,
                              float inv_tau, int N, int C, int K, void *stream) {
  CP2_REQUIRE(C == 128, "cp2_queue_sums: C=%d unsupported (128)", C);
  CP2_REQUIRE(N > 0 && K > 0, "cp2_queue_sums: bad shape N=%d K=%d", N, K);
  CP2_REQUIRE(inv_tau > 0.f && inv_tau <= 16.f, "cp2_queue_sums: 1/tau=%g outside (0, 16]", (double)inv_tau);
  CP2_REQUIRE(rows_bf && queue_bf && nl && zp && gacc, "cp2_queue_sums: null pointer");
  const int nsplit = cp2_queue_sums_split(N, K), ktiles = cp2_cdiv(K, 128), tps = cp2_cdiv(ktiles, nsplit);
  CUtensorMap tq, tk;
  int rc;
  if ((rc = make_feature_tmap(&tq, rows_bf, 1, C, N, TILE)) != CP2_OK) return rc;
  if ((rc = make_feature_tmap(&tk, queue_bf, 1, C, K, 128)) != CP2_OK) return rc;
  return launch_queue_sums(tq, tk, nl, zp, gacc, inv_tau * LOG2E, nsplit, N, cp2_dense_pad(N), tps, ktiles,
                           static_cast<cudaStream_t>(stream));
}

#ifdef CP2_DEBUG_BUILD
extern "C" int cp2_dbg_dense_trace(long long *buf) {   // [n_sms * 16] device counters or NULL (cp2_b200_debug.h)
  g_dense_trace = buf;
  return CP2_OK;
}
#endif

// row length of the nl scratch: whole key tiles of either backward variant (128 / 192 rows)
extern "C" int cp2_dense_pad(int P) { return P > 0 ? cp2_cdiv(P, 384) * 384 : 0; }

extern "C" int cp2_dense_lse(const void *qhat_t, const void *khat_t, const int *key_count, float *lse,
                             float inv_tau, int B, int C, int P, void *stream) {
  CP2_REQUIRE(C == 128 || C == 256, "cp2_dense_lse: C=%d unsupported (128 or 256)", C);
  CP2_REQUIRE(B >= 0 && P > 0 && B <= 65535, "cp2_dense_lse: bad shape B=%d P=%d", B, P);
  CP2_REQUIRE(inv_tau > 0.f && inv_tau <= 16.f, "cp2_dense_lse: 1/tau=%g outside (0, 16]", (double)inv_tau);
  if (B == 0) return CP2_OK;
  CP2_REQUIRE(qhat_t && khat_t && lse, "cp2_dense_lse: null pointer");
  CUtensorMap tq, tk;
  int rc;
  if ((rc = make_feature_tmap(&tq, qhat_t, B, C, P, TILE)) != CP2_OK) return rc;
  if ((rc = make_feature_tmap(&tk, khat_t, B, C, P, TILE)) != CP2_OK) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const float sl2 = inv_tau * LOG2E;
  // measured on B200 (dev/tune_dense.py, P=3136): C=128 best at 8 softmax warps + 1/4 of the exponentials on
  // the FMA pipe; at C=256 (one CTA per SM, MMA twice as long per exponential) the plain MUFU stream wins
  const int pf = poly_quarters("CP2_DENSE_PF_FWD", C == 128 ? 1 : 0);
  const bool nsw8 = fwd_nsw8();
#define CP2_LSE(CC, PFV) (nsw8 ? launch_lse<CC, PFV, 8>(tq, tk, key_count, lse, sl2, B, P, st) \
                               : launch_lse<CC, PFV, 4>(tq, tk, key_count, lse, sl2, B, P, st))
  if (C == 128) {
    switch (pf) { case 0: return CP2_LSE(128, 0); case 1: return CP2_LSE(128, 1); case 2: return CP2_LSE(128, 2);
                  default: return CP2_LSE(128, 3); }
  }
  switch (pf) { case 0: return CP2_LSE(256, 0); case 1: return CP2_LSE(256, 1); case 2: return CP2_LSE(256, 2);
                default: return CP2_LSE(256, 3); }
#undef CP2_LSE
}

extern "C" int cp2_dense_grad(const void *qhat_t, const void *khat_t, const int *key_count, const float *lse,
                              const float *mask_a, const float *mask_b, const float *inv_norm,
                              const float *alpha, const float *v1, const float *v2, const float *gscale,
                              float *dq, float *scratch_nl, float inv_tau, int B, int C, int P, void *stream) {
  CP2_REQUIRE(C == 128 || C == 256, "cp2_dense_grad: C=%d unsupported (128 or 256)", C);
  CP2_REQUIRE(B >= 0 && P > 0 && B <= 65535, "cp2_dense_grad: bad shape B=%d P=%d", B, P);
  CP2_REQUIRE(inv_tau > 0.f && inv_tau <= 16.f, "cp2_dense_grad: 1/tau=%g outside (0, 16]", (double)inv_tau);
  if (B == 0) return CP2_OK;
  CP2_REQUIRE(qhat_t && khat_t && lse && mask_a && inv_norm && alpha && v1 && dq && scratch_nl,
              "cp2_dense_grad: null pointer");
  // C = 128 variants: 0 = key tiles of 128 rows, 8 softmax warps, double-buffered accumulator;
  //                   1 = key tiles of 192 rows, 12 softmax warps (3 column groups of 64), single accumulator
  const int variant = (C == 128) ? bwd_variant(0) : 0;
  const int TY = (C == 128) ? (variant == 1 ? 192 : 128) : 64;
  CUtensorMap tq, tk;
  int rc;
  if ((rc = make_feature_tmap(&tq, qhat_t, B, C, P, TILE)) != CP2_OK) return rc;
  if ((rc = make_feature_tmap(&tk, khat_t, B, C, P, TY)) != CP2_OK) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int P_pad = cp2_dense_pad(P);
  dense_nl_kernel<<<dim3(cp2_cdiv(P_pad, 256), B), 256, 0, st>>>(lse, mask_b, key_count, scratch_nl, P, P_pad);
  CP2_LAUNCH_CHECK();
  const float sl2 = inv_tau * LOG2E;
  const int pf = poly_quarters("CP2_DENSE_PF_BWD", 1);
#define CP2_GRAD(CC, TYV, NSTV, NQV, NDV, PFV)                                                                     \
  launch_grad<CC, TYV, NSTV, NQV, NDV, PFV>(tq, tk, qhat_t, scratch_nl, key_count, mask_a, inv_norm, alpha, v1, \
                                         v2, gscale, dq, sl2, B, P, P_pad, st)
  if (C == 128 && variant == 1) {
    return pf == 0 ? launch_grad<128, 192, 3, 1, 1, 0, 3>(tq, tk, qhat_t, scratch_nl, key_count, mask_a, inv_norm, alpha,
                                                           v1, v2, gscale, dq, sl2, B, P, P_pad, st)
                   : launch_grad<128, 192, 3, 1, 1, 1, 3>(tq, tk, qhat_t, scratch_nl, key_count, mask_a, inv_norm, alpha,
                                                           v1, v2, gscale, dq, sl2, B, P, P_pad, st);
  }
  if (C == 128) {
    switch (pf) { case 0: return CP2_GRAD(128, 128, 5, 1, 2, 0); case 1: return CP2_GRAD(128, 128, 5, 1, 2, 1);
                  case 2: return CP2_GRAD(128, 128, 5, 1, 2, 2); default: return CP2_GRAD(128, 128, 5, 1, 2, 3); }
  }
  switch (pf) { case 0: return CP2_GRAD(256, 64, 4, 1, 1, 0); case 1: return CP2_GRAD(256, 64, 4, 1, 1, 1);
                case 2: return CP2_GRAD(256, 64, 4, 1, 1, 2); default: return CP2_GRAD(256, 64, 4, 1, 1, 3); }
#undef CP2_GRAD
}
