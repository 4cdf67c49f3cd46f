// K2-small — the dense pixel InfoNCE (reference builder.py:1289-1292, 1431-1437 and their autograd) for
// feature maps that fit the tensor memory of one SM: P <= 256 pixels (14x14 = 196 is the map CP2 trains at).
//
// ONE launch produces the column log-sum-exp of the forward AND the P^2-sized part of the gradient:
//   CTA = (image n, 128 query pixels xt), one CTA per SM, both bf16 operands of the image in shared memory
//   (two TMA loads), all 512 TMEM columns:
//   phase 1  S^T[yt] = Khat[yt] . Qhat^T for every 128-key tile yt (M = y -> TMEM lanes, N = x -> up to 256
//            columns, tcgen05.mma SS form): the softmax over the QUERY axis is a per-lane sum of exp2, no
//            cross-lane traffic, no running max (|S| <= 1/tau for unit vectors).  8 softmax warps share the
//            (key tile, 32-column chunk) units; lse and nl = -log2(sum) go to global / shared memory.
//   phase 2  S[xt] = Qhat[xt] . Khat^T (M = x, N = y) into the columns phase 1 has released, turned in place into
//            P~ = exp2(S*c + nl_y) as bf16 pairs, then acc[x,:] = sum_y P~[x,y] Khat[y,:] with A read from TMEM
//            (TS form) and B the same shared-memory key image addressed MN-major.
//   epilogue acc -> global [B,C,P] (coalesced along P).
// The loss is a scalar, so the rank-1 terms, the upstream gradient and the F.normalize backward are applied to
// `acc` by the backward's only kernel (dense_small_bwd_kernel, elementwise + one dot per pixel): the backward
// recomputes nothing.  The two CTAs of an image with 128 < P <= 256 both run phase 1 (cheaper than a
// cluster-wide exchange at this size).
#include "common.cuh"
#include "umma.cuh"
#include "dense_common.cuh"

using namespace umma;
using namespace dense;

namespace {

constexpr int SM_THREADS = 320;          // warp 0 TMA, warp 1 MMA + TMEM owner, warps 2-9 softmax / epilogue
constexpr int PF_SMALL = 1;              // exponentials evaluated on the FMA pipe: 1 of every 4 column pairs
constexpr float SEG_DELTA = 0.0078125f;  // acc_seg meter: candidate band below the bf16-operand maximum (2^-7)
constexpr int SEG_MAXC = 32;             // ... and the most candidates re-scored per image

template <int C>
struct SmallCfg {
  static size_t smem(int rows) {
    return 1024 + 2 * (size_t)(C / 64) * rows * 128 + 256 * 4 + 2 * 2 * 128 * 4 + 16 * 8 + 16 + 32 * 4 + 16 +
           SEG_MAXC * 4;
  }
};

// acc_seg meter (builder.py:1442-1448): arg-max over ALL pixel pairs of the image's similarity matrix, hit = label
// of that pair.  The CTA xt = 0 of an image finds the largest bf16-operand S while it is in TMEM, collects every
// pair within SEG_DELTA of it (the bound of the operands' bf16 rounding, both ways) and re-scores those exactly in
// fp32 from the encoder outputs, so the decision is the fp32 reference's.
struct SegMeter {
  const float *q_f32, *k_f32;        // [B,C,P] encoder outputs (before normalisation)
  const float *inv_q, *inv_k;        // [B,P] 1 / max(||.||, eps), pixel order
  const int *perm_k;                 // [B,P] pixel of each key row, or NULL (rows in pixel order)
  const float *mask_a, *mask_b;      // [B,P] labels' factors, pixel order
  float *hit;                        // [B] OUT, NULL = meter off
};

template <int C>
__global__ void __launch_bounds__(SM_THREADS, 1)
dense_small_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                   const int *__restrict__ key_count, const float *__restrict__ mask_b,
                   float *__restrict__ lse, float *__restrict__ acc, const SegMeter seg, float scale_log2, int P,
                   int rows) {
  const int box = rows * 128;                      // bytes of one (rows x 64 channels) swizzled TMA box
  const int oper = (C / 64) * box;
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = align1024(smem_raw);
  uint8_t *sK = smem;
  uint8_t *sQ = smem + oper;
  float *nl = reinterpret_cast<float *>(sQ + oper);            // [256]
  float *part = nl + 256;                                       // [2 halves][2 key tiles][128]
  uint64_t *bars = reinterpret_cast<uint64_t *>(part + 512);
  uint64_t *bar_k = bars, *bar_q = bars + 1, *bar_s1 = bars + 2 /* [2] */, *bar_free = bars + 4,
           *bar_s2 = bars + 5, *bar_p = bars + 6 /* [2] */, *bar_d = bars + 8, *bar_cand = bars + 9;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 16);
  float *smax = reinterpret_cast<float *>(tmem_slot + 4);       // [8] per-warp maxima / [8][4] per-warp bests
  int *ncand = reinterpret_cast<int *>(smax + 32);
  int *cand = ncand + 4;                                        // [SEG_MAXC] (x << 16) | key row

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.y, xt = blockIdx.x;
  // both operand loads are requested before anything else (they do not depend on key_count): the memory round
  // trip overlaps the key_count read, the TMEM allocation and the role dispatch
  if (threadIdx.x == 0) {
    mbar_init(bar_k, 1);
    mbar_init(bar_q, 1);
    mbar_init(&bar_s1[0], 1);
    mbar_init(&bar_s1[1], 1);
    mbar_init(bar_free, 8);
    mbar_init(bar_s2, 1);
    mbar_init(&bar_p[0], 4);
    mbar_init(&bar_p[1], 4);
    mbar_init(bar_d, 1);
    mbar_init(bar_cand, 8);
    *ncand = 0;
    fence_barrier_init();
    mbar_expect_tx(bar_k, oper);
    for (int kb = 0; kb < C / 64; ++kb) tma_load_3d(sK + kb * box, &tm_k, bar_k, kb * 64, 0, b);
    mbar_expect_tx(bar_q, oper);
    for (int kb = 0; kb < C / 64; ++kb) tma_load_3d(sQ + kb * box, &tm_q, bar_q, kb * 64, 0, b);
  }
  const int cnt = key_count ? min(max(key_count[b], 0), P) : P;
  const int nyt = (cnt + TILE - 1) / TILE;          // key tiles that carry loss terms (0, 1 or 2)
  const int nch = (P + 31) / 32;                    // 32-column chunks of the query axis

  if (nyt == 0) {                                   // no foreground key: no loss term, zero gradient part
    if (xt == 0)
      for (int y = threadIdx.x; y < P; y += SM_THREADS) lse[(long long)b * P + y] = 0.f;
    const int x = xt * TILE + (threadIdx.x & 127);
    if (threadIdx.x < 256 && x < P)
      for (int c = threadIdx.x >> 7; c < C; c += 2) acc[((long long)b * C + c) * P + x] = 0.f;
    if (threadIdx.x == 0) {                         // the CTA's shared memory must outlive the copies in flight
      if (seg.hit && xt == 0) seg.hit[b] = 0.f;     // every label of the image is 0
      mbar_wait(bar_k, 0);
      mbar_wait(bar_q, 0);
    }
    return;
  }

  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tmem_D = tmem + 256;
  const int W = nyt * 64;                           // S2 columns per softmax-warp half
  const bool meter = seg.hit != nullptr && xt == 0;
  const int nyt_s = meter ? rows / TILE : nyt;      // key tiles phase 1 looks at (the meter needs every key)

  if (warp == 0) {
    // (the two TMA loads were issued at the top of the kernel)
    // acc_seg: this otherwise idle warp scores the candidates exactly in fp32 while the others run phase 2
    if (meter) {
      mbar_wait(bar_cand, 0);                        // every softmax warp has finished its scan: the list is complete
      const int n = min(*reinterpret_cast<volatile int *>(ncand), SEG_MAXC);
      float bestv = -INFINITY;
      int bestf = 0x7fffffff, bestx = 0, besty = 0, bestr = 0;
      // four candidates per round: their index reads, then their 2 x C/32 feature reads per lane, are in flight
      // together (two memory round trips per FOUR candidates instead of per candidate)
      for (int i0 = 0; i0 < n; i0 += 4) {
        int xs[4], yrs[4], ys[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int i = min(i0 + u, n - 1);
          xs[u] = cand[i] >> 16;
          yrs[u] = cand[i] & 0xffff;
          ys[u] = seg.perm_k ? __ldg(seg.perm_k + (long long)b * P + yrs[u]) : yrs[u];
        }
        float qv[4][C / 32], kv[4][C / 32], iq[4], ik[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const float *qp = seg.q_f32 + (long long)b * C * P + xs[u], *kp = seg.k_f32 + (long long)b * C * P + ys[u];
#pragma unroll
          for (int j = 0; j < C / 32; ++j) {
            qv[u][j] = __ldg(qp + (long long)(lane + 32 * j) * P);
            kv[u][j] = __ldg(kp + (long long)(lane + 32 * j) * P);
          }
          iq[u] = __ldg(seg.inv_q + (long long)b * P + xs[u]);
          ik[u] = __ldg(seg.inv_k + (long long)b * P + ys[u]);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          float d = 0.f;
#pragma unroll
          for (int j = 0; j < C / 32; ++j) d = fmaf(qv[u][j] * iq[u], kv[u][j] * ik[u], d);
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
          const int flat = xs[u] * P + ys[u];
          if (i0 + u < n && (d > bestv || (d == bestv && flat < bestf))) {
            bestv = d; bestf = flat; bestx = xs[u]; besty = ys[u]; bestr = yrs[u];
          }
        }
      }
      if (lane == 0) {
        float h = 0.f;
        if (bestf != 0x7fffffff) {
          const float lb = seg.mask_b ? seg.mask_b[(long long)b * P + besty] : (bestr < cnt ? 1.f : 0.f);
          h = seg.mask_a[(long long)b * P + bestx] * lb;
        }
        seg.hit[b] = h;
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    const uint32_t idesc1 = make_idesc(TILE, rows, FMT_BF16, FMT_BF16, MAJOR_K, MAJOR_K);
    const uint32_t idesc2 = make_idesc(TILE, nyt * TILE, FMT_BF16, FMT_BF16, MAJOR_K, MAJOR_K);
    constexpr uint32_t idesc_d = make_idesc(TILE, C, FMT_BF16, FMT_BF16, MAJOR_K, MAJOR_MN);
    const uint64_t dK0 = make_sdesc(smem_u32(sK), 16, 1024);
    const uint64_t dQ0 = make_sdesc(smem_u32(sQ), 16, 1024);
    const uint64_t dKmn0 = make_sdesc(smem_u32(sK), box, 1024);
    mbar_wait(bar_k, 0);
    mbar_wait(bar_q, 0);
    tc_fence_after();
    if (elect_one()) {
      for (int yt = 0; yt < nyt_s; ++yt) {
#pragma unroll
        for (int kk = 0; kk < C / 16; ++kk) {
          const uint32_t off = (kk >> 2) * box + (kk & 3) * 32;
          mma_f16_ss(tmem + yt * 256, sdesc_advance(dK0, off + yt * TILE * 128), sdesc_advance(dQ0, off), idesc1, kk > 0);
        }
        mma_commit(&bar_s1[yt]);
      }
    }
    __syncwarp();
    mbar_wait(bar_free, 0);                          // every softmax warp is done with S^T[0]'s columns
    tc_fence_after();
    if (elect_one()) {
#pragma unroll
      for (int kk = 0; kk < C / 16; ++kk) {
        const uint32_t off = (kk >> 2) * box + (kk & 3) * 32;
        mma_f16_ss(tmem, sdesc_advance(dQ0, off + xt * TILE * 128), sdesc_advance(dK0, off), idesc2, kk > 0);
      }
      mma_commit(bar_s2);
    }
    __syncwarp();
    for (int h = 0; h < 2; ++h) {
      mbar_wait(&bar_p[h], 0);
      tc_fence_after();
      if (elect_one()) {
        for (int kk = 0; kk < W / 16; ++kk) {
          const int key0 = h * W + kk * 16;          // first key row of this K = 16 slice
          if (key0 >= cnt) break;
          mma_f16_ts(tmem_D, tmem + h * W + kk * 8, sdesc_advance(dKmn0, (key0 / 16) * 2048), idesc_d,
                     (h > 0 || kk > 0) ? 1u : 0u);      // half 0 always issues slice 0 (cnt >= 1)
        }
      }
      __syncwarp();
    }
    if (elect_one()) mma_commit(bar_d);
    __syncwarp();
  } else {
    // ------------------------------------------------------------------ softmax warps
    const int quad = warp & 3, half = (warp - 2) >> 2;
    const uint32_t lane_off = uint32_t(quad * 32) << 16;
    const f2 scale2 = f2_pack(scale_log2, scale_log2);
    const int row = quad * 32 + lane;
    // ---- phase 1: units u = yt * nch + ch, first half of the list to half 0
    // units of this half: the chunks ch = half, half + 2, ... of key tile 0, then those of key tile 1 (both halves
    // get the same share of every tile, so the exponentials stay balanced when tile 1 only feeds the meter)
    const int nch_h = (nch - half + 1) / 2;
    const int u_begin = 0, u_end = nyt_s * nch_h;
    float sum0 = 0.f, sum1 = 0.f, vmax = -INFINITY;
    {
      f2 acc_m = f2_pack(0.f, 0.f), acc_p = acc_m, acc_m2 = acc_m, acc_p2 = acc_m;
      float tail = 0.f;
      int cur = -1;
      bool freed = false;
      auto flush = [&]() {
        if (cur >= 0) {
          float s0, s1;
          f2_unpack(f2_add(f2_add(acc_m, acc_m2), f2_add(acc_p, acc_p2)), s0, s1);
          if (cur == 0) sum0 = s0 + s1 + tail; else sum1 = s0 + s1 + tail;
          acc_m = acc_p = acc_m2 = acc_p2 = f2_pack(0.f, 0.f);
          tail = 0.f;
        }
      };
      uint32_t r0[32], r1[32];
      auto issue = [&](int u, uint32_t (&dst)[32]) {
        const int yt = u / nch_h, ch = 2 * (u - yt * nch_h) + half;
        if (yt != cur) {
          flush();
          if (cur == 0 && !freed && !meter) {        // leaving key tile 0: its columns may be overwritten
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_free);
            freed = true;
          }
          cur = yt;
          mbar_wait(&bar_s1[yt], 0);
          tc_fence_after();
        }
        tmem_ld32(tmem + lane_off + yt * 256 + ch * 32, dst);
      };
      // one unit: `now` holds its chunk (load in flight), `nxt` receives the next unit's.  The next load is issued
      // before the exponentials when it stays in the same key tile (a tile switch flushes the running sums, so it
      // has to wait for this unit; it happens at most once per warp).
      auto step = [&](int u, uint32_t (&now)[32], uint32_t (&nxt)[32]) {
        const int yt = u / nch_h, ch = 2 * (u - yt * nch_h) + half;
        tmem_wait_ld();
        const bool same_next = (u + 1 < u_end) && ((u + 1) / nch_h == yt);
        if (same_next) issue(u + 1, nxt);
        const int nvalid = min(32, P - ch * 32);
        if (meter && yt * TILE + row < P) {          // running maximum of the raw similarities (valid pairs only)
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (j < nvalid) vmax = fmaxf(vmax, __uint_as_float(now[j]));
        }
        if (yt < nyt) {                              // tiles past the foreground keys only feed the meter
          if (nvalid == 32) {
            softmax_chunk_fwd<PF_SMALL>(now, scale2, acc_m, acc_m2, acc_p, acc_p2);
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (j < nvalid) tail += ex2_approx(__uint_as_float(now[j]) * scale_log2);
          }
        }
        if (!same_next && u + 1 < u_end) issue(u + 1, nxt);
      };
      if (!meter && u_begin >= u_end) {              // no unit at all in this warp's share (nch = 1, half 1)
        if (lane == 0) mbar_arrive(bar_free);
        freed = true;
      }
      if (u_begin < u_end) issue(u_begin, r0);
      for (int u = u_begin; u < u_end; u += 2) {
        step(u, r0, r1);
        if (u + 1 < u_end) step(u + 1, r1, r0);
      }
      flush();
      if (!freed && !meter) {                        // ended inside key tile 0, or had no unit of it
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_free);
      }
    }
    part[(half * 2 + 0) * 128 + row] = sum0;
    part[(half * 2 + 1) * 128 + row] = sum1;
    if (meter) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) vmax = fmaxf(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
      if (lane == 0) smax[warp - 2] = vmax;
    }
    asm volatile("bar.sync 1, 256;" ::: "memory");
    if (meter) {
      // second look at this warp's units (S^T is still in TMEM): every pair within SEG_DELTA of the maximum
      float M = smax[0];
#pragma unroll
      for (int w = 1; w < 8; ++w) M = fmaxf(M, smax[w]);
      const float thr = M - SEG_DELTA;
      for (int u = u_begin; u < u_end; ++u) {
        const int yt = u / nch_h, ch = 2 * (u - yt * nch_h) + half;
        uint32_t rc[32];
        tmem_ld32(tmem + lane_off + yt * 256 + ch * 32, rc);
        tmem_wait_ld();
        const int nvalid = min(32, P - ch * 32);
        if (yt * TILE + row < P) {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (j < nvalid && __uint_as_float(rc[j]) >= thr) {
              const int slot = atomicAdd(ncand, 1);
              if (slot < SEG_MAXC) cand[slot] = ((ch * 32 + j) << 16) | (yt * TILE + row);
            }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) { mbar_arrive(bar_free); mbar_arrive(bar_cand); }
    }
    if (half < nyt) {                                // half h finishes key tile h
      const int y = half * 128 + row;
      const float lg = log2f(part[(0 * 2 + half) * 128 + row] + part[(1 * 2 + half) * 128 + row]);
      bool on = y < cnt;
      if (on && mask_b) on = mask_b[(long long)b * P + y] != 0.f;
      nl[y] = on ? -lg : NL_MASKED;
      if (xt == 0 && y < P) lse[(long long)b * P + y] = (y < cnt) ? lg * LN2 : 0.f;
    }
    if (nyt == 1 && half == 1 && xt == 0) {          // rows of the skipped (all-background) key tile
      const int y = 128 + row;
      if (y < P) lse[(long long)b * P + y] = 0.f;
    }
    asm volatile("bar.sync 1, 256;" ::: "memory");
    // ---- phase 2: S[xt] -> P~ (bf16 pairs, in place); half h owns key columns [h*W, h*W + W)
    mbar_wait(bar_s2, 0);
    tc_fence_after();
    {
      const int mine = min(cnt - half * W, W);       // key columns of this half that carry gradient (may be <= 0)
      const int nchh = mine > 0 ? (mine + 31) / 32 : 0;
      const uint32_t tS = tmem + lane_off + half * W;
      const uint32_t nl_saddr = smem_u32(nl + half * W);
      uint32_t rb0[32], rb1[32];
      auto step2 = [&](int ch, uint32_t (&now)[32], uint32_t (&nxt)[32]) {
        uint32_t pk[16];
        tmem_wait_ld();
        if (ch + 1 < nchh) tmem_ld32(tS + (ch + 1) * 32, nxt);
        softmax_chunk_bwd<PF_SMALL>(now, nl_saddr + ch * 128, scale2, pk);
        tmem_st16(tS + ch * 16, pk);
      };
      if (nchh > 0) tmem_ld32(tS, rb0);
      for (int ch = 0; ch < nchh; ch += 2) {
        step2(ch, rb0, rb1);
        if (ch + 1 < nchh) step2(ch + 1, rb1, rb0);
      }
      tmem_wait_st();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bar_p[half]);
    }
    // ---- epilogue: acc[x, c] for the channel half of this warp, coalesced along P
    mbar_wait(bar_d, 0);
    tc_fence_after();
    {
      const int x = xt * TILE + row;
      float *dst = acc + (long long)b * C * P + x;
#pragma unroll 1
      for (int ch = 0; ch < C / 64; ++ch) {
        const int c0 = half * (C / 2) + ch * 32;
        uint32_t r[32];
        tmem_ld32(tmem_D + lane_off + c0, r);
        tmem_wait_ld();
        if (x < P) {
#pragma unroll
          for (int j = 0; j < 32; ++j) dst[(long long)(c0 + j) * P] = __uint_as_float(r[j]);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

// Backward of the small path: dq[b,c,x] = F.normalize backward of  g * (alpha_b acc[b,c,x] + a_x (v1-v2)[c] + v2[c]).
// CTA = PX pixels x G groups of 32 channels (512 threads); the per-pixel dot product goes through shared memory.
template <int C>
__global__ void __launch_bounds__(512)
dense_small_bwd_kernel(const float *__restrict__ acc, const __nv_bfloat16 *__restrict__ qhat,
                       const int *__restrict__ key_count, const float *__restrict__ mask_a,
                       const float *__restrict__ inv_norm, const float *__restrict__ alpha,
                       const float *__restrict__ v1, const float *__restrict__ v2,
                       const float *__restrict__ gscale, float *__restrict__ dq, int P) {
  constexpr int CG = 32, G = C / CG, PX = 512 / G;
  __shared__ float sva[C], svb[C];
  __shared__ float red[G][PX];
  const int b = blockIdx.y, px = threadIdx.x % PX, grp = threadIdx.x / PX;
  const int x = blockIdx.x * PX + px;
  const bool valid = x < P;
  // every global load of the thread is issued before the first barrier: one memory round trip
  const float *arow = acc + ((long long)b * C + grp * CG) * P + (valid ? x : 0);
  const uint4 *qrow = reinterpret_cast<const uint4 *>(qhat + ((long long)b * P + (valid ? x : 0)) * C + grp * CG);
  float av[CG];
  uint4 qv[CG / 8];
#pragma unroll
  for (int cl = 0; cl < CG; ++cl) av[cl] = __ldg(arow + (long long)cl * P);
#pragma unroll
  for (int u = 0; u < CG / 8; ++u) qv[u] = __ldg(qrow + u);
  const float gup = gscale ? __ldg(gscale) : 1.f;
  const int cnt = key_count ? min(max(__ldg(key_count + b), 0), P) : P;
  const float al_raw = __ldg(alpha + b);
  const float a = valid ? __ldg(mask_a + (long long)b * P + x) : 0.f;
  const float inr = valid ? __ldg(inv_norm + (long long)b * P + x) : 0.f;
  for (int c = threadIdx.x; c < C; c += 512) {
    const float b2 = v2 ? __ldg(v2 + (long long)b * C + c) * gup : 0.f;
    sva[c] = __ldg(v1 + (long long)b * C + c) * gup - b2;
    svb[c] = b2;
  }
  const float al = cnt > 0 ? al_raw * gup : 0.f;
  __syncthreads();
  float dot = 0.f;
#pragma unroll
  for (int u = 0; u < CG / 8; ++u) {
    const uint32_t qw[4] = {qv[u].x, qv[u].y, qv[u].z, qv[u].w};
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int cl = u * 8 + e, c = grp * CG + cl;
      const float qh = (e & 1) ? __uint_as_float(qw[e >> 1] & 0xffff0000u) : __uint_as_float(qw[e >> 1] << 16);
      const float g = fmaf(al, valid ? av[cl] : 0.f, fmaf(a, sva[c], svb[c]));
      av[cl] = g;                                   // acc value is dead: keep the combined gradient in its place
      dot = fmaf(qh, g, dot);
    }
  }
  red[grp][px] = dot;
  __syncthreads();
  float dsum = 0.f;
#pragma unroll
  for (int k = 0; k < G; ++k) dsum += red[k][px];
  const float ndot = -dsum;
  if (!valid) return;
  float *drow = dq + ((long long)b * C + grp * CG) * P + x;
#pragma unroll
  for (int u = 0; u < CG / 8; ++u) {
    const uint32_t qw[4] = {qv[u].x, qv[u].y, qv[u].z, qv[u].w};
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int cl = u * 8 + e;
      const float qh = (e & 1) ? __uint_as_float(qw[e >> 1] & 0xffff0000u) : __uint_as_float(qw[e >> 1] << 16);
      drow[(long long)cl * P] = fmaf(qh, ndot, av[cl]) * inr;
    }
  }
}

}  // namespace

// 1 when cp2_dense_small / cp2_dense_small_bwd take this shape (else use cp2_dense_lse + cp2_dense_grad)
extern "C" int cp2_dense_small_ok(int C, int P) {
  return (P > 0 && ((C == 128 && P <= 256) || (C == 256 && P <= 128))) ? 1 : 0;
}

extern "C" int cp2_dense_small(const void *qhat_t, const void *khat_t, const int *key_count, const float *mask_b,
                               float *lse, float *acc, const float *seg_q, const float *seg_k,
                               const float *seg_inv_q, const float *seg_inv_k, const int *seg_perm_k,
                               const float *seg_mask_a, const float *seg_mask_b, float *seg_hit, float inv_tau, int B,
                               int C, int P, void *stream) {
  CP2_REQUIRE(cp2_dense_small_ok(C, P), "cp2_dense_small: unsupported shape C=%d P=%d (C=128: P<=256, C=256: P<=128)", C, P);
  CP2_REQUIRE(B >= 0 && B <= 65535, "cp2_dense_small: bad batch B=%d", B);
  CP2_REQUIRE(inv_tau > 0.f && inv_tau <= 16.f, "cp2_dense_small: 1/tau=%g outside (0, 16]", (double)inv_tau);
  if (B == 0) return CP2_OK;
  CP2_REQUIRE(qhat_t && khat_t && lse && acc, "cp2_dense_small: null pointer");
  CP2_REQUIRE(seg_hit == nullptr || (seg_q && seg_k && seg_inv_q && seg_inv_k && seg_mask_a &&
                                     (seg_mask_b || key_count)),
              "cp2_dense_small: the acc_seg meter needs seg_q, seg_k, seg_inv_q, seg_inv_k, seg_mask_a and "
              "seg_mask_b (or key_count)");
  const int nxt = cp2_cdiv(P, TILE), rows = nxt * TILE;
  CUtensorMap tq, tk;
  int rc;
  if ((rc = make_feature_tmap(&tq, qhat_t, B, C, P, rows)) != CP2_OK) return rc;
  if ((rc = make_feature_tmap(&tk, khat_t, B, C, P, rows)) != CP2_OK) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const float sl2 = inv_tau * LOG2E;
  const SegMeter seg{seg_q, seg_k, seg_inv_q, seg_inv_k, seg_perm_k, seg_mask_a, seg_mask_b, seg_hit};
  dim3 grid(nxt, B);
  if (C == 128) {
    const size_t smem = SmallCfg<128>::smem(rows);
    CP2_SMEM_ONCE(dense_small_kernel<128>, SmallCfg<128>::smem(256));
    dense_small_kernel<128><<<grid, SM_THREADS, smem, st>>>(tq, tk, key_count, mask_b, lse, acc, seg, sl2, P, rows);
  } else {
    const size_t smem = SmallCfg<256>::smem(rows);
    CP2_SMEM_ONCE(dense_small_kernel<256>, SmallCfg<256>::smem(128));
    dense_small_kernel<256><<<grid, SM_THREADS, smem, st>>>(tq, tk, key_count, mask_b, lse, acc, seg, sl2, P, rows);
  }
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}

extern "C" int cp2_dense_small_bwd(const float *acc, const void *qhat_t, const int *key_count, const float *mask_a,
                                   const float *inv_norm, const float *alpha, const float *v1, const float *v2,
                                   const float *gscale, float *dq, int B, int C, int P, void *stream) {
  CP2_REQUIRE(C == 128 || C == 256, "cp2_dense_small_bwd: C=%d unsupported (128 or 256)", C);
  CP2_REQUIRE(B >= 0 && P > 0 && B <= 65535, "cp2_dense_small_bwd: bad shape B=%d P=%d", B, P);
  if (B == 0) return CP2_OK;
  CP2_REQUIRE(acc && qhat_t && mask_a && inv_norm && alpha && v1 && dq, "cp2_dense_small_bwd: null pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  dim3 grid(cp2_cdiv(P, C == 128 ? 128 : 64), B);
  if (C == 128)
    dense_small_bwd_kernel<128><<<grid, 512, 0, st>>>(acc, static_cast<const __nv_bfloat16 *>(qhat_t), key_count,
                                                      mask_a, inv_norm, alpha, v1, v2, gscale, dq, P);
  else
    dense_small_bwd_kernel<256><<<grid, 512, 0, st>>>(acc, static_cast<const __nv_bfloat16 *>(qhat_t), key_count,
                                                      mask_a, inv_norm, alpha, v1, v2, gscale, dq, P);
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}
