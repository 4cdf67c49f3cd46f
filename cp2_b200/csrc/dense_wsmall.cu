// f3 on the tensor cores — the dense loss of forward_cp2 in the PROPOSED modes (pair weights from the id maps,
// reference builder.py:1204-1243; NegativeType squash, builder.py:1332-1386; loss builder.py:1431-1437; acc_seg
// builder.py:1442-1448) for feature maps that fit the tensor memory of one SM (P <= 256 at C = 128, P <= 128 at
// C = 256: the 14x14 map the modes run at).  Same skeleton as dense_small.cu — CTA = (image, 128 query pixels),
// both bf16 operands of the image in shared memory, all 512 TMEM columns — with the pair weight and the squash
// evaluated by the softmax warps on the accumulator:
//   phase 1  S^T[yt] (M = y, N = x):  z = w[x,y] * t(S) / tau, per-lane online (max, sum) over the query axis ->
//            lse_y; label sum  pos_y = sum_x a_x z[x,y]; arg-max of z; per-image loss
//            dense_n = sum_y b_y (A lse_y - pos_y) / (A Bm)  and the label of the arg-max pair (hit).
//   phase 2  S[xt] (M = x, N = y) -> g[x,y] = b_y (exp(z - lse_y) - a_x/A) w t'(S) as bf16 pairs in place ->
//            acc[x,:] = sum_y g[x,y] khat[y,:]  (tcgen05.mma, A from TMEM, B = the key image MN-major).
//   epilogue acc -> global [B,C,P]; the backward is cp2_dense_small_bwd (alpha * acc + rank-1 terms + normalize).
// The ids of the image's P pixels (four int64 maps, read at the [OS//2::OS] lattice) and the masks sit in shared
// memory; the statistics pre-pass (wstats, MEDIAN select) is weighted.cu's.
#include "common.cuh"
#include "umma.cuh"
#include "dense_common.cuh"
#include "weighted_common.cuh"

using namespace umma;
using namespace dense;
using namespace weighted;

namespace {

constexpr int WS_THREADS = 320;          // warp 0 idle (loads issued by thread 0), warp 1 MMA, warps 2-9 softmax
constexpr int MAXP = 256;

struct WSmem {                           // after the two operand images
  float a_s[MAXP], b_s[MAXP], nl_s[MAXP];
  long long pa_s[MAXP], ra_s[MAXP], pb_s[MAXP], rb_s[MAXP];
  float red_m[2][MAXP], red_l[2][MAXP], red_pos[2][MAXP], red_bv[2][MAXP];
  int red_bi[2][MAXP];
  float cont[MAXP], fbv[MAXP];
  int fbf[MAXP];
  uint64_t bars[16];
  uint32_t tmem_slot[4];
};

template <int C>
size_t ws_smem(int rows) { return 1024 + 2 * (size_t)(C / 64) * rows * 128 + sizeof(WSmem) + 16; }

template <int C>
__global__ void __launch_bounds__(WS_THREADS, 1)
dense_wsmall_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                    const float *__restrict__ mask_a, const float *__restrict__ mask_b,
                    const long long *__restrict__ pid_a, const long long *__restrict__ pid_b,
                    const long long *__restrict__ rid_a, const long long *__restrict__ rid_b,
                    const float *__restrict__ wstats, float *__restrict__ lse, float *__restrict__ dense_n,
                    float *__restrict__ hit, float *__restrict__ acc, const WParams p, int rows) {
  const int P = p.P;
  const int box = rows * 128;
  const int oper = (C / 64) * box;
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = align1024(smem_raw);
  uint8_t *sK = smem;
  uint8_t *sQ = smem + oper;
  WSmem &sm = *reinterpret_cast<WSmem *>(sQ + oper);
  uint64_t *bar_k = sm.bars, *bar_q = sm.bars + 1, *bar_s1 = sm.bars + 2 /* [2] */, *bar_free = sm.bars + 4,
           *bar_s2 = sm.bars + 5, *bar_p = sm.bars + 6 /* [2] */, *bar_d = sm.bars + 8;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.y, xt = blockIdx.x;
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
    fence_barrier_init();
    mbar_expect_tx(bar_k, oper);
    for (int kb = 0; kb < C / 64; ++kb) tma_load_3d(sK + kb * box, &tm_k, bar_k, kb * 64, 0, b);
    mbar_expect_tx(bar_q, oper);
    for (int kb = 0; kb < C / 64; ++kb) tma_load_3d(sQ + kb * box, &tm_q, bar_q, kb * 64, 0, b);
  }
  // masks and ids of the image's pixels (zero past P)
  for (int i = threadIdx.x; i < MAXP; i += WS_THREADS) {
    const bool v = i < P;
    const long long gi = v ? id_index(p, b, i) : 0;
    sm.a_s[i] = v ? mask_a[(long long)b * P + i] : 0.f;
    sm.b_s[i] = v ? mask_b[(long long)b * P + i] : 0.f;
    sm.pa_s[i] = v ? pid_a[gi] : 0;
    sm.ra_s[i] = v ? rid_a[gi] : 0;
    sm.pb_s[i] = v ? pid_b[gi] : 0;
    sm.rb_s[i] = v ? rid_b[gi] : 0;
  }
  const float centre = p.neg_type >= 2 ? wstats[3 * b] : 0.f, A = wstats[3 * b + 1], Bm = wstats[3 * b + 2];
  if (warp == 1) tmem_alloc(sm.tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *sm.tmem_slot;
  const uint32_t tmem_D = tmem + 256;
  const int nyt = rows / TILE;                      // key tiles (all keys take part: columns of b_y = 0 feed the arg-max)
  const int nch = (P + 31) / 32;
  const int W = nyt * 64;
  const float sl2 = p.inv_tau * LOG2E;

  if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    const uint32_t idesc1 = make_idesc(TILE, rows, FMT_BF16, FMT_BF16, MAJOR_K, MAJOR_K);
    constexpr uint32_t idesc_d = make_idesc(TILE, C, FMT_BF16, FMT_BF16, MAJOR_K, MAJOR_MN);
    const uint64_t dK0 = make_sdesc(smem_u32(sK), 16, 1024);
    const uint64_t dQ0 = make_sdesc(smem_u32(sQ), 16, 1024);
    const uint64_t dKmn0 = make_sdesc(smem_u32(sK), box, 1024);
    mbar_wait(bar_k, 0);
    mbar_wait(bar_q, 0);
    tc_fence_after();
    if (elect_one()) {
      for (int yt = 0; yt < nyt; ++yt) {
#pragma unroll
        for (int kk = 0; kk < C / 16; ++kk) {
          const uint32_t off = (kk >> 2) * box + (kk & 3) * 32;
          mma_f16_ss(tmem + yt * 256, sdesc_advance(dK0, off + yt * TILE * 128), sdesc_advance(dQ0, off), idesc1, kk > 0);
        }
        mma_commit(&bar_s1[yt]);
      }
    }
    __syncwarp();
    mbar_wait(bar_free, 0);                          // phase 1 is over: S^T's columns may be overwritten
    tc_fence_after();
    if (elect_one()) {
#pragma unroll
      for (int kk = 0; kk < C / 16; ++kk) {
        const uint32_t off = (kk >> 2) * box + (kk & 3) * 32;
        mma_f16_ss(tmem, sdesc_advance(dQ0, off + xt * TILE * 128), sdesc_advance(dK0, off), idesc1, kk > 0);
      }
      mma_commit(bar_s2);
    }
    __syncwarp();
    for (int h = 0; h < 2; ++h) {
      mbar_wait(&bar_p[h], 0);
      tc_fence_after();
      if (elect_one()) {
        for (int kk = 0; kk < W / 16; ++kk) {
          const int key0 = h * W + kk * 16;
          if (key0 >= P) break;
          mma_f16_ts(tmem_D, tmem + h * W + kk * 8, sdesc_advance(dKmn0, (key0 / 16) * 2048), idesc_d,
                     (h > 0 || kk > 0) ? 1u : 0u);
        }
      }
      __syncwarp();
    }
    if (elect_one()) mma_commit(bar_d);
    __syncwarp();
  } else if (warp >= 2) {
    // ------------------------------------------------------------------ softmax warps
    const int quad = warp & 3, half = (warp - 2) >> 2;
    const uint32_t lane_off = uint32_t(quad * 32) << 16;
    const int row = quad * 32 + lane;
    // ---- phase 1: lanes = keys y; units u = yt * nch + ch shared between the two halves
    const int U = nyt * nch, U0 = (U + 1) / 2;
    const int u_begin = half == 0 ? 0 : U0, u_end = half == 0 ? U0 : U;
#pragma unroll
    for (int yt = 0; yt < 2; ++yt) {                 // neutral partial results for tiles this half never visits
      sm.red_m[half][yt * TILE + row] = -INFINITY;
      sm.red_l[half][yt * TILE + row] = 0.f;
      sm.red_pos[half][yt * TILE + row] = 0.f;
      sm.red_bv[half][yt * TILE + row] = -INFINITY;
      sm.red_bi[half][yt * TILE + row] = 0;
    }
    {
      float m = -INFINITY, l = 0.f, pos = 0.f, bv = -INFINITY;
      int bi = 0, cur = -1;
      long long pb = 0, rb = 0;
      float by = 0.f;
      auto flush = [&]() {
        if (cur >= 0) {
          const int y = cur * TILE + row;
          sm.red_m[half][y] = m;
          sm.red_l[half][y] = l;
          sm.red_pos[half][y] = pos;
          sm.red_bv[half][y] = bv;
          sm.red_bi[half][y] = bi;
        }
        m = -INFINITY; l = 0.f; pos = 0.f; bv = -INFINITY; bi = 0;
      };
      for (int u = u_begin; u < u_end; ++u) {
        const int yt = u / nch, ch = u - yt * nch;
        if (yt != cur) {
          flush();
          cur = yt;
          const int y = yt * TILE + row;
          pb = sm.pb_s[y]; rb = sm.rb_s[y]; by = sm.b_s[y];
          mbar_wait(&bar_s1[yt], 0);
          tc_fence_after();
        }
        uint32_t r[32];
        tmem_ld32(tmem + lane_off + yt * 256 + ch * 32, r);
        tmem_wait_ld();
        const int nvalid = min(32, P - ch * 32);
        float zc[32], cm = -INFINITY;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          zc[j] = -INFINITY;
          if (j < nvalid) {
            const int x = ch * 32 + j;
            const float ax = sm.a_s[x];
            float z2, wd;
            pair_eval(__uint_as_float(r[j]), ax != 0.f && by != 0.f, sm.pa_s[x], pb, sm.ra_s[x], rb, centre, p, sl2, z2, wd);
            pos = fmaf(ax, z2, pos);
            if (z2 > bv) { bv = z2; bi = x; }
            zc[j] = z2;
            cm = fmaxf(cm, z2);
          }
        }
        const float mn = fmaxf(m, cm);               // nvalid >= 1: finite
        float add = 0.f;
#pragma unroll
        for (int j = 0; j < 32; ++j) add += ex2_approx(zc[j] - mn);
        l = fmaf(l, ex2_approx(m - mn), add);
        m = mn;
      }
      flush();
    }
    asm volatile("bar.sync 1, 256;" ::: "memory");
    if (half < nyt) {                                // half h combines and finishes key tile h
      const int y = half * TILE + row;
      const float m0 = sm.red_m[0][y], m1 = sm.red_m[1][y];
      const float M = fmaxf(m0, m1);
      const float L = sm.red_l[0][y] * ex2_approx(m0 - M) + sm.red_l[1][y] * ex2_approx(m1 - M);
      const float lse2 = M + log2f(L);
      const float posn = (sm.red_pos[0][y] + sm.red_pos[1][y]) * LN2;
      const float by = sm.b_s[y];
      const bool vy = y < P;
      sm.nl_s[y] = -lse2;
      if (xt == 0 && vy) lse[(long long)b * P + y] = lse2 * LN2;
      sm.cont[y] = (vy && by != 0.f) ? by * (A * (lse2 * LN2) - posn) : 0.f;
      float v = sm.red_bv[0][y];
      int xi = sm.red_bi[0][y];
      if (sm.red_bv[1][y] > v) { v = sm.red_bv[1][y]; xi = sm.red_bi[1][y]; }   // half 1 holds the larger x: ties stay
      sm.fbv[y] = vy ? v : -INFINITY;
      sm.fbf[y] = xi * P + y;
    } else {
      const int y = half * TILE + row;               // tile absent (rows = 128): neutral entries
      sm.cont[y] = 0.f;
      sm.fbv[y] = -INFINITY;
      sm.fbf[y] = 0x7fffffff;
    }
    tc_fence_before();
    asm volatile("bar.sync 1, 256;" ::: "memory");
    if (lane == 0) mbar_arrive(bar_free);
    if (xt == 0 && warp == 2) {                      // per-image loss and arg-max label, fixed order
      float s = 0.f, best = -INFINITY;
      int bf = 0x7fffffff;
#pragma unroll
      for (int i = 0; i < MAXP / 32; ++i) {
        const int y = lane + 32 * i;
        s += sm.cont[y];
        const float v = sm.fbv[y];
        const int f = sm.fbf[y];
        if (v > best || (v == best && f < bf)) { best = v; bf = f; }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        s += __shfl_xor_sync(0xffffffffu, s, o);
        const float v = __shfl_xor_sync(0xffffffffu, best, o);
        const int f = __shfl_xor_sync(0xffffffffu, bf, o);
        if (v > best || (v == best && f < bf)) { best = v; bf = f; }
      }
      if (lane == 0) {
        dense_n[b] = s / (A * Bm);                   // 0/0 -> NaN as the reference
        float h = 0.f;
        if (bf != 0x7fffffff) { const int x = bf / P, y = bf - x * P; h = sm.a_s[x] * sm.b_s[y]; }
        if (hit) hit[b] = h;
      }
    }
    // ---- phase 2: lanes = queries x; half h owns key columns [h*W, h*W + W)
    mbar_wait(bar_s2, 0);
    tc_fence_after();
    {
      const int x = xt * TILE + row;
      const bool vx = x < P;
      const float ax = vx ? sm.a_s[x] : 0.f;
      const long long pa = vx ? sm.pa_s[x] : 0, ra = vx ? sm.ra_s[x] : 0;
      const float aoa = ax / A;
      const int mine = min(P - half * W, W);
      const int nchh = mine > 0 ? (mine + 31) / 32 : 0;
      const uint32_t tS = tmem + lane_off + half * W;
      for (int ch = 0; ch < nchh; ++ch) {
        uint32_t r[32], pk[16];
        tmem_ld32(tS + ch * 32, r);
        tmem_wait_ld();
#pragma unroll
        for (int j = 0; j < 32; j += 2) {
          float g2[2];
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            const int y = half * W + ch * 32 + j + e;
            const float by = sm.b_s[y];              // 0 past P
            float g = 0.f;
            if (by != 0.f && vx) {
              float z2, wd;
              pair_eval(__uint_as_float(r[j + e]), ax != 0.f, pa, sm.pb_s[y], ra, sm.rb_s[y], centre, p, sl2, z2, wd);
              g = by * (ex2_approx(z2 + sm.nl_s[y]) - aoa) * wd;
            }
            g2[e] = g;
          }
          pk[j >> 1] = pack_bf16x2(g2[0], g2[1]);
        }
        tmem_st16(tS + ch * 16, pk);
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

}  // namespace

extern "C" int cp2_dense_weighted_small(const void *qhat_t, const void *khat_t, const float *mask_a,
                                        const float *mask_b, const int64_t *pixel_ids_a, const int64_t *pixel_ids_b,
                                        const int64_t *region_ids_a, const int64_t *region_ids_b, const float *pos_q,
                                        const float *neg_q, const float *pos_k, const float *neg_k, float *lse,
                                        float *dense_n, float *hit, float *wstats, int32_t *key_count, float *scratch,
                                        float *acc, float w_pix, float w_reg, float w_not, int neg_type,
                                        float neg_scale, float inv_tau, int B, int C, int P, int ids_h, int ids_w,
                                        int os, void *stream) {
  CP2_REQUIRE(cp2_dense_small_ok(C, P), "cp2_dense_weighted_small: unsupported shape C=%d P=%d (C=128: P<=256, C=256: P<=128)", C, P);
  CP2_REQUIRE(B > 0 && B <= 65535, "cp2_dense_weighted_small: bad batch B=%d", B);
  CP2_REQUIRE(qhat_t && khat_t && mask_a && mask_b && pixel_ids_a && pixel_ids_b && region_ids_a && region_ids_b &&
              pos_q && neg_q && pos_k && neg_k && lse && dense_n && wstats && key_count && scratch && acc,
              "cp2_dense_weighted_small: null pointer");
  WParams p;
  if (int rc = make_params(p, w_pix, w_reg, w_not, neg_type, neg_scale, inv_tau, P, ids_h, ids_w, os)) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const auto *pa = reinterpret_cast<const long long *>(pixel_ids_a), *pb = reinterpret_cast<const long long *>(pixel_ids_b);
  const auto *ra = reinterpret_cast<const long long *>(region_ids_a), *rb = reinterpret_cast<const long long *>(region_ids_b);
  const auto *q = static_cast<const __nv_bfloat16 *>(qhat_t), *k = static_cast<const __nv_bfloat16 *>(khat_t);
  if (int rc = weighted::prepass(q, k, mask_a, mask_b, pos_q, neg_q, pos_k, neg_k, wstats, key_count, scratch, neg_type, B, C, P, st))
    return rc;
  const int nxt = cp2_cdiv(P, TILE), rows = nxt * TILE;
  CUtensorMap tq, tk;
  int rc;
  if ((rc = make_feature_tmap(&tq, qhat_t, B, C, P, rows)) != CP2_OK) return rc;
  if ((rc = make_feature_tmap(&tk, khat_t, B, C, P, rows)) != CP2_OK) return rc;
  dim3 grid(nxt, B);
  if (C == 128) {
    CP2_SMEM_ONCE(dense_wsmall_kernel<128>, ws_smem<128>(256));
    dense_wsmall_kernel<128><<<grid, WS_THREADS, ws_smem<128>(rows), st>>>(tq, tk, mask_a, mask_b, pa, pb, ra, rb, wstats,
                                                                            lse, dense_n, hit, acc, p, rows);
  } else {
    CP2_SMEM_ONCE(dense_wsmall_kernel<256>, ws_smem<256>(128));
    dense_wsmall_kernel<256><<<grid, WS_THREADS, ws_smem<256>(rows), st>>>(tq, tk, mask_a, mask_b, pa, pb, ra, rb, wstats,
                                                                            lse, dense_n, hit, acc, p, rows);
  }
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}
