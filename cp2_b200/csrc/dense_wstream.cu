// f3 on the tensor cores for maps that do NOT fit TMEM (P > 256, e.g. 28x28): the PROPOSED-mode dense loss
// (pair weights from the id maps, reference builder.py:1204-1243; NegativeType squash, builder.py:1332-1386; loss
// builder.py:1431-1437; acc_seg builder.py:1442-1448) on the streaming skeleton of dense_umma.cu.
//
// Forward  (wstream_lse_kernel): CTA = (image n, 128 key pixels y); S^T tile = Khat_y . Qhat_x^T (M = y -> TMEM
//   lanes, N = x -> columns, accumulator double-buffered), query tiles streamed by TMA.  The 8 softmax warps apply
//   z = w[x,y] t(S) / tau on the accumulator and keep, per key column (= per lane, in registers across all query
//   tiles), the online (max, sum) of the soft-max over the QUERY axis, the label sum pos_y = sum_x a_x z and the
//   arg-max of z.  Output: lse [B,P], and per (image, key tile) the partial loss  sum_y b_y (A lse_y - pos_y)  and
//   the best (z, flat index) — what weighted_lse_kernel writes per 32-key tile; weighted::finalize turns them
//   into dense_n [B] and acc_seg.
// Backward: dense_umma.cu's dense_grad_kernel in MODE 2 (cp2_dense_weighted_stream_grad below).
// The ids the weights need are gathered once per call at the [OS//2::OS] lattice into [4,B,P] int64
// (gather kernel), so both kernels read contiguous per-pixel arrays.
#include "common.cuh"
#include "umma.cuh"
#include "dense_common.cuh"
#include "weighted_common.cuh"

using namespace umma;
using namespace dense;
using namespace weighted;

namespace {

constexpr int WL_THREADS = 320;          // warp 0 TMA, warp 1 MMA + TMEM owner, warps 2-9 softmax
constexpr int WL_NST = 1;              // one query stage: two CTAs fit one SM up to P = 896 (16 softmax warps hide the
                                       // latency of the CUDA-core epilogue better than a second stage does)

// ids[0..3][b][pix] = pid_a, pid_b, rid_a, rid_b at the lattice point of pixel pix, b_pad[b][pix] = mask_b; rows are
// P_pad long (whole key tiles of the backward), zero past P
__global__ void __launch_bounds__(256)
wstream_gather_kernel(const long long *__restrict__ pid_a, const long long *__restrict__ pid_b,
                      const long long *__restrict__ rid_a, const long long *__restrict__ rid_b,
                      const float *__restrict__ mask_b, long long *__restrict__ ids, float *__restrict__ b_pad,
                      WParams p, int B, int P_pad) {
  const int b = blockIdx.y, pix = blockIdx.x * 256 + threadIdx.x;
  if (pix >= P_pad) return;
  const bool v = pix < p.P;
  const long long gi = v ? id_index(p, b, pix) : 0, o = (long long)b * P_pad + pix, plane = (long long)B * P_pad;
  ids[o] = v ? pid_a[gi] : 0;
  ids[plane + o] = v ? pid_b[gi] : 0;
  ids[2 * plane + o] = v ? rid_a[gi] : 0;
  ids[3 * plane + o] = v ? rid_b[gi] : 0;
  b_pad[o] = v ? mask_b[(long long)b * p.P + pix] : 0.f;
}

size_t wl_smem(int P) {
  const size_t pp = (size_t)cp2_cdiv(P, 128) * 128;
  return 1024 + (size_t)(1 + WL_NST) * 2 * TILE * 128 + pp * 20 + 5 * 128 * 4 + 3 * 4 * 4 + 16 * 8 + 32;
}

__global__ void __launch_bounds__(WL_THREADS, 2)
wstream_lse_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                   const float *__restrict__ mask_a, const float *__restrict__ b_pad,
                   const long long *__restrict__ ids, const float *__restrict__ wstats, float *__restrict__ lse,
                   float *__restrict__ partial, float *__restrict__ best_v, int *__restrict__ best_i,
                   const WParams p, int B, int P_pad) {
  constexpr int C = 128, BOX = TILE * 128, OPER = (C / 64) * BOX, NST = WL_NST;
  const int P = p.P, PP = (P + 127) / 128 * 128;
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = align1024(smem_raw);
  uint8_t *sK = smem;
  uint8_t *sQ = smem + OPER;
  long long *pa_s = reinterpret_cast<long long *>(sQ + NST * OPER);      // [PP] query-side pixel ids
  long long *ra_s = pa_s + PP;                                            // [PP] query-side region ids
  float *a_s = reinterpret_cast<float *>(ra_s + PP);                      // [PP] query-side mask
  float *comb = a_s + PP;                                                 // [5][128] upper half's (m, l, pos, bv, bi)
  float *red = comb + 5 * 128;                                            // [3][4] per-warp (loss, best v, best i)
  uint64_t *bars = reinterpret_cast<uint64_t *>(red + 12);
  uint64_t *bar_k = bars, *full = bars + 1, *empty = full + NST, *tfull = empty + NST, *tempty = tfull + 2;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.y, y0 = blockIdx.x * TILE;
  const int nx = (P + TILE - 1) / TILE;
  const long long plane = (long long)B * P_pad, rowo = (long long)b * P_pad;

  if (threadIdx.x == 0) {
    mbar_init(bar_k, 1);
    for (int s = 0; s < NST; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], 8); }
    fence_barrier_init();
    prefetch_tmap(&tm_q);
    prefetch_tmap(&tm_k);
  }
  for (int i = threadIdx.x; i < PP; i += WL_THREADS) {          // query-side ids / mask of the whole image
    const bool v = i < P;
    pa_s[i] = v ? ids[rowo + i] : 0;
    ra_s[i] = v ? ids[2 * plane + rowo + i] : 0;
    a_s[i] = v ? mask_a[(long long)b * P + i] : 0.f;
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
    // ------------------------------------------------------------------ softmax warps: lane = key pixel y
    const int quad = warp & 3, half = (warp - 2) >> 2, c0 = half * 64;
    const int row = quad * 32 + lane, y = y0 + row;
    const bool vy = y < P;
    const uint32_t lane_base = tmem + (uint32_t(quad * 32) << 16) + c0;
    const long long pb = vy ? ids[plane + rowo + y] : 0, rb = vy ? ids[3 * plane + rowo + y] : 0;
    const float by = vy ? b_pad[rowo + y] : 0.f;
    const float centre = p.neg_type >= 2 ? wstats[3 * b] : 0.f, A = wstats[3 * b + 1];
    const float sl2 = p.inv_tau * LOG2E;
    float m = -INFINITY, l = 0.f, pos = 0.f, bv = -INFINITY;
    int bi = 0;
    for (int i = 0; i < nx; ++i) {
      const int buf = i & 1, tph = (i >> 1) & 1;
      mbar_wait(&tfull[buf], tph);
      tc_fence_after();
#pragma unroll 1
      for (int ch = 0; ch < 2; ++ch) {
        const int xb = i * TILE + c0 + ch * 32;            // first query pixel of this chunk
        if (xb >= P) break;                                // warp-uniform
        uint32_t r[32];
        tmem_ld32(lane_base + buf * TILE + ch * 32, r);
        tmem_wait_ld();
        const int nvalid = min(32, P - xb);
        float zc[32], cm = -INFINITY;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          zc[j] = -INFINITY;
          if (j < nvalid) {
            const int x = xb + j;
            const float ax = a_s[x];
            float z2, wd;
            pair_eval(__uint_as_float(r[j]), ax != 0.f && by != 0.f, pa_s[x], pb, ra_s[x], rb, centre, p, sl2, z2, wd);
            pos = fmaf(ax, z2, pos);
            if (z2 > bv) { bv = z2; bi = x; }
            zc[j] = z2;
            cm = fmaxf(cm, z2);
          }
        }
        const float mn = fmaxf(m, cm);
        float add = 0.f;
#pragma unroll
        for (int j = 0; j < 32; ++j) add += ex2_approx(zc[j] - mn);
        l = fmaf(l, ex2_approx(m - mn), add);
        m = mn;
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[buf]);
    }
    // ---- the two column halves of a key meet in shared memory; half 0 finishes the key
    if (half == 1) {
      comb[0 * 128 + row] = m;
      comb[1 * 128 + row] = l;
      comb[2 * 128 + row] = pos;
      comb[3 * 128 + row] = bv;
      comb[4 * 128 + row] = __int_as_float(bi);
    }
    asm volatile("bar.sync 1, 256;" ::: "memory");
    if (half == 0) {
      const float m1 = comb[0 * 128 + row], l1 = comb[1 * 128 + row];
      const float M = fmaxf(m, m1);
      const float L = l * ex2_approx(m - M) + l1 * ex2_approx(m1 - M);
      const float lse_nat = (M + log2f(L)) * LN2;
      const float posn = (pos + comb[2 * 128 + row]) * LN2;
      // half 0 saw the smaller x of every tile, but tiles interleave: compare flat indices on ties
      float v = bv;
      int xi = bi;
      const float v1 = comb[3 * 128 + row];
      const int x1 = __float_as_int(comb[4 * 128 + row]);
      if (v1 > v || (v1 == v && x1 < xi)) { v = v1; xi = x1; }
      if (vy) lse[(long long)b * P + y] = lse_nat;
      float contrib = (vy && by != 0.f) ? by * (A * lse_nat - posn) : 0.f;
      float fv = vy ? v : -INFINITY;
      int ff = vy ? xi * P + y : 0x7fffffff;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        contrib += __shfl_xor_sync(0xffffffffu, contrib, o);
        const float ov = __shfl_xor_sync(0xffffffffu, fv, o);
        const int of = __shfl_xor_sync(0xffffffffu, ff, o);
        if (ov > fv || (ov == fv && of < ff)) { fv = ov; ff = of; }
      }
      if (lane == 0) { red[quad] = contrib; red[4 + quad] = fv; red[8 + quad] = __int_as_float(ff); }
      asm volatile("bar.sync 2, 128;" ::: "memory");
      if (quad == 0 && lane == 0) {                        // warp 4 (quad 0 of half 0): fixed order over the quadrants
        float s = 0.f, gv = -INFINITY;
        int gf = 0x7fffffff;
        for (int w = 0; w < 4; ++w) {
          s += red[w];
          const float ov = red[4 + w];
          const int of = __float_as_int(red[8 + w]);
          if (ov > gv || (ov == gv && of < gf)) { gv = ov; gf = of; }
        }
        const long long o = (long long)b * gridDim.x + blockIdx.x;
        partial[o] = s;
        best_v[o] = gv;
        best_i[o] = gf;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 256);
}

}  // namespace

extern "C" int64_t cp2_dense_weighted_stream_ids_bytes(int B, int P) { return (int64_t)4 * B * cp2_dense_pad(P) * 8; }

extern "C" int cp2_dense_weighted_stream(const void *qhat_t, const void *khat_t, const float *mask_a,
                                         const float *mask_b, const int64_t *pixel_ids_a, const int64_t *pixel_ids_b,
                                         const int64_t *region_ids_a, const int64_t *region_ids_b, const float *pos_q,
                                         const float *neg_q, const float *pos_k, const float *neg_k, float *lse,
                                         float *dense_n, float *acc_seg, float *wstats, int32_t *key_count,
                                         float *scratch, void *ids_pad, float *b_pad, float w_pix, float w_reg,
                                         float w_not, int neg_type, float neg_scale, float inv_tau, int B, int C, int P,
                                         int ids_h, int ids_w, int os, void *stream) {
  CP2_REQUIRE(C == 128, "cp2_dense_weighted_stream: C=%d unsupported (128)", C);
  CP2_REQUIRE(B > 0 && B <= 65535 && P > 0 && P <= 8192, "cp2_dense_weighted_stream: bad shape B=%d P=%d", B, P);
  CP2_REQUIRE(qhat_t && khat_t && mask_a && mask_b && pixel_ids_a && pixel_ids_b && region_ids_a && region_ids_b &&
              pos_q && neg_q && pos_k && neg_k && lse && dense_n && wstats && key_count && scratch && ids_pad && b_pad,
              "cp2_dense_weighted_stream: null pointer");
  WParams p;
  if (int rc = make_params(p, w_pix, w_reg, w_not, neg_type, neg_scale, inv_tau, P, ids_h, ids_w, os)) return rc;
  CP2_REQUIRE(p.inv_tau <= 16.f, "cp2_dense_weighted_stream: 1/tau=%g outside (0, 16]", (double)inv_tau);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const auto *q = static_cast<const __nv_bfloat16 *>(qhat_t), *k = static_cast<const __nv_bfloat16 *>(khat_t);
  if (int rc = weighted::prepass(q, k, mask_a, mask_b, pos_q, neg_q, pos_k, neg_k, wstats, key_count, scratch, neg_type, B, C, P, st))
    return rc;
  const int P_pad = cp2_dense_pad(P), tiles = cp2_cdiv(P, TILE);
  auto *ids = static_cast<long long *>(ids_pad);
  wstream_gather_kernel<<<dim3(cp2_cdiv(P_pad, 256), B), 256, 0, st>>>(
      reinterpret_cast<const long long *>(pixel_ids_a), reinterpret_cast<const long long *>(pixel_ids_b),
      reinterpret_cast<const long long *>(region_ids_a), reinterpret_cast<const long long *>(region_ids_b), mask_b, ids,
      b_pad, p, B, P_pad);
  CP2_LAUNCH_CHECK();
  CUtensorMap tq, tk;
  int rc;
  if ((rc = make_feature_tmap(&tq, qhat_t, B, C, P, TILE)) != CP2_OK) return rc;
  if ((rc = make_feature_tmap(&tk, khat_t, B, C, P, TILE)) != CP2_OK) return rc;
  const size_t smem = wl_smem(P);
  CP2_REQUIRE(smem <= 227 * 1024, "cp2_dense_weighted_stream: P=%d needs %zu bytes of shared memory", P, smem);
  CP2_CUDA(cudaFuncSetAttribute(wstream_lse_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  float *partial = scratch, *bv = scratch + (size_t)B * tiles;
  int *bi = reinterpret_cast<int *>(scratch + 2 * (size_t)B * tiles);
  wstream_lse_kernel<<<dim3(tiles, B), WL_THREADS, smem, st>>>(tq, tk, mask_a, b_pad, ids, wstats, lse, partial, bv, bi, p,
                                                               B, P_pad);
  CP2_LAUNCH_CHECK();
  return weighted::finalize(partial, bv, bi, wstats, mask_a, mask_b, dense_n, acc_seg, B, P, tiles, st);
}
