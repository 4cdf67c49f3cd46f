// f3 — pieces shared by the two implementations of the PROPOSED-mode dense loss (weighted.cu: fp32 SIMT, any P;
// dense_wsmall.cu: tcgen05, maps that fit TMEM): the pair weight of builder.py:1204-1243, the NegativeType squash
// of builder.py:1332-1375 and the statistics pre-pass.
#pragma once
#include "common.cuh"

namespace weighted {

struct WParams {
  float w_pix, w_reg, w_not, neg_scale, inv_tau;
  int neg_type;                         // 0 none, 1 fixed (centre 0), 2 average, 3 median (per-image centre)
  int P, ws, ids_w, os;
  long long ids_hw;
};

__device__ __forceinline__ long long id_index(const WParams &p, int b, int pix) {
  const int r = pix / p.ws, c = pix - r * p.ws;
  return (long long)b * p.ids_hw + (long long)(r * p.os + p.os / 2) * p.ids_w + c * p.os + p.os / 2;
}

// builder.py:1234-1243: region weight where the region ids agree and are both known (non-zero),
// overridden by the pixel weight where the pixel ids agree, the "not" weight where still zero.
__device__ __forceinline__ float pair_weight(long long pa, long long pb, long long ra, long long rb,
                                             const WParams &p) {
  float w = (ra == rb && ra != 0) ? p.w_reg : 0.f;
  if (pa == pb) w = p.w_pix;
  if (w == 0.f) w = p.w_not;
  return w;
}

// builder.py:1332-1375: t = s on label pairs, 2*sigmoid(scale*(s-centre))-1 elsewhere; dt = dt/ds.
__device__ __forceinline__ void squash(float s, bool label, float centre, const WParams &p, float &t, float &dt) {
  if (label || p.neg_type == 0) { t = s; dt = 1.f; return; }
  t = 2.f / (1.f + expf((s - centre) * -p.neg_scale)) - 1.f;
  dt = (1.f - t * t) * 0.5f * p.neg_scale;
}


// One pixel pair for the tensor-core implementations: base-2 logit z2 = w * t(s) * sl2 (sl2 = log2(e) / tau) and
// wd = w * dt/ds; the squash's exponential and division on the MUFU (ex2 / rcp approximations, rel. error ~2^-22).
__device__ __forceinline__ void pair_eval(float s, bool label, long long pa, long long pb, long long ra, long long rb,
                                          float centre, const WParams &p, float sl2, float &z2, float &wd) {
  float t = s, dt = 1.f;
  if (p.neg_type != 0 && !label) {
    float e;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"((s - centre) * (-p.neg_scale * 1.4426950408889634f)));
    t = __fdividef(2.f, 1.f + e) - 1.f;
    dt = (1.f - t * t) * 0.5f * p.neg_scale;
  }
  const float w = pair_weight(pa, pb, ra, rb, p);
  z2 = w * t * sl2;
  wd = w * dt;
}


inline int make_params(WParams &p, float w_pix, float w_reg, float w_not, int neg_type, float neg_scale, float inv_tau,
                int P, int ids_h, int ids_w, int os) {
  CP2_REQUIRE(os >= 1 && ids_h >= 1 && ids_w >= 1, "cp2_dense_weighted: bad id map %dx%d stride %d", ids_h, ids_w, os);
  const int hs = (ids_h - os / 2 + os - 1) / os, ws = (ids_w - os / 2 + os - 1) / os;
  CP2_REQUIRE(hs * ws == P, "cp2_dense_weighted: id maps %dx%d at stride %d give %dx%d pixels, P=%d", ids_h, ids_w,
              os, hs, ws, P);
  CP2_REQUIRE(neg_type >= 0 && neg_type <= 3,
              "cp2_dense_weighted: negative type %d (0 none, 1 fixed, 2 average, 3 median)", neg_type);
  CP2_REQUIRE(inv_tau > 0.f, "cp2_dense_weighted: bad temperature");
  p = WParams{w_pix, w_reg, w_not, neg_scale, inv_tau, neg_type, P, ws, ids_w, os, (long long)ids_h * ids_w};
  return CP2_OK;
}


// dense_n[b] / acc_seg from the per-(image, key tile) partial results of a forward kernel (weighted.cu)
int finalize(const float *partial, const float *best_v, const int *best_i, const float *wstats, const float *mask_a,
             const float *mask_b, float *dense_n, float *acc_seg, int B, int P, int tiles, cudaStream_t st);

int prepass(const __nv_bfloat16 *q, const __nv_bfloat16 *k, const float *mask_a, const float *mask_b,
            const float *pos_q, const float *neg_q, const float *pos_k, const float *neg_k, float *wstats,
            int32_t *key_count, float *scratch, int neg_type, int B, int C, int P, cudaStream_t st);

}  // namespace weighted
