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


int prepass(const __nv_bfloat16 *q, const __nv_bfloat16 *k, const float *mask_a, const float *mask_b,
            const float *pos_q, const float *neg_q, const float *pos_k, const float *neg_k, float *wstats,
            int32_t *key_count, float *scratch, int neg_type, int B, int C, int P, cudaStream_t st);

}  // namespace weighted
