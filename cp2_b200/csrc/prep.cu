// K2a — per-pixel L2 normalisation, masked pooling and bf16 transpose
// (reference builder.py:1261-1268 / 1279-1285).
// HBM-bound: reads feat [B,C,P] f32 once, writes fhat_t [B,P,C] bf16
// (algorithmic bytes 6*B*C*P).  A CTA owns 32 consecutive pixels of one image: the
// loads are coalesced along P (128 B per warp per channel), the tile is transposed
// through shared memory, and the stores are coalesced along C.  Pooled sums are
// reduced per tile into scratch and summed over tiles by the last CTA of the image in a
// fixed order (deterministic).
#include "common.cuh"

namespace {

constexpr int PT = 32;   // pixels per CTA

struct PrepSide {
  const float *feat, *mask;
  __nv_bfloat16 *fhat_t;
  float *inv_norm, *scratch, *pos_sum, *neg_sum, *pos_unit, *neg_unit, *norms;
  int32_t *key_count, *perm;
};
struct PrepSides { PrepSide s[2]; };        // blockIdx.z selects the side (query / key) of a fused launch

// Fixed-order reduction of the per-tile partial sums of image b -> pos_sum / neg_sum [B,C]; optionally also the
// outer F.normalize of builder.py:1264,1268 (x / max(||x||, 1e-12)) -> pos_unit / neg_unit [B,C] and the raw
// norms.  Run by the LAST tile CTA of the image to finish (done counter), 256 threads.
__device__ void prep_reduce_image(const PrepSide &sd, int b, int C, int tiles, float *sums /* [512] */,
                                  float (*red)[8]) {
  float sq_pos = 0.f, sq_neg = 0.f;
  for (int t = threadIdx.x; t < 2 * C; t += blockDim.x) {
    const float *p = sd.scratch + (long long)b * tiles * 2 * C + t;
    float s = 0.f;
#pragma unroll 8
    for (int tl = 0; tl < tiles; ++tl) s += __ldcg(p + (long long)tl * 2 * C);
    sums[t] = s;
    if (t < C) { sd.pos_sum[(long long)b * C + t] = s; sq_pos = fmaf(s, s, sq_pos); }
    else { sd.neg_sum[(long long)b * C + (t - C)] = s; sq_neg = fmaf(s, s, sq_neg); }
  }
  if (sd.pos_unit == nullptr) return;
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    sq_pos += __shfl_down_sync(0xffffffffu, sq_pos, off);
    sq_neg += __shfl_down_sync(0xffffffffu, sq_neg, off);
  }
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) { red[0][wid] = sq_pos; red[1][wid] = sq_neg; }
  __syncthreads();
  float np = 0.f, nn = 0.f;
#pragma unroll
  for (int w = 0; w < 8; ++w) { np += red[0][w]; nn += red[1][w]; }
  np = sqrtf(np);
  nn = sqrtf(nn);
  const float ip = 1.f / fmaxf(np, 1e-12f), in = 1.f / fmaxf(nn, 1e-12f);
  for (int t = threadIdx.x; t < 2 * C; t += blockDim.x) {
    if (t < C) sd.pos_unit[(long long)b * C + t] = sums[t] * ip;
    else sd.neg_unit[(long long)b * C + (t - C)] = sums[t] * in;
  }
  if (threadIdx.x == 0) { sd.norms[2 * b] = np; sd.norms[2 * b + 1] = nn; }
}

template <int C>
__global__ void __launch_bounds__(256)
prep_kernel(const PrepSides sides, unsigned int *__restrict__ counters, int P, int tiles) {
  __shared__ float tile[C][PT + 1];
  __shared__ float ssq[8][PT];
  __shared__ float den[PT];
  __shared__ float mk[PT];
  __shared__ int dest[PT];          // compact mode: destination row of each pixel of the tile
  __shared__ int cnt_s[2][8];
  __shared__ int is_last;
  const PrepSide &sd = sides.s[blockIdx.z];
  const float *__restrict__ feat = sd.feat;
  const float *__restrict__ mask = sd.mask;
  __nv_bfloat16 *__restrict__ fhat_t = sd.fhat_t;
  float *__restrict__ inv_norm = sd.inv_norm;
  float *__restrict__ scratch = sd.scratch;
  int32_t *__restrict__ key_count = sd.key_count;
  const int b = blockIdx.y, tl = blockIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int x = tl * PT + lane;
  const bool valid = x < P;
  const float *__restrict__ src = feat + (long long)b * C * P + x;
  if (key_count != nullptr) {
    // stable partition of the image's pixels: foreground rows first, background rows after
    int before = 0, total = 0;
    for (int i = threadIdx.x; i < P; i += 256) {
      const int on = mask[(long long)b * P + i] != 0.f;
      total += on;
      before += on && (i < tl * PT);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      before += __shfl_xor_sync(0xffffffffu, before, o);
      total += __shfl_xor_sync(0xffffffffu, total, o);
    }
    if (lane == 0) { cnt_s[0][warp] = before; cnt_s[1][warp] = total; }
    __syncthreads();
    if (warp == 0) {
      before = 0; total = 0;
#pragma unroll
      for (int w = 0; w < 8; ++w) { before += cnt_s[0][w]; total += cnt_s[1][w]; }
      const bool on = valid && mask[(long long)b * P + x] != 0.f;
      const unsigned fgm = __ballot_sync(0xffffffffu, on);
      const int fg_lt = __popc(fgm & ((1u << lane) - 1u));
      dest[lane] = on ? (before + fg_lt) : (total + (x - before - fg_lt));
      if (tl == 0 && lane == 0) key_count[b] = total;
    }
  }
  // all C/8 loads of a thread are issued before the first use: one DRAM round trip per CTA instead of four
  float sum = 0.f;
  float v[C / 8];
#pragma unroll
  for (int i = 0; i < C / 8; ++i) v[i] = valid ? __ldcs(src + (long long)(warp + 8 * i) * P) : 0.f;
#pragma unroll
  for (int i = 0; i < C / 8; ++i) {
    tile[warp + 8 * i][lane] = v[i];
    sum = fmaf(v[i], v[i], sum);
  }
  ssq[warp][lane] = sum;
  __syncthreads();
  if (warp == 0) {
    float tot = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) tot += ssq[w][lane];
    const float d = fmaxf(sqrtf(tot), 1e-12f);       // F.normalize eps
    den[lane] = d;
    mk[lane] = valid ? mask[(long long)b * P + x] : 0.f;
    if (valid) inv_norm[(long long)b * P + x] = 1.0f / d;
  }
  __syncthreads();
  {
    const float d = den[lane];
    for (int c = warp; c < C; c += 8) tile[c][lane] = tile[c][lane] / d;
  }
  __syncthreads();
  {
    // pooled sums of the tile: one thread per (channel, pos | neg) walks the 32 pixels in order (rows of `tile` are
    // 33 floats apart: conflict-free), no shuffles
    float *__restrict__ part = scratch + ((long long)(b * tiles + tl) * 2) * C;
    const int nv = min(PT, P - tl * PT);
    for (int o = threadIdx.x; o < 2 * C; o += 256) {
      const int c = o < C ? o : o - C;
      float s = 0.f;
      if (o < C) {
#pragma unroll 8
        for (int p = 0; p < nv; ++p) s = fmaf(tile[c][p], mk[p], s);
      } else {
#pragma unroll 8
        for (int p = 0; p < nv; ++p) s = fmaf(tile[c][p], 1.0f - mk[p], s);
      }
      part[o] = s;
    }
  }
  __threadfence();                       // this tile's partial sums are visible before the done counter moves
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int *cnt = counters + blockIdx.z * gridDim.y + b;
    const int last = atomicAdd(cnt, 1u) == (unsigned)tiles - 1u;
    if (last) *cnt = 0u;
    is_last = last;
  }
  // transposed store: warp w writes pixels w*4..w*4+3, 64 channels (128 B) per pass
#pragma unroll
  for (int pp = 0; pp < PT / 8; ++pp) {
    const int p = warp * (PT / 8) + pp;
    const int xx = tl * PT + p;
    if (xx < P) {
      const int row = key_count ? dest[p] : xx;
      if (sd.perm && lane == 0) sd.perm[(long long)b * P + row] = xx;
      __nv_bfloat162 *dst = reinterpret_cast<__nv_bfloat162 *>(fhat_t + ((long long)b * P + row) * C);
#pragma unroll
      for (int j = 0; j < C / 64; ++j) {
        const int c = j * 64 + lane * 2;
        dst[c >> 1] = __floats2bfloat162_rn(tile[c][p], tile[c + 1][p]);
      }
    }
  }
  __syncthreads();
  if (is_last) {                         // block-uniform: the image's last tile reduces the partial sums
    __threadfence();
    prep_reduce_image(sd, b, C, tiles, &tile[0][0], reinterpret_cast<float (*)[8]>(&ssq[0][0]));
  }
}

}  // namespace

extern "C" int cp2_prep_tiles(int P) { return cp2_cdiv(P, PT); }

namespace {

int check_side(const PrepSide &sd) {
  if (!(sd.feat && sd.mask && sd.fhat_t && sd.inv_norm && sd.pos_sum && sd.neg_sum && sd.scratch)) {
    cp2_set_error("cp2_prep_features: null pointer");
    return CP2_E_INVALID;
  }
  if (!((sd.pos_unit == nullptr) == (sd.neg_unit == nullptr) && (sd.pos_unit == nullptr) == (sd.norms == nullptr))) {
    cp2_set_error("cp2_prep_features: pos_unit, neg_unit and norms go together");
    return CP2_E_INVALID;
  }
  return CP2_OK;
}

int launch_prep(const PrepSides &sides, int nsides, unsigned int *counters, int B, int C, int P, cudaStream_t st) {
  const int tiles = cp2_prep_tiles(P);
  dim3 grid(tiles, B, nsides);
  switch (C) {
    case 64: prep_kernel<64><<<grid, 256, 0, st>>>(sides, counters, P, tiles); break;
    case 128: prep_kernel<128><<<grid, 256, 0, st>>>(sides, counters, P, tiles); break;
    case 192: prep_kernel<192><<<grid, 256, 0, st>>>(sides, counters, P, tiles); break;
    default: prep_kernel<256><<<grid, 256, 0, st>>>(sides, counters, P, tiles); break;
  }
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}

}  // namespace

extern "C" int cp2_prep_features(const float *feat, const float *mask, void *fhat_t, float *inv_norm,
                                 float *pos_sum, float *neg_sum, float *scratch, int32_t *key_count, int32_t *perm,
                                 float *pos_unit, float *neg_unit, float *norms, void *scratch_counters,
                                 int B, int C, int P, void *stream) {
  CP2_REQUIRE(B >= 0 && P > 0 && B <= 65535, "cp2_prep_features: bad shape B=%d P=%d", B, P);
  CP2_REQUIRE(C == 64 || C == 128 || C == 192 || C == 256, "cp2_prep_features: C=%d unsupported (64,128,192,256)", C);
  if (B == 0) return CP2_OK;
  CP2_REQUIRE(scratch_counters, "cp2_prep_features: null scratch_counters");
  PrepSides sides;
  CP2_REQUIRE(perm == nullptr || key_count != nullptr, "cp2_prep_features: perm needs key_count (compaction)");
  sides.s[0] = PrepSide{feat, mask, static_cast<__nv_bfloat16 *>(fhat_t), inv_norm, scratch, pos_sum, neg_sum,
                        pos_unit, neg_unit, norms, key_count, perm};
  sides.s[1] = sides.s[0];
  int rc = check_side(sides.s[0]);
  if (rc != CP2_OK) return rc;
  return launch_prep(sides, 1, static_cast<unsigned int *>(scratch_counters), B, C, P, static_cast<cudaStream_t>(stream));
}

extern "C" int cp2_prep_features2(const float *feat_q, const float *mask_q, void *qhat_t, float *inv_norm_q,
                                  float *pos_sum_q, float *neg_sum_q, float *scratch_q, float *pos_unit_q,
                                  float *neg_unit_q, float *norms_q, const float *feat_k, const float *mask_k,
                                  void *khat_t, float *inv_norm_k, float *pos_sum_k, float *neg_sum_k,
                                  float *scratch_k, int32_t *key_count, int32_t *perm_k, float *pos_unit_k,
                                  float *neg_unit_k, float *norms_k, void *scratch_counters, int B, int C, int P,
                                  void *stream) {
  CP2_REQUIRE(B >= 0 && P > 0 && B <= 65535, "cp2_prep_features2: bad shape B=%d P=%d", B, P);
  CP2_REQUIRE(C == 64 || C == 128 || C == 192 || C == 256, "cp2_prep_features2: C=%d unsupported (64,128,192,256)", C);
  if (B == 0) return CP2_OK;
  CP2_REQUIRE(scratch_counters, "cp2_prep_features2: null scratch_counters");
  PrepSides sides;
  sides.s[0] = PrepSide{feat_q, mask_q, static_cast<__nv_bfloat16 *>(qhat_t), inv_norm_q, scratch_q, pos_sum_q,
                        neg_sum_q, pos_unit_q, neg_unit_q, norms_q, nullptr, nullptr};
  sides.s[1] = PrepSide{feat_k, mask_k, static_cast<__nv_bfloat16 *>(khat_t), inv_norm_k, scratch_k, pos_sum_k,
                        neg_sum_k, pos_unit_k, neg_unit_k, norms_k, key_count, perm_k};
  int rc = check_side(sides.s[0]);
  if (rc == CP2_OK) rc = check_side(sides.s[1]);
  if (rc != CP2_OK) return rc;
  return launch_prep(sides, 2, static_cast<unsigned int *>(scratch_counters), B, C, P, static_cast<cudaStream_t>(stream));
}
