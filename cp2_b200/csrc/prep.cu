// K2a — per-pixel L2 normalisation, masked pooling and bf16 transpose
// (reference builder.py:1261-1268 / 1279-1285).
// HBM-bound: reads feat [B,C,P] f32 once, writes fhat_t [B,P,C] bf16
// (algorithmic bytes 6*B*C*P).  A CTA owns 32 consecutive pixels of one image: the
// loads are coalesced along P (128 B per warp per channel), the tile is transposed
// through shared memory, and the stores are coalesced along C.  Pooled sums are
// reduced per tile into scratch and summed over tiles by the last CTA of the image in a
// fixed order (deterministic).
#include "common.cuh"
#include <mutex>

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

// Thread (warp w, lane l) owns pixel l of the tile and the channel pairs (2j, 2j+1), j = w + 8i: every global load
// is a coalesced 128-byte row piece, the normalisation is one multiply per register, and adjacent channels convert to
// one packed bf16 word.  Two smem images of the tile: fp32 [C][33] for the pooled sums (thread per channel walks the
// pixels, conflict-free) and packed bf16 [32][C/2+1] for the transposed store (conflict-free both ways).
template <int C>
__global__ void __launch_bounds__(256)
prep_kernel(const PrepSides sides, unsigned int *__restrict__ counters, int P, int tiles, int nt) {
  constexpr int NP = C / 16;         // channel pairs per thread
  constexpr int TW = C / 2 + 1;      // words per pixel row of the packed tile (odd: conflict-free)
  extern __shared__ __align__(16) unsigned char prep_dyn[];
  float (*tile)[PT + 1] = reinterpret_cast<float (*)[PT + 1]>(prep_dyn);
  uint32_t (*tile_t)[TW] = reinterpret_cast<uint32_t (*)[TW]>(prep_dyn + sizeof(float) * C * (PT + 1));
  __shared__ float ssq[8][PT];
  __shared__ __align__(16) float2 wts[PT];   // (a_x, 1 - a_x) of the tile's pixels
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
  const int b = blockIdx.y;
  const int tl0 = blockIdx.x * nt, tl1 = min(tiles, tl0 + nt);      // this CTA's run of 32-pixel tiles
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // row pointer of this thread's first channel pair; pixels beyond P re-read pixel P-1 (their weights are zero and
  // their rows are never stored), so no load is predicated
  const float *__restrict__ src = feat + (long long)b * C * P + (long long)(2 * warp) * P;
  const long long pair_stride = 16LL * P;             // channel pairs j and j+8
  // all C/8 loads of a thread are issued before the first use (one DRAM round trip per tile), and the loads of
  // tile t+1 are issued before the reductions of tile t: a CTA always has 32*C*4 bytes in flight
  float v[2 * NP];
  {
    const float *r = src + min(tl0 * PT + lane, P - 1);
#pragma unroll
    for (int i = 0; i < NP; ++i) {
      v[2 * i] = __ldcs(r);
      v[2 * i + 1] = __ldcs(r + P);
      r += pair_stride;
    }
  }
  int before = 0, total = 0;        // warp 0: foreground pixels before the current tile / in the image
  if (key_count != nullptr) {
    // stable partition of the image's pixels: foreground rows first, background rows after
    for (int i = threadIdx.x; i < P; i += 256) {
      const int on = mask[(long long)b * P + i] != 0.f;
      total += on;
      before += on && (i < tl0 * PT);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      before += __shfl_xor_sync(0xffffffffu, before, o);
      total += __shfl_xor_sync(0xffffffffu, total, o);
    }
    if (lane == 0) { cnt_s[0][warp] = before; cnt_s[1][warp] = total; }
    __syncthreads();
    before = 0; total = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) { before += cnt_s[0][w]; total += cnt_s[1][w]; }
    if (blockIdx.x == 0 && threadIdx.x == 0) key_count[b] = total;
  }
  uint32_t *__restrict__ out_rows = reinterpret_cast<uint32_t *>(fhat_t + (long long)b * P * C);
  int32_t *__restrict__ perm_rows = sd.perm ? sd.perm + (long long)b * P : nullptr;
  for (int tl = tl0; tl < tl1; ++tl) {
    const int x = tl * PT + lane;
    const bool valid = x < P;
    float f[2 * NP];
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < 2 * NP; ++i) {
      f[i] = v[i];
      sum = fmaf(f[i], f[i], sum);
    }
    ssq[warp][lane] = sum;
    if (tl + 1 < tl1) {
      const float *r = src + min(x + PT, P - 1);
#pragma unroll
      for (int i = 0; i < NP; ++i) {
        v[2 * i] = __ldcs(r);
        v[2 * i + 1] = __ldcs(r + P);
        r += pair_stride;
      }
    }
    __syncthreads();
    {
      // every thread sums the eight partial squares of its own pixel (same order everywhere: same value)
      float tot = 0.f;
#pragma unroll
      for (int w = 0; w < 8; ++w) tot += ssq[w][lane];
      const float rd = tot >= 1e-24f ? rsqrtf(tot) : 1e12f;    // 1 / max(||x||, 1e-12): F.normalize eps
      if (warp == 0) {
        const float m = valid ? mask[(long long)b * P + x] : 0.f;
        wts[lane] = valid ? make_float2(m, 1.0f - m) : make_float2(0.f, 0.f);
        if (valid) inv_norm[(long long)b * P + x] = rd;
        if (key_count != nullptr) {
          const bool on = valid && m != 0.f;
          const unsigned fgm = __ballot_sync(0xffffffffu, on);
          const int fg_lt = __popc(fgm & ((1u << lane) - 1u));
          dest[lane] = on ? (before + fg_lt) : (total + (x - before - fg_lt));
          before += __popc(fgm);
        }
      }
#pragma unroll
      for (int i = 0; i < NP; ++i) {
        const int j = warp + 8 * i;
        const float a = f[2 * i] * rd, c = f[2 * i + 1] * rd;
        tile[2 * j][lane] = a;
        tile[2 * j + 1][lane] = c;
        const __nv_bfloat162 h = __floats2bfloat162_rn(a, c);
        tile_t[lane][j] = *reinterpret_cast<const uint32_t *>(&h);
      }
    }
    __syncthreads();
    // pooled sums of the tile: one thread per channel walks the 32 pixels in order for both sums (pixels beyond P
    // carry zero weights), no shuffles
    for (int c = threadIdx.x; c < C; c += 256) {
      float sp0 = 0.f, sn0 = 0.f, sp1 = 0.f, sn1 = 0.f;      // even / odd pixels: two short FMA chains each
#pragma unroll
      for (int p = 0; p < PT; p += 2) {
        const float4 w = *reinterpret_cast<const float4 *>(&wts[p]);
        const float x0 = tile[c][p], x1 = tile[c][p + 1];
        sp0 = fmaf(x0, w.x, sp0);
        sn0 = fmaf(x0, w.y, sn0);
        sp1 = fmaf(x1, w.z, sp1);
        sn1 = fmaf(x1, w.w, sn1);
      }
      float *__restrict__ part = scratch + ((long long)(b * tiles + tl) * 2) * C;
      part[c] = sp0 + sp1;
      part[C + c] = sn0 + sn1;
    }
    // transposed store: warp w writes pixels w*4..w*4+3, 64 channels (128 B) per pass
#pragma unroll
    for (int pp = 0; pp < PT / 8; ++pp) {
      const int p = warp * (PT / 8) + pp;
      const int xx = tl * PT + p;
      if (xx < P) {
        const int row = key_count ? dest[p] : xx;
        if (perm_rows && lane == 0) perm_rows[row] = xx;
        uint32_t *dst = out_rows + (size_t)row * (C / 2) + lane;
#pragma unroll
        for (int j = 0; j < C / 64; ++j) dst[j * 32] = tile_t[p][j * 32 + lane];
      }
    }
    // no barrier here: the next iteration writes ssq before its first barrier (last read before this iteration's
    // second one) and the tiles / wts / dest after it (last read above)
  }
  __threadfence();                       // this CTA's partial sums are visible before the done counter moves
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int *cnt = counters + blockIdx.z * gridDim.y + b;
    const int last = atomicAdd(cnt, 1u) == gridDim.x - 1u;
    if (last) *cnt = 0u;
    is_last = last;
  }
  __syncthreads();
  if (is_last) {                         // block-uniform: the image's last CTA reduces the partial sums
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

// fp32 tile [C][33] + packed bf16 tile [32][C/2+1]
size_t prep_smem(int C) { return sizeof(float) * C * (PT + 1) + sizeof(uint32_t) * PT * (C / 2 + 1); }

int launch_prep(const PrepSides &sides, int nsides, unsigned int *counters, int B, int C, int P, cudaStream_t st) {
  const int tiles = cp2_prep_tiles(P);
  // tiles per CTA: 1 while every CTA is resident at once (the pre-training step: latency-bound, widest grid);
  // beyond that a CTA walks a run of tiles with the next tile's loads in flight (HBM-bound sizes)
  const long long ctas = (long long)B * nsides * tiles, resident = 8LL * cp2_num_sms();
  int nt = 1;
  if (ctas > resident) nt = (int)std::min<long long>(8, std::max<long long>(2, ctas / (2 * resident)));
  dim3 grid(cp2_cdiv(tiles, nt), B, nsides);
  const size_t smem = prep_smem(C);
  switch (C) {
    case 64: prep_kernel<64><<<grid, 256, smem, st>>>(sides, counters, P, tiles, nt); break;
    case 128: prep_kernel<128><<<grid, 256, smem, st>>>(sides, counters, P, tiles, nt); break;
    case 192: prep_kernel<192><<<grid, 256, smem, st>>>(sides, counters, P, tiles, nt); break;
    default:
      CP2_SMEM_ONCE(prep_kernel<256>, smem);       // 50.3 KB: above the 48 KB default
      prep_kernel<256><<<grid, 256, smem, st>>>(sides, counters, P, tiles, nt);
      break;
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

// ------------------------------------------------------------------------------------------------------------
// cp2_prep_features_static — cp2_prep_features for callers that cannot allocate (CUDA graph capture, serving
// loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The work space cp2_prep_features takes from the caller (per-tile partial sums + per-image done
// counters) lives in ONE file-scope device array instead; every GPU that loads the library holds its
// CP2_PREP_STATIC_BYTES until the process exits.  All calls on a device share it: calls that overlap on the GPU
// (different streams, no ordering between them) race on it.
namespace {

constexpr size_t kStaticCounterBytes = (size_t)CP2_PREP_STATIC_MAX_B * sizeof(unsigned int);   // 16 KB, first
constexpr size_t kStaticSumBytes = (size_t)CP2_PREP_STATIC_BYTES - kStaticCounterBytes;        // partial sums after
static_assert(kStaticCounterBytes % 256 == 0, "the partial sums start on a 256-byte boundary");

static __device__ __align__(256) unsigned char g_prep_static[CP2_PREP_STATIC_BYTES];

// In host code `g_prep_static` is a host-side stand-in, not a device address: the address the kernel may use
// comes from cudaGetSymbolAddress, asked once per device with that device current.
int prep_static_base(unsigned char **base) {
  static std::mutex mu;
  static unsigned char *cached[64] = {nullptr};
  int dev = 0;
  CP2_CUDA(cudaGetDevice(&dev));
  CP2_REQUIRE(dev >= 0 && dev < 64, "cp2_prep_features_static: device ordinal %d unsupported (< 64)", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (cached[dev] == nullptr) {
    unsigned char *p = nullptr;
    CP2_CUDA(cudaGetSymbolAddress(reinterpret_cast<void **>(&p), g_prep_static));
    CP2_REQUIRE(p != nullptr, "cp2_prep_features_static: cudaGetSymbolAddress returned a null address");
    cached[dev] = p;
  }
  *base = cached[dev];
  return CP2_OK;
}

}  // namespace

extern "C" int cp2_prep_features_static(const float *feat, const float *mask, void *fhat_t, float *inv_norm,
                                        float *pos_sum, float *neg_sum, int32_t *key_count, int32_t *perm,
                                        float *pos_unit, float *neg_unit, float *norms, int B, int C, int P,
                                        void *stream) {
  CP2_REQUIRE(B >= 0 && P > 0, "cp2_prep_features_static: bad shape B=%d P=%d", B, P);
  CP2_REQUIRE(C == 64 || C == 128 || C == 192 || C == 256,
              "cp2_prep_features_static: C=%d unsupported (64,128,192,256)", C);
  if (B == 0) return CP2_OK;
  CP2_REQUIRE(B <= CP2_PREP_STATIC_MAX_B, "cp2_prep_features_static: B=%d exceeds the static limit of %d images",
              B, (int)CP2_PREP_STATIC_MAX_B);
  const unsigned long long need = 1ull * B * cp2_prep_tiles(P) * 2ull * C * sizeof(float);
  CP2_REQUIRE(need <= kStaticSumBytes,
              "cp2_prep_features_static: B=%d C=%d P=%d needs %llu bytes of work space, the static array holds %llu "
              "(use cp2_prep_features with caller-owned scratch)",
              B, C, P, need, (unsigned long long)kStaticSumBytes);
  CP2_REQUIRE(perm == nullptr || key_count != nullptr, "cp2_prep_features_static: perm needs key_count (compaction)");
  unsigned char *base = nullptr;
  int rc = prep_static_base(&base);
  if (rc != CP2_OK) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  PrepSides sides;
  sides.s[0] = PrepSide{feat, mask, static_cast<__nv_bfloat16 *>(fhat_t), inv_norm,
                        reinterpret_cast<float *>(base + kStaticCounterBytes), pos_sum, neg_sum,
                        pos_unit, neg_unit, norms, key_count, perm};
  sides.s[1] = sides.s[0];
  rc = check_side(sides.s[0]);
  if (rc != CP2_OK) return rc;
  // the array keeps the last call's contents.  The partial sums are all written before they are read; the done
  // counters must start at zero (a completed call leaves them zero, an aborted one might not): clear them in-stream.
  CP2_CUDA(cudaMemsetAsync(base, 0, (size_t)B * sizeof(unsigned int), st));
  return launch_prep(sides, 1, reinterpret_cast<unsigned int *>(base), B, C, P, st);
}
