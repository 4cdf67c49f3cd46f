// K3 — global MoCo InfoNCE against the negative queue, forward + backward in ONE pass
// over the queue (reference builder.py:1396, 1420-1428 and autograd; removes the
// queue.clone()).  HBM-bound: the 4*C*K-byte fp32 queue is read exactly once, by TMA.
//
// The queue is [C,K] row-major, so the logits GEMM contracts over the STRIDED axis and
// the gradient GEMM over the CONTIGUOUS one.  tcgen05 has no shared-memory layout that
// serves a 32-bit operand both ways (MN-major tf32 needs the 32-byte-atom swizzle, which
// K-major forbids), so four converter warps turn each fp32 tile into a bf16 [k][c] image
// with the ordinary 128-byte swizzle; that one image is the K-major A operand of GEMM1
// and the MN-major A operand of GEMM2 (the pattern of the dense backward kernel).
//
// Per tile of up to 128 queue entries (4 TMA boxes of 32 entries x 128 channels, 16 KB each):
//   convert  f32 [c][k] -> bf16 image [k][c]                                  (warps 6-9)
//   GEMM1    D1[k, b]  = sum_c img[k,c] * (q_hi[b,c] + q_lo[b,c])              (warp 1)
//   epilogue e = exp((D1 - 1)/tau); z[b] += e (fp32); E[b,k] = bf16(e) -> smem (warps 2-5)
//   GEMM2    D2[c, b] += sum_k img[k,c] * E[b,k]                               (warp 1)
// q_pos is split into two bf16 terms (hi + lo) so only the queue's own bf16 rounding
// (|dl| ~ 1e-4 rms) enters the logits.  Persistent CTAs own balanced contiguous ranges of
// 32-entry boxes; per-CTA partials are reduced by a second tiny kernel in a fixed order
// (deterministic, no atomics).
#include "common.cuh"
#include "umma.cuh"

using namespace umma;

namespace {

constexpr int KT = 128;                 // queue entries per tile (MMA M of GEMM1, K of GEMM2)
constexpr int CDIM = 128;               // feature dim
constexpr int FBOX = CDIM * 128;        // one fp32 TMA box: 128 c-rows x 32 k (128 B) = 16 KB
constexpr int FSTAGE = 4 * FBOX;        // 64 KB
constexpr int IBOX = KT * 128;          // one bf16 image box: 128 k-rows x 64 c (128 B) = 16 KB
constexpr int IMG = 2 * IBOX;           // 32 KB
constexpr int NST = 2;                  // fp32 stages
constexpr int NIMG = 2;                 // bf16 images
constexpr int NTHREADS = 320;
constexpr int ROWS = 32;                // batch rows per CTA chunk
constexpr float LOG2E = 1.4426950408889634f;

__device__ __forceinline__ uint8_t *align1024(uint8_t *p) {
  return reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(p) + 1023) & ~uintptr_t(1023));
}

// byte offset of element (row, col) of a [rows][128 cols] bf16 operand stored as two
// 128-byte-swizzled boxes of 64 columns (box_bytes each)
__device__ __forceinline__ uint32_t sw128_off(int row, int col, int box_bytes) {
  return (uint32_t)((col >> 6) * box_bytes + row * 128 + ((((col & 63) >> 3) ^ (row & 7)) << 4) + (col & 7) * 2);
}

template <int NB>
__global__ void __launch_bounds__(NTHREADS)
global_nce_kernel(const __grid_constant__ CUtensorMap tm_queue, const float *__restrict__ q_pos,
                  float *__restrict__ part_z, float *__restrict__ part_g, float scale_log2, int B, int K) {
  constexpr int QBOX = NB * 128;                                  // one q / E box: NB rows x 128 B
  constexpr uint32_t TMEM_COLS = (3 * NB <= 32) ? 32 : (3 * NB <= 64) ? 64 : 128;
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = align1024(smem_raw);
  uint8_t *sF = smem;                             // NST fp32 stages
  uint8_t *sI = sF + NST * FSTAGE;                // NIMG bf16 images
  uint8_t *sQh = sI + NIMG * IMG;                 // q_pos hi, 2 boxes
  uint8_t *sQl = sQh + 2 * QBOX;                  // q_pos lo, 2 boxes
  uint8_t *sE = sQl + 2 * QBOX;                   // E, 2 boxes
  float *zs = reinterpret_cast<float *>(sE + 2 * QBOX);           // [4][NB]
  uint64_t *bars = reinterpret_cast<uint64_t *>(zs + 4 * NB);
  uint64_t *ffull = bars, *fempty = ffull + NST, *ifull = fempty + NST, *iempty = ifull + NIMG,
           *d1_full = iempty + NIMG, *d1_empty = d1_full + 2, *e_full = d1_empty + 2, *e_empty = e_full + 1,
           *d2_full = e_empty + 1;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(d2_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int chunk = blockIdx.y, b0 = chunk * ROWS;
  const int cta = blockIdx.x, nctas = gridDim.x;
  // balanced contiguous range of 32-entry boxes
  const int nbox_all = (K + 31) >> 5;
  const int r0 = (int)((long long)cta * nbox_all / nctas);
  const int r1 = (int)((long long)(cta + 1) * nbox_all / nctas);
  const int my_tiles = (r1 - r0 + 3) >> 2;

  if (threadIdx.x == 0) {
    for (int s = 0; s < NST; ++s) { mbar_init(&ffull[s], 1); mbar_init(&fempty[s], 4); }
    for (int i = 0; i < NIMG; ++i) { mbar_init(&ifull[i], 4); mbar_init(&iempty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&d1_full[i], 1); mbar_init(&d1_empty[i], 4); }
    mbar_init(e_full, 4);
    mbar_init(e_empty, 1);
    mbar_init(d2_full, 1);
    fence_barrier_init();
    // the first NST stages are requested right here, by the thread that initialised the barriers: the DRAM
    // round trip then overlaps the rest of the prologue (q_pos conversion, TMEM allocation, CTA barrier)
    for (int i = 0; i < min(NST, my_tiles); ++i) {
      const int nb = min(4, r1 - r0 - 4 * i);
      const int k0 = (r0 + 4 * i) * 32;
      mbar_expect_tx(&ffull[i], nb * FBOX);
      for (int kq = 0; kq < nb; ++kq) tma_load_2d(sF + i * FSTAGE + kq * FBOX, &tm_queue, &ffull[i], k0 + kq * 32, 0);
    }
  }
  if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
  // q_pos -> bf16 hi / lo operands (K-major, rows past B are zero)
  for (int idx = threadIdx.x; idx < NB * (CDIM / 2); idx += NTHREADS) {
    const int row = idx / (CDIM / 2), c = (idx % (CDIM / 2)) * 2;
    float2 v = make_float2(0.f, 0.f);
    if (b0 + row < B) v = *reinterpret_cast<const float2 *>(q_pos + (long long)(b0 + row) * CDIM + c);
    const __nv_bfloat16 h0 = __float2bfloat16_rn(v.x), h1 = __float2bfloat16_rn(v.y);
    const float l0 = v.x - __bfloat162float(h0), l1 = v.y - __bfloat162float(h1);
    const uint32_t off = sw128_off(row, c, QBOX);
    *reinterpret_cast<uint32_t *>(sQh + off) = pack_bf16x2(__bfloat162float(h0), __bfloat162float(h1));
    *reinterpret_cast<uint32_t *>(sQl + off) = pack_bf16x2(l0, l1);
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tmem_D2 = tmem + 2 * NB;

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (lane == 0) {
      for (int i = NST; i < my_tiles; ++i) {           // tiles 0 .. NST-1 were requested in the prologue
        const int s = i % NST, ph = (i / NST) & 1;
        const int nb = min(4, r1 - r0 - 4 * i);
        const int k0 = (r0 + 4 * i) * 32;
        mbar_wait(&fempty[s], ph ^ 1);
        mbar_expect_tx(&ffull[s], nb * FBOX);
        for (int kq = 0; kq < nb; ++kq)
          tma_load_2d(sF + s * FSTAGE + kq * FBOX, &tm_queue, &ffull[s], k0 + kq * 32, 0);
      }
    }
  } else if (warp == 1) {
    // -------------------------------------------------------------- MMA issuer
    if (lane == 0 && my_tiles > 0) {
      constexpr uint32_t idesc1 = make_idesc(KT, NB, FMT_BF16, FMT_BF16, MAJOR_K, MAJOR_K);
      constexpr uint32_t idesc2 = make_idesc(CDIM, NB, FMT_BF16, FMT_BF16, MAJOR_MN, MAJOR_K);
      const uint32_t aQh = smem_u32(sQh), aQl = smem_u32(sQl), aE = smem_u32(sE);
      auto gemm1 = [&](int i) {
        const int ib = i % NIMG, iph = (i / NIMG) & 1, buf = i & 1, tph = (i >> 1) & 1;
        const uint32_t aI = smem_u32(sI + ib * IMG);
        mbar_wait(&ifull[ib], iph);
        mbar_wait(&d1_empty[buf], tph ^ 1);
        tc_fence_after();
#pragma unroll
        for (int kk = 0; kk < CDIM / 16; ++kk)      // K = c, 16 per slice (32 B inside a 128-B row)
          mma_f16_ss(tmem + buf * NB, make_sdesc(aI + (kk >> 2) * IBOX + (kk & 3) * 32, 16, 1024),
                     make_sdesc(aQh + (kk >> 2) * QBOX + (kk & 3) * 32, 16, 1024), idesc1, kk > 0);
#pragma unroll
        for (int kk = 0; kk < CDIM / 16; ++kk)
          mma_f16_ss(tmem + buf * NB, make_sdesc(aI + (kk >> 2) * IBOX + (kk & 3) * 32, 16, 1024),
                     make_sdesc(aQl + (kk >> 2) * QBOX + (kk & 3) * 32, 16, 1024), idesc1, 1u);
        mma_commit(&d1_full[buf]);
      };
      gemm1(0);
      for (int i = 0; i < my_tiles; ++i) {
        if (i + 1 < my_tiles) gemm1(i + 1);
        const int ib = i % NIMG;
        const uint32_t aI = smem_u32(sI + ib * IMG);
        mbar_wait(e_full, i & 1);
        tc_fence_after();
        // K = queue entry, 16 image rows (2 KB) per slice; A read MN-major: 64-channel chunks IBOX apart
#pragma unroll
        for (int kk = 0; kk < KT / 16; ++kk)
          mma_f16_ss(tmem_D2, make_sdesc(aI + kk * 2048, IBOX, 1024),
                     make_sdesc(aE + (kk >> 2) * QBOX + (kk & 3) * 32, 16, 1024), idesc2,
                     (i > 0 || kk > 0) ? 1u : 0u);
        mma_commit(&iempty[ib]);
        mma_commit(e_empty);
      }
      mma_commit(d2_full);
    }
  } else if (warp >= 6) {
    // ------------------------------------------ converters: fp32 [c][k] -> bf16 image [k][c]
    const int w = warp - 6;                         // this warp owns box w: image rows 32w .. 32w+31
    const int krow = w * 32 + lane;
    for (int i = 0; i < my_tiles; ++i) {
      const int s = i % NST, ph = (i / NST) & 1, ib = i % NIMG, iph = (i / NIMG) & 1;
      const int nb = min(4, r1 - r0 - 4 * i);
      mbar_wait(&ffull[s], ph);
      mbar_wait(&iempty[ib], iph ^ 1);
      const float *src = reinterpret_cast<const float *>(sF + s * FSTAGE + w * FBOX) + lane;
      uint8_t *dst = sI + ib * IMG + krow * 128;
      if (w < nb) {
#pragma unroll 4
        for (int j = 0; j < CDIM / 8; ++j) {        // 8 channels -> one 16-byte chunk of image row krow
          float v[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) v[u] = src[(j * 8 + u) * 32];
          uint4 o;
          o.x = pack_bf16x2(v[0], v[1]); o.y = pack_bf16x2(v[2], v[3]);
          o.z = pack_bf16x2(v[4], v[5]); o.w = pack_bf16x2(v[6], v[7]);
          *reinterpret_cast<uint4 *>(dst + (j >> 3) * IBOX + (((j & 7) ^ (krow & 7)) << 4)) = o;
        }
      } else {                                      // box beyond this CTA's range: zero rows
        const uint4 zero = make_uint4(0, 0, 0, 0);
#pragma unroll 4
        for (int j = 0; j < CDIM / 8; ++j)
          *reinterpret_cast<uint4 *>(dst + (j >> 3) * IBOX + (((j & 7) ^ (krow & 7)) << 4)) = zero;
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) { mbar_arrive(&ifull[ib]); mbar_arrive(&fempty[s]); }
    }
  } else {
    // ---------------------------------------------------- epilogue warps 2-5 (TMEM quadrant warp&3)
    const int quad = warp & 3;
    const int row = quad * 32 + lane;               // image row = queue entry within the tile
    const uint32_t lane_off = uint32_t(quad * 32) << 16;
    float zacc[NB];
#pragma unroll
    for (int j = 0; j < NB; ++j) zacc[j] = 0.f;
    for (int i = 0; i < my_tiles; ++i) {
      const int buf = i & 1, tph = (i >> 1) & 1;
      const int nb = min(4, r1 - r0 - 4 * i);
      const int kg = (r0 + 4 * i) * 32 + row;
      const bool kvalid = (row < nb * 32) && (kg < K);
      mbar_wait(&d1_full[buf], tph);
      tc_fence_after();
      mbar_wait(e_empty, (i & 1) ^ 1);          // GEMM2 of the previous tile has consumed sE
#pragma unroll
      for (int ch = 0; ch < NB / 16; ++ch) {
        uint32_t r[16];
        tmem_ld16(tmem + lane_off + buf * NB + ch * 16, r);
        tmem_wait_ld();
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const int bb = ch * 16 + j;
          float e = ex2_approx(fmaf(__uint_as_float(r[j]), scale_log2, -scale_log2));
          e = kvalid ? e : 0.f;
          zacc[bb] += e;
          *reinterpret_cast<__nv_bfloat16 *>(sE + sw128_off(bb, row, QBOX)) = __float2bfloat16_rn(e);
        }
      }
      fence_proxy_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) { mbar_arrive(e_full); mbar_arrive(&d1_empty[buf]); }
    }
    // ---- partial results of this CTA
#pragma unroll
    for (int j = 0; j < NB; ++j) {
      float v = zacc[j];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0) zs[quad * NB + j] = v;
    }
    asm volatile("bar.sync 1, 128;" ::: "memory");
    const int st = threadIdx.x - 64;
    const long long pbase = (long long)(chunk * nctas + cta) * NB;
    if (st < NB) part_z[pbase + st] = zs[st] + zs[NB + st] + zs[2 * NB + st] + zs[3 * NB + st];
    if (my_tiles > 0) {
      mbar_wait(d2_full, 0);
      tc_fence_after();
#pragma unroll
      for (int ch = 0; ch < NB / 16; ++ch) {
        uint32_t r[16];
        tmem_ld16(tmem_D2 + lane_off + ch * 16, r);
        tmem_wait_ld();
#pragma unroll
        for (int j = 0; j < 16; ++j) part_g[(pbase + ch * 16 + j) * CDIM + row] = __uint_as_float(r[j]);
      }
    } else {
#pragma unroll
      for (int j = 0; j < NB; ++j) part_g[(pbase + j) * CDIM + row] = 0.f;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, TMEM_COLS);
}

// Fixed-order reduction of the per-CTA partials: 8 slices of the CTA range are summed in parallel
// (thread = channel x slice), then combined in slice order -> deterministic, ~19 dependent loads deep.
constexpr int FIN_SLICES = 8;
__global__ void __launch_bounds__(CDIM * FIN_SLICES)
global_finalize_kernel(const float *__restrict__ part_z, const float *__restrict__ part_g,
                       float *__restrict__ z, float *__restrict__ g, int NB, int nctas) {
  __shared__ float sg[FIN_SLICES][CDIM];
  __shared__ float sz[FIN_SLICES];
  const int b = blockIdx.x, chunk = b / ROWS, bl = b % ROWS;
  const int c = threadIdx.x % CDIM, sl = threadIdx.x / CDIM;
  float s = 0.f, szv = 0.f;
#pragma unroll 4
  for (int cta = sl; cta < nctas; cta += FIN_SLICES) {
    const long long pb = (long long)(chunk * nctas + cta) * NB + bl;
    s += part_g[pb * CDIM + c];
    if (c == 0) szv += part_z[pb];
  }
  sg[sl][c] = s;
  if (c == 0) sz[sl] = szv;
  __syncthreads();
  if (sl == 0) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < FIN_SLICES; ++i) t += sg[i][c];
    g[(long long)b * CDIM + c] = t;
    if (c == 0) {
      float tz = 0.f;
#pragma unroll
      for (int i = 0; i < FIN_SLICES; ++i) tz += sz[i];
      z[b] = tz;
    }
  }
}

int pick_nb(int B) { return B <= 16 ? 16 : 32; }

int pick_nctas(int K) {
  const int nbox = cp2_cdiv(K, 32);
  const int tiles = cp2_cdiv(nbox, 4);
  return tiles < cp2_num_sms() ? tiles : cp2_num_sms();
}

int make_queue_tmap(CUtensorMap *tm, const void *base, uint64_t K) {
  EncodeTiledFn enc = get_encode_tiled();
  if (!enc) { cp2_set_error("cuTensorMapEncodeTiled unavailable in this driver"); return CP2_E_UNSUPPORTED; }
  const cuuint64_t dims[2] = {K, (cuuint64_t)CDIM};
  const cuuint64_t strides[1] = {K * 4};
  const cuuint32_t box[2] = {32, CDIM};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void *>(base), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                         CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { cp2_set_error("cuTensorMapEncodeTiled(queue) failed: CUresult %d", (int)r); return CP2_E_CUDA; }
  return CP2_OK;
}

template <int NB>
int launch_global(const CUtensorMap &tq, const float *q_pos, float *part_z, float *part_g, float scale_log2,
                  int B, int K, int nctas, int chunks, cudaStream_t st) {
  const size_t smem = 1024 + (size_t)NST * FSTAGE + (size_t)NIMG * IMG + 6 * (size_t)NB * 128 +
                      sizeof(float) * 4 * NB + 256;
  CP2_CUDA(cudaFuncSetAttribute(global_nce_kernel<NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  global_nce_kernel<NB><<<dim3(nctas, chunks), NTHREADS, smem, st>>>(tq, q_pos, part_z, part_g, scale_log2, B, K);
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}

}  // namespace

extern "C" int64_t cp2_global_scratch_floats(int B, int C) {
  if (B <= 0) return 0;
  const int chunks = cp2_cdiv(B, ROWS);
  return (int64_t)chunks * cp2_num_sms() * pick_nb(B) * (1 + C);
}

extern "C" int cp2_global_nce(const float *q_pos, const float *queue, float *z, float *g, float *scratch,
                              float inv_tau, int B, int C, int K, void *stream) {
  CP2_REQUIRE(C == CDIM, "cp2_global_nce: C=%d unsupported (128)", C);
  CP2_REQUIRE(B >= 0 && K > 0 && K % 4 == 0, "cp2_global_nce: bad shape B=%d K=%d (K %% 4 == 0)", B, K);
  if (B == 0) return CP2_OK;
  CP2_REQUIRE(q_pos && queue && z && g && scratch, "cp2_global_nce: null pointer");
  CP2_REQUIRE((reinterpret_cast<uintptr_t>(queue) & 15u) == 0 && (reinterpret_cast<uintptr_t>(q_pos) & 7u) == 0,
              "cp2_global_nce: queue must be 16-byte and q_pos 8-byte aligned");
  const int NB = pick_nb(B), chunks = cp2_cdiv(B, ROWS);
  const int nctas = pick_nctas(K);
  CUtensorMap tq;
  int rc;
  if ((rc = make_queue_tmap(&tq, queue, (uint64_t)K)) != CP2_OK) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float *part_z = scratch;
  float *part_g = scratch + (size_t)chunks * nctas * NB;
  const float scale_log2 = inv_tau * LOG2E;
  if (NB == 16) rc = launch_global<16>(tq, q_pos, part_z, part_g, scale_log2, B, K, nctas, chunks, st);
  else rc = launch_global<32>(tq, q_pos, part_z, part_g, scale_log2, B, K, nctas, chunks, st);
  if (rc != CP2_OK) return rc;
  global_finalize_kernel<<<B, CDIM * FIN_SLICES, 0, st>>>(part_z, part_g, z, g, NB, nctas);
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}
