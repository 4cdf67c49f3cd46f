// K3 — global MoCo InfoNCE against the negative queue, forward + backward in ONE pass
// over the queue (reference builder.py:1396, 1420-1428 and autograd; removes the
// queue.clone()).  HBM-bound: the 4*C*K-byte fp32 queue is read exactly once, by TMA,
// and consumed in place by the tensor cores as TF32 — no conversion pass, no bf16 shadow.
//
// Per 128-entry queue tile (64 KB, four [128 c][32 k] swizzled boxes):
//   GEMM1  D1[k, b] = sum_c queue[c,k] * q_pos[b,c]       A = queue tile read MN-major
//   epilogue warps: e = exp((D1 - 1)/tau); z[b] += e; E[b,k] = e -> shared memory (K-major)
//   GEMM2  D2[c, b] += sum_k queue[c,k] * E[b,k]           A = the same tile read K-major
// Persistent CTAs (one per SM) stride over the tiles; per-CTA partials are reduced by a
// second tiny kernel in a fixed order (deterministic).
#include "common.cuh"
#include "umma.cuh"

using namespace umma;

namespace {

constexpr int KT = 128;                 // queue entries per tile
constexpr int CDIM = 128;               // feature dim (MMA M of GEMM2)
constexpr int QBOX = CDIM * 128;        // one queue box: 128 c-rows x 32 k (128 B) = 16 KB
constexpr int QSTAGE = 4 * QBOX;        // 64 KB
constexpr int NST = 2;
constexpr int NTHREADS = 192;
constexpr float LOG2E = 1.4426950408889634f;

__device__ __forceinline__ uint8_t *align1024(uint8_t *p) {
  return reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(p) + 1023) & ~uintptr_t(1023));
}

template <int NB>
__global__ void __launch_bounds__(NTHREADS)
global_nce_kernel(const __grid_constant__ CUtensorMap tm_queue, const __grid_constant__ CUtensorMap tm_qpos,
                  float *__restrict__ part_z, float *__restrict__ part_g, float scale_log2, int K, int ntiles) {
  constexpr int BBOX = NB * 128;                                  // one q_pos / E box: NB rows x 128 B
  constexpr uint32_t TMEM_COLS = (3 * NB <= 32) ? 32 : (3 * NB <= 64) ? 64 : (3 * NB <= 128) ? 128 : 256;
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = align1024(smem_raw);
  uint8_t *sQu = smem;                            // NST stages of the queue tile
  uint8_t *sB1 = sQu + NST * QSTAGE;              // q_pos operand, 4 boxes
  uint8_t *sE = sB1 + 4 * BBOX;                   // E operand, 4 boxes
  float *zs = reinterpret_cast<float *>(sE + 4 * BBOX);           // [4][NB]
  uint64_t *bars = reinterpret_cast<uint64_t *>(zs + 4 * NB);
  uint64_t *bar_b1 = bars, *qfull = bars + 1, *qempty = qfull + NST, *d1_full = qempty + NST,
           *d1_empty = d1_full + 2, *e_full = d1_empty + 2, *e_empty = e_full + 1, *d2_full = e_empty + 1;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(d2_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int chunk = blockIdx.y, b0 = chunk * 64;
  const int cta = blockIdx.x;
  const int my_tiles = (ntiles - cta + (int)gridDim.x - 1) / (int)gridDim.x;

  if (threadIdx.x == 0) {
    mbar_init(bar_b1, 1);
    for (int s = 0; s < NST; ++s) { mbar_init(&qfull[s], 1); mbar_init(&qempty[s], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&d1_full[i], 1); mbar_init(&d1_empty[i], 4); }
    mbar_init(e_full, 4);
    mbar_init(e_empty, 1);
    mbar_init(d2_full, 1);
    fence_barrier_init();
    prefetch_tmap(&tm_queue);
    prefetch_tmap(&tm_qpos);
  }
  if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tmem_D2 = tmem + 2 * NB;

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(bar_b1, 4 * BBOX);
      for (int kb = 0; kb < 4; ++kb) tma_load_2d(sB1 + kb * BBOX, &tm_qpos, bar_b1, kb * 32, b0);
      for (int i = 0; i < my_tiles; ++i) {
        const int s = i % NST, ph = (i / NST) & 1;
        const int k0 = (cta + i * (int)gridDim.x) * KT;
        mbar_wait(&qempty[s], ph ^ 1);
        mbar_expect_tx(&qfull[s], QSTAGE);
        for (int kq = 0; kq < 4; ++kq) tma_load_2d(sQu + s * QSTAGE + kq * QBOX, &tm_queue, &qfull[s], k0 + kq * 32, 0);
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && my_tiles > 0) {
      constexpr uint32_t idesc1 = make_idesc(KT, NB, FMT_TF32, FMT_TF32, MAJOR_MN, MAJOR_K);
      constexpr uint32_t idesc2 = make_idesc(CDIM, NB, FMT_TF32, FMT_TF32, MAJOR_K, MAJOR_K);
      const uint32_t aB1 = smem_u32(sB1), aE = smem_u32(sE);
      auto gemm1 = [&](int i) {
        const int s = i % NST, ph = (i / NST) & 1, buf = i & 1, tph = (i >> 1) & 1;
        const uint32_t aQ = smem_u32(sQu + s * QSTAGE);
        mbar_wait(&qfull[s], ph);
        mbar_wait(&d1_empty[buf], tph ^ 1);
        tc_fence_after();
        // K = c in slices of 8 rows (1 KB); A MN-major: 32-entry chunks 16 KB apart
#pragma unroll
        for (int kk = 0; kk < CDIM / 8; ++kk)
          mma_tf32_ss(tmem + buf * NB, make_sdesc(aQ + kk * 1024, QBOX, 1024),
                      make_sdesc(aB1 + (kk >> 2) * BBOX + (kk & 3) * 32, 16, 1024), idesc1, kk > 0);
        mma_commit(&d1_full[buf]);
      };
      mbar_wait(bar_b1, 0);
      gemm1(0);
      for (int i = 0; i < my_tiles; ++i) {
        if (i + 1 < my_tiles) gemm1(i + 1);
        const int s = i % NST;
        const uint32_t aQ = smem_u32(sQu + s * QSTAGE);
        mbar_wait(e_full, i & 1);
        tc_fence_after();
        // K = queue entry in slices of 8 (32 B inside a box row)
#pragma unroll
        for (int kk = 0; kk < KT / 8; ++kk)
          mma_tf32_ss(tmem_D2, make_sdesc(aQ + (kk >> 2) * QBOX + (kk & 3) * 32, 16, 1024),
                      make_sdesc(aE + (kk >> 2) * BBOX + (kk & 3) * 32, 16, 1024), idesc2,
                      (i > 0 || kk > 0) ? 1u : 0u);
        mma_commit(&qempty[s]);
        mma_commit(e_empty);
      }
      mma_commit(d2_full);
    }
  } else {
    const int quad = warp & 3;
    const int row = quad * 32 + lane;
    const uint32_t lane_off = uint32_t(quad * 32) << 16;
    float zacc[NB];
#pragma unroll
    for (int j = 0; j < NB; ++j) zacc[j] = 0.f;
    uint8_t *erow = sE + quad * BBOX + (lane & 3) * 4;
    for (int i = 0; i < my_tiles; ++i) {
      const int buf = i & 1, tph = (i >> 1) & 1;
      const int kg = (cta + i * (int)gridDim.x) * KT + row;
      const bool kvalid = kg < K;
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
          *reinterpret_cast<float *>(erow + bb * 128 + (((lane >> 2) ^ (bb & 7)) << 4)) = e;
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
    const long long pbase = (long long)(chunk * gridDim.x + cta) * NB;
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

__global__ void global_finalize_kernel(const float *__restrict__ part_z, const float *__restrict__ part_g,
                                       float *__restrict__ z, float *__restrict__ g, int B, int NB, int nctas) {
  const int b = blockIdx.x, chunk = b / 64, bl = b % 64, c = threadIdx.x;
  float s = 0.f, sz = 0.f;
  for (int cta = 0; cta < nctas; ++cta) {
    const long long pb = (long long)(chunk * nctas + cta) * NB + bl;
    s += part_g[pb * CDIM + c];
    if (c == 0) sz += part_z[pb];
  }
  g[(long long)b * CDIM + c] = s;
  if (c == 0) z[b] = sz;
}

int pick_nb(int B) { return B <= 16 ? 16 : (B <= 32 ? 32 : 64); }

int make_f32_tmap_2d(CUtensorMap *tm, const void *base, uint64_t rows, uint64_t cols, uint32_t box_rows) {
  EncodeTiledFn enc = get_encode_tiled();
  if (!enc) { cp2_set_error("cuTensorMapEncodeTiled unavailable in this driver"); return CP2_E_UNSUPPORTED; }
  const cuuint64_t dims[2] = {cols, rows};
  const cuuint64_t strides[1] = {cols * 4};
  const cuuint32_t box[2] = {32, box_rows};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void *>(base), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                         CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { cp2_set_error("cuTensorMapEncodeTiled(f32 2d) failed: CUresult %d", (int)r); return CP2_E_CUDA; }
  return CP2_OK;
}

template <int NB>
int launch_global(const CUtensorMap &tq, const CUtensorMap &tp, float *part_z, float *part_g, float scale_log2,
                  int K, int ntiles, int nctas, int chunks, cudaStream_t st) {
  const size_t smem = 1024 + (size_t)NST * QSTAGE + 8 * (size_t)NB * 128 + sizeof(float) * 4 * NB + 256;
  CP2_CUDA(cudaFuncSetAttribute(global_nce_kernel<NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  global_nce_kernel<NB><<<dim3(nctas, chunks), NTHREADS, smem, st>>>(tq, tp, part_z, part_g, scale_log2, K, ntiles);
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}

}  // namespace

extern "C" int64_t cp2_global_scratch_floats(int B, int C) {
  if (B <= 0) return 0;
  const int chunks = cp2_cdiv(B, 64);
  return (int64_t)chunks * cp2_num_sms() * pick_nb(B) * (1 + C);
}

extern "C" int cp2_global_nce(const float *q_pos, const float *queue, float *z, float *g, float *scratch,
                              float inv_tau, int B, int C, int K, void *stream) {
  CP2_REQUIRE(q_pos && queue && z && g && scratch, "cp2_global_nce: null pointer");
  CP2_REQUIRE(C == CDIM, "cp2_global_nce: C=%d unsupported (128)", C);
  CP2_REQUIRE(B >= 0 && K > 0 && K % 4 == 0, "cp2_global_nce: bad shape B=%d K=%d (K %% 4 == 0)", B, K);
  if (B == 0) return CP2_OK;
  const int NB = pick_nb(B), chunks = cp2_cdiv(B, 64);
  const int ntiles = cp2_cdiv(K, KT);
  const int nctas = ntiles < cp2_num_sms() ? ntiles : cp2_num_sms();
  CUtensorMap tq, tp;
  int rc;
  if ((rc = make_f32_tmap_2d(&tq, queue, CDIM, (uint64_t)K, CDIM)) != CP2_OK) return rc;
  if ((rc = make_f32_tmap_2d(&tp, q_pos, (uint64_t)B, CDIM, NB)) != CP2_OK) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float *part_z = scratch;
  float *part_g = scratch + (size_t)chunks * nctas * NB;
  const float scale_log2 = inv_tau * LOG2E;
  if (NB == 16) rc = launch_global<16>(tq, tp, part_z, part_g, scale_log2, K, ntiles, nctas, chunks, st);
  else if (NB == 32) rc = launch_global<32>(tq, tp, part_z, part_g, scale_log2, K, ntiles, nctas, chunks, st);
  else rc = launch_global<64>(tq, tp, part_z, part_g, scale_log2, K, ntiles, nctas, chunks, st);
  if (rc != CP2_OK) return rc;
  global_finalize_kernel<<<B, CDIM, 0, st>>>(part_z, part_g, z, g, B, NB, nctas);
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}
