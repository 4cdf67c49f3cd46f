// Pieces shared by the dense-loss tensor-core kernels (dense_umma.cu: the streaming forward / backward pair for
// any P; dense_small.cu: the single-kernel path for maps that fit TMEM, P <= 256): tile constants, the
// software-pipelined exp2 chunks of the softmax warps and the TMA descriptor of a [B,P,C] bf16 operand.
#pragma once
#include "common.cuh"
#include "umma.cuh"

namespace dense {

using namespace umma;

constexpr int TILE = 128;                 // MMA M (TMEM lanes): key pixels (fwd) / query pixels (bwd)
constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;
constexpr float NL_MASKED = -80.0f;       // exp2 argument offset of a masked key column (2^-80 ~ 0)

__device__ __forceinline__ uint8_t *align1024(uint8_t *p) {
  return reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(p) + 1023) & ~uintptr_t(1023));
}

// K-major operand tile [rows][C] stored as C/64 swizzled boxes of `box_bytes`: descriptor of the
// kk-th K=16 slice (32 bytes inside the 128-byte row).
__device__ __forceinline__ uint64_t kmajor_desc(uint32_t tile_addr, int kk, int box_bytes) {
  return make_sdesc(tile_addr + (kk >> 2) * box_bytes + (kk & 3) * 32, 16, 1024);
}

// One 32-column chunk of the backward softmax: P~ = exp2(S*scale + nl) packed to bf16 pairs.  Software
// pipeline over four 8-column groups in explicit program order: X(k) = scale/offset (FFMA2), E(k) = the
// exponentials (MUFU, or the FMA-pipe polynomial for the first PF pairs of a group), C(k) = bf16 pack.
template <int PF, bool ZS = false>
__device__ __forceinline__ void softmax_chunk_bwd(const uint32_t (&r)[32], uint32_t nl_saddr, f2 scale2,
                                                  uint32_t (&pk)[16], f2 *zsum = nullptr) {
  float4 n[8];
#pragma unroll
  for (int u = 0; u < 8; ++u) n[u] = lds_f4(nl_saddr + u * 16);
  f2 x[16];
  auto X = [&](int k) {
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      const int j = 4 * k + p;
      const f2 nlp = (j & 1) ? f2_pack(n[j >> 1].z, n[j >> 1].w) : f2_pack(n[j >> 1].x, n[j >> 1].y);
      x[j] = f2_fma_v(f2_pack_u(r[2 * j], r[2 * j + 1]), scale2, nlp);
    }
  };
  auto E = [&](int k) {
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      const int j = 4 * k + p;
      if (p < PF) {
        x[j] = exp2_poly2<2>(x[j]);
      } else {
        float a, c;
        f2_unpack(x[j], a, c);
        a = ex2_approx_v(a);
        c = ex2_approx_v(c);
        x[j] = f2_pack(a, c);
      }
    }
  };
  auto Cv = [&](int k) {
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      const int j = 4 * k + p;
      float a, c;
      if constexpr (ZS) *zsum = f2_add(*zsum, x[j]);      // fp32 row sums of P~ (queue-sums mode)
      f2_unpack(x[j], a, c);
      pk[j] = pack_bf16x2_v(a, c);
    }
  };
  X(0);
  X(1);
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    E(k);
    if (k + 2 < 4) X(k + 2);
    if (k > 0) Cv(k - 1);
  }
  Cv(3);
}

// One full 32-column chunk of the forward: adds sum_j exp2(S_j*scale) into the running pair sums.
template <int PF>
__device__ __forceinline__ void softmax_chunk_fwd(const uint32_t (&r)[32], f2 scale2, f2 &acc_m, f2 &acc_m2,
                                                  f2 &acc_p, f2 &acc_p2) {
  f2 x[16];
  auto X = [&](int k) {
#pragma unroll
    for (int p = PF; p < 4; ++p) x[4 * k + p] = f2_mul_v(f2_pack_u(r[8 * k + 2 * p], r[8 * k + 2 * p + 1]), scale2);
  };
  auto E = [&](int k) {
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      const int j = 4 * k + p;
      if (p < PF) {
        x[j] = exp2_poly2_scaled<3>(f2_pack_u(r[2 * j], r[2 * j + 1]), scale2);
      } else {
        float a, c;
        f2_unpack(x[j], a, c);
        a = ex2_approx_v(a);
        c = ex2_approx_v(c);
        x[j] = f2_pack(a, c);
      }
    }
  };
  auto A = [&](int k) {
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      const int j = 4 * k + p;
      if (p < PF) {
        if (k & 1) acc_p = f2_add(acc_p, x[j]); else acc_p2 = f2_add(acc_p2, x[j]);
      } else {
        if (k & 1) acc_m = f2_add_v(acc_m, x[j]); else acc_m2 = f2_add_v(acc_m2, x[j]);
      }
    }
  };
  X(0);
  X(1);
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    E(k);
    if (k + 2 < 4) X(k + 2);
    if (k > 0) A(k - 1);
  }
  A(3);
}


// TMA descriptor of a [B,P,C] bf16 feature operand: boxes of 64 channels (128 B, 128-byte swizzle) x box_rows pixels
inline int make_feature_tmap(CUtensorMap *tm, const void *base, int B, int C, int P, int box_rows) {
  EncodeTiledFn enc = get_encode_tiled();
  if (!enc) { cp2_set_error("cuTensorMapEncodeTiled unavailable in this driver"); return CP2_E_UNSUPPORTED; }
  const cuuint64_t dims[3] = {(cuuint64_t)C, (cuuint64_t)P, (cuuint64_t)B};
  const cuuint64_t strides[2] = {(cuuint64_t)C * 2, (cuuint64_t)P * C * 2};
  const cuuint32_t box[3] = {64, (cuuint32_t)box_rows, 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  const CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void *>(base), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                         CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { cp2_set_error("cuTensorMapEncodeTiled(features) failed: CUresult %d", (int)r); return CP2_E_CUDA; }
  return CP2_OK;
}

}  // namespace dense
