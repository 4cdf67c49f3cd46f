// Hand-written sm_100a plumbing: mbarrier, TMA (cp.async.bulk.tensor), tcgen05
// (TMEM alloc / mma / commit / ld / st) and the shared-memory / instruction descriptor
// encodings.  Bit layouts follow the PTX ISA "tcgen05 matrix descriptor" /
// "instruction descriptor" tables (cross-checked against cute/arch/mma_sm100_desc.hpp).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>

namespace umma {

// ------------------------------------------------------------------ descriptors
constexpr uint32_t FMT_F16 = 0, FMT_BF16 = 1, FMT_TF32 = 2;
constexpr uint32_t MAJOR_K = 0, MAJOR_MN = 1;

// 32-bit instruction descriptor for tcgen05.mma kind::f16 / kind::tf32, fp32 accumulate.
__host__ __device__ constexpr uint32_t make_idesc(uint32_t M, uint32_t N, uint32_t a_fmt, uint32_t b_fmt,
                                                  uint32_t a_major, uint32_t b_major) {
  return (1u << 4)                 // c_format = F32
         | (a_fmt << 7) | (b_fmt << 10) | (a_major << 15) | (b_major << 16) |
         ((N >> 3) << 17) | ((M >> 4) << 24);
}

// 64-bit shared-memory matrix descriptor, 128-byte swizzle, version 1 (Blackwell).
// K-major operand : rows of 128 B (64 bf16 / 32 tf32), 8-row groups SBO apart; LBO unused (=16 B).
// MN-major operand: 128 B of contiguous MN elements per K row, 8 K-rows per group (SBO apart),
//                   successive 128-B MN chunks LBO apart.
__device__ __forceinline__ uint64_t make_sdesc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= (uint64_t)1 << 46;          // descriptor version
  d |= (uint64_t)2 << 61;          // SWIZZLE_128B
  return d;
}

__device__ __forceinline__ uint32_t smem_u32(const void *p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// One lane of a converged warp (the MMA / TMA issue pattern: the whole warp runs the loop so that loop
// state stays warp-uniform, only the elected lane issues the asynchronous instruction).
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "elect.sync _|p, 0xffffffff;\n"
      "selp.b32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(pred));
  return pred != 0;
}
// Descriptor of the same operand `bytes` further into shared memory (start-address field, 16-byte units).
__device__ __forceinline__ uint64_t sdesc_advance(uint64_t desc, uint32_t bytes) { return desc + (bytes >> 4); }

// ---------------------------------------------------------------------- mbarrier
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.b32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait: a protocol bug traps (launch failure) instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > 4000000000LL) {   // ~2 s at 2 GHz
      printf("cp2_b200: mbarrier timeout block=(%d,%d) thread=%d bar=%p parity=%u\n", blockIdx.x, blockIdx.y,
             threadIdx.x, (void *)bar, parity);
      __trap();
    }
  }
}

// --------------------------------------------------------------------------- TMA
__device__ __forceinline__ void prefetch_tmap(const CUtensorMap *m) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(m) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void *dst, const CUtensorMap *m, uint64_t *bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(dst)), "l"(m), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_3d(void *dst, const CUtensorMap *m, uint64_t *bar, int c0, int c1,
                                            int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(smem_u32(dst)), "l"(m), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}

// ----------------------------------------------------------------------- tcgen05
__device__ __forceinline__ void tmem_alloc(uint32_t *dst_smem, uint32_t ncols) {   // whole warp
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)),
               "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {      // whole warp
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// D[tmem] (+)= A[smem] * B[smem]            (one thread issues)
__device__ __forceinline__ void mma_f16_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                           uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem]
__device__ __forceinline__ void mma_f16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc,
                                           uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n"
      "}\n" ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void mma_tf32_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// mbarrier arrives once every previously issued tcgen05.mma of this thread has completed
// (implies tcgen05.fence::before_thread_sync).
__device__ __forceinline__ void mma_commit(uint64_t *bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}

// TMEM -> registers: this thread's lane, 32 / 16 consecutive 32-bit columns.
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
// registers -> TMEM: this thread's lane, 16 consecutive 32-bit columns.
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
      "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}

// 1-D bulk copy global -> shared (bytes % 16 == 0, both addresses 16-byte aligned), completes on an mbarrier
__device__ __forceinline__ void bulk_load_1d(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// ------------------------------------------------------- packed fp32x2 math (FFMA2 / FADD2 / FMUL2)
struct f2 { unsigned long long v; };
__device__ __forceinline__ f2 f2_pack(float lo, float hi) {
  f2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r.v) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ f2 f2_pack_u(uint32_t lo, uint32_t hi) {
  f2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r.v) : "r"(lo), "r"(hi));
  return r;
}
__device__ __forceinline__ void f2_unpack(f2 a, float &lo, float &hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(a.v));
}
__device__ __forceinline__ void f2_unpack_u(f2 a, uint32_t &lo, uint32_t &hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=r"(lo), "=r"(hi) : "l"(a.v));
}
__device__ __forceinline__ f2 f2_fma(f2 a, f2 b, f2 c) {
  f2 r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r.v) : "l"(a.v), "l"(b.v), "l"(c.v));
  return r;
}
__device__ __forceinline__ f2 f2_add(f2 a, f2 b) {
  f2 r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v));
  return r;
}
__device__ __forceinline__ f2 f2_sub(f2 a, f2 b) {
  f2 r;
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v));
  return r;
}
__device__ __forceinline__ f2 f2_mul(f2 a, f2 b) {
  f2 r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v));
  return r;
}

// p + (t << 23) as ONE integer multiply-add (ptxas emits SHF.L + IADD3 for the shift-add form)
__device__ __forceinline__ uint32_t exp_splice(uint32_t p, uint32_t t) {
  uint32_t r;
  asm("mad.lo.u32 %0, %1, 8388608, %2;" : "=r"(r) : "r"(t), "r"(p));
  return r;
}

// 2^x for a pair on the FMA/ALU pipes (no MUFU): Cody-Waite split x = n + f, |f| <= 0.5 by the
// 1.5*2^23 rounding trick, minimax polynomial of degree DEG for 2^f (max rel. error 1.7e-3 / 7.6e-5 /
// 2.7e-6 for DEG 2 / 3 / 4), exponent spliced in with one shift-add.  Requires -126 <= x <= 126.
template <int DEG>
__device__ __forceinline__ f2 exp2_poly2(f2 x) {
  const f2 magic = f2_pack(12582912.f, 12582912.f);
  const f2 t = f2_add(x, magic);
  const f2 f = f2_sub(x, f2_sub(t, magic));
  f2 p;
  if constexpr (DEG == 2) {
    p = f2_fma(f2_pack(0.23840933f, 0.23840933f), f, f2_pack(0.70340872f, 0.70340872f));
    p = f2_fma(p, f, f2_pack(1.00044143f, 1.00044143f));
  } else if constexpr (DEG == 3) {
    p = f2_fma(f2_pack(0.05516955f, 0.05516955f), f, f2_pack(0.24260588f, 0.24260588f));
    p = f2_fma(p, f, f2_pack(0.69326085f, 0.69326085f));
    p = f2_fma(p, f, f2_pack(0.99992841f, 0.99992841f));
  } else {
    p = f2_fma(f2_pack(0.0095698982f, 0.0095698982f), f, f2_pack(0.055917259f, 0.055917259f));
    p = f2_fma(p, f, f2_pack(0.24024744f, 0.24024744f));
    p = f2_fma(p, f, f2_pack(0.69312191f, 0.69312191f));
    p = f2_fma(p, f, f2_pack(0.99999928f, 0.99999928f));
  }
  uint32_t p0, p1, t0, t1;
  f2_unpack_u(p, p0, p1);
  f2_unpack_u(t, t0, t1);
  return f2_pack_u(exp_splice(p0, t0), exp_splice(p1, t1));
}

// 2^(r*scale) for a pair, the scale folded into the range reduction (saves the separate multiply).
template <int DEG>
__device__ __forceinline__ f2 exp2_poly2_scaled(f2 r, f2 scale) {
  const f2 magic = f2_pack(12582912.f, 12582912.f);
  const f2 t = f2_fma(r, scale, magic);
  const f2 f = f2_fma(r, scale, f2_sub(magic, t));
  f2 p;
  if constexpr (DEG == 2) {
    p = f2_fma(f2_pack(0.23840933f, 0.23840933f), f, f2_pack(0.70340872f, 0.70340872f));
    p = f2_fma(p, f, f2_pack(1.00044143f, 1.00044143f));
  } else if constexpr (DEG == 3) {
    p = f2_fma(f2_pack(0.05516955f, 0.05516955f), f, f2_pack(0.24260588f, 0.24260588f));
    p = f2_fma(p, f, f2_pack(0.69326085f, 0.69326085f));
    p = f2_fma(p, f, f2_pack(0.99992841f, 0.99992841f));
  } else {
    p = f2_fma(f2_pack(0.0095698982f, 0.0095698982f), f, f2_pack(0.055917259f, 0.055917259f));
    p = f2_fma(p, f, f2_pack(0.24024744f, 0.24024744f));
    p = f2_fma(p, f, f2_pack(0.69312191f, 0.69312191f));
    p = f2_fma(p, f, f2_pack(0.99999928f, 0.99999928f));
  }
  uint32_t p0, p1, t0, t1;
  f2_unpack_u(p, p0, p1);
  f2_unpack_u(t, t0, t1);
  return f2_pack_u(exp_splice(p0, t0), exp_splice(p1, t1));
}

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}

// "volatile" twins: the compiler keeps volatile asm statements in program order, which is how the softmax
// loops interleave MUFU work with FMA / conversion work of neighbouring column groups (the two softmax
// warps that share an SM sub-partition would otherwise run their MUFU phases in lock-step).
__device__ __forceinline__ float ex2_approx_v(float x) {
  float y;
  asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ f2 f2_fma_v(f2 a, f2 b, f2 c) {
  f2 r;
  asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r.v) : "l"(a.v), "l"(b.v), "l"(c.v));
  return r;
}
__device__ __forceinline__ f2 f2_mul_v(f2 a, f2 b) {
  f2 r;
  asm volatile("mul.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v));
  return r;
}
__device__ __forceinline__ f2 f2_add_v(f2 a, f2 b) {
  f2 r;
  asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v));
  return r;
}
__device__ __forceinline__ uint32_t pack_bf16x2_v(float lo, float hi) {
  uint32_t r;
  asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ float4 lds_f4(uint32_t saddr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(saddr));
  return v;
}

// ------------------------------------------------------- host: tensor-map encode
// Resolves cuTensorMapEncodeTiled through the runtime (no link-time libcuda dependency).
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);
EncodeTiledFn get_encode_tiled();   // defined in tmap.cu; nullptr if the driver lacks it

}  // namespace umma
