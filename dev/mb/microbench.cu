// Pipe-throughput probes for the K2 design (B200): TMEM read (tcgen05.ld), MUFU ex2, FFMA, FFMA2.
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I../../cp2_b200/csrc -I../../include -o microbench.bin microbench.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "umma.cuh"
using namespace umma;

template <int MODE>   // 0: x32 ld + wait each; 1: 4 x32 lds then one wait; 2: x16
__global__ void __launch_bounds__(512) tmem_ld_kernel(long long *cycles, unsigned *sink, int iters) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) tmem_alloc(&slot, 512);
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t base = slot + (uint32_t((warp & 3) * 32) << 16);
  unsigned acc = 0;
  __syncthreads();
  const long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
    if (MODE == 0) {
#pragma unroll
      for (int c = 0; c < 4; ++c) { uint32_t r[32]; tmem_ld32(base + ((i + c) & 3) * 32 + (warp >> 2) * 128 % 384, r); tmem_wait_ld(); acc ^= r[0] ^ r[31]; }
    } else if (MODE == 1) {
      uint32_t r0[32], r1[32], r2[32], r3[32];
      tmem_ld32(base + 0, r0); tmem_ld32(base + 32, r1); tmem_ld32(base + 64, r2); tmem_ld32(base + 96, r3);
      tmem_wait_ld(); acc ^= r0[0] ^ r1[1] ^ r2[2] ^ r3[3];
    } else {
#pragma unroll
      for (int c = 0; c < 8; ++c) { uint32_t r[16]; tmem_ld16(base + c * 16, r); tmem_wait_ld(); acc ^= r[0] ^ r[15]; }
    }
  }
  const long long t1 = clock64();
  __syncthreads();
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  if (acc == 0x12345678u) sink[0] = acc;
  tc_fence_before(); __syncthreads();
  if (warp == 0) tmem_dealloc(slot, 512);
}

template <int MODE>   // 0 MUFU ex2, 1 FFMA scalar, 2 FFMA2 packed, 3 MUFU + FFMA2 interleaved, 4 poly2<3>, 5 cvt bf16x2
__global__ void __launch_bounds__(512) alu_kernel(long long *cycles, float *sink, int iters) {
  float a[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) a[j] = 0.001f * (threadIdx.x + j);
  f2 p[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) p[j] = f2_pack(a[2 * j], a[2 * j + 1]);
  const f2 s2 = f2_pack(0.999f, 0.999f), c2 = f2_pack(0.001f, 0.001f);
  __syncthreads();
  const long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (MODE == 0) a[j] = ex2_approx(a[j]);
      if (MODE == 1) a[j] = fmaf(a[j], 0.999f, 0.001f);
      if (MODE == 3) a[j] = ex2_approx(a[j]);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (MODE == 2 || MODE == 3) { p[j] = f2_fma(p[j], s2, c2); p[j] = f2_fma(p[j], s2, c2); }
      if (MODE == 4) p[j] = exp2_poly2<3>(p[j]);
      if (MODE == 5) { float x, y; f2_unpack(p[j], x, y); uint32_t u = pack_bf16x2(x, y); p[j] = f2_pack_u(u, u ^ 1u); }
    }
  }
  const long long t1 = clock64();
  float r = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) r += a[j];
#pragma unroll
  for (int j = 0; j < 4; ++j) { float x, y; f2_unpack(p[j], x, y); r += x + y; }
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  if (r == 1.2345f) sink[0] = r;
}

// MODE 0: per 8 pairs: 8 FMUL2 + 16 MUFU + 8 FADD2 (the forward's MUFU path, packed fp32x2)
// MODE 1: same with scalar FMUL/FADD (16 + 16 + 16)
// MODE 2: MODE 0 in bursts (8 FMUL2, then 16 MUFU, then 8 FADD2)
template <int MODE>
__global__ void __launch_bounds__(512) mix_kernel(long long *cycles, float *sink, int iters, float scale) {
  f2 acc[4];
  float sa[8];
#pragma unroll
  for (int j = 0; j < 4; ++j) acc[j] = f2_pack(0.f, 0.f);
#pragma unroll
  for (int j = 0; j < 8; ++j) sa[j] = 0.f;
  f2 src[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) src[j] = f2_pack(0.001f * (threadIdx.x + j), 0.002f * (threadIdx.x + j));
  const f2 scale2 = f2_pack(scale, scale);
  __syncthreads();
  const long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int j = 0; j < 8; ++j) src[j].v += 0x0000000100000001ull;      // new inputs every iteration (ALU pipe)
    if (MODE == 0) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        f2 x = f2_mul_v(src[j], scale2);
        float a, c;
        f2_unpack(x, a, c);
        a = ex2_approx_v(a);
        c = ex2_approx_v(c);
        acc[j & 3] = f2_add_v(acc[j & 3], f2_pack(a, c));
      }
    } else if (MODE == 1) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float a, c;
        f2_unpack(src[j], a, c);
        a = ex2_approx_v(a * scale);
        c = ex2_approx_v(c * scale);
        sa[j] += a;
        sa[(j + 4) & 7] += c;
      }
    } else {
      f2 x[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) x[j] = f2_mul_v(src[j], scale2);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float a, c;
        f2_unpack(x[j], a, c);
        a = ex2_approx_v(a);
        c = ex2_approx_v(c);
        x[j] = f2_pack(a, c);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j & 3] = f2_add_v(acc[j & 3], x[j]);
    }
  }
  const long long t1 = clock64();
  float r = 0.f;
#pragma unroll
  for (int j = 0; j < 4; ++j) { float x, y; f2_unpack(acc[j], x, y); r += x + y; }
#pragma unroll
  for (int j = 0; j < 8; ++j) r += sa[j];
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  if (r == 1.2345f) sink[0] = r;
}

int main() {
  long long *cyc; unsigned *sink; cudaMalloc(&cyc, 148 * 8); cudaMalloc(&sink, 64);
  long long h[148];
  const int iters = 2000;
  auto report = [&](const char *name, double units_per_cta, const char *unit) {
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(e)); return; }
    cudaMemcpy(h, cyc, 148 * 8, cudaMemcpyDeviceToHost);
    double mx = 0; for (int i = 0; i < 148; ++i) mx = h[i] > mx ? h[i] : mx;
    printf("%-46s %10.0f cycles  %8.2f %s/clk/SM\n", name, mx, units_per_cta / mx, unit);
  };
  for (int warps : {4, 8, 16}) {
    char nm[96];
    tmem_ld_kernel<0><<<148, warps * 32>>>(cyc, sink, iters);
    snprintf(nm, 96, "tmem ld32+wait            warps=%d", warps); report(nm, (double)iters * 4 * 4096 * warps, "B");
    tmem_ld_kernel<1><<<148, warps * 32>>>(cyc, sink, iters);
    snprintf(nm, 96, "tmem 4xld32 then wait     warps=%d", warps); report(nm, (double)iters * 4 * 4096 * warps, "B");
    tmem_ld_kernel<2><<<148, warps * 32>>>(cyc, sink, iters);
    snprintf(nm, 96, "tmem ld16+wait            warps=%d", warps); report(nm, (double)iters * 8 * 2048 * warps, "B");
  }
  for (int warps : {4, 8, 16}) {
    char nm[96];
    alu_kernel<0><<<148, warps * 32>>>(cyc, (float *)sink, iters);
    snprintf(nm, 96, "MUFU ex2                  warps=%d", warps); report(nm, (double)iters * 8 * 32 * warps, "ex2");
    alu_kernel<1><<<148, warps * 32>>>(cyc, (float *)sink, iters);
    snprintf(nm, 96, "FFMA scalar               warps=%d", warps); report(nm, (double)iters * 8 * 32 * warps, "fma");
    alu_kernel<2><<<148, warps * 32>>>(cyc, (float *)sink, iters);
    snprintf(nm, 96, "FFMA2 packed              warps=%d", warps); report(nm, (double)iters * 16 * 32 * warps, "fma");
    alu_kernel<3><<<148, warps * 32>>>(cyc, (float *)sink, iters);
    snprintf(nm, 96, "MUFU + FFMA2 (8 ex2+16fma) warps=%d", warps); report(nm, (double)iters * 8 * 32 * warps, "ex2(+2fma)");
    alu_kernel<4><<<148, warps * 32>>>(cyc, (float *)sink, iters);
    snprintf(nm, 96, "exp2_poly2<3>             warps=%d", warps); report(nm, (double)iters * 8 * 32 * warps, "exp");
    alu_kernel<5><<<148, warps * 32>>>(cyc, (float *)sink, iters);
    snprintf(nm, 96, "cvt.bf16x2                warps=%d", warps); report(nm, (double)iters * 4 * 32 * warps, "cvt");
  }
  for (int warps : {4, 8, 16}) {
    char nm[96];
    mix_kernel<0><<<148, warps * 32>>>(cyc, (float *)sink, iters, 0.999f);
    snprintf(nm, 96, "mix FMUL2+2MUFU+FADD2 interleaved warps=%d", warps); report(nm, (double)iters * 16 * 32 * warps, "ex2");
    mix_kernel<1><<<148, warps * 32>>>(cyc, (float *)sink, iters, 0.999f);
    snprintf(nm, 96, "mix scalar FMUL+MUFU+FADD        warps=%d", warps); report(nm, (double)iters * 16 * 32 * warps, "ex2");
    mix_kernel<2><<<148, warps * 32>>>(cyc, (float *)sink, iters, 0.999f);
    snprintf(nm, 96, "mix bursts 8 FMUL2/16 MUFU/8 FADD2 warps=%d", warps); report(nm, (double)iters * 16 * 32 * warps, "ex2");
  }
  return 0;
}
