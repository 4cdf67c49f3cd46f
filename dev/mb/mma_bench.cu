// tcgen05.mma issue/throughput probe (B200): cycles per 128xNx16 bf16 MMA for the operand sources used by K2.
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I../../cp2_b200/csrc -I../../include -o mma_bench.bin mma_bench.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "umma.cuh"
using namespace umma;

// MODE 0: SS  S-GEMM (A = Q tile K-major, B = K tile K-major), N=128, 8 MMAs per "tile"
// MODE 1: TS  dQ-GEMM (A = TMEM, B = K tile MN-major), N=128, 8 MMAs per tile
// MODE 2: SS then TS alternating (the backward's mix), 16 MMAs per tile
// MODE 3: SS with N=256 (two key tiles side by side)
// BG: number of extra warps doing LDTM+STTM traffic on other TMEM columns (softmax stand-in)
template <int MODE, int BG>
__global__ void __launch_bounds__(64 + 32 * (BG > 0 ? BG : 1)) mma_kernel(long long *cycles, int tiles) {
  extern __shared__ uint8_t raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  __shared__ int stop;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 3 * 32768 / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); stop = 0; }
  if (warp == 0) tmem_alloc(&slot, 512);
  fence_proxy_async_smem();
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tmem = slot;
  if (warp == 0) {
    constexpr uint32_t id128 = make_idesc(128, 128, FMT_BF16, FMT_BF16, MAJOR_K, MAJOR_K);
    constexpr uint32_t id256 = make_idesc(128, 256, FMT_BF16, FMT_BF16, MAJOR_K, MAJOR_K);
    constexpr uint32_t idmn = make_idesc(128, 128, FMT_BF16, FMT_BF16, MAJOR_K, MAJOR_MN);
    const uint64_t dQ = make_sdesc(smem_u32(smem), 16, 1024);
    const uint64_t dK = make_sdesc(smem_u32(smem + 32768), 16, 1024);
    const uint64_t dKmn = make_sdesc(smem_u32(smem + 32768), 16384, 1024);
    const long long t0 = clock64();
    for (int t = 0; t < tiles; ++t) {
      if (elect_one()) {
        if (MODE == 0 || MODE == 2) {
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) {
            const uint32_t off = (kk >> 2) * 16384 + (kk & 3) * 32;
            mma_f16_ss(tmem + (t & 1) * 128, sdesc_advance(dQ, off), sdesc_advance(dK, off), id128, kk > 0);
          }
        }
        if (MODE == 1 || MODE == 2) {
#pragma unroll
          for (int kk = 0; kk < 8; ++kk)
            mma_f16_ts(tmem + 256, tmem + (t & 1) * 128 + kk * 8, sdesc_advance(dKmn, kk * 2048), idmn, 1u);
        }
        if (MODE == 3) {
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) {
            const uint32_t off = (kk >> 2) * 16384 + (kk & 3) * 32;
            mma_f16_ss(tmem, sdesc_advance(dQ, off), sdesc_advance(dK, off), id256, kk > 0);
          }
        }
        if (MODE == 4 || MODE == 5) {            // TS S-GEMM: A = stationary tile in TMEM (cols 320..383), B K-major
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) {
            const uint32_t off = (kk >> 2) * 16384 + (kk & 3) * 32;
            mma_f16_ts(tmem + (t & 1) * 128, tmem + 320 + kk * 8, sdesc_advance(dK, off), id128, kk > 0);
          }
        }
        if (MODE == 5) {                         // + TS dQ-GEMM (proposed backward mix)
#pragma unroll
          for (int kk = 0; kk < 8; ++kk)
            mma_f16_ts(tmem + 256, tmem + (t & 1) * 128 + kk * 8, sdesc_advance(dKmn, kk * 2048), idmn, 1u);
        }
        if (MODE == 6) {                         // TS S-GEMM with N=256
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) {
            const uint32_t off = (kk >> 2) * 16384 + (kk & 3) * 32;
            mma_f16_ts(tmem, tmem + 320 + kk * 8, sdesc_advance(dK, off), id256, kk > 0);
          }
        }
        if ((t & 7) == 7 || t == tiles - 1) mma_commit(&bar);
      }
      __syncwarp();
      if ((t & 7) == 7 || t == tiles - 1) mbar_wait(&bar, (t >> 3) & 1);
    }
    const long long t1 = clock64();
    if (lane == 0) { cycles[blockIdx.x] = t1 - t0; *(volatile int *)&stop = 1; }
  } else if (BG > 0 && warp >= 2) {
    // softmax stand-in: read 32 columns, write 16 back, on columns 384.. (not touched by the MMAs)
    const uint32_t base = tmem + 384 + (uint32_t((warp & 3) * 32) << 16) + ((warp >> 2) & 1) * 64;
    uint32_t r[32], pk[16];
    while (*(volatile int *)&stop == 0) {
      tmem_ld32(base, r);
      tmem_wait_ld();
#pragma unroll
      for (int i = 0; i < 16; ++i) pk[i] = r[2 * i] ^ r[2 * i + 1];
      tmem_st16(base + 32, pk);
      tmem_wait_st();
    }
  }
  tc_fence_before(); __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

template <int MODE, int BG>
void run(const char *name, int mmas_per_tile) {
  long long *cyc; cudaMalloc(&cyc, 148 * 8);
  const int tiles = 4000, smem = 3 * 32768 + 2048;
  cudaFuncSetAttribute(mma_kernel<MODE, BG>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  mma_kernel<MODE, BG><<<148, 64 + 32 * (BG > 0 ? BG : 1), smem>>>(cyc, tiles);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(e)); exit(1); }
  long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
  double mx = 0; for (int i = 0; i < 148; ++i) mx = h[i] > mx ? h[i] : mx;
  printf("%-52s %8.1f cycles/tile  %6.1f cycles/MMA\n", name, mx / tiles, mx / tiles / mmas_per_tile);
  cudaFree(cyc);
}

int main() {
  run<0, 0>("SS  128x128x16 x8", 8);
  run<1, 0>("TS  128x128x16 x8 (B MN-major)", 8);
  run<2, 0>("SS x8 + TS x8 (backward mix)", 16);
  run<3, 0>("SS  128x256x16 x8", 8);
  run<4, 0>("TS  128x128x16 x8 (B K-major; fwd proposal)", 8);
  run<5, 0>("TS S x8 + TS dQ x8 (bwd proposal)", 16);
  run<6, 0>("TS  128x256x16 x8 (B K-major)", 8);
  run<0, 8>("SS  128x128x16 x8   + 8 warps LDTM/STTM", 8);
  run<1, 8>("TS  128x128x16 x8   + 8 warps LDTM/STTM", 8);
  run<2, 8>("SS x8 + TS x8       + 8 warps LDTM/STTM", 16);
  return 0;
}
