// K5 — queue enqueue with wrap-around and a device-side pointer, ONE launch
// (reference builder.py:574-587).  keys_all [n,C] (row-major, the NCCL all-gather's
// output) is written transposed into columns (ptr+i) mod K of queue [C,K] and, when a
// bf16 shadow [K,C] is kept for cp2_global_nce, as rows (ptr+i) mod K of the shadow.
// A 32x32 tile goes through shared memory so both the read (along C) and the write
// (along the queue's K axis) are coalesced.  Latency-bound (262 KB at W=8, B=32).
// The pointer is advanced by the last CTA to have READ it (done counter in a persistent,
// self-resetting scratch word): no second launch, no host read of the pointer.
#include "common.cuh"

namespace {

__global__ void __launch_bounds__(256)
enqueue_kernel(const float *__restrict__ keys, float *__restrict__ queue, __nv_bfloat16 *__restrict__ shadow,
               int64_t *__restrict__ queue_ptr, unsigned int *__restrict__ counter, int n, int C, int K) {
  __shared__ float tile[32][33];
  __shared__ long long ptr_s;
  if (threadIdx.x == 0) ptr_s = *reinterpret_cast<volatile long long *>(queue_ptr);
  const int i0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 8 rows per pass
#pragma unroll
  for (int r = ty; r < 32; r += 8) {
    const int i = i0 + r, c = c0 + tx;
    tile[r][tx] = (i < n && c < C) ? keys[(long long)i * C + c] : 0.0f;
  }
  __syncthreads();
  const long long ptr = ptr_s;
  if (threadIdx.x == 0) {                 // this CTA holds the old pointer: the last one to say so advances it
    __threadfence();
    const unsigned int total = gridDim.x * gridDim.y;
    if (atomicAdd(counter, 1u) == total - 1u) {
      *queue_ptr = (ptr + n) % K;
      *counter = 0u;
    }
  }
#pragma unroll
  for (int r = ty; r < 32; r += 8) {
    const int c = c0 + r, i = i0 + tx;
    if (c < C && i < n) queue[(long long)c * K + (ptr + i) % K] = tile[tx][r];
  }
  if (shadow != nullptr) {
#pragma unroll
    for (int r = ty; r < 32; r += 8) {
      const int i = i0 + r, c = c0 + tx;
      if (i < n && c < C) shadow[((ptr + i) % K) * C + c] = __float2bfloat16_rn(tile[r][tx]);
    }
  }
}

}  // namespace

extern "C" int cp2_enqueue(const float *keys_all, float *queue, void *queue_bf, int64_t *queue_ptr,
                           void *scratch_counter, int n, int C, int K, void *stream) {
  CP2_REQUIRE(n >= 0 && C > 0 && K > 0 && n <= K, "cp2_enqueue: bad shape n=%d C=%d K=%d", n, C, K);
  if (n == 0) return CP2_OK;
  CP2_REQUIRE(keys_all && queue && queue_ptr && scratch_counter, "cp2_enqueue: null pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  dim3 grid(cp2_cdiv(n, 32), cp2_cdiv(C, 32));
  enqueue_kernel<<<grid, 256, 0, st>>>(keys_all, queue, static_cast<__nv_bfloat16 *>(queue_bf), queue_ptr,
                                       static_cast<unsigned int *>(scratch_counter), n, C, K);
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}
