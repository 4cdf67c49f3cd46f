// K5 — queue enqueue with wrap-around and a device-side pointer
// (reference builder.py:574-587).  keys_all [n,C] (row-major, the NCCL all-gather's
// output) is written transposed into columns (ptr+i) mod K of queue [C,K].
// A 32x32 tile goes through shared memory so both the read (along C) and the write
// (along the queue's K axis) are coalesced.  Latency-bound (262 KB at W=8, B=32).
#include "common.cuh"

namespace {

__global__ void __launch_bounds__(256)
enqueue_kernel(const float *__restrict__ keys, float *__restrict__ queue,
               const int64_t *__restrict__ queue_ptr, int n, int C, int K) {
  __shared__ float tile[32][33];
  const long long ptr = *queue_ptr;               // every block reads the same value
  const int i0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 8 rows per pass
#pragma unroll
  for (int r = ty; r < 32; r += 8) {
    const int i = i0 + r, c = c0 + tx;
    tile[r][tx] = (i < n && c < C) ? keys[(long long)i * C + c] : 0.0f;
  }
  __syncthreads();
#pragma unroll
  for (int r = ty; r < 32; r += 8) {
    const int c = c0 + r, i = i0 + tx;
    if (c < C && i < n) queue[(long long)c * K + (ptr + i) % K] = tile[tx][r];
  }
}

__global__ void advance_ptr_kernel(int64_t *queue_ptr, int n, int K) {
  *queue_ptr = (*queue_ptr + n) % K;
}

}  // namespace

extern "C" int cp2_enqueue(const float *keys_all, float *queue, int64_t *queue_ptr, int n, int C,
                           int K, void *stream) {
  CP2_REQUIRE(n >= 0 && C > 0 && K > 0 && n <= K, "cp2_enqueue: bad shape n=%d C=%d K=%d", n, C, K);
  if (n == 0) return CP2_OK;
  CP2_REQUIRE(keys_all && queue && queue_ptr, "cp2_enqueue: null pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  dim3 grid(cp2_cdiv(n, 32), cp2_cdiv(C, 32));
  enqueue_kernel<<<grid, 256, 0, st>>>(keys_all, queue, queue_ptr, n, C, K);
  CP2_LAUNCH_CHECK();
  advance_ptr_kernel<<<1, 1, 0, st>>>(queue_ptr, n, K);   // stream order: after every reader
  CP2_LAUNCH_CHECK();
  return CP2_OK;
}
