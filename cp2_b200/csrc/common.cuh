// Shared host-side plumbing for the cp2_b200 C ABI (include/cp2_b200.h).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cstdarg>
#include <cstdint>
#include <cstdio>

#include "cp2_b200.h"

void cp2_set_error(const char *fmt, ...);

#define CP2_CUDA(call)                                                                     \
  do {                                                                                     \
    cudaError_t e_ = (call);                                                               \
    if (e_ != cudaSuccess) {                                                               \
      cp2_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
      return CP2_E_CUDA;                                                                   \
    }                                                                                      \
  } while (0)

#define CP2_REQUIRE(cond, ...)      \
  do {                              \
    if (!(cond)) {                  \
      cp2_set_error(__VA_ARGS__);   \
      return CP2_E_INVALID;         \
    }                               \
  } while (0)

#define CP2_LAUNCH_CHECK() CP2_CUDA(cudaGetLastError())

// cudaFuncAttributeMaxDynamicSharedMemorySize is sticky per (function, device): set it on the first launch of a
// kernel instantiation on each device instead of on every call (it is a driver round trip on the launch path).
#define CP2_SMEM_ONCE(kern, bytes)                                                                        \
  do {                                                                                                    \
    static unsigned long long done_ = 0ull;                                                               \
    int dev_ = 0;                                                                                         \
    CP2_CUDA(cudaGetDevice(&dev_));                                                                       \
    if (!((done_ >> (dev_ & 63)) & 1ull)) {                                                               \
      CP2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(bytes)));    \
      done_ |= 1ull << (dev_ & 63);                                                                       \
    }                                                                                                     \
  } while (0)

static inline int cp2_cdiv(long long a, long long b) { return (int)((a + b - 1) / b); }

// SM count of the current device (cached per process; 148 on B200).
int cp2_num_sms();
