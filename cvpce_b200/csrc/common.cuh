// Shared host/device helpers for the cvpce_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include "../../include/cvpce_b200.h"

namespace cvpce {

// Thread-local last-error string (cvpce_last_error()).
char* err_buf();
int set_err(int code, const char* fmt, ...);

#define CVPCE_CUDA_OK(expr)                                                                         \
  do {                                                                                              \
    cudaError_t _e = (expr);                                                                        \
    if (_e != cudaSuccess)                                                                          \
      return cvpce::set_err(CVPCE_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                            __FILE__, __LINE__);                                                    \
  } while (0)

#define CVPCE_REQUIRE(cond, code, ...)                  \
  do {                                                  \
    if (!(cond)) return cvpce::set_err(code, __VA_ARGS__); \
  } while (0)

// Launch check that does not synchronise.
#define CVPCE_LAUNCH_OK(name)                                                                          \
  do {                                                                                                 \
    cudaError_t _e = cudaGetLastError();                                                               \
    if (_e != cudaSuccess)                                                                             \
      return cvpce::set_err(CVPCE_ERR_CUDA, "launch of %s failed: %s", name, cudaGetErrorString(_e));  \
  } while (0)

int require_sm100();  // CVPCE_OK or CVPCE_ERR_NO_DEVICE (cached per device)
int sm_count();       // SM count of the current device (148 on B200)

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

}  // namespace cvpce
