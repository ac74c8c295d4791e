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
int tuning(const char* name);  // development override set through cvpce_tuning_set (0 = automatic)
// Opt a kernel into `bytes` of dynamic shared memory on the CURRENT device (cached per (device, function); no-op while
// dynamic + static surely fit the 48 KB default).
int ensure_dynamic_smem(const void* func, size_t bytes);

// rmac_staged.cu: the persistent TMA-staged R-MAC kernel.  CVPCE_OK = launched; kRmacNotApplicable = the caller falls back
// to the cluster kernel of region_pool.cu; anything else is an error.  Arguments validated by the caller.
constexpr int kRmacNotApplicable = -1000;  // not a cvpce_status value
int rmac_staged_launch(const cvpce_feature_map* maps, int n_maps, int dtype, int B, const int32_t* regions, const int32_t* n_regions,
                       int per_region_l2, int final_l2, float eps, float* out, int64_t out_stride, void* stream);

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

// ---------------------------------------------------------------------------------------------------------------------
// Ascending bitonic sort of `cap` (a power of two >= 2) 64-bit keys in shared memory by the whole block.
// The network is the textbook one; what is tuned is where each compare-exchange step runs, because a step that goes
// through shared memory costs one full pass over the keys plus a block barrier:
//   * strides 32..1 of every stage run in registers, one warp per 64-key chunk (lane holds keys lane and lane+32;
//     stride 32 is in-thread, 16..1 are 64-bit warp shuffles) -- and the first six stages (k <= 64) entirely so;
//   * strides >= 64 are fused three at a time: a thread loads the 8 keys of a radix-8 butterfly, runs 3 steps in
//     registers and stores them (consecutive threads touch consecutive keys: conflict-free).
// 8192 keys: 20 passes + barriers instead of 91.  Every thread of the block must call it; keys must be visible
// (barrier) before the call and are visible to all threads after it.
// ---------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void sort_cx(unsigned long long& lo, unsigned long long& hi, const bool up) {
  if ((lo > hi) == up) {
    const unsigned long long t = lo;
    lo = hi;
    hi = t;
  }
}

// strides 32..1 of stage k (or, with first == true, all stages 2..64) for the 64-key chunks, one warp per chunk
__device__ __forceinline__ void sort_warp_pass(unsigned long long* keys, const int cap, const int k, const bool first) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  for (int c = warp; c < (cap >> 6); c += nwarps) {
    const int ia = (c << 6) + lane, ib = ia + 32;
    unsigned long long a = keys[ia], b = keys[ib];
    for (int kk = first ? 2 : k; kk <= (first ? 64 : k); kk <<= 1) {
      const bool up_a = (ia & kk) == 0, up_b = (ib & kk) == 0;
      if (kk >= 64) sort_cx(a, b, up_a);  // stride 32 (only from stage 64 on)
#pragma unroll
      for (int j = 16; j > 0; j >>= 1) {
        if (j < kk) {
          const unsigned long long pa = __shfl_xor_sync(0xffffffffu, a, j), pb = __shfl_xor_sync(0xffffffffu, b, j);
          const bool lower = (lane & j) == 0;
          // the lower partner keeps the minimum when the run ascends
          a = ((a < pa) == (lower == up_a)) ? a : pa;
          b = ((b < pb) == (lower == up_b)) ? b : pb;
        }
      }
    }
    keys[ia] = a;
    keys[ib] = b;
  }
}

template <int R>  // R fused steps: strides j, j/2, .., j >> (R-1) (all >= 64) of stage k
__device__ __forceinline__ void sort_high_pass(unsigned long long* keys, const int cap, const int k, const int j) {
  const int jl = j >> (R - 1);
  for (int t = threadIdx.x; t < (cap >> R); t += blockDim.x) {
    const int base = ((t & ~(jl - 1)) << R) | (t & (jl - 1));
    const bool up = (base & k) == 0;
    unsigned long long e[1 << R];
#pragma unroll
    for (int m = 0; m < (1 << R); ++m) e[m] = keys[base + m * jl];
#pragma unroll
    for (int s = R - 1; s >= 0; --s) {
#pragma unroll
      for (int m = 0; m < (1 << R); ++m)
        if (!(m & (1 << s))) sort_cx(e[m], e[m | (1 << s)], up);
    }
#pragma unroll
    for (int m = 0; m < (1 << R); ++m) keys[base + m * jl] = e[m];
  }
}

__device__ __forceinline__ void block_bitonic_sort_u64(unsigned long long* keys, const int cap) {
  if (cap < 64) {  // tiny inputs: plain network
    for (int k = 2; k <= cap; k <<= 1) {
      for (int j = k >> 1; j > 0; j >>= 1) {
        for (int t = threadIdx.x; t < (cap >> 1); t += blockDim.x) {
          const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
          sort_cx(keys[i], keys[i | j], (i & k) == 0);
        }
        __syncthreads();
      }
    }
    return;
  }
  sort_warp_pass(keys, cap, 64, true);
  __syncthreads();
  for (int k = 128; k <= cap; k <<= 1) {
    int j = k >> 1;
    while (j >= 64) {
      if (j >= 256) {
        sort_high_pass<3>(keys, cap, k, j);
        j >>= 3;
      } else if (j >= 128) {
        sort_high_pass<2>(keys, cap, k, j);
        j >>= 2;
      } else {
        sort_high_pass<1>(keys, cap, k, j);
        j >>= 1;
      }
      __syncthreads();
    }
    sort_warp_pass(keys, cap, k, false);
    __syncthreads();
  }
}

}  // namespace cvpce
