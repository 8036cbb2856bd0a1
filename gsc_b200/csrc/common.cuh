// Shared helpers for the gsc_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include "../../include/gsc_b200.h"

namespace gsc {

// Thread-local message of the last failure; exposed through gsc_last_error().
char* err_buf();
int fail(int code, const char* fmt, ...);

#define GSC_CUDA_TRY(expr)                                                                  \
  do {                                                                                      \
    cudaError_t _e = (expr);                                                                \
    if (_e != cudaSuccess)                                                                  \
      return ::gsc::fail(GSC_ECUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                         __FILE__, __LINE__);                                               \
  } while (0)

#define GSC_LAUNCH_CHECK(name)                                                               \
  do {                                                                                       \
    cudaError_t _e = cudaGetLastError();                                                     \
    if (_e != cudaSuccess)                                                                   \
      return ::gsc::fail(GSC_ECUDA, "launch of %s failed: %s", name, cudaGetErrorString(_e)); \
  } while (0)

#define GSC_REQUIRE(cond, ...)                                  \
  do {                                                          \
    if (!(cond)) return ::gsc::fail(GSC_EINVAL, __VA_ARGS__);   \
  } while (0)

constexpr int kNumSMs = 148;  // B200
// SMs that persistent kernels leave free for concurrently running communication kernels (gsc_set_sm_reserve)
int sm_reserve();

static inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline size_t align_up(size_t a, size_t b) { return (a + b - 1) / b * b; }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

}  // namespace gsc
