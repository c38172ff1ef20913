// Shared helpers for libndivplan (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <atomic>
#include "../../include/ndivplan.h"

namespace ndp {

void set_error(const char* fmt, ...);
extern std::atomic<unsigned long long> g_launches;
inline void count_launch(int n = 1) { g_launches.fetch_add((unsigned long long)n, std::memory_order_relaxed); }

#define NDP_CHECK_ARG(cond, ...)                 \
  do {                                           \
    if (!(cond)) {                               \
      ndp::set_error(__VA_ARGS__);               \
      return NDP_ERR_ARG;                        \
    }                                            \
  } while (0)

#define NDP_CUDA(call)                                                              \
  do {                                                                              \
    cudaError_t e__ = (call);                                                       \
    if (e__ != cudaSuccess) {                                                       \
      ndp::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__),       \
                     __FILE__, __LINE__);                                           \
      return NDP_ERR_CUDA;                                                          \
    }                                                                               \
  } while (0)

#define NDP_LAUNCH_CHECK()                                                          \
  do {                                                                              \
    cudaError_t e__ = cudaGetLastError();                                           \
    if (e__ != cudaSuccess) {                                                       \
      ndp::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(e__),   \
                     __FILE__, __LINE__);                                           \
      return NDP_ERR_CUDA;                                                          \
    }                                                                               \
    ndp::count_launch();                                                            \
  } while (0)

int device_check();   // 0 if current device is sm_100
int num_sms();

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is per device: returns true the first time `mask` sees the current
// device (bit = device ordinal, 64 devices), so a process driving several GPUs sets the attribute on each of them
inline bool first_use_on_device(std::atomic<unsigned long long>& mask) {
  int dev = 0;
  cudaGetDevice(&dev);
  const unsigned long long bit = 1ull << (dev & 63);
  return (mask.fetch_or(bit) & bit) == 0;
}

// fp32 -> order-preserving uint32 (ascending). NaN -> 0xFFFFFFFF (sorts last).
__host__ __device__ inline uint32_t ordered_key(float c) {
  if (c != c) return 0xFFFFFFFFu;
#ifdef __CUDA_ARCH__
  uint32_t b = __float_as_uint(c);
#else
  union { float f; uint32_t u; } cv; cv.f = c; uint32_t b = cv.u;
#endif
  return (b >> 31) ? ~b : (b | 0x80000000u);
}
__host__ __device__ inline float key_to_float(uint32_t k) {
  uint32_t b = (k >> 31) ? (k & 0x7FFFFFFFu) : ~k;
#ifdef __CUDA_ARCH__
  return __uint_as_float(b);
#else
  union { float f; uint32_t u; } cv; cv.u = b; return cv.f;
#endif
}

static constexpr unsigned long long kInvalidPacked = 0xFFFFFFFF00000000ull;

}  // namespace ndp
