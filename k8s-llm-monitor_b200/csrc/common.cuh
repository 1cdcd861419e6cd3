// Shared device/host helpers for libklm.so (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/klm.h"

namespace klm {

constexpr int kNumSMs = 148;  // B200: 2 dies x 74 SMs; grids are sized in multiples of this

// Thread-local error text behind klm_last_error().
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);

#define KLM_CUDA(call)                                     \
  do {                                                     \
    cudaError_t klm_e_ = (call);                           \
    if (klm_e_ != cudaSuccess) return ::klm::cuda_fail(klm_e_, #call); \
  } while (0)

#define KLM_LAUNCH_CHECK(name)                             \
  do {                                                     \
    cudaError_t klm_e_ = cudaGetLastError();               \
    if (klm_e_ != cudaSuccess) return ::klm::cuda_fail(klm_e_, name); \
  } while (0)

static inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

static inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }

// "Done once per device" flags for cudaFuncSetAttribute and the like: one bit per device
// ordinal, set and tested atomically (handles of several threads share the launchers).
// Devices beyond 63 simply repeat the (idempotent) set-up on every call.
struct PerDeviceOnce {
  unsigned long long mask = 0ull;
  bool done(int dev) const {
    return dev >= 0 && dev < 64 && ((__atomic_load_n(&mask, __ATOMIC_ACQUIRE) >> dev) & 1ull);
  }
  void set(int dev) {
    if (dev >= 0 && dev < 64) __atomic_fetch_or(&mask, 1ull << dev, __ATOMIC_RELEASE);
  }
};

#ifdef __CUDACC__

// Streaming 128-bit load: read-only path, do not allocate in L1 (each byte of
// a metric window or node column is touched once per launch).
__device__ __forceinline__ int4 ld_stream_v4(const void* p) {
  int4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

__device__ __forceinline__ float4 ld_stream_f4(const float* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}

__device__ __forceinline__ double2 ld_stream_d2(const double* p) {
  double2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0, %1}, [%2];"
               : "=d"(r.x), "=d"(r.y)
               : "l"(p));
  return r;
}

__device__ __forceinline__ int64_t warp_sum_i64(int64_t v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ int warp_sum_i32(int v) { return __reduce_add_sync(0xffffffffu, v); }

#endif  // __CUDACC__

}  // namespace klm
