// Helpers shared by the window-statistics kernels.
#pragma once

#include <stdint.h>

namespace klm {

// Nearest-rank 95th percentile: 0-based index ceil(0.95 n) - 1 into the
// ascending order, in integers ceil(19 n / 20) - 1 (n >= 1).
__host__ __device__ __forceinline__ int p95_rank(int n) { return (19 * n + 19) / 20 - 1; }

#ifdef __CUDACC__
// Order-preserving map f32 -> u32 (a < b  <=>  key(a) < key(b), NaN excluded).
__device__ __forceinline__ uint32_t f32_key(float x) {
  const uint32_t b = __float_as_uint(x);
  return b ^ ((b & 0x80000000u) ? 0xffffffffu : 0x80000000u);
}
__device__ __forceinline__ float key_f32(uint32_t k) {
  return __uint_as_float((k & 0x80000000u) ? (k ^ 0x80000000u) : ~k);
}
#endif

}  // namespace klm
