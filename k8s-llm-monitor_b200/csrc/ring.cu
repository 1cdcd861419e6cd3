// Metric ring kernels: K6 window_ingest, K7 window_stats (block variant).
//
// The reference keeps only the latest snapshot (internal/metrics/manager.go:29,
// :289-290) and has no window statistics, so these kernels have no reference
// arithmetic to match; the definition they implement is DESIGN.md §"window
// statistics", restated on the CPU by the checker under oracle/ - "parity unpinned".
// Producers of samples in the reference: Manager.UpdateUAVReport
// (manager.go:391-449) and the end of Manager.Collect (manager.go:286).
//
//   n     = min(window, seen[s])         valid slots of row s are [0, n)
//   min   = exact f32 minimum, max likewise
//   mean  = (f64 sum of the f32 samples) / n
//   p95   = sorted[ceil(0.95 n) - 1]     (nearest rank; an input element)
#include "common.cuh"
#include "window_stats.cuh"

namespace klm {

// ---------------------------------------------------------------- K6

// 8 B/sample algorithmic (4 read + 4 written); the write is a 4-byte scatter
// with a row pitch of ~14 KB, so DRAM sees one 32 B sector per sample.
__global__ void __launch_bounds__(256)
ring_ingest_kernel(float* __restrict__ ring, uint32_t* __restrict__ seen, int64_t n_series,
                   int32_t window, int32_t pitch, int64_t n,
                   const int64_t* __restrict__ series_ids, const float* __restrict__ values) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) {
    const int64_t s = series_ids ? series_ids[k] : k;
    if (s < 0 || s >= n_series) continue;
    const uint32_t c = seen[s];
    ring[s * (int64_t)pitch + (int64_t)(c % (uint32_t)window)] = values[k];
    if (c != 0xffffffffu) seen[s] = c + 1;
  }
}

// Bulk history load: rows [s0, s1) <- data[(s - s0) * count + t], seen = count.
__global__ void __launch_bounds__(256)
ring_load_kernel(float* __restrict__ ring, uint32_t* __restrict__ seen, int32_t pitch, int64_t s0,
                 int64_t n_rows, int32_t count, const float* __restrict__ data) {
  const int64_t total = n_rows * (int64_t)count;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int64_t r = i / count, t = i - r * count;
    ring[(s0 + r) * (int64_t)pitch + t] = data[i];
  }
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += stride)
    seen[s0 + r] = (uint32_t)count;
}

// ---------------------------------------------------------------- K7 (block)

constexpr int kWsbThreads = 256;

// One CTA per series. The row is staged once in shared memory as
// order-preserving u32 keys; min/max/sum come from the same pass over the
// loaded values; p95 is an exact MSB-first radix select (4 passes of 8 bits)
// over the staged keys. HBM traffic is the algorithmic 4 n + 20 B per series.
// Works for any window that fits shared memory; the warp variant
// (window_stats_warp.cu) is the fast path for windows <= 4096.
__global__ void __launch_bounds__(kWsbThreads)
window_stats_block_kernel(const float* __restrict__ ring, const uint32_t* __restrict__ seen,
                          int32_t window, int32_t pitch, int64_t s0, float* __restrict__ out_min,
                          float* __restrict__ out_max, double* __restrict__ out_mean,
                          float* __restrict__ out_p95) {
  extern __shared__ __align__(16) uint32_t s_keys[];  // [pitch]
  __shared__ unsigned int s_hist[256];
  __shared__ float s_mn[kWsbThreads / 32], s_mx[kWsbThreads / 32];
  __shared__ double s_sum[kWsbThreads / 32];
  __shared__ uint32_t s_prefix;
  __shared__ int s_k;

  const int64_t s = s0 + blockIdx.x;
  const uint32_t sn = seen[s];
  const int n = sn < (uint32_t)window ? (int)sn : window;
  const int64_t o = blockIdx.x;
  if (n == 0) {
    if (threadIdx.x == 0) {
      if (out_min) out_min[o] = 0.f;
      if (out_max) out_max[o] = 0.f;
      if (out_mean) out_mean[o] = 0.0;
      if (out_p95) out_p95[o] = 0.f;
    }
    return;
  }
  const float* row = ring + s * (int64_t)pitch;

  // pass 0: load, stage keys, min/max/sum
  float mn = __int_as_float(0x7f800000), mx = __int_as_float(0xff800000);
  double sum = 0.0;
  const int nv = n >> 2;
  for (int q = threadIdx.x; q < nv; q += kWsbThreads) {
    const float4 v = ld_stream_f4(row + 4 * q);
    const float x[4] = {v.x + 0.f, v.y + 0.f, v.z + 0.f, v.w + 0.f};  // -0.0 -> +0.0
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      mn = fminf(mn, x[c]);
      mx = fmaxf(mx, x[c]);
      sum += (double)x[c];
      s_keys[4 * q + c] = f32_key(x[c]);
    }
  }
  for (int t = 4 * nv + threadIdx.x; t < n; t += kWsbThreads) {
    const float x = row[t] + 0.f;
    mn = fminf(mn, x);
    mx = fmaxf(mx, x);
    sum += (double)x;
    s_keys[t] = f32_key(x);
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, off));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
    sum += __shfl_xor_sync(0xffffffffu, sum, off);
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) {
    s_mn[warp] = mn;
    s_mx[warp] = mx;
    s_sum[warp] = sum;
  }
  if (threadIdx.x == 0) {
    s_prefix = 0;
    s_k = n - p95_rank(n);  // k-th largest, 1-based
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float a = s_mn[0], b = s_mx[0];
    double t = s_sum[0];
    for (int w = 1; w < kWsbThreads / 32; ++w) {
      a = fminf(a, s_mn[w]);
      b = fmaxf(b, s_mx[w]);
      t += s_sum[w];
    }
    if (out_min) out_min[o] = a + 0.f;
    if (out_max) out_max[o] = b + 0.f;
    if (out_mean) out_mean[o] = t / (double)n;
  }
  if (!out_p95) return;

  // radix select, most significant byte first: after each pass s_prefix holds
  // the chosen high bytes and s_k the rank within the surviving keys
  for (int shift = 24; shift >= 0; shift -= 8) {
    s_hist[threadIdx.x] = 0;
    __syncthreads();
    const uint32_t prefix = s_prefix;
    const uint32_t himask = shift == 24 ? 0u : (0xffffffffu << (shift + 8));
    for (int t = threadIdx.x; t < n; t += kWsbThreads) {
      const uint32_t k = s_keys[t];
      if ((k & himask) == prefix) atomicAdd(&s_hist[(k >> shift) & 255u], 1u);
    }
    __syncthreads();
    // thread t owns bin 255 - t; inclusive scan from the top digit down
    const unsigned int mine = s_hist[255 - threadIdx.x];
    unsigned int incl = mine;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const unsigned int u = __shfl_up_sync(0xffffffffu, incl, off);
      if (lane >= off) incl += u;
    }
    __shared__ unsigned int s_wsum[kWsbThreads / 32];
    if (lane == 31) s_wsum[warp] = incl;
    __syncthreads();
    unsigned int before = 0;
    for (int w = 0; w < warp; ++w) before += s_wsum[w];
    incl += before;
    const unsigned int excl = incl - mine;
    const unsigned int k = (unsigned int)s_k;
    __syncthreads();  // everyone has read s_k
    if (excl < k && k <= incl) {
      s_prefix = prefix | ((uint32_t)(255 - threadIdx.x) << shift);
      s_k = (int)(k - excl);
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) out_p95[o] = key_f32(s_prefix) + 0.f;
}

}  // namespace klm

using namespace klm;

extern "C" int klm_dev_ring_ingest(float* ring, uint32_t* seen, int64_t n_series, int32_t window,
                                   int32_t pitch, int64_t n, const int64_t* series_ids,
                                   const float* values, void* stream) {
  if (!ring || !seen || n_series < 0 || window < 1 || pitch < window || n < 0 ||
      (n > 0 && !values) || (!series_ids && n != n_series)) {
    set_error("klm_dev_ring_ingest: bad argument");
    return KLM_EINVAL;
  }
  if (n == 0) return KLM_OK;
  int64_t blocks = ceil_div(n, 256);
  if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
  ring_ingest_kernel<<<(int)blocks, 256, 0, as_stream(stream)>>>(ring, seen, n_series, window,
                                                                 pitch, n, series_ids, values);
  KLM_LAUNCH_CHECK("ring_ingest_kernel");
  return KLM_OK;
}

namespace klm {
int launch_ring_load(float* ring, uint32_t* seen, int32_t pitch, int64_t s0, int64_t n_rows,
                     int32_t count, const float* data, cudaStream_t st) {
  if (n_rows <= 0) return KLM_OK;
  int64_t blocks = ceil_div(n_rows * (int64_t)(count > 0 ? count : 1), 256);
  if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
  ring_load_kernel<<<(int)blocks, 256, 0, st>>>(ring, seen, pitch, s0, n_rows, count, data);
  KLM_LAUNCH_CHECK("ring_load_kernel");
  return KLM_OK;
}

int launch_window_stats_block(const float* ring, const uint32_t* seen, int32_t window,
                              int32_t pitch, int64_t s0, int64_t s1, float* out_min,
                              float* out_max, double* out_mean, float* out_p95, cudaStream_t st) {
  const size_t smem = (size_t)pitch * sizeof(uint32_t);
  if (smem > 200 * 1024) {
    set_error("klm_dev_window_stats: window %d exceeds the shared-memory variant's limit (51200)",
              window);
    return KLM_ELIMIT;
  }
  static PerDeviceOnce attr_set;
  int dev = 0;
  KLM_CUDA(cudaGetDevice(&dev));
  // (the 48 KB default covers static + dynamic shared memory: opt in well below it)
  if (smem > 32 * 1024 && !attr_set.done(dev)) {
    KLM_CUDA(cudaFuncSetAttribute(window_stats_block_kernel,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    attr_set.set(dev);
  }
  // one CTA per series; grid.x is limited to 2^31 - 1, far above any ring here
  const int64_t rows = s1 - s0;
  for (int64_t done = 0; done < rows;) {
    const int64_t chunk = rows - done < 0x40000000 ? rows - done : 0x40000000;
    window_stats_block_kernel<<<(unsigned)chunk, kWsbThreads, smem, st>>>(
        ring, seen, window, pitch, s0 + done, out_min ? out_min + done : nullptr,
        out_max ? out_max + done : nullptr, out_mean ? out_mean + done : nullptr,
        out_p95 ? out_p95 + done : nullptr);
    KLM_LAUNCH_CHECK("window_stats_block_kernel");
    done += chunk;
  }
  return KLM_OK;
}
}  // namespace klm

namespace klm {
int launch_window_stats_warp(const float* ring, const uint32_t* seen, int32_t window,
                             int32_t pitch, int64_t s0, int64_t s1, float* out_min,
                             float* out_max, double* out_mean, float* out_p95, cudaStream_t st);
int launch_window_stats_stream(const float* ring, const uint32_t* seen, int32_t window,
                               int32_t pitch, int64_t s0, int64_t s1, float* out_min,
                               float* out_max, double* out_mean, float* out_p95, float* state,
                               cudaStream_t st);
}

constexpr int kHintedMaxWindow = 8192;  // the streaming kernel's lists hold ~10 % of a window this large

static int window_stats_args_ok(const float* ring, const uint32_t* seen, int32_t window,
                                int32_t pitch, int64_t s0, int64_t s1) {
  if (!ring || !seen || window < 1 || pitch < window || (pitch & 3) || s0 < 0 || s1 < s0 ||
      !aligned16(ring)) {
    set_error("klm_dev_window_stats: bad argument (pitch must be a multiple of 4 floats and "
              "the ring 16-byte aligned)");
    return KLM_EINVAL;
  }
  return KLM_OK;
}

extern "C" int klm_dev_window_stats(const float* ring, const uint32_t* seen, int32_t window,
                                    int32_t pitch, int64_t s0, int64_t s1, float* out_min,
                                    float* out_max, double* out_mean, float* out_p95,
                                    void* stream) {
  const int rc = window_stats_args_ok(ring, seen, window, pitch, s0, s1);
  if (rc != KLM_OK || s1 == s0) return rc;
#ifdef KLM_K7_STREAM_DEFAULT
  if (window <= 8192)
    return launch_window_stats_stream(ring, seen, window, pitch, s0, s1, out_min, out_max, out_mean,
                                      out_p95, nullptr, as_stream(stream));
#endif
  // windows <= 4096: the register-resident warp kernel (fastest measured);
  // larger windows fall through to the shared-memory block kernel
  const int w = launch_window_stats_warp(ring, seen, window, pitch, s0, s1, out_min, out_max,
                                         out_mean, out_p95, as_stream(stream));
  if (w != 1) return w;
  return launch_window_stats_block(ring, seen, window, pitch, s0, s1, out_min, out_max, out_mean,
                                   out_p95, as_stream(stream));
}

extern "C" int klm_dev_window_stats_warp(const float* ring, const uint32_t* seen, int32_t window,
                                         int32_t pitch, int64_t s0, int64_t s1, float* out_min,
                                         float* out_max, double* out_mean, float* out_p95,
                                         void* stream) {
  const int rc = window_stats_args_ok(ring, seen, window, pitch, s0, s1);
  if (rc != KLM_OK || s1 == s0) return rc;
  const int w = launch_window_stats_warp(ring, seen, window, pitch, s0, s1, out_min, out_max,
                                         out_mean, out_p95, as_stream(stream));
  if (w == 1) {
    set_error("klm_dev_window_stats_warp: window %d exceeds the register-resident variant's 4096", window);
    return KLM_ELIMIT;
  }
  return w;
}

extern "C" int klm_dev_window_stats_stream(const float* ring, const uint32_t* seen, int32_t window,
                                           int32_t pitch, int64_t s0, int64_t s1, float* out_min,
                                           float* out_max, double* out_mean, float* out_p95,
                                           void* stream) {
  const int rc = window_stats_args_ok(ring, seen, window, pitch, s0, s1);
  if (rc != KLM_OK || s1 == s0) return rc;
  return launch_window_stats_stream(ring, seen, window, pitch, s0, s1, out_min, out_max, out_mean,
                                    out_p95, nullptr, as_stream(stream));
}

extern "C" int klm_dev_window_stats_hinted(const float* ring, const uint32_t* seen, int32_t window,
                                           int32_t pitch, int64_t s0, int64_t s1, float* out_min,
                                           float* out_max, double* out_mean, float* out_p95,
                                           float* state, void* stream) {
  const int rc = window_stats_args_ok(ring, seen, window, pitch, s0, s1);
  if (rc != KLM_OK || s1 == s0) return rc;
  if (!state || (reinterpret_cast<uintptr_t>(state) & 7u)) {
    set_error("klm_dev_window_stats_hinted: state must be a non-null, 8-byte aligned array of 2 floats per series");
    return KLM_EINVAL;
  }
  if (window > kHintedMaxWindow) {
    set_error("klm_dev_window_stats_hinted: window %d exceeds %d (use klm_dev_window_stats)", window, kHintedMaxWindow);
    return KLM_ELIMIT;
  }
  return launch_window_stats_stream(ring, seen, window, pitch, s0, s1, out_min, out_max, out_mean,
                                    out_p95, state, as_stream(stream));
}

extern "C" int klm_dev_window_stats_block(const float* ring, const uint32_t* seen, int32_t window,
                                          int32_t pitch, int64_t s0, int64_t s1, float* out_min,
                                          float* out_max, double* out_mean, float* out_p95,
                                          void* stream) {
  const int rc = window_stats_args_ok(ring, seen, window, pitch, s0, s1);
  if (rc != KLM_OK || s1 == s0) return rc;
  return launch_window_stats_block(ring, seen, window, pitch, s0, s1, out_min, out_max, out_mean,
                                   out_p95, as_stream(stream));
}
