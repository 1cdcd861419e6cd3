// Element-wise leftovers of the metrics path: small, bandwidth-bound
// kernels with one thread per node / pod / probe.
//   NodeMetrics.GetAvailableResources   pkg/metrics/types.go:151-156
//   NetworkMetrics.GetQuality           pkg/metrics/types.go:187-199
//   pod restart count                   internal/metrics/sources/pod_metrics.go:174-177
//   node health from conditions         internal/metrics/sources/node_metrics.go:142-164
#include "common.cuh"

namespace klm {

// float64(cap - use) / 1000.0 and / 1024 / 1024 / 1024, left to right as Go evaluates them
// (every divisor but 1000 is a power of two, so only the first line ever rounds).
__global__ void __launch_bounds__(256)
node_available_kernel(int64_t n, const int64_t* __restrict__ cpu_cap, const int64_t* __restrict__ cpu_use,
                      const int64_t* __restrict__ mem_cap, const int64_t* __restrict__ mem_use,
                      const int64_t* __restrict__ disk_cap, const int64_t* __restrict__ disk_use,
                      double* __restrict__ cpu_cores, double* __restrict__ mem_gb,
                      double* __restrict__ disk_gb) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    if (cpu_cores) cpu_cores[i] = __ddiv_rn((double)(cpu_cap[i] - cpu_use[i]), 1000.0);  // :152
    if (mem_gb)
      mem_gb[i] = __ddiv_rn(__ddiv_rn(__ddiv_rn((double)(mem_cap[i] - mem_use[i]), 1024.0), 1024.0), 1024.0);  // :153
    if (disk_gb) {
      const int64_t du = disk_use ? disk_use[i] : 0;
      const int64_t dc = disk_cap ? disk_cap[i] : 0;
      disk_gb[i] = __ddiv_rn(__ddiv_rn(__ddiv_rn((double)(dc - du), 1024.0), 1024.0), 1024.0);  // :154
    }
  }
}

// 0 disconnected, 1 excellent (< 10 ms), 2 good (< 50), 3 fair (< 100), 4 poor; types.go:188-198
__global__ void __launch_bounds__(256)
network_quality_kernel(int64_t n, const uint8_t* __restrict__ connected, const double* __restrict__ rtt,
                       uint8_t* __restrict__ quality) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const double r = rtt[i];
    uint8_t q = 4;
    if (!connected[i]) q = 0;
    else if (r < 10) q = 1;
    else if (r < 50) q = 2;
    else if (r < 100) q = 3;
    quality[i] = q;
  }
}

// restarts += containerStatus.RestartCount over a pod's container statuses (a CSR segment);
// int32 arithmetic that wraps like Go's
__global__ void __launch_bounds__(256)
pod_restarts_kernel(int64_t n_pods, const int32_t* __restrict__ st_off,
                    const int32_t* __restrict__ restart_count, int32_t* __restrict__ restarts) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n_pods; p += stride) {
    uint32_t acc = 0;
    const int32_t e = st_off[p + 1];
    for (int32_t c = st_off[p]; c < e; ++c) acc += (uint32_t)restart_count[c];
    restarts[p] = (int32_t)acc;
  }
}

// healthy unless Ready is not True, or one of the four pressure / unavailability conditions is
// True (node_metrics.go:145-163). failing bit t is set for every condition type t that made the
// node unhealthy (the host builds the Conditions strings from it).
__global__ void __launch_bounds__(256)
node_health_kernel(int64_t n, const int32_t* __restrict__ cond_off, const uint8_t* __restrict__ cond_type,
                   const uint8_t* __restrict__ cond_status, uint8_t* __restrict__ healthy,
                   uint8_t* __restrict__ failing) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    uint32_t bad = 0;
    const int32_t e = cond_off[i + 1];
    for (int32_t c = cond_off[i]; c < e; ++c) {
      const uint32_t t = cond_type[c], st = cond_status[c];
      if (t == KLM_COND_READY) {
        if (st != KLM_COND_TRUE) bad |= 1u << KLM_COND_READY;  // :148
      } else if (st == KLM_COND_TRUE && t >= KLM_COND_MEMORY_PRESSURE && t <= KLM_COND_NETWORK_UNAVAILABLE) {
        bad |= 1u << t;                                        // :154-161
      }
    }
    if (healthy) healthy[i] = bad == 0;
    if (failing) failing[i] = (uint8_t)bad;
  }
}

static int grid1d(int64_t n) {
  int64_t b = ceil_div(n, 256);
  if (b > kNumSMs * 8) b = kNumSMs * 8;
  return (int)(b < 1 ? 1 : b);
}

}  // namespace klm

using namespace klm;

extern "C" int klm_dev_node_available(int64_t n, const int64_t* cpu_cap, const int64_t* cpu_use,
                                      const int64_t* mem_cap, const int64_t* mem_use,
                                      const int64_t* disk_cap, const int64_t* disk_use, double* cpu_cores,
                                      double* mem_gb, double* disk_gb, void* stream) {
  if (n < 0 || (n > 0 && ((cpu_cores && (!cpu_cap || !cpu_use)) || (mem_gb && (!mem_cap || !mem_use))))) {
    set_error("klm_dev_node_available: bad argument");
    return KLM_EINVAL;
  }
  if (n == 0) return KLM_OK;
  node_available_kernel<<<grid1d(n), 256, 0, as_stream(stream)>>>(n, cpu_cap, cpu_use, mem_cap, mem_use, disk_cap,
                                                                  disk_use, cpu_cores, mem_gb, disk_gb);
  KLM_LAUNCH_CHECK("node_available_kernel");
  return KLM_OK;
}

extern "C" int klm_dev_network_quality(int64_t n, const uint8_t* connected, const double* rtt,
                                       uint8_t* quality, void* stream) {
  if (n < 0 || (n > 0 && (!connected || !rtt || !quality))) {
    set_error("klm_dev_network_quality: bad argument");
    return KLM_EINVAL;
  }
  if (n == 0) return KLM_OK;
  network_quality_kernel<<<grid1d(n), 256, 0, as_stream(stream)>>>(n, connected, rtt, quality);
  KLM_LAUNCH_CHECK("network_quality_kernel");
  return KLM_OK;
}

extern "C" int klm_dev_pod_restarts(int64_t n_pods, const int32_t* st_off, const int32_t* restart_count,
                                    int32_t* restarts, void* stream) {
  if (n_pods < 0 || (n_pods > 0 && (!st_off || !restarts))) {
    set_error("klm_dev_pod_restarts: bad argument");
    return KLM_EINVAL;
  }
  if (n_pods == 0) return KLM_OK;
  pod_restarts_kernel<<<grid1d(n_pods), 256, 0, as_stream(stream)>>>(n_pods, st_off, restart_count, restarts);
  KLM_LAUNCH_CHECK("pod_restarts_kernel");
  return KLM_OK;
}

extern "C" int klm_dev_node_health(int64_t n, const int32_t* cond_off, const uint8_t* cond_type,
                                   const uint8_t* cond_status, uint8_t* healthy, uint8_t* failing,
                                   void* stream) {
  if (n < 0 || (n > 0 && !cond_off)) {
    set_error("klm_dev_node_health: bad argument");
    return KLM_EINVAL;
  }
  if (n == 0) return KLM_OK;
  node_health_kernel<<<grid1d(n), 256, 0, as_stream(stream)>>>(n, cond_off, cond_type, cond_status, healthy, failing);
  KLM_LAUNCH_CHECK("node_health_kernel");
  return KLM_OK;
}

// ------------------------------------------------- window-derived node features
//
// NOT IN THE REFERENCE: controller.go:202-203 parses last_update and drops it, manager.go:296-313
// records heartbeats nobody reads, and no score looks at history. This opt-in step connects the
// window statistics (K7) to the scheduler's inputs (K1 / K2): per UAV node i, from the node's ring
// rows (series i * 64 + channel, host/klm_wire.h channel map),
//   stale      heartbeat_age[i] > max_age_s                     -> node ineligible (status bit cleared)
//   too hot    p95 of battery temperature (channel 24) > temp_limit -> node ineligible
//   projected  battery[i] - horizon_s * drain, drain = (max - min of remaining_percent, channel 21)
//              / span_s, clamped to [0, battery[i]]             -> the battery the scheduler scores
// Parity: unpinned by construction (no reference semantics); the oracle restates this definition.
namespace klm {
__global__ void __launch_bounds__(256)
node_features_kernel(int64_t n, const double* __restrict__ battery, const uint8_t* __restrict__ flags,
                     const double* __restrict__ heartbeat_age, const float* __restrict__ win_min,
                     const float* __restrict__ win_max, const float* __restrict__ win_p95,
                     klm_feature_params prm, double* __restrict__ battery_out, uint8_t* __restrict__ flags_out) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    uint8_t f = flags[i];
    double b = battery[i];
    if (heartbeat_age && prm.max_age_s > 0 && heartbeat_age[i] > prm.max_age_s) f &= (uint8_t)~KLM_NODE_STATUS_OK;
    if (win_p95 && prm.temp_limit > 0 && (double)win_p95[i * 64 + 24] > prm.temp_limit) f &= (uint8_t)~KLM_NODE_STATUS_OK;
    if (win_min && win_max && prm.horizon_s > 0 && prm.span_s > 0) {
      const double drop = (double)win_max[i * 64 + 21] - (double)win_min[i * 64 + 21];  // exact in f64
      double proj = b - __dmul_rn(__ddiv_rn(drop, prm.span_s), prm.horizon_s);
      if (proj < 0) proj = 0;
      if (proj < b) b = proj;
    }
    battery_out[i] = b;
    flags_out[i] = f;
  }
}
}  // namespace klm

extern "C" int klm_dev_node_features(int64_t n, const double* battery, const uint8_t* flags,
                                     const double* heartbeat_age, const float* win_min, const float* win_max,
                                     const float* win_p95, const klm_feature_params* prm, double* battery_out,
                                     uint8_t* flags_out, void* stream) {
  if (n < 0 || !prm || (n > 0 && (!battery || !flags || !battery_out || !flags_out))) {
    set_error("klm_dev_node_features: bad argument");
    return KLM_EINVAL;
  }
  if (n == 0) return KLM_OK;
  node_features_kernel<<<grid1d(n), 256, 0, as_stream(stream)>>>(n, battery, flags, heartbeat_age, win_min, win_max,
                                                                 win_p95, *prm, battery_out, flags_out);
  KLM_LAUNCH_CHECK("node_features_kernel");
  return KLM_OK;
}
