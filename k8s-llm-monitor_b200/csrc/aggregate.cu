// Snapshot aggregation kernels: K4 cluster_aggregate, K5 node/pod rates and
// predicates, fleet scans.
//
// Reference semantics (Sabre94/k8s-llm-monitor):
//   internal/metrics/manager.go:493-565              calculateClusterMetrics
//   internal/metrics/sources/node_metrics.go:116-140 node rates, disk estimate
//   internal/metrics/sources/pod_metrics.go:104-171  pod sums over containers, rates
//   pkg/metrics/types.go:159-184                     IsUnderPressure, utilisation, IsOverLimit
//   internal/metrics/sources/uav_metrics.go:197-233  fleet scans
// Integer sums are int64 and order-independent, so they are bit-exact; the two
// cluster rates are one f64 divide and one multiply each, correctly rounded
// like Go's.
#include "common.cuh"

namespace klm {

// ---------------------------------------------------------------- K4

constexpr int kAggThreads = 256;
constexpr int kAggMaxBlocks = 1024;
constexpr int kAggSums = 10;  // the ten int64 members that open klm_cluster_out
// scratch: [0, 256) ticket (u32, self-resetting); then i64[kAggMaxBlocks][10].
constexpr int64_t kAggScratchBytes = 256 + (int64_t)kAggMaxBlocks * kAggSums * 8;

enum { S_HEALTHY = 0, S_RUNNING, S_TCPU, S_UCPU, S_TMEM, S_UMEM, S_GPUS, S_AVAIL, S_NUM };

__device__ __forceinline__ void cluster_finish_dev(const int64_t* s, klm_cluster_out* o) {
  klm_cluster_out r;
  r.total_nodes = s[0];
  r.healthy_nodes = s[1];
  r.total_pods = s[2];
  r.running_pods = s[3];
  r.total_cpu = s[4];
  r.used_cpu = s[5];
  r.total_memory = s[6];
  r.used_memory = s[7];
  r.total_gpus = s[8];
  r.available_gpus = s[9];
  r.cpu_usage_rate = 0.0;
  r.memory_usage_rate = 0.0;
  // manager.go:538-543; __ddiv_rn/__dmul_rn keep the two roundings separate
  if (r.total_cpu > 0)
    r.cpu_usage_rate = __dmul_rn(__ddiv_rn((double)r.used_cpu, (double)r.total_cpu), 100.0);
  if (r.total_memory > 0)
    r.memory_usage_rate =
        __dmul_rn(__ddiv_rn((double)r.used_memory, (double)r.total_memory), 100.0);
  uint32_t issues = 0;
  if (r.healthy_nodes < r.total_nodes) issues |= KLM_ISSUE_UNHEALTHY_NODES;  // :547
  if (r.cpu_usage_rate > 80) issues |= KLM_ISSUE_HIGH_CPU;                   // :550
  if (r.memory_usage_rate > 80) issues |= KLM_ISSUE_HIGH_MEMORY;             // :553
  r.issues = issues;
  if (issues == 0)
    r.health_status = KLM_HEALTH_HEALTHY;
  else if (r.cpu_usage_rate > 90 || r.memory_usage_rate > 90 ||
           r.healthy_nodes < r.total_nodes / 2)  // integer division as in Go, :560
    r.health_status = KLM_HEALTH_CRITICAL;
  else
    r.health_status = KLM_HEALTH_WARNING;
  *o = r;
}

struct NodeColsDev {
  const int64_t* cpu_cap;
  const int64_t* cpu_use;
  const int64_t* mem_cap;
  const int64_t* mem_use;
  const int32_t* gpu_count;
  const uint8_t* healthy;
  const double* gpu_usage;
  int64_t n_gpu_usage;
};

// 37 B/node + 8 B/GPU entry + 1 B/pod, each read once with 128-bit loads when
// the columns are 16-byte aligned. Per-thread int64 partials -> warp shuffle ->
// block partial row in scratch; the last block (ticket) adds the rows and
// either writes the ten sums (sharded use) or the finished klm_cluster_out.
__global__ void __launch_bounds__(kAggThreads)
cluster_aggregate_kernel(NodeColsDev c, int64_t n, const uint8_t* __restrict__ pod_running,
                         int64_t n_pods, int vec_nodes, int vec_pods, int vec_gpu,
                         unsigned int* __restrict__ ticket, int64_t* __restrict__ partials,
                         int64_t* __restrict__ sums_out, klm_cluster_out* __restrict__ out) {
  int64_t acc[S_NUM];
#pragma unroll
  for (int k = 0; k < S_NUM; ++k) acc[k] = 0;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;

  // nodes, 8 per trip
  int64_t done = 0;
  if (vec_nodes) {
    const int64_t nchunks = n / 8;
    for (int64_t ch = tid; ch < nchunks; ch += nthreads) {
      const int64_t i0 = ch * 8;
      int4 v[16];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        v[k] = ld_stream_v4(c.cpu_cap + i0 + 2 * k);
        v[4 + k] = ld_stream_v4(c.cpu_use + i0 + 2 * k);
        v[8 + k] = ld_stream_v4(c.mem_cap + i0 + 2 * k);
        v[12 + k] = ld_stream_v4(c.mem_use + i0 + 2 * k);
      }
      const int4 g0 = ld_stream_v4(c.gpu_count + i0);
      const int4 g1 = ld_stream_v4(c.gpu_count + i0 + 4);
      const uint2 h = *reinterpret_cast<const uint2*>(c.healthy + i0);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
#pragma unroll
        for (int col = 0; col < 4; ++col) {
          const int4 w = v[col * 4 + k];
          const int64_t a = (int64_t)(((uint64_t)(uint32_t)w.y << 32) | (uint32_t)w.x);
          const int64_t b = (int64_t)(((uint64_t)(uint32_t)w.w << 32) | (uint32_t)w.z);
          acc[S_TCPU + col] += a + b;  // S_TCPU, S_UCPU, S_TMEM, S_UMEM are consecutive
        }
      }
      acc[S_GPUS] += (int64_t)g0.x + g0.y + g0.z + g0.w + g1.x + g1.y + g1.z + g1.w;
      // healthy is "any non-zero byte": count the non-zero bytes of the 8
      unsigned int cnt = 0;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        cnt += ((h.x >> (8 * k)) & 0xffu) != 0;
        cnt += ((h.y >> (8 * k)) & 0xffu) != 0;
      }
      acc[S_HEALTHY] += cnt;
    }
    done = nchunks * 8;
  }
  for (int64_t i = done + tid; i < n; i += nthreads) {
    acc[S_TCPU] += c.cpu_cap[i];
    acc[S_UCPU] += c.cpu_use[i];
    acc[S_TMEM] += c.mem_cap[i];
    acc[S_UMEM] += c.mem_use[i];
    acc[S_GPUS] += c.gpu_count[i];
    acc[S_HEALTHY] += c.healthy[i] != 0;
  }

  // pods, 16 per trip
  done = 0;
  if (vec_pods) {
    const int64_t nchunks = n_pods / 16;
    for (int64_t ch = tid; ch < nchunks; ch += nthreads) {
      const int4 w = ld_stream_v4(pod_running + ch * 16);
      const unsigned int u[4] = {(unsigned)w.x, (unsigned)w.y, (unsigned)w.z, (unsigned)w.w};
      unsigned int cnt = 0;
#pragma unroll
      for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int k = 0; k < 4; ++k) cnt += ((u[q] >> (8 * k)) & 0xffu) != 0;
      acc[S_RUNNING] += cnt;
    }
    done = nchunks * 16;
  }
  for (int64_t p = done + tid; p < n_pods; p += nthreads) acc[S_RUNNING] += pod_running[p] != 0;

  // GPU usage entries, 2 per trip: available = usage < 50.0 (manager.go:530-534)
  done = 0;
  if (vec_gpu) {
    const int64_t nchunks = c.n_gpu_usage / 2;
    for (int64_t ch = tid; ch < nchunks; ch += nthreads) {
      const double2 u = ld_stream_d2(c.gpu_usage + ch * 2);
      acc[S_AVAIL] += (u.x < 50.0) + (u.y < 50.0);
    }
    done = nchunks * 2;
  }
  for (int64_t k = done + tid; k < c.n_gpu_usage; k += nthreads)
    acc[S_AVAIL] += c.gpu_usage[k] < 50.0;

  // block reduce
  __shared__ int64_t s_part[kAggThreads / 32][S_NUM];
  __shared__ bool s_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < S_NUM; ++k) {
    const int64_t v = warp_sum_i64(acc[k]);
    if (lane == 0) s_part[warp][k] = v;
  }
  __syncthreads();
  if (threadIdx.x < S_NUM) {
    int64_t v = 0;
#pragma unroll
    for (int w = 0; w < kAggThreads / 32; ++w) v += s_part[w][threadIdx.x];
    partials[(int64_t)blockIdx.x * kAggSums + threadIdx.x] = v;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int t = atomicAdd(ticket, 1u);
    s_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();

  // last block: add the rows (S_NUM columns, strided over threads)
  __shared__ int64_t s_tot[S_NUM];
  if (threadIdx.x < S_NUM) s_tot[threadIdx.x] = 0;
  __syncthreads();
  {
    const int col = threadIdx.x % 8;  // S_NUM == 8
    int64_t v = 0;
    for (int b = threadIdx.x / 8; b < (int)gridDim.x; b += kAggThreads / 8)
      v += __ldcg(partials + (int64_t)b * kAggSums + col);
    atomicAdd(reinterpret_cast<unsigned long long*>(&s_tot[col]), (unsigned long long)v);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int64_t s[kAggSums];
    s[0] = n;                 // TotalNodes, :497
    s[1] = s_tot[S_HEALTHY];
    s[2] = n_pods;            // TotalPods, :506
    s[3] = s_tot[S_RUNNING];
    s[4] = s_tot[S_TCPU];
    s[5] = s_tot[S_UCPU];
    s[6] = s_tot[S_TMEM];
    s[7] = s_tot[S_UMEM];
    s[8] = s_tot[S_GPUS];
    s[9] = s_tot[S_AVAIL];
    if (sums_out)
      for (int k = 0; k < kAggSums; ++k) sums_out[k] = s[k];
    if (out) cluster_finish_dev(s, out);
    *ticket = 0;
  }
}
static_assert(S_NUM == 8, "column mapping above assumes 8 partial sums");

__global__ void cluster_finish_kernel(const int64_t* __restrict__ sums, klm_cluster_out* out) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    int64_t s[kAggSums];
    for (int k = 0; k < kAggSums; ++k) s[k] = sums[k];
    cluster_finish_dev(s, out);
  }
}

// ---------------------------------------------------------------- K5

__device__ __forceinline__ double rate_pct(int64_t used, int64_t total) {
  // float64(used) / float64(total) * 100.0 when total > 0, else 0
  return total > 0 ? __dmul_rn(__ddiv_rn((double)used, (double)total), 100.0) : 0.0;
}

__global__ void __launch_bounds__(256)
node_rates_kernel(int64_t n, const int64_t* __restrict__ cpu_cap, const int64_t* __restrict__ cpu_use,
                  const int64_t* __restrict__ mem_cap, const int64_t* __restrict__ mem_use,
                  const int64_t* __restrict__ disk_cap, const int64_t* __restrict__ disk_alloc,
                  double* __restrict__ cpu_rate, double* __restrict__ mem_rate,
                  int64_t* __restrict__ disk_usage, double* __restrict__ disk_rate,
                  uint8_t* __restrict__ under_pressure) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const int64_t dc = disk_cap ? disk_cap[i] : 0;
    int64_t du = 0;
    if (disk_alloc) {  // node_metrics.go:117-124
      du = dc - disk_alloc[i];
      if (du < 0) du = 0;
    }
    const double cr = rate_pct(cpu_use[i], cpu_cap[i]);  // :127-130
    const double mr = rate_pct(mem_use[i], mem_cap[i]);  // :132-135
    const double dr = rate_pct(du, dc);                  // :137-140
    if (cpu_rate) cpu_rate[i] = cr;
    if (mem_rate) mem_rate[i] = mr;
    if (disk_usage) disk_usage[i] = du;
    if (disk_rate) disk_rate[i] = dr;
    if (under_pressure) under_pressure[i] = (cr > 80.0 || mr > 80.0 || dr > 90.0);  // types.go:161
  }
}

// Two nodes per thread and trip, every column read and written 16 bytes at a time (all columns
// 16-byte aligned, under_pressure 2-byte aligned): twelve 128-bit loads in flight per thread
// instead of six 64-bit ones, and a grid sized so that every thread makes the same number of trips.
__global__ void __launch_bounds__(256)
node_rates_pair_kernel(int64_t n, const int64_t* __restrict__ cpu_cap, const int64_t* __restrict__ cpu_use,
                       const int64_t* __restrict__ mem_cap, const int64_t* __restrict__ mem_use,
                       const int64_t* __restrict__ disk_cap, const int64_t* __restrict__ disk_alloc,
                       double* __restrict__ cpu_rate, double* __restrict__ mem_rate,
                       int64_t* __restrict__ disk_usage, double* __restrict__ disk_rate,
                       uint8_t* __restrict__ under_pressure) {
  const int64_t npairs = n >> 1;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  auto ld2 = [](const int64_t* a, int64_t p) { return __ldg(reinterpret_cast<const longlong2*>(a) + p); };
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < npairs; p += stride) {
    const longlong2 cc = ld2(cpu_cap, p), cu = ld2(cpu_use, p), mc = ld2(mem_cap, p), mu = ld2(mem_use, p);
    const longlong2 dc = disk_cap ? ld2(disk_cap, p) : make_longlong2(0, 0);
    longlong2 du = make_longlong2(0, 0);
    if (disk_alloc) {  // node_metrics.go:117-124
      const longlong2 da = ld2(disk_alloc, p);
      du.x = dc.x - da.x > 0 ? dc.x - da.x : 0;
      du.y = dc.y - da.y > 0 ? dc.y - da.y : 0;
    }
    const double2 cr = make_double2(rate_pct(cu.x, cc.x), rate_pct(cu.y, cc.y));  // :127-130
    const double2 mr = make_double2(rate_pct(mu.x, mc.x), rate_pct(mu.y, mc.y));  // :132-135
    const double2 dr = make_double2(rate_pct(du.x, dc.x), rate_pct(du.y, dc.y));  // :137-140
    if (cpu_rate) reinterpret_cast<double2*>(cpu_rate)[p] = cr;
    if (mem_rate) reinterpret_cast<double2*>(mem_rate)[p] = mr;
    if (disk_usage) reinterpret_cast<longlong2*>(disk_usage)[p] = du;
    if (disk_rate) reinterpret_cast<double2*>(disk_rate)[p] = dr;
    if (under_pressure) {  // types.go:161
      const unsigned short u = (unsigned short)((cr.x > 80.0 || mr.x > 80.0 || dr.x > 90.0) ? 1 : 0) |
                               (unsigned short)((cr.y > 80.0 || mr.y > 80.0 || dr.y > 90.0) ? 0x100 : 0);
      reinterpret_cast<unsigned short*>(under_pressure)[p] = u;
    }
  }
  if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {  // the odd last node
    const int64_t i = n - 1;
    const int64_t dc = disk_cap ? disk_cap[i] : 0;
    int64_t du = 0;
    if (disk_alloc) {
      du = dc - disk_alloc[i];
      if (du < 0) du = 0;
    }
    const double cr = rate_pct(cpu_use[i], cpu_cap[i]), mr = rate_pct(mem_use[i], mem_cap[i]), dr = rate_pct(du, dc);
    if (cpu_rate) cpu_rate[i] = cr;
    if (mem_rate) mem_rate[i] = mr;
    if (disk_usage) disk_usage[i] = du;
    if (disk_rate) disk_rate[i] = dr;
    if (under_pressure) under_pressure[i] = (cr > 80.0 || mr > 80.0 || dr > 90.0);
  }
}

// One thread per pod; its containers are a CSR segment (pod_metrics.go:105-127).
__global__ void __launch_bounds__(256)
pod_rates_kernel(int64_t n_pods, const int32_t* __restrict__ ctr_off,
                 const int64_t* __restrict__ c_cpu_use, const int64_t* __restrict__ c_mem_use,
                 const int64_t* __restrict__ c_cpu_req, const int64_t* __restrict__ c_cpu_lim,
                 const int64_t* __restrict__ c_mem_req, const int64_t* __restrict__ c_mem_lim,
                 int64_t* __restrict__ pod_totals, double* __restrict__ cpu_rate,
                 double* __restrict__ mem_rate, double* __restrict__ cpu_util,
                 double* __restrict__ mem_util, uint8_t* __restrict__ over_limit) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n_pods; p += stride) {
    int64_t cu = 0, mu = 0, cq = 0, cl = 0, mq = 0, ml = 0;
    const int32_t e = ctr_off[p + 1];
    for (int32_t c = ctr_off[p]; c < e; ++c) {
      cu += c_cpu_use[c];
      mu += c_mem_use[c];
      cq += c_cpu_req[c];
      cl += c_cpu_lim[c];
      mq += c_mem_req[c];
      ml += c_mem_lim[c];
    }
    if (pod_totals) {
      pod_totals[0 * n_pods + p] = cu;
      pod_totals[1 * n_pods + p] = mu;
      pod_totals[2 * n_pods + p] = cq;
      pod_totals[3 * n_pods + p] = cl;
      pod_totals[4 * n_pods + p] = mq;
      pod_totals[5 * n_pods + p] = ml;
    }
    if (cpu_rate) cpu_rate[p] = rate_pct(cu, cl);  // pod_metrics.go:162-165
    if (mem_rate) mem_rate[p] = rate_pct(mu, ml);  // :167-170
    if (cpu_util) cpu_util[p] = rate_pct(cu, cq);  // types.go:166-168
    if (mem_util) mem_util[p] = rate_pct(mu, mq);  // types.go:169-171
    if (over_limit) {
      // types.go:177/180: usage >= int64(float64(limit) * 0.9), truncation toward zero
      bool ov = false;
      if (cl > 0 && cu >= (int64_t)__dmul_rn((double)cl, 0.9)) ov = true;
      if (ml > 0 && mu >= (int64_t)__dmul_rn((double)ml, 0.9)) ov = true;
      over_limit[p] = ov;
    }
  }
}

// ---------------------------------------------------------------- fleet scans

constexpr int kScanThreads = 256;
constexpr int kScanPerThread = 4;
constexpr int kScanTile = kScanThreads * kScanPerThread;

// Pass 1: per-tile count of low-battery UAVs; healthy count by atomic.
__global__ void __launch_bounds__(kScanThreads)
fleet_count_kernel(int64_t n, const uint8_t* __restrict__ status_ok,
                   const double* __restrict__ pct, double threshold, int64_t* counters,
                   int32_t* __restrict__ tile_count) {
  const int64_t base = (int64_t)blockIdx.x * kScanTile;
  int low = 0, ok = 0;
#pragma unroll
  for (int k = 0; k < kScanPerThread; ++k) {
    const int64_t i = base + (int64_t)k * kScanThreads + threadIdx.x;
    if (i < n) {
      low += pct[i] < threshold;   // uav_metrics.go:226
      ok += status_ok[i] != 0;     // uav_metrics.go:207
    }
  }
  low = warp_sum_i32(low);
  ok = warp_sum_i32(ok);
  __shared__ int s_low[kScanThreads / 32], s_ok[kScanThreads / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) {
    s_low[warp] = low;
    s_ok[warp] = ok;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int tl = 0, to = 0;
    for (int w = 0; w < kScanThreads / 32; ++w) {
      tl += s_low[w];
      to += s_ok[w];
    }
    tile_count[blockIdx.x] = tl;
    if (to) atomicAdd(reinterpret_cast<unsigned long long*>(&counters[0]), (unsigned long long)to);
  }
}

// Pass 2: one block turns tile counts into exclusive offsets (in place).
__global__ void __launch_bounds__(1024)
fleet_offsets_kernel(int32_t* __restrict__ tile_count, int n_tiles, int64_t* counters) {
  __shared__ int s_warp[32];
  __shared__ int s_carry;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int t0 = 0; t0 < n_tiles; t0 += 1024) {
    const int i = t0 + threadIdx.x;
    const int v = i < n_tiles ? tile_count[i] : 0;
    int incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int u = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += u;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      int w = s_warp[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int u = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += u;
      }
      s_warp[lane] = w;  // inclusive over warps
    }
    __syncthreads();
    const int carry = s_carry;
    const int before = carry + (warp ? s_warp[warp - 1] : 0) + incl - v;
    if (i < n_tiles) tile_count[i] = before;
    __syncthreads();
    if (threadIdx.x == 1023) s_carry = carry + s_warp[31];
    __syncthreads();
  }
  if (threadIdx.x == 0) counters[1] = s_carry;
}

// Pass 3: write the indices in ascending order. Element e of a tile is
// i = base + e with thread t owning e = t*4 .. t*4+3, so ranks follow i.
__global__ void __launch_bounds__(kScanThreads)
fleet_write_kernel(int64_t n, const double* __restrict__ pct, double threshold,
                   const int32_t* __restrict__ tile_off, int32_t* __restrict__ low_idx) {
  const int64_t base = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanPerThread;
  bool f[kScanPerThread];
  int mine = 0;
#pragma unroll
  for (int k = 0; k < kScanPerThread; ++k) {
    f[k] = (base + k < n) && (pct[base + k] < threshold);
    mine += f[k];
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int incl = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int u = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += u;
  }
  __shared__ int s_warp[kScanThreads / 32];
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  int before = tile_off[blockIdx.x] + incl - mine;
  for (int w = 0; w < warp; ++w) before += s_warp[w];
#pragma unroll
  for (int k = 0; k < kScanPerThread; ++k)
    if (f[k]) low_idx[before++] = (int32_t)(base + k);
}

static int grid_for(int64_t items, int threads, int max_waves) {
  int64_t blocks = ceil_div(items, threads);
  const int64_t cap = (int64_t)kNumSMs * max_waves;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (int)blocks;
}

static int launch_cluster(const klm_node_cols* nodes, int64_t n, const uint8_t* pod_running,
                          int64_t n_pods, int64_t* sums, klm_cluster_out* out, void* scratch,
                          void* stream, const char* who) {
  if (!nodes || n < 0 || n_pods < 0 || !scratch || nodes->n_gpu_usage < 0 ||
      (n > 0 && (!nodes->cpu_capacity || !nodes->cpu_usage || !nodes->memory_capacity ||
                 !nodes->memory_usage || !nodes->gpu_count || !nodes->healthy)) ||
      (n_pods > 0 && !pod_running) || (nodes->n_gpu_usage > 0 && !nodes->gpu_usage)) {
    set_error("%s: bad argument", who);
    return KLM_EINVAL;
  }
  NodeColsDev c{nodes->cpu_capacity, nodes->cpu_usage, nodes->memory_capacity,
                nodes->memory_usage, nodes->gpu_count,  nodes->healthy,
                nodes->gpu_usage,    nodes->n_gpu_usage};
  const int vec_nodes = aligned16(c.cpu_cap) && aligned16(c.cpu_use) && aligned16(c.mem_cap) &&
                        aligned16(c.mem_use) && aligned16(c.gpu_count) &&
                        ((reinterpret_cast<uintptr_t>(c.healthy) & 7u) == 0);
  const int vec_pods = aligned16(pod_running);
  const int vec_gpu = aligned16(c.gpu_usage);
  char* s = static_cast<char*>(scratch);
  // enough threads that each reads ~8 nodes / 16 pods once; never more blocks than partial rows
  int64_t items = n / 8 + n_pods / 16 + c.n_gpu_usage / 2 + 1;
  int blocks = grid_for(items, kAggThreads, 4);
  if (blocks > kAggMaxBlocks) blocks = kAggMaxBlocks;
  cluster_aggregate_kernel<<<blocks, kAggThreads, 0, as_stream(stream)>>>(
      c, n, pod_running, n_pods, vec_nodes, vec_pods, vec_gpu,
      reinterpret_cast<unsigned int*>(s), reinterpret_cast<int64_t*>(s + 256), sums, out);
  KLM_LAUNCH_CHECK("cluster_aggregate_kernel");
  return KLM_OK;
}

}  // namespace klm

using namespace klm;

extern "C" int64_t klm_dev_cluster_scratch_bytes(void) { return kAggScratchBytes; }

extern "C" int klm_dev_cluster_aggregate(const klm_node_cols* nodes, int64_t n_nodes,
                                         const uint8_t* pod_running, int64_t n_pods,
                                         klm_cluster_out* out, void* scratch, void* stream) {
  if (!out) {
    set_error("klm_dev_cluster_aggregate: out is null");
    return KLM_EINVAL;
  }
  return launch_cluster(nodes, n_nodes, pod_running, n_pods, nullptr, out, scratch, stream,
                        "klm_dev_cluster_aggregate");
}

extern "C" int klm_dev_cluster_sums(const klm_node_cols* nodes, int64_t n_nodes,
                                    const uint8_t* pod_running, int64_t n_pods, int64_t* sums,
                                    void* scratch, void* stream) {
  if (!sums) {
    set_error("klm_dev_cluster_sums: sums is null");
    return KLM_EINVAL;
  }
  return launch_cluster(nodes, n_nodes, pod_running, n_pods, sums, nullptr, scratch, stream,
                        "klm_dev_cluster_sums");
}

extern "C" int klm_dev_cluster_finish(const int64_t* sums, klm_cluster_out* out, void* stream) {
  if (!sums || !out) {
    set_error("klm_dev_cluster_finish: null argument");
    return KLM_EINVAL;
  }
  cluster_finish_kernel<<<1, 32, 0, as_stream(stream)>>>(sums, out);
  KLM_LAUNCH_CHECK("cluster_finish_kernel");
  return KLM_OK;
}

extern "C" int klm_dev_node_rates(int64_t n, const int64_t* cpu_cap, const int64_t* cpu_use,
                                  const int64_t* mem_cap, const int64_t* mem_use,
                                  const int64_t* disk_cap, const int64_t* disk_alloc,
                                  double* cpu_rate, double* mem_rate, int64_t* disk_usage,
                                  double* disk_rate, uint8_t* under_pressure, void* stream) {
  if (n < 0 || (n > 0 && (!cpu_cap || !cpu_use || !mem_cap || !mem_use))) {
    set_error("klm_dev_node_rates: bad argument");
    return KLM_EINVAL;
  }
  if (n == 0) return KLM_OK;
  const bool vec = aligned16(cpu_cap) && aligned16(cpu_use) && aligned16(mem_cap) && aligned16(mem_use) &&
                   aligned16(disk_cap) && aligned16(disk_alloc) && aligned16(cpu_rate) && aligned16(mem_rate) &&
                   aligned16(disk_usage) && aligned16(disk_rate) && (reinterpret_cast<uintptr_t>(under_pressure) & 1) == 0;
  if (vec && n >= 2) {
    // equal trips for every thread: the fewest trips that fit 8 CTAs per SM, then just enough CTAs
    const int64_t pairs = n >> 1, per_trip = (int64_t)kNumSMs * 8 * 256;
    const int64_t trips = ceil_div(pairs, per_trip);
    const int64_t blocks = ceil_div(ceil_div(pairs, trips), 256);
    node_rates_pair_kernel<<<(unsigned)blocks, 256, 0, as_stream(stream)>>>(
        n, cpu_cap, cpu_use, mem_cap, mem_use, disk_cap, disk_alloc, cpu_rate, mem_rate, disk_usage, disk_rate,
        under_pressure);
    KLM_LAUNCH_CHECK("node_rates_pair_kernel");
    return KLM_OK;
  }
  node_rates_kernel<<<grid_for(n, 256, 8), 256, 0, as_stream(stream)>>>(
      n, cpu_cap, cpu_use, mem_cap, mem_use, disk_cap, disk_alloc, cpu_rate, mem_rate, disk_usage,
      disk_rate, under_pressure);
  KLM_LAUNCH_CHECK("node_rates_kernel");
  return KLM_OK;
}

extern "C" int klm_dev_pod_rates(int64_t n_pods, const int32_t* ctr_off, const int64_t* c_cpu_use,
                                 const int64_t* c_mem_use, const int64_t* c_cpu_req,
                                 const int64_t* c_cpu_lim, const int64_t* c_mem_req,
                                 const int64_t* c_mem_lim, int64_t* pod_totals, double* cpu_rate,
                                 double* mem_rate, double* cpu_util, double* mem_util,
                                 uint8_t* over_limit, void* stream) {
  if (n_pods < 0 || (n_pods > 0 && !ctr_off)) {
    set_error("klm_dev_pod_rates: bad argument");
    return KLM_EINVAL;
  }
  if (n_pods == 0) return KLM_OK;
  pod_rates_kernel<<<grid_for(n_pods, 256, 8), 256, 0, as_stream(stream)>>>(
      n_pods, ctr_off, c_cpu_use, c_mem_use, c_cpu_req, c_cpu_lim, c_mem_req, c_mem_lim,
      pod_totals, cpu_rate, mem_rate, cpu_util, mem_util, over_limit);
  KLM_LAUNCH_CHECK("pod_rates_kernel");
  return KLM_OK;
}

extern "C" int64_t klm_dev_fleet_scan_scratch_bytes(int64_t n) {
  const int64_t tiles = n > 0 ? ceil_div(n, kScanTile) : 1;
  return ((tiles * 4 + 255) / 256) * 256;
}

extern "C" int klm_dev_fleet_scan(int64_t n, const uint8_t* status_ok, const double* pct,
                                  double threshold, int64_t* counters, int32_t* low_idx,
                                  void* scratch, void* stream) {
  if (n < 0 || !counters || !scratch || (n > 0 && (!status_ok || !pct || !low_idx))) {
    set_error("klm_dev_fleet_scan: bad argument");
    return KLM_EINVAL;
  }
  cudaStream_t st = as_stream(stream);
  KLM_CUDA(cudaMemsetAsync(counters, 0, 2 * sizeof(int64_t), st));
  if (n == 0) return KLM_OK;
  const int64_t tiles = ceil_div(n, kScanTile);
  if (tiles > 0x7fffffff) {
    set_error("klm_dev_fleet_scan: n too large");
    return KLM_ELIMIT;
  }
  int32_t* tile_count = static_cast<int32_t*>(scratch);
  fleet_count_kernel<<<(unsigned)tiles, kScanThreads, 0, st>>>(n, status_ok, pct, threshold,
                                                               counters, tile_count);
  KLM_LAUNCH_CHECK("fleet_count_kernel");
  fleet_offsets_kernel<<<1, 1024, 0, st>>>(tile_count, (int)tiles, counters);
  KLM_LAUNCH_CHECK("fleet_offsets_kernel");
  fleet_write_kernel<<<(unsigned)tiles, kScanThreads, 0, st>>>(n, pct, threshold, tile_count,
                                                               low_idx);
  KLM_LAUNCH_CHECK("fleet_write_kernel");
  return KLM_OK;
}
