// Scheduling kernels: K1 node_best_reduce, K2 schedule (decomposed),
// K2b schedule_bruteforce, K3 merge_partials.
//
// Reference semantics (Sabre94/k8s-llm-monitor):
//   internal/scheduler/controller.go:174-221  buildCandidates (filters + score)
//   internal/scheduler/controller.go:160-161  stable sort desc, take [0]
// Selection rule everywhere below: highest score; on equal scores the lowest
// list index (what a stable descending sort followed by [0] yields).
// All score arithmetic is IEEE f64, as Go's float64; comparisons are on
// values, never on bit patterns (so -0.0 == +0.0 like in Go).
#include "common.cuh"

namespace klm {

// ---------------------------------------------------------------- K1

constexpr int kBestThreads = 256;
constexpr int kBestNodesPerThread = 16;  // 8 x 16 B of battery + 16 B of flags per trip
constexpr int kBestMaxBlocks = 2048;
// scratch layout: [0, 256) ticket (u32, self-resetting); then partial
// batteries (f64[kBestMaxBlocks]); then partial indices (i64[kBestMaxBlocks]).
constexpr int64_t kBestScratchBytes = 256 + (int64_t)kBestMaxBlocks * 16;

struct Best {
  double bat;
  int64_t idx;  // -1: none
};

__device__ __forceinline__ Best better(Best a, Best b) {
  // a wins unless b has a larger battery, or the same battery and a lower index
  if (b.idx < 0) return a;
  if (a.idx < 0) return b;
  if (b.bat > a.bat || (b.bat == a.bat && b.idx < a.idx)) return b;
  return a;
}

__device__ __forceinline__ Best warp_best(Best v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    Best w;
    w.bat = __shfl_xor_sync(0xffffffffu, v.bat, o);
    w.idx = __shfl_xor_sync(0xffffffffu, v.idx, o);
    v = better(v, w);
  }
  return v;
}

__device__ __forceinline__ Best block_best(Best v) {
  __shared__ double s_bat[kBestThreads / 32];
  __shared__ int64_t s_idx[kBestThreads / 32];
  v = warp_best(v);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) {
    s_bat[warp] = v.bat;
    s_idx[warp] = v.idx;
  }
  __syncthreads();
  if (warp == 0) {
    Best w;
    w.bat = lane < kBestThreads / 32 ? s_bat[lane] : 0.0;
    w.idx = lane < kBestThreads / 32 ? s_idx[lane] : -1;
    v = warp_best(w);
  }
  return v;  // valid in warp 0
}

// One pass over the node table: 9 B/node (f64 battery + u8 flags), 128-bit
// loads when both columns are 16-byte aligned. Block partials go to scratch;
// the last block to finish (ticket) folds them, so one launch suffices.
__global__ void __launch_bounds__(kBestThreads)
node_best_kernel(const double* __restrict__ battery, const uint8_t* __restrict__ flags,
                 int64_t n, int64_t node_base, int vec_ok, int64_t* __restrict__ best,
                 unsigned int* __restrict__ ticket, double* __restrict__ part_bat,
                 int64_t* __restrict__ part_idx) {
  // A kernel launched behind this one with programmatic stream serialisation
  // (klm_dev_schedule_after_best) may start now: K1 occupies a fraction of the SMs and its
  // result is needed only at the end of K2. Without such a dependent this is a no-op.
  asm volatile("griddepcontrol.launch_dependents;");
  Best me{0.0, -1};
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
  int64_t done = 0;
  if (vec_ok) {
    const int64_t nchunks = n / kBestNodesPerThread;
    for (int64_t c = tid; c < nchunks; c += nthreads) {
      const int64_t i0 = c * kBestNodesPerThread;
      const int4 fw = ld_stream_v4(flags + i0);
      double2 b[kBestNodesPerThread / 2];
#pragma unroll
      for (int k = 0; k < kBestNodesPerThread / 2; ++k) b[k] = ld_stream_d2(battery + i0 + 2 * k);
      const unsigned int f[4] = {(unsigned)fw.x, (unsigned)fw.y, (unsigned)fw.z, (unsigned)fw.w};
#pragma unroll
      for (int k = 0; k < kBestNodesPerThread; ++k) {
        const unsigned int fl = (f[k >> 2] >> (8 * (k & 3))) & 3u;
        const double v = (k & 1) ? b[k >> 1].y : b[k >> 1].x;
        // indices grow within a thread, so a strict > keeps the first maximum
        if (fl == 3u && (me.idx < 0 || v > me.bat)) {
          me.bat = v;
          me.idx = i0 + k;
        }
      }
    }
    done = nchunks * kBestNodesPerThread;
  }
  for (int64_t i = done + tid; i < n; i += nthreads) {
    const double v = battery[i];
    if ((flags[i] & 3u) == 3u && (me.idx < 0 || v > me.bat)) {
      me.bat = v;
      me.idx = i;
    }
  }
  me = block_best(me);

  __shared__ bool s_last;
  if (threadIdx.x == 0) {
    part_bat[blockIdx.x] = me.bat;
    part_idx[blockIdx.x] = me.idx;
    __threadfence();
    const unsigned int t = atomicAdd(ticket, 1u);
    s_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  Best acc{0.0, -1};
  for (int b = threadIdx.x; b < (int)gridDim.x; b += blockDim.x) {
    Best p;
    p.bat = __ldcg(part_bat + b);
    p.idx = __ldcg(part_idx + b);
    acc = better(acc, p);
  }
  acc = block_best(acc);
  if (threadIdx.x == 0) {
    best[0] = acc.idx < 0 ? -1 : acc.idx + node_base;
    best[1] = __double_as_longlong(acc.idx < 0 ? 0.0 : acc.bat);
    *ticket = 0;  // ready for the next launch on this scratch
  }
}

// ---------------------------------------------------------------- K2

constexpr int kSchedThreads = 256;
constexpr int kSchedUnroll = 4;  // requests per thread per pass
constexpr int kSchedSlots = 2;   // preferred ids fetched up front (the common list lengths)

// One thread per request. Winner = best of {g} U {preferred nodes that pass
// every filter}, where g is K1's result: a non-preferred node other than g
// scores battery <= battery[g] <= score(g) and on equality has a higher index
// (proved equal to the literal scan by the tests and by K2b).
// If g fails the battery gate every node does (g has the largest battery).
// AFTER_BEST: launched with programmatic stream serialisation right behind K1; `best` may be
// read only after griddepcontrol.wait (K1 complete and visible), everything else before.
template <bool AFTER_BEST>
__global__ void __launch_bounds__(kSchedThreads)
schedule_kernel(const double* __restrict__ battery, const uint8_t* __restrict__ flags,
                int64_t n, int64_t node_base, const int64_t* __restrict__ best,
                int64_t n_requests, const double* __restrict__ min_battery,
                const int32_t* __restrict__ pref_off, const int32_t* __restrict__ pref_ids,
                int32_t* __restrict__ out_idx, double* __restrict__ out_score) {
  int64_t g = 0;
  double gb = 0.0;
  if (!AFTER_BEST) {
    g = best[0];
    gb = __longlong_as_double(best[1]);
  }
  bool have_best = !AFTER_BEST;
  // A request is a chain of dependent loads (offsets -> ids -> node columns), each level an
  // L2 or DRAM round trip, so one request per thread leaves the kernel waiting on latency.
  // Every thread therefore carries kSchedUnroll requests and issues each level's loads for
  // all of them (and for the first kSchedSlots preferred ids of each) before using any.
  const int64_t tile = (int64_t)blockDim.x * kSchedUnroll;
  for (int64_t base = (int64_t)blockIdx.x * tile; base < n_requests; base += (int64_t)gridDim.x * tile) {
    double mb[kSchedUnroll];
    int32_t k0[kSchedUnroll], k1[kSchedUnroll];
#pragma unroll
    for (int j = 0; j < kSchedUnroll; ++j) {
      const int64_t r = base + (int64_t)j * blockDim.x + threadIdx.x;
      const bool live = r < n_requests;
      mb[j] = live ? min_battery[r] : 0.0;
      k0[j] = live ? pref_off[r] : 0;
      k1[j] = live ? pref_off[r + 1] : 0;
    }
    int32_t p[kSchedUnroll][kSchedSlots];
#pragma unroll
    for (int j = 0; j < kSchedUnroll; ++j)
#pragma unroll
      for (int q = 0; q < kSchedSlots; ++q) p[j][q] = k0[j] + q < k1[j] ? pref_ids[k0[j] + q] : -1;
    uint8_t f[kSchedUnroll][kSchedSlots];
    double b[kSchedUnroll][kSchedSlots];
#pragma unroll
    for (int j = 0; j < kSchedUnroll; ++j)
#pragma unroll
      for (int q = 0; q < kSchedSlots; ++q) {
        const int64_t lp = (int64_t)p[j][q] - node_base;
        const bool in = p[j][q] >= 0 && lp >= 0 && lp < n;
        f[j][q] = in ? flags[lp] : (uint8_t)0;  // 0 fails the name/status test below
        b[j][q] = in ? battery[lp] : 0.0;
      }
    if (!have_best) {  // first tile only: every load of this tile is in flight by now
      asm volatile("griddepcontrol.wait;" ::: "memory");
      g = __ldcg(best);
      gb = __longlong_as_double(__ldcg(best + 1));
      have_best = true;
    }
#pragma unroll
    for (int j = 0; j < kSchedUnroll; ++j) {
      const int64_t r = base + (int64_t)j * blockDim.x + threadIdx.x;
      if (r >= n_requests) continue;
      const bool gate = mb[j] > 0;  // controller.go:194
      int64_t bi = -1;
      double bs = 0.0;
      if (g >= 0 && !(gate && gb < mb[j])) {
        bi = g;
        bs = gb;
#pragma unroll
        for (int q = 0; q < kSchedSlots; ++q) {
          if ((f[j][q] & 3u) != 3u) continue;     // :190, :198
          if (gate && b[j][q] < mb[j]) continue;  // :194
          const double s = b[j][q] + 10;          // :205-208
          if (s > bs || (s == bs && p[j][q] < bi)) {
            bs = s;
            bi = p[j][q];
          }
        }
        for (int32_t k = k0[j] + kSchedSlots; k < k1[j]; ++k) {  // longer preferred lists
          const int64_t pk = pref_ids[k];
          const int64_t lp = pk - node_base;
          if (lp < 0 || lp >= n) continue;
          if ((flags[lp] & 3u) != 3u) continue;
          const double bk = battery[lp];
          if (gate && bk < mb[j]) continue;
          const double s = bk + 10;
          if (s > bs || (s == bs && pk < bi)) {
            bs = s;
            bi = pk;
          }
        }
      }
      out_idx[r] = (int32_t)bi;
      out_score[r] = bs;
    }
  }
}

// ---------------------------------------------------------------- K2b

constexpr int kBruteThreads = 256;
constexpr int kBruteTile = 2048;  // nodes per shared-memory tile: 16 KB + 2 KB
constexpr int kBruteRegPrefs = 4;

// The literal loop of controller.go:181-218 for every (request, node) pair:
// one request per thread, node tiles staged in shared memory and read as
// warp-wide broadcasts. First strict maximum in index order == stable sort
// then [0]. On-chip bound by construction (R*N*9 B of shared-memory reads).
__global__ void __launch_bounds__(kBruteThreads)
schedule_bruteforce_kernel(const double* __restrict__ battery, const uint8_t* __restrict__ flags,
                           int64_t n, int64_t node_base, int64_t n_requests,
                           const double* __restrict__ min_battery,
                           const int32_t* __restrict__ pref_off,
                           const int32_t* __restrict__ pref_ids, int32_t* __restrict__ out_idx,
                           double* __restrict__ out_score) {
  __shared__ double s_bat[kBruteTile];
  __shared__ uint8_t s_flag[kBruteTile];

  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = r < n_requests;
  double mb = 0.0;
  int32_t k0 = 0, k1 = 0;
  if (live) {
    mb = min_battery[r];
    k0 = pref_off[r];
    k1 = pref_off[r + 1];
  }
  const bool gate = mb > 0;
  const int npref = k1 - k0;
  int64_t p[kBruteRegPrefs];
#pragma unroll
  for (int k = 0; k < kBruteRegPrefs; ++k) p[k] = (k < npref) ? (int64_t)pref_ids[k0 + k] : -2;

  int64_t bi = -1;
  double bs = 0.0;
  for (int64_t t0 = 0; t0 < n; t0 += kBruteTile) {
    const int tn = (int)min((int64_t)kBruteTile, n - t0);
    __syncthreads();
    for (int i = threadIdx.x; i < tn; i += blockDim.x) {
      s_bat[i] = battery[t0 + i];
      s_flag[i] = flags[t0 + i];
    }
    __syncthreads();
    if (!live) continue;
#pragma unroll 4
    for (int i = 0; i < tn; ++i) {
      if ((s_flag[i] & 3u) != 3u) continue;  // :190, :198
      const double b = s_bat[i];
      if (gate && b < mb) continue;          // :194
      const int64_t gi = node_base + t0 + i;
      bool pref = false;
#pragma unroll
      for (int k = 0; k < kBruteRegPrefs; ++k) pref |= (gi == p[k]);
      if (npref > kBruteRegPrefs)
        for (int32_t k = k0 + kBruteRegPrefs; k < k1 && !pref; ++k) pref = (pref_ids[k] == gi);
      const double s = pref ? b + 10 : b;    // :205-208
      if (bi < 0 || s > bs) {                // first maximum in list order
        bs = s;
        bi = gi;
      }
    }
  }
  if (live) {
    out_idx[r] = (int32_t)bi;
    out_score[r] = bs;
  }
}

// ---------------------------------------------------------------- K3

// Per request, fold the partial winners of the node-range shards. Global
// indices make "lowest index on ties" independent of which rank held the node.
constexpr int kMergeChunk = 8;  // parts whose loads are in flight together
__global__ void __launch_bounds__(256)
merge_partials_kernel(int n_parts, int64_t n_requests, const int32_t* __restrict__ part_idx,
                      const double* __restrict__ part_score, int32_t* __restrict__ out_idx,
                      double* __restrict__ out_score) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n_requests; r += stride) {
    int32_t bi = -1;
    double bs = 0.0;
    // all loads of a chunk of parts are issued before the first compare (the score is read
    // whether or not the index is valid), so a request costs one memory round trip per chunk
    for (int g0 = 0; g0 < n_parts; g0 += kMergeChunk) {
      int32_t pi[kMergeChunk];
      double ps[kMergeChunk];
#pragma unroll
      for (int j = 0; j < kMergeChunk; ++j) {
        const bool live = g0 + j < n_parts;
        pi[j] = live ? __ldcs(part_idx + (int64_t)(g0 + j) * n_requests + r) : -1;
        ps[j] = live ? __ldcs(part_score + (int64_t)(g0 + j) * n_requests + r) : 0.0;
      }
#pragma unroll
      for (int j = 0; j < kMergeChunk; ++j) {
        if (pi[j] >= 0 && (bi < 0 || ps[j] > bs || (ps[j] == bs && pi[j] < bi))) {
          bs = ps[j];
          bi = pi[j];
        }
      }
    }
    out_idx[r] = bi;
    out_score[r] = bs;
  }
}

// K3 over peer memory: rank-local merge of requests [r0, r1) reading every rank's partial
// arrays where they live (peer pointers over NVLink / NVSwitch, or local ones). Each rank
// merges only its own slice, so every partial crosses the fabric at most once and no
// staging copy (all-gather) is made: exchange and merge are one kernel.
constexpr int kMaxPeerParts = 16;
struct PeerParts {
  const int32_t* idx[kMaxPeerParts];
  const double* score[kMaxPeerParts];
};

__global__ void __launch_bounds__(256)
merge_partials_peer_kernel(PeerParts parts, int n_parts, int64_t r0, int64_t r1,
                           int32_t* __restrict__ out_idx, double* __restrict__ out_score) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t r = r0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < r1; r += stride) {
    // issue every rank's loads before comparing: the peer reads overlap
    int32_t pi[kMaxPeerParts];
    double ps[kMaxPeerParts];
#pragma unroll
    for (int g = 0; g < kMaxPeerParts; ++g) {
      if (g < n_parts) {
        pi[g] = parts.idx[g][r];
        ps[g] = parts.score[g][r];
      }
    }
    int32_t bi = -1;
    double bs = 0.0;
#pragma unroll
    for (int g = 0; g < kMaxPeerParts; ++g) {
      if (g < n_parts && pi[g] >= 0 && (bi < 0 || ps[g] > bs || (ps[g] == bs && pi[g] < bi))) {
        bs = ps[g];
        bi = pi[g];
      }
    }
    out_idx[r - r0] = bi;
    out_score[r - r0] = bs;
  }
}

static int grid_for(int64_t work_items, int threads, int max_waves = 8) {
  int64_t blocks = ceil_div(work_items, threads);
  const int64_t cap = (int64_t)kNumSMs * max_waves;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (int)blocks;
}

}  // namespace klm

using namespace klm;

extern "C" int64_t klm_dev_node_best_scratch_bytes(void) { return kBestScratchBytes; }

extern "C" int klm_dev_node_best(const double* battery, const uint8_t* flags, int64_t n_nodes,
                                 int64_t node_base, int64_t* best, void* scratch, void* stream) {
  if (n_nodes < 0 || !best || !scratch || (n_nodes > 0 && (!battery || !flags))) {
    set_error("klm_dev_node_best: bad argument");
    return KLM_EINVAL;
  }
  char* s = static_cast<char*>(scratch);
  unsigned int* ticket = reinterpret_cast<unsigned int*>(s);
  double* part_bat = reinterpret_cast<double*>(s + 256);
  int64_t* part_idx = reinterpret_cast<int64_t*>(s + 256 + (int64_t)kBestMaxBlocks * 8);
  int64_t blocks = ceil_div(n_nodes, (int64_t)kBestThreads * kBestNodesPerThread);
  if (blocks < 1) blocks = 1;
  if (blocks > kBestMaxBlocks) blocks = kBestMaxBlocks;
  if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
  const int vec_ok = aligned16(battery) && aligned16(flags);
  node_best_kernel<<<(int)blocks, kBestThreads, 0, as_stream(stream)>>>(
      battery, flags, n_nodes, node_base, vec_ok, best, ticket, part_bat, part_idx);
  KLM_LAUNCH_CHECK("node_best_kernel");
  return KLM_OK;
}

static int launch_schedule(bool after_best, const double* battery, const uint8_t* flags, int64_t n_nodes,
                           int64_t node_base, const int64_t* best, int64_t n_requests,
                           const double* min_battery, const int32_t* pref_off, const int32_t* pref_ids,
                           int32_t* out_idx, double* out_score, void* stream, const char* who) {
  if (n_nodes < 0 || n_requests < 0 || !best) {
    set_error("%s: bad argument", who);
    return KLM_EINVAL;
  }
  if (n_requests == 0) return KLM_OK;
  if (!min_battery || !pref_off || !out_idx || !out_score) {
    set_error("%s: null request array", who);
    return KLM_EINVAL;
  }
  const int grid = grid_for(ceil_div(n_requests, (int64_t)kSchedUnroll), kSchedThreads, 8);
  if (!after_best) {
    schedule_kernel<false><<<grid, kSchedThreads, 0, as_stream(stream)>>>(
        battery, flags, n_nodes, node_base, best, n_requests, min_battery, pref_off, pref_ids, out_idx,
        out_score);
  } else {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(kSchedThreads);
    cfg.stream = as_stream(stream);
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    KLM_CUDA(cudaLaunchKernelEx(&cfg, schedule_kernel<true>, battery, flags, n_nodes, node_base, best,
                                n_requests, min_battery, pref_off, pref_ids, out_idx, out_score));
  }
  KLM_LAUNCH_CHECK("schedule_kernel");
  return KLM_OK;
}

extern "C" int klm_dev_schedule(const double* battery, const uint8_t* flags, int64_t n_nodes,
                                int64_t node_base, const int64_t* best, int64_t n_requests,
                                const double* min_battery, const int32_t* pref_off,
                                const int32_t* pref_ids, int32_t* out_idx, double* out_score,
                                void* stream) {
  return launch_schedule(false, battery, flags, n_nodes, node_base, best, n_requests, min_battery,
                         pref_off, pref_ids, out_idx, out_score, stream, "klm_dev_schedule");
}

extern "C" int klm_dev_schedule_after_best(const double* battery, const uint8_t* flags, int64_t n_nodes,
                                           int64_t node_base, const int64_t* best, int64_t n_requests,
                                           const double* min_battery, const int32_t* pref_off,
                                           const int32_t* pref_ids, int32_t* out_idx,
                                           double* out_score, void* stream) {
  return launch_schedule(true, battery, flags, n_nodes, node_base, best, n_requests, min_battery,
                         pref_off, pref_ids, out_idx, out_score, stream, "klm_dev_schedule_after_best");
}

extern "C" int klm_dev_schedule_bruteforce(const double* battery, const uint8_t* flags,
                                           int64_t n_nodes, int64_t node_base, int64_t n_requests,
                                           const double* min_battery, const int32_t* pref_off,
                                           const int32_t* pref_ids, int32_t* out_idx,
                                           double* out_score, void* stream) {
  if (n_nodes < 0 || n_requests < 0) {
    set_error("klm_dev_schedule_bruteforce: bad argument");
    return KLM_EINVAL;
  }
  if (n_requests == 0) return KLM_OK;
  if (!min_battery || !pref_off || !out_idx || !out_score) {
    set_error("klm_dev_schedule_bruteforce: null request array");
    return KLM_EINVAL;
  }
  const int64_t blocks = ceil_div(n_requests, kBruteThreads);
  if (blocks > 0x7fffffff) {
    set_error("klm_dev_schedule_bruteforce: too many requests for one launch");
    return KLM_ELIMIT;
  }
  schedule_bruteforce_kernel<<<(unsigned)blocks, kBruteThreads, 0, as_stream(stream)>>>(
      battery, flags, n_nodes, node_base, n_requests, min_battery, pref_off, pref_ids, out_idx,
      out_score);
  KLM_LAUNCH_CHECK("schedule_bruteforce_kernel");
  return KLM_OK;
}

extern "C" int klm_dev_merge_partials(int n_parts, int64_t n_requests, const int32_t* part_idx,
                                      const double* part_score, int32_t* out_idx,
                                      double* out_score, void* stream) {
  if (n_parts < 1 || n_requests < 0) {
    set_error("klm_dev_merge_partials: bad argument");
    return KLM_EINVAL;
  }
  if (n_requests == 0) return KLM_OK;
  merge_partials_kernel<<<grid_for(n_requests, 256, 1 << 16), 256, 0, as_stream(stream)>>>(
      n_parts, n_requests, part_idx, part_score, out_idx, out_score);
  KLM_LAUNCH_CHECK("merge_partials_kernel");
  return KLM_OK;
}

extern "C" int klm_dev_merge_partials_peer(int n_parts, const int32_t* const* part_idx,
                                           const double* const* part_score, int64_t r0, int64_t r1,
                                           int32_t* out_idx, double* out_score, void* stream) {
  if (n_parts < 1 || n_parts > kMaxPeerParts || !part_idx || !part_score || r0 < 0 || r1 < r0) {
    set_error("klm_dev_merge_partials_peer: bad argument (1 <= n_parts <= %d)", kMaxPeerParts);
    return n_parts > kMaxPeerParts ? KLM_ELIMIT : KLM_EINVAL;
  }
  if (r1 == r0) return KLM_OK;
  if (!out_idx || !out_score) {
    set_error("klm_dev_merge_partials_peer: null output");
    return KLM_EINVAL;
  }
  PeerParts p{};
  for (int g = 0; g < n_parts; ++g) {
    if (!part_idx[g] || !part_score[g]) {
      set_error("klm_dev_merge_partials_peer: null partial array for part %d", g);
      return KLM_EINVAL;
    }
    p.idx[g] = part_idx[g];
    p.score[g] = part_score[g];
  }
  merge_partials_peer_kernel<<<grid_for(r1 - r0, 256, 16), 256, 0, as_stream(stream)>>>(
      p, n_parts, r0, r1, out_idx, out_score);
  KLM_LAUNCH_CHECK("merge_partials_peer_kernel");
  return KLM_OK;
}
