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
#include <cstdlib>

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
// Node-range sharding: every rank owns one shard of the node table and an array
// of exchange slots, [2 epoch parities][kMaxShardParts] of 32 bytes, that its peers write into
// over NVLink (peer pointers: cudaDeviceEnablePeerAccess in one process, CUDA IPC / symmetric
// memory across processes). Rank r publishes its shard's K1 result into slot [epoch & 1][r] of
// EVERY rank -- payload first, then the epoch with release semantics at system scope -- and the
// sharded K2 of each rank waits (acquire) until all slots of its own array carry the epoch.
// Alternating parities let step e + 1 publish while a slower rank still reads step e; a rank
// cannot be two steps ahead, because its own K2 of step e + 1 needs every rank's K1 of e + 1.
constexpr int kMaxShardParts = 16;
struct alignas(32) ShardSlot {
  int64_t idx;                // global index of the shard's best node, -1: none eligible
  int64_t bat_bits;           // its battery (f64 bits)
  unsigned long long epoch;   // written last
  unsigned long long pad;
};
constexpr int64_t kShardSlotsBytes = 2 * kMaxShardParts * (int64_t)sizeof(ShardSlot) + 64;  // + error word
struct ShardPublish {
  ShardSlot* slots[kMaxShardParts];  // slot array of every rank, as seen from this device
  int n_parts, my_part;
  unsigned long long epoch;
};
// K1 with SPREAD also pushes the shard it is reading into every peer's replica of the node table
// (plain stores through peer pointers: an all-gather fused into the scan, the shard crosses NVLink
// once per peer in full lines while K1 would be reading it anyway). rep_* are the replica BASES;
// the shard goes to offset node_base. The rank's own replica is skipped (the shard already is
// that part of it).
struct ShardSpread {
  double* rep_bat[kMaxShardParts];
  uint8_t* rep_flg[kMaxShardParts];
};

// The scan of K1: this CTA's share of the nodes (tiles strided by the grid), per-thread first
// maximum; with SPREAD every tile is also pushed into the peers' replicas.
template <bool SPREAD>
__device__ __forceinline__ Best scan_nodes(const double* __restrict__ battery, const uint8_t* __restrict__ flags,
                                           int64_t n, int64_t node_base, int vec_ok, const ShardPublish& pub,
                                           const ShardSpread& spr) {
  Best me{0.0, -1};
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
  int64_t done = 0;
  if (vec_ok) {
    // Tiles of 256 x 8 node pairs: in every load and store the lanes of a warp touch consecutive
    // 16 bytes (battery pairs) / 2 bytes (flag pairs), i.e. whole 512-byte and 64-byte runs --
    // which is what lets the pushes into the peers' replicas (SPREAD) cross NVLink as full lines.
    constexpr int kPairs = kBestNodesPerThread / 2;  // node pairs per thread per tile
    const int64_t npairs = n >> 1;
    const int64_t tile_pairs = (int64_t)kBestThreads * kPairs;
    for (int64_t p0 = (int64_t)blockIdx.x * tile_pairs; p0 < npairs; p0 += (int64_t)gridDim.x * tile_pairs) {
      double2 b[kPairs];
      unsigned short f[kPairs];
#pragma unroll
      for (int u = 0; u < kPairs; ++u) {
        const int64_t p = p0 + (int64_t)u * kBestThreads + threadIdx.x;
        b[u] = make_double2(0.0, 0.0);
        f[u] = 0;
        if (p < npairs) {
          b[u] = ld_stream_d2(battery + 2 * p);
          f[u] = __ldg(reinterpret_cast<const unsigned short*>(flags) + p);
        }
      }
      if (SPREAD) {
        for (int g = 0; g < pub.n_parts; ++g) {
          if (g == pub.my_part) continue;
          double2* db = reinterpret_cast<double2*>(spr.rep_bat[g] + node_base);
          unsigned short* df = reinterpret_cast<unsigned short*>(spr.rep_flg[g] + node_base);
#pragma unroll
          for (int u = 0; u < kPairs; ++u) {
            const int64_t p = p0 + (int64_t)u * kBestThreads + threadIdx.x;
            if (p < npairs) {
              db[p] = b[u];
              df[p] = f[u];
            }
          }
        }
      }
#pragma unroll
      for (int u = 0; u < kPairs; ++u) {
        const int64_t p = p0 + (int64_t)u * kBestThreads + threadIdx.x;
        // indices grow within a thread (u, then the tile), so a strict > keeps the first maximum;
        // a pair beyond the end carries flags 0 and is never eligible
        if ((f[u] & 3u) == 3u && (me.idx < 0 || b[u].x > me.bat)) {
          me.bat = b[u].x;
          me.idx = 2 * p;
        }
        if (((f[u] >> 8) & 3u) == 3u && (me.idx < 0 || b[u].y > me.bat)) {
          me.bat = b[u].y;
          me.idx = 2 * p + 1;
        }
      }
    }
    done = npairs * 2;
  }
  for (int64_t i = done + tid; i < n; i += nthreads) {
    const double v = battery[i];
    if (SPREAD) {
      const uint8_t fl = flags[i];
      for (int g = 0; g < pub.n_parts; ++g) {
        if (g == pub.my_part) continue;
        spr.rep_bat[g][node_base + i] = v;
        spr.rep_flg[g][node_base + i] = fl;
      }
    }
    if ((flags[i] & 3u) == 3u && (me.idx < 0 || v > me.bat)) {
      me.bat = v;
      me.idx = i;
    }
  }
  return me;
}

// The publication at the end of K1: lane g of warp 0 writes rank g's copy of this rank's slot.
__device__ __forceinline__ void publish_best(const ShardPublish& pub, int64_t g_idx, int64_t g_bits) {
  if ((int)threadIdx.x < pub.n_parts) {
    ShardSlot* sl = pub.slots[threadIdx.x] + (pub.epoch & 1ull) * kMaxShardParts + pub.my_part;
    asm volatile("st.relaxed.sys.global.v2.s64 [%0], {%1, %2};" :: "l"(sl), "l"(g_idx), "l"(g_bits) : "memory");
    asm volatile("st.release.sys.global.u64 [%0], %1;" :: "l"(&sl->epoch), "l"(pub.epoch) : "memory");
  }
}

template <bool PUBLISH, bool SPREAD = false>
__global__ void __launch_bounds__(kBestThreads)
node_best_kernel(const double* __restrict__ battery, const uint8_t* __restrict__ flags,
                 int64_t n, int64_t node_base, int vec_ok, int64_t* __restrict__ best,
                 unsigned int* __restrict__ ticket, double* __restrict__ part_bat,
                 int64_t* __restrict__ part_idx, ShardPublish pub, ShardSpread spr) {
  // A kernel launched behind this one with programmatic stream serialisation
  // (klm_dev_schedule_after_best) may start now: K1 occupies a fraction of the SMs and its
  // result is needed only at the end of K2. Without such a dependent this is a no-op.
  asm volatile("griddepcontrol.launch_dependents;");
  Best me = scan_nodes<SPREAD>(battery, flags, n, node_base, vec_ok, pub, spr);
  me = block_best(me);

  __shared__ bool s_last;
  if (threadIdx.x == 0) {
    part_bat[blockIdx.x] = me.bat;
    part_idx[blockIdx.x] = me.idx;
    // SPREAD: the block's stores into the peers' replicas (made visible to this thread by the
    // barrier inside block_best) must be performed system-wide before the ticket says "done"
    if (SPREAD) __threadfence_system(); else __threadfence();
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
  acc = block_best(acc);  // warp 0 holds the result in every lane
  const int64_t g_idx = acc.idx < 0 ? -1 : acc.idx + node_base;
  const int64_t g_bits = __double_as_longlong(acc.idx < 0 ? 0.0 : acc.bat);
  if (threadIdx.x == 0) {
    if (best) {
      best[0] = g_idx;
      best[1] = g_bits;
    }
    *ticket = 0;  // ready for the next launch on this scratch
  }
  if (PUBLISH) publish_best(pub, g_idx, g_bits);  // (n_parts <= 16: all in warp 0, which holds the result)
}

// K1 as ONE thread-block cluster (tables up to a few million nodes): the CTAs' partial winners
// meet in CTA 0's shared memory through distributed shared memory and a cluster barrier -- no
// global partials, no __threadfence + ticket, no scratch. Same result as node_best_kernel.
constexpr int kBestClusterMax = 16;
template <bool PUBLISH, bool SPREAD>
__global__ void __launch_bounds__(kBestThreads)
node_best_cluster_kernel(const double* __restrict__ battery, const uint8_t* __restrict__ flags, int64_t n,
                         int64_t node_base, int vec_ok, int64_t* __restrict__ best, ShardPublish pub,
                         ShardSpread spr) {
  __shared__ double c_bat[kBestClusterMax];
  __shared__ int64_t c_idx[kBestClusterMax];
  asm volatile("griddepcontrol.launch_dependents;");
  Best me = scan_nodes<SPREAD>(battery, flags, n, node_base, vec_ok, pub, spr);
  me = block_best(me);
  unsigned int rank, n_cta;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(n_cta));
  if (threadIdx.x == 0) {
    // slot `rank` of CTA 0's arrays
    const uint32_t lb = (uint32_t)__cvta_generic_to_shared(&c_bat[rank]);
    const uint32_t li = (uint32_t)__cvta_generic_to_shared(&c_idx[rank]);
    uint32_t rb, ri;
    asm volatile("mapa.shared::cluster.u32 %0, %1, 0;" : "=r"(rb) : "r"(lb));
    asm volatile("mapa.shared::cluster.u32 %0, %1, 0;" : "=r"(ri) : "r"(li));
    asm volatile("st.shared::cluster.f64 [%0], %1;" :: "r"(rb), "d"(me.bat) : "memory");
    asm volatile("st.shared::cluster.s64 [%0], %1;" :: "r"(ri), "l"(me.idx) : "memory");
  }
  // SPREAD: this CTA's pushes into the peers' replicas must be performed system-wide before CTA 0
  // may publish; the cluster barrier orders them only at cluster scope
  if (SPREAD) __threadfence_system();
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
  if (rank != 0 || threadIdx.x >= 32) return;
  Best w{0.0, -1};
  if (threadIdx.x < n_cta) {
    w.bat = c_bat[threadIdx.x];
    w.idx = c_idx[threadIdx.x];
  }
  w = warp_best(w);
  const int64_t g_idx = w.idx < 0 ? -1 : w.idx + node_base;
  const int64_t g_bits = __double_as_longlong(w.idx < 0 ? 0.0 : w.bat);
  if (threadIdx.x == 0 && best) {
    best[0] = g_idx;
    best[1] = g_bits;
  }
  if (PUBLISH) publish_best(pub, g_idx, g_bits);
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

// K1 and K2 in ONE launch (klm_dev_schedule_fused). Two dependent kernels cost two cold starts and
// K1's block -> ticket -> last-block chain before any request can be decided; here every CTA of a
// one-wave grid (all co-resident: cooperative launch) scans its share of the node table, leaves its
// partial winner in scratch and arrives on a counter, issues the loads of its first request tile
// (they do not depend on the winner), and only then waits for the counter; each CTA folds the
// partial winners itself (a few KB from L2) and decides its requests. The counters are
// self-resetting: the last CTA to leave zeroes them. Same results as K1 + K2.
constexpr int kFusedMinBlocks = 4;  // CTAs per SM the register budget is set for
__global__ void __launch_bounds__(kSchedThreads, kFusedMinBlocks)
schedule_fused_kernel(const double* __restrict__ battery, const uint8_t* __restrict__ flags, int64_t n,
                      int64_t node_base, int vec_ok, int nscan, unsigned int* __restrict__ arrive,
                      unsigned int* __restrict__ depart, double* __restrict__ part_bat,
                      int64_t* __restrict__ part_idx, int64_t* __restrict__ best_out, int64_t n_requests,
                      const double* __restrict__ min_battery, const int32_t* __restrict__ pref_off,
                      const int32_t* __restrict__ pref_ids, int32_t* __restrict__ out_idx,
                      double* __restrict__ out_score) {
  static_assert(kSchedThreads == kBestThreads, "scan_nodes and block_best assume the K1 block size");
  __shared__ int64_t s_g;
  __shared__ double s_gb;
  int64_t g = -1;
  double gb = 0.0;
  bool have_best = false;
  // all partial winners are in: fold them (every CTA for itself) and leave the barrier
  auto wait_best = [&]() {
    if (threadIdx.x == 0) {
      unsigned int a;
      do {
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(a) : "l"(arrive) : "memory");
      } while (a < (unsigned int)nscan);
    }
    __syncthreads();
    Best acc{0.0, -1};
    for (int b = threadIdx.x; b < nscan; b += kSchedThreads) {
      Best q;
      q.bat = __ldcg(part_bat + b);
      q.idx = __ldcg(part_idx + b);
      acc = better(acc, q);
    }
    acc = block_best(acc);  // warp 0 holds the result
    if (threadIdx.x == 0) {
      s_g = acc.idx < 0 ? -1 : acc.idx + node_base;
      s_gb = acc.idx < 0 ? 0.0 : acc.bat;
      if (blockIdx.x == 0 && best_out) {
        best_out[0] = s_g;
        best_out[1] = __double_as_longlong(s_gb);
      }
      if (atomicAdd(depart, 1u) == gridDim.x - 1) {  // everybody is past the wait: ready for the next launch
        *arrive = 0;
        *depart = 0;
      }
    }
    __syncthreads();
    g = s_g;
    gb = s_gb;
    have_best = true;
  };
  if ((int)blockIdx.x < nscan) {
    Best me = scan_nodes<false>(battery, flags, n, node_base, vec_ok, ShardPublish{}, ShardSpread{});
    me = block_best(me);
    if (threadIdx.x == 0) {
      part_bat[blockIdx.x] = me.bat;
      part_idx[blockIdx.x] = me.idx;
      __threadfence();
      atomicAdd(arrive, 1u);
    }
  }
  // (Streaming the request tiles through shared memory with cp.async one tile ahead -- a tile's
  // offsets, thresholds and CSR ids are contiguous -- was tried here and is slower, 27.2 against
  // 25.2 us at C4: the kernel is bound by its ~140 instructions per request and the three dependent
  // levels of one or two tiles per CTA, not by bytes.)
  // Node indices are 32-bit here (n + node_base < 2^31, checked by the caller: preferred ids and
  // out_idx are int32 anyway), the battery gate is folded into one threshold (-inf: no gate), and a
  // request is a straight line of predicated selects.
  const int n32 = (int)n, nb32 = (int)node_base;
  const int64_t tile = (int64_t)kSchedThreads * kSchedUnroll;
  for (int64_t base = (int64_t)blockIdx.x * tile; base < n_requests; base += (int64_t)gridDim.x * tile) {
    const int rem = (int)(n_requests - base < tile ? n_requests - base : tile);
    const double* __restrict__ t_mb = min_battery + base;
    const int32_t* __restrict__ t_off = pref_off + base;
    double mb[kSchedUnroll];
    int32_t k0[kSchedUnroll], k1[kSchedUnroll];
#pragma unroll
    for (int j = 0; j < kSchedUnroll; ++j) {
      const int i = j * kSchedThreads + (int)threadIdx.x;
      const bool live = i < rem;
      mb[j] = live ? t_mb[i] : 0.0;
      k0[j] = live ? t_off[i] : 0;
      k1[j] = live ? t_off[i + 1] : 0;
    }
    int32_t p[kSchedUnroll][kSchedSlots];
#pragma unroll
    for (int j = 0; j < kSchedUnroll; ++j)
#pragma unroll
      for (int q = 0; q < kSchedSlots; ++q) p[j][q] = k0[j] + q < k1[j] ? pref_ids[k0[j] + q] : -1;
    uint32_t f[kSchedUnroll][kSchedSlots];
    double b[kSchedUnroll][kSchedSlots];
#pragma unroll
    for (int j = 0; j < kSchedUnroll; ++j)
#pragma unroll
      for (int q = 0; q < kSchedSlots; ++q) {
        const int lp = p[j][q] - nb32;
        const bool in = p[j][q] >= 0 && (unsigned)lp < (unsigned)n32;
        f[j][q] = in ? (uint32_t)flags[lp] : 0u;  // 0 fails the name/status test below
        b[j][q] = in ? battery[lp] : 0.0;
      }
    if (!have_best) wait_best();  // first tile only: every load of this tile is in flight by now
    const int g32 = (int)g;
#pragma unroll
    for (int j = 0; j < kSchedUnroll; ++j) {
      const int i = j * kSchedThreads + (int)threadIdx.x;
      if (i >= rem) continue;
      const double thr = mb[j] > 0 ? mb[j] : __longlong_as_double(0xfff0000000000000ll);  // controller.go:194
      const bool okg = g32 >= 0 && !(gb < thr);  // g has the largest battery: if it fails the gate, all do
      double bs = gb;
      int bi = g32;
#pragma unroll
      for (int q = 0; q < kSchedSlots; ++q) {
        const double sc = b[j][q] + 10;                                // :205-208
        const bool e = (f[j][q] & 3u) == 3u && !(b[j][q] < thr);       // :190, :198, :194
        const bool take = e && (sc > bs || (sc == bs && p[j][q] < bi));
        bs = take ? sc : bs;
        bi = take ? p[j][q] : bi;
      }
      if (okg) {
        for (int32_t k = k0[j] + kSchedSlots; k < k1[j]; ++k) {  // longer preferred lists
          const int pk = pref_ids[k];
          const int lp = pk - nb32;
          if (pk < 0 || (unsigned)lp >= (unsigned)n32) continue;
          if ((flags[lp] & 3u) != 3u) continue;
          const double bk = battery[lp];
          if (bk < thr) continue;
          const double sc = bk + 10;
          if (sc > bs || (sc == bs && pk < bi)) {
            bs = sc;
            bi = pk;
          }
        }
      }
      out_idx[base + i] = okg ? bi : -1;
      out_score[base + i] = okg ? bs : 0.0;
    }
  }
  if (!have_best) wait_best();  // a CTA without requests still folds (CTA 0 writes best_out) and leaves
}

// Rendezvous of the ranks on their streams: publish `epoch` into this rank's slot of every rank's
// array, then wait until this rank's own array carries it from everybody. One warp. Used to line
// the ranks up in front of a timed region (bench.py) or a snapshot change; the data path itself
// (K1 publish -> sharded K2) needs no barrier. `slots` arrays are separate from the K1 ones.
__global__ void __launch_bounds__(32)
shard_rendezvous_kernel(ShardPublish pub, const ShardSlot* __restrict__ my_slots, int* __restrict__ err) {
  const int g = threadIdx.x;
  if (g >= pub.n_parts) return;
  ShardSlot* out = pub.slots[g] + (pub.epoch & 1ull) * kMaxShardParts + pub.my_part;
  asm volatile("st.release.sys.global.u64 [%0], %1;" :: "l"(&out->epoch), "l"(pub.epoch) : "memory");
  const ShardSlot* in = my_slots + (pub.epoch & 1ull) * kMaxShardParts + g;
  unsigned long long e = 0, t0 = 0, now = 0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  while (true) {
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(e) : "l"(&in->epoch) : "memory");
    if (e == pub.epoch) break;
    __nanosleep(40);
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
    if (now - t0 > 4000000000ull) {  // 4 s: a peer never arrived
      if (err) *err = 1;
      break;
    }
  }
}

// ---------------------------------------------------------------- K2, node table in shards

// The node table as G contiguous shards of `shard_len` nodes (the last ones shorter or empty),
// each readable from this device: its own memory, or a peer's mapped over NVLink / NVSwitch.
// Node i lives in shard i / shard_len at offset i % shard_len.
struct ShardView {
  const double* bat[kMaxShardParts];
  const uint8_t* flg[kMaxShardParts];
  int64_t n_total;
  int32_t shard_len;
  float inv_len;  // 1 / shard_len, for the owner estimate
  int n_parts;
};

__device__ __forceinline__ bool shard_lookup(const ShardView& v, int32_t id, uint8_t& f, double& b) {
  if (id < 0 || (int64_t)id >= v.n_total) return false;
  int g = __float2int_rz(__int2float_rz(id) * v.inv_len);  // within one of id / shard_len
  g = min(g, v.n_parts - 1);
  int32_t off = id - g * v.shard_len;
  if (off < 0) { --g; off += v.shard_len; }
  else if (off >= v.shard_len) { ++g; off -= v.shard_len; }
  f = v.flg[g][off];
  b = v.bat[g][off];
  return true;
}

// Wait until this rank's slot array holds every rank's K1 result of `epoch`, fold them into the
// global best (largest battery, lowest global index). Block-wide.
__device__ __forceinline__ void shard_wait_best(const ShardView& v, const ShardSlot* __restrict__ my_slots,
                                                unsigned long long epoch, int* __restrict__ err, int64_t& g,
                                                double& gb) {
  __shared__ int64_t s_g;
  __shared__ double s_gb;
  asm volatile("griddepcontrol.wait;" ::: "memory");
  if (threadIdx.x < 32) {
    Best w{0.0, -1};
    if ((int)threadIdx.x < v.n_parts) {
      const ShardSlot* sl = my_slots + (epoch & 1ull) * kMaxShardParts + threadIdx.x;
      unsigned long long e = 0, t0 = 0, now = 0;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
      bool ok = true;
      while (true) {
        asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(e) : "l"(&sl->epoch) : "memory");
        if (e == epoch) break;
        __nanosleep(40);
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
        if (now - t0 > 4000000000ull) { ok = false; break; }  // 4 s: a peer never published
      }
      if (ok) {
        int64_t pi, pb;
        asm volatile("ld.relaxed.sys.global.v2.s64 {%0, %1}, [%2];" : "=l"(pi), "=l"(pb) : "l"(sl) : "memory");
        w.idx = pi;
        w.bat = __longlong_as_double(pb);
      } else if (err) {
        *err = 1;
      }
    }
    w = warp_best(w);
    if (threadIdx.x == 0) {
      s_g = w.idx;
      s_gb = w.bat;
    }
  }
  __syncthreads();
  g = s_g;
  gb = s_gb;
}

// Sharded form of schedule_kernel: this rank decides ITS slice of the requests. The global best
// node is the fold of the G shard winners the ranks published into `my_slots` (waited for with
// acquire loads: no host barrier, no K3); preferred nodes are probed where they live, through
// the pointers of the view -- a 9-byte read over NVLink for a node another rank owns.
// Results are identical to schedule_kernel on the whole table (global indices; lowest index on
// ties, controller.go:160).
// WAIT_FIRST: the node columns themselves are written by this epoch's K1s (klm_dev_node_best_spread
// fills the replicas), so nothing of them may be read before the slots say that every rank is
// done; otherwise the columns were complete before K1 started and the first tile's probes are
// issued ahead of the wait.
template <bool WAIT_FIRST>
__global__ void __launch_bounds__(kSchedThreads, 4)  // 64 registers: a rank's slice of C4 / C5 is one wave
schedule_sharded_kernel(ShardView v, const ShardSlot* __restrict__ my_slots, unsigned long long epoch,
                        int* __restrict__ err, int64_t n_requests,
                        const double* __restrict__ min_battery, const int32_t* __restrict__ pref_off,
                        const int32_t* __restrict__ pref_ids, int32_t* __restrict__ out_idx,
                        double* __restrict__ out_score) {
  int64_t g = -1;
  double gb = 0.0;
  bool have_best = false;
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
    if (WAIT_FIRST && !have_best) {  // the request loads above are in flight; the node columns come after
      shard_wait_best(v, my_slots, epoch, err, g, gb);
      have_best = true;
    }
    uint8_t f[kSchedUnroll][kSchedSlots];
    double b[kSchedUnroll][kSchedSlots];
#pragma unroll
    for (int j = 0; j < kSchedUnroll; ++j)
#pragma unroll
      for (int q = 0; q < kSchedSlots; ++q) {
        f[j][q] = 0;  // 0 fails the name/status test below
        b[j][q] = 0.0;
        shard_lookup(v, p[j][q], f[j][q], b[j][q]);
      }
    if (!WAIT_FIRST && !have_best) {  // first tile only: every load of this tile is in flight by now
      shard_wait_best(v, my_slots, epoch, err, g, gb);
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
          const double sc = b[j][q] + 10;         // :205-208
          if (sc > bs || (sc == bs && p[j][q] < bi)) {
            bs = sc;
            bi = p[j][q];
          }
        }
        for (int32_t k = k0[j] + kSchedSlots; k < k1[j]; ++k) {  // longer preferred lists
          const int32_t pk = pref_ids[k];
          uint8_t fk = 0;
          double bk = 0.0;
          if (!shard_lookup(v, pk, fk, bk)) continue;
          if ((fk & 3u) != 3u) continue;
          if (gate && bk < mb[j]) continue;
          const double sc = bk + 10;
          if (sc > bs || (sc == bs && pk < bi)) {
            bs = sc;
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

// Tables (shards) up to this many nodes take the one-cluster form of K1: a CTA works through its
// tiles one memory round trip at a time, so 16 CTAs win only while each has two or three tiles
// (100 k nodes: 12.3 us against 14.3 us for the ticket chain, cold; 500 k: 22.5 against 14.3 --
// profiles/r02_c5_phases_n2_cluster_ab.jsonl).
constexpr int64_t kBestClusterNodes = 160 * 1024;

// Launches K1 as one cluster of 1..16 CTAs. Returns KLM_OK, or 1 if the cluster launch is not
// possible here (the caller then takes the multi-block kernel), or an error.
template <bool PUBLISH, bool SPREAD>
static int launch_best_cluster(const double* battery, const uint8_t* flags, int64_t n, int64_t node_base,
                               int vec_ok, int64_t* best, const ShardPublish& pub, const ShardSpread& spr,
                               cudaStream_t st) {
  static PerDeviceOnce attr_set, unusable;
  static const bool disabled = std::getenv("KLM_K1_NO_CLUSTER") != nullptr;  // A/B measurements
  int dev = 0;
  KLM_CUDA(cudaGetDevice(&dev));
  if (disabled || unusable.done(dev)) return 1;
  auto kern = node_best_cluster_kernel<PUBLISH, SPREAD>;
  if (!attr_set.done(dev)) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) {
      (void)cudaGetLastError();
      unusable.set(dev);
      return 1;
    }
    attr_set.set(dev);
  }
  int csize = 1;
  while (csize < kBestClusterMax && (int64_t)csize * kBestThreads * kBestNodesPerThread < n) csize *= 2;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)csize);
  cfg.blockDim = dim3(kBestThreads);
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)csize;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  const cudaError_t e = cudaLaunchKernelEx(&cfg, kern, battery, flags, n, node_base, vec_ok, best, pub, spr);
  if (e == cudaErrorInvalidValue || e == cudaErrorInvalidConfiguration || e == cudaErrorLaunchOutOfResources ||
      e == cudaErrorNotSupported) {
    (void)cudaGetLastError();  // a 16-CTA cluster cannot be placed on this device / partition
    unusable.set(dev);
    return 1;
  }
  if (e != cudaSuccess) return cuda_fail(e, "node_best_cluster_kernel");
  return KLM_OK;
}

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
  if (n_nodes <= kBestClusterNodes) {
    const int rc = launch_best_cluster<false, false>(battery, flags, n_nodes, node_base, vec_ok, best, ShardPublish{},
                                                     ShardSpread{}, as_stream(stream));
    if (rc != 1) return rc;
  }
  node_best_kernel<false><<<(int)blocks, kBestThreads, 0, as_stream(stream)>>>(
      battery, flags, n_nodes, node_base, vec_ok, best, ticket, part_bat, part_idx, ShardPublish{}, ShardSpread{});
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

// ---------------------------------------------------------------- node-range sharding

extern "C" int64_t klm_dev_shard_slots_bytes(void) { return kShardSlotsBytes; }

extern "C" void klm_shard_range(int64_t n_nodes, int n_parts, int part, int64_t* lo, int64_t* hi) {
  // shards are multiples of 16 nodes, so that every shard's columns keep the 16-byte alignment
  // of the table's (K1 reads them with 128-bit loads)
  const int64_t len = n_parts > 0 ? (ceil_div(n_nodes > 0 ? n_nodes : 0, n_parts) + 15) / 16 * 16 : 0;
  int64_t a = (int64_t)part * len, b = a + len;
  if (a > n_nodes) a = n_nodes;
  if (b > n_nodes) b = n_nodes;
  if (a < 0) a = 0;
  if (b < a) b = a;
  if (lo) *lo = a;
  if (hi) *hi = b;
}

static int launch_publish(const char* who, const double* battery, const uint8_t* flags, int64_t n_nodes,
                          int64_t node_base, int n_parts, int my_part, void* const* peer_slots,
                          double* const* peer_bat, uint8_t* const* peer_flg, uint64_t epoch, int64_t* best,
                          void* scratch, void* stream) {
  if (n_nodes < 0 || !scratch || (n_nodes > 0 && (!battery || !flags)) || n_parts < 1 ||
      n_parts > kMaxShardParts || my_part < 0 || my_part >= n_parts || !peer_slots || epoch == 0) {
    set_error("%s: bad argument (1 <= n_parts <= %d, epoch >= 1)", who, kMaxShardParts);
    return n_parts > kMaxShardParts ? KLM_ELIMIT : KLM_EINVAL;
  }
  ShardPublish pub{};
  ShardSpread spr{};
  const bool spread = peer_bat != nullptr;
  for (int g = 0; g < n_parts; ++g) {
    if (!peer_slots[g] || (spread && g != my_part && n_nodes > 0 && (!peer_bat[g] || !peer_flg || !peer_flg[g]))) {
      set_error("%s: null slot array or replica for part %d", who, g);
      return KLM_EINVAL;
    }
    pub.slots[g] = static_cast<ShardSlot*>(peer_slots[g]);
    if (spread) {
      spr.rep_bat[g] = peer_bat[g];
      spr.rep_flg[g] = peer_flg[g];
    }
  }
  pub.n_parts = n_parts;
  pub.my_part = my_part;
  pub.epoch = epoch;
  char* s = static_cast<char*>(scratch);
  unsigned int* ticket = reinterpret_cast<unsigned int*>(s);
  double* part_bat = reinterpret_cast<double*>(s + 256);
  int64_t* part_idx = reinterpret_cast<int64_t*>(s + 256 + (int64_t)kBestMaxBlocks * 8);
  int64_t blocks = ceil_div(n_nodes, (int64_t)kBestThreads * kBestNodesPerThread);
  if (blocks < 1) blocks = 1;
  if (blocks > kBestMaxBlocks) blocks = kBestMaxBlocks;
  if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
  int vec_ok = aligned16(battery) && aligned16(flags);
  if (spread) {
    // the 128-bit stores land at replica + node_base: every replica base and node_base must keep
    // the alignment too (klm_shard_range hands out multiples of 16 nodes)
    vec_ok = vec_ok && (node_base % 16 == 0);
    for (int g = 0; g < n_parts && vec_ok; ++g)
      if (g != my_part) vec_ok = aligned16(spr.rep_bat[g]) && aligned16(spr.rep_flg[g]);
    if (n_nodes <= kBestClusterNodes) {
      const int rc = launch_best_cluster<true, true>(battery, flags, n_nodes, node_base, vec_ok, best, pub, spr,
                                                     as_stream(stream));
      if (rc != 1) return rc;
    }
    node_best_kernel<true, true><<<(int)blocks, kBestThreads, 0, as_stream(stream)>>>(
        battery, flags, n_nodes, node_base, vec_ok, best, ticket, part_bat, part_idx, pub, spr);
  } else {
    if (n_nodes <= kBestClusterNodes) {
      const int rc = launch_best_cluster<true, false>(battery, flags, n_nodes, node_base, vec_ok, best, pub, spr,
                                                      as_stream(stream));
      if (rc != 1) return rc;
    }
    node_best_kernel<true, false><<<(int)blocks, kBestThreads, 0, as_stream(stream)>>>(
        battery, flags, n_nodes, node_base, vec_ok, best, ticket, part_bat, part_idx, pub, spr);
  }
  KLM_LAUNCH_CHECK("node_best_kernel<publish>");
  return KLM_OK;
}

extern "C" int klm_dev_node_best_publish(const double* battery, const uint8_t* flags, int64_t n_nodes,
                                         int64_t node_base, int n_parts, int my_part,
                                         void* const* peer_slots, uint64_t epoch, int64_t* best,
                                         void* scratch, void* stream) {
  return launch_publish("klm_dev_node_best_publish", battery, flags, n_nodes, node_base, n_parts, my_part,
                        peer_slots, nullptr, nullptr, epoch, best, scratch, stream);
}

extern "C" int klm_dev_node_best_spread(const double* battery, const uint8_t* flags, int64_t n_nodes,
                                        int64_t node_base, int n_parts, int my_part,
                                        void* const* peer_slots, double* const* peer_battery,
                                        uint8_t* const* peer_flags, uint64_t epoch, int64_t* best,
                                        void* scratch, void* stream) {
  if (!peer_battery || !peer_flags) {
    set_error("klm_dev_node_best_spread: null replica arrays");
    return KLM_EINVAL;
  }
  return launch_publish("klm_dev_node_best_spread", battery, flags, n_nodes, node_base, n_parts, my_part,
                        peer_slots, peer_battery, peer_flags, epoch, best, scratch, stream);
}

extern "C" int klm_dev_schedule_sharded(int n_parts, int64_t n_nodes_total,
                                        const double* const* battery_parts,
                                        const uint8_t* const* flags_parts, void* my_slots,
                                        uint64_t epoch, int64_t n_requests, const double* min_battery,
                                        const int32_t* pref_off, const int32_t* pref_ids,
                                        int32_t* out_idx, double* out_score, int table_from_k1, void* stream) {
  if (n_parts < 1 || n_parts > kMaxShardParts || n_nodes_total < 0 || n_nodes_total > 0x7fffffff ||
      !battery_parts || !flags_parts || !my_slots || epoch == 0 || n_requests < 0) {
    set_error("klm_dev_schedule_sharded: bad argument (1 <= n_parts <= %d, epoch >= 1)", kMaxShardParts);
    return n_parts > kMaxShardParts ? KLM_ELIMIT : KLM_EINVAL;
  }
  if (n_requests == 0) return KLM_OK;
  if (!min_battery || !pref_off || !out_idx || !out_score) {
    set_error("klm_dev_schedule_sharded: null request array");
    return KLM_EINVAL;
  }
  ShardView v{};
  int64_t lo = 0, hi = 0;
  klm_shard_range(n_nodes_total, n_parts, 0, &lo, &hi);
  v.shard_len = (int32_t)(hi - lo > 0 ? hi - lo : 1);
  v.inv_len = 1.0f / (float)v.shard_len;
  v.n_total = n_nodes_total;
  v.n_parts = n_parts;
  for (int g = 0; g < n_parts; ++g) {
    klm_shard_range(n_nodes_total, n_parts, g, &lo, &hi);
    if (hi > lo && (!battery_parts[g] || !flags_parts[g])) {
      set_error("klm_dev_schedule_sharded: null node columns for part %d", g);
      return KLM_EINVAL;
    }
    v.bat[g] = battery_parts[g];
    v.flg[g] = flags_parts[g];
  }
  ShardSlot* slots = static_cast<ShardSlot*>(my_slots);
  int* err = reinterpret_cast<int*>(slots + 2 * kMaxShardParts);
  const int grid = grid_for(ceil_div(n_requests, (int64_t)kSchedUnroll), kSchedThreads, 8);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3(kSchedThreads);
  cfg.stream = as_stream(stream);
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  if (table_from_k1) {
    KLM_CUDA(cudaLaunchKernelEx(&cfg, schedule_sharded_kernel<true>, v, (const ShardSlot*)slots,
                                (unsigned long long)epoch, err, n_requests, min_battery, pref_off, pref_ids,
                                out_idx, out_score));
  } else {
    KLM_CUDA(cudaLaunchKernelEx(&cfg, schedule_sharded_kernel<false>, v, (const ShardSlot*)slots,
                                (unsigned long long)epoch, err, n_requests, min_battery, pref_off, pref_ids,
                                out_idx, out_score));
  }
  KLM_LAUNCH_CHECK("schedule_sharded_kernel");
  return KLM_OK;
}

extern "C" int klm_dev_shard_rendezvous(int n_parts, int my_part, void* const* peer_slots, void* my_slots,
                                        uint64_t epoch, void* stream) {
  if (n_parts < 1 || n_parts > kMaxShardParts || my_part < 0 || my_part >= n_parts || !peer_slots || !my_slots ||
      epoch == 0) {
    set_error("klm_dev_shard_rendezvous: bad argument (1 <= n_parts <= %d, epoch >= 1)", kMaxShardParts);
    return n_parts > kMaxShardParts ? KLM_ELIMIT : KLM_EINVAL;
  }
  ShardPublish pub{};
  for (int g = 0; g < n_parts; ++g) {
    if (!peer_slots[g]) {
      set_error("klm_dev_shard_rendezvous: null slot array for part %d", g);
      return KLM_EINVAL;
    }
    pub.slots[g] = static_cast<ShardSlot*>(peer_slots[g]);
  }
  pub.n_parts = n_parts;
  pub.my_part = my_part;
  pub.epoch = epoch;
  ShardSlot* slots = static_cast<ShardSlot*>(my_slots);
  shard_rendezvous_kernel<<<1, 32, 0, as_stream(stream)>>>(pub, slots, reinterpret_cast<int*>(slots + 2 * kMaxShardParts));
  KLM_LAUNCH_CHECK("shard_rendezvous_kernel");
  return KLM_OK;
}

extern "C" int klm_dev_shard_error(const void* my_slots, int* out, void* stream) {
  if (!my_slots || !out) {
    set_error("klm_dev_shard_error: bad argument");
    return KLM_EINVAL;
  }
  const ShardSlot* slots = static_cast<const ShardSlot*>(my_slots);
  KLM_CUDA(cudaMemcpyAsync(out, slots + 2 * kMaxShardParts, sizeof(int), cudaMemcpyDeviceToHost, as_stream(stream)));
  KLM_CUDA(cudaStreamSynchronize(as_stream(stream)));
  return KLM_OK;
}

// ---------------------------------------------------------------- K1 + K2 in one launch

extern "C" int klm_dev_schedule_fused(const double* battery, const uint8_t* flags, int64_t n_nodes,
                                      int64_t node_base, int64_t* best, int64_t n_requests,
                                      const double* min_battery, const int32_t* pref_off,
                                      const int32_t* pref_ids, int32_t* out_idx, double* out_score,
                                      void* scratch, void* stream) {
  if (n_nodes < 0 || n_requests < 0 || !scratch || (n_nodes > 0 && (!battery || !flags))) {
    set_error("klm_dev_schedule_fused: bad argument");
    return KLM_EINVAL;
  }
  if (n_requests > 0 && (!min_battery || !pref_off || !out_idx || !out_score)) {
    set_error("klm_dev_schedule_fused: null request array");
    return KLM_EINVAL;
  }
  if (node_base < 0 || n_nodes + node_base > 0x7fffffffll) {
    set_error("klm_dev_schedule_fused: node indices must fit int32 (n_nodes + node_base <= 2^31 - 1)");
    return KLM_ELIMIT;
  }
  // one wave: as many CTAs as can be resident at once (asked once per device), never more than
  // the work needs
  static PerDeviceOnce cap_known;
  static int cap_blocks[64];
  int dev = 0;
  KLM_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) {
    set_error("klm_dev_schedule_fused: device ordinal %d out of range", dev);
    return KLM_ELIMIT;
  }
  if (!cap_known.done(dev)) {
    int per_sm = 0, sms = 0, coop = 0;
    KLM_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, schedule_fused_kernel, kSchedThreads, 0));
    KLM_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    KLM_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev));
    cap_blocks[dev] = coop ? per_sm * sms : 0;
    cap_known.set(dev);
  }
  int64_t cap = cap_blocks[dev];
  if (cap > kBestMaxBlocks) cap = kBestMaxBlocks;
  if (cap < 1) {
    set_error("klm_dev_schedule_fused: cooperative launch is not available on device %d", dev);
    return KLM_ECUDA;
  }
  const int vec_ok = aligned16(battery) && aligned16(flags);
  const int64_t node_tiles = vec_ok ? ceil_div(n_nodes >> 1, (int64_t)kBestThreads * (kBestNodesPerThread / 2))
                                    : ceil_div(n_nodes, (int64_t)kBestThreads);
  const int64_t req_tiles = ceil_div(n_requests, (int64_t)kSchedThreads * kSchedUnroll);
  int64_t grid = node_tiles > req_tiles ? node_tiles : req_tiles;
  if (grid < 1) grid = 1;
  if (grid > cap) grid = cap;
  int nscan = (int)(node_tiles < grid ? node_tiles : grid);
  if (nscan < 1) nscan = 1;  // (CTA 0 also takes the odd last node)
  char* s = static_cast<char*>(scratch);
  unsigned int* arrive = reinterpret_cast<unsigned int*>(s);
  unsigned int* depart = arrive + 1;
  double* part_bat = reinterpret_cast<double*>(s + 256);
  int64_t* part_idx = reinterpret_cast<int64_t*>(s + 256 + (int64_t)kBestMaxBlocks * 8);
  void* args[] = {&battery, &flags, &n_nodes, &node_base, const_cast<int*>(&vec_ok), &nscan, &arrive, &depart,
                  &part_bat, &part_idx, &best, &n_requests, &min_battery, &pref_off, &pref_ids, &out_idx, &out_score};
  KLM_CUDA(cudaLaunchCooperativeKernel(reinterpret_cast<const void*>(schedule_fused_kernel), dim3((unsigned)grid),
                                       dim3(kSchedThreads), args, 0, as_stream(stream)));
  return KLM_OK;
}
