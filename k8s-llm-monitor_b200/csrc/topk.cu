// Top-k scheduling: the k best candidates of every request in the order
// a stable descending sort leaves them (score desc, list index asc) -- controller.go:160 sorts
// the whole candidate list and then keeps only [0] (:161); examples/multi-pod-request.yaml asks
// for two GPUs that the reference never assigns. k = 1 is exactly klm_dev_schedule. No reference
// semantics beyond [0]: parity unpinned; the oracle is the literal candidate list + stable sort
// cut at k.
//
// Decomposition (as for k = 1): a node that is not preferred by the request scores its battery,
// so if it is not among the k best ELIGIBLE nodes of the whole table (battery desc, index asc)
// it has k nodes ahead of it for every request. The k best of a request are therefore among
// T = that global top-k list (request-independent: one reduction per snapshot) and the
// request's preferred nodes (+10).
#include "common.cuh"

namespace klm {

constexpr int kTopkMax = 8;
constexpr int kTkThreads = 256;
constexpr int kTkPerThread = 16;
constexpr int kTkMaxBlocks = 2048;
// scratch: [0,256) ticket; then per block kTopkMax (battery f64, index i64) pairs
constexpr int64_t kTkScratchBytes = 256 + (int64_t)kTkMaxBlocks * kTopkMax * 16;

struct Cand {
  double bat;
  int64_t idx;  // -1: none
};
// strict order of the global list: larger battery first, lower index on ties
__device__ __forceinline__ bool cand_before(double ab, int64_t ai, double bb, int64_t bi) {
  return ab > bb || (ab == bb && ai < bi);
}
__device__ __forceinline__ Cand cand_better(Cand a, Cand b) {
  if (b.idx < 0) return a;
  if (a.idx < 0) return b;
  return cand_before(b.bat, b.idx, a.bat, a.idx) ? b : a;
}
__device__ __forceinline__ Cand block_cand_best(Cand v) {
  __shared__ double s_bat[kTkThreads / 32];
  __shared__ int64_t s_idx[kTkThreads / 32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    Cand w{__shfl_xor_sync(0xffffffffu, v.bat, o), __shfl_xor_sync(0xffffffffu, v.idx, o)};
    v = cand_better(v, w);
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();  // the arrays may still be read from the previous round
  if (lane == 0) {
    s_bat[warp] = v.bat;
    s_idx[warp] = v.idx;
  }
  __syncthreads();
  Cand w{lane < kTkThreads / 32 ? s_bat[lane] : 0.0, lane < kTkThreads / 32 ? s_idx[lane] : -1};
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    Cand u{__shfl_xor_sync(0xffffffffu, w.bat, o), __shfl_xor_sync(0xffffffffu, w.idx, o)};
    w = cand_better(w, u);
  }
  return w;  // the same in every thread
}

// k rounds of "best candidate strictly after the previous pick" over this thread's values
template <int NVAL>
__device__ __forceinline__ void block_topk(const double (&bat)[NVAL], const int64_t (&idx)[NVAL], int k,
                                           double* out_bat, int64_t* out_idx) {
  double pb = 0.0;
  int64_t pi = -1;  // previous pick (none yet)
  for (int r = 0; r < k; ++r) {
    Cand me{0.0, -1};
#pragma unroll
    for (int j = 0; j < NVAL; ++j) {
      const bool after = pi < 0 || cand_before(pb, pi, bat[j], idx[j]);
      if (idx[j] >= 0 && after && (me.idx < 0 || cand_before(bat[j], idx[j], me.bat, me.idx))) {
        me.bat = bat[j];
        me.idx = idx[j];
      }
    }
    const Cand w = block_cand_best(me);
    if (threadIdx.x == 0) {
      out_bat[r] = w.bat;
      out_idx[r] = w.idx;
    }
    if (w.idx < 0) {  // fewer than k candidates: the rest stays empty
      for (int q = r + 1; q < k && threadIdx.x == 0; ++q) {
        out_bat[q] = 0.0;
        out_idx[q] = -1;
      }
      return;
    }
    pb = w.bat;
    pi = w.idx;
  }
}

// K1 for top-k: top[0..k) = the k best eligible nodes, (battery, global index) pairs as
// top[2 j] = index (int64, -1: none), top[2 j + 1] = battery bits.
__global__ void __launch_bounds__(kTkThreads)
node_topk_kernel(const double* __restrict__ battery, const uint8_t* __restrict__ flags, int64_t n,
                 int64_t node_base, int k, int64_t* __restrict__ top, unsigned int* __restrict__ ticket,
                 double* __restrict__ part_bat, int64_t* __restrict__ part_idx) {
  double b[kTkPerThread];
  int64_t id[kTkPerThread];
  const int64_t base = (int64_t)blockIdx.x * kTkThreads * kTkPerThread;
#pragma unroll
  for (int j = 0; j < kTkPerThread; ++j) {
    const int64_t i = base + (int64_t)j * kTkThreads + threadIdx.x;  // coalesced
    const bool ok = i < n && (flags[i] & 3u) == 3u;
    b[j] = ok ? battery[i] : 0.0;
    id[j] = ok ? i : -1;
  }
  __shared__ double s_ob[kTopkMax];
  __shared__ int64_t s_oi[kTopkMax];
  block_topk(b, id, k, s_ob, s_oi);
  __shared__ bool s_last;
  __syncthreads();
  if (threadIdx.x < k) {
    part_bat[blockIdx.x * kTopkMax + threadIdx.x] = s_ob[threadIdx.x];
    part_idx[blockIdx.x * kTopkMax + threadIdx.x] = s_oi[threadIdx.x];
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    s_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  // merge the partial lists: again k rounds, every thread walking a strided share of the entries
  // (at most 16 K of them, in L2)
  const int total = (int)gridDim.x * kTopkMax;
  double pb = 0.0;
  int64_t pi = -1;
  for (int r = 0; r < k; ++r) {
    Cand me{0.0, -1};
    for (int e = threadIdx.x; e < total; e += kTkThreads) {
      if ((e % kTopkMax) >= k) continue;
      const int64_t ci = __ldcg(part_idx + e);
      const double cb = __ldcg(part_bat + e);
      const bool after = pi < 0 || cand_before(pb, pi, cb, ci);
      if (ci >= 0 && after && (me.idx < 0 || cand_before(cb, ci, me.bat, me.idx))) {
        me.bat = cb;
        me.idx = ci;
      }
    }
    const Cand w = block_cand_best(me);
    if (threadIdx.x == 0) {
      s_ob[r] = w.bat;
      s_oi[r] = w.idx;
    }
    pb = w.bat;
    pi = w.idx;
    if (w.idx < 0) {
      for (int q = r + 1; q < k && threadIdx.x == 0; ++q) s_oi[q] = -1;
      break;
    }
  }
  __syncthreads();
  if (threadIdx.x < k) {
    top[2 * threadIdx.x] = s_oi[threadIdx.x] < 0 ? -1 : s_oi[threadIdx.x] + node_base;
    top[2 * threadIdx.x + 1] = __double_as_longlong(s_oi[threadIdx.x] < 0 ? 0.0 : s_ob[threadIdx.x]);
  }
  if (threadIdx.x == 0) *ticket = 0;
}

// K2 for top-k: one thread per request; a sorted list of at most k (score, index) entries in
// registers, fed first by the request's preferred nodes (+10) and then by the global list T.
__global__ void __launch_bounds__(256)
schedule_topk_kernel(const double* __restrict__ battery, const uint8_t* __restrict__ flags, int64_t n,
                     int64_t node_base, const int64_t* __restrict__ top, int k, int64_t n_requests,
                     const double* __restrict__ min_battery, const int32_t* __restrict__ pref_off,
                     const int32_t* __restrict__ pref_ids, int32_t* __restrict__ out_idx,
                     double* __restrict__ out_score) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n_requests; r += stride) {
    const double mb = min_battery[r];
    const bool gate = mb > 0;  // controller.go:194
    const int32_t k0 = pref_off[r], k1 = pref_off[r + 1];
    double ls[kTopkMax];
    int64_t li[kTopkMax];
    int cnt = 0;
#pragma unroll
    for (int j = 0; j < kTopkMax; ++j) {
      ls[j] = 0.0;
      li[j] = -1;
    }
    // insert (s, i) keeping (score desc, index asc); an entry already present is skipped
    auto insert = [&](double sc, int64_t i) {
      bool dup = false;
#pragma unroll
      for (int j = 0; j < kTopkMax; ++j) dup |= (j < cnt && li[j] == i);
      if (dup) return;
      if (cnt == k && !(sc > ls[k - 1] || (sc == ls[k - 1] && i < li[k - 1]))) return;
      double cs = sc;
      int64_t ci = i;
#pragma unroll
      for (int j = 0; j < kTopkMax; ++j) {
        if (j < k) {
          const bool here = j >= cnt || cs > ls[j] || (cs == ls[j] && ci < li[j]);
          if (here) {  // put the carried entry here and carry the old one on
            const double ts = ls[j];
            const int64_t ti = li[j];
            ls[j] = cs;
            li[j] = ci;
            cs = ts;
            ci = ti;
            if (j >= cnt) break;
          }
        }
      }
      if (cnt < k) ++cnt;
    };
    for (int32_t q = k0; q < k1; ++q) {  // preferred nodes: :205-208
      const int64_t pk = pref_ids[q];
      const int64_t lp = pk - node_base;
      if (lp < 0 || lp >= n) continue;
      if ((flags[lp] & 3u) != 3u) continue;  // :190, :198
      const double bk = battery[lp];
      if (gate && bk < mb) continue;         // :194
      insert(bk + 10, pk);
    }
    for (int j = 0; j < k; ++j) {  // the global list, best first
      const int64_t gi = top[2 * j];
      if (gi < 0) break;
      const double gb = __longlong_as_double(top[2 * j + 1]);
      if (gate && gb < mb) break;  // sorted by battery: the rest fails the gate too
      bool pref = false;
      for (int32_t q = k0; q < k1 && !pref; ++q) pref = pref_ids[q] == gi;
      if (pref) continue;  // it went in above, with its +10
      insert(gb, gi);
    }
    for (int j = 0; j < k; ++j) {
      out_idx[r * k + j] = j < cnt ? (int32_t)li[j] : -1;
      out_score[r * k + j] = j < cnt ? ls[j] : 0.0;
    }
  }
}

}  // namespace klm

using namespace klm;

extern "C" int klm_topk_max(void) { return kTopkMax; }
extern "C" int64_t klm_dev_node_topk_scratch_bytes(void) { return kTkScratchBytes; }

extern "C" int klm_dev_node_topk(const double* battery, const uint8_t* flags, int64_t n_nodes, int64_t node_base,
                                 int k, int64_t* top, void* scratch, void* stream) {
  if (n_nodes < 0 || k < 1 || k > kTopkMax || !top || !scratch || (n_nodes > 0 && (!battery || !flags))) {
    set_error("klm_dev_node_topk: bad argument (1 <= k <= %d)", kTopkMax);
    return k > kTopkMax ? KLM_ELIMIT : KLM_EINVAL;
  }
  int64_t blocks = ceil_div(n_nodes, (int64_t)kTkThreads * kTkPerThread);
  if (blocks < 1) blocks = 1;
  if (blocks > kTkMaxBlocks) {
    set_error("klm_dev_node_topk: %lld nodes exceed %lld", (long long)n_nodes,
              (long long)kTkMaxBlocks * kTkThreads * kTkPerThread);
    return KLM_ELIMIT;
  }
  char* s = static_cast<char*>(scratch);
  node_topk_kernel<<<(int)blocks, kTkThreads, 0, as_stream(stream)>>>(
      battery, flags, n_nodes, node_base, k, top, reinterpret_cast<unsigned int*>(s),
      reinterpret_cast<double*>(s + 256), reinterpret_cast<int64_t*>(s + 256 + (int64_t)kTkMaxBlocks * kTopkMax * 8));
  KLM_LAUNCH_CHECK("node_topk_kernel");
  return KLM_OK;
}

extern "C" int klm_dev_schedule_topk(const double* battery, const uint8_t* flags, int64_t n_nodes, int64_t node_base,
                                     const int64_t* top, int k, int64_t n_requests, const double* min_battery,
                                     const int32_t* pref_off, const int32_t* pref_ids, int32_t* out_idx,
                                     double* out_score, void* stream) {
  if (n_nodes < 0 || n_requests < 0 || k < 1 || k > kTopkMax || !top) {
    set_error("klm_dev_schedule_topk: bad argument (1 <= k <= %d)", kTopkMax);
    return k > kTopkMax ? KLM_ELIMIT : KLM_EINVAL;
  }
  if (n_requests == 0) return KLM_OK;
  if (!min_battery || !pref_off || !out_idx || !out_score) {
    set_error("klm_dev_schedule_topk: null request array");
    return KLM_EINVAL;
  }
  int64_t blocks = ceil_div(n_requests, 256);
  if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
  schedule_topk_kernel<<<(int)blocks, 256, 0, as_stream(stream)>>>(battery, flags, n_nodes, node_base, top, k,
                                                                  n_requests, min_battery, pref_off, pref_ids,
                                                                  out_idx, out_score);
  KLM_LAUNCH_CHECK("schedule_topk_kernel");
  return KLM_OK;
}
