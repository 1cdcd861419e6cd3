// K7 window_stats, warp variant: one warp per series, the whole window held in
// registers (windows <= 4096 samples), selection without sorting.
//
// No reference code exists for this (the reference keeps only the latest
// snapshot, internal/metrics/manager.go:29, :289-290); definition in
// window_stats.cuh / DESIGN.md, "parity unpinned".
//
// Per series (n valid samples, k = n - rank95 = how many samples are >= p95):
//   1. 128-bit streaming loads: lane l takes float4 l + 32 j (j < NV), so a warp
//      reads its row as NV fully coalesced 512-byte requests. Pads are -inf.
//   2. 128 samples (4 rows spread over the window, one group of 32 per row):
//      the three largest of each group by three rounds of warp-wide max (redux),
//      the four groups interleaved. Pivot = the smallest of the four "third
//      largest" values: about 14 % of the window lies above it and fewer than k
//      samples do with probability ~0.2 %.
//   3. ONE fused pass over the registers: min / max / f64 sum, and every element
//      >= the pivot is appended to a per-lane list in shared memory.
//   4. The lists are packed into one contiguous array and narrowed by counting
//      passes that test two pivots at a time (first the sample's own order
//      statistics, then interpolation on the counts) until <= 32 candidates
//      remain, which are ranked exactly against each other. Brackets that cannot
//      shrink because of ties are peeled one distinct value at a time.
//   Heavy ties at the pivot (quantised channels, mostly-zero counters) are settled
//   by counting passes over the registers at the tied value; anything the
//   shortcuts cannot place (list overflow, unlucky pivot) falls back to an exact
//   bisection over the order-preserving key space.
// Everything is exact: every branch ends on an input element whose rank has
// been counted, so the result is bit-identical to sorting.
#include "common.cuh"
#include "window_stats.cuh"

namespace klm {

#ifndef KLM_WSW_MIN_BLOCKS
#define KLM_WSW_MIN_BLOCKS 3
#endif
constexpr int kWswWarps = 4;               // warps (series) per CTA; registers are allocated in
                                           // units of 4 warps, so smaller CTAs waste them
constexpr int kWswThreads = kWswWarps * 32;
constexpr int kSampledMinRows = 7;         // below 7 full rows (896 samples) list everything
constexpr int kLaneCap = 64;               // per-lane list entries (wraps beyond: overflow -> bisect)
constexpr int kLaneListBytes = kLaneCap * 128;  // [entry][lane] f32; power of two, self-aligned
// shared memory: first the lane lists (8 KB per warp, each on an 8 KB boundary), then per
// warp 64 candidates + a counter (512 B)
constexpr int kWarpAux = 512;
constexpr size_t kWswSmemBytes =
    (size_t)kWswWarps * (kLaneListBytes + kWarpAux) + kLaneListBytes;  // + alignment slack

// if (x > lox) { lane_list[c / 128 mod 64] = x; c += 128; }
// lb = 8 KB-aligned list base | lane * 4, so the slot address is (c & 0x1F80) | lb.
// The address lives in a temporary: the store never holds up the counter.
__device__ __forceinline__ void append_if_gt(uint32_t& c, uint32_t lb, float x, float lox) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b32 t;\n\t"
      "setp.gt.f32 p, %2, %3;\n\t"
      "lop3.b32 t, %0, 0x1F80, %1, 0xEA;\n\t"
      "@p st.shared.f32 [t], %2;\n\t"
      "@p add.u32 %0, %0, 128;\n\t}"
      : "+r"(c)
      : "r"(lb), "f"(x), "f"(lox));
}

template <int NV>
struct Window {
  static constexpr int NE = NV * 4;
  float v[NE];   // element 4 j + c = sample 4 (lane + 32 j) + c, or -inf
  float tail;    // lane c < (n & 3) holds sample 4 (n / 4) + c, else -inf

  // number of samples >= p (all lanes get the warp total)
  __device__ __forceinline__ int count_ge(float p) const {
    int c0 = 0, c1 = 0, c2 = 0, c3 = 0;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      inc_if_ge(c0, v[4 * j + 0], p);
      inc_if_ge(c1, v[4 * j + 1], p);
      inc_if_ge(c2, v[4 * j + 2], p);
      inc_if_ge(c3, v[4 * j + 3], p);
    }
    inc_if_ge(c0, tail, p);
    return warp_sum_i32((c0 + c1) + (c2 + c3));
  }

  // exact k-th largest by bisection over keys; needs count_ge(key_f32(lo)) >= k >
  // count_ge(key_f32(hi)) where hi may be one past the largest key present. `first`
  // (if strictly inside) is probed before the midpoints. Inlined at ONE call site: the
  // window must stay in registers.
  __device__ __forceinline__ float select_bisect(uint32_t lo, uint32_t hi, int k,
                                                 uint32_t first) const {
    uint32_t mid = (first > lo && first < hi) ? first : lo + ((hi - lo) >> 1);
#pragma unroll 1
    while (hi - lo > 1u) {
      if (count_ge(key_f32(mid)) >= k) lo = mid; else hi = mid;
      mid = lo + ((hi - lo) >> 1);
    }
    return key_f32(lo);
  }

  // append(x > lox) only: the retry after a pivot that listed fewer than k elements
  __device__ __forceinline__ void append_pass(float lox, uint32_t lb, uint32_t& c) const {
#pragma unroll
    for (int i = 0; i < NE; ++i) append_if_gt(c, lb, v[i], lox);
    append_if_gt(c, lb, tail, lox);
  }

  // One pass over the registers: min / max / f64 sum of the valid samples, and append
  // (x > lox) to the lane's list. FULL: every row but the last is valid in every lane,
  // so the pass is straight-line code the scheduler can interleave freely.
  template <bool FULL>
  __device__ __forceinline__ void fused_pass(int lane, int n, int nfull, int jr, float lox,
                                             uint32_t lb, uint32_t& c, float& mn, float& mx,
                                             double& sum) const {
    float mn0 = POS_INF, mn1 = POS_INF, mx0 = NEG_INF, mx1 = NEG_INF;
    double sum0 = 0.0, sum1 = 0.0;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const float a = v[4 * j], b = v[4 * j + 1], cc = v[4 * j + 2], d = v[4 * j + 3];
      const bool uniform_valid = FULL ? (j < NV - 1) : (j < jr);
      if (uniform_valid) {
        if (j & 1) {
          mn1 = fminf(fminf(mn1, a), fminf(b, fminf(cc, d)));
          mx1 = fmaxf(fmaxf(mx1, a), fmaxf(b, fmaxf(cc, d)));
          sum1 += ((double)a + (double)b) + ((double)cc + (double)d);
        } else {
          mn0 = fminf(fminf(mn0, a), fminf(b, fminf(cc, d)));
          mx0 = fmaxf(fmaxf(mx0, a), fmaxf(b, fmaxf(cc, d)));
          sum0 += ((double)a + (double)b) + ((double)cc + (double)d);
        }
      } else if (lane + 32 * j < nfull) {  // boundary row: some lanes hold data
        mn0 = fminf(fminf(mn0, a), fminf(b, fminf(cc, d)));
        mx0 = fmaxf(fmaxf(mx0, a), fmaxf(b, fmaxf(cc, d)));
        sum0 += ((double)a + (double)b) + ((double)cc + (double)d);
      }
      append_if_gt(c, lb, a, lox);  // pads are -inf: never appended
      append_if_gt(c, lb, b, lox);
      append_if_gt(c, lb, cc, lox);
      append_if_gt(c, lb, d, lox);
    }
    if (lane < (n & 3)) {
      mn0 = fminf(mn0, tail);
      mx0 = fmaxf(mx0, tail);
      sum0 += (double)tail;
    }
    append_if_gt(c, lb, tail, lox);
    mn = fminf(mn0, mn1);
    mx = fmaxf(mx0, mx1);
    sum = sum0 + sum1;
  }
};

template <int NV>
__global__ void __launch_bounds__(kWswThreads, KLM_WSW_MIN_BLOCKS)
window_stats_warp_kernel(const float* __restrict__ ring, const uint32_t* __restrict__ seen,
                         int32_t window, int32_t pitch, int64_t s0, int64_t n_series,
                         float* __restrict__ out_min, float* __restrict__ out_max,
                         double* __restrict__ out_mean, float* __restrict__ out_p95) {
  extern __shared__ unsigned char s_raw[];

  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int64_t o = (int64_t)blockIdx.x * kWswWarps + wib;  // output slot
  if (o >= n_series) return;                                  // no block-wide barrier below
  const int64_t s = s0 + o;
  const float* row = ring + s * (int64_t)pitch;
  const uint32_t sn = seen[s];  // consumed only after the row loads below are in flight

  // shared memory of this warp; the lane lists must start on their own size boundary
  const uint32_t raw_s = (uint32_t)__cvta_generic_to_shared(s_raw);
  const uint32_t base_s = (raw_s + (uint32_t)kLaneListBytes - 1u) & ~((uint32_t)kLaneListBytes - 1u);
  unsigned char* base_g = s_raw + (base_s - raw_s);
  float* lst = reinterpret_cast<float*>(base_g + wib * kLaneListBytes) + lane;  // entry e: lst[e * 32]
  unsigned char* aux = base_g + kWswWarps * kLaneListBytes + wib * kWarpAux;
  float* fin = reinterpret_cast<float*>(aux);
  const uint32_t lb = (base_s + (uint32_t)wib * kLaneListBytes) | ((uint32_t)lane << 2);

  // ---- 1. load. The addresses do not depend on seen[s]: every float4 inside the row
  // pitch is fetched (stale slots beyond n are masked below), so the row loads and the
  // seen[] load share one memory round trip.
  Window<NV> w;
  {
    const int pfull = pitch >> 2;
    const float* p = row + 4 * lane;
    if (pfull >= 32 * (NV - 1)) {  // usual case: all rows but the last lie inside the pitch
#pragma unroll
      for (int j = 0; j < NV - 1; ++j) {
        const float4 t = ld_stream_f4(p + 128 * j);
        w.v[4 * j + 0] = t.x; w.v[4 * j + 1] = t.y; w.v[4 * j + 2] = t.z; w.v[4 * j + 3] = t.w;
      }
      float4 t = make_float4(NEG_INF, NEG_INF, NEG_INF, NEG_INF);
      if (lane + 32 * (NV - 1) < pfull) t = ld_stream_f4(p + 128 * (NV - 1));
      w.v[4 * NV - 4] = t.x; w.v[4 * NV - 3] = t.y; w.v[4 * NV - 2] = t.z; w.v[4 * NV - 1] = t.w;
    } else {
#pragma unroll
      for (int j = 0; j < NV; ++j) {
        float4 t = make_float4(NEG_INF, NEG_INF, NEG_INF, NEG_INF);
        if (lane + 32 * j < pfull) t = ld_stream_f4(p + 128 * j);
        w.v[4 * j + 0] = t.x; w.v[4 * j + 1] = t.y; w.v[4 * j + 2] = t.z; w.v[4 * j + 3] = t.w;
      }
    }
  }
  // While the loads are in flight: fill the first kListPrefill entries of this lane's list with -inf,
  // so the counting passes need no padding later (longer lists are padded when they occur).
#pragma unroll
  for (int e = 0; e < kListPrefill; ++e) lst[e * 32] = NEG_INF;
  const int n = sn < (uint32_t)window ? (int)sn : window;
  if (n == 0) {
    if (lane == 0) {
      if (out_min) out_min[o] = 0.f;
      if (out_max) out_max[o] = 0.f;
      if (out_mean) out_mean[o] = 0.0;
      if (out_p95) out_p95[o] = 0.f;
    }
    return;
  }
  // Ask L2 for the row that a warp will want about one residency later (12 warps per SM
  // are in flight): one bulk-prefetch instruction covers the whole row.
  if (lane == 0) {
    const int64_t later = o + (int64_t)kNumSMs * 12;
    if (later < n_series) {
      const uint32_t bytes = ((uint32_t)n * 4u + 15u) & ~15u;
      asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;"
                   :: "l"(ring + (s0 + later) * (int64_t)pitch), "r"(bytes) : "memory");
    }
  }
  const int nfull = n >> 2;   // complete float4 groups
  const int jr = nfull >> 5;  // rows valid in every lane
  const bool full = jr >= NV - 1;
  // mask the slots at and beyond n: rows below jr are valid in every lane
  if (full) {
    if (!(lane + 32 * (NV - 1) < nfull))
      w.v[4 * NV - 4] = w.v[4 * NV - 3] = w.v[4 * NV - 2] = w.v[4 * NV - 1] = NEG_INF;
  } else {
#pragma unroll
    for (int j = 0; j < NV; ++j)
      if (j >= jr && !(lane + 32 * j < nfull))
        w.v[4 * j] = w.v[4 * j + 1] = w.v[4 * j + 2] = w.v[4 * j + 3] = NEG_INF;
  }
  w.tail = (lane < (n & 3)) ? row[4 * nfull + lane] : NEG_INF;

  const int k = n - p95_rank(n);
  bool bisect = false;             // finish with the exact bisection over the registers
  uint32_t bis_lo = f32_key(NEG_INF), bis_hi = 0, bis_first = 0;
  float lox = POS_INF;             // list predicate: x > lox (+inf: list nothing)
  float hint_lo = 0.f, hint_hi = 0.f;  // first pair of pivots (sample order statistics)
  float lox_retry = POS_INF;       // lower pivot for one retry of the append (+inf: none)
  bool tie_mode = false, retried = false;
  bool tie_hit = false;            // the heavy-tie test already found the answer
  float tie_val = 0.f;
  const bool want_p95 = out_p95 != nullptr;
  const bool sampled = want_p95 && jr >= kSampledMinRows;

  // ---- 2. sample pivot: four groups of 32 (one sample row each), top three of each
  if (sampled) {
    float sm[4];
    if (full) {  // the rows are known at compile time
#pragma unroll
      for (int q = 0; q < 4; ++q) sm[q] = w.v[4 * (((2 * q + 1) * (NV - 1)) >> 3) + q];
    } else {
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int jq = ((2 * q + 1) * jr) >> 3;
        float t = NEG_INF;
#pragma unroll
        for (int j = 0; j < NV; ++j) t = (j == jq) ? w.v[4 * j + q] : t;
        sm[q] = t;
      }
    }
    uint32_t h[4];
#pragma unroll
    for (int g = 0; g < 4; ++g) h[g] = f32_key(sm[g]);
    const uint32_t lane_bit = 1u << lane;
    uint32_t t1 = 0xffffffffu, t2 = 0xffffffffu, t3 = 0xffffffffu;  // min over groups of rank 1/2/3
#pragma unroll
    for (int r = 1; r <= 3; ++r) {
      uint32_t m[4];
#pragma unroll
      for (int g = 0; g < 4; ++g) m[g] = warp_max_u32(h[g]);
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        const unsigned who = __ballot_sync(0xffffffffu, h[g] == m[g]);
        if ((who & (0u - who)) == lane_bit) h[g] = 0u;  // retire one copy
      }
      const uint32_t mm = min(min(m[0], m[1]), min(m[2], m[3]));
      if (r == 1) t1 = mm; else if (r == 2) t2 = mm; else t3 = mm;
    }
    // pivot: the smallest "second largest" (about 10 % of the window above it; fewer than k
    // with probability ~5 %, in which case the append is redone from the "third largest")
    const float V = key_f32(t2);
    hint_lo = key_f32(t2);
    hint_hi = key_f32(t1);
    if (t3 < t2) lox_retry = next_down(key_f32(t3));
    if (t1 == t2 || t2 == t3) {
      tie_mode = true;
      // heavy tie at V: place k against the block of copies of V (two counts, one code copy)
      int c_ge = 0, c_gt = 0;
#pragma unroll 1
      for (int t = 0; t < 2; ++t) {
        const int c = w.count_ge(t ? next_up(V) : V);
        if (t) c_gt = c; else c_ge = c;
      }
      if (c_gt >= k) {
        lox = V;                       // the answer is above V: list the elements > V
      } else if (c_ge >= k) {
        tie_hit = true; tie_val = V;   // the answer is V itself
      } else if (t3 < t2) {
        lox = lox_retry;               // below V: list from the next sample statistic down
        lox_retry = POS_INF;
      } else {
        bisect = true; bis_hi = t2;    // the answer is below V
      }
    } else {
      lox = next_down(V);  // x > lox  <=>  x >= V
    }
  } else if (want_p95) {
    lox = NEG_INF;  // short window: list every sample
  }

  // ---- 3. fused pass: min / max / f64 sum + append(x > lox)
  float mn, mx;
  double sum;
  uint32_t c128 = 0;
  if (full) w.template fused_pass<true>(lane, n, nfull, jr, lox, lb, c128, mn, mx, sum);
  else w.template fused_pass<false>(lane, n, nfull, jr, lox, lb, c128, mn, mx, sum);
  asm volatile("" ::: "memory");  // the appends above are plain shared-memory stores
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, off));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
    sum += __shfl_xor_sync(0xffffffffu, sum, off);
  }
  if (lane == 0) {
    if (out_min) out_min[o] = mn;
    if (out_max) out_max[o] = mx;
    if (out_mean) out_mean[o] = sum / (double)n;
  }
  if (!want_p95) return;

  // ---- 4. p95 = k-th largest
  float result = 0.f;
  do {  // single-trip block: `break` = result (or the bisect request) is final
    if (bisect) break;
    if (tie_hit) { result = tie_val; break; }
    if (mn == mx) { result = mx; break; }  // constant series
    int cnt = (int)(c128 >> 7);
    int total = warp_sum_i32(cnt);
    int maxcnt = warp_max_i32(cnt);
    if (total < k && !sampled) { result = NEG_INF; break; }  // lox = -inf: only -inf samples are unlisted
    if (total < k && !tie_mode && lox_retry != POS_INF) {
      // the pivot was too high: redo the append (only) from the next sample statistic down
#pragma unroll
      for (int e = 0; e < kListPrefill; ++e) lst[e * 32] = NEG_INF;
      lox = lox_retry;
      c128 = 0;
      w.append_pass(lox, lb, c128);
      asm volatile("" ::: "memory");
      cnt = (int)(c128 >> 7);
      total = warp_sum_i32(cnt);
      maxcnt = warp_max_i32(cnt);
      retried = true;
    }
    if (total < k || maxcnt > kLaneCap) {
      // unlucky pivot or list overflow: exact bisection over the registers
      bisect = true; bis_hi = f32_key(mx) + 1u; bis_first = 0u;
      break;
    }
    result = select_from_lists(lst, fin, lane, cnt, total, maxcnt, k, sampled ? next_up(lox) : mn, mx,
                               retried, hint_lo, hint_hi);
  } while (false);
  if (bisect) result = w.select_bisect(bis_lo, bis_hi, k, bis_first);
  if (lane == 0) out_p95[o] = result;
}

template <int NV>
static int launch_warp(const float* ring, const uint32_t* seen, int32_t window, int32_t pitch,
                       int64_t s0, int64_t n_series, float* out_min, float* out_max,
                       double* out_mean, float* out_p95, cudaStream_t st) {
  const int64_t blocks = ceil_div(n_series, kWswWarps);
  if (blocks > 0x7fffffff) {
    set_error("klm_dev_window_stats: too many series for one launch");
    return KLM_ELIMIT;
  }
  static bool attr_set[64] = {};
  int dev = 0;
  KLM_CUDA(cudaGetDevice(&dev));
  if (!(dev < 64 && attr_set[dev])) {
    // 3 CTAs x 57 KB per SM: opt in above 48 KB and ask for the largest shared-memory carve-out
    KLM_CUDA(cudaFuncSetAttribute(window_stats_warp_kernel<NV>,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kWswSmemBytes));
    KLM_CUDA(cudaFuncSetAttribute(window_stats_warp_kernel<NV>,
                                  cudaFuncAttributePreferredSharedMemoryCarveout,
                                  (int)cudaSharedmemCarveoutMaxShared));
    if (dev < 64) attr_set[dev] = true;
  }
  window_stats_warp_kernel<NV><<<(unsigned)blocks, kWswThreads, kWswSmemBytes, st>>>(
      ring, seen, window, pitch, s0, n_series, out_min, out_max, out_mean, out_p95);
  KLM_LAUNCH_CHECK("window_stats_warp_kernel");
  return KLM_OK;
}

// Windows up to 4096 samples take the warp kernel; returns 1 if it did not handle the call.
int launch_window_stats_warp(const float* ring, const uint32_t* seen, int32_t window,
                             int32_t pitch, int64_t s0, int64_t s1, float* out_min,
                             float* out_max, double* out_mean, float* out_p95, cudaStream_t st) {
  static_assert(KLM_WSW_MIN_BLOCKS * (kWswSmemBytes + 1024) <= 227 * 1024, "the CTAs per SM must fit");
  const int64_t n = s1 - s0;
  if (window <= 8 * 128) return launch_warp<8>(ring, seen, window, pitch, s0, n, out_min, out_max, out_mean, out_p95, st);
  if (window <= 16 * 128) return launch_warp<16>(ring, seen, window, pitch, s0, n, out_min, out_max, out_mean, out_p95, st);
  if (window <= 29 * 128) return launch_warp<29>(ring, seen, window, pitch, s0, n, out_min, out_max, out_mean, out_p95, st);
  if (window <= 32 * 128) return launch_warp<32>(ring, seen, window, pitch, s0, n, out_min, out_max, out_mean, out_p95, st);
  return 1;
}

}  // namespace klm
