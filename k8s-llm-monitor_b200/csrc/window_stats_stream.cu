// K7 window_stats, streaming variant: one warp per series, the window is NOT kept on chip.
//
// No reference code exists for this (the reference keeps only the latest snapshot,
// internal/metrics/manager.go:29, :289-290); definition in window_stats.cuh / DESIGN.md,
// "parity unpinned".
//
// The register-resident warp kernel (window_stats_warp.cu) holds a whole 14.4 KB window per
// warp, which caps an SM at 16-20 warps; it then runs at the pace its few warps can hide
// latency, not at the pace of HBM ("K7 in detail"). This kernel keeps only what the
// statistics need:
//   1. 128 samples (4 rows spread over the window, one element per lane and row) are fetched
//      first, together with the first rows of the pass; the pivot is the smallest "second
//      largest" of the four groups of 32, exactly as in the warp kernel.
//   2. ONE pass streams the rows through two register buffers of four rows (the loads of one
//      buffer are in flight while the other is consumed): min / max / exact f64 sum, and every
//      element > the pivot is appended to the lane's list in shared memory.
//   3. select_from_lists (window_stats.cuh) finds the k-th largest among the listed elements.
//   The rare paths that need the window again (heavy ties at the pivot, a pivot that listed
//   fewer than k elements, list overflow) re-read the row -- it is still in L2 -- in small
//   rolled loops kept out of line.
// 64-80 registers per thread instead of 128, no window-size template (any window, one
// instantiation), and 24-32 warps per SM. Exact: every exit is an input element whose rank has
// been counted, so the result is bit-identical to sorting.
#include "common.cuh"
#include "window_stats.cuh"

namespace klm {

#ifdef KLM_WSS_STATS
// tuning builds only: how many series took which way through the kernel
// [0] tie value held  [1] bracket held (one count + sort)  [2] bracket listed but missed (full select)
// [3] sampled pivot  [4] counting fallback (narrow)  [5] short window / list everything  [6] retried append
__device__ unsigned long long g_wss_paths[8];
#define KLM_WSS_COUNT(i) do { if (lane == 0) atomicAdd(&g_wss_paths[i], 1ull); } while (0)
#else
#define KLM_WSS_COUNT(i) do { } while (0)
#endif

constexpr int kWssWarps = 4;  // warps (series) per CTA
constexpr int kWssThreads = kWssWarps * 32;
constexpr int kWssAux = 256;  // per warp: the 64 candidates of the final sort
constexpr int kWssSampledMinRows = 7;  // below 7 full rows (896 samples) list everything

template <int CAP, int RS = 0, int D = 0>
struct WssCfg {
  static constexpr int kListBytes = CAP * 128;  // [entry][lane] f32
  static constexpr int kClamp = CAP - 4;        // a lane is held at this count (four appends may follow)
  static constexpr int kPrefill = CAP >= kListPrefill + 8 ? kListPrefill : 16;  // entries pre-filled with -inf
  static constexpr int kRingBytes = RS * D * 512;  // per warp: D stages of RS rows (cp.async landing ring)
  static constexpr size_t kSmem = (size_t)kWssWarps * (kListBytes + kWssAux + kRingBytes);
  static_assert(CAP >= kPrefill + 8 && kPrefill % 4 == 0, "the prefilled part must fit");
};

// 16 bytes global -> shared, asynchronously (LDGSTS), past L1; groups are per thread, and every
// lane reads back only what it copied itself, so no warp barrier is involved.
__device__ __forceinline__ void cp_async16(uint32_t dst, const float* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ float4 lds_f4(uint32_t a) {
  float4 r;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "r"(a)
               : "memory");
  return r;
}

// if (x > lox) { list[c / 128] = x; c += 128; }  c = byte offset of the lane's next slot; the
// slot address lives in a temporary so that the store never holds up the counter.
__device__ __forceinline__ void wss_append(uint32_t& c, uint32_t lb, float x, float lox) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b32 t;\n\t"
      "setp.gt.f32 p, %2, %3;\n\t"
      "add.u32 t, %0, %1;\n\t"
      "@p st.shared.f32 [t], %2;\n\t"
      "@p add.u32 %0, %0, 128;\n\t}"
      : "+r"(c)
      : "r"(lb), "f"(x), "f"(lox));
}

struct WssAcc {
  float mn0, mn1, mx0, mx1;
  double sum0, sum1;
};

// four valid elements of one lane: min / max / f64 sum into the accumulators of parity `odd`,
// append(x > lox)
// COUNT: instead of the append, c += (x >= lox) and lb += (x > lox) -- the pass of a series whose
// p95 sat on a block of equal samples last time (`lox` is that value): two counts settle whether
// it still does, with no list and no select.
template <int CLAMP, bool COUNT = false>
__device__ __forceinline__ void wss_row(WssAcc& a, uint32_t& c, uint32_t& lb, float lox, float4 t,
                                        bool odd) {
  if (odd) {
    a.mn1 = fminf(fminf(a.mn1, t.x), fminf(t.y, fminf(t.z, t.w)));
    a.mx1 = fmaxf(fmaxf(a.mx1, t.x), fmaxf(t.y, fmaxf(t.z, t.w)));
    a.sum1 += ((double)t.x + (double)t.y) + ((double)t.z + (double)t.w);
  } else {
    a.mn0 = fminf(fminf(a.mn0, t.x), fminf(t.y, fminf(t.z, t.w)));
    a.mx0 = fmaxf(fmaxf(a.mx0, t.x), fmaxf(t.y, fmaxf(t.z, t.w)));
    a.sum0 += ((double)t.x + (double)t.y) + ((double)t.z + (double)t.w);
  }
  if (COUNT) {
    int ge = (int)c, gt = (int)lb;
    inc_if_ge(ge, t.x, lox); inc_if_ge(ge, t.y, lox); inc_if_ge(ge, t.z, lox); inc_if_ge(ge, t.w, lox);
    inc_if_gt(gt, t.x, lox); inc_if_gt(gt, t.y, lox); inc_if_gt(gt, t.z, lox); inc_if_gt(gt, t.w, lox);
    c = (uint32_t)ge;
    lb = (uint32_t)gt;
    return;
  }
  c = min(c, (uint32_t)(CLAMP * 128));
  wss_append(c, lb, t.x, lox);
  wss_append(c, lb, t.y, lox);
  wss_append(c, lb, t.z, lox);
  wss_append(c, lb, t.w, lox);
}

// ---- out-of-line helpers of the rare paths: they re-read the row (L2) in rolled loops

// visits every valid sample of the row: f(x). Four 128-bit loads per lane are in flight per
// trip (the row is in L2: a trip costs one round trip, not four).
template <typename F>
__device__ __forceinline__ void wss_for_each(const float* __restrict__ row, int lane, int n, F&& f) {
  const int nfull = n >> 2;
#pragma unroll 1
  for (int g0 = lane; g0 < nfull; g0 += 128) {
    float4 t[4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (g0 + 32 * u < nfull) t[u] = *reinterpret_cast<const float4*>(row + 4 * (g0 + 32 * u));
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (g0 + 32 * u < nfull) { f(t[u].x); f(t[u].y); f(t[u].z); f(t[u].w); }
  }
  if (lane < (n & 3)) f(row[4 * nfull + lane]);
}

// counts of valid samples >= p1 and >= p2 (warp totals)
__device__ __noinline__ int2 wss_count2(const float* __restrict__ row, int lane, int n, float p1, float p2) {
  int c1 = 0, c2 = 0;
  wss_for_each(row, lane, n, [&](float x) {
    inc_if_ge(c1, x, p1);
    inc_if_ge(c2, x, p2);
  });
  return make_int2(warp_sum_i32(c1), warp_sum_i32(c2));
}

// append(x > lox) over the whole window; returns the lane's slot offset
template <int CLAMP>
__device__ __noinline__ uint32_t wss_append_pass(const float* __restrict__ row, int lane, int n,
                                                 float lox, uint32_t lb) {
  uint32_t c = 0;
  wss_for_each(row, lane, n, [&](float x) {
    c = min(c, (uint32_t)(CLAMP * 128));
    wss_append(c, lb, x, lox);
  });
  return c;
}

// Fallback when the sampled pivots listed fewer than k elements, or a lane's list overflowed:
// narrow a bracket [a, b) with count_ge(a) = ca >= k > cb = count_ge(b) by counting passes over
// the row (two pivots per pass, interpolated on the counts; thirds of the key space when that
// stalls) until ca - cb elements fit the lists comfortably; the caller then lists the bracket
// and selects from it. Returns a and ca; if every element of the bracket equals a (adjacent
// keys), ca = -1 and a is the answer itself.
struct WssBracket {
  float a;
  int ca;  // -1: a is the answer
};
__device__ __noinline__ WssBracket wss_narrow(const float* __restrict__ row, int lane, int n, int k, float mn,
                                              float mx, int fit) {
  float a = mn, b = POS_INF;
  int ca = n, cb = 0;
  int it = 0;
  while (ca - cb > fit) {
    const uint32_t ka = f32_key(a), kb = (cb == 0) ? f32_key(mx) + 1u : f32_key(b);
    if (kb - ka <= 1u) return WssBracket{a, -1};
    float p1, p2;
    const float hi = (cb == 0) ? mx : b;
    const int span = ca - cb;
    const int dl = fit / 4;
    const float inv = 1.0f / (float)span;
    p1 = a + (hi - a) * ((float)max(ca - k - dl, 0) + 0.5f) * inv;
    p2 = a + (hi - a) * ((float)min(ca - k + dl, span) - 0.5f) * inv;
    uint32_t k1 = f32_key(p1), k2 = f32_key(p2);
    if (!(p1 == p1) || !(p2 == p2) || (it & 1) || k1 <= ka || k2 >= kb || k1 > k2) {
      // every other pass (and whenever the interpolation leaves the bracket): thirds of the key
      // space, which bounds the number of passes whatever the distribution
      k1 = ka + max((kb - ka) / 3u, 1u);
      k2 = max(kb - max((kb - ka) / 3u, 1u), k1);
    }
    p1 = key_f32(k1); p2 = key_f32(k2);
    const int2 cc = wss_count2(row, lane, n, p1, p2);
    if (cc.y >= k) { a = p2; ca = cc.y; }
    else if (cc.x >= k) { a = p1; ca = cc.x; b = p2; cb = cc.y; }
    else { b = p1; cb = cc.x; }
    ++it;
  }
  return WssBracket{a, ca};
}

// The largest sample strictly below `cur` and how many samples equal it (warp totals; -inf and 0 if
// there is none): one pass over the row (L2). A carried floor only ever moves up, so after enough
// ticks that evict a sample above it and bring one below it, fewer than k samples are left at or
// above it; the answer is then the first or second distinct value under the floor, and stepping
// down from the floor finds it in one or two passes instead of the counting fallback's five.
struct WssBelow {
  float v;
  int count;
};
__device__ __noinline__ WssBelow wss_step_below(const float* __restrict__ row, int lane, int n, float cur) {
  float m = NEG_INF;
  int c = 0;
  wss_for_each(row, lane, n, [&](float x) {
    if (x < cur) {
      if (x > m) { m = x; c = 1; }
      else if (x == m) ++c;
    }
  });
  const float M = warp_max_f32(m);
  return WssBelow{M, warp_sum_i32(m == M ? c : 0)};
}

// select_from_lists for the fallback path, out of line (one inlined copy serves the hot path)
__device__ __noinline__ float wss_select_cold(float* lst, float* fin, int lane, int cnt, int total, int maxcnt,
                                              int k, float a, float mx, int prefilled, float2* next_hint) {
  return select_from_lists(lst, fin, lane, cnt, total, maxcnt, k, a, mx, false, 0.f, 0.f, prefilled, next_hint);
}

// RS, D > 0: the rows do not pass through register buffers but through a per-warp ring of D stages
// of RS rows in shared memory, filled by cp.async: D - 1 stages are in flight whatever the
// register budget, and the 32 buffer registers are free, which is what lets 28-32 warps share an
// SM (the lists must then be short, CAP 24-32: made for the hinted pass, which lists ~200 samples).
template <int CAP, int MINB, int NBUF, int RS = 0, int D = 0>
__global__ void __launch_bounds__(kWssThreads, MINB)
window_stats_stream_kernel(const float* __restrict__ ring, const uint32_t* __restrict__ seen,
                           int32_t window, int32_t pitch, int64_t s0, int64_t n_series,
                           float* __restrict__ out_min, float* __restrict__ out_max,
                           double* __restrict__ out_mean, float* __restrict__ out_p95,
                           float2* __restrict__ state) {
  using Cfg = WssCfg<CAP, RS, D>;
  constexpr bool kRing = D > 0;
  extern __shared__ __align__(16) unsigned char s_raw[];

  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int64_t o = (int64_t)blockIdx.x * kWssWarps + wib;  // output slot
  if (o >= n_series) return;                                  // no block-wide barrier below
  const int64_t s = s0 + o;
  const float* __restrict__ row = ring + s * (int64_t)pitch;
  const uint32_t sn = seen[s];  // consumed only after the first row loads are in flight
#ifndef KLM_WSS_PF
#define KLM_WSS_PF 0
#endif
  // KLM_WSS_PF == 2: one bulk-prefetch instruction asks L2 for this warp's whole row right away, so
  // that the register buffers, which hold only a few rows, are refilled from L2 and not from DRAM
  if (KLM_WSS_PF == 2 && lane == 0) {
    const uint32_t bytes = ((uint32_t)window * 4u + 15u) & ~15u;
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" :: "l"(row), "r"(bytes) : "memory");
  }

  unsigned char* list_g = s_raw + wib * Cfg::kListBytes;
  float* lst = reinterpret_cast<float*>(list_g) + lane;  // entry e: lst[e * 32]
  float4* const lst4 = reinterpret_cast<float4*>(list_g) + lane;
  const uint32_t lb = (uint32_t)__cvta_generic_to_shared(list_g) + ((uint32_t)lane << 2);
  float* fin = reinterpret_cast<float*>(s_raw + kWssWarps * Cfg::kListBytes + wib * kWssAux);
  // this lane's 16 bytes of stage 0, row 0 of the warp's landing ring
  const uint32_t ring_s = (uint32_t)__cvta_generic_to_shared(s_raw + kWssWarps * (Cfg::kListBytes + kWssAux) +
                                                             wib * Cfg::kRingBytes) + ((uint32_t)lane << 4);

  // ---- 1. first loads: the sample rows and the first NBUF - 1 chunks of four rows, at
  // addresses that depend only on the kernel arguments (a full window is the steady state; a
  // ring that is still filling takes the generic sample below). All lie inside the row pitch.
  const float* p = row + 4 * lane;
  const int wjr = (window >> 2) >> 5;  // full rows of a full window
  const bool want_p95 = out_p95 != nullptr;
  // (the landing ring first: its copies must not wait for the bracket to arrive)
  auto fetch_stage = [&](int slot, int stage) {  // rows stage * RS .. + RS - 1 -> ring slot
#pragma unroll
    for (int r = 0; r < RS; ++r) cp_async16(ring_s + (uint32_t)((slot * RS + r) * 512), p + 128 * (stage * RS + r));
  };
  if constexpr (kRing) {
#pragma unroll
    for (int b = 0; b < D - 1; ++b) {
      if ((b + 1) * RS <= wjr) fetch_stage(b, b);
      cp_async_commit();  // (one group per stage, empty or not: the waits below count groups)
    }
    // the row that only some lanes hold is read last, straight from global memory: ask L2 for it now
    if (lane + 32 * wjr < (window >> 2)) asm volatile("prefetch.global.L2 [%0];" ::"l"(p + 128 * wjr));
  }
  // Streaming windows: `state` holds, per series, the bracket (floor, ceil) around the p95 that
  // the previous call left (window_stats.cuh, hint_from_sorted). A usable bracket replaces the
  // sampling: the pass lists everything >= floor, about k + 12 elements, and the finish is one
  // counting pass and one sort. The bracket is only a pivot: stale, wrong or garbage values cost
  // time, never exactness.
  float2 hint = make_float2(0.f, 0.f);
  bool hinted = false, tie_hinted = false;
  if (state != nullptr && want_p95) {
    hint = state[o];
    hinted = wjr >= kWssSampledMinRows && hint.x < hint.y && hint.x > NEG_INF;  // (NaN fails every test)
    // floor == ceil: last time the p95 sat on a block of equal samples (quantised channels, flags,
    // constants); the pass then only counts around that value
    tie_hinted = hint.x == hint.y && hint.x > NEG_INF && hint.x < POS_INF;
  }
  float sm[4] = {NEG_INF, NEG_INF, NEG_INF, NEG_INF};
  if (!hinted && !tie_hinted && want_p95 && wjr >= kWssSampledMinRows) {
#pragma unroll
    for (int q = 0; q < 4; ++q) sm[q] = __ldg(p + 128 * (((2 * q + 1) * wjr) >> 3) + q);
  }
  float4 buf[kRing ? 1 : NBUF][4];
  auto load_chunk = [&](float4 (&dst)[4], int c) {
#pragma unroll
    for (int i = 0; i < 4; ++i) dst[i] = ld_stream_f4(p + 128 * (4 * c + i));
  };
  if constexpr (!kRing) {
#pragma unroll
    for (int b = 0; b < NBUF - 1; ++b)
      if (4 * b + 4 <= wjr) load_chunk(buf[b], b);
  }
  // While the loads are in flight: fill the first Cfg::kPrefill entries of this lane's list with
  // -inf (128-bit stores across the warp), so the counting passes need no padding later.
  auto prefill = [&]() {
#pragma unroll
    for (int e = 0; e < Cfg::kPrefill / 4; ++e)
      lst4[e * 32] = make_float4(NEG_INF, NEG_INF, NEG_INF, NEG_INF);
    __syncwarp();  // lanes append into slots another lane has just filled
  };
  prefill();
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
  // (tuning builds, KLM_WSS_PF == 1: ask L2 for the row that a warp will want about one residency
  // later; measured slower than no prefetch)
  if (KLM_WSS_PF == 1 && lane == 0) {
    const int64_t later = o + (int64_t)kNumSMs * MINB * kWssWarps;
    if (later < n_series) {
      const uint32_t bytes = ((uint32_t)n * 4u + 15u) & ~15u;
      asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;"
                   :: "l"(ring + (s0 + later) * (int64_t)pitch), "r"(bytes) : "memory");
    }
  }
  const int nfull = n >> 2;   // complete float4 groups
  const int jr = nfull >> 5;  // rows valid in every lane
  const int k = n - p95_rank(n);
  hinted = hinted && n == window;  // (a ring that is still filling changes too fast for a bracket)
  tie_hinted = tie_hinted && n >= 1;
  const bool sampled = want_p95 && jr >= kWssSampledMinRows && !hinted && !tie_hinted;
  if (sampled && (n != window || hint.x < hint.y)) {  // partly filled ring, or a bracket dropped for it: sample the rows that hold data
#pragma unroll
    for (int q = 0; q < 4; ++q) sm[q] = __ldg(p + 128 * (((2 * q + 1) * jr) >> 3) + q);
  }
  float2 next_hint = make_float2(__int_as_float(0x7fc00000), __int_as_float(0x7fc00000));  // NaN: no bracket

  bool narrow = false;             // settle by counting passes over the row (wss_narrow)
  float lox = POS_INF;             // list predicate: x > lox (+inf: list nothing)
  float hint_lo = 0.f, hint_hi = 0.f;  // first pair of pivots (sample order statistics)
  float lox_retry = POS_INF;       // lower pivot for one retry of the append (+inf: none)
  bool tie_mode = false, retried = false;
  bool tie_hit = false;            // the heavy-tie test already found the answer
  float tie_val = 0.f;

  // ---- 2. sample pivot: four groups of 32 (one sample row each), top three of each
  if (sampled) {
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
      // heavy tie at V: place k against the block of copies of V (one pass over the row)
      const int2 cc = wss_count2(row, lane, n, V, next_up(V));
      if (cc.y >= k) {
        lox = V;                       // the answer is above V: list the elements > V
      } else if (cc.x >= k) {
        tie_hit = true; tie_val = V;   // the answer is V itself
      } else if (t3 < t2) {
        lox = lox_retry;               // below V: list from the next sample statistic down
        lox_retry = POS_INF;
      } else {
        narrow = true;                 // the answer is below V
      }
    } else {
      lox = next_down(V);  // x > lox  <=>  x >= V
    }
  } else if (hinted) {
    lox = next_down(hint.x);  // x > lox  <=>  x >= floor
  } else if (tie_hinted) {
    lox = POS_INF;            // (the tie pass lists nothing)
  } else if (want_p95) {
    lox = NEG_INF;  // short window: list every sample
  }

  // ---- 3. the pass: min / max / f64 sum + append(x > lox). Chunks of four rows rotate through
  // NBUF register buffers: the loads of NBUF - 1 chunks are in flight while one is consumed.
  WssAcc acc{POS_INF, POS_INF, NEG_INF, NEG_INF, 0.0, 0.0};
  uint32_t c128 = 0;
  uint32_t tie_gt = 0;  // tie pass: samples > the tied value (c128 then counts those >= it)
  const int nch = kRing ? 0 : jr >> 2;
  const int nst = kRing ? jr / (RS > 0 ? RS : 1) : 0;  // full stages
  const int jdone = kRing ? nst * RS : 4 * nch;        // rows the main loop covers
  auto pass = [&](auto row_fn, auto tail_fn) {
    if constexpr (kRing) {
      // Stage g lives in slot g mod D. A trip refills the slot consumed one trip earlier with stage
      // g + D - 1, waits for stage g (all but the D - 1 youngest copy groups) and consumes it; the
      // last D - 1 trips only drain (empty groups keep the count). Slot addresses and the fetch
      // pointer are carried in registers: ~12 instructions of ring handling per stage.
      constexpr uint32_t kStage = RS * 512;
      const uint32_t ring_end = ring_s + D * kStage;
      uint32_t cs = ring_s, fs = ring_s + (D - 1) * kStage;
      const float* pf = p + 128 * RS * (D - 1);
      auto consume = [&]() {
        float4 t[RS > 0 ? RS : 1];
#pragma unroll
        for (int r = 0; r < RS; ++r) t[r] = lds_f4(cs + (uint32_t)(r * 512));
#pragma unroll
        for (int r = 0; r < RS; ++r) row_fn(t[r], r & 1);
        fs = cs;
        cs += kStage;
        if (cs == ring_end) cs = ring_s;
      };
      int g = 0;
#pragma unroll 1
      for (; g + D - 1 < nst; ++g) {
#pragma unroll
        for (int r = 0; r < RS; ++r) cp_async16(fs + (uint32_t)(r * 512), pf + 128 * r);
        pf += 128 * RS;
        cp_async_commit();
        cp_async_wait<D - 1>();
        consume();
      }
#pragma unroll 1
      for (; g < nst; ++g) {
        cp_async_commit();
        cp_async_wait<D - 1>();
        consume();
      }
    } else if (nch > 0) {
#pragma unroll
      for (int b = 0; b < NBUF - 1; ++b)  // chunks the early loads did not cover (cannot happen for
        if (4 * b + 4 > wjr && b < nch) load_chunk(buf[b], b);  // n <= window; keeps buf defined)
#pragma unroll 1
      for (int c0 = 0; c0 < nch; c0 += NBUF) {
#pragma unroll
        for (int b = 0; b < NBUF; ++b) {
          const int c = c0 + b;
          if (c < nch) {
            if (c + NBUF - 1 < nch) load_chunk(buf[(b + NBUF - 1) % NBUF], c + NBUF - 1);
#pragma unroll
            for (int i = 0; i < 4; ++i) row_fn(buf[b][i], i & 1);
          }
        }
      }
    }
    // the rows left over (fewer than a chunk / a stage), then the row that only some lanes hold
#pragma unroll 1
    for (int j = jdone; j < jr; ++j) row_fn(ld_stream_f4(p + 128 * j), false);
    if (lane + 32 * jr < nfull) row_fn(ld_stream_f4(p + 128 * jr), false);
    if (lane < (n & 3)) {
      const float x = row[4 * nfull + lane];
      acc.mn0 = fminf(acc.mn0, x);
      acc.mx0 = fmaxf(acc.mx0, x);
      acc.sum0 += (double)x;
      tail_fn(x);
    }
  };
  if (tie_hinted) {
    const float V = hint.x;
    pass([&](float4 t, bool odd) { wss_row<Cfg::kClamp, true>(acc, c128, tie_gt, V, t, odd); },
         [&](float x) {
           c128 += x >= V;
           tie_gt += x > V;
         });
  } else {
    uint32_t lbv = lb;
    pass([&](float4 t, bool odd) { wss_row<Cfg::kClamp, false>(acc, c128, lbv, lox, t, odd); },
         [&](float x) {
           c128 = min(c128, (uint32_t)(Cfg::kClamp * 128));
           wss_append(c128, lb, x, lox);
         });
  }
  __syncwarp();
  asm volatile("" ::: "memory");  // the appends above are plain shared-memory stores
  float mn = fminf(acc.mn0, acc.mn1), mx = fmaxf(acc.mx0, acc.mx1);
  double sum = acc.sum0 + acc.sum1;
  mn = warp_min_f32(mn);
  mx = warp_max_f32(mx);
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, off);
  if (lane == 0) {
    if (out_min) out_min[o] = mn + 0.f;  // -0.0 -> +0.0: one canonical zero in the outputs
    if (out_max) out_max[o] = mx + 0.f;
    if (out_mean) out_mean[o] = sum / (double)n;
  }
  if (!want_p95) return;

  // ---- 4. p95 = k-th largest
  float result = 0.f;
  do {  // single-trip block: `break` = result (or the request to narrow) is final
    if (tie_hinted) {
      const int c_ge = warp_sum_i32((int)c128), c_gt = warp_sum_i32((int)tie_gt);
      if (c_gt < k && k <= c_ge) { KLM_WSS_COUNT(0); result = hint.x; next_hint = hint; break; }  // still on the same block
      narrow = true;  // the p95 moved off that value: settle it from the row
      break;
    }
    if (narrow) break;
    if (tie_hit) { result = tie_val; next_hint = make_float2(tie_val, tie_val); break; }
    if (mn == mx) { result = mx; next_hint = make_float2(mx, mx); break; }  // constant series
    int cnt = (int)(c128 >> 7);
    int total = warp_sum_i32(cnt);
    int maxcnt = warp_max_i32(cnt);
    if (hinted) {
      if (total < k && maxcnt < Cfg::kClamp) {
        // the floor has drifted above the answer: step down from it, one distinct value per pass
        float cur = hint.x;
        int have = total;
        bool found = false;
#pragma unroll 1
        for (int step = 0; step < 3 && !found; ++step) {
          const WssBelow bl = wss_step_below(row, lane, n, cur);
          if (bl.count == 0) break;  // nothing below: only -inf / NaN samples are left
          if (have + bl.count >= k) {
            result = bl.v;
            found = true;
            // a block of equal samples: carry it as a tie value; else no bracket (the next call samples afresh)
            if (bl.count > kMaxCand / 2) next_hint = make_float2(bl.v, bl.v);
          }
          have += bl.count;
          cur = bl.v;
        }
        KLM_WSS_COUNT(7);
        if (found) break;
        narrow = true;
        break;
      }
      // far too many listed (the floor sits on a plateau that has grown): settle by counting
      if (total < k || maxcnt >= Cfg::kClamp) { narrow = true; break; }
      if (select_hinted(lst, fin, lane, cnt, total, maxcnt, k, hint.x, hint.y, &result, &next_hint, Cfg::kPrefill)) { KLM_WSS_COUNT(1); break; }
      KLM_WSS_COUNT(2);
      // the answer left the bracket, or the bracket is wide: the lists still hold everything >= floor
      result = select_from_lists(lst, fin, lane, cnt, total, maxcnt, k, hint.x, mx, false, 0.f, 0.f, Cfg::kPrefill,
                                 &next_hint);
      break;
    }
    if (total < k && !sampled) { result = NEG_INF; break; }  // lox = -inf: only -inf samples are unlisted
    if (total < k && !tie_mode && lox_retry != POS_INF) {
      // the pivot was too high: redo the append (only) from the next sample statistic down
      __syncwarp();
      prefill();
      lox = lox_retry;
      c128 = wss_append_pass<Cfg::kClamp>(row, lane, n, lox, lb);
      __syncwarp();
      asm volatile("" ::: "memory");
      cnt = (int)(c128 >> 7);
      total = warp_sum_i32(cnt);
      maxcnt = warp_max_i32(cnt);
      retried = true;
      KLM_WSS_COUNT(6);
    }
    if (total < k || maxcnt >= Cfg::kClamp) {  // unlucky pivots, or a lane's list overflowed
      narrow = true;
      break;
    }
    KLM_WSS_COUNT(sampled ? 3 : 5);
    result = select_from_lists(lst, fin, lane, cnt, total, maxcnt, k, sampled ? next_up(lox) : mn, mx,
                               retried, hint_lo, hint_hi, Cfg::kPrefill, &next_hint);
  } while (false);
  if (narrow) {
    KLM_WSS_COUNT(4);
    if (mn == mx) {
      result = mx;
      next_hint = make_float2(mx, mx);
    } else {
      // first try: a bracket of at most kFit elements on top of the < k above it, listed and
      // selected from; if a single lane's share still overflows, settle by counting alone
#pragma unroll 1
      for (int attempt = 0; attempt < 2; ++attempt) {
        // (the lists must then hold the < k samples above the bracket and the bracket itself)
        constexpr int kFit = CAP >= 64 ? 16 * 32 : 0;
        const int fit = kFit > 0 ? kFit : max(Cfg::kClamp * 16 - k, 32);
        const WssBracket br = wss_narrow(row, lane, n, k, mn, mx, attempt == 0 ? fit : 0);
        const float a = br.a;
        if (br.ca < 0) { result = a; next_hint = make_float2(a, a); break; }  // a block of equal samples
        __syncwarp();
        prefill();
        lox = a == NEG_INF ? NEG_INF : next_down(a);
        c128 = wss_append_pass<Cfg::kClamp>(row, lane, n, lox, lb);
        __syncwarp();
        asm volatile("" ::: "memory");
        const int cnt = (int)(c128 >> 7);
        const int total = warp_sum_i32(cnt);
        const int maxcnt = warp_max_i32(cnt);
        if (maxcnt >= Cfg::kClamp) continue;
        if (total < k) { result = NEG_INF; break; }  // a = -inf: only -inf samples are unlisted
        result = wss_select_cold(lst, fin, lane, cnt, total, maxcnt, k, a, mx, Cfg::kPrefill, &next_hint);
        break;
      }
    }
  }
  if (lane == 0) {
    out_p95[o] = result + 0.f;
    if (state != nullptr) state[o] = next_hint;  // NaN pair when this path produced no bracket
  }
}

template <int CAP, int MINB, int NBUF, int RS = 0, int D = 0>
static int launch_stream(const float* ring, const uint32_t* seen, int32_t window, int32_t pitch,
                         int64_t s0, int64_t n_series, float* out_min, float* out_max,
                         double* out_mean, float* out_p95, float2* state, cudaStream_t st) {
  using Cfg = WssCfg<CAP, RS, D>;
  static_assert(MINB * (Cfg::kSmem + 1024) <= 228 * 1024, "the CTAs per SM must fit");
  const int64_t blocks = ceil_div(n_series, kWssWarps);
  if (blocks > 0x7fffffff) {
    set_error("klm_dev_window_stats: too many series for one launch");
    return KLM_ELIMIT;
  }
  static PerDeviceOnce attr_set;
  int dev = 0;
  KLM_CUDA(cudaGetDevice(&dev));
  auto kern = window_stats_stream_kernel<CAP, MINB, NBUF, RS, D>;
  if (!attr_set.done(dev)) {
    // opt in above 48 KB and ask for the largest shared-memory carve-out
    KLM_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::kSmem));
    KLM_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout,
                                  (int)cudaSharedmemCarveoutMaxShared));
    attr_set.set(dev);
  }
  kern<<<(unsigned)blocks, kWssThreads, Cfg::kSmem, st>>>(ring, seen, window, pitch, s0, n_series, out_min, out_max,
                                                          out_mean, out_p95, state);
  KLM_LAUNCH_CHECK("window_stats_stream_kernel");
  return KLM_OK;
}

#ifndef KLM_WSS_CAP
#define KLM_WSS_CAP 64
#endif
#ifndef KLM_WSS_MINB
#define KLM_WSS_MINB 6
#endif
#ifndef KLM_WSS_NBUF
#define KLM_WSS_NBUF 2
#endif

int launch_window_stats_stream(const float* ring, const uint32_t* seen, int32_t window, int32_t pitch,
                               int64_t s0, int64_t s1, float* out_min, float* out_max,
                               double* out_mean, float* out_p95, float* state, cudaStream_t st) {
  float2* st2 = reinterpret_cast<float2*>(state);
#define KLM_WSS_GO(...) \
  return launch_stream<__VA_ARGS__>(ring, seen, window, pitch, s0, s1 - s0, out_min, out_max, out_mean, out_p95, st2, st)
#ifdef KLM_WSS_EXPERIMENTS
  // tuning builds: KLM_WSS_VARIANT picks a landing-ring shape at run time
  static const int variant = [] { const char* e = getenv("KLM_WSS_VARIANT"); return e ? atoi(e) : 0; }();
  if (state != nullptr && window <= 4096) {
    switch (variant) {
      case 5: KLM_WSS_GO(32, 7, 2, 1, 7);
      case 6: KLM_WSS_GO(32, 6, 2, 2, 4);
      case 7: KLM_WSS_GO(32, 6, 2, 1, 9);
      case 9: KLM_WSS_GO(32, 6, 2, 2, 5);
      case 10: KLM_WSS_GO(32, 6, 2, 3, 3);
      case 11: KLM_WSS_GO(32, 6, 2, 4, 2);
      case 12: KLM_WSS_GO(32, 5, 2, 2, 6);
      case 13: KLM_WSS_GO(32, 7, 2, 2, 3);
      case 14: KLM_WSS_GO(40, 6, 2, 2, 4);
      default: break;
    }
  }
  if (window <= 8192) {
    switch (variant) {
      case 20: KLM_WSS_GO(64, 4, 2, 2, 4);
      case 21: KLM_WSS_GO(48, 5, 2, 2, 4);
      case 22: KLM_WSS_GO(64, 4, 2, 2, 5);
      default: break;
    }
  }
#endif
  // carried brackets: the pass lists ~k + 12 samples, so short lists leave shared memory for a
  // landing ring of five 1 KB stages per warp (77 us at C2 against 96 us through register buffers)
  if (state != nullptr && window <= 8192) KLM_WSS_GO(32, 6, 2, 2, 5);
  KLM_WSS_GO(KLM_WSS_CAP, KLM_WSS_MINB, KLM_WSS_NBUF);
#undef KLM_WSS_GO
}

}  // namespace klm

#ifdef KLM_WSS_STATS
extern "C" int klm_dbg_wss_paths(unsigned long long* out8, int reset) {
  unsigned long long z[8] = {0};
  if (out8 && cudaMemcpyFromSymbol(out8, klm::g_wss_paths, sizeof(z)) != cudaSuccess) return -1;
  if (reset && cudaMemcpyToSymbol(klm::g_wss_paths, z, sizeof(z)) != cudaSuccess) return -1;
  return 0;
}
#endif
