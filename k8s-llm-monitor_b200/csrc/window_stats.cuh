// Helpers shared by the window-statistics kernels.
#pragma once

#include <stdint.h>

#include "common.cuh"

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

#define NEG_INF __int_as_float(0xff800000)
#define POS_INF __int_as_float(0x7f800000)

// Neighbours in VALUE order: -0.0 and +0.0 are adjacent keys but one value, so stepping off a zero
// must skip the other zero -- `x > next_down(0)` has to hold for every zero sample, and
// `x >= next_up(0)` for none (a p95 that sits on a plateau of zeros next to a few spikes is an
// everyday series: idle flags, a drained battery).
__device__ __forceinline__ float next_up(float x) {   // smallest float > x (x finite or -inf)
  return x == 0.f ? __uint_as_float(0x00000001u) : key_f32(f32_key(x) + 1u);
}
__device__ __forceinline__ float next_down(float x) { // largest float < x (x finite or +inf)
  return x == 0.f ? __uint_as_float(0x80000001u) : key_f32(f32_key(x) - 1u);
}
__device__ __forceinline__ int warp_max_i32(int v) { return __reduce_max_sync(0xffffffffu, v); }
__device__ __forceinline__ uint32_t warp_max_u32(uint32_t v) { return __reduce_max_sync(0xffffffffu, v); }

// c += (x >= p) as FSETP + predicated add (ptxas has no cheaper form on sm_100a: a
// `set` instruction lowers to FSETP + SEL, and `c += (x >= p)` to FSETP + add + move)
__device__ __forceinline__ void inc_if_ge(int& c, float x, float p) {
  asm("{\n\t.reg .pred q;\n\tsetp.ge.f32 q, %1, %2;\n\t@q add.s32 %0, %0, 1;\n\t}"
      : "+r"(c) : "f"(x), "f"(p));
}
__device__ __forceinline__ void inc_if_gt(int& c, float x, float p) {
  asm("{\n\t.reg .pred q;\n\tsetp.gt.f32 q, %1, %2;\n\t@q add.s32 %0, %0, 1;\n\t}"
      : "+r"(c) : "f"(x), "f"(p));
}

// -1 if a >= b else 0
__device__ __forceinline__ int set_ge(float a, float b) {
  int d;
  asm("set.ge.s32.f32 %0, %1, %2;" : "=r"(d) : "f"(a), "f"(b));
  return d;
}
__device__ __forceinline__ int set_gt(float a, float b) {
  int d;
  asm("set.gt.s32.f32 %0, %1, %2;" : "=r"(d) : "f"(a), "f"(b));
  return d;
}

// Counts of list entries >= p1 and >= p2 (p1 <= p2) in one pass over the lane lists
// ([entry][lane], every lane padded with -inf up to m8, a multiple of 8). The eight
// loads of a batch are issued together so their latencies overlap. Returns c1 | c2 << 16.
__device__ __forceinline__ uint32_t list_count2(const float* lst, int m8, float p1, float p2,
                                                int& lane_c1, int& lane_c2) {
  int a0 = 0, a1 = 0, b0 = 0, b1 = 0;
#pragma unroll 1
  for (int e = 0; e < m8; e += 8) {
    float x[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) x[u] = lst[(e + u) * 32];
#pragma unroll
    for (int u = 0; u < 8; u += 2) {
      inc_if_ge(a0, x[u], p1); inc_if_ge(a1, x[u + 1], p1);
      inc_if_ge(b0, x[u], p2); inc_if_ge(b1, x[u + 1], p2);
    }
  }
  lane_c1 = a0 + a1;
  lane_c2 = b0 + b1;
  return (uint32_t)warp_sum_i32(lane_c1 + (lane_c2 << 16));
}

// Largest list entry < b (or the largest entry if `open`), as a key (0 if none).
__device__ __forceinline__ uint32_t list_max_below(const float* lst, int m8, float b, bool open) {
  uint32_t m = 0;
#pragma unroll 1
  for (int e = 0; e < m8; e += 8) {
    float x[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) x[u] = lst[(e + u) * 32];
#pragma unroll
    for (int u = 0; u < 8; ++u)
      if (open || x[u] < b) m = max(m, f32_key(x[u]));
  }
  return warp_max_u32(m);
}

// Bitonic sort of 64 keys, descending, two per lane: element 2 * lane + r is k[r].
__device__ __forceinline__ void sort64_desc(uint32_t& k0, uint32_t& k1, int lane) {
#pragma unroll
  for (int size = 2; size <= 64; size <<= 1) {
    const bool desc = ((2 * lane) & size) == 0;  // size == 64: every lane descending
#pragma unroll
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      if (stride == 1) {
        const uint32_t hi = max(k0, k1), lo = min(k0, k1);
        k0 = desc ? hi : lo;
        k1 = desc ? lo : hi;
      } else {
        const int ls = stride >> 1;
        const bool keepmax = (((lane & ls) == 0) == desc);
        const uint32_t o0 = __shfl_xor_sync(0xffffffffu, k0, ls);
        const uint32_t o1 = __shfl_xor_sync(0xffffffffu, k1, ls);
        k0 = keepmax ? max(k0, o0) : min(k0, o0);
        k1 = keepmax ? max(k1, o1) : min(k1, o1);
      }
    }
  }
}

// Bitonic sort of 32 keys, descending, one per lane: element `lane` is k.
__device__ __forceinline__ void sort32_desc(uint32_t& k, int lane) {
#pragma unroll
  for (int size = 2; size <= 32; size <<= 1) {
    const bool desc = (lane & size) == 0;  // size == 32: every lane descending
#pragma unroll
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      const bool keepmax = ((lane & stride) == 0) == desc;
      const uint32_t o = __shfl_xor_sync(0xffffffffu, k, stride);
      k = keepmax ? max(k, o) : min(k, o);
    }
  }
}

// warp-wide min / max of f32 in one instruction each (CREDUX on sm_100a; NaN inputs are passed over,
// like fminf / fmaxf)
__device__ __forceinline__ float warp_min_f32(float v) {
  float r;
  asm volatile("redux.sync.min.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
  return r;
}
__device__ __forceinline__ float warp_max_f32(float v) {
  float r;
  asm volatile("redux.sync.max.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
  return r;
}

constexpr int kMaxCand = 64;  // candidates sorted exactly at the end of select_from_lists
constexpr int kHintMargin = 12;    // the hint brackets ranks k - 12 .. k + 18 of the window:
constexpr int kHintMarginLo = 18;  // wider below, where it cannot be re-centred; <= 31 candidates
// Elements below the floor are never listed, so a carried floor can only move up: ticks that evict a
// sample above it and bring one below it walk it towards the answer one rank at a time, and once
// fewer than k samples are left at or above it the call needs extra passes over the row -- a serial
// chain of L2 round trips that lengthens the tail of the whole launch (a handful of such series per
// launch cost C2 about 3 us of 77). So a bracket is dropped when fewer than k + kHintRefresh samples
// remain above its floor, and the next call takes a fresh one from the sampled pass: about a third
// more work for that series once in ~2000 ticks, and no straggler.
constexpr int kHintRefresh = 3;

// The bracket handed to the next call on this series (one tick later at most one sample has
// entered and one has left, so ranks move by at most one): from the sorted candidates (descending,
// element e in lane e / 2, register e % 2; n_cand of them; the answer at 1-based position kk),
// floor = the candidate kHintMargin places below the answer, ceil = the one kHintMargin places
// above (or `above`, a value with fewer than k samples at or over it, when the candidates do not
// reach that far). A hint is only ever a pivot: any value gives exact results.
__device__ __forceinline__ float2 hint_from_sorted(uint32_t k0, uint32_t k1, int kk, int n_cand, float above) {
  const int pl = min(kk - 1 + kHintMarginLo, n_cand - 1), ph = kk - 1 - kHintMargin;
  const uint32_t lo = __shfl_sync(0xffffffffu, (pl & 1) ? k1 : k0, pl >> 1);
  const uint32_t hi = __shfl_sync(0xffffffffu, (max(ph, 0) & 1) ? k1 : k0, max(ph, 0) >> 1);
  return make_float2(key_f32(lo), ph >= 0 ? key_f32(hi) : above);
}
// the same for 32 sorted candidates held one per lane
__device__ __forceinline__ float2 hint_from_sorted32(uint32_t key, int kk, int n_cand, float above) {
  const int pl = min(kk - 1 + kHintMarginLo, n_cand - 1), ph = kk - 1 - kHintMargin;
  const uint32_t lo = __shfl_sync(0xffffffffu, key, pl);
  const uint32_t hi = __shfl_sync(0xffffffffu, key, max(ph, 0));
  return make_float2(key_f32(lo), ph >= 0 ? key_f32(hi) : above);
}
constexpr int kListPrefill = 24;  // list entries per lane the kernels pre-fill with -inf

// The k-th largest of the elements held in the lane lists (lst = this lane's list, entry e at
// lst[e * 32]; entries below kListPrefill beyond the lane's count already hold -inf). `a` is a value
// with count_ge(a) == total >= k (every listed element is >= a); mx is the window maximum.
// `fin` is 64 floats of scratch owned by the warp. Counting passes test two pivots at a
// time (first the sample hints, then interpolation on the counts) until <= 64 candidates
// remain; each lane knows its share, an exclusive scan hands out slots, a bitonic sort
// finishes. Brackets held wide by ties are peeled one distinct value at a time; after 12
// passes plain key-space bisection bounds the work. Exact: the result is a listed element
// whose rank has been counted.
__device__ __forceinline__ float select_from_lists(float* lst, float* fin, int lane, int cnt,
                                                   int total, int maxcnt, int k, float a, float mx,
                                                   bool have_hints, float hint_lo, float hint_hi,
                                                   int prefilled, float2* next_hint = nullptr) {
  const int m8 = (maxcnt + 7) & ~7;
  // entries below `prefilled` beyond the lane's count already hold -inf (pre-filled by the
  // caller); pad the rest up to m8 here. prefilled == 0: the usual lengths as straight-line
  // predicated stores, the rare longer lists in a loop.
  if (prefilled == 0) {
#pragma unroll
    for (int e = 0; e < kListPrefill; ++e)
      if (e >= cnt && e < m8) lst[e * 32] = NEG_INF;
    prefilled = kListPrefill;
  }
  for (int e = max(cnt, prefilled); e < m8; ++e) lst[e * 32] = NEG_INF;
  fin[lane] = NEG_INF;
  fin[lane + 32] = NEG_INF;
  __syncwarp();

  // narrow [a, b): count_ge(a) = ca >= k > cb = count_ge(b); b = +inf with cb = 0 at first
  float b = POS_INF;
  int ca = total, cb = 0;
  int lane_ca = cnt, lane_cb = 0;  // this lane's share of ca and cb
  int it = 0;
#pragma unroll 1
  while (ca - cb > kMaxCand) {
    const uint32_t ka = f32_key(a), kb = (cb == 0) ? f32_key(mx) + 1u : f32_key(b);
    if (kb - ka <= 1u) {  // every candidate equals a: the answer sits on a block of equal samples
      if (next_hint) *next_hint = make_float2(a, a);
      return a;
    }
    if (it >= 4 && it < 12) {
      // ties keep the bracket wide: peel its largest distinct value
      const float m = key_f32(list_max_below(lst, m8, b, cb == 0));
      int l1, l2;
      const int c = (int)(list_count2(lst, m8, m, m, l1, l2) & 0xffffu);
      if (c >= k) {  // everything above m is >= b
        if (next_hint) *next_hint = make_float2(m, m);
        return m;
      }
      b = m; cb = c; lane_cb = l1;
    } else {
      float p1, p2;
      if (it == 0 && have_hints && hint_lo > a && hint_hi > hint_lo) {
        p1 = hint_lo; p2 = hint_hi;       // sample order statistics around the 5 % point
      } else {
        // interpolate on the counts towards ranks k + dl and k - dl (value-linear inside)
        const float hi = (cb == 0) ? mx : b;
        const int span = ca - cb;
        const int dl = min(max(span / 7, 12), 24);
        const float inv = 1.0f / (float)span;
        p1 = a + (hi - a) * ((float)max(ca - k - dl, 0) + 0.5f) * inv;
        p2 = a + (hi - a) * ((float)min(ca - k + dl, span) - 0.5f) * inv;
      }
      // keep both pivots strictly inside (a, b) in key order, p1 <= p2
      uint32_t k1 = f32_key(p1), k2 = f32_key(p2);
      if (!(p1 == p1) || !(p2 == p2) || it >= 12 || k1 <= ka || k2 >= kb || k1 > k2) {
        k1 = ka + max((kb - ka) / 3u, 1u);
        k2 = max(kb - max((kb - ka) / 3u, 1u), k1);
        if (it >= 12) k1 = k2 = ka + ((kb - ka) >> 1);  // plain bisection: bounded
      }
      p1 = key_f32(k1); p2 = key_f32(k2);
      int l1, l2;
      const uint32_t cc = list_count2(lst, m8, p1, p2, l1, l2);
      const int c1 = (int)(cc & 0xffffu), c2 = (int)(cc >> 16);
      if (c2 >= k) { a = p2; ca = c2; lane_ca = l2; }
      else if (c1 >= k) { a = p1; ca = c1; lane_ca = l1; b = p2; cb = c2; lane_cb = l2; }
      else { b = p1; cb = c1; lane_cb = l1; }
    }
    ++it;
  }

  // gather the <= 64 candidates in [a, b): each lane already knows how many of them it
  // holds (lane_ca - lane_cb), so an exclusive scan gives it a private range of slots
  const bool open = cb == 0;  // nothing lies at or above b (b may still be +inf)
  const int mine = lane_ca - lane_cb;
  int incl = mine;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const int u = __shfl_up_sync(0xffffffffu, incl, off);
    if (lane >= off) incl += u;
  }
  float* dst = fin + (incl - mine);
  if (mine > 0) {
#pragma unroll 1
    for (int e = 0; e < m8; e += 8) {
      float x[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) x[u] = lst[(e + u) * 32];
      // (e + u < cnt: the -inf padding beyond the lane's entries is not a candidate -- it passes
      // `>= a` when a itself is -inf, i.e. for a window that holds -inf samples and lists everything)
#pragma unroll
      for (int u = 0; u < 8; ++u)
        if (e + u < cnt && x[u] >= a && (open || x[u] < b)) *dst++ = x[u];
    }
  }
  __syncwarp();
  const int kk = k - cb;       // rank wanted among the candidates, 1-based from the top
  const float2 y = *reinterpret_cast<const float2*>(fin + 2 * lane);  // -inf beyond ca - cb
  uint32_t k0 = f32_key(y.x), k1 = f32_key(y.y);
  sort64_desc(k0, k1, lane);
  if (next_hint) *next_hint = hint_from_sorted(k0, k1, kk, ca - cb, open ? POS_INF : b);
  const uint32_t pick = ((kk - 1) & 1) ? k1 : k0;
  return key_f32(__shfl_sync(0xffffffffu, pick, (kk - 1) >> 1));
}

// The hinted finish (streaming windows): `fl` < `ce` bracket the answer as left by the previous
// call on this series -- every element >= fl is in the lists (total of them, total >= k). One
// counting pass at `ce` leaves the candidates in [fl, ce); if they are at most 64 and the answer is
// among them they are gathered and sorted at once. Returns false (nothing decided) when the
// bracket no longer holds the answer or is too wide: the caller then runs select_from_lists on
// the same lists. Exact like select_from_lists: the result is a listed element whose rank has
// been counted.
__device__ __forceinline__ bool select_hinted(float* lst, float* fin, int lane, int cnt, int total,
                                              int maxcnt, int k, float fl, float ce, float* result,
                                              float2* next_hint, int prefilled = kListPrefill) {
  const int m8 = (maxcnt + 7) & ~7;
  for (int e = max(cnt, prefilled); e < m8; ++e) lst[e * 32] = NEG_INF;  // (entries < prefilled are pre-filled)
  fin[lane] = NEG_INF;
  fin[lane + 32] = NEG_INF;
  __syncwarp();
  int lane_cb, dummy;
  const int cb = (int)(list_count2(lst, m8, ce, ce, lane_cb, dummy) & 0xffffu);
  if (cb >= k || total - cb > kMaxCand) return false;
  const int mine = cnt - lane_cb;
  int incl = mine;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const int u = __shfl_up_sync(0xffffffffu, incl, off);
    if (lane >= off) incl += u;
  }
  float* dst = fin + (incl - mine);
  if (mine > 0) {
#pragma unroll 1
    for (int e = 0; e < m8; e += 8) {
      float x[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) x[u] = lst[(e + u) * 32];
#pragma unroll
      for (int u = 0; u < 8; ++u)
        if (x[u] >= fl && x[u] < ce) *dst++ = x[u];
    }
  }
  __syncwarp();
  const int kk = k - cb;
  if (total - cb <= 32) {  // the usual case (a bracket spans 2 * kHintMargin + 1 ranks): a 32-key sort
    uint32_t key = f32_key(fin[lane]);
    sort32_desc(key, lane);
    *next_hint = hint_from_sorted32(key, kk, total - cb, ce);
    if (total - k < kHintRefresh)  // the floor has drifted close: sample afresh next time
      *next_hint = make_float2(__int_as_float(0x7fc00000), __int_as_float(0x7fc00000));
    *result = key_f32(__shfl_sync(0xffffffffu, key, kk - 1));
    return true;
  }
  const float2 y = *reinterpret_cast<const float2*>(fin + 2 * lane);
  uint32_t k0 = f32_key(y.x), k1 = f32_key(y.y);
  sort64_desc(k0, k1, lane);
  *next_hint = hint_from_sorted(k0, k1, kk, total - cb, ce);
  if (total - k < kHintRefresh)
    *next_hint = make_float2(__int_as_float(0x7fc00000), __int_as_float(0x7fc00000));
  const uint32_t pick = ((kk - 1) & 1) ? k1 : k0;
  *result = key_f32(__shfl_sync(0xffffffffu, pick, (kk - 1) >> 1));
  return true;
}

#endif  // __CUDACC__

}  // namespace klm
