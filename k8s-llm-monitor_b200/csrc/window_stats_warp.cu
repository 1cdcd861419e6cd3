// K7 window_stats, warp variant: one warp per series, the whole window resident on chip
// (windows <= 4096 samples), selection without sorting.
//
// No reference code exists for this (the reference keeps only the latest
// snapshot, internal/metrics/manager.go:29, :289-290); definition in
// window_stats.cuh / DESIGN.md, "parity unpinned".
//
// Per series (n valid samples, k = n - rank95 = how many samples are >= p95):
//   1. 128-bit streaming loads: lane l takes float4 l + 32 j (j < NV), so a warp
//      reads its row as NV fully coalesced 512-byte requests. Pads are -inf. The
//      first NR rows land in registers, the other NV - NR in shared memory
//      (cp.async, each lane its own 16 bytes): at NV = 29 the split 19 + 10 fits
//      128 registers and a fourth CTA per SM; this kernel is bound by the latency
//      its resident warps can hide ("K7 in detail").
//   2. 128 samples (4 register rows, one group of 32 per row): the three largest of
//      each group by three rounds of warp-wide max (redux), the four groups
//      interleaved. Pivot = the smallest of the four "second largest" values:
//      about 10 % of the window lies above it; fewer than k samples do with
//      probability ~5 %, and then the append alone is redone from the smallest
//      "third largest" (~14.5 % above).
//   3. ONE fused pass over the window: min / max / f64 sum, and every element
//      >= the pivot is appended to a per-lane list in shared memory.
//   4. select_from_lists (window_stats.cuh): counting passes over the lists, two
//      pivots at a time, until <= 64 candidates remain; those are gathered (each
//      lane knows its share, an exclusive scan hands out slots) and sorted by a
//      64-key warp bitonic network. Brackets that cannot shrink because of ties
//      are peeled one distinct value at a time.
//   Heavy ties at the pivot (quantised channels, mostly-zero counters) are settled
//   by counting passes over the window at the tied value; anything the
//   shortcuts cannot place (list overflow, unlucky pivot) falls back to an exact
//   bisection over the order-preserving key space.
// Everything is exact: every branch ends on an input element whose rank has
// been counted, so the result is bit-identical to sorting.
#include "common.cuh"
#include "window_stats.cuh"

namespace klm {

#ifndef KLM_WSW_NR  // register rows of the 29-row split instantiation (tuning builds)
#define KLM_WSW_NR 19
#endif
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
// warp 64 candidates (256 B)
constexpr int kWarpAux = 256;                  // per warp: the 64 candidates of the final sort
#ifndef KLM_WSW_EXTRA_SMEM  // experiments: unused shared memory per CTA, to lower the occupancy
#define KLM_WSW_EXTRA_SMEM 0
#endif
constexpr size_t kWswSmemBytes =
    (size_t)kWswWarps * (kLaneListBytes + kWarpAux) + kLaneListBytes + KLM_WSW_EXTRA_SMEM;  // + alignment slack

// if (x > lox) { lane_list[c / 128 mod 64] = x; c += 128; }
// lb = 8 KB-aligned list base | lane * 4, so the slot address is (c & 0x1F80) | lb.
// The address lives in a temporary: the store never holds up the counter.
__device__ __forceinline__ void append_if_gt(uint32_t& c, uint32_t lb, float x, float lox) {
#ifdef KLM_ABL_NOAPPEND
  return;
#endif
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b32 t;\n\t"
      "setp.gt.f32 p, %2, %3;\n\t"
      "lop3.b32 t, %0, 0x1F80, %1, 0xEA;\n\t"
      "@p st.shared.f32 [t], %2;\n\t"
      "@p add.u32 %0, %0, 128;\n\t}"
      : "+r"(c)
      : "r"(lb), "f"(x), "f"(lox));
}

// Ablation switches (tools/build_variant.sh; timing experiments only, the results are then
// wrong): each removes one phase so that its cost can be read off as a difference.
#if defined(KLM_ABL_NOSUM)
#define KLM_SUM4(acc, a, b, c, d) ((void)0)
#elif defined(KLM_ABL_F32SUM)
#define KLM_SUM4(acc, a, b, c, d) acc += (double)(((a) + (b)) + ((c) + (d)))
#else
#define KLM_SUM4(acc, a, b, c, d) acc += ((double)(a) + (double)(b)) + ((double)(c) + (double)(d))
#endif

// The window of one series, spread over a warp: row j (j < NV) is float4 l + 32 j of lane
// l. Rows below NR live in registers, the other NS = NV - NR rows in shared memory, each
// lane reading back only the float4 it loaded itself (no cross-lane traffic, so no
// barrier). NR == NV keeps everything in registers (166 registers at NV = 29: 3 CTAs per
// SM); NR = 19 fits 128 registers, i.e. a fourth CTA per SM, and the kernel is bound by the
// latency the resident warps can hide, not by issue slots ("K7 in detail").
// OVL: the lane lists overlay the shared-memory rows (NS * 512 == kLaneListBytes). The passes
// then visit the shared rows first and in order: after row r has been read, the 4 (r + 1)
// elements seen so far can have produced at most that many list entries, i.e. 512 (r + 1)
// bytes from the start of the block -- only rows that every lane has already read (a warp's
// shared-memory instructions execute in order). Anything that needs the rows again
// afterwards (append retry, bisection) reloads them.
template <int NV, int NR, bool OVL = false>
struct Window {
  static constexpr int NS = NV - NR;
  float v[4 * NR];  // element 4 j + c = sample 4 (lane + 32 j) + c, or -inf
  float tail;       // lane c < (n & 3) holds sample 4 (n / 4) + c, else -inf
  uint32_t srow;    // shared address of this lane's float4 of row NR (row j: + 512 (j - NR))

  __device__ __forceinline__ float4 row(int j) const {
    if (j < NR) return make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
    float4 t;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(t.x), "=f"(t.y), "=f"(t.z), "=f"(t.w)
                 : "r"(srow + (uint32_t)((j - NR) * 512)));
    return t;
  }
  __device__ __forceinline__ void set_row_neg_inf(int j) {
    if (j < NR) {
      v[4 * j] = v[4 * j + 1] = v[4 * j + 2] = v[4 * j + 3] = NEG_INF;
    } else {
      asm volatile("st.shared.v4.f32 [%0], {%1, %1, %1, %1};"
                   :: "r"(srow + (uint32_t)((j - NR) * 512)), "f"(NEG_INF) : "memory");
    }
  }

  // number of samples >= p (all lanes get the warp total)
  __device__ __forceinline__ int count_ge(float p) const {
    int c0 = 0, c1 = 0, c2 = 0, c3 = 0;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const float4 t = row(j);
      inc_if_ge(c0, t.x, p);
      inc_if_ge(c1, t.y, p);
      inc_if_ge(c2, t.z, p);
      inc_if_ge(c3, t.w, p);
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

  __device__ __forceinline__ float4 shared_row(int r) const {  // r may be a run-time value
    float4 t;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(t.x), "=f"(t.y), "=f"(t.z), "=f"(t.w)
                 : "r"(srow + (uint32_t)(r * 512)));
    return t;
  }

  // calls f(j, row j, j & 1) for every row. OVL: shared rows first, the next one fetched
  // before the current one is used (its slot is not yet written by the appends of the
  // current one). Fully unrolled: rolling the shared rows into a loop of four per trip
  // shrinks the code by 1300 instructions but runs 10 % slower (111 vs 101 us at C2).
  template <typename F>
  __device__ __forceinline__ void for_rows(F&& f) const {
    if (OVL) {
      float4 cur = shared_row(0);
#pragma unroll
      for (int r = 0; r < NS; ++r) {
        float4 nxt = cur;
        if (r + 1 < NS) nxt = shared_row(r + 1);
        f(NR + r, cur, (NR + r) & 1);
        cur = nxt;
      }
#pragma unroll
      for (int j = 0; j < NR; ++j) f(j, row(j), j & 1);
    } else {
#pragma unroll
      for (int j = 0; j < NV; ++j) f(j, row(j), j & 1);
    }
  }

  // append(x > lox) only: the retry after a pivot that listed fewer than k elements
  __device__ __forceinline__ void append_pass(float lox, uint32_t lb, uint32_t& c) const {
    for_rows([&](int, float4 t, int) {
      append_if_gt(c, lb, t.x, lox);
      append_if_gt(c, lb, t.y, lox);
      append_if_gt(c, lb, t.z, lox);
      append_if_gt(c, lb, t.w, lox);
    });
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
    for_rows([&](int j, float4 t4, int odd) {
      if (!FULL && j > jr) return;  // rows beyond the boundary row hold only pads (warp-uniform)
      const float a = t4.x, b = t4.y, cc = t4.z, d = t4.w;
      const bool uniform_valid = FULL ? (j < NV - 1) : (j < jr);
      if (uniform_valid) {
        if (odd) {
          mn1 = fminf(fminf(mn1, a), fminf(b, fminf(cc, d)));
          mx1 = fmaxf(fmaxf(mx1, a), fmaxf(b, fmaxf(cc, d)));
          KLM_SUM4(sum1, a, b, cc, d);
        } else {
          mn0 = fminf(fminf(mn0, a), fminf(b, fminf(cc, d)));
          mx0 = fmaxf(fmaxf(mx0, a), fmaxf(b, fmaxf(cc, d)));
          KLM_SUM4(sum0, a, b, cc, d);
        }
      } else if (lane + 32 * j < nfull) {  // boundary row: some lanes hold data
        mn0 = fminf(fminf(mn0, a), fminf(b, fminf(cc, d)));
        mx0 = fmaxf(fmaxf(mx0, a), fmaxf(b, fmaxf(cc, d)));
        KLM_SUM4(sum0, a, b, cc, d);
      }
      append_if_gt(c, lb, a, lox);  // pads are -inf: never appended
      append_if_gt(c, lb, b, lox);
      append_if_gt(c, lb, cc, lox);
      append_if_gt(c, lb, d, lox);
    });
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

template <int NV, int NR, int MINB, bool OVL>
__global__ void __launch_bounds__(kWswThreads, MINB)
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
  constexpr int NS = NV - NR;
  const uint32_t raw_s = (uint32_t)__cvta_generic_to_shared(s_raw);
  const uint32_t base_s = (raw_s + (uint32_t)kLaneListBytes - 1u) & ~((uint32_t)kLaneListBytes - 1u);
  unsigned char* base_g = s_raw + (base_s - raw_s);
  float* lst = reinterpret_cast<float*>(base_g + wib * kLaneListBytes) + lane;  // entry e: lst[e * 32]
  const uint32_t lb = (base_s + (uint32_t)wib * kLaneListBytes) | ((uint32_t)lane << 2);
  unsigned char* aux;
  uint32_t rows_s = 0;  // shared address of this warp's window rows (NS > 0)
  if (NS == 0) {
    aux = base_g + kWswWarps * kLaneListBytes + wib * kWarpAux;
  } else if (OVL) {
    // the rows share the 8 KB list block of the warp; the aux areas sit below the first block
    static_assert(!OVL || NS * 512 == kLaneListBytes, "overlay: the shared rows are exactly the list block");
    if (base_s - raw_s < (uint32_t)(kWswWarps * kWarpAux)) __trap();  // host checked
    aux = s_raw + wib * kWarpAux;
    rows_s = base_s + (uint32_t)(wib * kLaneListBytes);
  } else {
    // Dynamic shared memory starts 1 KB into the CTA's window (the system reserve), so 7 KB
    // precede the first 8 KB boundary: they take the four aux areas and warp 0's rows, the
    // rows of warps 1..3 follow the lists. Nothing is lost to alignment, 4 CTAs fit per SM.
    if (base_s - raw_s < (uint32_t)(kWswWarps * kWarpAux + NS * 512)) __trap();  // host checked
    aux = s_raw + wib * kWarpAux;
    rows_s = wib == 0 ? raw_s + (uint32_t)(kWswWarps * kWarpAux)
                      : base_s + (uint32_t)(kWswWarps * kLaneListBytes + (wib - 1) * NS * 512);
  }
  float* fin = reinterpret_cast<float*>(aux);

  // ---- 1. load. The addresses do not depend on seen[s]: every float4 inside the row
  // pitch is fetched (stale slots beyond n are masked below), so the row loads and the
  // seen[] load share one memory round trip.
  Window<NV, NR, OVL> w;
  w.srow = rows_s + ((uint32_t)lane << 4);
  const int pfull = pitch >> 2;
  const float* p = row + 4 * lane;
  {
    // rows kept in shared memory arrive by asynchronous 16-byte copies, every lane its own
    // float4 (slots outside the pitch stay unwritten and are overwritten with -inf below)
    auto fetch = [&](int j, bool pred) {
      if (j < NR) {
        float4 t = make_float4(NEG_INF, NEG_INF, NEG_INF, NEG_INF);
        if (pred) t = ld_stream_f4(p + 128 * j);
        w.v[4 * j + 0] = t.x; w.v[4 * j + 1] = t.y; w.v[4 * j + 2] = t.z; w.v[4 * j + 3] = t.w;
      } else if (pred) {
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;"
                     :: "r"(w.srow + (uint32_t)((j - NR) * 512)), "l"(p + 128 * j) : "memory");
      }
    };
    if (pfull >= 32 * (NV - 1)) {  // usual case: all rows but the last lie inside the pitch
#pragma unroll
      for (int j = 0; j < NV - 1; ++j) fetch(j, true);
      fetch(NV - 1, lane + 32 * (NV - 1) < pfull);
    } else {
#pragma unroll
      for (int j = 0; j < NV; ++j) fetch(j, lane + 32 * j < pfull);
    }
    if (NS > 0) asm volatile("cp.async.commit_group;" ::: "memory");
  }
  // While the loads are in flight: fill the first kListPrefill entries of this lane's list with -inf,
  // so the counting passes need no padding later (longer lists are padded when they occur).
  // The first kListPrefill entries of all 32 lists are one contiguous block, written as
  // 128-bit stores across the warp (a quarter of the store instructions; the memory
  // instruction queue is what throttles this phase).
  static_assert(kListPrefill % 4 == 0, "whole 512-byte warp stores");
  float4* const lst4 = reinterpret_cast<float4*>(base_g + wib * kLaneListBytes) + lane;
  if (!OVL) {  // (overlay: the block still holds the rows; the lists are padded after the pass)
#pragma unroll
    for (int e = 0; e < kListPrefill / 4; ++e)
      lst4[e * 32] = make_float4(NEG_INF, NEG_INF, NEG_INF, NEG_INF);
    __syncwarp();  // lanes append into slots another lane has just filled
  }
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
  // Ask L2 for the row that a warp will want about one residency later (MINB CTAs of
  // kWswWarps warps per SM are in flight): one bulk-prefetch instruction covers the whole row.
  if (lane == 0) {
    const int64_t later = o + (int64_t)kNumSMs * MINB * kWswWarps;
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
  if (NS > 0) asm volatile("cp.async.wait_group 0;" ::: "memory");  // this lane's own copies
  if (full) {
    if (!(lane + 32 * (NV - 1) < nfull)) w.set_row_neg_inf(NV - 1);
  } else {
#pragma unroll
    for (int j = 0; j < NV; ++j)
      if (j >= jr && !(lane + 32 * j < nfull)) w.set_row_neg_inf(j);
  }
  // overlay only: the appends overwrite the shared rows, so the rare paths that need the
  // window again fetch and mask those rows once more (they are still in L2)
  auto reload_shared_rows = [&]() {
#pragma unroll
    for (int j = NR; j < NV; ++j) {
      if (lane + 32 * j < pfull)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;"
                     :: "r"(w.srow + (uint32_t)((j - NR) * 512)), "l"(p + 128 * j) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
#pragma unroll
    for (int j = NR; j < NV; ++j)
      if (j >= jr && !(lane + 32 * j < nfull)) w.set_row_neg_inf(j);
    __syncwarp();
  };
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
#ifdef KLM_ABL_NOPIVOT
  const bool sampled = false;
#else
  const bool sampled = want_p95 && jr >= kSampledMinRows;
#endif

  // ---- 2. sample pivot: four groups of 32 (one sample row each), top three of each
  if (sampled) {
    float sm[4];
    if (full) {  // the rows are known at compile time
#pragma unroll
      for (int q = 0; q < 4; ++q) sm[q] = w.v[4 * (((2 * q + 1) * (NR - 1)) >> 3) + q];  // register rows
    } else {
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int jq = ((2 * q + 1) * jr) >> 3;
        float t = NEG_INF;
#pragma unroll
        for (int j = 0; j < NV; ++j) {
          if (j == jq) {  // warp-uniform
            const float4 r4 = w.row(j);
            t = q == 0 ? r4.x : q == 1 ? r4.y : q == 2 ? r4.z : r4.w;
          }
        }
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
    if (out_min) out_min[o] = mn + 0.f;  // -0.0 -> +0.0: one canonical zero in the outputs
    if (out_max) out_max[o] = mx + 0.f;
    if (out_mean) out_mean[o] = sum / (double)n;
  }
  if (!want_p95) return;

  // ---- 4. p95 = k-th largest
  float result = 0.f;
  do {  // single-trip block: `break` = result (or the bisect request) is final
    if (bisect) break;
    if (tie_hit) { result = tie_val; break; }
    if (mn == mx) { result = mx; break; }  // constant series
#ifdef KLM_ABL_NOSELECT
    result = lox; break;
#endif
    int cnt = (int)(c128 >> 7);
    int total = warp_sum_i32(cnt);
    int maxcnt = warp_max_i32(cnt);
    if (total < k && !sampled) { result = NEG_INF; break; }  // lox = -inf: only -inf samples are unlisted
    if (total < k && !tie_mode && lox_retry != POS_INF) {
      // the pivot was too high: redo the append (only) from the next sample statistic down
      __syncwarp();
      if (OVL) {
        reload_shared_rows();
      } else {
#pragma unroll
        for (int e = 0; e < kListPrefill / 4; ++e)
          lst4[e * 32] = make_float4(NEG_INF, NEG_INF, NEG_INF, NEG_INF);
        __syncwarp();
      }
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
                               retried, hint_lo, hint_hi, OVL ? 0 : kListPrefill);
  } while (false);
  if (bisect) {
    if (OVL) reload_shared_rows();
    result = w.select_bisect(bis_lo, bis_hi, k, bis_first);
  }
  if (lane == 0) out_p95[o] = result + 0.f;
}

// dynamic shared memory of one CTA: NR == NV: lists (with alignment slack) then aux; else the
// layout described in the kernel, which counts on the 1 KB system reserve before it
template <int NV, int NR, bool OVL>
constexpr size_t wsw_smem_bytes() {
  return NR == NV ? kWswSmemBytes
         : OVL    ? (size_t)(kLaneListBytes - 1024) + (size_t)kWswWarps * kLaneListBytes + KLM_WSW_EXTRA_SMEM
                  : (size_t)(kLaneListBytes - 1024) + (size_t)kWswWarps * kLaneListBytes +
                        (size_t)(kWswWarps - 1) * (NV - NR) * 512 + KLM_WSW_EXTRA_SMEM;
}

template <int NV, int NR, int MINB, bool OVL = false>
static int launch_warp(const float* ring, const uint32_t* seen, int32_t window, int32_t pitch,
                       int64_t s0, int64_t n_series, float* out_min, float* out_max,
                       double* out_mean, float* out_p95, cudaStream_t st) {
  static_assert(KLM_WSW_EXTRA_SMEM != 0 || MINB * (wsw_smem_bytes<NV, NR, OVL>() + 1024) <= 227 * 1024,
                "the CTAs per SM must fit");
  static_assert(NR == NV || OVL || kWswWarps * kWarpAux + (NV - NR) * 512 <= kLaneListBytes - 1024,
                "aux areas + one warp's rows must fit below the first list boundary");
  const int64_t blocks = ceil_div(n_series, kWswWarps);
  if (blocks > 0x7fffffff) {
    set_error("klm_dev_window_stats: too many series for one launch");
    return KLM_ELIMIT;
  }
  static PerDeviceOnce attr_set;
  int dev = 0;
  KLM_CUDA(cudaGetDevice(&dev));
  if (!attr_set.done(dev)) {
    // opt in above 48 KB and ask for the largest shared-memory carve-out
    KLM_CUDA(cudaFuncSetAttribute(window_stats_warp_kernel<NV, NR, MINB, OVL>,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)wsw_smem_bytes<NV, NR, OVL>()));
    KLM_CUDA(cudaFuncSetAttribute(window_stats_warp_kernel<NV, NR, MINB, OVL>,
                                  cudaFuncAttributePreferredSharedMemoryCarveout,
                                  (int)cudaSharedmemCarveoutMaxShared));
    attr_set.set(dev);
  }
  window_stats_warp_kernel<NV, NR, MINB, OVL>
      <<<(unsigned)blocks, kWswThreads, wsw_smem_bytes<NV, NR, OVL>(), st>>>(
          ring, seen, window, pitch, s0, n_series, out_min, out_max, out_mean, out_p95);
  KLM_LAUNCH_CHECK("window_stats_warp_kernel");
  return KLM_OK;
}

// The split layout needs dynamic shared memory to start exactly 1 KB into the CTA's window.
static bool split_layout_ok() {
  static int cached = -1;
  if (cached < 0) {
    int dev = 0, reserved = 0;
    cached = cudaGetDevice(&dev) == cudaSuccess &&
             cudaDeviceGetAttribute(&reserved, cudaDevAttrReservedSharedMemoryPerBlock, dev) == cudaSuccess &&
             reserved == 1024;
  }
  return cached == 1;
}

// Windows up to 4096 samples take the warp kernel; returns 1 if it did not handle the call.
int launch_window_stats_warp(const float* ring, const uint32_t* seen, int32_t window,
                             int32_t pitch, int64_t s0, int64_t s1, float* out_min,
                             float* out_max, double* out_mean, float* out_p95, cudaStream_t st) {
  const int64_t n = s1 - s0;
#define KLM_WSW_GO(NV, NR, MINB) \
  return launch_warp<NV, NR, MINB>(ring, seen, window, pitch, s0, n, out_min, out_max, out_mean, out_p95, st)
  // The pass is straight-line only when the window fills all rows but the last of its
  // instantiation, so there is one per common size class (300, 600, 900, 1500, 1800, 2048,
  // 2500, 3000, 3600 and 4096 samples land in the top row of theirs); other sizes take the
  // generic pass of the next one up. Up to 16 rows fit 128 registers (4 CTAs per SM; 5 at
  // <= 8 rows); from 20 rows on the window is split between registers and shared memory.
#define KLM_WSW_SPLIT_OR_3(NV)                              \
  do {                                                      \
    if (split_layout_ok()) KLM_WSW_GO(NV, KLM_WSW_NR, 4);   \
    KLM_WSW_GO(NV, NV, KLM_WSW_MIN_BLOCKS);                 \
  } while (0)
  if (window <= 3 * 128) KLM_WSW_GO(3, 3, 5);
  if (window <= 5 * 128) KLM_WSW_GO(5, 5, 5);
  if (window <= 8 * 128) KLM_WSW_GO(8, 8, 5);
  if (window <= 12 * 128) KLM_WSW_GO(12, 12, 4);
  if (window <= 15 * 128) KLM_WSW_GO(15, 15, 4);
  if (window <= 16 * 128) KLM_WSW_GO(16, 16, 4);
#ifndef KLM_WSW_NO_SPLIT
  if (window <= 20 * 128) KLM_WSW_SPLIT_OR_3(20);
  if (window <= 24 * 128) KLM_WSW_SPLIT_OR_3(24);
#ifdef KLM_WSW_OVERLAY  // 13 rows in registers, 16 in the list block: 96 registers, 5 CTAs per SM
  if (window <= 29 * 128 && split_layout_ok())
    return launch_warp<29, 13, 5, true>(ring, seen, window, pitch, s0, n, out_min, out_max, out_mean, out_p95, st);
#endif
  if (window <= 29 * 128) KLM_WSW_SPLIT_OR_3(29);
#else
  if (window <= 29 * 128) KLM_WSW_GO(29, 29, KLM_WSW_MIN_BLOCKS);
#endif
  if (window <= 32 * 128) KLM_WSW_GO(32, 32, KLM_WSW_MIN_BLOCKS);
#undef KLM_WSW_SPLIT_OR_3
#undef KLM_WSW_GO
  return 1;
}

}  // namespace klm
