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
//   2. min / max / f64 sum in one pass over the registers, warp-shuffle reduce.
//   3. p95 = k-th largest, found by narrowing, never by sorting the window:
//      a. 128 samples (4 rows spread over the window) -> their 16 largest, by
//         16 rounds of warp-wide max (redux) over each lane's sorted four;
//      b. every element >= the 16th largest sample (about 12 % of the window;
//         P[fewer than k] ~ 0.5 %) is appended to a per-lane list in shared memory;
//      c. a few counting probes over that list (guided by the sample's ranks,
//         then interpolation) narrow the bracket to <= 32 candidates;
//      d. the candidates are ranked exactly against each other.
//      Heavy ties (quantised channels, mostly-zero counters) are settled by two
//      counting passes at the tied value; anything the shortcuts cannot place
//      falls back to an exact bisection over the order-preserving key space.
// Everything is exact: every branch ends on an input element whose rank has
// been counted, so the result is bit-identical to sorting.
#include "common.cuh"
#include "window_stats.cuh"

namespace klm {

constexpr int kWswWarps = 4;               // warps (series) per CTA
constexpr int kWswThreads = kWswWarps * 32;
constexpr int kListCap = 32;               // list entries per lane (4 KB per warp)
constexpr int kTopR = 16;                  // sample ranks kept
constexpr int kTieR = 8;                   // t16 == t8 -> treat as a heavy tie
constexpr int kSampledMinRows = 7;         // below 7 full rows (896 samples) list everything

#define NEG_INF __int_as_float(0xff800000)
#define POS_INF __int_as_float(0x7f800000)

__device__ __forceinline__ float next_up(float x) {   // smallest float > x (x finite or -inf)
  return key_f32(f32_key(x) + 1u);
}
__device__ __forceinline__ float next_down(float x) { // largest float < x (x finite or +inf)
  return key_f32(f32_key(x) - 1u);
}
__device__ __forceinline__ int warp_max_i32(int v) { return __reduce_max_sync(0xffffffffu, v); }
__device__ __forceinline__ uint32_t warp_max_u32(uint32_t v) { return __reduce_max_sync(0xffffffffu, v); }

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
      c0 += v[4 * j + 0] >= p;
      c1 += v[4 * j + 1] >= p;
      c2 += v[4 * j + 2] >= p;
      c3 += v[4 * j + 3] >= p;
    }
    c0 += tail >= p;
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
};

// count of list entries >= p; list is [entry][lane], padded with -inf up to maxcnt
__device__ __forceinline__ int list_count_ge(const float* lst, int maxcnt, float p) {
  int c = 0;
  for (int e = 0; e < maxcnt; ++e) c += lst[e * 32] >= p;
  return warp_sum_i32(c);
}

template <int NV>
__global__ void __launch_bounds__(kWswThreads, 3)
window_stats_warp_kernel(const float* __restrict__ ring, const uint32_t* __restrict__ seen,
                         int32_t window, int32_t pitch, int64_t s0, int64_t n_series,
                         float* __restrict__ out_min, float* __restrict__ out_max,
                         double* __restrict__ out_mean, float* __restrict__ out_p95) {
  __shared__ float s_list[kWswWarps][kListCap * 32];
  __shared__ float s_fin[kWswWarps][32];

  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int64_t o = (int64_t)blockIdx.x * kWswWarps + wib;  // output slot
  if (o >= n_series) return;                                  // no block-wide barrier below
  const int64_t s = s0 + o;
  const uint32_t sn = seen[s];
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
  const float* row = ring + s * (int64_t)pitch;
  const int nfull = n >> 2;   // complete float4 groups
  const int jr = nfull >> 5;  // rows valid in every lane

  // ---- 1. load
  Window<NV> w;
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const int q = lane + 32 * j;
    float4 t = make_float4(NEG_INF, NEG_INF, NEG_INF, NEG_INF);
    if (q < nfull) t = ld_stream_f4(row + 4 * q);
    w.v[4 * j + 0] = t.x;
    w.v[4 * j + 1] = t.y;
    w.v[4 * j + 2] = t.z;
    w.v[4 * j + 3] = t.w;
  }
  w.tail = (lane < (n & 3)) ? row[4 * nfull + lane] : NEG_INF;

  // ---- 2. min / max / sum
  float mn = POS_INF, mx = NEG_INF;
  double sum = 0.0;
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    if (lane + 32 * j < nfull) {
      const float a = w.v[4 * j], b = w.v[4 * j + 1], c = w.v[4 * j + 2], d = w.v[4 * j + 3];
      mn = fminf(fminf(mn, a), fminf(b, fminf(c, d)));
      mx = fmaxf(fmaxf(mx, a), fmaxf(b, fmaxf(c, d)));
      sum += ((double)a + (double)b) + ((double)c + (double)d);
    }
  }
  if (lane < (n & 3)) {
    mn = fminf(mn, w.tail);
    mx = fmaxf(mx, w.tail);
    sum += (double)w.tail;
  }
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
  if (!out_p95) return;

  // ---- 3. p95 = k-th largest
  const int k = n - p95_rank(n);
  float result = 0.f;
  bool bisect = false;             // finish with the exact bisection over the registers
  uint32_t bis_lo = 0, bis_hi = 0, bis_first = 0;
  float* lst = &s_list[wib][lane];
  do {  // single-trip block: `break` = result is final
    if (mn == mx) { result = mx; break; }  // constant series

    float lox = NEG_INF;  // list predicate: x > lox
    uint32_t tkey = 0;    // lane r (1..16) holds the r-th largest sample key
    const bool sampled = jr >= kSampledMinRows;
    if (sampled) {
      // 3a. four sample rows spread over the valid rows (all 32 lanes valid there)
      float sm[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int jq = ((2 * q + 1) * jr) >> 3;
        float t = NEG_INF;
#pragma unroll
        for (int j = 0; j < NV; ++j) t = (j == jq) ? w.v[4 * j + q] : t;
        sm[q] = t;
      }
      uint32_t h0 = f32_key(sm[0]), h1 = f32_key(sm[1]), h2 = f32_key(sm[2]), h3 = f32_key(sm[3]);
      // sort the lane's four keys descending (5 compare-exchanges)
#define KLM_CE(a, b) { const uint32_t hi_ = max(a, b), lo_ = min(a, b); a = hi_; b = lo_; }
      KLM_CE(h0, h1) KLM_CE(h2, h3) KLM_CE(h0, h2) KLM_CE(h1, h3) KLM_CE(h1, h2)
#undef KLM_CE
#pragma unroll 1
      for (int r = 1; r <= kTopR; ++r) {
        const uint32_t m = warp_max_u32(h0);
        const unsigned who = __ballot_sync(0xffffffffu, h0 == m);
        if (lane == __ffs(who) - 1) { h0 = h1; h1 = h2; h2 = h3; h3 = 0u; }  // pop one copy
        if (lane == r) tkey = m;
      }
      const uint32_t t16 = __shfl_sync(0xffffffffu, tkey, kTopR);
      const uint32_t t8 = __shfl_sync(0xffffffffu, tkey, kTieR);
      const float V = key_f32(t16);
      if (t16 == t8) {
        // heavy tie at V: is the answer above the block of copies of V?
        const int c_gt = w.count_ge(next_up(V));
        if (c_gt < k) {
          // no: it is V itself (found by the first probe) or below it
          bisect = true; bis_lo = f32_key(mn); bis_hi = t16 + 1u; bis_first = t16;
          break;
        }
        lox = V;  // yes: list the c_gt elements > V
      } else {
        lox = next_down(V);  // x > lox  <=>  x >= V
      }
    }

    // 3b. append every element > lox to this lane's list: entry e at lst[e * 32]
    int cnt = 0;
#pragma unroll
    for (int i = 0; i < Window<NV>::NE; ++i) {
      const float x = w.v[i];
      if (x > lox) {
        if (cnt < kListCap) lst[cnt * 32] = x;
        ++cnt;
      }
    }
    if (w.tail > lox) {
      if (cnt < kListCap) lst[cnt * 32] = w.tail;
      ++cnt;
    }
    const int total = warp_sum_i32(cnt);
    const int maxcnt = warp_max_i32(cnt);
    if (maxcnt > kListCap || total < k) {
      // list overflow, or the sample pivot was too high: exact bisection over the registers.
      // (total < k with lox = -inf means the answer is a -inf sample.)
      if (!sampled && total < k) { result = NEG_INF; break; }
      bisect = true; bis_lo = f32_key(mn); bis_hi = f32_key(mx) + 1u; bis_first = 0u;
      break;
    }
    for (int e = cnt; e < maxcnt; ++e) lst[e * 32] = NEG_INF;  // pad to the common length
    __syncwarp();

    // 3c. narrow [a, b): count_ge(a) = ca >= k > cb = count_ge(b)
    float a = sampled ? next_up(lox) : mn, b = POS_INF;
    int ca = total, cb = 0;
    if (!sampled) a = mn;            // lox = -inf: every sample is listed, a = min
    if (sampled) {
      // sample-guided probes: the answer's expected sample rank is k * 129 / n
      const int rt = (k * 129) / n;
      const int r_hi = max(1, rt - 1), r_lo = min(kTopR - 1, rt + 2);
      const float pb = key_f32(__shfl_sync(0xffffffffu, tkey, r_hi));
      const float pa = key_f32(__shfl_sync(0xffffffffu, tkey, r_lo));
      if (pb > a) {
        const int c = list_count_ge(lst, maxcnt, pb);
        if (c >= k) { a = pb; ca = c; } else { b = pb; cb = c; }
      }
      if (pa > a && pa < b) {
        const int c = list_count_ge(lst, maxcnt, pa);
        if (c >= k) { a = pa; ca = c; } else { b = pa; cb = c; }
      }
    }
    int it = 0;
    while (ca - cb > 32) {
      const uint32_t ka = f32_key(a), kb = (b == POS_INF) ? f32_key(mx) + 1u : f32_key(b);
      if (kb - ka <= 1u) break;  // no key strictly between: every candidate equals a
      float p;
      if (it < 3 || (it < 12 && (it & 1))) {
        // interpolate on the counts towards rank k (value-linear inside the bracket)
        const float hi = (b == POS_INF) ? mx : b;
        const float frac = ((float)(ca - k) + 0.5f) / (float)(ca - cb);
        p = a + (hi - a) * frac;
        if (!(p > a)) p = next_up(a);
        if (!(p < b) || !(p <= mx)) p = key_f32(ka + ((kb - ka) >> 1));
      } else {
        p = key_f32(ka + ((kb - ka) >> 1));  // guaranteed progress: halve the key range
      }
      const int c = list_count_ge(lst, maxcnt, p);
      if (c >= k) { a = p; ca = c; } else { b = p; cb = c; }
      ++it;
    }
    if (ca - cb > 32) { result = a; break; }  // all candidates are copies of a

    // 3d. gather the <= 32 candidates in [a, b) and rank them against each other
    float* fin = s_fin[wib];
    int base = 0;
    for (int e = 0; e < maxcnt; ++e) {
      const float x = lst[e * 32];
      const bool in = (x >= a) && (x < b);
      const unsigned m = __ballot_sync(0xffffffffu, in);
      if (in) fin[base + __popc(m & ((1u << lane) - 1u))] = x;
      base += __popc(m);
    }
    __syncwarp();
    const int mcand = base;      // == ca - cb
    const int kk = k - cb;       // rank wanted among the candidates, 1-based from the top
    const float y = lane < mcand ? fin[lane] : NEG_INF;
    int gt = 0, ge = 0;
    for (int i = 0; i < mcand; ++i) {
      const float z = fin[i];
      gt += z > y;
      ge += z >= y;
    }
    const unsigned okm = __ballot_sync(0xffffffffu, lane < mcand && gt < kk && kk <= ge);
    result = __shfl_sync(0xffffffffu, y, __ffs(okm) - 1);
    __syncwarp();  // fin / list are reused by nobody else, but keep lanes together
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
  window_stats_warp_kernel<NV><<<(unsigned)blocks, kWswThreads, 0, st>>>(
      ring, seen, window, pitch, s0, n_series, out_min, out_max, out_mean, out_p95);
  KLM_LAUNCH_CHECK("window_stats_warp_kernel");
  return KLM_OK;
}

// Windows up to 4096 samples take the warp kernel; returns 1 if it did not handle the call.
int launch_window_stats_warp(const float* ring, const uint32_t* seen, int32_t window,
                             int32_t pitch, int64_t s0, int64_t s1, float* out_min,
                             float* out_max, double* out_mean, float* out_p95, cudaStream_t st) {
  const int64_t n = s1 - s0;
  if (window <= 8 * 128) return launch_warp<8>(ring, seen, window, pitch, s0, n, out_min, out_max, out_mean, out_p95, st);
  if (window <= 16 * 128) return launch_warp<16>(ring, seen, window, pitch, s0, n, out_min, out_max, out_mean, out_p95, st);
  if (window <= 29 * 128) return launch_warp<29>(ring, seen, window, pitch, s0, n, out_min, out_max, out_mean, out_p95, st);
  if (window <= 32 * 128) return launch_warp<32>(ring, seen, window, pitch, s0, n, out_min, out_max, out_mean, out_p95, st);
  return 1;
}

}  // namespace klm
