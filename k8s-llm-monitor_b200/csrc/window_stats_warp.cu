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
//   2. 128 samples (4 rows spread over the window) -> their 16 largest, by 16
//      rounds of warp-wide max (redux) over each lane's sorted four.
//   3. ONE fused pass over the registers: min / max / f64 sum, and every element
//      >= the 16th largest sample (about 12 % of the window; P[fewer than k]
//      ~ 0.5 %) is appended to a per-lane list in shared memory. The list area
//      holds a whole window, so the append needs no bounds check.
//   4. A few counting probes over that list (guided by the sample's ranks, then
//      interpolation) narrow the bracket to <= 32 candidates, which are ranked
//      exactly against each other.
//   Heavy ties (quantised channels, mostly-zero counters) are settled by counting
//   passes at the tied value; anything the shortcuts cannot place falls back to
//   an exact bisection over the order-preserving key space.
// Everything is exact: every branch ends on an input element whose rank has
// been counted, so the result is bit-identical to sorting.
#include "common.cuh"
#include "window_stats.cuh"

namespace klm {

constexpr int kWswWarps = 4;               // warps (series) per CTA
constexpr int kWswThreads = kWswWarps * 32;
constexpr int kTopR = 16;                  // sample ranks kept
constexpr int kTieR = 8;                   // t16 == t8 -> treat as a heavy tie
constexpr int kSampledMinRows = 7;         // below 7 full rows (896 samples) list everything
constexpr int kMaxCand = 32;               // candidates ranked exactly at the end

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

// -1 if a >= b else 0 (FSET); summed with IADD3, two compares per add
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

// if (x > lox) { *addr = x; addr += 128; } as three predicated instructions
__device__ __forceinline__ void append_if_gt(uint32_t& addr, float x, float lox) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.gt.f32 p, %1, %2;\n\t"
      "@p st.shared.f32 [%0], %1;\n\t"
      "@p add.u32 %0, %0, 128;\n\t}"
      : "+r"(addr)
      : "f"(x), "f"(lox)
      : "memory");
}

template <int NV>
struct Window {
  static constexpr int NE = NV * 4;
  float v[NE];   // element 4 j + c = sample 4 (lane + 32 j) + c, or -inf
  float tail;    // lane c < (n & 3) holds sample 4 (n / 4) + c, else -inf

  // number of samples >= p (all lanes get the warp total)
  __device__ __forceinline__ int count_ge(float p) const {
    int c0 = 0, c1 = 0;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      c0 -= set_ge(v[4 * j + 0], p) + set_ge(v[4 * j + 1], p);
      c1 -= set_ge(v[4 * j + 2], p) + set_ge(v[4 * j + 3], p);
    }
    c0 -= set_ge(tail, p);
    return warp_sum_i32(c0 + c1);
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

// count of list entries >= p; list is [entry][lane], padded with -inf up to m4 (multiple of 4)
__device__ __forceinline__ int list_count_ge(const float* lst, int m4, float p) {
  int c0 = 0, c1 = 0;
#pragma unroll 2
  for (int e = 0; e < m4; e += 4) {
    const float x0 = lst[(e + 0) * 32], x1 = lst[(e + 1) * 32];
    const float x2 = lst[(e + 2) * 32], x3 = lst[(e + 3) * 32];
    c0 -= set_ge(x0, p) + set_ge(x1, p);
    c1 -= set_ge(x2, p) + set_ge(x3, p);
  }
  return warp_sum_i32(c0 + c1);
}

template <int NV>
constexpr size_t wsw_smem_bytes() {
  // per warp: a list area of NE + 4 entries x 32 lanes, then 32 candidates + a counter
  return (size_t)kWswWarps * ((size_t)(NV * 4 + 4) * 128 + 256);
}

template <int NV>
__global__ void __launch_bounds__(kWswThreads, 3)
window_stats_warp_kernel(const float* __restrict__ ring, const uint32_t* __restrict__ seen,
                         int32_t window, int32_t pitch, int64_t s0, int64_t n_series,
                         float* __restrict__ out_min, float* __restrict__ out_max,
                         double* __restrict__ out_mean, float* __restrict__ out_p95) {
  extern __shared__ __align__(128) unsigned char s_raw[];
  constexpr int kListEntries = NV * 4 + 4;
  constexpr int kWarpBytes = kListEntries * 128 + 256;

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
  float* lst = reinterpret_cast<float*>(s_raw + (size_t)wib * kWarpBytes) + lane;  // entry e: lst[e * 32]
  float* fin = reinterpret_cast<float*>(s_raw + (size_t)wib * kWarpBytes + kListEntries * 128);
  int* fin_cnt = reinterpret_cast<int*>(fin + 32);

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

  const int k = n - p95_rank(n);
  bool bisect = false;             // finish with the exact bisection over the registers
  uint32_t bis_lo = f32_key(NEG_INF), bis_hi = 0, bis_first = 0;
  float lox = POS_INF;             // list predicate: x > lox (+inf: list nothing)
  uint32_t tkey = 0;               // lane r (1..16) holds the r-th largest sample key
  const bool want_p95 = out_p95 != nullptr;
  const bool sampled = want_p95 && jr >= kSampledMinRows;

  // ---- 2. sample pivot
  if (sampled) {
    // four sample rows spread over the valid rows (all 32 lanes hold data there)
    float sm[4];
    if (jr >= NV - 1) {  // window (almost) full: the rows are known at compile time
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
    uint32_t h0 = f32_key(sm[0]), h1 = f32_key(sm[1]), h2 = f32_key(sm[2]), h3 = f32_key(sm[3]);
    // sort the lane's four keys descending (5 compare-exchanges)
#define KLM_CE(a, b) { const uint32_t hi_ = max(a, b), lo_ = min(a, b); a = hi_; b = lo_; }
    KLM_CE(h0, h1) KLM_CE(h2, h3) KLM_CE(h0, h2) KLM_CE(h1, h3) KLM_CE(h1, h2)
#undef KLM_CE
    const uint32_t lane_bit = 1u << lane;
#pragma unroll 1
    for (int r = 1; r <= kTopR; ++r) {
      const uint32_t m = warp_max_u32(h0);
      const unsigned who = __ballot_sync(0xffffffffu, h0 == m);
      if ((who & (0u - who)) == lane_bit) { h0 = h1; h1 = h2; h2 = h3; h3 = 0u; }  // pop one copy
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
        bisect = true; bis_hi = t16 + 1u; bis_first = t16;
      } else {
        lox = V;  // yes: list the c_gt elements > V
      }
    } else {
      lox = next_down(V);  // x > lox  <=>  x >= V
    }
  } else if (want_p95) {
    lox = NEG_INF;  // short window: list every sample
  }

  // ---- 3. fused pass: min / max / f64 sum + append(x > lox)
  float mn0 = POS_INF, mn1 = POS_INF, mx0 = NEG_INF, mx1 = NEG_INF;
  double sum0 = 0.0, sum1 = 0.0;
  const uint32_t list_base = (uint32_t)__cvta_generic_to_shared(lst);
  uint32_t addr = list_base;
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const float a = w.v[4 * j], b = w.v[4 * j + 1], c = w.v[4 * j + 2], d = w.v[4 * j + 3];
    if (j < jr) {  // row valid in every lane (warp-uniform)
      if (j & 1) {
        mn1 = fminf(fminf(mn1, a), fminf(b, fminf(c, d)));
        mx1 = fmaxf(fmaxf(mx1, a), fmaxf(b, fmaxf(c, d)));
        sum1 += ((double)a + (double)b) + ((double)c + (double)d);
      } else {
        mn0 = fminf(fminf(mn0, a), fminf(b, fminf(c, d)));
        mx0 = fmaxf(fmaxf(mx0, a), fmaxf(b, fmaxf(c, d)));
        sum0 += ((double)a + (double)b) + ((double)c + (double)d);
      }
    } else if (lane + 32 * j < nfull) {  // boundary row: some lanes hold data
      mn0 = fminf(fminf(mn0, a), fminf(b, fminf(c, d)));
      mx0 = fmaxf(fmaxf(mx0, a), fmaxf(b, fmaxf(c, d)));
      sum0 += ((double)a + (double)b) + ((double)c + (double)d);
    }
    append_if_gt(addr, a, lox);  // pads are -inf: never appended
    append_if_gt(addr, b, lox);
    append_if_gt(addr, c, lox);
    append_if_gt(addr, d, lox);
  }
  if (lane < (n & 3)) {
    mn0 = fminf(mn0, w.tail);
    mx0 = fmaxf(mx0, w.tail);
    sum0 += (double)w.tail;
  }
  append_if_gt(addr, w.tail, lox);
  float mn = fminf(mn0, mn1), mx = fmaxf(mx0, mx1);
  double sum = sum0 + sum1;
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
    if (mn == mx) { result = mx; break; }  // constant series
    const int cnt = (int)((addr - list_base) >> 7);
    const int total = warp_sum_i32(cnt);
    const int maxcnt = warp_max_i32(cnt);
    if (total < k) {
      // the sample pivot was too high (or, with lox = -inf, the answer is a -inf sample)
      if (!sampled) { result = NEG_INF; break; }
      bisect = true; bis_hi = f32_key(mx) + 1u; bis_first = 0u;
      break;
    }
    const int m4 = (maxcnt + 3) & ~3;
    for (int e = cnt; e < m4; ++e) lst[e * 32] = NEG_INF;  // pad to the common length
    if (lane == 0) *fin_cnt = 0;
    fin[lane] = NEG_INF;
    __syncwarp();

    // narrow [a, b): count_ge(a) = ca >= k > cb = count_ge(b)
    float a = sampled ? next_up(lox) : mn, b = POS_INF;
    int ca = total, cb = 0;
    if (sampled) {
      // sample-guided probes: the answer's expected sample rank is k * 129 / n
      const int rt = (k * 129) / n;
      const int r_hi = max(1, rt - 1), r_lo = min(kTopR - 1, rt + 2);
      const float pb = key_f32(__shfl_sync(0xffffffffu, tkey, r_hi));
      const float pa = key_f32(__shfl_sync(0xffffffffu, tkey, r_lo));
      if (pb > a) {
        const int c = list_count_ge(lst, m4, pb);
        if (c >= k) { a = pb; ca = c; } else { b = pb; cb = c; }
      }
      if (pa > a && pa < b) {
        const int c = list_count_ge(lst, m4, pa);
        if (c >= k) { a = pa; ca = c; } else { b = pa; cb = c; }
      }
    }
    int it = 0;
#pragma unroll 1
    while (ca - cb > kMaxCand) {
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
      const int c = list_count_ge(lst, m4, p);
      if (c >= k) { a = p; ca = c; } else { b = p; cb = c; }
      ++it;
    }
    if (ca - cb > kMaxCand) { result = a; break; }  // all candidates are copies of a

    // gather the <= 32 candidates in [a, b) (a smem counter hands out the slots;
    // only ~1 entry per lane qualifies) and rank them against each other
#pragma unroll 1
    for (int e = 0; e < maxcnt; ++e) {
      const float x = lst[e * 32];
      // cb == 0: nothing lies at or above b (b may still be the initial +inf)
      if (x >= a && (cb == 0 || x < b)) fin[atomicAdd(fin_cnt, 1)] = x;
    }
    __syncwarp();
    const int mcand = ca - cb;   // == *fin_cnt
    const int kk = k - cb;       // rank wanted among the candidates, 1-based from the top
    const float y = fin[lane];   // -inf beyond mcand
    int gt = 0, ge = 0;
    const int mc4 = (mcand + 3) & ~3;
#pragma unroll 1
    for (int i = 0; i < mc4; i += 4) {
      const float4 z = *reinterpret_cast<const float4*>(fin + i);
      gt -= set_gt(z.x, y) + set_gt(z.y, y);
      ge -= set_ge(z.x, y) + set_ge(z.y, y);
      gt -= set_gt(z.z, y) + set_gt(z.w, y);
      ge -= set_ge(z.z, y) + set_ge(z.w, y);
    }
    // lanes >= mcand hold y = -inf: every pad compares >= to them, so they can never
    // satisfy kk <= ge with gt < kk unless kk > mcand, which cannot happen
    const unsigned okm = __ballot_sync(0xffffffffu, lane < mcand && gt < kk && kk <= ge);
    result = __shfl_sync(0xffffffffu, y, __ffs(okm) - 1);
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
  constexpr size_t smem = wsw_smem_bytes<NV>();
  static bool attr_set[64] = {};
  int dev = 0;
  KLM_CUDA(cudaGetDevice(&dev));
  if (smem > 48 * 1024 && !(dev < 64 && attr_set[dev])) {
    KLM_CUDA(cudaFuncSetAttribute(window_stats_warp_kernel<NV>,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (dev < 64) attr_set[dev] = true;
  }
  window_stats_warp_kernel<NV><<<(unsigned)blocks, kWswThreads, smem, st>>>(
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
