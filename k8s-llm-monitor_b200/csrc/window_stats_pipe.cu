// K7 window_stats, pipelined variant (default for windows <= 4096 samples).
//
// One persistent CTA per SM, 16 warps. Warp 0 is a producer: for every series the CTA
// owns it issues ONE cp.async.bulk (TMA bulk copy, global -> shared) of the whole row
// into a ring of shared-memory slots; completion is signalled on a per-slot mbarrier
// (`full`), reuse is gated by a second one (`empty`). Warps 1..15 are consumers, one
// series each: they wait on `full`, work on the row where it lies in shared memory
// (128-bit LDS, conflict-free), release the slot as soon as the single pass over the
// row is done, and finish the percentile from a small candidate list. The window is
// never held in registers, so the kernel needs ~1/2 the registers of the
// register-resident variant (window_stats_warp.cu) and keeps 15 series per SM in
// flight with no exposed global-load latency.
//
// Per series (n valid samples, k = n - rank95 samples are >= p95) a consumer does:
//   1. 128 samples evenly spaced over the window, four groups of 32; the three largest
//      of each group by redux.sync.max. Pivot = the smallest "second largest" (about
//      10 % of the window lies above it).
//   2. ONE pass over the row: min / max / f64 sum + every element >= pivot appended to a
//      per-lane list (32 entries/lane). If fewer than k qualified (~5 %) the append is
//      redone with the smallest "third largest"; if a lane overflowed (~1 %) with the
//      smallest "largest". Ties at the pivot are settled by one counting pass.
//   3. slot released; select_from_lists() narrows the list to <= 64 candidates and sorts.
//   Whatever these shortcuts cannot place is found by bisection over the key space on
//   the row in global memory (L2-resident; rare).
// Exact by construction, like the other variants. No reference code exists for window
// statistics (internal/metrics/manager.go:29, :289-290 keep one snapshot): parity unpinned.
#include "common.cuh"
#include "window_stats.cuh"

namespace klm {

constexpr int kPipeConsumers = 15;
constexpr int kPipeThreads = (kPipeConsumers + 1) * 32;
constexpr int kPipeLaneCap = 32;                   // list entries per lane
constexpr int kPipeListBytes = kPipeLaneCap * 128;  // [entry][lane] f32, 4 KB, self-aligned
constexpr int kPipeFinBytes = 512;                 // 64 candidates (+ spare)
constexpr int kPipeMaxSlots = 16;
constexpr int kPipeFixedBytes = kPipeListBytes /* alignment slack */ +
                                kPipeConsumers * (kPipeListBytes + kPipeFinBytes) + 512 /* barriers + counters */;
constexpr int kPipeMinRowsSampled = 7;             // below 896 samples list everything
constexpr long long kSpinCycles = 4000000000LL;   // ~2 s: a stuck barrier traps instead of hanging

// ---- mbarrier / bulk-copy primitives (shared-window addresses)
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
#ifdef KLM_PIPE_DEBUG
__device__ int g_pipe_dbg[8];
#define KLM_PIPE_FAIL(site)                                              \
  do {                                                                   \
    if (atomicCAS(&g_pipe_dbg[0], 0, (site)) == 0) {                     \
      g_pipe_dbg[1] = (int)blockIdx.x; g_pipe_dbg[2] = (int)threadIdx.x; \
      g_pipe_dbg[3] = (int)bar; g_pipe_dbg[4] = (int)parity;             \
    }                                                                    \
    asm volatile("exit;");                                               \
  } while (0)
#else
#define KLM_PIPE_FAIL(site) __trap()
#endif

__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity, int site = 1) {
  const long long t0 = clock64();
  for (;;) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    if (ok) return;
    __nanosleep(20);  // back off: a polling warp must not take issue slots from working ones
    if (clock64() - t0 > kSpinCycles) KLM_PIPE_FAIL(site);
  }
}
// whole-warp wait: every lane polls the same barrier, so the warp stays converged (a
// lane-0-only wait makes ptxas emit the slow non-converged form of every later collective)
__device__ __forceinline__ void mbar_wait_warp(uint32_t bar, uint32_t parity, int lane, int site) {
  (void)lane;
  mbar_wait(bar, parity, site);
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
      "l"(src), "r"(bytes), "r"(bar)
      : "memory");
}

// if (x > lox) { lane_list[(c / 128) mod 32] = x; c += 128; }   lb = 4 KB-aligned base | lane * 4
__device__ __forceinline__ void pipe_append(uint32_t& c, uint32_t lb, float x, float lox) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b32 t;\n\t"
      "setp.gt.f32 p, %2, %3;\n\t"
      "lop3.b32 t, %0, 0xF80, %1, 0xEA;\n\t"
      "@p st.shared.f32 [t], %2;\n\t"
      "@p add.u32 %0, %0, 128;\n\t}"
      : "+r"(c)
      : "r"(lb), "f"(x), "f"(lox));
}

struct RowView {
  const float* f;    // the row in shared memory
  const float4* v4;  // (float4*)f + lane
  int lane, n, nfull, jr;
};

// One pass over the row in shared memory: optional min / max / f64 sum, and append(x > lox).
template <bool STATS>
__device__ __forceinline__ void pipe_pass(const RowView& r, float lox, uint32_t lb, uint32_t& c,
                                          float& mn, float& mx, double& sum) {
  float mn0 = POS_INF, mn1 = POS_INF, mx0 = NEG_INF, mx1 = NEG_INF;
  double sum0 = 0.0, sum1 = 0.0;
  int j = 0;
#pragma unroll 2
  for (; j + 2 <= r.jr; j += 2) {  // rows valid in every lane, two per trip
    const float4 x = r.v4[j * 32], y = r.v4[(j + 1) * 32];
    if (STATS) {
      mn0 = fminf(fminf(mn0, x.x), fminf(x.y, fminf(x.z, x.w)));
      mx0 = fmaxf(fmaxf(mx0, x.x), fmaxf(x.y, fmaxf(x.z, x.w)));
      sum0 += ((double)x.x + (double)x.y) + ((double)x.z + (double)x.w);
      mn1 = fminf(fminf(mn1, y.x), fminf(y.y, fminf(y.z, y.w)));
      mx1 = fmaxf(fmaxf(mx1, y.x), fmaxf(y.y, fmaxf(y.z, y.w)));
      sum1 += ((double)y.x + (double)y.y) + ((double)y.z + (double)y.w);
    }
    pipe_append(c, lb, x.x, lox); pipe_append(c, lb, x.y, lox);
    pipe_append(c, lb, x.z, lox); pipe_append(c, lb, x.w, lox);
    pipe_append(c, lb, y.x, lox); pipe_append(c, lb, y.y, lox);
    pipe_append(c, lb, y.z, lox); pipe_append(c, lb, y.w, lox);
  }
  // odd full row and the boundary row (only lanes below nfull - 32 jr hold data there)
  for (; j <= r.jr; ++j) {
    if (r.lane + 32 * j < r.nfull) {
      const float4 x = r.v4[j * 32];
      if (STATS) {
        mn0 = fminf(fminf(mn0, x.x), fminf(x.y, fminf(x.z, x.w)));
        mx0 = fmaxf(fmaxf(mx0, x.x), fmaxf(x.y, fmaxf(x.z, x.w)));
        sum0 += ((double)x.x + (double)x.y) + ((double)x.z + (double)x.w);
      }
      pipe_append(c, lb, x.x, lox); pipe_append(c, lb, x.y, lox);
      pipe_append(c, lb, x.z, lox); pipe_append(c, lb, x.w, lox);
    }
  }
  if (r.lane < (r.n & 3)) {  // tail: n not a multiple of 4
    const float t = r.f[4 * r.nfull + r.lane];
    if (STATS) {
      mn0 = fminf(mn0, t);
      mx0 = fmaxf(mx0, t);
      sum0 += (double)t;
    }
    pipe_append(c, lb, t, lox);
  }
  asm volatile("" ::: "memory");  // the appends are plain shared-memory stores
  if (STATS) {
    mn = fminf(mn0, mn1);
    mx = fmaxf(mx0, mx1);
    sum = sum0 + sum1;
  }
}

// Counts of samples >= p1 and >= p2 over the row in shared memory: c1 | c2 << 16.
__device__ __forceinline__ uint32_t pipe_count2(const RowView& r, float p1, float p2) {
  int a0 = 0, a1 = 0, b0 = 0, b1 = 0;
#pragma unroll 2
  for (int j = 0; j <= r.jr; ++j) {
    if (r.lane + 32 * j < r.nfull) {
      const float4 x = r.v4[j * 32];
      inc_if_ge(a0, x.x, p1); inc_if_ge(a1, x.y, p1); inc_if_ge(a0, x.z, p1); inc_if_ge(a1, x.w, p1);
      inc_if_ge(b0, x.x, p2); inc_if_ge(b1, x.y, p2); inc_if_ge(b0, x.z, p2); inc_if_ge(b1, x.w, p2);
    }
  }
  if (r.lane < (r.n & 3)) {
    const float t = r.f[4 * r.nfull + r.lane];
    inc_if_ge(a0, t, p1);
    inc_if_ge(b0, t, p2);
  }
  return (uint32_t)warp_sum_i32((a0 + a1) + ((b0 + b1) << 16));
}

// Exact k-th largest by bisection over keys, counting on the row in GLOBAL memory (the slot
// has been released; the row was just read, so these loads hit L2). Needs
// count_ge(key_f32(lo)) >= k > count_ge(key_f32(hi)). Rare path.
__device__ __noinline__ float pipe_bisect_global(const float* row, int n, int lane, uint32_t lo,
                                                 uint32_t hi, int k) {
  while (hi - lo > 1u) {
    const uint32_t mid = lo + ((hi - lo) >> 1);
    const float p = key_f32(mid);
    int c = 0;
    for (int t = lane; t < n; t += 32) c += row[t] >= p;
    if (warp_sum_i32(c) >= k) lo = mid; else hi = mid;
  }
  return key_f32(lo);
}

__global__ void __launch_bounds__(kPipeThreads, 1)
window_stats_pipe_kernel(const float* __restrict__ ring, const uint32_t* __restrict__ seen,
                         int32_t window, int32_t pitch, int64_t s0, int64_t n_series,
                         float* __restrict__ out_min, float* __restrict__ out_max,
                         double* __restrict__ out_mean, float* __restrict__ out_p95, int n_slots,
                         int slot_bytes) {
  extern __shared__ unsigned char s_raw[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  // shared layout (from a 4 KB boundary): lane lists [15][4 KB] | candidates [15][512 B] |
  // barriers full[16], empty[16] | slots [n_slots][slot_bytes]
  const uint32_t raw_s = (uint32_t)__cvta_generic_to_shared(s_raw);
  const uint32_t base_s = (raw_s + (uint32_t)kPipeListBytes - 1u) & ~((uint32_t)kPipeListBytes - 1u);
  unsigned char* base_g = s_raw + (base_s - raw_s);
  constexpr uint32_t kFinOff = kPipeConsumers * kPipeListBytes;
  constexpr uint32_t kBarOff = kFinOff + kPipeConsumers * kPipeFinBytes;
  constexpr uint32_t kSlotOff = kBarOff + 512;
  const uint32_t full_s = base_s + kBarOff, empty_s = full_s + kPipeMaxSlots * 8;
  // drained_s[i]: how many rounds of slot i have been drained. An mbarrier wait names its
  // phase only by parity, so a waiter must never be two phases away from the barrier: a
  // consumer first sees drained[slot] == its round, and only then waits on full[slot].
  const uint32_t drained_s = empty_s + kPipeMaxSlots * 8;
  // next_s: the next series (sequence number within this CTA) nobody has claimed yet. Any
  // idle consumer takes it, so a filled slot never waits for one particular warp.
  const uint32_t next_s = drained_s + kPipeMaxSlots * 4;

  if (tid == 0) {
    for (int i = 0; i < n_slots; ++i) {
      mbar_init(full_s + 8 * i, 1);   // one arrive.expect_tx by the producer (+ the copy's bytes)
      mbar_init(empty_s + 8 * i, 1);  // one arrive by the consumer that drained the slot
      asm volatile("st.shared.u32 [%0], %1;" ::"r"(drained_s + 4 * i), "r"(0u) : "memory");
    }
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(next_s), "r"(0u) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();  // the only block-wide barrier: warps part ways below

  const int64_t G = gridDim.x;
  const int64_t my_count = blockIdx.x < n_series ? (n_series - blockIdx.x + G - 1) / G : 0;
  const uint32_t row_bytes = (uint32_t)pitch * 4u;

  if (warp == 0) {
    // ---------------------------------------------------------------- producer
    if (lane == 0) {
      for (int64_t q = 0; q < my_count; ++q) {
        const int slot = (int)(q % n_slots);
        const uint32_t round = (uint32_t)(q / n_slots);
        mbar_wait(empty_s + 8 * slot, (round & 1u) ^ 1u, 100 + (int)q);  // first round: passes at once
        mbar_expect_tx(full_s + 8 * slot, row_bytes);
        bulk_g2s(base_s + kSlotOff + (uint32_t)slot * (uint32_t)slot_bytes,
                 ring + (s0 + blockIdx.x + q * G) * (int64_t)pitch, row_bytes, full_s + 8 * slot);
      }
    }
    return;
  }

  // ------------------------------------------------------------------ consumers
  const int c_id = warp - 1;
  float* lst = reinterpret_cast<float*>(base_g + c_id * kPipeListBytes) + lane;  // entry e: lst[e * 32]
  float* fin = reinterpret_cast<float*>(base_g + kFinOff + c_id * kPipeFinBytes);
  const uint32_t lb = (base_s + (uint32_t)c_id * kPipeListBytes) | ((uint32_t)lane << 2);
  const bool want_p95 = out_p95 != nullptr;

  for (;;) {
    uint32_t q32 = 0;
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.eq.u32 p, %2, 0;\n\t"
        "@p atom.shared.add.u32 %0, [%1], 1;\n\t}"
        : "+r"(q32) : "r"(next_s), "r"((uint32_t)lane) : "memory");
    const int64_t q = __shfl_sync(0xffffffffu, q32, 0);
    if (q >= my_count) break;
    const int64_t o = blockIdx.x + q * G;  // output slot
    const int64_t s = s0 + o;
    const uint32_t sn = seen[s];
    // while the row is on its way: fill this lane's list with -inf (entries a pass does not
    // overwrite must compare below every pivot)
#pragma unroll
    for (int e = 0; e < kListPrefill; ++e) lst[e * 32] = NEG_INF;

    const int slot = (int)(q % n_slots);
    const uint32_t round = (uint32_t)(q / n_slots);
    if (round > 0) {  // the previous round of this slot must have been drained
      const long long t0 = clock64();
      for (;;) {
        uint32_t d;
        asm volatile("ld.acquire.cta.shared.u32 %0, [%1];" : "=r"(d) : "r"(drained_s + 4 * slot) : "memory");
        if (d >= round) break;
        __nanosleep(20);
        if (clock64() - t0 > kSpinCycles) {
          const uint32_t bar = drained_s + 4 * slot, parity = round;
          (void)bar; (void)parity;
          KLM_PIPE_FAIL(300 + (int)q);
        }
      }
    }
    mbar_wait_warp(full_s + 8 * slot, round & 1u, lane, 200 + (int)q);
    const float* sf = reinterpret_cast<const float*>(base_g + kSlotOff + (size_t)slot * slot_bytes);

    const int n = sn < (uint32_t)window ? (int)sn : window;
    if (n == 0) {
      __syncwarp();
      if (lane == 0) {
        asm volatile("st.release.cta.shared.u32 [%0], %1;" ::"r"(drained_s + 4 * slot), "r"(round + 1u) : "memory");
        mbar_arrive(empty_s + 8 * slot);
        if (out_min) out_min[o] = 0.f;
        if (out_max) out_max[o] = 0.f;
        if (out_mean) out_mean[o] = 0.0;
        if (out_p95) out_p95[o] = 0.f;
      }
      continue;
    }
    RowView r;
    r.f = sf;
    r.v4 = reinterpret_cast<const float4*>(sf) + lane;
    r.lane = lane;
    r.n = n;
    r.nfull = n >> 2;
    r.jr = r.nfull >> 5;
    const int k = n - p95_rank(n);

    // ---- 1. pivot from 128 samples
    enum { kList, kDone, kBisect };
    int state = kList;
    float result = 0.f;
    float lox = POS_INF;          // list predicate x > lox (+inf: list nothing)
    uint32_t t1 = 0, t2 = 0, t3 = 0;
    uint32_t bis_hi = 0;          // exclusive upper key for the bisection fallback (0: max + 1)
    const bool sampled = want_p95 && r.jr >= kPipeMinRowsSampled;
    bool tie_mode = false;
    if (sampled) {
      uint32_t h[4];
#pragma unroll
      for (int g = 0; g < 4; ++g) h[g] = f32_key(sf[((lane * 4 + g) * n) >> 7]);
      const uint32_t lane_bit = 1u << lane;
#pragma unroll
      for (int rr = 1; rr <= 3; ++rr) {
        uint32_t m[4];
#pragma unroll
        for (int g = 0; g < 4; ++g) m[g] = warp_max_u32(h[g]);
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const unsigned who = __ballot_sync(0xffffffffu, h[g] == m[g]);
          if ((who & (0u - who)) == lane_bit) h[g] = 0u;  // retire one copy
        }
        const uint32_t mm = min(min(m[0], m[1]), min(m[2], m[3]));
        if (rr == 1) t1 = mm; else if (rr == 2) t2 = mm; else t3 = mm;
      }
      const float V = key_f32(t2);
      if (t1 == t2 || t2 == t3) {
        // ties around the pivot: place k against the block of copies of V
        tie_mode = true;
        const uint32_t cc = pipe_count2(r, V, next_up(V));
        const int c_ge = (int)(cc & 0xffffu), c_gt = (int)(cc >> 16);
        if (c_gt >= k) lox = V;                                   // the answer is above V
        else if (c_ge >= k) { result = V; state = kDone; }        // the answer is V
        else if (t3 < t2) lox = next_down(key_f32(t3));           // below V: list from the next sample down
        else { state = kBisect; bis_hi = t2; }
      } else {
        lox = next_down(V);  // x > lox  <=>  x >= V
      }
    } else if (want_p95) {
      lox = NEG_INF;  // short window: list every sample
    }

    // ---- 2. the pass: statistics + candidate list
    float mn, mx;
    double sum;
    uint32_t c128 = 0;
    pipe_pass<true>(r, state == kList ? lox : POS_INF, lb, c128, mn, mx, sum);
    int cnt = (int)(c128 >> 7);
    int total = 0, maxcnt = 0;
    if (want_p95 && state == kList) {
      total = warp_sum_i32(cnt);
      maxcnt = warp_max_i32(cnt);
      if (!sampled) {
        if (total < k) { result = NEG_INF; state = kDone; }  // only -inf samples are unlisted
      } else if (total < k || maxcnt > kPipeLaneCap) {
        // one retry with a neighbouring sample statistic, then the bisection fallback
        float lox2 = POS_INF;
        if (!tie_mode && maxcnt > kPipeLaneCap && total >= k && t1 > t2) lox2 = next_down(key_f32(t1));
        else if (!tie_mode && total < k && t3 < t2) lox2 = next_down(key_f32(t3));
        if (lox2 != POS_INF) {
#pragma unroll
          for (int e = 0; e < kListPrefill; ++e) lst[e * 32] = NEG_INF;
          lox = lox2;
          c128 = 0;
          float d0, d1;
          double d2;
          pipe_pass<false>(r, lox, lb, c128, d0, d1, d2);
          cnt = (int)(c128 >> 7);
          total = warp_sum_i32(cnt);
          maxcnt = warp_max_i32(cnt);
        }
        if (total < k || maxcnt > kPipeLaneCap) state = kBisect;
      }
    }
    // ---- slot no longer needed: hand it back to the producer
    __syncwarp();
    if (lane == 0) {
      asm volatile("st.release.cta.shared.u32 [%0], %1;" ::"r"(drained_s + 4 * slot), "r"(round + 1u) : "memory");
      mbar_arrive(empty_s + 8 * slot);
    }

#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, off));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
      sum += __shfl_xor_sync(0xffffffffu, sum, off);
    }
    if (lane == 0) {
      if (out_min) out_min[o] = mn + 0.f;  // -0.0 -> +0.0
      if (out_max) out_max[o] = mx + 0.f;
      if (out_mean) out_mean[o] = sum / (double)n;
    }
    if (!want_p95) continue;

    // ---- 3. p95 from the list
    if (state == kList) {
      if (mn == mx) {
        result = mx;  // constant series
      } else {
        // t1 (about 6 % of the window above it) is the hint for the first counting pass
        const float hint = sampled ? key_f32(t1) : 0.f;
        result = select_from_lists(lst, fin, lane, cnt, total, maxcnt, k,
                                   sampled ? next_up(lox) : mn, mx, sampled && !tie_mode, hint, hint,
                                   kListPrefill);
      }
    } else if (state == kBisect) {
      result = pipe_bisect_global(ring + s * (int64_t)pitch, n, lane, f32_key(NEG_INF),
                                  bis_hi ? bis_hi : f32_key(mx) + 1u, k);
    }
    if (lane == 0) out_p95[o] = result + 0.f;
  }
}

// Returns 1 if this variant cannot take the call (window too large for the slot ring).
int launch_window_stats_pipe(const float* ring, const uint32_t* seen, int32_t window, int32_t pitch,
                             int64_t s0, int64_t s1, float* out_min, float* out_max,
                             double* out_mean, float* out_p95, cudaStream_t st) {
  if (window > 4096) return 1;
  const int slot_bytes = ((pitch * 4) + 127) & ~127;
  int dev = 0, smem_max = 0, n_sm = 0;
  KLM_CUDA(cudaGetDevice(&dev));
  static int cached_smem[64] = {}, cached_sm[64] = {};
  if (dev < 64 && cached_smem[dev]) {
    smem_max = cached_smem[dev];
    n_sm = cached_sm[dev];
  } else {
    KLM_CUDA(cudaDeviceGetAttribute(&smem_max, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    KLM_CUDA(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
    KLM_CUDA(cudaFuncSetAttribute(window_stats_pipe_kernel,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max));
    if (dev < 64) {
      cached_smem[dev] = smem_max;
      cached_sm[dev] = n_sm;
    }
  }
  int n_slots = (smem_max - kPipeFixedBytes) / slot_bytes;
  if (n_slots > kPipeMaxSlots) n_slots = kPipeMaxSlots;
  if (n_slots < 4) return 1;
  const size_t smem = (size_t)kPipeFixedBytes + (size_t)n_slots * slot_bytes;
  const int64_t n = s1 - s0;
  const int grid = (int)(n < n_sm ? n : n_sm);
  window_stats_pipe_kernel<<<grid, kPipeThreads, smem, st>>>(ring, seen, window, pitch, s0, n,
                                                             out_min, out_max, out_mean, out_p95,
                                                             n_slots, slot_bytes);
  KLM_LAUNCH_CHECK("window_stats_pipe_kernel");
  return KLM_OK;
}

#ifdef KLM_PIPE_DEBUG
extern "C" int klm_pipe_debug_read(int* out) {
  return (int)cudaMemcpyFromSymbol(out, g_pipe_dbg, sizeof(int) * 8);
}
#endif

}  // namespace klm
