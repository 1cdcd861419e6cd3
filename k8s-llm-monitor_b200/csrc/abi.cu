// Host-buffer half of the C ABI (include/klm.h): a handle that owns device
// memory and a stream, copies caller arrays in, launches the klm_dev_* kernels
// and copies results out before returning. No caller pointer is retained.
//
// Callers in the reference (Sabre94/k8s-llm-monitor) this layer is shaped for:
//   internal/scheduler/controller.go:88-110   reconcile: one UAV snapshot, then
//                                             every pending request (klm_nodes_set
//                                             + klm_schedule_batch)
//   internal/metrics/manager.go:195-319       Collect -> calculateClusterMetrics
//   internal/metrics/manager.go:391-449       UpdateUAVReport (ring ingest)
// There is no CPU fallback: every entry point needs a live CUDA device.
#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <new>
#include <vector>

#include "common.cuh"

namespace klm {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
  set_error("CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
  // reset the runtime's per-thread last error: a recoverable failure (an allocation that
  // did not fit) must not be reported again by the next launch check
  (void)cudaGetLastError();
  return KLM_ECUDA;
}

int launch_ring_load(float* ring, uint32_t* seen, int32_t pitch, int64_t s0, int64_t n_rows,
                     int32_t count, const float* data, cudaStream_t st);

// Grow-only device buffer.
struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  int ensure(size_t bytes) {
    if (bytes <= cap) return KLM_OK;
    size_t want = bytes + bytes / 2;
    want = (want + 255) & ~size_t(255);
    if (p) {
      KLM_CUDA(cudaFree(p));
      p = nullptr;
      cap = 0;
    }
    KLM_CUDA(cudaMalloc(&p, want));
    cap = want;
    return KLM_OK;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
  template <typename T>
  T* as() const { return static_cast<T*>(p); }
};

// Grow-only page-locked host buffer: staging for the small-copy paths below. A call that
// moves only a few KB per array pays ~5 us of copy-engine latency per cudaMemcpyAsync, so
// such arrays are gathered here with memcpy and cross PCIe as one transfer each way.
struct PinBuf {
  void* p = nullptr;
  size_t cap = 0;
  int ensure(size_t bytes) {
    if (bytes <= cap) return KLM_OK;
    release();
    KLM_CUDA(cudaHostAlloc(&p, bytes, cudaHostAllocDefault));
    cap = bytes;
    return KLM_OK;
  }
  void release() {
    if (p) cudaFreeHost(p);
    p = nullptr;
    cap = 0;
  }
  char* at(size_t off) const { return static_cast<char*>(p) + off; }
};

// Arrays up to this many bytes in total go through the staging block; above it the
// memcpy would cost more than the extra copy submissions and the arrays are sent directly.
constexpr size_t kStageMax = 256 << 10;
// klm_collect_begin stages every snapshot (it must return without waiting): its own limit
constexpr size_t kAsyncSnapMax = 1 << 20;

// klm_schedule_batch pipelines batches of at least 2 * kChunkRequests requests over two
// streams, in at most kMaxChunks chunks of at least kChunkRequests requests.
constexpr int64_t kChunkRequests = 128 << 10;
constexpr int64_t kMaxChunks = 8;

}  // namespace klm

using namespace klm;

struct klm_handle {
  int device = 0;
  cudaStream_t stream = nullptr;
  cudaStream_t stream2 = nullptr;  // klm_collect: node-snapshot branch beside the ring branch
  std::mutex mu;

  // staging (host, page-locked) and the events that say the last transfer out of each
  // block has been read, i.e. the block may be overwritten
  PinBuf sched_pin, snap_pin;
  cudaEvent_t sched_read = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_chain[2] = {nullptr, nullptr};  // chunked klm_schedule_batch

  // scheduling
  int64_t n_nodes = -1;  // -1: no snapshot installed yet
  DevBuf node_tab;       // battery f64[n] | flags u8[n] (256-byte aligned parts)
  double* d_battery = nullptr;
  uint8_t* d_flags = nullptr;
  DevBuf req_blk, res_blk;  // small batches: min | off | ids and score | idx as single blocks
  DevBuf rec_blk;           // klm_reconcile, small case: node table | min | off | ids
  DevBuf best, best_scratch;
  DevBuf req_min, req_off, req_ids, res_idx, res_score;
  DevBuf snap_out;  // klm_collect, small snapshots: cluster_out | the five rate columns
  DevBuf topk_top, topk_scratch;  // klm_schedule_batch_topk: the global top-k list; its scratch

  // aggregation
  DevBuf agg_cols, agg_pods, agg_gpu, agg_out, agg_scratch;
  DevBuf rate_in, rate_out, pod_tot;
  DevBuf fleet_in, fleet_out, fleet_scratch;

  // ring
  DevBuf ring, seen, ing_ids, ing_vals, stats_out;
  DevBuf win_state;  // per series, the bracket K7 carries from one call to the next (2 floats; NaN: none)
  int64_t n_series = 0;
  int32_t window = 0, pitch = 0;

  // klm_collect_begin / klm_collect_end: two cycles in flight. Uploads and the snapshot branch
  // run on stream2, K6 + K7 on stream, the result download on stream3; what two overlapping
  // cycles could both touch exists once per slot.
  struct CollectSlot {
    DevBuf tick_dev, stats;   // this cycle's samples; K7's packed results
    PinBuf pin;               // tick | snapshot columns | small results
    cudaEvent_t up = nullptr, k7 = nullptr, snap = nullptr, win = nullptr;
    klm_collect_out out{};    // where klm_collect_end delivers the small results
    int64_t n = 0;
    size_t col8 = 0, out_at = 0;
    bool rates = false, snapshot = false;
  } slot[2];
  cudaStream_t stream3 = nullptr, stream_bulk = nullptr;  // downloads; K6 + K7 of the async cycles
  uint64_t begun = 0, ended = 0;
  // Samples pushed (klm_ring_ingest / klm_ring_ingest_tick) while cycles are in flight: kept in
  // arrival order and applied by the next klm_collect_begin before its own tick, or by the next
  // synchronous ring call once the pipeline is idle. One entry per call (ids of one call are
  // distinct; calls may repeat a series, so every call is its own K6 launch, in order).
  struct Pushed {
    int64_t n;         // samples of the call
    size_t ids_at;     // first id in q_ids, or SIZE_MAX for a whole tick (n == n_series)
    size_t vals_at;    // first value in q_vals
  };
  std::vector<Pushed> q_calls;
  std::vector<int64_t> q_ids;
  std::vector<float> q_vals;
  PinBuf q_pin[2];     // staging of the queue, one per cycle slot
  DevBuf q_dev[2];

  // ---- more than one device (klm_create with n_dev > 1): the scheduler's batch is split
  // over the devices inside the library; `device` above is devs[0].device.
  struct Dev {
    int device = 0;
    cudaStream_t stream = nullptr, stream2 = nullptr;  // the chunks of a large slice alternate between them
    cudaEvent_t ev_k1 = nullptr, ev_chain[2] = {nullptr, nullptr};
    DevBuf tab;       // node columns this device holds: the whole table (request-shard mode)
                      // or its own shard (node-shard mode); battery | flags
    DevBuf slots, scratch;  // exchange slots peers write into; K1 scratch
    DevBuf req_min, req_off, req_ids, res_idx, res_score;
    double* d_battery = nullptr;  // base of the columns in `tab`
    uint8_t* d_flags = nullptr;
  };
  std::vector<Dev> devs;
  PinBuf md_tab;             // the snapshot of klm_nodes_set (page-locked copy: the layout is
                             // only known when the batch arrives)
  int64_t md_nodes = -1;     // -1: no snapshot
  bool md_tab_valid = false; // md_tab holds the current snapshot (klm_nodes_set, not klm_reconcile)
  int md_uploaded_mode = -1; // which layout the devices hold for this snapshot (-1: none)
  bool md_spread = false;    // node-shard layout with replicas filled over NVLink by K1 (else: peer probes)
  uint64_t md_epoch = 0;     // exchange epoch of the last K1
};

namespace {

struct Guard {
  std::lock_guard<std::mutex> lk;
  int rc;
  explicit Guard(klm_handle* h) : lk(h->mu), rc(KLM_OK) {
    cudaError_t e = cudaSetDevice(h->device);
    if (e != cudaSuccess) rc = cuda_fail(e, "cudaSetDevice");
  }
};

#define KLM_ENTER(h)                       \
  if (!(h)) {                              \
    set_error("null handle");              \
    return KLM_EINVAL;                     \
  }                                        \
  Guard klm_guard_(h);                     \
  if (klm_guard_.rc != KLM_OK) return klm_guard_.rc

#define KLM_TRY(expr)              \
  do {                             \
    int klm_rc_ = (expr);          \
    if (klm_rc_ != KLM_OK) return klm_rc_; \
  } while (0)

int h2d(void* dst, const void* src, size_t bytes, cudaStream_t st) {
  if (bytes == 0) return KLM_OK;
  KLM_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, st));
  return KLM_OK;
}
int d2h(void* dst, const void* src, size_t bytes, cudaStream_t st) {
  if (bytes == 0) return KLM_OK;
  KLM_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, st));
  return KLM_OK;
}
int sync(cudaStream_t st) {
  KLM_CUDA(cudaStreamSynchronize(st));
  return KLM_OK;
}

size_t align256(size_t b) { return (b + 255) & ~size_t(255); }

// CSR offsets off[0..n] must not decrease (with off[0] >= 0 and off[n] <= total checked by the
// caller, every inner offset is then in range too): a kernel indexes pref_ids / container
// columns with them. One pass over memory the call is about to copy anyway.
bool csr_non_decreasing(const int32_t* off, int64_t n) {
  int bad = 0;
  for (int64_t i = 0; i < n; ++i) bad |= off[i + 1] < off[i];
  return !bad;
}

// K7 over series [s0, s1) of the handle's ring. Rings whose windows the streaming kernel takes
// (<= 8192 samples) run it with the per-series brackets of win_state: between two collection
// ticks a window changes by one sample, so the previous bracket stands in for the pivot sampling
// (klm.h, klm_dev_window_stats_hinted; exact whatever the bracket holds). p95 not wanted, or
// longer windows: the stateless kernels.
int ring_window_stats(klm_handle* h, int64_t s0, int64_t s1, float* d_min, float* d_max, double* d_mean,
                      float* d_p95, cudaStream_t st) {
  if (d_p95 != nullptr && h->window <= 8192 && h->win_state.p != nullptr)
    return klm_dev_window_stats_hinted(h->ring.as<float>(), h->seen.as<uint32_t>(), h->window, h->pitch, s0, s1,
                                       d_min, d_max, d_mean, d_p95, h->win_state.as<float>() + 2 * s0, st);
  return klm_dev_window_stats(h->ring.as<float>(), h->seen.as<uint32_t>(), h->window, h->pitch, s0, s1, d_min,
                              d_max, d_mean, d_p95, st);
}

// Apply the queued pushes in arrival order on `st` (staging block `which`). The host copy is
// consumed before this returns; the device work is ordered by the stream.
int flush_pushed(klm_handle* h, int which, cudaStream_t st) {
  if (h->q_calls.empty()) return KLM_OK;
  const size_t ids_bytes = align256(h->q_ids.size() * sizeof(int64_t)), vals_bytes = h->q_vals.size() * sizeof(float);
  PinBuf& pin = h->q_pin[which];
  DevBuf& dev = h->q_dev[which];
  KLM_TRY(pin.ensure(ids_bytes + vals_bytes + 256));
  KLM_TRY(dev.ensure(ids_bytes + vals_bytes + 256));
  if (!h->q_ids.empty()) std::memcpy(pin.at(0), h->q_ids.data(), h->q_ids.size() * sizeof(int64_t));
  std::memcpy(pin.at(ids_bytes), h->q_vals.data(), vals_bytes);
  KLM_TRY(h2d(dev.p, pin.p, ids_bytes + vals_bytes, st));
  const int64_t* d_ids = dev.as<int64_t>();
  const float* d_vals = reinterpret_cast<const float*>(dev.as<char>() + ids_bytes);
  for (const klm_handle::Pushed& c : h->q_calls)
    KLM_TRY(klm_dev_ring_ingest(h->ring.as<float>(), h->seen.as<uint32_t>(), h->n_series, h->window, h->pitch,
                                c.n, c.ids_at == SIZE_MAX ? nullptr : d_ids + c.ids_at, d_vals + c.vals_at, st));
  h->q_calls.clear();
  h->q_ids.clear();
  h->q_vals.clear();
  return KLM_OK;
}

// the synchronous ring / collect / aggregate calls share buffers and streams with the cycles
// klm_collect_begin has in flight
int collect_idle(klm_handle* h, const char* who) {
  if (h->begun != h->ended) {
    set_error("%s: klm_collect_begin cycles are in flight; finish them with klm_collect_end first", who);
    return KLM_ESTATE;
  }
  return KLM_OK;
}

}  // namespace


// ------------------------------------------------- scheduling over several devices
//
// One process, one handle, G devices (what the Go scheduler is: cmd/scheduler/main.go:55-62
// constructs one controller). Both modes split the REQUESTS over the devices; they differ in
// where the node table lives:
//   request-shard (KLM_SHARD_REQUESTS): every device holds the whole table and probes locally;
//   node-shard (KLM_SHARD_NODES): device g holds nodes [g L, (g + 1) L), L = ceil(N / G) rounded
//     up to 16, and the others read a preferred node that lives there through peer pointers.
// In both, device g runs K1 over its N / G range only and publishes the 16-byte result into every
// device's exchange slots; K2 folds the G results. No all-gather, no merge kernel, no host
// barrier between K1 and K2 (schedule.cu, "node-range sharding").
namespace {

int multi_init(klm_handle* h, const int* ids, int n) {
  h->devs.resize((size_t)n);
  for (int i = 0; i < n; ++i) {
    cudaDeviceProp prop;
    KLM_CUDA(cudaGetDeviceProperties(&prop, ids[i]));
    if (prop.major != 10) {
      set_error("klm_create: device %d is sm_%d%d; libklm is built for sm_100a (B200) only", ids[i],
                prop.major, prop.minor);
      return KLM_ENODEV;
    }
    h->devs[(size_t)i].device = ids[i];
  }
  for (int i = 0; i < n; ++i) {
    klm_handle::Dev& d = h->devs[(size_t)i];
    KLM_CUDA(cudaSetDevice(d.device));
    for (int j = 0; j < n; ++j) {
      if (j == i) continue;
      int can = 0;
      KLM_CUDA(cudaDeviceCanAccessPeer(&can, d.device, ids[j]));
      if (!can) {
        set_error("klm_create: device %d cannot map device %d's memory (no peer access)", d.device, ids[j]);
        return KLM_ENODEV;
      }
      const cudaError_t e = cudaDeviceEnablePeerAccess(ids[j], 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return cuda_fail(e, "cudaDeviceEnablePeerAccess");
      (void)cudaGetLastError();
    }
    int lo = 0, hi = 0;
    KLM_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    KLM_CUDA(cudaStreamCreateWithPriority(&d.stream, cudaStreamNonBlocking, hi));
    KLM_CUDA(cudaStreamCreateWithPriority(&d.stream2, cudaStreamNonBlocking, hi));
    KLM_CUDA(cudaEventCreateWithFlags(&d.ev_k1, cudaEventDisableTiming));
    KLM_CUDA(cudaEventCreateWithFlags(&d.ev_chain[0], cudaEventDisableTiming));
    KLM_CUDA(cudaEventCreateWithFlags(&d.ev_chain[1], cudaEventDisableTiming));
    KLM_TRY(d.slots.ensure((size_t)klm_dev_shard_slots_bytes()));
    KLM_TRY(d.scratch.ensure((size_t)klm_dev_node_best_scratch_bytes()));
    KLM_CUDA(cudaMemsetAsync(d.slots.p, 0, d.slots.cap, d.stream));
    KLM_CUDA(cudaMemsetAsync(d.scratch.p, 0, d.scratch.cap, d.stream));
    KLM_CUDA(cudaStreamSynchronize(d.stream));
  }
  KLM_CUDA(cudaSetDevice(h->device));
  return KLM_OK;
}

// Make every device hold the snapshot, then K1 + publication. Layouts:
//   request-shard: the whole table goes to every device over its own PCIe link;
//   node-shard:    device g receives only its shard from the host. With `spread` K1 then pushes
//                  the shard into every peer's replica over NVLink while it scans it, and K2 probes
//                  locally; without, K2 probes the peers' shards in place (fewer bytes when the
//                  batch is small).
int multi_upload(klm_handle* h, bool node_shard, bool spread, int64_t n, const double* battery,
                 const uint8_t* flags) {
  const int G = (int)h->devs.size();
  ++h->md_epoch;
  const bool whole = !node_shard || spread;  // every device has room for the whole table
  std::vector<void*> peer_slots((size_t)G);
  std::vector<double*> rep_bat((size_t)G);
  std::vector<uint8_t*> rep_flg((size_t)G);
  for (int g = 0; g < G; ++g) {
    klm_handle::Dev& d = h->devs[(size_t)g];
    KLM_CUDA(cudaSetDevice(d.device));
    int64_t lo = 0, hi = 0;
    klm_shard_range(n, G, g, &lo, &hi);
    const int64_t cnt = whole ? n : hi - lo;
    const size_t flag_off = align256((size_t)cnt * sizeof(double));
    KLM_TRY(d.tab.ensure(flag_off + (size_t)cnt + 256));
    d.d_battery = d.tab.as<double>();
    d.d_flags = d.tab.as<uint8_t>() + flag_off;
    peer_slots[(size_t)g] = d.slots.p;
    rep_bat[(size_t)g] = d.d_battery;
    rep_flg[(size_t)g] = d.d_flags;
  }
  for (int g = 0; g < G; ++g) {
    klm_handle::Dev& d = h->devs[(size_t)g];
    KLM_CUDA(cudaSetDevice(d.device));
    int64_t lo = 0, hi = 0;
    klm_shard_range(n, G, g, &lo, &hi);
    const int64_t a = node_shard ? lo : 0, b = node_shard ? hi : n;  // what the host sends here
    const int64_t at = whole ? a : 0;                               // where it lands in `tab`
    KLM_TRY(h2d(d.d_battery + at, battery + a, (size_t)(b - a) * sizeof(double), d.stream));
    KLM_TRY(h2d(d.d_flags + at, flags + a, (size_t)(b - a), d.stream));
    // K1 over this device's RANGE in every layout; global indices through node_base
    const int64_t k1_at = whole ? lo : 0;
    if (node_shard && spread) {
      KLM_TRY(klm_dev_node_best_spread(d.d_battery + k1_at, d.d_flags + k1_at, hi - lo, lo, G, g, peer_slots.data(),
                                       rep_bat.data(), rep_flg.data(), h->md_epoch, nullptr, d.scratch.p, d.stream));
    } else {
      KLM_TRY(klm_dev_node_best_publish(d.d_battery + k1_at, d.d_flags + k1_at, hi - lo, lo, G, g,
                                        peer_slots.data(), h->md_epoch, nullptr, d.scratch.p, d.stream));
    }
    KLM_CUDA(cudaEventRecord(d.ev_k1, d.stream));
    KLM_CUDA(cudaStreamWaitEvent(d.stream2, d.ev_k1, 0));  // K2 chunks on stream2 start behind this K1
  }
  h->md_uploaded_mode = node_shard ? 1 : 0;
  h->md_spread = node_shard && spread;
  return KLM_OK;
}

int multi_sync(klm_handle* h) {
  int rc = KLM_OK;
  for (auto& d : h->devs) {
    cudaSetDevice(d.device);
    for (cudaStream_t st : {d.stream, d.stream2}) {
      const cudaError_t e = cudaStreamSynchronize(st);
      if (e != cudaSuccess && rc == KLM_OK) rc = cuda_fail(e, "cudaStreamSynchronize");
    }
  }
  cudaSetDevice(h->device);
  return rc;
}

int multi_schedule(klm_handle* h, int64_t n, const double* battery, const uint8_t* flags, int64_t R,
                   const double* min_battery, const int32_t* pref_off, const int32_t* pref_ids,
                   int32_t* out_idx, double* out_score, int algo, bool node_shard) {
  const int G = (int)h->devs.size();
  if (algo == KLM_ALGO_BRUTEFORCE && node_shard) {
    set_error("klm_schedule_batch: the brute-force verifier scans whole tables; use it with KLM_SHARD_REQUESTS");
    return KLM_EINVAL;
  }
  if (R < 0 || (algo != KLM_ALGO_DECOMPOSED && algo != KLM_ALGO_BRUTEFORCE)) {
    set_error("klm_schedule_batch: bad argument");
    return KLM_EINVAL;
  }
  if (R > 0 && (!min_battery || !pref_off || !out_idx || !out_score)) {
    set_error("klm_schedule_batch: null array");
    return KLM_EINVAL;
  }
  const int64_t n_ids = R > 0 ? pref_off[R] : 0;
  if (R > 0 && (pref_off[0] != 0 || n_ids < 0 || (n_ids > 0 && !pref_ids))) {
    set_error("klm_schedule_batch: pref_off must start at 0 and pref_ids be non-null");
    return KLM_EINVAL;
  }
  for (int g = 1; g <= G && R > 0; ++g) {  // slice boundaries, before anything is in flight
    const int64_t r = R * g / G, rp = R * (g - 1) / G;
    if (pref_off[r] < pref_off[rp] || pref_off[r] > n_ids) {
      set_error("klm_schedule_batch: pref_off is not non-decreasing within [0, pref_off[R]]");
      return KLM_EINVAL;
    }
  }
  if (h->md_uploaded_mode != (node_shard ? 1 : 0)) {
    // node-shard: replicate over NVLink (K1 spreads its shard) once the batch probes more preferred
    // nodes than about 0.08 G N -- a remote probe is two 32-byte packets at ~11 G/s, a bulk
    // copy runs at several hundred GB/s (measured); a small batch probes in place
    const bool spread = node_shard && n_ids * 12 > n * (int64_t)G;
    KLM_TRY(multi_upload(h, node_shard, spread, n, battery, flags));
  }
  if (R == 0) return multi_sync(h);
  const bool local_table = !node_shard || h->md_spread;  // K2 probes this device's own memory
  std::vector<const double*> bat_parts((size_t)G);
  std::vector<const uint8_t*> flg_parts((size_t)G);
  // Every device's slice is cut into chunks that alternate between its two streams, so that the
  // download of one chunk overlaps the upload of the next (PCIe is full duplex and is what bounds
  // a large batch); the chunk loop is the outer one, so all devices' copy engines fill together.
  int64_t n_chunks = (R / G) / kChunkRequests;
  if (n_chunks > kMaxChunks) n_chunks = kMaxChunks;
  if (n_chunks < 1 || algo == KLM_ALGO_BRUTEFORCE) n_chunks = 1;
  for (int g = 0; g < G; ++g) {
    klm_handle::Dev& d = h->devs[(size_t)g];
    KLM_CUDA(cudaSetDevice(d.device));
    const int64_t r0 = R * g / G, r1 = R * (g + 1) / G, rl = r1 - r0;
    if (rl == 0) continue;
    const int64_t i0 = pref_off[r0], i1 = pref_off[r1];
    KLM_TRY(d.req_min.ensure((size_t)rl * 8));
    KLM_TRY(d.req_off.ensure((size_t)(rl + 1) * 4));
    KLM_TRY(d.req_ids.ensure((size_t)(i1 - i0 > 0 ? i1 - i0 : 1) * 4));
    KLM_TRY(d.res_idx.ensure((size_t)rl * 4));
    KLM_TRY(d.res_score.ensure((size_t)rl * 8));
  }
  for (int64_t c = 0; c < n_chunks; ++c) {
    for (int g = 0; g < G; ++g) {
      klm_handle::Dev& d = h->devs[(size_t)g];
      const int64_t r0 = R * g / G, r1 = R * (g + 1) / G, rl = r1 - r0;
      if (rl == 0) continue;
      KLM_CUDA(cudaSetDevice(d.device));
      const int64_t q0 = r0 + rl * c / n_chunks, q1 = r0 + rl * (c + 1) / n_chunks;  // this chunk's requests
      if (q1 == q0) continue;
      if (!csr_non_decreasing(pref_off + q0, q1 - q0)) {
        set_error("klm_schedule_batch: pref_off is not non-decreasing");
        return KLM_EINVAL;
      }
      const int64_t i0 = pref_off[r0], j0 = pref_off[q0], j1 = pref_off[q1];
      cudaStream_t st = (c & 1) ? d.stream2 : d.stream;
      // offsets (q0, q1]: entry q0 came with the previous chunk (the slice's first with chunk 0)
      const int64_t o0 = c == 0 ? q0 : q0 + 1;
      KLM_TRY(h2d(d.req_min.as<double>() + (q0 - r0), min_battery + q0, (size_t)(q1 - q0) * 8, st));
      KLM_TRY(h2d(d.req_off.as<int32_t>() + (o0 - r0), pref_off + o0, (size_t)(q1 + 1 - o0) * 4, st));
      KLM_TRY(h2d(d.req_ids.as<int32_t>() + (j0 - i0), pref_ids + j0, (size_t)(j1 - j0) * 4, st));
      KLM_CUDA(cudaEventRecord(d.ev_chain[c & 1], st));
      // ... which is why this chunk's kernel waits for the previous chunk's upload
      if (c > 0) KLM_CUDA(cudaStreamWaitEvent(st, d.ev_chain[(c - 1) & 1], 0));
      // the offsets stay absolute: hand the kernel the ids pointer moved back by i0 entries
      const int32_t* ids_abs = d.req_ids.as<int32_t>() - i0;
      const double* c_min = d.req_min.as<double>() + (q0 - r0);
      const int32_t* c_off = d.req_off.as<int32_t>() + (q0 - r0);
      int32_t* c_idx = d.res_idx.as<int32_t>() + (q0 - r0);
      double* c_score = d.res_score.as<double>() + (q0 - r0);
      if (algo == KLM_ALGO_BRUTEFORCE) {
        KLM_TRY(klm_dev_schedule_bruteforce(d.d_battery, d.d_flags, n, 0, q1 - q0, c_min, c_off, ids_abs, c_idx,
                                            c_score, st));
      } else {
        // the table as this device sees it: its own columns, or the peers' shards
        for (int q = 0; q < G; ++q) {
          int64_t lo = 0, hi = 0;
          klm_shard_range(n, G, q, &lo, &hi);
          const klm_handle::Dev& owner = h->devs[(size_t)(local_table ? g : q)];
          bat_parts[(size_t)q] = owner.d_battery + (local_table ? lo : 0);
          flg_parts[(size_t)q] = owner.d_flags + (local_table ? lo : 0);
        }
        KLM_TRY(klm_dev_schedule_sharded(G, n, bat_parts.data(), flg_parts.data(), d.slots.p, h->md_epoch, q1 - q0,
                                         c_min, c_off, ids_abs, c_idx, c_score, h->md_spread ? 1 : 0, st));
      }
      KLM_TRY(d2h(out_idx + q0, c_idx, (size_t)(q1 - q0) * 4, st));
      KLM_TRY(d2h(out_score + q0, c_score, (size_t)(q1 - q0) * 8, st));
    }
  }
  KLM_TRY(multi_sync(h));
  for (auto& d : h->devs) {  // a K2 that waited in vain for a peer's K1 says so
    int err = 0;
    KLM_CUDA(cudaSetDevice(d.device));
    KLM_TRY(klm_dev_shard_error(d.slots.p, &err, d.stream));
    if (err) {
      KLM_CUDA(cudaMemsetAsync(static_cast<char*>(d.slots.p) + klm_dev_shard_slots_bytes() - 64, 0, 4, d.stream));
      cudaSetDevice(h->device);
      set_error("klm_schedule_batch: device %d timed out waiting for a peer's node_best result", d.device);
      return KLM_ESTATE;
    }
  }
  KLM_CUDA(cudaSetDevice(h->device));
  return KLM_OK;
}

// `mode` = KLM_ALGO_* | KLM_SHARD_*; anything else is rejected
bool split_mode(int mode, int* algo, bool* node_shard) {
  *algo = mode & 0xff;
  *node_shard = (mode & KLM_SHARD_NODES) != 0;
  return (mode & ~(0xff | KLM_SHARD_NODES)) == 0;
}

// several devices: keep a page-locked copy; the devices are loaded when the first batch says
// in which layout (the caller's arrays may be gone by then)
int multi_nodes_set(klm_handle* h, int64_t n, const double* battery, const uint8_t* flags) {
  if (n < 0 || n > 0x7fffffff || (n > 0 && (!battery || !flags))) {
    set_error("klm_nodes_set: bad argument (0 <= n_nodes < 2^31, arrays non-null)");
    return KLM_EINVAL;
  }
  h->md_nodes = -1;
  h->md_uploaded_mode = -1;
  h->md_tab_valid = false;
  KLM_TRY(multi_sync(h));  // nothing may still read the previous copy
  const size_t bat_bytes = (size_t)n * sizeof(double);
  KLM_TRY(h->md_tab.ensure(align256(bat_bytes) + (size_t)n + 256));
  if (n > 0) {
    std::memcpy(h->md_tab.at(0), battery, bat_bytes);
    std::memcpy(h->md_tab.at(align256(bat_bytes)), flags, (size_t)n);
  }
  h->md_nodes = n;
  h->md_tab_valid = true;
  return KLM_OK;
}

}  // namespace

extern "C" const char* klm_last_error(void) { return g_err; }
extern "C" int klm_abi_version(void) { return 1004; }

extern "C" int klm_host_alloc(size_t bytes, void** out) {
  if (!out || bytes == 0) {
    set_error("klm_host_alloc: bad argument");
    return KLM_EINVAL;
  }
  *out = nullptr;
  // portable: usable with every handle of the process, whichever device it drives
  KLM_CUDA(cudaHostAlloc(out, bytes, cudaHostAllocPortable));
  return KLM_OK;
}

extern "C" int klm_host_free(void* p) {
  if (!p) return KLM_OK;
  KLM_CUDA(cudaFreeHost(p));
  return KLM_OK;
}

extern "C" int klm_create(const int* device_ids, int n_dev, klm_handle** out) {
  if (!out) {
    set_error("klm_create: out is null");
    return KLM_EINVAL;
  }
  *out = nullptr;
  if (n_dev < 0 || n_dev > 16 || (n_dev == 0 && device_ids) || (n_dev > 0 && !device_ids)) {
    set_error("klm_create: n_dev must be 0 (device 0) or 1..16 with device_ids");
    return n_dev > 16 ? KLM_ELIMIT : KLM_EINVAL;
  }
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0) {
    set_error("klm_create: no CUDA device (%s); libklm has no CPU fallback",
              e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
    return KLM_ENODEV;
  }
  const int dev = device_ids ? device_ids[0] : 0;
  if (dev < 0 || dev >= count) {
    set_error("klm_create: device %d out of range [0, %d)", dev, count);
    return KLM_EINVAL;
  }
  for (int i = 0; i < n_dev; ++i) {
    if (device_ids[i] < 0 || device_ids[i] >= count) {
      set_error("klm_create: device %d out of range [0, %d)", device_ids[i], count);
      return KLM_EINVAL;
    }
    for (int j = 0; j < i; ++j)
      if (device_ids[j] == device_ids[i]) {
        set_error("klm_create: device %d listed twice", device_ids[i]);
        return KLM_EINVAL;
      }
  }
  KLM_CUDA(cudaSetDevice(dev));
  cudaDeviceProp prop;
  KLM_CUDA(cudaGetDeviceProperties(&prop, dev));
  if (prop.major != 10) {
    set_error("klm_create: device %d is sm_%d%d; libklm is built for sm_100a (B200) only", dev,
              prop.major, prop.minor);
    return KLM_ENODEV;
  }
  klm_handle* h = new (std::nothrow) klm_handle();
  if (!h) {
    set_error("klm_create: out of host memory");
    return KLM_EINVAL;
  }
  h->device = dev;
  // The main stream carries the short synchronous calls (a reconcile is two tiny kernels): it
  // gets the highest priority so that its CTAs are placed ahead of another handle's pending
  // bulk work (the window statistics of klm_collect_begin run on a default-priority stream).
  int prio_least = 0, prio_greatest = 0;
  e = cudaDeviceGetStreamPriorityRange(&prio_least, &prio_greatest);
  if (e == cudaSuccess) e = cudaStreamCreateWithPriority(&h->stream, cudaStreamNonBlocking, prio_greatest);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->stream2, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->sched_read, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ev_chain[0], cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ev_chain[1], cudaEventDisableTiming);
  if (e != cudaSuccess) {
    klm_destroy(h);
    return cuda_fail(e, "cudaStreamCreateWithFlags / cudaEventCreateWithFlags");
  }
  // scratch areas carry self-resetting tickets that must start at zero
  int rc = h->best_scratch.ensure((size_t)klm_dev_node_best_scratch_bytes());
  if (rc == KLM_OK) rc = h->agg_scratch.ensure((size_t)klm_dev_cluster_scratch_bytes());
  if (rc == KLM_OK) rc = h->best.ensure(2 * sizeof(int64_t));
  if (rc == KLM_OK) rc = h->agg_out.ensure(sizeof(klm_cluster_out));
  if (rc == KLM_OK) {
    e = cudaMemsetAsync(h->best_scratch.p, 0, h->best_scratch.cap, h->stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(h->agg_scratch.p, 0, h->agg_scratch.cap, h->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    if (e != cudaSuccess) rc = cuda_fail(e, "scratch memset");
  }
  if (rc != KLM_OK) {
    klm_destroy(h);
    return rc;
  }
  if (n_dev > 1) {
    rc = multi_init(h, device_ids, n_dev);
    if (rc != KLM_OK) {
      klm_destroy(h);
      return rc;
    }
  }
  *out = h;
  return KLM_OK;
}

extern "C" void klm_destroy(klm_handle* h) {
  if (!h) return;
  for (auto& d : h->devs) {
    cudaSetDevice(d.device);
    if (d.stream) cudaStreamSynchronize(d.stream);
    if (d.stream2) cudaStreamSynchronize(d.stream2);
    for (cudaEvent_t ev : {d.ev_k1, d.ev_chain[0], d.ev_chain[1]})
      if (ev) cudaEventDestroy(ev);
    if (d.stream2) cudaStreamDestroy(d.stream2);
    for (DevBuf* b : {&d.tab, &d.slots, &d.scratch, &d.req_min, &d.req_off, &d.req_ids, &d.res_idx, &d.res_score})
      b->release();
    if (d.stream) cudaStreamDestroy(d.stream);
  }
  h->md_tab.release();
  cudaSetDevice(h->device);
  for (int i = 0; i < 2; ++i) {
    h->q_pin[i].release();
    h->q_dev[i].release();
  }
  if (h->stream) cudaStreamSynchronize(h->stream);
  if (h->stream2) cudaStreamSynchronize(h->stream2);
  DevBuf* bufs[] = {&h->node_tab, &h->req_blk,  &h->res_blk,    &h->snap_out,   &h->rec_blk,
                    &h->best,     &h->best_scratch, &h->req_min,
                    &h->req_off,  &h->req_ids,  &h->res_idx,    &h->res_score,    &h->agg_cols,
                    &h->agg_pods, &h->agg_gpu,  &h->agg_out,    &h->agg_scratch,  &h->rate_in,
                    &h->rate_out, &h->pod_tot, &h->fleet_in, &h->fleet_out,  &h->fleet_scratch, &h->ring,
                    &h->seen,     &h->ing_ids,  &h->ing_vals,   &h->stats_out, &h->win_state,
                    &h->topk_top, &h->topk_scratch};
  for (DevBuf* b : bufs) b->release();
  if (h->stream_bulk) cudaStreamSynchronize(h->stream_bulk);
  if (h->stream3) cudaStreamSynchronize(h->stream3);
  for (auto& sl : h->slot) {
    sl.tick_dev.release();
    sl.stats.release();
    sl.pin.release();
    for (cudaEvent_t ev : {sl.up, sl.k7, sl.snap, sl.win})
      if (ev) cudaEventDestroy(ev);
  }
  if (h->stream3) cudaStreamDestroy(h->stream3);
  if (h->stream_bulk) cudaStreamDestroy(h->stream_bulk);
  h->sched_pin.release();
  h->snap_pin.release();
  if (h->sched_read) cudaEventDestroy(h->sched_read);
  if (h->ev_fork) cudaEventDestroy(h->ev_fork);
  for (cudaEvent_t ev : h->ev_chain)
    if (ev) cudaEventDestroy(ev);
  if (h->stream2) cudaStreamDestroy(h->stream2);
  if (h->stream) cudaStreamDestroy(h->stream);
  delete h;
}

// ------------------------------------------------------------ scheduling

static int nodes_set_impl(klm_handle* h, int64_t n, const double* battery, const uint8_t* flags) {
  if (n < 0 || n > 0x7fffffff || (n > 0 && (!battery || !flags))) {
    set_error("klm_nodes_set: bad argument (0 <= n_nodes < 2^31, arrays non-null)");
    return KLM_EINVAL;
  }
  const size_t bat_bytes = (size_t)n * sizeof(double), flag_off = align256(bat_bytes);
  const size_t tab_bytes = flag_off + (size_t)n;
  h->n_nodes = -1;  // no snapshot until this one is in place
  KLM_TRY(h->node_tab.ensure(tab_bytes + 256));
  h->d_battery = h->node_tab.as<double>();
  h->d_flags = h->node_tab.as<uint8_t>() + flag_off;
  const bool staged = tab_bytes <= kStageMax;
  if (staged) {
    // Small table: gather both columns in the staging block and send one transfer. The
    // caller's arrays are consumed by the memcpy, so the call returns without waiting
    // for the GPU; a device-side failure is reported by the next call that synchronises.
    KLM_TRY(h->sched_pin.ensure(kStageMax));
    KLM_CUDA(cudaEventSynchronize(h->sched_read));  // previous transfer out of the block
    if (n > 0) {
      std::memcpy(h->sched_pin.at(0), battery, bat_bytes);
      std::memcpy(h->sched_pin.at(flag_off), flags, (size_t)n);
    }
    KLM_TRY(h2d(h->node_tab.p, h->sched_pin.p, tab_bytes, h->stream));
    KLM_CUDA(cudaEventRecord(h->sched_read, h->stream));
  } else {
    KLM_TRY(h2d(h->d_battery, battery, bat_bytes, h->stream));
    KLM_TRY(h2d(h->d_flags, flags, (size_t)n, h->stream));
  }
  // K1 runs once per snapshot, not once per batch
  KLM_TRY(klm_dev_node_best(h->d_battery, h->d_flags, n, 0, h->best.as<int64_t>(),
                            h->best_scratch.p, h->stream));
  if (!staged) KLM_TRY(sync(h->stream));  // caller may free its arrays after return
  h->n_nodes = n;
  return KLM_OK;
}

extern "C" int klm_nodes_set(klm_handle* h, int64_t n, const double* battery,
                             const uint8_t* flags) {
  KLM_ENTER(h);
  if (!h->devs.empty()) return multi_nodes_set(h, n, battery, flags);
  return nodes_set_impl(h, n, battery, flags);
}

static int schedule_batch_impl(klm_handle* h, int64_t R, const double* min_battery,
                               const int32_t* pref_off, const int32_t* pref_ids, int32_t* out_idx,
                               double* out_score, int algo) {
  if (h->n_nodes < 0) {
    set_error("klm_schedule_batch: no node snapshot; call klm_nodes_set first");
    return KLM_ESTATE;
  }
  if (R < 0 || (algo != KLM_ALGO_DECOMPOSED && algo != KLM_ALGO_BRUTEFORCE)) {
    set_error("klm_schedule_batch: bad argument");
    return KLM_EINVAL;
  }
  if (R == 0) return KLM_OK;
  if (!min_battery || !pref_off || !out_idx || !out_score) {
    set_error("klm_schedule_batch: null array");
    return KLM_EINVAL;
  }
  const int64_t n_ids = pref_off[R];
  if (pref_off[0] != 0 || n_ids < 0 || (n_ids > 0 && !pref_ids)) {
    set_error("klm_schedule_batch: pref_off must start at 0 and pref_ids be non-null");
    return KLM_EINVAL;
  }
  const size_t min_bytes = (size_t)R * sizeof(double), off_bytes = (size_t)(R + 1) * sizeof(int32_t);
  const size_t ids_bytes = (size_t)n_ids * sizeof(int32_t), idx_bytes = (size_t)R * sizeof(int32_t);
  // block layouts of the small path (each part 256-byte aligned)
  const size_t in_off = align256(min_bytes), in_ids = in_off + align256(off_bytes);
  const size_t in_bytes = in_ids + align256(ids_bytes), out_idx_at = align256(min_bytes);
  const size_t out_bytes = out_idx_at + idx_bytes;
  const bool staged = in_bytes + out_bytes <= kStageMax;
  double *d_min, *d_score;
  int32_t *d_off, *d_ids, *d_idx;
  if (staged) {
    // A reconcile with a handful of pending requests: three arrays in and two out would be
    // five copy submissions for a few hundred bytes; stage them and move one block each way.
    KLM_TRY(h->req_blk.ensure(in_bytes + 256));
    KLM_TRY(h->res_blk.ensure(out_bytes + 256));
    KLM_TRY(h->sched_pin.ensure(kStageMax));
    KLM_CUDA(cudaEventSynchronize(h->sched_read));
    char* blk = h->req_blk.as<char>();
    d_min = reinterpret_cast<double*>(blk);
    d_off = reinterpret_cast<int32_t*>(blk + in_off);
    d_ids = reinterpret_cast<int32_t*>(blk + in_ids);
    d_score = h->res_blk.as<double>();
    d_idx = reinterpret_cast<int32_t*>(h->res_blk.as<char>() + out_idx_at);
    if (!csr_non_decreasing(pref_off, R)) {
      set_error("klm_schedule_batch: pref_off is not non-decreasing");
      return KLM_EINVAL;
    }
    std::memcpy(h->sched_pin.at(0), min_battery, min_bytes);
    std::memcpy(h->sched_pin.at(in_off), pref_off, off_bytes);
    if (ids_bytes) std::memcpy(h->sched_pin.at(in_ids), pref_ids, ids_bytes);
    KLM_TRY(h2d(blk, h->sched_pin.p, in_bytes, h->stream));
  } else {
    KLM_TRY(h->req_min.ensure(min_bytes));
    KLM_TRY(h->req_off.ensure(off_bytes));
    KLM_TRY(h->req_ids.ensure(ids_bytes ? ids_bytes : sizeof(int32_t)));
    KLM_TRY(h->res_idx.ensure(idx_bytes));
    KLM_TRY(h->res_score.ensure(min_bytes));
    d_min = h->req_min.as<double>();
    d_off = h->req_off.as<int32_t>();
    d_ids = h->req_ids.as<int32_t>();
    d_idx = h->res_idx.as<int32_t>();
    d_score = h->res_score.as<double>();
    if (algo == KLM_ALGO_DECOMPOSED && R >= 2 * kChunkRequests) {
      // A large batch is PCIe time, not kernel time (C4: 17 MB in, 12 MB out, 16 us of K2),
      // so it is cut into chunks that alternate between the two streams: chunk c's download
      // runs while chunk c+1 uploads (separate copy engines) and K2 disappears between them.
      // With pageable caller arrays the driver stages every copy synchronously and the
      // chunks simply run back to back.
      const int64_t n_chunks = std::min<int64_t>(kMaxChunks, R / kChunkRequests);
      for (int64_t c = 0; c < n_chunks; ++c) {  // before anything is in flight
        const int64_t i0 = pref_off[R * c / n_chunks], i1 = pref_off[R * (c + 1) / n_chunks];
        if (i0 < 0 || i1 < i0 || i1 > n_ids) {
          set_error("klm_schedule_batch: pref_off is not non-decreasing within [0, pref_off[R]]");
          return KLM_EINVAL;
        }
      }
      cudaStream_t ss[2] = {h->stream, h->stream2};
      auto pipeline = [&]() -> int {
        // stream2 has not seen the node table or K1's result (a staged klm_nodes_set does not
        // synchronise)
        KLM_CUDA(cudaEventRecord(h->ev_fork, h->stream));
        KLM_CUDA(cudaStreamWaitEvent(h->stream2, h->ev_fork, 0));
        for (int64_t c = 0; c < n_chunks; ++c) {
          const int64_t r0 = R * c / n_chunks, r1 = R * (c + 1) / n_chunks;
          const int64_t i0 = pref_off[r0], i1 = pref_off[r1];
          cudaStream_t st = ss[c & 1];
          // (checked chunk by chunk: the previous chunk's transfers are in flight meanwhile)
          if (!csr_non_decreasing(pref_off + r0, r1 - r0)) {
            set_error("klm_schedule_batch: pref_off is not non-decreasing");
            return KLM_EINVAL;
          }
          // offsets (r0, r1]: entry r0 came with the previous chunk (entry 0 with the first)
          const int64_t o0 = c == 0 ? 0 : r0 + 1;
          KLM_TRY(h2d(d_min + r0, min_battery + r0, (size_t)(r1 - r0) * sizeof(double), st));
          KLM_TRY(h2d(d_off + o0, pref_off + o0, (size_t)(r1 + 1 - o0) * sizeof(int32_t), st));
          KLM_TRY(h2d(d_ids + i0, pref_ids + i0, (size_t)(i1 - i0) * sizeof(int32_t), st));
          KLM_CUDA(cudaEventRecord(h->ev_chain[c & 1], st));
          // ... which is why this chunk's kernel waits for the previous chunk's upload
          if (c > 0) KLM_CUDA(cudaStreamWaitEvent(st, h->ev_chain[(c - 1) & 1], 0));
          KLM_TRY(klm_dev_schedule(h->d_battery, h->d_flags, h->n_nodes, 0, h->best.as<int64_t>(), r1 - r0,
                                   d_min + r0, d_off + r0, d_ids, d_idx + r0, d_score + r0, st));
          KLM_TRY(d2h(out_idx + r0, d_idx + r0, (size_t)(r1 - r0) * sizeof(int32_t), st));
          KLM_TRY(d2h(out_score + r0, d_score + r0, (size_t)(r1 - r0) * sizeof(double), st));
        }
        KLM_TRY(sync(h->stream2));
        return sync(h->stream);
      };
      const int rc = pipeline();
      if (rc != KLM_OK) {  // earlier chunks may still be writing the caller's arrays
        cudaStreamSynchronize(h->stream);
        cudaStreamSynchronize(h->stream2);
      }
      return rc;
    }
    if (!csr_non_decreasing(pref_off, R)) {
      set_error("klm_schedule_batch: pref_off is not non-decreasing");
      return KLM_EINVAL;
    }
    KLM_TRY(h2d(d_min, min_battery, min_bytes, h->stream));
    KLM_TRY(h2d(d_off, pref_off, off_bytes, h->stream));
    KLM_TRY(h2d(d_ids, pref_ids, ids_bytes, h->stream));
  }
  if (algo == KLM_ALGO_DECOMPOSED) {
    KLM_TRY(klm_dev_schedule(h->d_battery, h->d_flags, h->n_nodes, 0, h->best.as<int64_t>(), R, d_min,
                             d_off, d_ids, d_idx, d_score, h->stream));
  } else {
    KLM_TRY(klm_dev_schedule_bruteforce(h->d_battery, h->d_flags, h->n_nodes, 0, R, d_min, d_off, d_ids,
                                        d_idx, d_score, h->stream));
  }
  if (staged) {
    // results come back into the same staging block (the inputs have been read by then:
    // same stream), then out to the caller's arrays
    KLM_TRY(d2h(h->sched_pin.p, d_score, out_bytes, h->stream));
    KLM_TRY(sync(h->stream));
    std::memcpy(out_score, h->sched_pin.at(0), min_bytes);
    std::memcpy(out_idx, h->sched_pin.at(out_idx_at), idx_bytes);
    return KLM_OK;
  }
  KLM_TRY(d2h(out_idx, d_idx, idx_bytes, h->stream));
  KLM_TRY(d2h(out_score, d_score, min_bytes, h->stream));
  return sync(h->stream);
}

extern "C" int klm_schedule_batch(klm_handle* h, int64_t R, const double* min_battery,
                                  const int32_t* pref_off, const int32_t* pref_ids,
                                  int32_t* out_idx, double* out_score, int mode) {
  KLM_ENTER(h);
  int algo = 0;
  bool node_shard = false;
  if (!split_mode(mode, &algo, &node_shard)) {
    set_error("klm_schedule_batch: bad mode %d", mode);
    return KLM_EINVAL;
  }
  if (h->devs.empty())  // one device: the two sharding modes are the same thing
    return schedule_batch_impl(h, R, min_battery, pref_off, pref_ids, out_idx, out_score, algo);
  if (h->md_nodes < 0) {
    set_error("klm_schedule_batch: no node snapshot; call klm_nodes_set first");
    return KLM_ESTATE;
  }
  const int64_t n = h->md_nodes;
  if (h->md_uploaded_mode != (node_shard ? 1 : 0) && !h->md_tab_valid) {
    set_error("klm_schedule_batch: the snapshot installed by klm_reconcile is held in the other sharding "
              "layout; install it with klm_nodes_set to use both");
    return KLM_ESTATE;
  }
  const double* hb = h->md_tab_valid ? reinterpret_cast<const double*>(h->md_tab.p) : nullptr;
  const uint8_t* hf = h->md_tab_valid ? reinterpret_cast<const uint8_t*>(h->md_tab.at(align256((size_t)n * 8))) : nullptr;
  const int rc = multi_schedule(h, n, hb, hf, R, min_battery, pref_off, pref_ids, out_idx, out_score, algo,
                                node_shard);
  if (rc != KLM_OK) multi_sync(h);  // nothing may still write the caller's arrays
  return rc;
}

// The k best candidates of every request (the reference keeps only [0] of its
// sorted list, controller.go:160-161).
extern "C" int klm_schedule_batch_topk(klm_handle* h, int64_t R, const double* min_battery,
                                       const int32_t* pref_off, const int32_t* pref_ids, int k,
                                       int32_t* out_idx, double* out_score) {
  KLM_ENTER(h);
  if (!h->devs.empty()) {
    set_error("klm_schedule_batch_topk: not available on a handle over several devices");
    return KLM_EINVAL;
  }
  if (h->n_nodes < 0) {
    set_error("klm_schedule_batch_topk: no node snapshot; call klm_nodes_set first");
    return KLM_ESTATE;
  }
  if (R < 0 || k < 1 || k > klm_topk_max()) {
    set_error("klm_schedule_batch_topk: bad argument (1 <= k <= %d)", klm_topk_max());
    return k > klm_topk_max() ? KLM_ELIMIT : KLM_EINVAL;
  }
  if (R == 0) return KLM_OK;
  if (!min_battery || !pref_off || !out_idx || !out_score) {
    set_error("klm_schedule_batch_topk: null array");
    return KLM_EINVAL;
  }
  const int64_t n_ids = pref_off[R];
  if (pref_off[0] != 0 || n_ids < 0 || (n_ids > 0 && !pref_ids) || !csr_non_decreasing(pref_off, R)) {
    set_error("klm_schedule_batch_topk: pref_off must start at 0 and not decrease; pref_ids non-null");
    return KLM_EINVAL;
  }
  if (!h->topk_scratch.p) {
    KLM_TRY(h->topk_scratch.ensure((size_t)klm_dev_node_topk_scratch_bytes()));
    KLM_TRY(h->topk_top.ensure(2 * 8 * sizeof(int64_t)));
    KLM_CUDA(cudaMemsetAsync(h->topk_scratch.p, 0, 256, h->stream));
  }
  const size_t min_bytes = (size_t)R * 8, off_bytes = (size_t)(R + 1) * 4, ids_bytes = (size_t)n_ids * 4;
  KLM_TRY(h->req_min.ensure(min_bytes));
  KLM_TRY(h->req_off.ensure(off_bytes));
  KLM_TRY(h->req_ids.ensure(ids_bytes ? ids_bytes : 4));
  KLM_TRY(h->res_idx.ensure((size_t)R * k * 4));
  KLM_TRY(h->res_score.ensure((size_t)R * k * 8));
  cudaStream_t st = h->stream;
  KLM_TRY(h2d(h->req_min.p, min_battery, min_bytes, st));
  KLM_TRY(h2d(h->req_off.p, pref_off, off_bytes, st));
  KLM_TRY(h2d(h->req_ids.p, pref_ids, ids_bytes, st));
  KLM_TRY(klm_dev_node_topk(h->d_battery, h->d_flags, h->n_nodes, 0, k, h->topk_top.as<int64_t>(),
                            h->topk_scratch.p, st));
  KLM_TRY(klm_dev_schedule_topk(h->d_battery, h->d_flags, h->n_nodes, 0, h->topk_top.as<int64_t>(), k, R,
                                h->req_min.as<double>(), h->req_off.as<int32_t>(), h->req_ids.as<int32_t>(),
                                h->res_idx.as<int32_t>(), h->res_score.as<double>(), st));
  KLM_TRY(d2h(out_idx, h->res_idx.p, (size_t)R * k * 4, st));
  KLM_TRY(d2h(out_score, h->res_score.p, (size_t)R * k * 8, st));
  return sync(st);
}

// One reconcile (controller.go:88-110): install the snapshot and decide every pending request.
// Small reconciles (everything within the staging block) are one upload, K1, K2 launched
// programmatically dependent on K1, one download and one synchronisation; anything larger
// is klm_nodes_set followed by klm_schedule_batch under one lock.
extern "C" int klm_reconcile(klm_handle* h, int64_t n, const double* battery, const uint8_t* flags,
                             int64_t R, const double* min_battery, const int32_t* pref_off,
                             const int32_t* pref_ids, int32_t* out_idx, double* out_score, int mode) {
  KLM_ENTER(h);
  int algo = 0;
  bool node_shard = false;
  if (!split_mode(mode, &algo, &node_shard)) {
    set_error("klm_reconcile: bad mode %d", mode);
    return KLM_EINVAL;
  }
  if (!h->devs.empty()) {
    // straight from the caller's arrays (they are valid for the whole call): no host copy
    if (n < 0 || n > 0x7fffffff || (n > 0 && (!battery || !flags))) {
      set_error("klm_nodes_set: bad argument (0 <= n_nodes < 2^31, arrays non-null)");
      return KLM_EINVAL;
    }
    h->md_nodes = -1;
    h->md_uploaded_mode = -1;
    h->md_tab_valid = false;
    const int rc = multi_schedule(h, n, battery, flags, R, min_battery, pref_off, pref_ids, out_idx, out_score,
                                  algo, node_shard);
    if (rc != KLM_OK) {
      multi_sync(h);
      h->md_uploaded_mode = -1;
      return rc;
    }
    h->md_nodes = n;
    return KLM_OK;
  }
  const bool args_ok = n >= 0 && n <= 0x7fffffff && (n == 0 || (battery && flags)) && R >= 0 &&
                       (algo == KLM_ALGO_DECOMPOSED || algo == KLM_ALGO_BRUTEFORCE) &&
                       (R == 0 || (min_battery && pref_off && out_idx && out_score));
  const int64_t n_ids = (args_ok && R > 0) ? pref_off[R] : 0;
  const size_t bat_bytes = (size_t)(n > 0 ? n : 0) * sizeof(double), flag_off = align256(bat_bytes);
  const size_t tab_bytes = align256(flag_off + (size_t)(n > 0 ? n : 0));
  const size_t min_bytes = (size_t)R * sizeof(double), off_bytes = (size_t)(R + 1) * sizeof(int32_t);
  const size_t ids_bytes = (size_t)(n_ids > 0 ? n_ids : 0) * sizeof(int32_t), idx_bytes = (size_t)R * sizeof(int32_t);
  const size_t q_off = tab_bytes + align256(min_bytes), q_ids = q_off + align256(off_bytes);
  const size_t in_bytes = q_ids + align256(ids_bytes), out_idx_at = align256(min_bytes);
  const size_t out_bytes = out_idx_at + idx_bytes;
  if (!args_ok || R == 0 || n_ids < 0 || pref_off[0] != 0 || (n_ids > 0 && !pref_ids) ||
      in_bytes + out_bytes > kStageMax || !csr_non_decreasing(pref_off, R)) {
    // let the two calls validate and handle it (same error texts as calling them separately)
    KLM_TRY(nodes_set_impl(h, n, battery, flags));
    return schedule_batch_impl(h, R, min_battery, pref_off, pref_ids, out_idx, out_score, algo);
  }
  h->n_nodes = -1;
  KLM_TRY(h->rec_blk.ensure(in_bytes + 256));
  KLM_TRY(h->res_blk.ensure(out_bytes + 256));
  KLM_TRY(h->sched_pin.ensure(kStageMax));
  KLM_CUDA(cudaEventSynchronize(h->sched_read));
  char* blk = h->rec_blk.as<char>();
  h->d_battery = reinterpret_cast<double*>(blk);
  h->d_flags = reinterpret_cast<uint8_t*>(blk + flag_off);
  double* d_min = reinterpret_cast<double*>(blk + tab_bytes);
  int32_t* d_off = reinterpret_cast<int32_t*>(blk + q_off);
  int32_t* d_ids = reinterpret_cast<int32_t*>(blk + q_ids);
  double* d_score = h->res_blk.as<double>();
  int32_t* d_idx = reinterpret_cast<int32_t*>(h->res_blk.as<char>() + out_idx_at);
  if (n > 0) {
    std::memcpy(h->sched_pin.at(0), battery, bat_bytes);
    std::memcpy(h->sched_pin.at(flag_off), flags, (size_t)n);
  }
  std::memcpy(h->sched_pin.at(tab_bytes), min_battery, min_bytes);
  std::memcpy(h->sched_pin.at(q_off), pref_off, off_bytes);
  if (ids_bytes) std::memcpy(h->sched_pin.at(q_ids), pref_ids, ids_bytes);
  KLM_TRY(h2d(blk, h->sched_pin.p, in_bytes, h->stream));
  KLM_TRY(klm_dev_node_best(h->d_battery, h->d_flags, n, 0, h->best.as<int64_t>(), h->best_scratch.p,
                            h->stream));
  h->n_nodes = n;
  if (algo == KLM_ALGO_DECOMPOSED) {
    // K1 (just enqueued) writes nothing K2 reads except `best`: the contract of the dependent launch
    KLM_TRY(klm_dev_schedule_after_best(h->d_battery, h->d_flags, n, 0, h->best.as<int64_t>(), R, d_min, d_off,
                                        d_ids, d_idx, d_score, h->stream));
  } else {
    KLM_TRY(klm_dev_schedule_bruteforce(h->d_battery, h->d_flags, n, 0, R, d_min, d_off, d_ids, d_idx, d_score,
                                        h->stream));
  }
  KLM_TRY(d2h(h->sched_pin.p, d_score, out_bytes, h->stream));
  KLM_TRY(sync(h->stream));
  std::memcpy(out_score, h->sched_pin.at(0), min_bytes);
  std::memcpy(out_idx, h->sched_pin.at(out_idx_at), idx_bytes);
  return KLM_OK;
}

// ----------------------------------------------------- cluster aggregate

extern "C" int klm_cluster_aggregate(klm_handle* h, const klm_node_cols* c, int64_t n,
                                     const uint8_t* pod_running, int64_t n_pods,
                                     klm_cluster_out* out) {
  KLM_ENTER(h);
  KLM_TRY(collect_idle(h, "klm_cluster_aggregate"));
  if (!c || !out || n < 0 || n_pods < 0 || c->n_gpu_usage < 0 ||
      (n > 0 && (!c->cpu_capacity || !c->cpu_usage || !c->memory_capacity || !c->memory_usage ||
                 !c->gpu_count || !c->healthy)) ||
      (n_pods > 0 && !pod_running) || (c->n_gpu_usage > 0 && !c->gpu_usage)) {
    set_error("klm_cluster_aggregate: bad argument");
    return KLM_EINVAL;
  }
  // one device block, columns at 256-byte aligned offsets
  const size_t col8 = align256((size_t)n * 8), col4 = align256((size_t)n * 4),
               col1 = align256((size_t)n);
  KLM_TRY(h->agg_cols.ensure(4 * col8 + col4 + col1 + 256));
  KLM_TRY(h->agg_pods.ensure((size_t)n_pods + 256));
  KLM_TRY(h->agg_gpu.ensure((size_t)c->n_gpu_usage * 8 + 256));
  char* base = h->agg_cols.as<char>();
  klm_node_cols d;
  d.cpu_capacity = reinterpret_cast<int64_t*>(base);
  d.cpu_usage = reinterpret_cast<int64_t*>(base + col8);
  d.memory_capacity = reinterpret_cast<int64_t*>(base + 2 * col8);
  d.memory_usage = reinterpret_cast<int64_t*>(base + 3 * col8);
  d.gpu_count = reinterpret_cast<int32_t*>(base + 4 * col8);
  d.healthy = reinterpret_cast<uint8_t*>(base + 4 * col8 + col4);
  d.gpu_usage = h->agg_gpu.as<double>();
  d.n_gpu_usage = c->n_gpu_usage;
  cudaStream_t st = h->stream;
  KLM_TRY(h2d((void*)d.cpu_capacity, c->cpu_capacity, (size_t)n * 8, st));
  KLM_TRY(h2d((void*)d.cpu_usage, c->cpu_usage, (size_t)n * 8, st));
  KLM_TRY(h2d((void*)d.memory_capacity, c->memory_capacity, (size_t)n * 8, st));
  KLM_TRY(h2d((void*)d.memory_usage, c->memory_usage, (size_t)n * 8, st));
  KLM_TRY(h2d((void*)d.gpu_count, c->gpu_count, (size_t)n * 4, st));
  KLM_TRY(h2d((void*)d.healthy, c->healthy, (size_t)n, st));
  KLM_TRY(h2d((void*)d.gpu_usage, c->gpu_usage, (size_t)c->n_gpu_usage * 8, st));
  KLM_TRY(h2d(h->agg_pods.p, pod_running, (size_t)n_pods, st));
  KLM_TRY(klm_dev_cluster_aggregate(&d, n, h->agg_pods.as<uint8_t>(), n_pods,
                                    h->agg_out.as<klm_cluster_out>(), h->agg_scratch.p, st));
  KLM_TRY(d2h(out, h->agg_out.p, sizeof(klm_cluster_out), st));
  return sync(st);
}

extern "C" int klm_node_rates(klm_handle* h, int64_t n, const int64_t* cpu_cap,
                              const int64_t* cpu_use, const int64_t* mem_cap,
                              const int64_t* mem_use, const int64_t* disk_cap,
                              const int64_t* disk_alloc, double* cpu_rate, double* mem_rate,
                              int64_t* disk_usage, double* disk_rate, uint8_t* under_pressure) {
  KLM_ENTER(h);
  if (n < 0 || (n > 0 && (!cpu_cap || !cpu_use || !mem_cap || !mem_use))) {
    set_error("klm_node_rates: bad argument");
    return KLM_EINVAL;
  }
  if (n == 0) return KLM_OK;
  const size_t col = align256((size_t)n * 8);
  KLM_TRY(h->rate_in.ensure(6 * col));
  KLM_TRY(h->rate_out.ensure(5 * col));
  char* in = h->rate_in.as<char>();
  char* outp = h->rate_out.as<char>();
  cudaStream_t st = h->stream;
  const int64_t* src[6] = {cpu_cap, cpu_use, mem_cap, mem_use, disk_cap, disk_alloc};
  const int64_t* dev[6];
  for (int k = 0; k < 6; ++k) {
    dev[k] = src[k] ? reinterpret_cast<const int64_t*>(in + k * col) : nullptr;
    if (src[k]) KLM_TRY(h2d(in + k * col, src[k], (size_t)n * 8, st));
  }
  double* d_cr = reinterpret_cast<double*>(outp);
  double* d_mr = reinterpret_cast<double*>(outp + col);
  int64_t* d_du = reinterpret_cast<int64_t*>(outp + 2 * col);
  double* d_dr = reinterpret_cast<double*>(outp + 3 * col);
  uint8_t* d_up = reinterpret_cast<uint8_t*>(outp + 4 * col);
  KLM_TRY(klm_dev_node_rates(n, dev[0], dev[1], dev[2], dev[3], dev[4], dev[5], d_cr, d_mr, d_du,
                             d_dr, d_up, st));
  if (cpu_rate) KLM_TRY(d2h(cpu_rate, d_cr, (size_t)n * 8, st));
  if (mem_rate) KLM_TRY(d2h(mem_rate, d_mr, (size_t)n * 8, st));
  if (disk_usage) KLM_TRY(d2h(disk_usage, d_du, (size_t)n * 8, st));
  if (disk_rate) KLM_TRY(d2h(disk_rate, d_dr, (size_t)n * 8, st));
  if (under_pressure) KLM_TRY(d2h(under_pressure, d_up, (size_t)n, st));
  return sync(st);
}

extern "C" int klm_pod_rates(klm_handle* h, int64_t n_pods, const int32_t* ctr_off,
                             const int64_t* c_cpu_use, const int64_t* c_mem_use,
                             const int64_t* c_cpu_req, const int64_t* c_cpu_lim,
                             const int64_t* c_mem_req, const int64_t* c_mem_lim,
                             int64_t* pod_totals, double* cpu_rate, double* mem_rate,
                             double* cpu_util, double* mem_util, uint8_t* over_limit) {
  KLM_ENTER(h);
  if (n_pods < 0 || (n_pods > 0 && !ctr_off)) {
    set_error("klm_pod_rates: bad argument");
    return KLM_EINVAL;
  }
  if (n_pods == 0) return KLM_OK;
  const int64_t n_ctr = ctr_off[n_pods];
  if (ctr_off[0] != 0 || n_ctr < 0 ||
      (n_ctr > 0 && (!c_cpu_use || !c_mem_use || !c_cpu_req || !c_cpu_lim || !c_mem_req ||
                     !c_mem_lim))) {
    set_error("klm_pod_rates: ctr_off must start at 0 and container columns be non-null");
    return KLM_EINVAL;
  }
  if (!csr_non_decreasing(ctr_off, n_pods)) {
    set_error("klm_pod_rates: ctr_off is not non-decreasing");
    return KLM_EINVAL;
  }
  const size_t ccol = align256((size_t)n_ctr * 8), off_b = align256((size_t)(n_pods + 1) * 4);
  const size_t pcol = align256((size_t)n_pods * 8);
  KLM_TRY(h->rate_in.ensure(6 * ccol + off_b));
  KLM_TRY(h->rate_out.ensure(11 * pcol));
  char* in = h->rate_in.as<char>();
  char* outp = h->rate_out.as<char>();
  cudaStream_t st = h->stream;
  const int64_t* src[6] = {c_cpu_use, c_mem_use, c_cpu_req, c_cpu_lim, c_mem_req, c_mem_lim};
  for (int k = 0; k < 6; ++k) KLM_TRY(h2d(in + k * ccol, src[k], (size_t)n_ctr * 8, st));
  KLM_TRY(h2d(in + 6 * ccol, ctr_off, (size_t)(n_pods + 1) * 4, st));
  auto C = [&](int k) { return reinterpret_cast<const int64_t*>(in + k * ccol); };
  DevBuf& tot = h->pod_tot;  // [6][P] int64, contiguous like the caller's
  KLM_TRY(tot.ensure((size_t)6 * n_pods * 8));
  double* d_cr = reinterpret_cast<double*>(outp);
  double* d_mr = reinterpret_cast<double*>(outp + pcol);
  double* d_cu = reinterpret_cast<double*>(outp + 2 * pcol);
  double* d_mu = reinterpret_cast<double*>(outp + 3 * pcol);
  uint8_t* d_ov = reinterpret_cast<uint8_t*>(outp + 4 * pcol);
  KLM_TRY(klm_dev_pod_rates(n_pods, reinterpret_cast<const int32_t*>(in + 6 * ccol), C(0), C(1),
                            C(2), C(3), C(4), C(5), tot.as<int64_t>(), d_cr, d_mr, d_cu, d_mu,
                            d_ov, st));
  if (pod_totals) KLM_TRY(d2h(pod_totals, tot.p, (size_t)6 * n_pods * 8, st));
  if (cpu_rate) KLM_TRY(d2h(cpu_rate, d_cr, (size_t)n_pods * 8, st));
  if (mem_rate) KLM_TRY(d2h(mem_rate, d_mr, (size_t)n_pods * 8, st));
  if (cpu_util) KLM_TRY(d2h(cpu_util, d_cu, (size_t)n_pods * 8, st));
  if (mem_util) KLM_TRY(d2h(mem_util, d_mu, (size_t)n_pods * 8, st));
  if (over_limit) KLM_TRY(d2h(over_limit, d_ov, (size_t)n_pods, st));
  return sync(st);
}

extern "C" int klm_fleet_scan(klm_handle* h, int64_t n, const uint8_t* status_ok,
                              const double* pct, double threshold, int64_t* healthy_count,
                              int32_t* low_idx, int64_t* n_low) {
  KLM_ENTER(h);
  if (n < 0 || n > 0x7fffffff || !healthy_count || !n_low ||
      (n > 0 && (!status_ok || !pct || !low_idx))) {
    set_error("klm_fleet_scan: bad argument");
    return KLM_EINVAL;
  }
  const size_t pb = align256((size_t)n * 8), sb = align256((size_t)n);
  KLM_TRY(h->fleet_in.ensure(pb + sb + 256));
  KLM_TRY(h->fleet_out.ensure(256 + (size_t)n * 4));
  KLM_TRY(h->fleet_scratch.ensure((size_t)klm_dev_fleet_scan_scratch_bytes(n)));
  char* in = h->fleet_in.as<char>();
  char* outp = h->fleet_out.as<char>();
  cudaStream_t st = h->stream;
  KLM_TRY(h2d(in, pct, (size_t)n * 8, st));
  KLM_TRY(h2d(in + pb, status_ok, (size_t)n, st));
  int64_t* d_cnt = reinterpret_cast<int64_t*>(outp);
  int32_t* d_idx = reinterpret_cast<int32_t*>(outp + 256);
  KLM_TRY(klm_dev_fleet_scan(n, reinterpret_cast<uint8_t*>(in + pb),
                             reinterpret_cast<double*>(in), threshold, d_cnt, d_idx,
                             h->fleet_scratch.p, st));
  int64_t cnt[2] = {0, 0};
  KLM_TRY(d2h(cnt, d_cnt, sizeof(cnt), st));
  KLM_TRY(sync(st));
  *healthy_count = cnt[0];
  *n_low = cnt[1];
  if (cnt[1] > 0) {
    KLM_TRY(d2h(low_idx, d_idx, (size_t)cnt[1] * 4, st));
    KLM_TRY(sync(st));
  }
  return KLM_OK;
}


// ------------------------------------------------ element-wise leftovers (A8 / A10)

namespace {
// upload `bytes` of src into blk at off (256-aligned parts); returns the device pointer or NULL
template <typename T>
const T* up(klm_handle* h, char* blk, size_t& off, const T* src, size_t count, int* rc) {
  if (!src || count == 0 || *rc != KLM_OK) return nullptr;
  char* d = blk + off;
  off += align256(count * sizeof(T));
  *rc = h2d(d, src, count * sizeof(T), h->stream);
  return reinterpret_cast<const T*>(d);
}
}  // namespace

extern "C" int klm_node_available(klm_handle* h, int64_t n, const int64_t* cpu_cap, const int64_t* cpu_use,
                                  const int64_t* mem_cap, const int64_t* mem_use, const int64_t* disk_cap,
                                  const int64_t* disk_use, double* cpu_cores, double* mem_gb, double* disk_gb) {
  KLM_ENTER(h);
  if (n < 0 || (n > 0 && ((cpu_cores && (!cpu_cap || !cpu_use)) || (mem_gb && (!mem_cap || !mem_use))))) {
    set_error("klm_node_available: bad argument");
    return KLM_EINVAL;
  }
  if (n == 0) return KLM_OK;
  const size_t col = align256((size_t)n * 8);
  KLM_TRY(h->rate_in.ensure(6 * col));
  KLM_TRY(h->rate_out.ensure(3 * col));
  size_t off = 0;
  int rc = KLM_OK;
  char* in = h->rate_in.as<char>();
  const int64_t* d[6];
  const int64_t* src[6] = {cpu_cap, cpu_use, mem_cap, mem_use, disk_cap, disk_use};
  for (int k = 0; k < 6; ++k) d[k] = up(h, in, off, src[k], (size_t)n, &rc);
  KLM_TRY(rc);
  double* o = h->rate_out.as<double>();
  double* o1 = reinterpret_cast<double*>(h->rate_out.as<char>() + col);
  double* o2 = reinterpret_cast<double*>(h->rate_out.as<char>() + 2 * col);
  KLM_TRY(klm_dev_node_available(n, d[0], d[1], d[2], d[3], d[4], d[5], cpu_cores ? o : nullptr,
                                 mem_gb ? o1 : nullptr, disk_gb ? o2 : nullptr, h->stream));
  if (cpu_cores) KLM_TRY(d2h(cpu_cores, o, (size_t)n * 8, h->stream));
  if (mem_gb) KLM_TRY(d2h(mem_gb, o1, (size_t)n * 8, h->stream));
  if (disk_gb) KLM_TRY(d2h(disk_gb, o2, (size_t)n * 8, h->stream));
  return sync(h->stream);
}

extern "C" int klm_network_quality(klm_handle* h, int64_t n, const uint8_t* connected, const double* rtt,
                                   uint8_t* quality) {
  KLM_ENTER(h);
  if (n < 0 || (n > 0 && (!connected || !rtt || !quality))) {
    set_error("klm_network_quality: bad argument");
    return KLM_EINVAL;
  }
  if (n == 0) return KLM_OK;
  KLM_TRY(h->rate_in.ensure(align256((size_t)n * 8) + align256((size_t)n)));
  KLM_TRY(h->rate_out.ensure((size_t)n));
  size_t off = 0;
  int rc = KLM_OK;
  char* in = h->rate_in.as<char>();
  const double* d_rtt = up(h, in, off, rtt, (size_t)n, &rc);
  const uint8_t* d_con = up(h, in, off, connected, (size_t)n, &rc);
  KLM_TRY(rc);
  KLM_TRY(klm_dev_network_quality(n, d_con, d_rtt, h->rate_out.as<uint8_t>(), h->stream));
  KLM_TRY(d2h(quality, h->rate_out.p, (size_t)n, h->stream));
  return sync(h->stream);
}

extern "C" int klm_pod_restarts(klm_handle* h, int64_t n_pods, const int32_t* st_off,
                                const int32_t* restart_count, int32_t* restarts) {
  KLM_ENTER(h);
  if (n_pods < 0 || (n_pods > 0 && (!st_off || !restarts))) {
    set_error("klm_pod_restarts: bad argument");
    return KLM_EINVAL;
  }
  if (n_pods == 0) return KLM_OK;
  const int64_t n_st = st_off[n_pods];
  if (st_off[0] != 0 || n_st < 0 || (n_st > 0 && !restart_count) || !csr_non_decreasing(st_off, n_pods)) {
    set_error("klm_pod_restarts: st_off must start at 0 and not decrease; restart_count non-null");
    return KLM_EINVAL;
  }
  KLM_TRY(h->rate_in.ensure(align256((size_t)(n_pods + 1) * 4) + align256((size_t)n_st * 4) + 256));
  KLM_TRY(h->rate_out.ensure((size_t)n_pods * 4));
  size_t off = 0;
  int rc = KLM_OK;
  char* in = h->rate_in.as<char>();
  const int32_t* d_off = up(h, in, off, st_off, (size_t)n_pods + 1, &rc);
  const int32_t* d_cnt = up(h, in, off, restart_count, (size_t)n_st, &rc);
  KLM_TRY(rc);
  KLM_TRY(klm_dev_pod_restarts(n_pods, d_off, d_cnt, h->rate_out.as<int32_t>(), h->stream));
  KLM_TRY(d2h(restarts, h->rate_out.p, (size_t)n_pods * 4, h->stream));
  return sync(h->stream);
}

extern "C" int klm_node_health(klm_handle* h, int64_t n, const int32_t* cond_off, const uint8_t* cond_type,
                               const uint8_t* cond_status, uint8_t* healthy, uint8_t* failing) {
  KLM_ENTER(h);
  if (n < 0 || (n > 0 && !cond_off)) {
    set_error("klm_node_health: bad argument");
    return KLM_EINVAL;
  }
  if (n == 0) return KLM_OK;
  const int64_t nc = cond_off[n];
  if (cond_off[0] != 0 || nc < 0 || (nc > 0 && (!cond_type || !cond_status)) || !csr_non_decreasing(cond_off, n)) {
    set_error("klm_node_health: cond_off must start at 0 and not decrease; condition columns non-null");
    return KLM_EINVAL;
  }
  KLM_TRY(h->rate_in.ensure(align256((size_t)(n + 1) * 4) + 2 * align256((size_t)nc) + 256));
  KLM_TRY(h->rate_out.ensure(2 * align256((size_t)n)));
  size_t off = 0;
  int rc = KLM_OK;
  char* in = h->rate_in.as<char>();
  const int32_t* d_off = up(h, in, off, cond_off, (size_t)n + 1, &rc);
  const uint8_t* d_t = up(h, in, off, cond_type, (size_t)nc, &rc);
  const uint8_t* d_s = up(h, in, off, cond_status, (size_t)nc, &rc);
  KLM_TRY(rc);
  uint8_t* o_h = h->rate_out.as<uint8_t>();
  uint8_t* o_f = o_h + align256((size_t)n);
  KLM_TRY(klm_dev_node_health(n, d_off, d_t, d_s, o_h, o_f, h->stream));
  if (healthy) KLM_TRY(d2h(healthy, o_h, (size_t)n, h->stream));
  if (failing) KLM_TRY(d2h(failing, o_f, (size_t)n, h->stream));
  return sync(h->stream);
}

// ------------------------------------------------------------------ ring

extern "C" int klm_ring_create(klm_handle* h, int64_t n_series, int32_t window) {
  KLM_ENTER(h);
  KLM_TRY(collect_idle(h, "klm_ring_create"));
  if (n_series < 1 || window < 1) {
    set_error("klm_ring_create: n_series and window must be positive");
    return KLM_EINVAL;
  }
  const int32_t pitch = (window + 3) & ~3;
  const size_t bytes = (size_t)n_series * pitch * sizeof(float);
  KLM_TRY(h->ring.ensure(bytes));
  KLM_TRY(h->seen.ensure((size_t)n_series * sizeof(uint32_t)));
  KLM_TRY(h->win_state.ensure((size_t)n_series * 2 * sizeof(float)));
  KLM_CUDA(cudaMemsetAsync(h->win_state.p, 0xFF, (size_t)n_series * 2 * sizeof(float), h->stream));  // NaN: no bracket
  KLM_CUDA(cudaMemsetAsync(h->ring.p, 0, bytes, h->stream));
  KLM_CUDA(cudaMemsetAsync(h->seen.p, 0, (size_t)n_series * sizeof(uint32_t), h->stream));
  KLM_TRY(sync(h->stream));
  h->n_series = n_series;
  h->window = window;
  h->pitch = pitch;
  h->q_calls.clear();  // pushes queued for the previous ring
  h->q_ids.clear();
  h->q_vals.clear();
  return KLM_OK;
}

static int ring_ready(klm_handle* h, const char* who) {
  if (h->n_series <= 0) {
    set_error("%s: no ring; call klm_ring_create first", who);
    return KLM_ESTATE;
  }
  return KLM_OK;
}

extern "C" int klm_ring_ingest(klm_handle* h, int64_t n, const int64_t* series_ids,
                               const float* values) {
  KLM_ENTER(h);
  KLM_TRY(ring_ready(h, "klm_ring_ingest"));
  if (n < 0 || (n > 0 && (!series_ids || !values))) {
    set_error("klm_ring_ingest: bad argument");
    return KLM_EINVAL;
  }
  if (n == 0) return KLM_OK;
  if (h->begun != h->ended) {
    // collect cycles are in flight (the pipelined collector always has one): the push is queued
    // and applied, in arrival order, by the next klm_collect_begin before its own tick --
    // UpdateUAVReport (manager.go:437-446) serialises with Collect under one mutex the same way
    try {
      h->q_calls.push_back({n, h->q_ids.size(), h->q_vals.size()});
      h->q_ids.insert(h->q_ids.end(), series_ids, series_ids + n);
      h->q_vals.insert(h->q_vals.end(), values, values + n);
    } catch (const std::bad_alloc&) {
      set_error("klm_ring_ingest: out of host memory");
      return KLM_EINVAL;
    }
    return KLM_OK;
  }
  KLM_TRY(flush_pushed(h, 0, h->stream));  // (pushes queued behind a pipeline that has drained)
  KLM_TRY(h->ing_ids.ensure((size_t)n * 8));
  KLM_TRY(h->ing_vals.ensure((size_t)n * 4));
  KLM_TRY(h2d(h->ing_ids.p, series_ids, (size_t)n * 8, h->stream));
  KLM_TRY(h2d(h->ing_vals.p, values, (size_t)n * 4, h->stream));
  KLM_TRY(klm_dev_ring_ingest(h->ring.as<float>(), h->seen.as<uint32_t>(), h->n_series,
                              h->window, h->pitch, n, h->ing_ids.as<int64_t>(),
                              h->ing_vals.as<float>(), h->stream));
  return sync(h->stream);
}

extern "C" int klm_ring_ingest_tick(klm_handle* h, const float* values) {
  KLM_ENTER(h);
  KLM_TRY(ring_ready(h, "klm_ring_ingest_tick"));
  if (!values) {
    set_error("klm_ring_ingest_tick: values is null");
    return KLM_EINVAL;
  }
  if (h->begun != h->ended) {  // queued like klm_ring_ingest, as one whole tick
    try {
      h->q_calls.push_back({h->n_series, SIZE_MAX, h->q_vals.size()});
      h->q_vals.insert(h->q_vals.end(), values, values + h->n_series);
    } catch (const std::bad_alloc&) {
      set_error("klm_ring_ingest_tick: out of host memory");
      return KLM_EINVAL;
    }
    return KLM_OK;
  }
  KLM_TRY(flush_pushed(h, 0, h->stream));
  KLM_TRY(h->ing_vals.ensure((size_t)h->n_series * 4));
  KLM_TRY(h2d(h->ing_vals.p, values, (size_t)h->n_series * 4, h->stream));
  KLM_TRY(klm_dev_ring_ingest(h->ring.as<float>(), h->seen.as<uint32_t>(), h->n_series,
                              h->window, h->pitch, h->n_series, nullptr,
                              h->ing_vals.as<float>(), h->stream));
  return sync(h->stream);
}

extern "C" int klm_ring_load(klm_handle* h, int64_t s0, int64_t s1, int32_t count,
                             const float* data) {
  KLM_ENTER(h);
  KLM_TRY(ring_ready(h, "klm_ring_load"));
  KLM_TRY(collect_idle(h, "klm_ring_load"));
  KLM_TRY(flush_pushed(h, 0, h->stream));
  if (s0 < 0 || s1 < s0 || s1 > h->n_series || count < 0 || count > h->window ||
      (s1 > s0 && count > 0 && !data)) {
    set_error("klm_ring_load: bad argument");
    return KLM_EINVAL;
  }
  if (s1 == s0) return KLM_OK;
  const int64_t rows = s1 - s0;
  cudaStream_t st = h->stream;
  // the rows change wholesale: their brackets would only cost time
  KLM_CUDA(cudaMemsetAsync(h->win_state.as<float>() + 2 * s0, 0xFF, (size_t)rows * 2 * sizeof(float), st));
  // stage through device memory in chunks of <= 64 Mi samples (256 MiB)
  int64_t rows_per_chunk = rows;
  if (count > 0) {
    rows_per_chunk = ((int64_t)64 << 20) / count;
    if (rows_per_chunk < 1) rows_per_chunk = 1;
  }
  KLM_TRY(h->ing_vals.ensure((size_t)(rows_per_chunk < rows ? rows_per_chunk : rows) *
                             (count > 0 ? count : 1) * 4));
  for (int64_t r = 0; r < rows; r += rows_per_chunk) {
    const int64_t m = rows - r < rows_per_chunk ? rows - r : rows_per_chunk;
    KLM_TRY(h2d(h->ing_vals.p, data + r * count, (size_t)m * count * 4, st));
    KLM_TRY(launch_ring_load(h->ring.as<float>(), h->seen.as<uint32_t>(), h->pitch, s0 + r, m,
                             count, h->ing_vals.as<float>(), st));
    KLM_TRY(sync(st));  // the staging buffer is reused by the next chunk
  }
  return KLM_OK;
}

extern "C" int klm_ring_stats(klm_handle* h, int64_t s0, int64_t s1, float* out_min,
                              float* out_max, double* out_mean, float* out_p95) {
  KLM_ENTER(h);
  KLM_TRY(ring_ready(h, "klm_ring_stats"));
  KLM_TRY(collect_idle(h, "klm_ring_stats"));
  KLM_TRY(flush_pushed(h, 0, h->stream));
  if (s0 < 0 || s1 < s0 || s1 > h->n_series) {
    set_error("klm_ring_stats: series range out of bounds");
    return KLM_EINVAL;
  }
  if (s1 == s0) return KLM_OK;
  const int64_t rows = s1 - s0;
  const size_t c4 = align256((size_t)rows * 4), c8 = align256((size_t)rows * 8);
  KLM_TRY(h->stats_out.ensure(3 * c4 + c8));
  char* o = h->stats_out.as<char>();
  float* d_min = reinterpret_cast<float*>(o);
  float* d_max = reinterpret_cast<float*>(o + c4);
  float* d_p95 = reinterpret_cast<float*>(o + 2 * c4);
  double* d_mean = reinterpret_cast<double*>(o + 3 * c4);
  cudaStream_t st = h->stream;
  KLM_TRY(ring_window_stats(h, s0, s1, d_min, d_max, d_mean, out_p95 ? d_p95 : nullptr, st));
  if (out_min) KLM_TRY(d2h(out_min, d_min, (size_t)rows * 4, st));
  if (out_max) KLM_TRY(d2h(out_max, d_max, (size_t)rows * 4, st));
  if (out_mean) KLM_TRY(d2h(out_mean, d_mean, (size_t)rows * 8, st));
  if (out_p95) KLM_TRY(d2h(out_p95, d_p95, (size_t)rows * 4, st));
  return sync(st);
}

// ------------------------------------------------------------- collect cycle

namespace {

// The two branches of a collect cycle share no data, so they run on two streams: the ring
// branch (tick upload, K6, K7, statistics download) on `stream`, the node-snapshot branch
// (column upload, K5, K4, small downloads) on `stream2`, where its copies and kernels hide
// behind K7. Both streams are joined before the call returns.
int collect_impl(klm_handle* h, const klm_collect_in* in, const klm_collect_out* out) {
  if (!in || !out || in->n_nodes < 0 || in->n_pods < 0 || (in->n_nodes > 0 && !in->nodes) ||
      (in->n_pods > 0 && !in->pod_running)) {
    set_error("klm_collect: bad argument");
    return KLM_EINVAL;
  }
  cudaStream_t st = h->stream, st2 = h->stream2;
  const bool want_win = out->win_min || out->win_max || out->win_mean || out->win_p95 || out->win_packed;
  if (out->win_packed && (out->win_min || out->win_max || out->win_mean || out->win_p95)) {
    set_error("klm_collect: win_packed excludes the separate win_* arrays");
    return KLM_EINVAL;
  }
  if ((in->tick_values || want_win) && h->n_series <= 0) {
    set_error("klm_collect: no ring; call klm_ring_create first");
    return KLM_ESTATE;
  }
  const int64_t n = in->n_nodes, np = in->n_pods;
  const klm_node_cols* c = in->nodes;
  const int64_t ng = (n > 0 && c) ? c->n_gpu_usage : 0;
  if (n > 0 && (!c->cpu_capacity || !c->cpu_usage || !c->memory_capacity || !c->memory_usage ||
                !c->gpu_count || !c->healthy || ng < 0 || (ng > 0 && !c->gpu_usage))) {
    set_error("klm_collect: null node column");
    return KLM_EINVAL;
  }
  // ---- ring branch: tick ingest, then the window statistics of every series
  const int64_t S = h->n_series;
  float *d_min = nullptr, *d_max = nullptr, *d_p95 = nullptr;
  double* d_mean = nullptr;
  if (in->tick_values) {
    KLM_TRY(h->ing_vals.ensure((size_t)S * 4));
    KLM_TRY(h2d(h->ing_vals.p, in->tick_values, (size_t)S * 4, st));
    KLM_TRY(klm_dev_ring_ingest(h->ring.as<float>(), h->seen.as<uint32_t>(), S, h->window, h->pitch, S,
                                nullptr, h->ing_vals.as<float>(), st));
  }
  if (want_win) {
    const size_t c4 = align256((size_t)S * 4), c8 = align256((size_t)S * 8);
    KLM_TRY(h->stats_out.ensure(3 * c4 + c8));
    char* o = h->stats_out.as<char>();
    d_min = reinterpret_cast<float*>(o);
    d_max = reinterpret_cast<float*>(o + c4);
    d_p95 = reinterpret_cast<float*>(o + 2 * c4);
    d_mean = reinterpret_cast<double*>(o + 3 * c4);
    KLM_TRY(ring_window_stats(h, 0, S, d_min, d_max, d_mean, (out->win_p95 || out->win_packed) ? d_p95 : nullptr, st));
  }
  // ---- snapshot branch: one upload of the columns serves the rates and the aggregate
  const size_t col8 = align256((size_t)n * 8), col4 = align256((size_t)n * 4), col1 = align256((size_t)n);
  const size_t o_gc = 6 * col8, o_hl = o_gc + col4, o_pod = o_hl + col1;
  const size_t o_gu = o_pod + align256((size_t)np), in_bytes = o_gu + align256((size_t)ng * 8);
  const size_t out_bytes = 256 + 5 * col8;  // cluster_out | cpu, memory, disk usage, disk rate, pressure
  const bool want_rates = (out->cpu_rate || out->memory_rate || out->disk_usage || out->disk_rate ||
                           out->under_pressure) && n > 0;
  const bool staged = in_bytes <= kStageMax && out_bytes <= kStageMax;
  klm_node_cols d{};
  char* rout = nullptr;
  klm_cluster_out* d_cluster = h->agg_out.as<klm_cluster_out>();
  if (n > 0 || np > 0) {
    uint8_t* d_pods;
    KLM_TRY(h->agg_cols.ensure((staged ? in_bytes : o_pod) + 256));
    char* base = h->agg_cols.as<char>();
    if (staged) {
      d_pods = reinterpret_cast<uint8_t*>(base + o_pod);
      d.gpu_usage = reinterpret_cast<double*>(base + o_gu);
    } else {
      KLM_TRY(h->agg_pods.ensure((size_t)np + 256));
      KLM_TRY(h->agg_gpu.ensure((size_t)ng * 8 + 256));
      d_pods = h->agg_pods.as<uint8_t>();
      d.gpu_usage = h->agg_gpu.as<double>();
    }
    d.cpu_capacity = reinterpret_cast<int64_t*>(base);
    d.cpu_usage = reinterpret_cast<int64_t*>(base + col8);
    d.memory_capacity = reinterpret_cast<int64_t*>(base + 2 * col8);
    d.memory_usage = reinterpret_cast<int64_t*>(base + 3 * col8);
    int64_t* d_dcap = reinterpret_cast<int64_t*>(base + 4 * col8);
    int64_t* d_dalloc = reinterpret_cast<int64_t*>(base + 5 * col8);
    d.gpu_count = reinterpret_cast<int32_t*>(base + o_gc);
    d.healthy = reinterpret_cast<uint8_t*>(base + o_hl);
    d.n_gpu_usage = ng;
    // the same layout on the host side of the staged path; `put` is a memcpy there and a
    // copy submission otherwise
    if (staged) KLM_TRY(h->snap_pin.ensure(2 * kStageMax));
    auto put = [&](size_t off, void* dev, const void* src, size_t bytes) -> int {
      if (bytes == 0 || !src) return KLM_OK;
      if (staged) {
        std::memcpy(h->snap_pin.at(off), src, bytes);
        return KLM_OK;
      }
      return h2d(dev, src, bytes, st2);
    };
    if (n > 0) {
      KLM_TRY(put(0, (void*)d.cpu_capacity, c->cpu_capacity, (size_t)n * 8));
      KLM_TRY(put(col8, (void*)d.cpu_usage, c->cpu_usage, (size_t)n * 8));
      KLM_TRY(put(2 * col8, (void*)d.memory_capacity, c->memory_capacity, (size_t)n * 8));
      KLM_TRY(put(3 * col8, (void*)d.memory_usage, c->memory_usage, (size_t)n * 8));
      KLM_TRY(put(4 * col8, d_dcap, in->disk_capacity, (size_t)n * 8));
      KLM_TRY(put(5 * col8, d_dalloc, in->disk_allocatable, (size_t)n * 8));
      KLM_TRY(put(o_gc, (void*)d.gpu_count, c->gpu_count, (size_t)n * 4));
      KLM_TRY(put(o_hl, (void*)d.healthy, c->healthy, (size_t)n));
      KLM_TRY(put(o_gu, (void*)d.gpu_usage, c->gpu_usage, (size_t)ng * 8));
    }
    KLM_TRY(put(o_pod, d_pods, in->pod_running, (size_t)np));
    if (staged) KLM_TRY(h2d(base, h->snap_pin.p, in_bytes, st2));
    if (staged) {
      KLM_TRY(h->snap_out.ensure(out_bytes));
      d_cluster = h->snap_out.as<klm_cluster_out>();
    }
    if (want_rates) {
      if (staged) {
        rout = h->snap_out.as<char>() + 256;
      } else {
        KLM_TRY(h->rate_out.ensure(5 * col8));
        rout = h->rate_out.as<char>();
      }
      KLM_TRY(klm_dev_node_rates(n, d.cpu_capacity, d.cpu_usage, d.memory_capacity, d.memory_usage,
                                 in->disk_capacity ? d_dcap : nullptr,
                                 in->disk_allocatable ? d_dalloc : nullptr,
                                 reinterpret_cast<double*>(rout), reinterpret_cast<double*>(rout + col8),
                                 reinterpret_cast<int64_t*>(rout + 2 * col8),
                                 reinterpret_cast<double*>(rout + 3 * col8),
                                 reinterpret_cast<uint8_t*>(rout + 4 * col8), st2));
    }
    if (out->cluster) {
      KLM_TRY(klm_dev_cluster_aggregate(&d, n, d_pods, np, d_cluster, h->agg_scratch.p, st2));
    }
  } else if (out->cluster) {
    KLM_TRY(klm_dev_cluster_aggregate(&d, 0, nullptr, 0, d_cluster, h->agg_scratch.p, st2));
  }
  const bool staged_out = staged && (n > 0 || np > 0);
  char* hout = staged_out ? h->snap_pin.at(kStageMax) : nullptr;
  if (staged_out) {
    KLM_TRY(d2h(hout, h->snap_out.p, rout ? out_bytes : 256, st2));
  } else {
    if (out->cluster) KLM_TRY(d2h(out->cluster, d_cluster, sizeof(klm_cluster_out), st2));
    if (rout) {
      if (out->cpu_rate) KLM_TRY(d2h(out->cpu_rate, rout, (size_t)n * 8, st2));
      if (out->memory_rate) KLM_TRY(d2h(out->memory_rate, rout + col8, (size_t)n * 8, st2));
      if (out->disk_usage) KLM_TRY(d2h(out->disk_usage, rout + 2 * col8, (size_t)n * 8, st2));
      if (out->disk_rate) KLM_TRY(d2h(out->disk_rate, rout + 3 * col8, (size_t)n * 8, st2));
      if (out->under_pressure) KLM_TRY(d2h(out->under_pressure, rout + 4 * col8, (size_t)n, st2));
    }
  }
  // ---- ring branch, second half: the statistics go straight to the caller's arrays
  if (out->win_packed) {  // the device block already has the packed layout
    KLM_TRY(d2h(out->win_packed, d_min, (size_t)klm_collect_packed_bytes(S), st));
  } else if (want_win) {
    if (out->win_min) KLM_TRY(d2h(out->win_min, d_min, (size_t)S * 4, st));
    if (out->win_max) KLM_TRY(d2h(out->win_max, d_max, (size_t)S * 4, st));
    if (out->win_mean) KLM_TRY(d2h(out->win_mean, d_mean, (size_t)S * 8, st));
    if (out->win_p95) KLM_TRY(d2h(out->win_p95, d_p95, (size_t)S * 4, st));
  }
  // ---- join; the snapshot branch finishes first, so its host-side scatter overlaps K7
  KLM_TRY(sync(st2));
  if (staged_out) {
    if (out->cluster) std::memcpy(out->cluster, hout, sizeof(klm_cluster_out));
    if (rout) {
      const char* r = hout + 256;
      if (out->cpu_rate) std::memcpy(out->cpu_rate, r, (size_t)n * 8);
      if (out->memory_rate) std::memcpy(out->memory_rate, r + col8, (size_t)n * 8);
      if (out->disk_usage) std::memcpy(out->disk_usage, r + 2 * col8, (size_t)n * 8);
      if (out->disk_rate) std::memcpy(out->disk_rate, r + 3 * col8, (size_t)n * 8);
      if (out->under_pressure) std::memcpy(out->under_pressure, r + 4 * col8, (size_t)n);
    }
  }
  return sync(st);
}

}  // namespace

// ---------------------------------------------- collect cycle, two in flight

namespace {

int collect_begin_impl(klm_handle* h, const klm_collect_in* in, const klm_collect_out* out) {
  if (!in || !out || in->n_nodes < 0 || in->n_pods < 0 || (in->n_nodes > 0 && !in->nodes) ||
      (in->n_pods > 0 && !in->pod_running)) {
    set_error("klm_collect_begin: bad argument");
    return KLM_EINVAL;
  }
  if (out->win_min || out->win_max || out->win_mean || out->win_p95) {
    set_error("klm_collect_begin: window statistics are delivered through win_packed only");
    return KLM_EINVAL;
  }
  if (h->begun - h->ended >= 2) {
    set_error("klm_collect_begin: two cycles are already in flight");
    return KLM_ESTATE;
  }
  const bool want_win = out->win_packed != nullptr;
  if ((in->tick_values || want_win) && h->n_series <= 0) {
    set_error("klm_collect_begin: no ring; call klm_ring_create first");
    return KLM_ESTATE;
  }
  const int64_t n = in->n_nodes, np = in->n_pods, S = h->n_series;
  const klm_node_cols* c = in->nodes;
  const int64_t ng = (n > 0 && c) ? c->n_gpu_usage : 0;
  if (n > 0 && (!c->cpu_capacity || !c->cpu_usage || !c->memory_capacity || !c->memory_usage ||
                !c->gpu_count || !c->healthy || ng < 0 || (ng > 0 && !c->gpu_usage))) {
    set_error("klm_collect_begin: null node column");
    return KLM_EINVAL;
  }
  const size_t col8 = align256((size_t)n * 8), col4 = align256((size_t)n * 4), col1 = align256((size_t)n);
  const size_t o_gc = 6 * col8, o_hl = o_gc + col4, o_pod = o_hl + col1;
  const size_t o_gu = o_pod + align256((size_t)np), in_bytes = o_gu + align256((size_t)ng * 8);
  const size_t out_bytes = 256 + 5 * col8;
  if (in_bytes > kAsyncSnapMax || out_bytes > kAsyncSnapMax) {
    set_error("klm_collect_begin: the node snapshot exceeds the %zu-byte staging block; use klm_collect",
              kAsyncSnapMax);
    return KLM_ELIMIT;
  }
  if (!h->stream3) KLM_CUDA(cudaStreamCreateWithFlags(&h->stream3, cudaStreamNonBlocking));
  if (!h->stream_bulk) KLM_CUDA(cudaStreamCreateWithFlags(&h->stream_bulk, cudaStreamNonBlocking));
  klm_handle::CollectSlot& sl = h->slot[h->begun & 1];
  for (cudaEvent_t* ev : {&sl.up, &sl.k7, &sl.snap, &sl.win})
    if (!*ev) KLM_CUDA(cudaEventCreateWithFlags(ev, cudaEventDisableTiming));
  cudaStream_t st = h->stream_bulk, st2 = h->stream2, st3 = h->stream3;
  const size_t tick_al = align256((size_t)(S > 0 ? S : 0) * 4);
  KLM_TRY(sl.pin.ensure(tick_al + 2 * kAsyncSnapMax));
  const bool snapshot = n > 0 || np > 0 || out->cluster;
  // ---- uploads + snapshot branch on stream2
  if (in->tick_values) {
    KLM_TRY(sl.tick_dev.ensure((size_t)S * 4));
    std::memcpy(sl.pin.at(0), in->tick_values, (size_t)S * 4);
    KLM_TRY(h2d(sl.tick_dev.p, sl.pin.at(0), (size_t)S * 4, st2));
  }
  KLM_CUDA(cudaEventRecord(sl.up, st2));
  const bool want_rates = (out->cpu_rate || out->memory_rate || out->disk_usage || out->disk_rate ||
                           out->under_pressure) && n > 0;
  if (snapshot) {
    KLM_TRY(h->agg_cols.ensure(in_bytes + 256));
    KLM_TRY(h->snap_out.ensure(out_bytes));
    char* base = h->agg_cols.as<char>();
    char* hin = sl.pin.at(tick_al);
    klm_node_cols d{};
    d.cpu_capacity = reinterpret_cast<int64_t*>(base);
    d.cpu_usage = reinterpret_cast<int64_t*>(base + col8);
    d.memory_capacity = reinterpret_cast<int64_t*>(base + 2 * col8);
    d.memory_usage = reinterpret_cast<int64_t*>(base + 3 * col8);
    int64_t* d_dcap = reinterpret_cast<int64_t*>(base + 4 * col8);
    int64_t* d_dalloc = reinterpret_cast<int64_t*>(base + 5 * col8);
    d.gpu_count = reinterpret_cast<int32_t*>(base + o_gc);
    d.healthy = reinterpret_cast<uint8_t*>(base + o_hl);
    uint8_t* d_pods = reinterpret_cast<uint8_t*>(base + o_pod);
    d.gpu_usage = reinterpret_cast<double*>(base + o_gu);
    d.n_gpu_usage = ng;
    auto put = [&](size_t off, const void* src, size_t bytes) {
      if (bytes && src) std::memcpy(hin + off, src, bytes);
    };
    if (n > 0) {
      put(0, c->cpu_capacity, (size_t)n * 8);
      put(col8, c->cpu_usage, (size_t)n * 8);
      put(2 * col8, c->memory_capacity, (size_t)n * 8);
      put(3 * col8, c->memory_usage, (size_t)n * 8);
      put(4 * col8, in->disk_capacity, (size_t)n * 8);
      put(5 * col8, in->disk_allocatable, (size_t)n * 8);
      put(o_gc, c->gpu_count, (size_t)n * 4);
      put(o_hl, c->healthy, (size_t)n);
      put(o_gu, c->gpu_usage, (size_t)ng * 8);
    }
    put(o_pod, in->pod_running, (size_t)np);
    KLM_TRY(h2d(base, hin, in_bytes, st2));
    char* rout = h->snap_out.as<char>() + 256;
    if (want_rates)
      KLM_TRY(klm_dev_node_rates(n, d.cpu_capacity, d.cpu_usage, d.memory_capacity, d.memory_usage,
                                 in->disk_capacity ? d_dcap : nullptr,
                                 in->disk_allocatable ? d_dalloc : nullptr,
                                 reinterpret_cast<double*>(rout), reinterpret_cast<double*>(rout + col8),
                                 reinterpret_cast<int64_t*>(rout + 2 * col8),
                                 reinterpret_cast<double*>(rout + 3 * col8),
                                 reinterpret_cast<uint8_t*>(rout + 4 * col8), st2));
    if (out->cluster)
      KLM_TRY(klm_dev_cluster_aggregate(&d, n, d_pods, np, h->snap_out.as<klm_cluster_out>(),
                                        h->agg_scratch.p, st2));
    KLM_TRY(d2h(sl.pin.at(tick_al + kAsyncSnapMax), h->snap_out.p, want_rates ? out_bytes : 256, st2));
  }
  KLM_CUDA(cudaEventRecord(sl.snap, st2));
  // ---- K6 + K7 on the main stream, as soon as the tick is on the device
  KLM_CUDA(cudaStreamWaitEvent(st, sl.up, 0));
  // pushes that arrived since the last cycle was begun go in first, in arrival order
  if (h->n_series > 0) KLM_TRY(flush_pushed(h, (int)(h->begun & 1), st));
  if (in->tick_values)
    KLM_TRY(klm_dev_ring_ingest(h->ring.as<float>(), h->seen.as<uint32_t>(), S, h->window, h->pitch, S,
                                nullptr, sl.tick_dev.as<float>(), st));
  if (want_win) {
    const size_t c4 = align256((size_t)S * 4);
    KLM_TRY(sl.stats.ensure((size_t)klm_collect_packed_bytes(S) + 256));
    char* o = sl.stats.as<char>();
    KLM_TRY(ring_window_stats(h, 0, S, reinterpret_cast<float*>(o), reinterpret_cast<float*>(o + c4),
                              reinterpret_cast<double*>(o + 3 * c4), reinterpret_cast<float*>(o + 2 * c4), st));
  }
  KLM_CUDA(cudaEventRecord(sl.k7, st));
  // ---- the one result download, behind K7, on its own stream
  KLM_CUDA(cudaStreamWaitEvent(st3, sl.k7, 0));
  if (want_win) KLM_TRY(d2h(out->win_packed, sl.stats.p, (size_t)klm_collect_packed_bytes(S), st3));
  KLM_CUDA(cudaEventRecord(sl.win, st3));
  sl.out = *out;
  sl.n = n;
  sl.col8 = col8;
  sl.out_at = tick_al + kAsyncSnapMax;
  sl.rates = want_rates;
  sl.snapshot = snapshot;
  ++h->begun;
  return KLM_OK;
}

}  // namespace

extern "C" int klm_collect_begin(klm_handle* h, const klm_collect_in* in, const klm_collect_out* out) {
  KLM_ENTER(h);
  const int rc = collect_begin_impl(h, in, out);
  if (rc != KLM_OK && h->stream3) {  // nothing half-enqueued may outlive the failed call
    if (h->stream_bulk) cudaStreamSynchronize(h->stream_bulk);
    cudaStreamSynchronize(h->stream2);
    cudaStreamSynchronize(h->stream3);
  }
  return rc;
}

extern "C" int klm_collect_end(klm_handle* h) {
  KLM_ENTER(h);
  if (h->begun == h->ended) {
    set_error("klm_collect_end: no cycle in flight");
    return KLM_ESTATE;
  }
  klm_handle::CollectSlot& sl = h->slot[h->ended & 1];
  ++h->ended;  // the slot is free again whatever happens below
  KLM_CUDA(cudaEventSynchronize(sl.snap));
  if (sl.snapshot) {
    const char* hout = sl.pin.at(sl.out_at);
    const klm_collect_out& o = sl.out;
    const size_t n = (size_t)sl.n, col8 = sl.col8;
    if (o.cluster) std::memcpy(o.cluster, hout, sizeof(klm_cluster_out));
    if (sl.rates) {
      const char* r = hout + 256;
      if (o.cpu_rate) std::memcpy(o.cpu_rate, r, n * 8);
      if (o.memory_rate) std::memcpy(o.memory_rate, r + col8, n * 8);
      if (o.disk_usage) std::memcpy(o.disk_usage, r + 2 * col8, n * 8);
      if (o.disk_rate) std::memcpy(o.disk_rate, r + 3 * col8, n * 8);
      if (o.under_pressure) std::memcpy(o.under_pressure, r + 4 * col8, n);
    }
  }
  KLM_CUDA(cudaEventSynchronize(sl.win));
  return KLM_OK;
}

extern "C" int64_t klm_collect_packed_bytes(int64_t n_series) {
  if (n_series <= 0) return 0;
  return (int64_t)(3 * align256((size_t)n_series * 4) + (size_t)n_series * 8);
}

extern "C" int klm_collect(klm_handle* h, const klm_collect_in* in, const klm_collect_out* out) {
  KLM_ENTER(h);
  KLM_TRY(collect_idle(h, "klm_collect"));
  if (h->n_series > 0) KLM_TRY(flush_pushed(h, 0, h->stream));
  const int rc = collect_impl(h, in, out);
  if (rc != KLM_OK) {  // leave nothing in flight that still reads or writes the staging blocks
    cudaStreamSynchronize(h->stream);
    cudaStreamSynchronize(h->stream2);
  }
  return rc;
}
