/*
 * klm.h — C ABI of libklm.so: the B200 (sm_100a) data path behind
 * k8s-llm-monitor's scheduler and metrics manager.
 *
 * The reference (Sabre94/k8s-llm-monitor, pure Go) has no plugin or FFI
 * interface around this path (internal/scheduler/controller.go:174 is an
 * unexported method), so this header DEFINES the boundary a cgo shim binds.
 * Every entry point cites the reference code whose arithmetic it replaces.
 * INTEGRATION.md shows the Go side.
 *
 * Conventions
 *  - extern "C", plain pointers and sizes; no C++ or torch types.
 *  - Return 0 on success, a negative KLM_E* code otherwise; the message is
 *    in klm_last_error() (thread-local).
 *  - Host-buffer API (klm_*): the caller owns every array. Inputs are copied
 *    to device memory owned by the handle before the call returns; outputs
 *    are completely written before the call returns. No pointer is retained
 *    (cgo pointer rules).
 *  - Device-pointer API (klm_dev_*): arrays are device pointers on the
 *    current CUDA device, `stream` is a cudaStream_t passed as void*; calls
 *    only enqueue work. This is what a caller that keeps its data in HBM
 *    (bench.py's `value` leg, or a multi-process host) uses.
 *  - There is no CPU fallback anywhere: without a CUDA device klm_create
 *    fails with KLM_ENODEV and the klm_dev_* calls return the CUDA error.
 *  - Inputs must not contain NaN (JSON cannot carry it and Go's
 *    sort.SliceStable is unspecified for it).
 */
#ifndef KLM_H_
#define KLM_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KLM_OK 0
#define KLM_EINVAL (-1)  /* bad argument */
#define KLM_ECUDA (-2)   /* CUDA runtime error; text in klm_last_error() */
#define KLM_ENODEV (-3)  /* no usable CUDA device */
#define KLM_ESTATE (-4)  /* call order (e.g. schedule before nodes_set) */
#define KLM_ELIMIT (-5)  /* size beyond a documented maximum */

/* node flags (klm_nodes_set / klm_dev_*): which of the reference's per-item
 * filters the node passes. controller.go:190 and :198. */
#define KLM_NODE_HAS_NAME 0x1u  /* spec.node_name != "" */
#define KLM_NODE_STATUS_OK 0x2u /* lower(status.collection_status) in {"", "active"} */
#define KLM_NODE_ELIGIBLE 0x3u

/* klm_schedule_batch / klm_reconcile `mode` = one algorithm | one sharding mode.
 * Algorithms; both give identical results. */
#define KLM_ALGO_DECOMPOSED 0 /* node_best_reduce + O(K) probe per request */
#define KLM_ALGO_BRUTEFORCE 1 /* every request scans every node (verifier) */
/* Sharding over the devices of the handle; identical results; on a
 * one-device handle both are the same thing. The requests are split over the devices
 * in both; the modes differ in where the node table lives. */
#define KLM_SHARD_REQUESTS 0x000 /* every device holds the whole node table */
#define KLM_SHARD_NODES 0x100    /* device g holds nodes [g L, (g+1) L), L = ceil(N / G) rounded up
                                  * to 16; the others read them through peer pointers (NVLink) */

/* cluster health (manager.go:558-564) */
#define KLM_HEALTH_HEALTHY 0
#define KLM_HEALTH_WARNING 1
#define KLM_HEALTH_CRITICAL 2
/* cluster issue bits (manager.go:547-555), in the order Go appends them */
#define KLM_ISSUE_UNHEALTHY_NODES 0x1u
#define KLM_ISSUE_HIGH_CPU 0x2u
#define KLM_ISSUE_HIGH_MEMORY 0x4u

/* node condition types and statuses for klm_node_health (corev1.NodeConditionType /
 * ConditionStatus as node_metrics.go:145-163 tests them) */
#define KLM_COND_READY 0
#define KLM_COND_MEMORY_PRESSURE 1
#define KLM_COND_DISK_PRESSURE 2
#define KLM_COND_PID_PRESSURE 3
#define KLM_COND_NETWORK_UNAVAILABLE 4
#define KLM_COND_OTHER 5
#define KLM_COND_FALSE 0
#define KLM_COND_TRUE 1
#define KLM_COND_UNKNOWN 2
/* NetworkMetrics.GetQuality (types.go:187-199) */
#define KLM_QUALITY_DISCONNECTED 0
#define KLM_QUALITY_EXCELLENT 1
#define KLM_QUALITY_GOOD 2
#define KLM_QUALITY_FAIR 3
#define KLM_QUALITY_POOR 4

typedef struct klm_handle klm_handle;

/* Columns of the node snapshot that calculateClusterMetrics reads
 * (pkg/metrics/types.go:8-47; manager.go:499-535). gpu_usage is the
 * concatenation of every node's GPUUsage slice (n_gpu_usage entries). */
typedef struct klm_node_cols {
  const int64_t* cpu_capacity;    /* [N] millicores */
  const int64_t* cpu_usage;       /* [N] */
  const int64_t* memory_capacity; /* [N] bytes */
  const int64_t* memory_usage;    /* [N] */
  const int32_t* gpu_count;       /* [N] */
  const uint8_t* healthy;         /* [N] 0/1 */
  const double* gpu_usage;        /* [n_gpu_usage] percent */
  int64_t n_gpu_usage;
} klm_node_cols;

/* pkg/metrics/types.go:114-139 ClusterMetrics, numeric part. */
typedef struct klm_cluster_out {
  int64_t total_nodes;
  int64_t healthy_nodes;
  int64_t total_pods;
  int64_t running_pods;
  int64_t total_cpu;
  int64_t used_cpu;
  int64_t total_memory;
  int64_t used_memory;
  int64_t total_gpus;
  int64_t available_gpus;
  double cpu_usage_rate;
  double memory_usage_rate;
  int32_t health_status; /* KLM_HEALTH_* */
  uint32_t issues;       /* KLM_ISSUE_* bits */
} klm_cluster_out;

/* ---------------------------------------------------------------- handle */

/* A handle drives the n_dev (1..16) devices listed in device_ids (NULL, 0: device 0).
 * With more than one device -- one process driving the whole box, which is what the
 * reference's scheduler is (cmd/scheduler/main.go:55-62 constructs one controller) --
 * klm_nodes_set / klm_schedule_batch / klm_reconcile spread the batch over all of them
 * (see KLM_SHARD_*); peer access between the devices is required and enabled here. Every
 * other call of such a handle runs on device_ids[0] (metric rings shard by series with no
 * exchange: use one handle per GPU for those).
 * Thread-safe: every call takes the handle's mutex and makes its device
 * current, so cgo may enter on any OS thread. */
int klm_create(const int* device_ids, int n_dev, klm_handle** out);
void klm_destroy(klm_handle* h);
const char* klm_last_error(void);
/* ABI version of this header: major*1000 + minor. */
int klm_abi_version(void);

/* Page-locked host memory for the caller's arrays (no reference counterpart).
 * Every klm_* call accepts ordinary pageable arrays, but the CUDA driver
 * stages those synchronously; with arrays from klm_host_alloc the copies are
 * asynchronous, which is what lets klm_collect overlap its two branches and
 * klm_schedule_batch pipeline a large batch over PCIe. From Go: wrap the
 * pointer with unsafe.Slice; it is C memory, so the cgo pointer rules do not
 * restrict it. Free with klm_host_free (NULL is a no-op). */
int klm_host_alloc(size_t bytes, void** out);
int klm_host_free(void* p);

/* ------------------------------------------------------------ scheduling */

/* Install the UAV snapshot of one reconcile (controller.go:96-98): one entry
 * per UAVMetric item in list order. battery = readFloat(spec, "battery",
 * "remaining_percent") (controller.go:187, :252-272; missing -> 0).
 * The arrays are consumed before the call returns. For tables up to 256 KB
 * the call does not wait for the GPU (the table is staged in page-locked
 * memory owned by the handle), so a device fault raised by this upload is
 * reported by the next call that synchronises, klm_schedule_batch included. */
int klm_nodes_set(klm_handle* h, int64_t n_nodes, const double* battery,
                  const uint8_t* flags);

/* Replaces buildCandidates + sort.SliceStable + candidates[0]
 * (controller.go:152-161, :174-221) for a batch of R independent requests
 * against the snapshot installed by klm_nodes_set.
 *   min_battery[r]             spec.MinBatteryPercent (gate only when > 0, :194)
 *   pref_ids[pref_off[r]..pref_off[r+1])  indices of the nodes whose
 *       lower-cased name is in lower(spec.PreferredNodes) (:175-178, :206);
 *       ids outside [0, n_nodes) are ignored.
 *   out_idx[r]   index of the chosen node, -1 when no candidate (:153-158)
 *   out_score[r] chosen.Score (:167); 0 when out_idx[r] == -1
 * Ties go to the lowest index (stable sort, :160). */
int klm_schedule_batch(klm_handle* h, int64_t n_requests,
                       const double* min_battery, const int32_t* pref_off,
                       const int32_t* pref_ids, int32_t* out_idx,
                       double* out_score, int mode);

/* The k best candidates of every request, in the order the reference's
 * stable descending sort leaves them (controller.go:160: score desc, list index asc); the
 * reference keeps only [0] (:161), so k = 1 gives exactly klm_schedule_batch and anything
 * beyond has no reference semantics (examples/multi-pod-request.yaml asks for two GPUs nobody
 * assigns). out_idx / out_score are [n_requests][k], padded with -1 / 0 when a request has fewer
 * than k candidates. 1 <= k <= klm_topk_max() (8). One-device handles. */
int klm_topk_max(void);
int klm_schedule_batch_topk(klm_handle* h, int64_t n_requests,
                            const double* min_battery, const int32_t* pref_off,
                            const int32_t* pref_ids, int k, int32_t* out_idx,
                            double* out_score);

/* One reconcile in one call (controller.go:88-110: one UAV snapshot, then every
 * pending request): klm_nodes_set + klm_schedule_batch with the same arguments,
 * results and errors, under one lock. A reconcile whose arrays fit the 256 KB
 * staging block costs one upload, K1, K2 overlapping K1, one download and one
 * synchronisation. The snapshot stays installed for later klm_schedule_batch calls. */
int klm_reconcile(klm_handle* h, int64_t n_nodes, const double* battery,
                  const uint8_t* flags, int64_t n_requests,
                  const double* min_battery, const int32_t* pref_off,
                  const int32_t* pref_ids, int32_t* out_idx, double* out_score,
                  int mode);

/* ----------------------------------------------------- cluster aggregate */

/* Replaces Manager.calculateClusterMetrics (manager.go:493-565).
 * pod_running[p] = (Phase == "Running") (:509). */
int klm_cluster_aggregate(klm_handle* h, const klm_node_cols* nodes,
                          int64_t n_nodes, const uint8_t* pod_running,
                          int64_t n_pods, klm_cluster_out* out);

/* Per-node arithmetic of buildNodeMetrics (sources/node_metrics.go:116-140)
 * and NodeMetrics.IsUnderPressure (pkg/metrics/types.go:159-162).
 * disk_allocatable may be NULL (then disk_usage = 0, node_metrics.go:117).
 * Outputs (any may be NULL): rates in percent, disk_usage bytes,
 * under_pressure 0/1. */
int klm_node_rates(klm_handle* h, int64_t n_nodes, const int64_t* cpu_capacity,
                   const int64_t* cpu_usage, const int64_t* memory_capacity,
                   const int64_t* memory_usage, const int64_t* disk_capacity,
                   const int64_t* disk_allocatable, double* cpu_rate,
                   double* memory_rate, int64_t* disk_usage, double* disk_rate,
                   uint8_t* under_pressure);

/* Per-pod arithmetic of buildPodMetrics (sources/pod_metrics.go:104-171):
 * containers of pod p are rows ctr_off[p]..ctr_off[p+1] of the six container
 * columns. Outputs: pod_totals[6][P] in the order cpu_usage, memory_usage,
 * cpu_request, cpu_limit, memory_request, memory_limit; rates vs limit
 * (:162-171); utilisation vs request (types.go:165-173); over_limit
 * (types.go:176-184, with its int64(float64(limit)*0.9) truncation). */
int klm_pod_rates(klm_handle* h, int64_t n_pods, const int32_t* ctr_off,
                  const int64_t* ctr_cpu_usage, const int64_t* ctr_memory_usage,
                  const int64_t* ctr_cpu_request, const int64_t* ctr_cpu_limit,
                  const int64_t* ctr_memory_request,
                  const int64_t* ctr_memory_limit, int64_t* pod_totals,
                  double* cpu_rate, double* memory_rate, double* cpu_util,
                  double* memory_util, uint8_t* over_limit);

/* NodeMetrics.GetAvailableResources (pkg/metrics/types.go:151-156): free CPU cores
 * (millicores / 1000), memory and disk in GiB (/ 1024 / 1024 / 1024, left to right).
 * disk_capacity / disk_usage may be NULL (0); any output may be NULL. */
int klm_node_available(klm_handle* h, int64_t n_nodes, const int64_t* cpu_capacity,
                       const int64_t* cpu_usage, const int64_t* memory_capacity,
                       const int64_t* memory_usage, const int64_t* disk_capacity,
                       const int64_t* disk_usage, double* cpu_cores, double* memory_gb,
                       double* disk_gb);
/* NetworkMetrics.GetQuality (pkg/metrics/types.go:187-199): KLM_QUALITY_* per probe
 * from Connected and RTT (ms): < 10 excellent, < 50 good, < 100 fair, else poor. */
int klm_network_quality(klm_handle* h, int64_t n, const uint8_t* connected,
                        const double* rtt_ms, uint8_t* quality);
/* Pod restart count (sources/pod_metrics.go:174-177): the sum of RestartCount over the
 * pod's container statuses, rows st_off[p]..st_off[p+1] of restart_count; int32 that
 * wraps as Go's does. */
int klm_pod_restarts(klm_handle* h, int64_t n_pods, const int32_t* st_off,
                     const int32_t* restart_count, int32_t* restarts);
/* Node health from conditions (sources/node_metrics.go:142-164): node i's conditions are
 * rows cond_off[i]..cond_off[i+1] of cond_type (KLM_COND_*) / cond_status. healthy[i] = 0
 * if Ready is not True or a pressure / network-unavailable condition is True; failing[i]
 * has bit t set for every condition type t that did it (the host formats the
 * "NotReady: ..." strings from it). Either output may be NULL. */
int klm_node_health(klm_handle* h, int64_t n_nodes, const int32_t* cond_off,
                    const uint8_t* cond_type, const uint8_t* cond_status,
                    uint8_t* healthy, uint8_t* failing);

/* Fleet scans of sources/uav_metrics.go:197-233: number of UAVs whose
 * Health.SystemStatus == "OK" (status_ok[i] != 0), and the indices, in
 * ascending order, of UAVs with RemainingPercent < threshold. low_idx has
 * room for n entries; *n_low receives how many were written. */
int klm_fleet_scan(klm_handle* h, int64_t n, const uint8_t* status_ok,
                   const double* remaining_percent, double threshold,
                   int64_t* healthy_count, int32_t* low_idx, int64_t* n_low);

/* --------------------------------------------------- metric window rings */

/* Allocate the metric ring: n_series rows of `window` f32 samples, one row
 * per (node|pod|UAV, channel) series. The reference keeps only the latest
 * snapshot (manager.go:29, :289-290) and lists time-series storage as future
 * work, so the ring and its statistics have no reference implementation:
 * their definition is in DESIGN.md and their parity is "unpinned". */
int klm_ring_create(klm_handle* h, int64_t n_series, int32_t window);
/* Append one sample to each listed series (UpdateUAVReport, manager.go:391-449,
 * and the end of Collect, :286, are the producers). series_ids must be
 * distinct within one call. While klm_collect_begin cycles are in flight the call
 * does not fail: the samples are copied, queued in arrival order and applied by the
 * next klm_collect_begin before its own tick (or by the next synchronous ring call
 * once the pipeline has drained), which is the order manager.go:437-446 gets from
 * its mutex. klm_ring_ingest_tick queues the same way. */
int klm_ring_ingest(klm_handle* h, int64_t n, const int64_t* series_ids,
                    const float* values);
/* Append one sample to EVERY series: values[n_series] (one collection tick). */
int klm_ring_ingest_tick(klm_handle* h, const float* values);
/* Bulk-load history: rows [s0, s1) each get `count` (<= window) samples from
 * data[(s - s0) * count + t], replacing what the rows held. */
int klm_ring_load(klm_handle* h, int64_t s0, int64_t s1, int32_t count,
                  const float* data);
/* Window statistics of series [s0, s1) over the last n = min(window, seen)
 * samples of each: min, max (exact f32), mean (f64 sum / n), p95 =
 * nearest-rank sorted[ceil(0.95 n) - 1] (an input element). Series with no
 * sample give min = max = p95 = 0 and mean = 0. Any output may be NULL. */
int klm_ring_stats(klm_handle* h, int64_t s0, int64_t s1, float* out_min,
                   float* out_max, double* out_mean, float* out_p95);

/* One collection cycle in one call — Manager.Collect (manager.go:195-319): append the
 * tick to the ring (if tick_values != NULL), compute the window statistics of every
 * series, the per-node rates (node_metrics.go:116-140) and calculateClusterMetrics
 * (manager.go:493-565) from the same node columns. Everything is enqueued on the
 * handle's stream and synchronised once, so a cycle costs one host round trip
 * instead of four. Any output pointer may be NULL; disk_capacity / disk_allocatable
 * may be NULL (rates then use 0). Results are identical to the separate calls. */
typedef struct klm_collect_in {
  const float* tick_values;        /* [n_series] or NULL: no ingest this cycle */
  const klm_node_cols* nodes;      /* may be NULL with n_nodes == 0 */
  int64_t n_nodes;
  const int64_t* disk_capacity;    /* [n_nodes] or NULL */
  const int64_t* disk_allocatable; /* [n_nodes] or NULL */
  const uint8_t* pod_running;      /* [n_pods] */
  int64_t n_pods;
} klm_collect_in;
typedef struct klm_collect_out {
  klm_cluster_out* cluster;
  double* cpu_rate;      /* [n_nodes] */
  double* memory_rate;   /* [n_nodes] */
  int64_t* disk_usage;   /* [n_nodes] */
  double* disk_rate;     /* [n_nodes] */
  uint8_t* under_pressure; /* [n_nodes] */
  float* win_min;        /* [n_series] */
  float* win_max;
  double* win_mean;
  float* win_p95;
  /* Optional: all four window statistics as ONE block of klm_collect_packed_bytes(n_series)
   * bytes, fetched with a single device-to-host copy (each separate array above costs its
   * own copy, ~6 us of fixed latency): with c = n_series * 4 rounded up to 256,
   * min f32[n_series] at byte 0, max at c, p95 at 2 c, mean f64[n_series] at 3 c.
   * When non-NULL the four win_* pointers above must be NULL. */
  void* win_packed;
} klm_collect_out;
int klm_collect(klm_handle* h, const klm_collect_in* in, const klm_collect_out* out);
/* size of the win_packed block for n_series series (0 if n_series <= 0) */
int64_t klm_collect_packed_bytes(int64_t n_series);

/* The same cycle, asynchronously, with up to TWO cycles in flight (no reference
 * counterpart): klm_collect_begin enqueues one cycle and returns; klm_collect_end
 * completes the OLDEST cycle begun. The tick upload and the result download of one
 * cycle then overlap the window statistics of the next (three streams, everything
 * two cycles could both touch exists once per cycle), so a collector loop
 *     begin(0); for (i = 0; ; ++i) { begin(i + 1); end(); use results of cycle i; }
 * runs at the pace of the kernels instead of kernels + copies.
 * - `in` arrays are consumed by klm_collect_begin (staged in page-locked memory).
 * - `out` arrays must stay valid until the matching klm_collect_end; their contents
 *   are defined only after it returns. Window statistics come through win_packed
 *   only (win_min/max/mean/p95 must be NULL).
 * - The node snapshot must fit a 1 MB staging block (about 20 000 nodes plus
 *   their pods), else KLM_ELIMIT: use klm_collect.
 * - While cycles are in flight klm_ring_ingest / klm_ring_ingest_tick queue their
 *   samples for the next klm_collect_begin (HTTP handlers push UAV reports at any
 *   time, cmd/server/main.go:607); the other synchronous ring / collect / aggregate
 *   calls on this handle return KLM_ESTATE. A third klm_collect_begin returns KLM_ESTATE. */
int klm_collect_begin(klm_handle* h, const klm_collect_in* in, const klm_collect_out* out);
int klm_collect_end(klm_handle* h);

/* ================================================= device-pointer layer */
/* All pointers below are DEVICE pointers; nothing is allocated or freed;
 * work is enqueued on `stream` and the call returns without synchronising.
 * `scratch` arguments are caller-provided device memory of at least the
 * stated size, 256-byte aligned, zero-filled ONCE after allocation (they hold
 * self-resetting tickets); a kernel leaves its scratch ready for the next
 * launch. Two launches that share a scratch area must not overlap. */

/* Bytes of scratch the two scheduling kernels need. */
int64_t klm_dev_node_best_scratch_bytes(void);
/* K1 node_best_reduce: (battery, index) of the lowest-index node with the
 * largest battery among nodes with (flags & 3) == 3; best[0] = index as
 * int64 (-1 if none), best[1] = bits of the f64 battery. controller.go:190,
 * :198 (filters) + :160-161 (first maximum). node_base is added to the
 * index (node-range sharding). */
int klm_dev_node_best(const double* battery, const uint8_t* flags,
                      int64_t n_nodes, int64_t node_base, int64_t* best,
                      void* scratch, void* stream);
/* K1 without caller scratch, for hosts that must not allocate device memory (a
 * process whose HBM is fully reserved by a pool or an MPS memory limit, or a
 * reconcile loop that forbids cudaMalloc because it synchronises the device).
 * No counterpart in the reference (it has no GPU path and no allocation contract).
 * Same result as klm_dev_node_best. The scratch is a 19 200-byte __device__ array
 * inside libklm.so, held by every GPU that loads the library until exit; it
 * limits one call to klm_dev_node_best_static_max_nodes() = 4 849 664 nodes
 * (larger tables: KLM_ELIMIT). All calls on a device share that array, so calls
 * that could overlap on the GPU race: serialise them. The first call on a device
 * looks up the array's address; make it before any stream capture. The address is kept
 * for the life of the process and does not survive cudaDeviceReset(). Never calls
 * cudaMalloc / cudaFree. */
int64_t klm_dev_node_best_static_max_nodes(void);
int klm_dev_node_best_static(const double* battery, const uint8_t* flags,
                             int64_t n_nodes, int64_t node_base, int64_t* best,
                             void* stream);
/* K2 schedule_batch (decomposed form of controller.go:174-221 + :160-161).
 * Node arrays hold global indices [node_base, node_base + n_nodes);
 * pref_ids are global; ids outside the local range are skipped, so with
 * node-range sharding each rank yields its partial winner and
 * klm_dev_merge_partials finishes the job. */
int klm_dev_schedule(const double* battery, const uint8_t* flags,
                     int64_t n_nodes, int64_t node_base, const int64_t* best,
                     int64_t n_requests, const double* min_battery,
                     const int32_t* pref_off, const int32_t* pref_ids,
                     int32_t* out_idx, double* out_score, void* stream);
/* Top-k forms of K1 / K2 (klm_schedule_batch_topk): top = the k best eligible nodes of the
 * table as k (index int64, battery bits) pairs; scratch as for klm_dev_node_best
 * (klm_dev_node_topk_scratch_bytes() bytes, zero-filled once). */
int64_t klm_dev_node_topk_scratch_bytes(void);
int klm_dev_node_topk(const double* battery, const uint8_t* flags, int64_t n_nodes,
                      int64_t node_base, int k, int64_t* top, void* scratch, void* stream);
int klm_dev_schedule_topk(const double* battery, const uint8_t* flags, int64_t n_nodes,
                          int64_t node_base, const int64_t* top, int k, int64_t n_requests,
                          const double* min_battery, const int32_t* pref_off,
                          const int32_t* pref_ids, int32_t* out_idx, double* out_score,
                          void* stream);
/* K2 launched to overlap the kernel before it on `stream` (no reference
 * counterpart; same result as klm_dev_schedule). Meant for the pair
 * klm_dev_node_best -> this call: K2 starts while K1 is still running, issues
 * all of its loads, and waits for K1 only before it reads `best` (programmatic
 * dependent launch). CONTRACT: if the operation enqueued on `stream` just
 * before this call is a kernel, it must write nothing this call reads except
 * `best` -- the node columns and the request arrays must have been complete
 * before that kernel started. Otherwise use klm_dev_schedule. */
int klm_dev_schedule_after_best(const double* battery, const uint8_t* flags,
                                int64_t n_nodes, int64_t node_base,
                                const int64_t* best, int64_t n_requests,
                                const double* min_battery, const int32_t* pref_off,
                                const int32_t* pref_ids, int32_t* out_idx,
                                double* out_score, void* stream);
/* K1 and K2 as ONE launch (no reference counterpart; same results as
 * klm_dev_node_best followed by klm_dev_schedule on the whole table). Every
 * CTA of a one-wave grid scans its share of the node table, the CTAs meet on
 * a counter in `scratch`, each folds the partial winners and decides its
 * requests: one cold start instead of two and no block -> ticket -> last
 * block chain in front of the first decision. `best` (2 x int64, as
 * klm_dev_node_best writes it) may be NULL. `scratch`: at least
 * klm_dev_node_best_scratch_bytes() bytes, zero-filled once; overlapping
 * calls need separate scratch. Launched cooperatively: the grid never exceeds
 * what the device holds at once, and a device without cooperative launch
 * returns KLM_ECUDA (callers then take the two-kernel path). */
int klm_dev_schedule_fused(const double* battery, const uint8_t* flags, int64_t n_nodes,
                           int64_t node_base, int64_t* best, int64_t n_requests,
                           const double* min_battery, const int32_t* pref_off,
                           const int32_t* pref_ids, int32_t* out_idx, double* out_score,
                           void* scratch, void* stream);
/* K2b schedule_bruteforce: same result, every request scans every node
 * through shared-memory tiles — the literal loop of controller.go:181-218. */
int klm_dev_schedule_bruteforce(const double* battery, const uint8_t* flags,
                                int64_t n_nodes, int64_t node_base,
                                int64_t n_requests, const double* min_battery,
                                const int32_t* pref_off,
                                const int32_t* pref_ids, int32_t* out_idx,
                                double* out_score, void* stream);
/* K3 merge_partials: part_idx/part_score are [n_parts][n_requests] partial
 * winners (global indices, -1 = none) as gathered from the ranks; picks the
 * highest score, lowest global index on ties (controller.go:160). */
int klm_dev_merge_partials(int n_parts, int64_t n_requests,
                           const int32_t* part_idx, const double* part_score,
                           int32_t* out_idx, double* out_score, void* stream);

/* ---- node-range sharding without an all-gather (no reference counterpart).
 * Rank g of G owns nodes [lo, hi) = klm_shard_range(N, G, g) and one array of exchange slots
 * (klm_dev_shard_slots_bytes() bytes, zero-filled once) that every rank can WRITE: in one
 * process after cudaDeviceEnablePeerAccess, across processes through CUDA IPC or symmetric
 * memory. Per snapshot, with a fresh epoch >= 1 that all ranks agree on and that grows by one:
 *   klm_dev_node_best_publish  K1 over the rank's shard; the 16-byte result goes into the
 *                              rank's slot of EVERY rank's array (release store, system scope);
 *   klm_dev_schedule_sharded   K2 over the rank's slice of the REQUESTS: waits (acquire) until
 *                              its own slot array holds all G results of the epoch, folds them,
 *                              and probes preferred nodes where they live -- battery_parts[q] /
 *                              flags_parts[q] are rank q's shard as seen from this device (peer
 *                              pointers; or slices of a local replica: request-shard mode).
 *                              table_from_k1 != 0: the columns are filled by this epoch's K1s
 *                              (klm_dev_node_best_spread), so they are read only after the
 *                              wait; 0: they were complete before K1 started and the first
 *                              probes are issued ahead of the wait.
 * No host barrier, no gather of partial winners, no merge kernel: the exchange is 16 bytes per
 * rank pair and per-rank work is N / G + R / G. Results equal klm_dev_schedule on the whole
 * table (global indices, lowest index on ties: controller.go:160). Both kernels of a rank go
 * on one stream, K1 first; every rank must launch its K1 (a K2 that waits 4 s for a missing
 * peer gives up: klm_dev_shard_error then reports 1 and the affected results are void). Epochs
 * alternate between two sets of slots, so a rank may publish epoch e + 1 while a slower rank
 * still reads e; the node columns themselves must not be overwritten until every rank's K2 of
 * the epoch has finished (double-buffer them, or synchronise once per snapshot). Never run
 * ranks that wait for each other as separate launches on ONE GPU unless every K1 is enqueued
 * before any K2 on the same stream. */
int64_t klm_dev_shard_slots_bytes(void);
void klm_shard_range(int64_t n_nodes, int n_parts, int part, int64_t* lo, int64_t* hi);
int klm_dev_node_best_publish(const double* battery, const uint8_t* flags,
                              int64_t n_nodes, int64_t node_base, int n_parts,
                              int my_part, void* const* peer_slots /* host array */,
                              uint64_t epoch, int64_t* best /* may be NULL */,
                              void* scratch, void* stream);
/* klm_dev_node_best_publish that also pushes the shard into every peer's REPLICA of the node
 * table while it scans it (an all-gather fused into K1: plain stores through peer pointers, the
 * shard crosses NVLink once per peer, in full lines). peer_battery[q] / peer_flags[q] are the
 * BASES of rank q's replica as seen from this device (entry my_part is not used: the rank's own
 * shard already is that part of its replica); the shard lands at offset node_base. A
 * klm_dev_schedule_sharded whose parts are slices of the local replica then probes locally: the
 * right trade once a batch probes more than about 0.08 G N preferred nodes -- a remote probe
 * costs two 32-byte NVLink packets, about 11 G/s per GPU measured, against a bulk copy at several
 * hundred GB/s; for a handful of requests klm_dev_node_best_publish + peer probes moves less.
 * The replicas must not be rewritten before every rank's K2 of the epoch has finished. */
int klm_dev_node_best_spread(const double* battery, const uint8_t* flags,
                             int64_t n_nodes, int64_t node_base, int n_parts,
                             int my_part, void* const* peer_slots,
                             double* const* peer_battery, uint8_t* const* peer_flags,
                             uint64_t epoch, int64_t* best, void* scratch, void* stream);
int klm_dev_schedule_sharded(int n_parts, int64_t n_nodes_total,
                             const double* const* battery_parts /* host array */,
                             const uint8_t* const* flags_parts /* host array */,
                             void* my_slots, uint64_t epoch, int64_t n_requests,
                             const double* min_battery, const int32_t* pref_off,
                             const int32_t* pref_ids, int32_t* out_idx,
                             double* out_score, int table_from_k1, void* stream);
/* Line the ranks up on their streams (no reference counterpart): publishes `epoch` into this
 * rank's slot of every rank's array and waits until its own array carries it from all
 * n_parts ranks. Needs slot arrays of its own (not the ones K1 publishes into), zero-filled
 * once, epochs growing by one. The data path needs no barrier; this is for a caller that wants
 * its ranks to start a region together (bench.py does, in front of every timed step). Every
 * rank must call it, each on its own GPU. */
int klm_dev_shard_rendezvous(int n_parts, int my_part, void* const* peer_slots,
                             void* my_slots, uint64_t epoch, void* stream);
/* *out = 1 if a klm_dev_schedule_sharded on these slots gave up waiting; synchronises `stream`. */
int klm_dev_shard_error(const void* my_slots, int* out, void* stream);

/* K3 over peer memory: part_idx[g] / part_score[g] (HOST arrays of n_parts DEVICE
 * pointers, at most 16) are rank g's partial winners for ALL requests, readable from this
 * device (its own memory, or a peer's mapped over NVLink / NVSwitch: CUDA IPC, symmetric
 * memory). Merges requests [r0, r1) only and writes out[r - r0]: with one slice per rank
 * the exchange and the merge are a single kernel and nothing is staged through an
 * all-gather. The caller orders it after every rank's K2 (a cross-rank barrier). */
int klm_dev_merge_partials_peer(int n_parts, const int32_t* const* part_idx,
                                const double* const* part_score, int64_t r0,
                                int64_t r1, int32_t* out_idx, double* out_score,
                                void* stream);

int64_t klm_dev_cluster_scratch_bytes(void);
/* K4 cluster_aggregate (manager.go:493-565). `out` is a device
 * klm_cluster_out. node_cols members are device pointers. */
int klm_dev_cluster_aggregate(const klm_node_cols* nodes, int64_t n_nodes,
                              const uint8_t* pod_running, int64_t n_pods,
                              klm_cluster_out* out, void* scratch,
                              void* stream);
/* Partial sums only (for node/pod sharding across ranks): writes the ten
 * int64 counters of klm_cluster_out to sums[10]; klm_dev_cluster_finish
 * turns (all-reduced) sums into rates and health. */
int klm_dev_cluster_sums(const klm_node_cols* nodes, int64_t n_nodes,
                         const uint8_t* pod_running, int64_t n_pods,
                         int64_t* sums, void* scratch, void* stream);
int klm_dev_cluster_finish(const int64_t* sums, klm_cluster_out* out,
                           void* stream);

/* K5 element-wise rates; same arithmetic as klm_node_rates/klm_pod_rates. */
int klm_dev_node_rates(int64_t n_nodes, const int64_t* cpu_capacity,
                       const int64_t* cpu_usage, const int64_t* memory_capacity,
                       const int64_t* memory_usage,
                       const int64_t* disk_capacity,
                       const int64_t* disk_allocatable, double* cpu_rate,
                       double* memory_rate, int64_t* disk_usage,
                       double* disk_rate, uint8_t* under_pressure,
                       void* stream);
int klm_dev_pod_rates(int64_t n_pods, const int32_t* ctr_off,
                      const int64_t* ctr_cpu_usage,
                      const int64_t* ctr_memory_usage,
                      const int64_t* ctr_cpu_request,
                      const int64_t* ctr_cpu_limit,
                      const int64_t* ctr_memory_request,
                      const int64_t* ctr_memory_limit, int64_t* pod_totals,
                      double* cpu_rate, double* memory_rate, double* cpu_util,
                      double* memory_util, uint8_t* over_limit, void* stream);
/* Window-derived node features (NO reference semantics -- the reference
 * parses last_update and drops it, controller.go:202-203, and scores the latest battery only;
 * opt-in, parity unpinned). Turns the node table of a snapshot and the window statistics of
 * the UAV ring (series node * 64 + channel; channel 21 remaining_percent, 24 battery
 * temperature) into the table the scheduler then sees:
 *   heartbeat_age[i] > max_age_s        -> status bit cleared (a stale node is no candidate)
 *   p95(temperature) > temp_limit       -> status bit cleared
 *   battery_out[i] = clamp(battery[i] - horizon_s * (max - min of remaining_percent) / span_s,
 *                          0, battery[i])   (the charge projected horizon_s ahead at the window's drain)
 * A parameter <= 0, or a NULL input array, switches its rule off. Feed battery_out / flags_out
 * to klm_dev_node_best / klm_dev_schedule (or copy them out for klm_nodes_set). */
typedef struct klm_feature_params {
  double max_age_s;   /* staleness limit for heartbeat_age, seconds */
  double temp_limit;  /* limit on the window's p95 battery temperature */
  double horizon_s;   /* how far ahead the battery is projected */
  double span_s;      /* time the window spans (window * sampling period) */
} klm_feature_params;
int klm_dev_node_features(int64_t n_nodes, const double* battery, const uint8_t* flags,
                          const double* heartbeat_age, const float* win_min,
                          const float* win_max, const float* win_p95,
                          const klm_feature_params* params, double* battery_out,
                          uint8_t* flags_out, void* stream);
/* Device-pointer forms of klm_node_available / klm_network_quality / klm_pod_restarts /
 * klm_node_health; same arithmetic. */
int klm_dev_node_available(int64_t n_nodes, const int64_t* cpu_capacity,
                           const int64_t* cpu_usage, const int64_t* memory_capacity,
                           const int64_t* memory_usage, const int64_t* disk_capacity,
                           const int64_t* disk_usage, double* cpu_cores,
                           double* memory_gb, double* disk_gb, void* stream);
int klm_dev_network_quality(int64_t n, const uint8_t* connected, const double* rtt_ms,
                            uint8_t* quality, void* stream);
int klm_dev_pod_restarts(int64_t n_pods, const int32_t* st_off,
                         const int32_t* restart_count, int32_t* restarts, void* stream);
int klm_dev_node_health(int64_t n_nodes, const int32_t* cond_off, const uint8_t* cond_type,
                        const uint8_t* cond_status, uint8_t* healthy, uint8_t* failing,
                        void* stream);
/* counters[0] = healthy count, counters[1] = n_low; both are overwritten by
 * the call. low_idx is written in ascending order. */
int64_t klm_dev_fleet_scan_scratch_bytes(int64_t n);
int klm_dev_fleet_scan(int64_t n, const uint8_t* status_ok,
                       const double* remaining_percent, double threshold,
                       int64_t* counters, int32_t* low_idx, void* scratch,
                       void* stream);

/* Ring layout: ring[s * pitch + slot], pitch a multiple of 4 floats and the
 * base 16-byte aligned; seen[s] = samples ever appended to series s (u32,
 * saturating); the valid slots of a row are [0, min(seen, window)). */
/* K6 window_ingest: scatter (series_ids, values), or one value per series
 * when series_ids == NULL (then n == n_series). */
int klm_dev_ring_ingest(float* ring, uint32_t* seen, int64_t n_series,
                        int32_t window, int32_t pitch, int64_t n,
                        const int64_t* series_ids, const float* values,
                        void* stream);
/* K7 window_stats over series [s0, s1). Outputs are indexed from s0
 * (out[s - s0]); any may be NULL. */
int klm_dev_window_stats(const float* ring, const uint32_t* seen,
                         int32_t window, int32_t pitch, int64_t s0, int64_t s1,
                         float* out_min, float* out_max, double* out_mean,
                         float* out_p95, void* stream);
/* The streaming variant of K7: one warp per series, the window is not kept on chip (rows pass
 * through two small register buffers), so 24 warps fit an SM; any window size; same results. */
int klm_dev_window_stats_stream(const float* ring, const uint32_t* seen,
                                int32_t window, int32_t pitch, int64_t s0,
                                int64_t s1, float* out_min, float* out_max,
                                double* out_mean, float* out_p95, void* stream);
/* K7 for streaming windows (no reference counterpart): the streaming kernel with a per-series
 * bracket around the p95 carried from call to call in `state` (2 floats per series, indexed from
 * s0 like the outputs; fill it with 0xFF bytes -- NaN -- before the first call). Between two
 * collection ticks one sample enters a window and one leaves, so the previous bracket still
 * holds the answer: the kernel skips the pivot sampling, lists about k + 12 samples instead of
 * 10 % of the window, and finishes with one counting pass and one sort. The bracket is only a
 * pivot: results are exact and identical to klm_dev_window_stats whatever `state` contains
 * (stale, another ring's, garbage); a useless bracket costs the time of the unhinted kernel.
 * Windows up to 8192 (KLM_ELIMIT beyond). Calls on the same series must not overlap. */
int klm_dev_window_stats_hinted(const float* ring, const uint32_t* seen,
                                int32_t window, int32_t pitch, int64_t s0,
                                int64_t s1, float* out_min, float* out_max,
                                double* out_mean, float* out_p95, float* state,
                                void* stream);
/* The register-resident (one warp per series) variant of K7, which
 * klm_dev_window_stats uses for windows up to 4096 (KLM_ELIMIT beyond). */
int klm_dev_window_stats_warp(const float* ring, const uint32_t* seen,
                              int32_t window, int32_t pitch, int64_t s0,
                              int64_t s1, float* out_min, float* out_max,
                              double* out_mean, float* out_p95, void* stream);
/* The shared-memory (one CTA per series) variant of K7 that
 * klm_dev_window_stats uses for windows above 4096 samples, callable directly
 * for any window up to 51200; same results. */
int klm_dev_window_stats_block(const float* ring, const uint32_t* seen,
                               int32_t window, int32_t pitch, int64_t s0,
                               int64_t s1, float* out_min, float* out_max,
                               double* out_mean, float* out_p95, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* KLM_H_ */
