/*
 * klm_oracle.c — CPU restatement of k8s-llm-monitor's scheduling and metric
 * aggregation arithmetic. TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library, and only as the checker or the
 * timed CPU baseline. Nothing under k8s-llm-monitor_b200/ links or calls it.
 *
 * PARITY STATUS
 *  - The reference is pure Go with no _test.go files, no golden vectors and no
 *    Go toolchain on this machine (its k8s.io dependencies are not vendored
 *    either), so this oracle cannot be checked against the reference running.
 *    Scheduling / cluster-aggregate / rate functions are line-by-line
 *    restatements pinned by hand-derived known-answer tests taken from the
 *    reference's own fixtures (the three YAML requests under examples/ and deployments/uav-simulator.yaml;
 *    see tests/test_oracle_kat.py): "derived, not reference-pinned".
 *  - Window statistics (klmo_window_stats*) have NO reference implementation
 *    (the reference keeps only the latest snapshot, manager.go:29,289-290):
 *    PARITY UNPINNED. Their definition is this file's.
 *
 * Reference paths are relative to the reference repository root.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define KLMO_API __attribute__((visibility("default")))

/* ------------------------------------------------------------------ */
/* Scheduling: literal form.                                           */
/* internal/scheduler/controller.go:174-221 buildCandidates, then      */
/* :160 sort.SliceStable(desc by Score), :161 candidates[0].           */
/* ------------------------------------------------------------------ */

typedef struct {
  int32_t idx;
  double battery;
  double score;
} klmo_candidate;

/* Stable merge sort, descending by score: less(i, j) = score[i] > score[j]
 * exactly as the closure at controller.go:160. Equal scores keep list order. */
static void merge_sort_desc(klmo_candidate* a, klmo_candidate* tmp, int64_t n) {
  if (n < 2) return;
  int64_t h = n / 2;
  merge_sort_desc(a, tmp, h);
  merge_sort_desc(a + h, tmp, n - h);
  int64_t i = 0, j = h, k = 0;
  while (i < h && j < n) {
    /* take from the right run only when it is strictly "less" (greater score) */
    if (a[j].score > a[i].score)
      tmp[k++] = a[j++];
    else
      tmp[k++] = a[i++];
  }
  while (i < h) tmp[k++] = a[i++];
  while (j < n) tmp[k++] = a[j++];
  memcpy(a, tmp, (size_t)n * sizeof(*a));
}

/* One request against the whole snapshot, the way the reference does it:
 * build the candidate list in item order, stable-sort it, take [0].
 * pref_mask[i] != 0 <=> lower(node_name[i]) is in the preferred set
 * (controller.go:175-178, :206). Returns the number of candidates. */
static int64_t schedule_one_literal(int64_t n, const double* battery,
                                    const uint8_t* flags,
                                    const uint8_t* pref_mask, double min_battery,
                                    klmo_candidate* cand, klmo_candidate* tmp,
                                    int32_t* out_idx, double* out_score) {
  int64_t m = 0;
  for (int64_t i = 0; i < n; ++i) {
    if (!(flags[i] & 1u)) continue;                              /* :190 */
    if (min_battery > 0 && battery[i] < min_battery) continue;   /* :194 */
    if (!(flags[i] & 2u)) continue;                              /* :198 */
    double score = battery[i];                                   /* :205 */
    if (pref_mask[i]) score += 10;                               /* :206-208 */
    cand[m].idx = (int32_t)i;
    cand[m].battery = battery[i];
    cand[m].score = score;
    ++m;
  }
  if (m == 0) { /* :153-158 */
    *out_idx = -1;
    *out_score = 0.0;
    return 0;
  }
  merge_sort_desc(cand, tmp, m); /* :160 */
  *out_idx = cand[0].idx;        /* :161 */
  *out_score = cand[0].score;
  return m;
}

/* Batch of requests, literal algorithm. pref_ids outside [0, n) are ignored
 * (a preferred name that matches no node). Returns 0, or -1 on allocation
 * failure. */
KLMO_API int klmo_schedule_literal(int64_t n, const double* battery,
                                   const uint8_t* flags, int64_t n_requests,
                                   const double* min_battery,
                                   const int32_t* pref_off,
                                   const int32_t* pref_ids, int32_t* out_idx,
                                   double* out_score) {
  klmo_candidate* cand = malloc((size_t)(n > 0 ? n : 1) * sizeof(*cand));
  klmo_candidate* tmp = malloc((size_t)(n > 0 ? n : 1) * sizeof(*tmp));
  uint8_t* mask = calloc((size_t)(n > 0 ? n : 1), 1);
  if (!cand || !tmp || !mask) {
    free(cand); free(tmp); free(mask);
    return -1;
  }
  for (int64_t r = 0; r < n_requests; ++r) {
    for (int32_t k = pref_off[r]; k < pref_off[r + 1]; ++k) {
      int32_t id = pref_ids[k];
      if (id >= 0 && id < n) mask[id] = 1;
    }
    schedule_one_literal(n, battery, flags, mask, min_battery[r], cand, tmp,
                         &out_idx[r], &out_score[r]);
    for (int32_t k = pref_off[r]; k < pref_off[r + 1]; ++k) {
      int32_t id = pref_ids[k];
      if (id >= 0 && id < n) mask[id] = 0;
    }
  }
  free(cand); free(tmp); free(mask);
  return 0;
}

/* Same selection without the sort: first strict maximum in item order.
 * sort.SliceStable + [0] == first maximum (stability, controller.go:160).
 * Cheaper O(N) per request; used to cross-check the literal form and as the
 * timed CPU baseline for the "scan every node" algorithm. */
KLMO_API int klmo_schedule_scan(int64_t n, const double* battery,
                                const uint8_t* flags, int64_t n_requests,
                                const double* min_battery,
                                const int32_t* pref_off, const int32_t* pref_ids,
                                int32_t* out_idx, double* out_score) {
  uint8_t* mask = calloc((size_t)(n > 0 ? n : 1), 1);
  if (!mask) return -1;
  for (int64_t r = 0; r < n_requests; ++r) {
    for (int32_t k = pref_off[r]; k < pref_off[r + 1]; ++k) {
      int32_t id = pref_ids[k];
      if (id >= 0 && id < n) mask[id] = 1;
    }
    double mb = min_battery[r];
    int32_t bi = -1;
    double bs = 0.0;
    for (int64_t i = 0; i < n; ++i) {
      if ((flags[i] & 3u) != 3u) continue;
      if (mb > 0 && battery[i] < mb) continue;
      double s = battery[i];
      if (mask[i]) s += 10;
      if (bi < 0 || s > bs) { bi = (int32_t)i; bs = s; }
    }
    out_idx[r] = bi;
    out_score[r] = bi < 0 ? 0.0 : bs;
    for (int32_t k = pref_off[r]; k < pref_off[r + 1]; ++k) {
      int32_t id = pref_ids[k];
      if (id >= 0 && id < n) mask[id] = 0;
    }
  }
  free(mask);
  return 0;
}

/* Decomposed form: g = first maximum of battery over
 * nodes passing the name+status filters; the winner of a request is the best
 * of {g} U {eligible preferred nodes passing the gate}. O(N + R*K). This is
 * the algorithm the GPU's production path uses; here it exists so the tests
 * can prove it equal to the literal form on the CPU, and as the CPU
 * baseline that gives the CPU every algorithmic advantage the GPU has. */
KLMO_API int klmo_schedule_decomposed(int64_t n, const double* battery,
                                      const uint8_t* flags, int64_t n_requests,
                                      const double* min_battery,
                                      const int32_t* pref_off,
                                      const int32_t* pref_ids, int32_t* out_idx,
                                      double* out_score) {
  int64_t g = -1;
  for (int64_t i = 0; i < n; ++i) {
    if ((flags[i] & 3u) != 3u) continue;
    if (g < 0 || battery[i] > battery[g]) g = i;
  }
  for (int64_t r = 0; r < n_requests; ++r) {
    double mb = min_battery[r];
    int32_t bi = -1;
    double bs = 0.0;
    if (g >= 0 && !(mb > 0 && battery[g] < mb)) {
      bi = (int32_t)g;
      bs = battery[g];
      for (int32_t k = pref_off[r]; k < pref_off[r + 1]; ++k)
        if (pref_ids[k] == g) { bs = battery[g] + 10; break; }
      for (int32_t k = pref_off[r]; k < pref_off[r + 1]; ++k) {
        int32_t p = pref_ids[k];
        if (p < 0 || p >= n) continue;
        if ((flags[p] & 3u) != 3u) continue;
        if (mb > 0 && battery[p] < mb) continue;
        double s = battery[p] + 10;
        if (s > bs || (s == bs && p < bi)) { bi = p; bs = s; }
      }
    }
    out_idx[r] = bi;
    out_score[r] = bi < 0 ? 0.0 : bs;
  }
  return 0;
}

/* ------------------------------------------------------------------ */
/* Cluster aggregate: internal/metrics/manager.go:493-565.             */
/* ------------------------------------------------------------------ */

typedef struct {
  int64_t total_nodes, healthy_nodes, total_pods, running_pods;
  int64_t total_cpu, used_cpu, total_memory, used_memory;
  int64_t total_gpus, available_gpus;
  double cpu_usage_rate, memory_usage_rate;
  int32_t health_status;
  uint32_t issues;
} klmo_cluster_out;

KLMO_API void klmo_cluster_aggregate(
    int64_t n, const int64_t* cpu_cap, const int64_t* cpu_use,
    const int64_t* mem_cap, const int64_t* mem_use, const int32_t* gpu_count,
    const uint8_t* healthy, const double* gpu_usage, int64_t n_gpu_usage,
    const uint8_t* pod_running, int64_t n_pods, klmo_cluster_out* o) {
  memset(o, 0, sizeof(*o));
  o->total_nodes = n;                                       /* :497 */
  for (int64_t i = 0; i < n; ++i)
    if (healthy[i]) o->healthy_nodes++;                     /* :499-503 */
  o->total_pods = n_pods;                                   /* :506 */
  for (int64_t p = 0; p < n_pods; ++p)
    if (pod_running[p]) o->running_pods++;                  /* :508-512 */
  for (int64_t i = 0; i < n; ++i) {                         /* :522-535 */
    o->total_cpu += cpu_cap[i];
    o->used_cpu += cpu_use[i];
    o->total_memory += mem_cap[i];
    o->used_memory += mem_use[i];
    o->total_gpus += gpu_count[i];
  }
  for (int64_t k = 0; k < n_gpu_usage; ++k)
    if (gpu_usage[k] < 50.0) o->available_gpus++;           /* :530-534 */
  if (o->total_cpu > 0)                                     /* :538-540 */
    o->cpu_usage_rate = (double)o->used_cpu / (double)o->total_cpu * 100.0;
  if (o->total_memory > 0)                                  /* :541-543 */
    o->memory_usage_rate =
        (double)o->used_memory / (double)o->total_memory * 100.0;
  if (o->healthy_nodes < o->total_nodes) o->issues |= 1u;   /* :547-549 */
  if (o->cpu_usage_rate > 80) o->issues |= 2u;              /* :550-552 */
  if (o->memory_usage_rate > 80) o->issues |= 4u;           /* :553-555 */
  if (o->issues == 0)                                       /* :558-564 */
    o->health_status = 0;
  else if (o->cpu_usage_rate > 90 || o->memory_usage_rate > 90 ||
           o->healthy_nodes < o->total_nodes / 2) /* integer division, :560 */
    o->health_status = 2;
  else
    o->health_status = 1;
}

/* ------------------------------------------------------------------ */
/* Node rates: internal/metrics/sources/node_metrics.go:116-140 and     */
/* pkg/metrics/types.go:159-162 IsUnderPressure.                        */
/* ------------------------------------------------------------------ */
KLMO_API void klmo_node_rates(int64_t n, const int64_t* cpu_cap,
                              const int64_t* cpu_use, const int64_t* mem_cap,
                              const int64_t* mem_use, const int64_t* disk_cap,
                              const int64_t* disk_alloc, double* cpu_rate,
                              double* mem_rate, int64_t* disk_usage,
                              double* disk_rate, uint8_t* under_pressure) {
  for (int64_t i = 0; i < n; ++i) {
    int64_t du = 0;
    if (disk_alloc) {                       /* :117-124 */
      du = disk_cap[i] - disk_alloc[i];
      if (du < 0) du = 0;
    }
    double cr = 0.0, mr = 0.0, dr = 0.0;
    if (cpu_cap[i] > 0) cr = (double)cpu_use[i] / (double)cpu_cap[i] * 100.0;
    if (mem_cap[i] > 0) mr = (double)mem_use[i] / (double)mem_cap[i] * 100.0;
    if (disk_cap[i] > 0) dr = (double)du / (double)disk_cap[i] * 100.0;
    if (cpu_rate) cpu_rate[i] = cr;
    if (mem_rate) mem_rate[i] = mr;
    if (disk_usage) disk_usage[i] = du;
    if (disk_rate) disk_rate[i] = dr;
    if (under_pressure)                      /* types.go:159-162 */
      under_pressure[i] = (cr > 80.0 || mr > 80.0 || dr > 90.0) ? 1 : 0;
  }
}

/* NodeMetrics.GetAvailableResources, pkg/metrics/types.go:151-156. */
KLMO_API void klmo_node_available(int64_t n, const int64_t* cpu_cap,
                                  const int64_t* cpu_use, const int64_t* mem_cap,
                                  const int64_t* mem_use, const int64_t* disk_cap,
                                  const int64_t* disk_use, double* cpu_cores,
                                  double* mem_gb, double* disk_gb) {
  for (int64_t i = 0; i < n; ++i) {
    cpu_cores[i] = (double)(cpu_cap[i] - cpu_use[i]) / 1000.0;
    mem_gb[i] = (double)(mem_cap[i] - mem_use[i]) / 1024 / 1024 / 1024;
    disk_gb[i] = (double)(disk_cap[i] - disk_use[i]) / 1024 / 1024 / 1024;
  }
}

/* ------------------------------------------------------------------ */
/* Pod rates: internal/metrics/sources/pod_metrics.go:104-171;          */
/* pkg/metrics/types.go:165-184.                                        */
/* ------------------------------------------------------------------ */
KLMO_API void klmo_pod_rates(int64_t n_pods, const int32_t* ctr_off,
                             const int64_t* c_cpu_use, const int64_t* c_mem_use,
                             const int64_t* c_cpu_req, const int64_t* c_cpu_lim,
                             const int64_t* c_mem_req, const int64_t* c_mem_lim,
                             int64_t* pod_totals, double* cpu_rate,
                             double* mem_rate, double* cpu_util,
                             double* mem_util, uint8_t* over_limit) {
  for (int64_t p = 0; p < n_pods; ++p) {
    int64_t cu = 0, mu = 0, cq = 0, cl = 0, mq = 0, ml = 0;
    for (int32_t c = ctr_off[p]; c < ctr_off[p + 1]; ++c) {
      cq += c_cpu_req[c]; cl += c_cpu_lim[c];   /* :105-119 */
      mq += c_mem_req[c]; ml += c_mem_lim[c];
      cu += c_cpu_use[c]; mu += c_mem_use[c];   /* :124-127 */
    }
    if (pod_totals) {
      pod_totals[0 * n_pods + p] = cu; pod_totals[1 * n_pods + p] = mu;
      pod_totals[2 * n_pods + p] = cq; pod_totals[3 * n_pods + p] = cl;
      pod_totals[4 * n_pods + p] = mq; pod_totals[5 * n_pods + p] = ml;
    }
    double cr = 0.0, mr = 0.0, cut = 0.0, mut = 0.0;
    if (cl > 0) cr = (double)cu / (double)cl * 100.0;   /* :162-165 */
    if (ml > 0) mr = (double)mu / (double)ml * 100.0;   /* :167-170 */
    if (cq > 0) cut = (double)cu / (double)cq * 100.0;  /* types.go:166-168 */
    if (mq > 0) mut = (double)mu / (double)mq * 100.0;  /* types.go:169-171 */
    if (cpu_rate) cpu_rate[p] = cr;
    if (mem_rate) mem_rate[p] = mr;
    if (cpu_util) cpu_util[p] = cut;
    if (mem_util) mem_util[p] = mut;
    if (over_limit) {                                    /* types.go:176-184 */
      uint8_t ov = 0;
      if (cl > 0 && cu >= (int64_t)((double)cl * 0.9)) ov = 1;
      if (ml > 0 && mu >= (int64_t)((double)ml * 0.9)) ov = 1;
      over_limit[p] = ov;
    }
  }
}

/* Pod restart count, internal/metrics/sources/pod_metrics.go:174-177:
 *   var restarts int32; for each containerStatus: restarts += RestartCount
 * (int32 that wraps). st_off is the CSR over a pod's container statuses. */
KLMO_API void klmo_pod_restarts(int64_t n_pods, const int32_t* st_off,
                                const int32_t* restart_count, int32_t* restarts) {
  for (int64_t p = 0; p < n_pods; ++p) {
    uint32_t acc = 0;
    for (int32_t c = st_off[p]; c < st_off[p + 1]; ++c) acc += (uint32_t)restart_count[c];
    restarts[p] = (int32_t)acc;
  }
}

/* Node health from conditions, internal/metrics/sources/node_metrics.go:142-164.
 * type: 0 Ready, 1 MemoryPressure, 2 DiskPressure, 3 PIDPressure, 4 NetworkUnavailable,
 * 5 anything else; status: 0 False, 1 True, 2 Unknown. failing: bit per type that made the
 * node unhealthy. */
KLMO_API void klmo_node_health(int64_t n, const int32_t* cond_off, const uint8_t* cond_type,
                               const uint8_t* cond_status, uint8_t* healthy, uint8_t* failing) {
  for (int64_t i = 0; i < n; ++i) {
    int ok = 1;                                               /* :143 healthy := true */
    unsigned bad = 0;
    for (int32_t c = cond_off[i]; c < cond_off[i + 1]; ++c) { /* :145 */
      const unsigned t = cond_type[c], st = cond_status[c];
      if (t == 0) {                                           /* :147 NodeReady */
        if (st != 1) { ok = 0; bad |= 1u; }                   /* :148-151 */
      } else if (st == 1) {                                   /* :154 */
        if (t == 1 || t == 2 || t == 3 || t == 4) { ok = 0; bad |= 1u << t; }  /* :155-161 */
      }
    }
    healthy[i] = (uint8_t)ok;
    failing[i] = (uint8_t)bad;
  }
}

/* Top-k: the literal candidate list of controller.go:174-221 in list order,
 * sorted by a STABLE descending sort on Score (controller.go:160), cut at k. The reference takes
 * [0] only (:161); beyond that there are no reference semantics (parity unpinned).
 * out_idx / out_score are [R][k], padded with -1 / 0. */
KLMO_API int klmo_schedule_topk(int64_t n, const double* battery, const uint8_t* flags, int64_t r,
                                const double* min_battery, const int32_t* pref_off,
                                const int32_t* pref_ids, int k, int32_t* out_idx,
                                double* out_score) {
  double* cs = (double*)malloc((size_t)(n > 0 ? n : 1) * sizeof(double));
  int32_t* ci = (int32_t*)malloc((size_t)(n > 0 ? n : 1) * sizeof(int32_t));
  if (!cs || !ci) { free(cs); free(ci); return -1; }
  for (int64_t q = 0; q < r; ++q) {
    int64_t m = 0;
    const double mb = min_battery[q];
    for (int64_t i = 0; i < n; ++i) {                         /* :181 items in list order */
      if (!(flags[i] & 1u)) continue;                         /* :190 */
      if (mb > 0 && battery[i] < mb) continue;                /* :194 */
      if (!(flags[i] & 2u)) continue;                         /* :198 */
      double sc = battery[i];                                 /* :205 */
      for (int32_t p = pref_off[q]; p < pref_off[q + 1]; ++p)
        if (pref_ids[p] == i) { sc += 10; break; }            /* :206-208 */
      cs[m] = sc; ci[m] = (int32_t)i; ++m;
    }
    /* stable descending insertion sort of the first k places (what SliceStable + [0:k] yields) */
    for (int j = 0; j < k; ++j) {
      int64_t best = -1;
      for (int64_t c = 0; c < m; ++c) {
        if (ci[c] < 0) continue;                              /* already taken */
        if (best < 0 || cs[c] > cs[best]) best = c;           /* strict >: the earlier one stays on ties */
      }
      if (best < 0) { out_idx[q * k + j] = -1; out_score[q * k + j] = 0.0; continue; }
      out_idx[q * k + j] = ci[best]; out_score[q * k + j] = cs[best];
      ci[best] = -1;
    }
  }
  free(cs); free(ci);
  return 0;
}

/* Window-derived node features: NOT in the reference (controller.go:202-203
 * drops last_update; no score reads history) -- the build's own definition, parity unpinned:
 * stale heartbeat or a hot battery (p95 of channel 24) clears the status bit; the battery is
 * projected horizon_s ahead at the window's drain of channel 21, clamped to [0, battery]. */
KLMO_API void klmo_node_features(int64_t n, const double* battery, const uint8_t* flags,
                                 const double* age, const float* wmin, const float* wmax,
                                 const float* wp95, double max_age_s, double temp_limit,
                                 double horizon_s, double span_s, double* battery_out,
                                 uint8_t* flags_out) {
  for (int64_t i = 0; i < n; ++i) {
    uint8_t f = flags[i];
    double b = battery[i];
    if (age && max_age_s > 0 && age[i] > max_age_s) f &= (uint8_t)~2u;
    if (wp95 && temp_limit > 0 && (double)wp95[i * 64 + 24] > temp_limit) f &= (uint8_t)~2u;
    if (wmin && wmax && horizon_s > 0 && span_s > 0) {
      const double drop = (double)wmax[i * 64 + 21] - (double)wmin[i * 64 + 21];
      double proj = b - drop / span_s * horizon_s;
      if (proj < 0) proj = 0;
      if (proj < b) b = proj;
    }
    battery_out[i] = b;
    flags_out[i] = f;
  }
}

/* NetworkMetrics.GetQuality, pkg/metrics/types.go:187-199:
 * 0 disconnected, 1 excellent, 2 good, 3 fair, 4 poor. */
KLMO_API void klmo_network_quality(int64_t n, const uint8_t* connected,
                                   const double* rtt, uint8_t* quality) {
  for (int64_t i = 0; i < n; ++i) {
    if (!connected[i]) quality[i] = 0;
    else if (rtt[i] < 10) quality[i] = 1;
    else if (rtt[i] < 50) quality[i] = 2;
    else if (rtt[i] < 100) quality[i] = 3;
    else quality[i] = 4;
  }
}

/* ------------------------------------------------------------------ */
/* Fleet scans: internal/metrics/sources/uav_metrics.go:197-233.        */
/* The reference returns the low-battery names in Go map order          */
/* (unspecified); the build fixes ascending index order.                */
/* ------------------------------------------------------------------ */
KLMO_API void klmo_fleet_scan(int64_t n, const uint8_t* status_ok,
                              const double* remaining_percent, double threshold,
                              int64_t* healthy_count, int32_t* low_idx,
                              int64_t* n_low) {
  int64_t h = 0, l = 0;
  for (int64_t i = 0; i < n; ++i) {
    if (status_ok[i]) ++h;                                       /* :206-210 */
    if (remaining_percent[i] < threshold) low_idx[l++] = (int32_t)i; /* :226 */
  }
  *healthy_count = h;
  *n_low = l;
}

/* ------------------------------------------------------------------ */
/* Metric ring + window statistics. NO REFERENCE CODE: parity unpinned. */
/* ------------------------------------------------------------------ */

/* Append one sample per listed series (series_ids == NULL: every series). */
KLMO_API void klmo_ring_ingest(float* ring, uint32_t* seen, int64_t n_series,
                               int32_t window, int32_t pitch, int64_t n,
                               const int64_t* series_ids, const float* values) {
  for (int64_t k = 0; k < n; ++k) {
    int64_t s = series_ids ? series_ids[k] : k;
    if (s < 0 || s >= n_series) continue;
    uint32_t c = seen[s];
    ring[s * (int64_t)pitch + (int64_t)(c % (uint32_t)window)] = values[k];
    if (c != UINT32_MAX) seen[s] = c + 1;
  }
}

static int cmp_f32(const void* a, const void* b) {
  float x = *(const float*)a, y = *(const float*)b;
  return (x > y) - (x < y);
}

/* Canonical definition: sort a copy, index it. */
KLMO_API int klmo_window_stats(const float* ring, const uint32_t* seen,
                               int32_t window, int32_t pitch, int64_t s0,
                               int64_t s1, float* out_min, float* out_max,
                               double* out_mean, float* out_p95) {
  float* buf = malloc((size_t)(window > 0 ? window : 1) * sizeof(float));
  if (!buf) return -1;
  for (int64_t s = s0; s < s1; ++s) {
    int64_t n = seen[s] < (uint32_t)window ? (int64_t)seen[s] : window;
    float mn = 0.f, mx = 0.f, p95 = 0.f;
    double mean = 0.0;
    if (n > 0) {
      const float* row = ring + s * (int64_t)pitch;
      double sum = 0.0;
      mn = mx = row[0];
      for (int64_t t = 0; t < n; ++t) {
        float v = row[t];
        if (v < mn) mn = v;
        if (v > mx) mx = v;
        sum += (double)v;
      }
      mean = sum / (double)n;
      memcpy(buf, row, (size_t)n * sizeof(float));
      qsort(buf, (size_t)n, sizeof(float), cmp_f32);
      /* nearest rank: ceil(0.95 n) - 1, in integers: ceil(19 n / 20) - 1 */
      int64_t rank = (19 * n + 19) / 20 - 1;
      p95 = buf[rank];
    }
    /* one canonical zero: -0.0 and +0.0 compare equal, so which of them a sort or a minimum
     * returns is arbitrary; every implementation reports +0.0 (x + 0.0f in round-to-nearest) */
    if (out_min) out_min[s - s0] = mn + 0.0f;
    if (out_max) out_max[s - s0] = mx + 0.0f;
    if (out_mean) out_mean[s - s0] = mean;
    if (out_p95) out_p95[s - s0] = p95 + 0.0f;
  }
  free(buf);
  return 0;
}

/* Quickselect with median-of-3 (what a careful CPU implementation would run:
 * expected O(n) instead of the sort's O(n log n)). Same results as
 * klmo_window_stats; tests assert that. This is the timed CPU baseline.
 * Parallel over series with OpenMP when built with -fopenmp. */
static float quickselect(float* a, int64_t n, int64_t k) {
  int64_t lo = 0, hi = n - 1;
  while (lo < hi) {
    int64_t mid = lo + (hi - lo) / 2;
    float x = a[lo], y = a[mid], z = a[hi];
    float pivot = (x < y) ? ((y < z) ? y : (x < z ? z : x))
                          : ((x < z) ? x : (y < z ? z : y));
    int64_t i = lo, j = hi;
    while (i <= j) {
      while (a[i] < pivot) ++i;
      while (a[j] > pivot) --j;
      if (i <= j) {
        float t = a[i]; a[i] = a[j]; a[j] = t;
        ++i; --j;
      }
    }
    if (k <= j) hi = j;
    else if (k >= i) lo = i;
    else return a[k];
  }
  return a[k];
}

KLMO_API int klmo_window_stats_fast(const float* ring, const uint32_t* seen,
                                    int32_t window, int32_t pitch, int64_t s0,
                                    int64_t s1, float* out_min, float* out_max,
                                    double* out_mean, float* out_p95,
                                    int n_threads) {
  int fail = 0;
#ifdef _OPENMP
#pragma omp parallel num_threads(n_threads > 0 ? n_threads : 1)
#endif
  {
    float* buf = malloc((size_t)(window > 0 ? window : 1) * sizeof(float));
    if (!buf) {
#ifdef _OPENMP
#pragma omp atomic write
#endif
      fail = 1;
    }
#ifdef _OPENMP
#pragma omp for schedule(static)
#endif
    for (int64_t s = s0; s < s1; ++s) {
      if (!buf) continue;
      int64_t n = seen[s] < (uint32_t)window ? (int64_t)seen[s] : window;
      float mn = 0.f, mx = 0.f, p95 = 0.f;
      double mean = 0.0;
      if (n > 0) {
        const float* row = ring + s * (int64_t)pitch;
        double sum = 0.0;
        mn = mx = row[0];
        for (int64_t t = 0; t < n; ++t) {
          float v = row[t];
          buf[t] = v;
          if (v < mn) mn = v;
          if (v > mx) mx = v;
          sum += (double)v;
        }
        mean = sum / (double)n;
        p95 = quickselect(buf, n, (19 * n + 19) / 20 - 1);
      }
      if (out_min) out_min[s - s0] = mn + 0.0f;
      if (out_max) out_max[s - s0] = mx + 0.0f;
      if (out_mean) out_mean[s - s0] = mean;
      if (out_p95) out_p95[s - s0] = p95 + 0.0f;
    }
    free(buf);
  }
  (void)n_threads;
  return fail ? -1 : 0;
}

/* Multi-threaded wrappers for the timed CPU baseline (requests are
 * independent, controller.go:103 loops them sequentially; giving the CPU all
 * cores is the generous reading). */
KLMO_API int klmo_schedule_scan_mt(int64_t n, const double* battery,
                                   const uint8_t* flags, int64_t n_requests,
                                   const double* min_battery,
                                   const int32_t* pref_off,
                                   const int32_t* pref_ids, int32_t* out_idx,
                                   double* out_score, int n_threads) {
  if (n_threads < 1) n_threads = 1;
  int fail = 0;
#ifdef _OPENMP
#pragma omp parallel for num_threads(n_threads) schedule(static)
#endif
  for (int t = 0; t < n_threads; ++t) {
    int64_t r0 = n_requests * t / n_threads, r1 = n_requests * (t + 1) / n_threads;
    /* pref_off is absolute, so pass shifted views of the per-request arrays */
    if (r1 > r0 &&
        klmo_schedule_scan(n, battery, flags, r1 - r0, min_battery + r0,
                           pref_off + r0, pref_ids, out_idx + r0,
                           out_score + r0) != 0) {
#ifdef _OPENMP
#pragma omp atomic write
#endif
      fail = 1;
    }
  }
  return fail ? -1 : 0;
}

KLMO_API int klmo_have_openmp(void) {
#ifdef _OPENMP
  return 1;
#else
  return 0;
#endif
}
