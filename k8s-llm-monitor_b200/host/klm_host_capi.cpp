// Flat C view of the C++ host layer so the (ctypes) tests can drive it. A session holds the
// inputs of one reconcile / one Collect and the outputs of the last call.
#include <cstring>
#include <string>
#include <vector>

#include "klm_host.h"
#include "klm_wire.h"

using namespace klmhost;

namespace {
struct Session {
  explicit Session(klm_handle* handle) : h(handle), uav_manager(handle) {}
  klm_handle* h;
  std::vector<UAVMetricItem> uavs;
  std::vector<SchedulingRequest> requests;
  std::vector<SchedulingRequestStatus> statuses;
  std::vector<NodeMetrics> nodes;
  std::vector<PodMetrics> pods;
  std::vector<UAVHealth> fleet;
  ClusterMetrics cluster;
  Manager uav_manager{nullptr};   // the UAV ingestion half (UpdateUAVReport) keeps state between calls
  std::vector<std::string> names;
  std::string error;
};
void put(char* dst, int cap, const std::string& s) {
  if (!dst || cap <= 0) return;
  const size_t n = s.size() < static_cast<size_t>(cap - 1) ? s.size() : static_cast<size_t>(cap - 1);
  std::memcpy(dst, s.data(), n);
  dst[n] = 0;
}
}  // namespace

extern "C" {

// h may be NULL for a decode / encode only session (no GPU call is possible then)
void* klmh_open(klm_handle* h) { return new Session(h); }
void klmh_close(void* s) { delete static_cast<Session*>(s); }
const char* klmh_error(void* s) { return static_cast<Session*>(s)->error.c_str(); }

void klmh_clear(void* sv) {
  Session* s = static_cast<Session*>(sv);
  s->uavs.clear(); s->requests.clear(); s->statuses.clear();
  s->nodes.clear(); s->pods.clear(); s->fleet.clear(); s->names.clear();
}

// node_name / status may be NULL (field absent); battery_kind: 0 missing, 1 number
void klmh_add_uav(void* sv, const char* node_name, const char* uav_id, int battery_kind,
                  double battery, const char* status) {
  UAVMetricItem u;
  u.HasNodeName = node_name != nullptr;
  if (node_name) u.NodeName = node_name;
  if (uav_id) u.UAVID = uav_id;
  u.HasBattery = battery_kind != 0;
  u.Battery = battery;
  u.HasStatus = status != nullptr;
  if (status) u.CollectionStatus = status;
  static_cast<Session*>(sv)->uavs.push_back(u);
}

void klmh_add_request(void* sv, const char* phase, const char* wl_name, const char* wl_ns,
                      double min_battery, int n_pref, const char** pref) {
  SchedulingRequest q;
  q.HasPhase = phase != nullptr;
  if (phase) q.Phase = phase;
  if (wl_name) q.Spec.Workload.Name = wl_name;
  if (wl_ns) q.Spec.Workload.Namespace = wl_ns;
  q.Spec.MinBatteryPercent = min_battery;
  for (int i = 0; i < n_pref; ++i) q.Spec.PreferredNodes.push_back(pref[i]);
  static_cast<Session*>(sv)->requests.push_back(q);
}

// wire format in: replace the session's UAV items / requests with the decoded lists
int klmh_load_uavs_json(void* sv, const char* json) {
  Session* s = static_cast<Session*>(sv);
  try { s->uavs = DecodeUAVMetricList(json); }
  catch (const std::exception& e) { s->error = e.what(); return -1; }
  return static_cast<int>(s->uavs.size());
}
int klmh_load_requests_json(void* sv, const char* json, int apimachinery_ints) {
  Session* s = static_cast<Session*>(sv);
  try { s->requests = DecodeSchedulingRequestList(json, apimachinery_ints != 0); }
  catch (const std::exception& e) { s->error = e.what(); return -1; }
  return static_cast<int>(s->requests.size());
}
// decoded request i: min battery, number of preferred nodes, workload name (for the tests)
int klmh_request_info(void* sv, int i, double* min_battery, int* n_pref, char* wl_name, char* phase, int cap) {
  Session* s = static_cast<Session*>(sv);
  if (i < 0 || i >= static_cast<int>(s->requests.size())) return -1;
  const SchedulingRequest& q = s->requests[i];
  if (min_battery) *min_battery = q.Spec.MinBatteryPercent;
  if (n_pref) *n_pref = static_cast<int>(q.Spec.PreferredNodes.size());
  put(wl_name, cap, q.Spec.Workload.Name);
  put(phase, cap, q.Phase);
  return q.HasPhase ? 1 : 0;
}
int klmh_uav_info(void* sv, int i, double* battery, int* has_battery, char* name, char* status, int cap) {
  Session* s = static_cast<Session*>(sv);
  if (i < 0 || i >= static_cast<int>(s->uavs.size())) return -1;
  const UAVMetricItem& u = s->uavs[i];
  if (battery) *battery = u.Battery;
  if (has_battery) *has_battery = u.HasBattery ? 1 : 0;
  put(name, cap, u.NodeName);
  put(status, cap, u.CollectionStatus);
  return (u.HasNodeName ? 1 : 0) | (u.HasStatus ? 2 : 0);
}
// wire format out: status i of the last reconcile as the JSON map controller.go:229-236 writes
int klmh_status_json(void* sv, int i, const char* last_updated, char* dst, int cap) {
  Session* s = static_cast<Session*>(sv);
  if (i < 0 || i >= static_cast<int>(s->statuses.size())) return -1;
  put(dst, cap, EncodeStatus(s->statuses[i], last_updated ? last_updated : ""));
  return s->statuses[i].Updated ? 1 : 0;
}

// every status of the last reconcile encoded in one call (timing: tools/host_bench.py); returns the
// number of bytes of JSON produced
int64_t klmh_status_json_all(void* sv, const char* last_updated) {
  Session* s = static_cast<Session*>(sv);
  const std::string stamp = last_updated ? last_updated : "";
  int64_t total = 0;
  for (const SchedulingRequestStatus& st : s->statuses) total += static_cast<int64_t>(EncodeStatus(st, stamp).size());
  return total;
}

// a float64 in Go's encoding/json form (what EncodeStatus writes for `score`)
void klmh_format_float(double v, char* dst, int cap) { put(dst, cap, FormatJSONFloat(v)); }

// decode one UAVReport (models.go:182-192) and flatten its state: out64 gets the 64 channels.
// Returns 1 with state, 0 without, -1 on a decode error (klmh_error).
int klmh_report_decode(void* sv, const char* json, float* out64, char* node, char* status, char* source,
                       double* remaining_percent, int cap) {
  Session* s = static_cast<Session*>(sv);
  try {
    UAVReport r = DecodeUAVReport(json);
    const bool kept = ApplyUAVReportDefaults(r, "now");
    for (int c = 0; c < kUAVChannels; ++c) out64[c] = r.Channels[c];
    put(node, cap, kept ? r.NodeName : "");
    put(status, cap, r.Status);
    put(source, cap, r.Source);
    if (remaining_percent) *remaining_percent = r.RemainingPercent;
    return r.HasState ? 1 : 0;
  } catch (const std::exception& e) { s->error = e.what(); return -1; }
}
int klmh_enable_uav_ring(void* sv, int64_t max_nodes, int window) {
  Session* s = static_cast<Session*>(sv);
  try { s->uav_manager.EnableUAVRing(max_nodes, window); }
  catch (const std::exception& e) { s->error = e.what(); return -1; }
  return 0;
}
// Manager.UpdateUAVReport on a JSON report: node index, -1 dropped, -2 error
int64_t klmh_update_uav_report(void* sv, const char* json, const char* now) {
  Session* s = static_cast<Session*>(sv);
  try { return s->uav_manager.UpdateUAVReport(DecodeUAVReport(json), now ? now : ""); }
  catch (const std::exception& e) { s->error = e.what(); return -2; }
}
// entry i of the UAV snapshot: battery, status, last heartbeat
int klmh_uav_entry(void* sv, int64_t i, double* remaining_percent, char* node, char* status, char* heartbeat, int cap) {
  Session* s = static_cast<Session*>(sv);
  const auto& snap = s->uav_manager.UAVSnapshot();
  if (i < 0 || i >= static_cast<int64_t>(snap.size())) return -1;
  if (remaining_percent) *remaining_percent = snap[i].RemainingPercent;
  put(node, cap, snap[i].NodeName);
  put(status, cap, snap[i].Status);
  put(heartbeat, cap, snap[i].LastHeartbeat);
  return snap[i].HasState ? 1 : 0;
}

// strings.ToLower as the host layer applies it to node names and statuses
void klmh_to_lower(const char* in, char* dst, int cap) { put(dst, cap, ToLower(in ? in : "")); }

// status write-back: the statuses of the last reconcile go to `cb` (ns, name, status JSON; nonzero
// = failure) from `workers` threads; names are "req-<i>" in namespace "default" unless set with
// klmh_name_request. Returns how many calls failed, -1 on error.
typedef int (*klmh_update_cb)(const char* ns, const char* name, const char* status_json);
int klmh_write_statuses(void* sv, const char* now, int workers, klmh_update_cb cb) {
  Session* s = static_cast<Session*>(sv);
  try {
    std::vector<StatusUpdate> ups(s->statuses.size());
    for (size_t i = 0; i < ups.size(); ++i) {
      ups[i].Namespace = "default";
      ups[i].Name = "req-" + std::to_string(i);
      ups[i].Status = s->statuses[i];
    }
    const auto res = WriteStatuses(ups, now ? now : "", [&](const std::string& ns, const std::string& name,
                                                            const std::string& js, std::string* err) {
      if (cb(ns.c_str(), name.c_str(), js.c_str()) != 0) { *err = "update refused"; return false; }
      return true;
    }, workers);
    int failed = 0;
    for (const auto& r : res) failed += r.Ok ? 0 : 1;
    return failed;
  } catch (const std::exception& e) { s->error = e.what(); return -1; }
}
// a status without a reconcile (tests of the write-back path alone)
void klmh_add_status(void* sv, int updated, const char* phase, const char* node, const char* uav, double score,
                     const char* message) {
  SchedulingRequestStatus st;
  st.Updated = updated != 0;
  st.Phase = phase ? phase : "";
  st.AssignedNode = node ? node : "";
  st.AssignedUAV = uav ? uav : "";
  st.Score = score;
  st.Message = message ? message : "";
  static_cast<Session*>(sv)->statuses.push_back(st);
}

int klmh_reconcile_capacity(void* sv, const int32_t* capacity, int n) {
  Session* s = static_cast<Session*>(sv);
  if (!s->h) { s->error = "no GPU handle in this session"; return -1; }
  try {
    s->statuses = Controller(s->h).ReconcileWithCapacity(s->requests, s->uavs,
                                                         std::vector<int32_t>(capacity, capacity + n));
  } catch (const std::exception& e) {
    s->error = e.what();
    return -1;
  }
  return static_cast<int>(s->statuses.size());
}

int klmh_reconcile(void* sv) {
  Session* s = static_cast<Session*>(sv);
  if (!s->h) { s->error = "no GPU handle in this session"; return -1; }
  try {
    s->statuses = Controller(s->h).Reconcile(s->requests, s->uavs);
  } catch (const std::exception& e) {
    s->error = e.what();
    return -1;
  }
  return static_cast<int>(s->statuses.size());
}

int klmh_status(void* sv, int i, char* phase, char* node, char* uav, double* score, char* message,
                int cap) {
  Session* s = static_cast<Session*>(sv);
  if (i < 0 || i >= static_cast<int>(s->statuses.size())) return -1;
  const SchedulingRequestStatus& st = s->statuses[i];
  put(phase, cap, st.Phase); put(node, cap, st.AssignedNode); put(uav, cap, st.AssignedUAV);
  put(message, cap, st.Message);
  if (score) *score = st.Score;
  return st.Updated ? 1 : 0;
}

void klmh_add_node(void* sv, int64_t cpu_cap, int64_t cpu_use, int64_t mem_cap, int64_t mem_use,
                   int healthy, int gpu_count, int n_gpu_usage, const double* gpu_usage) {
  NodeMetrics n;
  n.CPUCapacity = cpu_cap; n.CPUUsage = cpu_use; n.MemoryCapacity = mem_cap; n.MemoryUsage = mem_use;
  n.Healthy = healthy != 0; n.GPUCount = gpu_count;
  n.GPUUsage.assign(gpu_usage, gpu_usage + n_gpu_usage);
  static_cast<Session*>(sv)->nodes.push_back(n);
}
void klmh_add_pod(void* sv, const char* phase) {
  PodMetrics p;
  p.Phase = phase ? phase : "";
  static_cast<Session*>(sv)->pods.push_back(p);
}

// numeric part -> *out; returns the number of issue strings, or -1
int klmh_calculate_cluster_metrics(void* sv, klm_cluster_out* out, char* health, int cap) {
  Session* s = static_cast<Session*>(sv);
  try {
    s->cluster = Manager(s->h).CalculateClusterMetrics(s->nodes, s->pods);
  } catch (const std::exception& e) {
    s->error = e.what();
    return -1;
  }
  const ClusterMetrics& m = s->cluster;
  if (out) {
    *out = klm_cluster_out{m.TotalNodes, m.HealthyNodes, m.TotalPods, m.RunningPods, m.TotalCPU, m.UsedCPU,
                           m.TotalMemory, m.UsedMemory, m.TotalGPUs, m.AvailableGPUs, m.CPUUsageRate,
                           m.MemoryUsageRate, 0, 0};
  }
  put(health, cap, m.HealthStatus);
  return static_cast<int>(m.Issues.size());
}
int klmh_issue(void* sv, int i, char* dst, int cap) {
  Session* s = static_cast<Session*>(sv);
  if (i < 0 || i >= static_cast<int>(s->cluster.Issues.size())) return -1;
  put(dst, cap, s->cluster.Issues[i]);
  return 0;
}

void klmh_add_fleet_uav(void* sv, const char* node_name, const char* system_status, double pct) {
  UAVHealth u;
  u.NodeName = node_name ? node_name : "";
  u.SystemStatus = system_status ? system_status : "";
  u.RemainingPercent = pct;
  static_cast<Session*>(sv)->fleet.push_back(u);
}
int64_t klmh_healthy_uav_count(void* sv) {
  Session* s = static_cast<Session*>(sv);
  try { return Manager(s->h).GetHealthyUAVCount(s->fleet); }
  catch (const std::exception& e) { s->error = e.what(); return -1; }
}
int klmh_low_battery_uavs(void* sv, double threshold) {
  Session* s = static_cast<Session*>(sv);
  try { s->names = Manager(s->h).GetLowBatteryUAVs(s->fleet, threshold); }
  catch (const std::exception& e) { s->error = e.what(); return -1; }
  return static_cast<int>(s->names.size());
}
int klmh_name(void* sv, int i, char* dst, int cap) {
  Session* s = static_cast<Session*>(sv);
  if (i < 0 || i >= static_cast<int>(s->names.size())) return -1;
  put(dst, cap, s->names[i]);
  return 0;
}

}  // extern "C"
