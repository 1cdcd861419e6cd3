#include "klm_host.h"
#include "klm_wire.h"
#include <algorithm>

#include <atomic>
#include <cstdio>
#include <thread>
#include <stdexcept>
#include <unordered_map>

namespace klmhost {

// unicode.ToLower for the scripts a node name realistically uses: ASCII, Latin-1, Latin
// Extended-A, Greek, Cyrillic (Go's tables cover every script; anything else passes through).
static uint32_t lower_rune(uint32_t r) {
  if (r < 0x80) return (r >= 'A' && r <= 'Z') ? r + 32 : r;
  if (r >= 0xC0 && r <= 0xDE) return r == 0xD7 ? r : r + 32;
  if (r >= 0x100 && r <= 0x17F) {
    if (r == 0x130) return 'i';            // LATIN CAPITAL LETTER I WITH DOT ABOVE
    if (r == 0x178) return 0xFF;           // Y WITH DIAERESIS
    if (r <= 0x137 || (r >= 0x14A && r <= 0x177)) return (r & 1) ? r : r + 1;    // even capitals
    if ((r >= 0x139 && r <= 0x148) || (r >= 0x179 && r <= 0x17E)) return (r & 1) ? r + 1 : r;  // odd capitals
    return r;
  }
  if (r >= 0x391 && r <= 0x3A9) return r == 0x3A2 ? r : r + 32;  // Greek
  if (r == 0x386) return 0x3AC;
  if (r >= 0x388 && r <= 0x38A) return r + 37;
  if (r == 0x38C) return 0x3CC;
  if (r == 0x38E || r == 0x38F) return r + 63;
  if (r >= 0x410 && r <= 0x42F) return r + 32;                   // Cyrillic
  if (r >= 0x400 && r <= 0x40F) return r + 80;
  return r;
}

static void put_rune(std::string& o, uint32_t r) {
  if (r < 0x80) o += static_cast<char>(r);
  else if (r < 0x800) { o += static_cast<char>(0xC0 | (r >> 6)); o += static_cast<char>(0x80 | (r & 0x3F)); }
  else if (r < 0x10000) {
    o += static_cast<char>(0xE0 | (r >> 12)); o += static_cast<char>(0x80 | ((r >> 6) & 0x3F));
    o += static_cast<char>(0x80 | (r & 0x3F));
  } else {
    o += static_cast<char>(0xF0 | (r >> 18)); o += static_cast<char>(0x80 | ((r >> 12) & 0x3F));
    o += static_cast<char>(0x80 | ((r >> 6) & 0x3F)); o += static_cast<char>(0x80 | (r & 0x3F));
  }
}

// strings.ToLower (controller.go:177, :188, :206): rune by rune over UTF-8; like Go's strings.Map
// an invalid byte becomes U+FFFD once the string is not pure ASCII.
std::string ToLower(const std::string& s) {
  bool ascii = true;
  for (unsigned char c : s) ascii = ascii && c < 0x80;
  std::string o;
  o.reserve(s.size());
  if (ascii) {
    for (char c : s) o += (c >= 'A' && c <= 'Z') ? static_cast<char>(c + 32) : c;
    return o;
  }
  const size_t n = s.size();
  for (size_t i = 0; i < n;) {
    const unsigned char c = static_cast<unsigned char>(s[i]);
    uint32_t r = 0xFFFD;
    size_t len = 1;
    auto cont = [&](size_t k) { return i + k < n && (static_cast<unsigned char>(s[i + k]) & 0xC0) == 0x80; };
    if (c < 0x80) r = c;
    else if (c >= 0xC2 && c <= 0xDF && cont(1)) { r = ((c & 0x1Fu) << 6) | (s[i + 1] & 0x3F); len = 2; }
    else if (c >= 0xE0 && c <= 0xEF && cont(1) && cont(2)) {
      const uint32_t v = ((c & 0x0Fu) << 12) | ((s[i + 1] & 0x3Fu) << 6) | (s[i + 2] & 0x3F);
      if (v >= 0x800 && !(v >= 0xD800 && v <= 0xDFFF)) { r = v; len = 3; }
    } else if (c >= 0xF0 && c <= 0xF4 && cont(1) && cont(2) && cont(3)) {
      const uint32_t v = ((c & 0x07u) << 18) | ((s[i + 1] & 0x3Fu) << 12) | ((s[i + 2] & 0x3Fu) << 6) | (s[i + 3] & 0x3F);
      if (v >= 0x10000 && v <= 0x10FFFF) { r = v; len = 4; }
    }
    put_rune(o, lower_rune(r));
    i += len;
  }
  return o;
}

static void check(int rc, const char* what) {
  if (rc != KLM_OK) throw std::runtime_error(std::string(what) + ": " + klm_last_error());
}

static std::string sprintf1(const char* fmt, const char* s, double v) {
  char buf[512];
  std::snprintf(buf, sizeof(buf), fmt, s, v);
  return buf;
}

std::vector<SchedulingRequestStatus> Controller::Reconcile(
    const std::vector<SchedulingRequest>& requests, const std::vector<UAVMetricItem>& uavs) {
  // ---- flatten the snapshot once per reconcile (controller.go:181-200 per item)
  const int64_t n = static_cast<int64_t>(uavs.size());
  std::vector<double> battery(n);
  std::vector<uint8_t> flags(n);
  // lower(name) -> every index that carries it. Names are almost always unique, so the map holds
  // the first index and a link to the next item of the same name (no vector per name).
  std::unordered_map<std::string, int32_t> by_name;
  by_name.reserve(static_cast<size_t>(n) * 2);
  std::vector<int32_t> next_same(n, -1), last_same(n, -1);
  static const std::string kActive = "active";
  for (int64_t i = 0; i < n; ++i) {
    const UAVMetricItem& u = uavs[i];
    const bool named = u.HasNodeName && !u.NodeName.empty();
    battery[i] = u.HasBattery ? u.Battery : 0.0;                                   // :187
    uint8_t f = 0;
    if (named) f |= KLM_NODE_HAS_NAME;                                             // :190
    if (!u.HasStatus || u.CollectionStatus.empty() || u.CollectionStatus == kActive ||
        ToLower(u.CollectionStatus) == kActive)
      f |= KLM_NODE_STATUS_OK;                                                     // :188, :198
    flags[i] = f;
    if (named) {                                                                   // :206
      auto ins = by_name.emplace(ToLower(u.NodeName), static_cast<int32_t>(i));
      if (!ins.second) {  // a duplicate name: append to the chain of its first item
        const int32_t head = ins.first->second;
        const int32_t tail = last_same[head] < 0 ? head : last_same[head];
        next_same[tail] = static_cast<int32_t>(i);
        last_same[head] = static_cast<int32_t>(i);
      }
    }
  }

  std::vector<SchedulingRequestStatus> out(requests.size());
  std::vector<size_t> live;  // requests that reach buildCandidates
  std::vector<double> min_battery;
  std::vector<int32_t> pref_off{0}, pref_ids;
  std::vector<std::string> seen;  // (a request names a handful of nodes: a scan beats a set)
  live.reserve(requests.size());
  min_battery.reserve(requests.size());
  pref_off.reserve(requests.size() + 1);
  for (size_t r = 0; r < requests.size(); ++r) {
    const SchedulingRequest& q = requests[r];
    if (q.HasPhase && !q.Phase.empty() && q.Phase != "Pending") continue;          // :117-119
    out[r].Updated = true;
    if (q.Spec.Workload.Name.empty() || q.Spec.Workload.Namespace.empty()) {       // :145-150
      out[r].Phase = "Failed";
      out[r].Message = "workload name/namespace 不能为空";
      continue;
    }
    live.push_back(r);
    min_battery.push_back(q.Spec.MinBatteryPercent);
    seen.clear();                                                                  // preferredSet, :175-178
    for (const std::string& p : q.Spec.PreferredNodes) {
      std::string lp = ToLower(p);
      if (std::find(seen.begin(), seen.end(), lp) != seen.end()) continue;
      auto it = by_name.find(lp);
      if (it != by_name.end())
        for (int32_t i = it->second; i >= 0; i = next_same[i]) pref_ids.push_back(i);
      seen.push_back(std::move(lp));
    }
    pref_off.push_back(static_cast<int32_t>(pref_ids.size()));
  }
  if (live.empty()) return out;

  // ---- the GPU path: one snapshot, one batch
  check(klm_nodes_set(h_, n, battery.data(), flags.data()), "klm_nodes_set");
  std::vector<int32_t> idx(live.size());
  std::vector<double> score(live.size());
  if (pref_ids.empty()) pref_ids.push_back(0);
  check(klm_schedule_batch(h_, static_cast<int64_t>(live.size()), min_battery.data(), pref_off.data(),
                           pref_ids.data(), idx.data(), score.data(), KLM_ALGO_DECOMPOSED),
        "klm_schedule_batch");

  for (size_t k = 0; k < live.size(); ++k) {
    SchedulingRequestStatus& s = out[live[k]];
    if (idx[k] < 0) {                                                              // :153-158
      s.Phase = "Failed";
      s.Message = "无满足要求的 UAV 节点";
      continue;
    }
    const UAVMetricItem& c = uavs[idx[k]];
    s.Phase = "Assigned";                                                          // :163-169
    s.AssignedNode = c.NodeName;
    s.AssignedUAV = c.UAVID;
    s.Score = score[k];
    s.Message = sprintf1("选中节点 %s (电量 %.1f%%)", c.NodeName.c_str(), battery[idx[k]]);
  }
  return out;
}

// ---- multi-assignment with per-node capacity (NO reference semantics: the
// reference assigns every request independently and ignores e.g. gpu-count). Requests are taken in
// list order; a node can be chosen `capacity[i]` times in this reconcile and is no candidate once it
// is full. One GPU call gives every request its k best candidates on the snapshot; walking them in
// order, the first candidate that still has room is exactly what a sequential buildCandidates over
// the remaining nodes would choose (every node ranked before it is in the list and full). Only when
// a whole list of k is full are the remaining requests re-batched against the snapshot without
// the full nodes.
std::vector<SchedulingRequestStatus> Controller::ReconcileWithCapacity(
    const std::vector<SchedulingRequest>& requests, const std::vector<UAVMetricItem>& uavs,
    std::vector<int32_t> capacity) {
  const int64_t n = static_cast<int64_t>(uavs.size());
  if (static_cast<int64_t>(capacity.size()) != n) throw std::runtime_error("capacity: one entry per UAV item");
  std::vector<double> battery(n);
  std::vector<uint8_t> flags(n);
  std::unordered_map<std::string, std::vector<int32_t>> by_name;
  for (int64_t i = 0; i < n; ++i) {
    const UAVMetricItem& u = uavs[i];
    const std::string name = u.HasNodeName ? u.NodeName : std::string();
    battery[i] = u.HasBattery ? u.Battery : 0.0;
    const std::string st = u.HasStatus ? ToLower(u.CollectionStatus) : std::string();
    uint8_t f = 0;
    if (!name.empty()) f |= KLM_NODE_HAS_NAME;
    if (st.empty() || st == "active") f |= KLM_NODE_STATUS_OK;
    flags[i] = f;
    if (!name.empty()) by_name[ToLower(name)].push_back(static_cast<int32_t>(i));
  }
  std::vector<SchedulingRequestStatus> out(requests.size());
  std::vector<size_t> live;
  std::vector<double> min_battery;
  std::vector<int32_t> pref_off{0}, pref_ids;
  for (size_t r = 0; r < requests.size(); ++r) {
    const SchedulingRequest& q = requests[r];
    if (q.HasPhase && !q.Phase.empty() && q.Phase != "Pending") continue;
    out[r].Updated = true;
    if (q.Spec.Workload.Name.empty() || q.Spec.Workload.Namespace.empty()) {
      out[r].Phase = "Failed";
      out[r].Message = "workload name/namespace 不能为空";
      continue;
    }
    live.push_back(r);
    min_battery.push_back(q.Spec.MinBatteryPercent);
    std::unordered_map<std::string, bool> seen;
    for (const std::string& p : q.Spec.PreferredNodes) {
      const std::string lp = ToLower(p);
      if (seen.emplace(lp, true).second) {
        auto it = by_name.find(lp);
        if (it != by_name.end()) pref_ids.insert(pref_ids.end(), it->second.begin(), it->second.end());
      }
    }
    pref_off.push_back(static_cast<int32_t>(pref_ids.size()));
  }
  if (live.empty()) return out;
  if (pref_ids.empty()) pref_ids.push_back(0);
  const int k = klm_topk_max();
  std::vector<int32_t> idx;
  std::vector<double> score;
  size_t batch_from = 0;  // the candidate lists in idx / score start at this live request
  auto rebatch = [&](size_t from) {
    // the snapshot without the nodes that are full; offsets re-based to the first remaining request
    std::vector<uint8_t> f2(flags);
    for (int64_t i = 0; i < n; ++i)
      if (capacity[i] <= 0) f2[i] = static_cast<uint8_t>(f2[i] & ~KLM_NODE_STATUS_OK);
    const size_t m = live.size() - from;
    std::vector<int32_t> off2(m + 1);
    for (size_t j = 0; j <= m; ++j) off2[j] = pref_off[from + j] - pref_off[from];
    idx.assign(m * k, -1);
    score.assign(m * k, 0.0);
    check(klm_nodes_set(h_, n, battery.data(), f2.data()), "klm_nodes_set");
    check(klm_schedule_batch_topk(h_, static_cast<int64_t>(m), min_battery.data() + from, off2.data(),
                                  pref_ids.data() + pref_off[from], k, idx.data(), score.data()),
          "klm_schedule_batch_topk");
    batch_from = from;
  };
  rebatch(0);
  for (size_t q = 0; q < live.size(); ++q) {
    SchedulingRequestStatus& s = out[live[q]];
    int32_t chosen = -1;
    double chosen_score = 0;
    for (int attempt = 0; attempt < 2 && chosen < 0; ++attempt) {
      const int32_t* li = idx.data() + (q - batch_from) * k;
      const double* ls = score.data() + (q - batch_from) * k;
      int listed = 0;
      for (int j = 0; j < k && li[j] >= 0; ++j) {
        ++listed;
        if (capacity[li[j]] > 0) {
          chosen = li[j];
          chosen_score = ls[j];
          break;
        }
      }
      if (chosen >= 0 || listed < k || attempt == 1) break;  // a short list names every candidate there is
      rebatch(q);  // all k are full: look further down, for this request and the ones after it
    }
    if (chosen < 0) {
      s.Phase = "Failed";
      s.Message = "无满足要求的 UAV 节点";
      continue;
    }
    --capacity[chosen];
    const UAVMetricItem& c = uavs[chosen];
    s.Phase = "Assigned";
    s.AssignedNode = c.NodeName;
    s.AssignedUAV = c.UAVID;
    s.Score = chosen_score;
    s.Message = sprintf1("选中节点 %s (电量 %.1f%%)", c.NodeName.c_str(), battery[chosen]);
  }
  return out;
}

ClusterMetrics Manager::CalculateClusterMetrics(const std::vector<NodeMetrics>& nodes,
                                                const std::vector<PodMetrics>& pods) {
  const int64_t n = static_cast<int64_t>(nodes.size()), p = static_cast<int64_t>(pods.size());
  std::vector<int64_t> cc(n), cu(n), mc(n), mu(n);
  std::vector<int32_t> gc(n);
  std::vector<uint8_t> healthy(n), running(p);
  std::vector<double> gpu;
  for (int64_t i = 0; i < n; ++i) {
    cc[i] = nodes[i].CPUCapacity; cu[i] = nodes[i].CPUUsage;
    mc[i] = nodes[i].MemoryCapacity; mu[i] = nodes[i].MemoryUsage;
    gc[i] = nodes[i].GPUCount;
    healthy[i] = nodes[i].Healthy ? 1 : 0;
    gpu.insert(gpu.end(), nodes[i].GPUUsage.begin(), nodes[i].GPUUsage.end());
  }
  for (int64_t i = 0; i < p; ++i) running[i] = pods[i].Phase == "Running" ? 1 : 0;  // manager.go:509
  klm_node_cols cols{cc.data(), cu.data(), mc.data(), mu.data(), gc.data(), healthy.data(),
                     gpu.data(), static_cast<int64_t>(gpu.size())};
  klm_cluster_out o{};
  check(klm_cluster_aggregate(h_, &cols, n, running.data(), p, &o), "klm_cluster_aggregate");

  ClusterMetrics m;
  m.TotalNodes = o.total_nodes; m.HealthyNodes = o.healthy_nodes;
  m.TotalPods = o.total_pods; m.RunningPods = o.running_pods;
  m.TotalCPU = o.total_cpu; m.UsedCPU = o.used_cpu;
  m.TotalMemory = o.total_memory; m.UsedMemory = o.used_memory;
  m.TotalGPUs = o.total_gpus; m.AvailableGPUs = o.available_gpus;
  m.CPUUsageRate = o.cpu_usage_rate; m.MemoryUsageRate = o.memory_usage_rate;
  char buf[128];
  if (o.issues & KLM_ISSUE_UNHEALTHY_NODES) {                                        // manager.go:547-549
    std::snprintf(buf, sizeof(buf), "%lld nodes are unhealthy",
                  static_cast<long long>(m.TotalNodes - m.HealthyNodes));
    m.Issues.push_back(buf);
  }
  if (o.issues & KLM_ISSUE_HIGH_CPU) {                                               // :550-552
    std::snprintf(buf, sizeof(buf), "High CPU usage: %.1f%%", m.CPUUsageRate);
    m.Issues.push_back(buf);
  }
  if (o.issues & KLM_ISSUE_HIGH_MEMORY) {                                            // :553-555
    std::snprintf(buf, sizeof(buf), "High memory usage: %.1f%%", m.MemoryUsageRate);
    m.Issues.push_back(buf);
  }
  static const char* kStatus[] = {"healthy", "warning", "critical"};                 // :558-564
  m.HealthStatus = kStatus[o.health_status];
  return m;
}

// ---- status write-back (controller.go:223-250): one UpdateStatus per request in the reference,
// issued one after the other from the reconcile goroutine; with scoring down to microseconds the
// round trips are the whole reconcile, so they are issued from `workers` threads at once. The
// status objects are independent API objects, so their order does not matter.
std::vector<WriteResult> WriteStatuses(const std::vector<StatusUpdate>& updates, const std::string& now_rfc3339,
                                       const UpdateStatusFn& update_status, int workers) {
  std::vector<WriteResult> res(updates.size());
  std::atomic<size_t> next{0};
  auto run = [&]() {
    for (size_t i = next.fetch_add(1); i < updates.size(); i = next.fetch_add(1)) {
      const StatusUpdate& u = updates[i];
      if (!u.Status.Updated) {  // skipped by the phase gate (:117-119): nothing to write
        res[i].Ok = true;
        continue;
      }
      std::string err;
      res[i].Written = true;
      res[i].Ok = update_status(u.Namespace, u.Name, EncodeStatus(u.Status, now_rfc3339), &err);  // :229-248
      if (!res[i].Ok) res[i].Error = err;
    }
  };
  const int w = workers < 1 ? 1 : (workers > 256 ? 256 : workers);
  std::vector<std::thread> pool;
  for (int t = 1; t < w && static_cast<size_t>(t) < updates.size(); ++t) pool.emplace_back(run);
  run();
  for (std::thread& t : pool) t.join();
  return res;
}

void Manager::EnableUAVRing(int64_t max_nodes, int32_t window) {
  check(klm_ring_create(h_, max_nodes * kUAVChannels, window), "klm_ring_create");
  ring_nodes_ = max_nodes;
}

int64_t Manager::UpdateUAVReport(UAVReport report, const std::string& now_rfc3339) {
  if (!ApplyUAVReportDefaults(report, now_rfc3339)) return -1;  // manager.go:392-409
  // m.uavSnapshot[report.NodeName] = entry (:437-442): one entry per node name
  auto it = std::lower_bound(uav_index_.begin(), uav_index_.end(), report.NodeName,
                             [](const std::pair<std::string, int64_t>& a, const std::string& k) { return a.first < k; });
  int64_t idx;
  if (it != uav_index_.end() && it->first == report.NodeName) {
    idx = it->second;
  } else {
    idx = static_cast<int64_t>(uav_.size());
    uav_index_.insert(it, {report.NodeName, idx});
    uav_.emplace_back();
  }
  UAVEntry& e = uav_[static_cast<size_t>(idx)];
  e.NodeName = report.NodeName; e.UAVID = report.UAVID; e.Status = report.Status; e.Source = report.Source;
  e.Timestamp = report.Timestamp; e.LastHeartbeat = report.Timestamp;   // :417-418, :445
  e.NodeIP = report.NodeIP;                                              // :420-422
  e.HeartbeatIntervalSeconds = report.HeartbeatIntervalSeconds > 0 ? report.HeartbeatIntervalSeconds : 0;  // :424-426
  e.HasState = report.HasState;                                          // :432-435
  if (report.HasState) {
    e.RemainingPercent = report.RemainingPercent;
    e.SystemStatus = report.SystemStatus;
    if (idx < ring_nodes_) {
      int64_t ids[kUAVChannels];
      for (int c = 0; c < kUAVChannels; ++c) ids[c] = idx * kUAVChannels + c;
      check(klm_ring_ingest(h_, kUAVChannels, ids, report.Channels), "klm_ring_ingest");
    }
  }
  return idx;
}

static void fleet_columns(const std::vector<UAVHealth>& fleet, std::vector<uint8_t>& ok,
                          std::vector<double>& pct) {
  ok.resize(fleet.size());
  pct.resize(fleet.size());
  for (size_t i = 0; i < fleet.size(); ++i) {
    ok[i] = fleet[i].SystemStatus == "OK" ? 1 : 0;  // uav_metrics.go:207
    pct[i] = fleet[i].RemainingPercent;
  }
}

int64_t Manager::GetHealthyUAVCount(const std::vector<UAVHealth>& fleet) {
  std::vector<uint8_t> ok;
  std::vector<double> pct;
  fleet_columns(fleet, ok, pct);
  std::vector<int32_t> low(fleet.size() + 1);
  int64_t healthy = 0, n_low = 0;
  check(klm_fleet_scan(h_, static_cast<int64_t>(fleet.size()), ok.data(), pct.data(), 0.0, &healthy,
                       low.data(), &n_low), "klm_fleet_scan");
  return healthy;
}

std::vector<std::string> Manager::GetLowBatteryUAVs(const std::vector<UAVHealth>& fleet, double threshold) {
  std::vector<uint8_t> ok;
  std::vector<double> pct;
  fleet_columns(fleet, ok, pct);
  std::vector<int32_t> low(fleet.size() + 1);
  int64_t healthy = 0, n_low = 0;
  check(klm_fleet_scan(h_, static_cast<int64_t>(fleet.size()), ok.data(), pct.data(), threshold,
                       &healthy, low.data(), &n_low), "klm_fleet_scan");
  std::vector<std::string> names;
  for (int64_t i = 0; i < n_low; ++i) names.push_back(fleet[low[i]].NodeName);
  return names;
}

}  // namespace klmhost
