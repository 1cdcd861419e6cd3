// C++ host layer above the C ABI: the reference's scheduler controller and metrics
// manager logic that sits around the GPU path, with the reference's names, argument
// meaning, status strings and failure behaviour (the reference is Go; no Go toolchain
// exists in this build environment, so the host side is C++ — see INTEGRATION.md for the
// cgo equivalent). All numeric work goes through libklm.so; nothing here scores,
// sorts or sums.
//
// Reference: internal/scheduler/controller.go, internal/metrics/manager.go,
// internal/metrics/sources/uav_metrics.go, pkg/models/scheduler.go, pkg/metrics/types.go
// of Sabre94/k8s-llm-monitor.
#pragma once

#include <cstdint>
#include <functional>
#include <string>
#include <utility>
#include <vector>

#include "../../include/klm.h"

namespace klmhost {

// ---- pkg/models/scheduler.go:6-29
struct SchedulingWorkload {
  std::string Name, Namespace, Type;
};
struct SchedulingRequestSpec {
  SchedulingWorkload Workload;
  double MinBatteryPercent = 0;  // as decoded at controller.go:133 (0 if the assertion failed)
  std::vector<std::string> PreferredNodes;
};
struct SchedulingRequestStatus {
  std::string Phase, AssignedNode, AssignedUAV;
  double Score = 0;
  std::string Message;
  bool Updated = false;  // false: request skipped by the phase gate (controller.go:117-119)
};

// One UAVMetric item as buildCandidates reads it (controller.go:182-188, :252-288).
struct UAVMetricItem {
  bool HasNodeName = false;    // spec.node_name present and a string
  std::string NodeName, UAVID;
  bool HasBattery = false;     // spec.battery.remaining_percent is float64 / int / int64
  double Battery = 0;          // readFloat result (0 when missing)
  bool HasStatus = false;      // status.collection_status present and a string
  std::string CollectionStatus;
};

struct SchedulingRequest {
  bool HasPhase = false;       // status.phase found
  std::string Phase;
  SchedulingRequestSpec Spec;
};

// Go's strings.ToLower over UTF-8 for ASCII, Latin-1, Latin Extended-A, Greek and Cyrillic
// (other scripts pass through unchanged); invalid bytes of a non-ASCII string become U+FFFD as in Go.
std::string ToLower(const std::string& s);

// internal/scheduler/controller.go Controller: reconcile over one snapshot.
class Controller {
 public:
  explicit Controller(klm_handle* h) : h_(h) {}
  // controller.go:88-172 for every request against `uavs`: phase gate, workload check,
  // ONE batched klm_schedule_batch for everything else, status build. Throws
  // std::runtime_error with klm_last_error() on a library failure ("reconcile failed",
  // retried next tick by the caller like controller.go:75-77).
  std::vector<SchedulingRequestStatus> Reconcile(const std::vector<SchedulingRequest>& requests,
                                                 const std::vector<UAVMetricItem>& uavs);
  // Multi-assignment (opt-in, no reference semantics): requests in list order,
  // node i can be assigned capacity[i] times in this reconcile. Built on klm_schedule_batch_topk.
  std::vector<SchedulingRequestStatus> ReconcileWithCapacity(const std::vector<SchedulingRequest>& requests,
                                                             const std::vector<UAVMetricItem>& uavs,
                                                             std::vector<int32_t> capacity);

 private:
  klm_handle* h_;
};

// ---- status write-back (controller.go:223-250)
struct StatusUpdate {
  std::string Namespace, Name;  // req.GetNamespace() / req.GetName()
  SchedulingRequestStatus Status;
};
struct WriteResult {
  bool Ok = false;       // the update succeeded, or there was nothing to write
  bool Written = false;  // update_status was called
  std::string Error;
};
// The apiserver call itself (dynamic.Resource(...).Namespace(ns).UpdateStatus, :246-248) belongs to
// the host; it is injected: returns false and fills *err on failure.
using UpdateStatusFn = std::function<bool(const std::string& ns, const std::string& name,
                                          const std::string& status_json, std::string* err)>;
// Encodes every updated status (EncodeStatus: the map of :229-236, "Pending" for an empty phase,
// lastUpdated = now_rfc3339) and issues the updates from `workers` threads at once.
std::vector<WriteResult> WriteStatuses(const std::vector<StatusUpdate>& updates, const std::string& now_rfc3339,
                                       const UpdateStatusFn& update_status, int workers);

// ---- pkg/metrics/types.go (fields calculateClusterMetrics reads)
struct NodeMetrics {
  std::string NodeName;
  int64_t CPUCapacity = 0, CPUUsage = 0, MemoryCapacity = 0, MemoryUsage = 0;
  int32_t GPUCount = 0;
  std::vector<double> GPUUsage;
  bool Healthy = false;
};
struct PodMetrics {
  std::string Phase;
};
struct ClusterMetrics {  // types.go:114-139
  int64_t TotalNodes = 0, HealthyNodes = 0, TotalPods = 0, RunningPods = 0;
  int64_t TotalCPU = 0, UsedCPU = 0, TotalMemory = 0, UsedMemory = 0, TotalGPUs = 0, AvailableGPUs = 0;
  double CPUUsageRate = 0, MemoryUsageRate = 0;
  std::string HealthStatus;
  std::vector<std::string> Issues;
};

// One UAV state as the fleet scans read it (uav_metrics.go:197-233).
struct UAVHealth {
  std::string NodeName, SystemStatus;
  double RemainingPercent = 0;
};

// What Manager.UpdateUAVReport keeps per node (the map entry of manager.go:411-435).
struct UAVEntry {
  std::string NodeName, UAVID, Status, Source, Timestamp, LastHeartbeat, NodeIP, SystemStatus;
  int64_t HeartbeatIntervalSeconds = 0;
  bool HasState = false;
  double RemainingPercent = 0;  // state.battery.remaining_percent: the battery the scheduler scores
};
struct UAVReport;  // klm_wire.h

// internal/metrics/manager.go Manager: the aggregation half.
class Manager {
 public:
  explicit Manager(klm_handle* h) : h_(h) {}
  // The 64-channel metric ring behind UpdateUAVReport: room for max_nodes UAV nodes (series
  // node * 64 + channel), `window` samples each. No reference counterpart (F5: the reference keeps
  // only the latest state).
  void EnableUAVRing(int64_t max_nodes, int32_t window);
  // manager.go:391-449: defaulting, latest entry per node, last heartbeat; and -- new -- the
  // state's 64 channels appended to the node's ring rows (klm_ring_ingest; queued by the library
  // while collect cycles are in flight). Returns the node's index (first-seen order), or -1 when
  // the report is dropped (nil / empty node_name, :392-394).
  int64_t UpdateUAVReport(UAVReport report, const std::string& now_rfc3339);
  const std::vector<UAVEntry>& UAVSnapshot() const { return uav_; }
  // manager.go:493-565
  ClusterMetrics CalculateClusterMetrics(const std::vector<NodeMetrics>& nodes,
                                         const std::vector<PodMetrics>& pods);
  // uav_metrics.go:197-214
  int64_t GetHealthyUAVCount(const std::vector<UAVHealth>& fleet);
  // uav_metrics.go:216-233; names in fleet order (Go returns map order)
  std::vector<std::string> GetLowBatteryUAVs(const std::vector<UAVHealth>& fleet, double threshold);

 private:
  klm_handle* h_;
  std::vector<UAVEntry> uav_;            // m.uavSnapshot, in first-seen order
  std::vector<std::string> uav_names_;   // index -> node name (lookup is linear over a sorted copy below)
  std::vector<std::pair<std::string, int64_t>> uav_index_;  // sorted by name
  int64_t ring_nodes_ = 0;
};

}  // namespace klmhost
