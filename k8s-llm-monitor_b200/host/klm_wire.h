// Wire format on either side of the scheduling path:
//   in : the item lists the controller gets from the API server — UAVMetric CRs as
//        written by internal/k8s/client.go:355-392, SchedulingRequest CRs as in
//        examples/*.yaml / deployments/scheduling-crd.yaml — as JSON text;
//   out: the status map controller.go:229-236 writes back.
// A small JSON reader (objects, arrays, strings with escapes, numbers, true/false/null)
// keeps the distinction Go's unstructured decoder makes between integral numbers (int64)
// and the rest (float64): `spec["minBatteryPercent"].(float64)` at controller.go:133 fails
// for `50` and succeeds for `50.0` / `5e1` — the quirk is reproduced, not fixed.
#pragma once

#include <string>
#include <vector>

#include "klm_host.h"

namespace klmhost {

// {"items":[{...UAVMetric...}, ...]} or a bare array. Throws std::runtime_error on bad JSON.
std::vector<UAVMetricItem> DecodeUAVMetricList(const std::string& json);

// Same for SchedulingRequest items: status.phase (controller.go:113), spec.workload,
// spec.minBatteryPercent (float64 only, :133), spec.preferredNodes (strings only, :137-143).
// apimachinery_ints = false treats every number as float64 (plain encoding/json).
std::vector<SchedulingRequest> DecodeSchedulingRequestList(const std::string& json,
                                                           bool apimachinery_ints = true);

// ---- UAV push ingestion
// models.UAVReport (pkg/models/models.go:182-192) as the agent POSTs it to /api/v1/uav/report
// (cmd/uav-agent/main.go:373-386). `Channels` is the UAVState (pkg/uav/mavlink_simulator.go:11-106)
// flattened into the 64-wide metric ring row of the node, in the order of SURVEY.md 8a A11:
//   0-8   gps: latitude, longitude, altitude, relative_altitude, hdop, satellite_count, fix_type,
//         ground_speed, course_over_ground
//   9-14  attitude: roll, pitch, yaw, roll_rate, pitch_rate, yaw_rate
//   15-18 flight: airspeed, ground_speed, vertical_speed, throttle_percent
//   19-26 battery: voltage, current, remaining_percent, remaining_capacity, total_capacity,
//         temperature, cell_count, time_remaining
//   27-30 mission: current_waypoint, total_waypoints, distance_to_wp, eta_to_wp
//   31-32 health: error_count, warning_count
//   33    flight.armed (0 / 1)
//   34-39 health.sensors_health: accelerometer, barometer, battery, compass, gps, gyroscope (0 / 1)
//   40-42 flight.mode (MANUAL 0, STABILIZE 1, LOITER 2, AUTO 3, RTL 4, LAND 5), mission.mission_state
//         (IDLE 0, ACTIVE 1, PAUSED 2, COMPLETED 3), health.system_status (OK 0, WARNING 1,
//         CRITICAL 2, ERROR 3); an unknown string gives -1
//   43-63 reserved, 0
// Channel 21 (battery.remaining_percent) is the battery the scheduler scores.
constexpr int kUAVChannels = 64;
constexpr int kUAVBatteryChannel = 21;
struct UAVReport {
  std::string NodeName, NodeIP, UAVID, Source, Status, Timestamp;  // Timestamp: RFC3339 text as sent ("" = zero)
  int64_t HeartbeatIntervalSeconds = 0;
  bool HasState = false;
  double RemainingPercent = 0;       // state.battery.remaining_percent as float64 (node table column)
  std::string SystemStatus;          // state.health.system_status (fleet scans, uav_metrics.go:207)
  float Channels[kUAVChannels] = {};
};
// encoding/json into the typed struct: absent fields keep their zero value; a field of the wrong
// JSON type is an error (Go: "json: cannot unmarshal ..."; the handler answers 400,
// cmd/server/main.go:580-583). Throws std::runtime_error.
UAVReport DecodeUAVReport(const std::string& json);
// The defaulting of Manager.UpdateUAVReport (manager.go:392-409): returns false when the report is
// dropped (empty node_name); otherwise Status "" -> "active", Source "" -> "agent", and an empty
// Timestamp becomes `now_rfc3339`.
bool ApplyUAVReportDefaults(UAVReport& r, const std::string& now_rfc3339);

// The status map of controller.go:229-236 as JSON; empty phase becomes "Pending" (:238-240);
// last_updated is the RFC3339 UTC string the caller took from its clock.
std::string EncodeStatus(const SchedulingRequestStatus& st, const std::string& last_updated);
// a float64 as Go's encoding/json writes it (shortest round-trip digits, 'f' form for exponents in [-6, 21))
std::string FormatJSONFloat(double v);

}  // namespace klmhost
