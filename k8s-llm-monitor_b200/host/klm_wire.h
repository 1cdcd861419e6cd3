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

// The status map of controller.go:229-236 as JSON; empty phase becomes "Pending" (:238-240);
// last_updated is the RFC3339 UTC string the caller took from its clock.
std::string EncodeStatus(const SchedulingRequestStatus& st, const std::string& last_updated);
// a float64 as Go's encoding/json writes it (shortest round-trip digits, 'f' form for exponents in [-6, 21))
std::string FormatJSONFloat(double v);

}  // namespace klmhost
