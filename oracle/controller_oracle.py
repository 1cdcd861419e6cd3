"""String-level CPU restatement of the reference scheduler. TEST INFRASTRUCTURE ONLY.

Follows internal/scheduler/controller.go of Sabre94/k8s-llm-monitor statement by
statement on the same data shapes the Go code sees (unstructured maps decoded
from JSON), so that the host-side work that sits ABOVE the C ABI — lower-casing,
name -> index resolution, status formatting — can be checked too. Pure-Python
loops: small cases only. Only tests/ may import this module.

Parity status: the reference has no tests or golden vectors for this path and no
Go toolchain exists here, so the known answers in tests/test_oracle_kat.py are
derived by hand from the reference's fixtures ("derived, not reference-pinned").
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional


@dataclass
class SchedulingWorkload:  # pkg/models/scheduler.go:6-10
    name: str = ""
    namespace: str = ""
    type: str = ""


@dataclass
class SchedulingRequestSpec:  # pkg/models/scheduler.go:13-19
    workload: SchedulingWorkload = field(default_factory=SchedulingWorkload)
    min_battery_percent: float = 0.0
    preferred_nodes: List[str] = field(default_factory=list)


@dataclass
class SchedulingRequestStatus:  # pkg/models/scheduler.go:22-29
    phase: str = ""
    assigned_node: str = ""
    assigned_uav: str = ""
    score: float = 0.0
    message: str = ""


@dataclass
class SchedulingCandidate:  # pkg/models/scheduler.go:32-38
    node_name: str
    uav_id: str
    battery: float
    score: float
    index: int  # position in the UAV list (not in the Go struct; for the tests)


def read_float(m: Any, *fields: str) -> float:
    """controller.go:252-272. Go's type switch accepts float64, int, int64."""
    cur = m
    for f in fields:
        if isinstance(cur, dict):
            cur = cur.get(f)
        else:
            return 0.0
    if isinstance(cur, bool):  # a JSON bool is not a number in Go's switch
        return 0.0
    if isinstance(cur, (int, float)):
        return float(cur)
    return 0.0


def read_string(m: Any, *fields: str) -> str:
    """controller.go:274-288."""
    cur = m
    for f in fields:
        if isinstance(cur, dict):
            cur = cur.get(f)
        else:
            return ""
    return cur if isinstance(cur, str) else ""


def decode_spec(spec: Dict[str, Any], apimachinery_ints: bool = True) -> SchedulingRequestSpec:
    """controller.go:126-143.

    apimachinery_ints=True models k8s.io/apimachinery v0.34.1 (go.mod:10), whose
    unstructured JSON decoder yields int64 for integral numbers, so the
    `.(float64)` assertion at :133 fails for `minBatteryPercent: 50` and the
    gate stays 0 (dependency source not available here, so
    this switch is "parity unpinned"). False models a plain encoding/json decode
    (every number a float64).
    """
    out = SchedulingRequestSpec()
    wl = spec.get("workload")
    if isinstance(wl, dict):
        out.workload.name = wl["name"] if isinstance(wl.get("name"), str) else ""
        out.workload.namespace = wl["namespace"] if isinstance(wl.get("namespace"), str) else ""
        out.workload.type = wl["type"] if isinstance(wl.get("type"), str) else ""
    v = spec.get("minBatteryPercent")
    if isinstance(v, float) or (not apimachinery_ints and isinstance(v, int) and not isinstance(v, bool)):
        out.min_battery_percent = float(v)
    lst = spec.get("preferredNodes")
    if isinstance(lst, list):
        out.preferred_nodes = [s for s in lst if isinstance(s, str)]
    return out


def build_candidates(spec: SchedulingRequestSpec, uav_items: List[Dict[str, Any]]) -> List[SchedulingCandidate]:
    """controller.go:174-221."""
    preferred = {n.lower() for n in spec.preferred_nodes}  # :175-178
    out: List[SchedulingCandidate] = []
    for index, item in enumerate(uav_items):  # :181
        uav_spec = item.get("spec") if isinstance(item.get("spec"), dict) else {}
        uav_status = item.get("status") if isinstance(item.get("status"), dict) else {}
        node_name = uav_spec.get("node_name") if isinstance(uav_spec.get("node_name"), str) else ""
        uav_id = uav_spec.get("uav_id") if isinstance(uav_spec.get("uav_id"), str) else ""
        battery = read_float(uav_spec, "battery", "remaining_percent")  # :187
        status = read_string(uav_status, "collection_status").lower()  # :188
        if node_name == "":  # :190
            continue
        if spec.min_battery_percent > 0 and battery < spec.min_battery_percent:  # :194
            continue
        if status != "" and status != "active":  # :198
            continue
        score = battery  # :205
        if node_name.lower() in preferred:  # :206
            score += 10
        out.append(SchedulingCandidate(node_name, uav_id, battery, score, index))
    return out


def process_request(spec: SchedulingRequestSpec, uav_items: List[Dict[str, Any]]) -> SchedulingRequestStatus:
    """controller.go:145-171 (after the phase gate and the decode)."""
    if spec.workload.name == "" or spec.workload.namespace == "":  # :145
        return SchedulingRequestStatus(phase="Failed", message="workload name/namespace 不能为空")
    cands = build_candidates(spec, uav_items)  # :152
    if not cands:  # :153
        return SchedulingRequestStatus(phase="Failed", message="无满足要求的 UAV 节点")
    # :160 sort.SliceStable descending by score; Python's sort is stable too.
    cands = sorted(cands, key=lambda c: -c.score)
    chosen = cands[0]  # :161
    return SchedulingRequestStatus(  # :163-169
        phase="Assigned",
        assigned_node=chosen.node_name,
        assigned_uav=chosen.uav_id,
        score=chosen.score,
        message="选中节点 %s (电量 %.1f%%)" % (chosen.node_name, chosen.battery),
    )


def chosen_index(spec: SchedulingRequestSpec, uav_items: List[Dict[str, Any]]) -> int:
    cands = build_candidates(spec, uav_items)
    if not cands:
        return -1
    return sorted(cands, key=lambda c: -c.score)[0].index


def uav_item(node_name: Optional[str], battery: Any = None, status: Optional[str] = "active",
             uav_id: str = "") -> Dict[str, Any]:
    """One UAVMetric item in the wire format internal/k8s/client.go:355-392 writes."""
    spec: Dict[str, Any] = {}
    if node_name is not None:
        spec["node_name"] = node_name
    spec["uav_id"] = uav_id
    if battery is not None:
        spec["battery"] = {"remaining_percent": battery}
    st: Dict[str, Any] = {"last_update": "2025-10-17T00:00:00Z"}
    if status is not None:
        st["collection_status"] = status
    return {"apiVersion": "monitoring.io/v1", "kind": "UAVMetric", "spec": spec, "status": st}
