"""Writes scheduler_kat.json: hand-derived known answers for the scheduler.

The reference has no tests for this path and cannot run here (no Go toolchain),
so nothing below is an output of the reference. Inputs are taken from the
reference's own fixtures; expected answers are worked out by hand from
internal/scheduler/controller.go:145-221 and recorded with their derivation.
  - UAV fleets: deployments/uav-simulator.yaml:33,59,101,127,169,195 (Fixture A,
    batteries 85/75/90 on server-0/agent-0/agent-1) and three Go agents at their
    initial 100 % (pkg/uav/mavlink_simulator.go:145; list order by CR name,
    internal/k8s/client.go:332) (Fixture B).
  - Requests: examples/test-scheduling-request.yaml, examples/multi-pod-request.yaml,
    examples/edge-compute-task.yaml (spec fields copied by value).
Run: python tests/golden/make_scheduler_kat.py
"""
import json
import os

S0, A0, A1 = ("k3d-k8s-llm-monitor-server-0", "k3d-k8s-llm-monitor-agent-0", "k3d-k8s-llm-monitor-agent-1")


def item(name, battery, status="active", uav_id=None):
    spec = {"uav_id": uav_id if uav_id is not None else ("uav-" + name if name else "uav-x")}
    if name is not None:
        spec["node_name"] = name
    if battery is not None:
        spec["battery"] = {"remaining_percent": battery}
    st = {"last_update": "2025-10-17T00:00:00Z"}
    if status is not None:
        st["collection_status"] = status
    return {"spec": spec, "status": st}


FIXTURE_A = [item(S0, 85.0), item(A0, 75.0), item(A1, 90.0)]
FIXTURE_B = [item(A0, 100.0), item(A1, 100.0), item(S0, 100.0)]

WL = {"name": "w", "namespace": "default", "type": "t"}
REQ_TEST = {"workload": {"name": "ai-training-job", "namespace": "default", "type": "compute-intensive"},
            "minBatteryPercent": 50.0, "preferredNodes": [A0]}
REQ_MULTI = {"workload": {"name": "multi-gpu-training", "namespace": "default", "type": "ml-training"},
             "minBatteryPercent": 60.0, "preferredNodes": [A0, A1]}
REQ_EDGE = {"workload": {"name": "image-recognition", "namespace": "default", "type": "ai-inference"},
            "minBatteryPercent": 80.0}


def assigned(idx, node, score, battery):
    return {"index": idx, "phase": "Assigned", "node": node, "score": score,
            "message": "选中节点 %s (电量 %.1f%%)" % (node, battery)}


NONE = {"index": -1, "phase": "Failed", "node": "", "score": 0.0, "message": "无满足要求的 UAV 节点"}

cases = [
    # ---- C1: the six fixture x example KATs
    dict(name="A/test-request", items=FIXTURE_A, spec=REQ_TEST, expect=assigned(2, A1, 90.0, 90.0),
         why="scores 85, 75+10=85, 90 -> agent-1 wins with 90"),
    dict(name="A/multi-pod", items=FIXTURE_A, spec=REQ_MULTI, expect=assigned(2, A1, 100.0, 90.0),
         why="scores 85, 85, 90+10=100"),
    dict(name="A/edge-task", items=FIXTURE_A, spec=REQ_EDGE, expect=assigned(2, A1, 90.0, 90.0),
         why="agent-0 (75) fails min 80; 85 vs 90"),
    dict(name="B/test-request", items=FIXTURE_B, spec=REQ_TEST, expect=assigned(0, A0, 110.0, 100.0),
         why="110, 100, 100"),
    dict(name="B/multi-pod", items=FIXTURE_B, spec=REQ_MULTI, expect=assigned(0, A0, 110.0, 100.0),
         why="110, 110, 100: tie -> stable sort keeps list order -> agent-0"),
    dict(name="B/edge-task", items=FIXTURE_B, spec=REQ_EDGE, expect=assigned(0, A0, 100.0, 100.0),
         why="three-way tie at 100 -> first in list"),
    # ---- edge cases read off the source
    dict(name="empty-node-name-skipped", items=[item("", 99.0), item("n1", 10.0)], spec={"workload": WL},
         expect=assigned(1, "n1", 10.0, 10.0), why="controller.go:190"),
    dict(name="missing-node-name-skipped", items=[item(None, 99.0), item("n1", 10.0)], spec={"workload": WL},
         expect=assigned(1, "n1", 10.0, 10.0), why="type assertion gives \"\" (:185), then :190"),
    dict(name="missing-battery-is-zero-no-gate", items=[item("n0", None)], spec={"workload": WL},
         expect=assigned(0, "n0", 0.0, 0.0), why="readFloat -> 0 (:258/:270); min 0 disables the gate (:194)"),
    dict(name="missing-battery-fails-gate", items=[item("n0", None)],
         spec={"workload": WL, "minBatteryPercent": 0.5}, expect=NONE, why="0 < 0.5"),
    dict(name="status-inactive-filtered", items=[item("n0", 90.0, "Inactive"), item("n1", 5.0, "ERROR"),
                                                 item("n2", 1.0, "ACTIVE")], spec={"workload": WL},
         expect=assigned(2, "n2", 1.0, 1.0), why="lower-cased status must be \"\" or \"active\" (:188, :198)"),
    dict(name="status-missing-accepted", items=[item("n0", 42.5, None)], spec={"workload": WL},
         expect=assigned(0, "n0", 42.5, 42.5), why="readString -> \"\" (:198)"),
    dict(name="preferred-case-insensitive", items=[item("Node-A", 50.0), item("node-b", 55.0)],
         spec={"workload": WL, "preferredNodes": ["NODE-a"]}, expect=assigned(0, "Node-A", 60.0, 50.0),
         why="both sides lower-cased (:177, :206): 60 vs 55"),
    dict(name="negative-min-disables-gate", items=[item("n0", 3.0)],
         spec={"workload": WL, "minBatteryPercent": -5.0}, expect=assigned(0, "n0", 3.0, 3.0), why=":194 min > 0"),
    dict(name="all-filtered", items=[item("n0", 10.0), item("n1", 20.0)],
         spec={"workload": WL, "minBatteryPercent": 50.0}, expect=NONE, why=":153-158"),
    dict(name="no-uavs", items=[], spec={"workload": WL}, expect=NONE, why="empty list"),
    dict(name="duplicate-names-distinct", items=[item("dup", 70.0), item("dup", 70.0), item("x", 79.9)],
         spec={"workload": WL, "preferredNodes": ["dup"]}, expect=assigned(0, "dup", 80.0, 70.0),
         why="both dups get +10 = 80 > 79.9; tie -> first"),
    dict(name="integer-battery", items=[item("n0", 77), item("n1", 76.5)], spec={"workload": WL},
         expect=assigned(0, "n0", 77.0, 77.0), why="readFloat accepts int64 (:267)"),
    dict(name="preferred-loses-on-index-tie", items=[item("a", 60.0), item("b", 50.0)],
         spec={"workload": WL, "preferredNodes": ["b"]}, expect=assigned(0, "a", 60.0, 60.0),
         why="60 vs 50+10=60: tie -> lower index a"),
    dict(name="preferred-wins-on-index-tie", items=[item("b", 50.0), item("a", 60.0)],
         spec={"workload": WL, "preferredNodes": ["b"]}, expect=assigned(0, "b", 60.0, 50.0),
         why="50+10=60 vs 60: tie -> lower index b"),
    dict(name="preferred-below-gate", items=[item("a", 40.0), item("b", 30.0)],
         spec={"workload": WL, "minBatteryPercent": 35.0, "preferredNodes": ["b"]},
         expect=assigned(0, "a", 40.0, 40.0), why="gate tests battery, not score: b (30) is out"),
    dict(name="empty-workload", items=FIXTURE_A, spec={"workload": {"name": "", "namespace": "default"}},
         expect={"index": -1, "phase": "Failed", "node": "", "score": 0.0,
                 "message": "workload name/namespace 不能为空"}, why=":145-150"),
    dict(name="message-rounding", items=[item("n0", 66.66)], spec={"workload": WL},
         expect=assigned(0, "n0", 66.66, 66.66), why="%.1f -> 66.7"),
]

out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "scheduler_kat.json")
with open(out, "w", encoding="utf-8") as f:
    json.dump({"note": "hand-derived from controller.go; NOT reference outputs (see make_scheduler_kat.py)",
               "cases": cases}, f, ensure_ascii=False, indent=1)
print("wrote", out, len(cases), "cases")
