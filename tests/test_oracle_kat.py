"""CPU: pins the oracle (string-level and flat) on the hand-derived KATs and proves the
three flat algorithms equal. No GPU."""
import numpy as np
import pytest

import controller_oracle as co
import helpers
import oracle_binding as ob

KATS = helpers.load_kats()


@pytest.mark.parametrize("case", KATS, ids=[c["name"] for c in KATS])
def test_string_oracle_kat(case):
    spec = co.decode_spec(case["spec"], apimachinery_ints=False)
    st = co.process_request(spec, case["items"])
    e = case["expect"]
    assert st.phase == e["phase"]
    assert st.assigned_node == e["node"]
    assert st.score == e["score"]
    assert st.message == e["message"]
    if e["phase"] == "Assigned":
        assert co.chosen_index(spec, case["items"]) == e["index"]


@pytest.mark.parametrize("case", [c for c in KATS if c["name"] != "empty-workload"],
                         ids=[c["name"] for c in KATS if c["name"] != "empty-workload"])
def test_flat_oracles_kat(case):
    spec = co.decode_spec(case["spec"], apimachinery_ints=False)
    bat, flags, names = helpers.flatten_items(case["items"])
    mb, off, ids = helpers.flatten_specs([spec], names)
    for fn in (ob.schedule_literal, ob.schedule_scan, ob.schedule_decomposed):
        idx, score = fn(bat, flags, mb, off, ids)
        assert idx[0] == case["expect"]["index"], fn.__name__
        assert score[0] == case["expect"]["score"], fn.__name__


def test_apimachinery_int_quirk():
    """integral minBatteryPercent decodes as int64 under apimachinery and the
    .(float64) assertion at controller.go:133 fails -> gate off. (Dependency source absent:
    this models documented behaviour, parity unpinned.)"""
    spec = {"workload": {"name": "w", "namespace": "d"}, "minBatteryPercent": 80}
    assert co.decode_spec(spec, apimachinery_ints=True).min_battery_percent == 0.0
    assert co.decode_spec(spec, apimachinery_ints=False).min_battery_percent == 80.0
    spec["minBatteryPercent"] = 80.5
    assert co.decode_spec(spec, apimachinery_ints=True).min_battery_percent == 80.5


@pytest.mark.parametrize("seed", range(8))
def test_literal_scan_decomposed_agree(seed):
    rng = np.random.default_rng(seed)
    n = int(rng.integers(1, 400))
    bat, flags, mb, off, ids = helpers.random_problem(rng, n, 300, grid=rng.choice([0.2, 1.0, 10.0]))
    a = ob.schedule_literal(bat, flags, mb, off, ids)
    b = ob.schedule_scan(bat, flags, mb, off, ids)
    c = ob.schedule_decomposed(bat, flags, mb, off, ids)
    d = ob.schedule_scan(bat, flags, mb, off, ids, threads=3)
    for other in (b, c, d):
        assert np.array_equal(a[0], other[0])
        assert np.array_equal(a[1], other[1])


def test_flat_matches_string_oracle_random():
    rng = np.random.default_rng(7)
    names_pool = ["Node-%d" % i for i in range(12)] + ["", "dup", "DUP"]
    for _ in range(60):
        n = int(rng.integers(0, 9))
        items = []
        for _ in range(n):
            nm = names_pool[int(rng.integers(len(names_pool)))]
            bt = [None, float(np.round(rng.random() * 100, 1)), int(rng.integers(0, 100))][int(rng.integers(3))]
            stt = [None, "active", "ACTIVE", "Active", "inactive", ""][int(rng.integers(6))]
            items.append(co.uav_item(nm, bt, stt))
        spec = co.SchedulingRequestSpec(co.SchedulingWorkload("w", "d"),
                                        float(rng.choice([0.0, 30.0, 50.5, 90.0])),
                                        [names_pool[int(i)].upper() for i in rng.integers(0, 15, size=int(rng.integers(0, 3)))])
        bat, flags, names = helpers.flatten_items(items)
        mb, off, ids = helpers.flatten_specs([spec], names)
        idx, score = ob.schedule_literal(bat, flags, mb, off, ids)
        assert idx[0] == co.chosen_index(spec, items)
        st = co.process_request(spec, items)
        assert score[0] == st.score


# ---------------------------------------------------------------- cluster aggregate KATs

def _snap(cpu_cap, cpu_use, mem_cap, mem_use, healthy, gpu_count=None, gpu_usage=(), pods=()):
    n = len(cpu_cap)
    return {"cpu_capacity": np.array(cpu_cap, np.int64), "cpu_usage": np.array(cpu_use, np.int64),
            "memory_capacity": np.array(mem_cap, np.int64), "memory_usage": np.array(mem_use, np.int64),
            "gpu_count": np.array(gpu_count if gpu_count is not None else [0] * n, np.int32),
            "healthy": np.array(healthy, np.uint8), "gpu_usage": np.array(gpu_usage, np.float64),
            "pod_running": np.array(pods, np.uint8)}


G = 1 << 30


def test_cluster_aggregate_kat_warning():
    # by hand from manager.go:493-565
    o = ob.cluster_aggregate(_snap([4000, 4000, 8000, 8000], [1000, 2000, 3000, 4000], [8 * G] * 4,
                                   [1 * G, 2 * G, 3 * G, 4 * G], [1, 1, 1, 0], [2, 0, 1, 0],
                                   [10.0, 50.0, 49.9], [1, 1, 0]))
    assert (o["total_nodes"], o["healthy_nodes"], o["total_pods"], o["running_pods"]) == (4, 3, 3, 2)
    assert (o["total_cpu"], o["used_cpu"]) == (24000, 10000)
    assert (o["total_memory"], o["used_memory"]) == (32 * G, 10 * G)
    assert (o["total_gpus"], o["available_gpus"]) == (3, 2)  # 50.0 is not < 50.0
    assert o["cpu_usage_rate"] == 10000 / 24000 * 100.0
    assert o["memory_usage_rate"] == 31.25
    assert o["issues"] == 1 and o["health_status"] == 1  # 3 < 4//2 is false -> warning


def test_cluster_aggregate_kat_integer_division():
    # 5 nodes: Total/2 == 2 (integer division, manager.go:560)
    base = dict(cpu_cap=[1000] * 5, cpu_use=[100] * 5, mem_cap=[G] * 5, mem_use=[0] * 5)
    o = ob.cluster_aggregate(_snap(healthy=[1, 1, 0, 0, 0], **base))
    assert o["health_status"] == 1  # 2 < 2 false -> warning
    o = ob.cluster_aggregate(_snap(healthy=[1, 0, 0, 0, 0], **base))
    assert o["health_status"] == 2  # 1 < 2 -> critical


def test_cluster_aggregate_kat_thresholds():
    o = ob.cluster_aggregate(_snap([1000], [800], [G], [0], [1]))
    assert o["cpu_usage_rate"] == 80.0 and o["issues"] == 0 and o["health_status"] == 0  # > 80 is strict
    o = ob.cluster_aggregate(_snap([1000], [850], [G], [0], [1]))
    assert o["issues"] == 2 and o["health_status"] == 1
    o = ob.cluster_aggregate(_snap([1000], [950], [G], [G], [1]))
    assert o["issues"] == 6 and o["health_status"] == 2
    o = ob.cluster_aggregate(_snap([], [], [], [], []))
    assert o["total_nodes"] == 0 and o["cpu_usage_rate"] == 0.0 and o["health_status"] == 0


def test_rates_kat():
    snap = {"cpu_capacity": [4000, 0], "cpu_usage": [1000, 5], "memory_capacity": [8 * G, 8 * G],
            "memory_usage": [7 * G, G], "disk_capacity": [100, 100], "disk_allocatable": [5, 200]}
    o = ob.node_rates(snap)
    assert list(o["cpu_rate"]) == [25.0, 0.0]            # capacity 0 -> 0 (node_metrics.go:127)
    assert list(o["memory_rate"]) == [87.5, 12.5]
    assert list(o["disk_usage"]) == [95, 0]              # clamped at 0 (:121-123)
    assert list(o["disk_rate"]) == [95.0, 0.0]
    assert list(o["under_pressure"]) == [1, 0]
    pods = {"ctr_off": [0, 2, 3, 3], "cpu_usage": [100, 200, 90], "memory_usage": [10, 20, 5],
            "cpu_request": [100, 100, 0], "cpu_limit": [200, 200, 100], "memory_request": [10, 0, 0],
            "memory_limit": [0, 0, 10]}
    p = ob.pod_rates(pods)
    assert p["pod_totals"][:, 0].tolist() == [300, 30, 200, 400, 10, 0]
    assert p["cpu_rate"].tolist() == [75.0, 90.0, 0.0]
    assert p["memory_rate"].tolist() == [0.0, 50.0, 0.0]     # limit 0 -> 0 (pod_metrics.go:167)
    assert p["cpu_util"].tolist() == [150.0, 0.0, 0.0]
    assert p["over_limit"].tolist() == [0, 1, 0]             # 90 >= int64(100*0.9)=90
    h, low = ob.fleet_scan([1, 0, 1, 1], [19.9, 20.0, 5.0, 50.0], 20.0)
    assert h == 3 and low.tolist() == [0, 2]


# ---------------------------------------------------------------- window stats (definition; parity unpinned)

def test_window_stats_kat():
    t = 3600
    rows = np.zeros((5, t), np.float32)
    rows[0, :20] = np.arange(20, 0, -1)          # 1..20 reversed, n = 20
    rows[1, :1] = -7.5                            # n = 1
    rows[2, :100] = np.arange(100)                # n = 100
    rows[3] = np.arange(t)[::-1]                  # n = 3600
    seen = np.array([20, 1, 100, t + 5, 0], np.uint32)
    for fast in (False, True):
        o = ob.window_stats(rows, seen, t, fast=fast, threads=2)
        assert o["p95"].tolist() == [19.0, -7.5, 94.0, 3419.0, 0.0]  # sorted[ceil(.95 n) - 1]
        assert o["min"].tolist() == [1.0, -7.5, 0.0, 0.0, 0.0]
        assert o["max"].tolist() == [20.0, -7.5, 99.0, 3599.0, 0.0]
        assert o["mean"].tolist() == [10.5, -7.5, 49.5, 1799.5, 0.0]


def test_window_stats_fast_equals_canonical():
    rng = np.random.default_rng(3)
    t = 257
    rows = np.round(rng.standard_normal((64, t)) * 8).astype(np.float32)  # many ties
    seen = rng.integers(0, 2 * t, size=64).astype(np.uint32)
    a = ob.window_stats(rows, seen, t)
    b = ob.window_stats(rows, seen, t, fast=True, threads=4)
    for k in a:
        assert np.array_equal(a[k], b[k]), k


def test_ring_ingest_wraps():
    ring = np.zeros((3, 4), np.float32)
    seen = np.zeros(3, np.uint32)
    for tick in range(6):
        ob.ring_ingest(ring, seen, 4, np.array([tick, 10 + tick, 20 + tick], np.float32))
    assert ring[0].tolist() == [4, 5, 2, 3] and seen.tolist() == [6, 6, 6]
    ob.ring_ingest(ring, seen, 4, np.array([99], np.float32), series_ids=[2])
    assert ring[2].tolist() == [24, 25, 99, 23] and seen.tolist() == [6, 6, 7]


def test_predicate_leftovers_hand_derived():
    """Known answers derived by hand from pkg/metrics/types.go:151-156 and :187-199,
    sources/pod_metrics.go:174-177 and sources/node_metrics.go:142-164 (derived, not
    reference-pinned: the reference has no tests)."""
    import oracle_binding as ob
    cores, mem, disk = ob.node_available([4000], [1500], [8 * 2**30], [2 * 2**30], [100 * 2**30], [25 * 2**30])
    assert (cores[0], mem[0], disk[0]) == (2.5, 6.0, 75.0)
    # disconnected, excellent (< 10), good (10 is not < 10), good, fair (50 is not < 50), poor
    assert list(ob.network_quality([0, 1, 1, 1, 1, 1], [1, 9.9, 10, 49.9, 50, 100])) == [0, 1, 2, 2, 3, 4]
    assert list(ob.pod_restarts([0, 2, 2, 3], [1, 2, 7])) == [3, 0, 7]
    assert list(ob.pod_restarts([0, 2], [2**31 - 1, 1])) == [-2**31]            # int32 wraps
    # node 0: Ready=True; node 1: Ready=True + MemoryPressure=True; node 2: no conditions;
    # node 3: Ready=Unknown + an unrelated condition that is True
    healthy, failing = ob.node_health([0, 1, 3, 3, 5], [0, 0, 1, 0, 5], [1, 1, 1, 2, 1])
    assert list(healthy) == [1, 0, 1, 0] and list(failing) == [0, 2, 0, 1]


def test_topk_oracle_hand_derived():
    """Top-k restatement pinned on the reference's own fixture (derived by hand from
    controller.go:160, :174-221; the reference itself keeps only [0])."""
    import oracle_binding as ob
    idx, score = ob.schedule_topk([85.0, 75.0, 90.0], [3, 3, 3], [60.0], [0, 2], [1, 2], 3)
    assert idx.tolist() == [[2, 0, 1]] and score.tolist() == [[100.0, 85.0, 85.0]]
    idx, score = ob.schedule_topk([100.0, 100.0, 100.0], [3, 3, 3], [50.0], [0, 1], [0], 2)   # Fixture B
    assert idx.tolist() == [[0, 1]] and score.tolist() == [[110.0, 100.0]]
    idx, _ = ob.schedule_topk([10.0, 20.0], [3, 1], [0.0], [0, 0], [], 2)
    assert idx.tolist() == [[0, -1]]
