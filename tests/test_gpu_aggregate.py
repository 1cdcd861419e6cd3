"""GPU parity: cluster aggregate, node/pod rates, fleet scans vs the oracle. Integers
bit-exact; the f64 rates are single correctly-rounded operations, so bit-exact too."""
import numpy as np
import pytest

import oracle_binding as ob

pytestmark = pytest.mark.gpu
COLS = ("cpu_capacity", "cpu_usage", "memory_capacity", "memory_usage", "gpu_count", "healthy", "gpu_usage",
        "pod_running")


@pytest.mark.parametrize("n,p,g", [(0, 0, 0), (1, 0, 1), (7, 3, 0), (8, 16, 2), (9, 17, 3), (1000, 10_000, 0),
                                   (5000, 20_000, 2), (100_003, 1_000_001, 1)])
def test_cluster_aggregate(handle, pkg, n, p, g):
    snap = pkg.synth.cluster_snapshot(2, n, p, gpus_per_node=g)
    got = handle.cluster_aggregate(*(snap[k] for k in COLS))
    assert got == ob.cluster_aggregate(snap)


def test_cluster_aggregate_health_cases(handle):
    G = 1 << 30
    def run(cap, use, healthy):
        n = len(cap)
        return handle.cluster_aggregate(np.array(cap, np.int64), np.array(use, np.int64), np.full(n, G, np.int64),
                                        np.zeros(n, np.int64), np.zeros(n, np.int32), np.array(healthy, np.uint8),
                                        np.zeros(0), np.zeros(0, np.uint8))
    assert run([1000], [800], [1])["health_status"] == 0
    assert run([1000], [850], [1])["health_status"] == 1
    assert run([1000], [950], [1])["health_status"] == 2
    assert run([1000] * 5, [100] * 5, [1, 1, 0, 0, 0])["health_status"] == 1  # 2 < 5//2 false
    assert run([1000] * 5, [100] * 5, [1, 0, 0, 0, 0])["health_status"] == 2
    assert run([1000] * 5, [100] * 5, [2, 0, 0, 0, 255])["healthy_nodes"] == 2   # any non-zero byte is true


def test_cluster_sums_sharded(dev, pkg):
    """Node/pod ranges summed per shard then added (what the all-reduce does) == unsharded."""
    import torch
    n, p = 10_000, 77_777
    snap = pkg.synth.cluster_snapshot(4, n, p, gpus_per_node=1)
    want = ob.cluster_aggregate(snap)
    tot = torch.zeros(10, dtype=torch.int64, device="cuda")
    sums = torch.empty(10, dtype=torch.int64, device="cuda")
    g = 4
    for k in range(g):
        lo, hi, plo, phi = n * k // g, n * (k + 1) // g, p * k // g, p * (k + 1) // g
        cols = {c: torch.from_numpy(snap[c][lo:hi].copy()).cuda() for c in COLS[:7]}
        pods = torch.from_numpy(snap["pod_running"][plo:phi].copy()).cuda()
        dev.cluster_sums(cols, pods, sums)
        tot += sums
    out = torch.zeros(104, dtype=torch.uint8, device="cuda")
    dev.cluster_finish(tot, out)
    torch.cuda.synchronize()
    rec = out.cpu().numpy().view(pkg.binding.CLUSTER_OUT_DTYPE)[0]
    assert {k: rec[k].item() for k in rec.dtype.names} == want


@pytest.mark.parametrize("n", [0, 1, 1000, 100_001])
def test_node_rates(handle, pkg, n):
    snap = pkg.synth.cluster_snapshot(3, n, 0)
    if n:
        snap["cpu_capacity"][::7] = 0     # capacity 0 -> rate 0
        snap["disk_allocatable"][::5] += 1 << 40  # allocatable > capacity -> usage clamps to 0
    got = handle.node_rates(*(snap[k] for k in ("cpu_capacity", "cpu_usage", "memory_capacity", "memory_usage",
                                                "disk_capacity", "disk_allocatable")))
    want = ob.node_rates(snap)
    for k in want:
        assert np.array_equal(got[k], want[k]), k
    got = handle.node_rates(*(snap[k] for k in ("cpu_capacity", "cpu_usage", "memory_capacity", "memory_usage",
                                                "disk_capacity")))
    want = ob.node_rates(snap, with_alloc=False)
    for k in want:
        assert np.array_equal(got[k], want[k]), k


@pytest.mark.parametrize("n,shift", [(1, 0), (2, 0), (4097, 0), (1_000_001, 0), (5000, 1), (303_105, 0)])
def test_dev_node_rates_pairs_and_unaligned(dev, pkg, n, shift):
    """klm_dev_node_rates on device columns: the two-nodes-per-thread kernel (odd lengths, every
    grid shape) and the scalar kernel that columns off 16-byte alignment fall back to."""
    import torch
    snap = pkg.synth.cluster_snapshot(5, n + shift, 0)
    snap["cpu_capacity"][::7] = 0
    snap["disk_allocatable"][::5] += 1 << 40
    cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()[shift:]
    cols = {k: cu(snap[k]) for k in ("cpu_capacity", "cpu_usage", "memory_capacity", "memory_usage")}
    dc, da = cu(snap["disk_capacity"]), cu(snap["disk_allocatable"])
    out = {"cpu_rate": torch.empty(n, dtype=torch.float64, device="cuda"),
           "memory_rate": torch.empty(n, dtype=torch.float64, device="cuda"),
           "disk_usage": torch.empty(n, dtype=torch.int64, device="cuda"),
           "disk_rate": torch.empty(n, dtype=torch.float64, device="cuda"),
           "under_pressure": torch.empty(n, dtype=torch.uint8, device="cuda")}
    dev.node_rates(cols, dc, da, out)
    torch.cuda.synchronize()
    want = ob.node_rates({k: v[shift:] for k, v in snap.items() if k != "pod_running" and k != "gpu_usage"})
    for k in want:
        assert np.array_equal(out[k].cpu().numpy(), want[k]), k


@pytest.mark.parametrize("p", [1, 10_000, 250_000])
def test_pod_rates(handle, pkg, p):
    c = pkg.synth.pod_containers(2, p)
    got = handle.pod_rates(*(c[k] for k in ("ctr_off", "cpu_usage", "memory_usage", "cpu_request", "cpu_limit",
                                            "memory_request", "memory_limit")))
    want = ob.pod_rates(c)
    for k in want:
        assert np.array_equal(got[k], want[k]), k


@pytest.mark.parametrize("n", [0, 1, 1023, 1024, 1025, 5000, 1_000_003])
def test_fleet_scan(handle, pkg, n):
    rng = np.random.default_rng(n)
    ok = (rng.random(n) < 0.9).astype(np.uint8)
    pct = np.round(rng.random(n) * 100, 1)
    for thr in (20.0, 0.0, 101.0):
        h, low = handle.fleet_scan(ok, pct, thr)
        wh, wlow = ob.fleet_scan(ok, pct, thr)
        assert h == wh and np.array_equal(low, wlow)


@pytest.mark.parametrize("n", [1, 257, 100_003])
def test_available_quality_restarts_health(handle, n):
    """GetAvailableResources / GetQuality (types.go:151-156, :187-199), the pod restart sum
    (pod_metrics.go:174-177) and node health from conditions (node_metrics.go:142-164): bit-exact
    against the oracle, including the band edges, an int32 sum that wraps, and nodes without
    conditions."""
    rng = np.random.default_rng(n)
    cap = rng.integers(1, 2**40, size=(3, n))
    use = (cap * rng.random((3, n)) * 1.1).astype(np.int64)   # sometimes above capacity: negative free
    got = handle.node_available(cap[0], use[0], cap[1], use[1], cap[2], use[2])
    want = ob.node_available(cap[0], use[0], cap[1], use[1], cap[2], use[2])
    for g, w in zip(got, want):
        assert np.array_equal(g.view(np.uint64), w.view(np.uint64))
    rtt = rng.choice([0.0, 9.999, 10.0, 49.99, 50.0, 99.9, 100.0, 1e6, -1.0], size=n) + rng.random(n) * 1e-9
    rtt[: min(n, 7)] = [10.0, 50.0, 100.0, 9.0, 49.0, 99.0, 0.0][: min(n, 7)]
    con = (rng.random(n) < 0.9).astype(np.uint8)
    assert np.array_equal(handle.network_quality(con, rtt), ob.network_quality(con, rtt))
    k = rng.integers(0, 5, size=n)
    off = np.zeros(n + 1, np.int32)
    off[1:] = np.cumsum(k)
    cnt = rng.integers(0, 50, size=int(off[-1])).astype(np.int32)
    if cnt.size >= 2:
        cnt[:2] = 2**31 - 1                                   # wraps like Go's int32
    assert np.array_equal(handle.pod_restarts(off, cnt), ob.pod_restarts(off, cnt))
    kc = rng.integers(0, 7, size=n)
    coff = np.zeros(n + 1, np.int32)
    coff[1:] = np.cumsum(kc)
    ctype = rng.integers(0, 6, size=int(coff[-1])).astype(np.uint8)
    cstat = rng.choice([0, 1, 1, 1, 2], size=int(coff[-1])).astype(np.uint8)
    cstat[ctype != 0] = rng.choice([0, 0, 0, 0, 1, 2], size=int((ctype != 0).sum())).astype(np.uint8)
    gh, gf = handle.node_health(coff, ctype, cstat)
    wh, wf = ob.node_health(coff, ctype, cstat)
    assert np.array_equal(gh, wh) and np.array_equal(gf, wf)
    assert 0 < gh.sum() < n or n < 10
