"""GPU parity, scheduling: CUDA path through the C ABI vs the oracle. Bit-exact on both
the chosen index and the f64 score."""
import numpy as np
import pytest

import controller_oracle as co
import helpers
import oracle_binding as ob

pytestmark = pytest.mark.gpu
KATS = [c for c in helpers.load_kats() if c["name"] != "empty-workload"]


def _both(handle, pkg, bat, flags, mb, off, ids):
    handle.nodes_set(bat, flags)
    a = handle.schedule_batch(mb, off, ids, algo=pkg.binding.ALGO_DECOMPOSED)
    b = handle.schedule_batch(mb, off, ids, algo=pkg.binding.ALGO_BRUTEFORCE)
    return a, b


def _eq(got, want, what):
    assert np.array_equal(got[0], want[0]), f"{what}: index mismatch at {np.flatnonzero(got[0] != want[0])[:5]}"
    assert np.array_equal(got[1], want[1]), f"{what}: score mismatch"


@pytest.mark.parametrize("case", KATS, ids=[c["name"] for c in KATS])
def test_c1_kats(handle, pkg, case):
    """Config C1: the reference's example requests against its fixture fleets."""
    spec = co.decode_spec(case["spec"], apimachinery_ints=False)
    bat, flags, names = helpers.flatten_items(case["items"])
    mb, off, ids = helpers.flatten_specs([spec], names)
    for got in _both(handle, pkg, bat, flags, mb, off, ids):
        assert got[0][0] == case["expect"]["index"]
        assert got[1][0] == case["expect"]["score"]


def test_c1_batched(handle, pkg):
    """All KAT requests that share fixture A in one batch (reconcile batches, controller.go:103)."""
    cases = [c for c in KATS if c["name"].startswith("A/")]
    bat, flags, names = helpers.flatten_items(cases[0]["items"])
    specs = [co.decode_spec(c["spec"], apimachinery_ints=False) for c in cases]
    mb, off, ids = helpers.flatten_specs(specs, names)
    for got in _both(handle, pkg, bat, flags, mb, off, ids):
        assert got[0].tolist() == [c["expect"]["index"] for c in cases]
        assert got[1].tolist() == [c["expect"]["score"] for c in cases]


@pytest.mark.parametrize("seed", range(6))
def test_random_tie_heavy(handle, pkg, seed):
    rng = np.random.default_rng(100 + seed)
    n = int(rng.integers(1, 5000))
    bat, flags, mb, off, ids = helpers.random_problem(rng, n, 2000, kmax=6, grid=rng.choice([0.2, 1.0, 10.0]))
    want = ob.schedule_literal(bat, flags, mb, off, ids)
    dec, bru = _both(handle, pkg, bat, flags, mb, off, ids)
    _eq(dec, want, "decomposed")
    _eq(bru, want, "bruteforce")


@pytest.mark.parametrize("n,r", [(300, 40), (300, 30_000), (60_000, 40), (60_000, 30_000)],
                         ids=["staged/staged", "staged/direct", "direct/staged", "direct/direct"])
def test_staged_and_direct_copy_paths(handle, pkg, n, r):
    """Small node tables and small batches cross PCIe as one staged block each way, large
    ones array by array (abi.cu, kStageMax); every combination gives the oracle's answer.
    Two snapshots are installed back to back first: a staged klm_nodes_set returns before
    the GPU has read the staging block, and the next one must wait for that."""
    rng = np.random.default_rng(n * 7 + r)
    stale = helpers.random_problem(rng, n, 1, kmax=1, grid=1.0)
    handle.nodes_set(stale[0], stale[1])
    bat, flags, mb, off, ids = helpers.random_problem(rng, n, r, kmax=3, grid=0.5)
    want = ob.schedule_scan(bat, flags, mb, off, ids, threads=8)  # the literal loop, threaded
    handle.nodes_set(bat, flags)
    for _ in range(2):
        _eq(handle.schedule_batch(mb, off, ids), want, "decomposed")


@pytest.mark.parametrize("r", [262_144, 300_001, 1_100_000])
def test_chunked_batch_with_pinned_arrays(pkg, r):
    """Batches of >= 256 k requests are cut into chunks that alternate between two streams
    (abi.cu, kChunkRequests). With page-locked caller arrays every copy is truly asynchronous,
    so this is the case where a missing dependency between the chunks would show."""
    import torch
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()
    rng = np.random.default_rng(r)
    n = 20_000
    bat, flags, mb, off, ids = helpers.random_problem(rng, n, r, kmax=4, grid=0.5)
    want = ob.schedule_decomposed(bat, flags, mb, off, ids)
    lit = ob.schedule_scan(bat, flags, mb[:2000], off[:2001], ids[: off[2000]], threads=8)
    assert np.array_equal(lit[0], want[0][:2000]) and np.array_equal(lit[1], want[1][:2000])
    bat, flags, mb, off, ids = map(pin, (bat, flags, mb, off, ids))
    res_i, res_s = pin(np.empty(r, np.int32)), pin(np.empty(r))
    with pkg.Handle(0) as h:
        call = h.reconcile_prepared(bat, flags, mb, off, ids, res_i, res_s)
        for _ in range(3):
            res_i.fill(-7)
            res_s.fill(np.nan)
            call()
            _eq((res_i, res_s), want, f"chunked R={r}")
        bad = off.copy()
        bad[r // 2] = bad[-1] + 5          # a chunk boundary beyond pref_off[R]
        with pytest.raises(pkg.binding.KlmError):
            h.schedule_batch(mb, bad, ids)


def test_host_alloc_arrays(pkg):
    """klm_host_alloc gives page-locked arrays a caller can fill and pass to any klm_* call."""
    b = pkg.binding
    rng = np.random.default_rng(3)
    n, r = 4000, 300_000
    prob = helpers.random_problem(rng, n, r, kmax=2, grid=1.0)
    want = ob.schedule_decomposed(*prob)
    arrs = []
    for a in prob:
        ha = b.host_array(a.shape, a.dtype)
        ha[...] = a
        arrs.append(ha)
    res_i, res_s = b.host_array(r, np.int32), b.host_array(r, np.float64)
    with pkg.Handle(0) as h:
        call = h.reconcile_prepared(*arrs, res_i, res_s)
        call()
        _eq((res_i, res_s), want, "host_alloc arrays")
        assert b.load().klm_host_free(None) == 0
        with pytest.raises(b.KlmError):
            b.host_array(1 << 50, np.uint8)   # beyond host memory: a clean error, not a crash
        res_i.fill(-7)
        call()                                 # and it does not poison the calls that follow
        _eq((res_i, res_s), want, "after a failed allocation")


@pytest.mark.parametrize("n,r", [(1, 1), (300, 40), (300, 30_000), (60_000, 40), (5000, 0)])
def test_reconcile_equals_two_calls(handle, pkg, n, r):
    """klm_reconcile == klm_nodes_set + klm_schedule_batch (both algorithms), in its fused small
    form and in its two-call form; the snapshot it installs serves later klm_schedule_batch calls."""
    rng = np.random.default_rng(n + 3 * r)
    bat, flags, mb, off, ids = helpers.random_problem(rng, n, r, kmax=3, grid=0.5)
    want = ob.schedule_scan(bat, flags, mb, off, ids, threads=8)
    other = helpers.random_problem(rng, n, 1, kmax=1, grid=1.0)
    for algo in (pkg.binding.ALGO_DECOMPOSED, pkg.binding.ALGO_BRUTEFORCE):
        handle.nodes_set(other[0], other[1])            # a different snapshot first
        _eq(handle.reconcile(bat, flags, mb, off, ids, algo=algo), want, f"reconcile algo {algo}")
        if r:
            _eq(handle.schedule_batch(mb, off, ids), want, "batch on the snapshot reconcile installed")
    with pytest.raises(pkg.KlmError):
        handle.reconcile(bat, flags, mb, off, ids, algo=9)


def test_edge_shapes(handle, pkg):
    # no nodes at all
    e = np.zeros(0)
    mb, off, ids = np.array([0.0, 50.0]), np.array([0, 1, 1], np.int32), np.array([0], np.int32)
    for got in _both(handle, pkg, e, e.astype(np.uint8), mb, off, ids):
        assert got[0].tolist() == [-1, -1] and got[1].tolist() == [0.0, 0.0]
    # zero requests
    handle.nodes_set(np.array([1.0]), np.array([3], np.uint8))
    i, s = handle.schedule_batch(np.zeros(0), np.zeros(1, np.int32), np.zeros(0, np.int32))
    assert i.size == 0 and s.size == 0
    # nobody eligible; unaligned sizes around the 16-node vector width
    for n in (1, 15, 16, 17, 31, 33, 4097):
        bat = np.linspace(0, 99, n)
        flags = np.full(n, 1, np.uint8)
        for got in _both(handle, pkg, bat, flags, np.array([0.0]), np.array([0, 0], np.int32), np.zeros(0, np.int32)):
            assert got[0][0] == -1
        flags[n - 1] = 3
        for got in _both(handle, pkg, bat, flags, np.array([0.0]), np.array([0, 0], np.int32), np.zeros(0, np.int32)):
            assert got[0][0] == n - 1 and got[1][0] == bat[n - 1]
    # negative and zero batteries, -0.0 == +0.0 ties go to the lowest index
    bat = np.array([-5.0, -0.0, 0.0, -1.0])
    flags = np.full(4, 3, np.uint8)
    for got in _both(handle, pkg, bat, flags, np.array([0.0, -3.0]), np.array([0, 0, 1], np.int32),
                     np.array([3], np.int32)):
        assert got[0].tolist() == [1, 3] and got[1].tolist() == [0.0, 9.0]


def test_state_and_errors(pkg):
    with pkg.Handle(0) as h:
        with pytest.raises(pkg.KlmError) as ei:
            h.schedule_batch(np.zeros(1), np.zeros(2, np.int32), np.zeros(0, np.int32))
        assert ei.value.code == pkg.binding.KLM_ESTATE
        h.nodes_set(np.array([1.0]), np.array([3], np.uint8))
        with pytest.raises(pkg.KlmError):
            h.schedule_batch(np.zeros(1), np.zeros(2, np.int32), np.zeros(0, np.int32), algo=7)


@pytest.mark.parametrize("cfg,n,r,k", [(2, 1000, 1, None), (3, 5000, 5000, 2)])
def test_config_shapes_vs_literal(handle, pkg, cfg, n, r, k):
    """C2 (1k nodes, single request) and C3 (5k UAV nodes, 5k requests, K=2) at full size."""
    bat, flags = pkg.synth.node_table(cfg, n)
    mb, off, ids = pkg.synth.requests(cfg, r, n, k_fixed=k)
    want = ob.schedule_literal(bat, flags, mb, off, ids)
    dec, bru = _both(handle, pkg, bat, flags, mb, off, ids)
    _eq(dec, want, "decomposed")
    _eq(bru, want, "bruteforce")


def test_c4_full_size(handle, pkg):
    """C4: 100k nodes x 1M requests. The literal oracle checks a 2000-request sample
    (10^11 evaluations are out of reach on a CPU); the whole batch is checked against the
    O(N + RK) oracle (proved equal to the literal one in test_oracle_kat.py) and, on the
    GPU, the brute-force scan of a 200k-request slice must equal the decomposed kernel."""
    n, r = 100_000, 1_000_000
    bat, flags = pkg.synth.node_table(4, n)
    mb, off, ids = pkg.synth.requests(4, r, n)
    handle.nodes_set(bat, flags)
    dec = handle.schedule_batch(mb, off, ids)
    _eq(dec, ob.schedule_decomposed(bat, flags, mb, off, ids), "decomposed vs O(N+RK) oracle")
    m = 2000
    sub_off = (off[: m + 1] - off[0]).astype(np.int32)
    lit = ob.schedule_literal(bat, flags, mb[:m], sub_off, ids[: off[m]])
    assert np.array_equal(dec[0][:m], lit[0]) and np.array_equal(dec[1][:m], lit[1])
    m = 200_000
    bru = handle.schedule_batch(mb[:m], off[: m + 1], ids[: off[m]], algo=pkg.binding.ALGO_BRUTEFORCE)
    assert np.array_equal(dec[0][:m], bru[0]) and np.array_equal(dec[1][:m], bru[1])
    # size-independent properties: the winner is eligible, passes the gate, and scores
    # battery or battery + 10
    w = dec[0]
    ok = w >= 0
    assert ((flags[w[ok]] & 3) == 3).all()
    gate = mb[ok] > 0
    assert (bat[w[ok]][gate] >= mb[ok][gate]).all()
    d = dec[1][ok] - bat[w[ok]]
    assert np.isin(d, (0.0, 10.0)).all() or np.allclose(d[~np.isin(d, (0.0, 10.0))], 10.0, atol=1e-9)


@pytest.mark.parametrize("g", [2, 4, 8])
def test_c5_node_sharded_merge(dev, pkg, g):
    """C5: 1M nodes split into g contiguous ranges (one per rank); every rank's partial
    winners + K3 merge must equal the unsharded result. The ranks are emulated on one GPU
    (the NCCL all-gather itself is covered by the gloo test and bench.py)."""
    import torch
    n, r = 1_000_000, 200_000
    bat, flags = pkg.synth.node_table(5, n)
    mb, off, ids = pkg.synth.requests(5, r, n)
    want = ob.schedule_decomposed(bat, flags, mb, off, ids)
    t = lambda a: torch.from_numpy(a).cuda()
    d_bat, d_flags, d_mb, d_off, d_ids = t(bat), t(flags), t(mb), t(off), t(ids)
    part_idx = torch.empty((g, r), dtype=torch.int32, device="cuda")
    part_score = torch.empty((g, r), dtype=torch.float64, device="cuda")
    best = torch.empty(2, dtype=torch.int64, device="cuda")
    for k in range(g):
        lo, hi = n * k // g, n * (k + 1) // g
        dev.node_best(d_bat[lo:hi], d_flags[lo:hi], best, node_base=lo)
        dev.schedule(d_bat[lo:hi], d_flags[lo:hi], best, d_mb, d_off, d_ids, part_idx[k], part_score[k],
                     node_base=lo)
    out_idx = torch.empty(r, dtype=torch.int32, device="cuda")
    out_score = torch.empty(r, dtype=torch.float64, device="cuda")
    dev.merge_partials(part_idx, part_score, out_idx, out_score)
    torch.cuda.synchronize()
    assert np.array_equal(out_idx.cpu().numpy(), want[0])
    assert np.array_equal(out_score.cpu().numpy(), want[1])
    # the peer-memory form: every "rank" merges only its slice of the requests, reading the
    # partial arrays through per-rank pointers (all local here; peers over NVLink in bench.py)
    ip = [part_idx[k].data_ptr() for k in range(g)]
    sp = [part_score[k].data_ptr() for k in range(g)]
    for k in range(g):
        r0, r1 = r * k // g, r * (k + 1) // g
        oi = torch.full((r1 - r0,), -5, dtype=torch.int32, device="cuda")
        os_ = torch.zeros(r1 - r0, dtype=torch.float64, device="cuda")
        dev.merge_partials_peer(ip, sp, r0, r1, oi, os_)
        torch.cuda.synchronize()
        assert np.array_equal(oi.cpu().numpy(), want[0][r0:r1])
        assert np.array_equal(os_.cpu().numpy(), want[1][r0:r1])


@pytest.mark.parametrize("n,r", [(1000, 5), (100_000, 300_000), (1_000_000, 1_000_000)])
def test_schedule_after_best_overlaps_k1(dev, pkg, n, r):
    """klm_dev_schedule_after_best starts while the node_best launched just before it is still
    running and may read `best` only once that kernel is complete. Two node tables with
    different winners alternate, back to back and without synchronisation, so a K2 that read
    `best` early would answer with the previous table's winner."""
    import torch
    rng = np.random.default_rng(n + r)
    mb, off, ids = pkg.synth.requests(4, r, n)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    d_mb, d_off, d_ids = t(mb), t(off), t(ids if ids.size else np.zeros(1, np.int32))
    tables, wants = [], []
    for k in range(2):
        bat, flags = pkg.synth.node_table(4 + k, n)
        bat = bat.copy()
        bat[rng.integers(0, n)] = 99.95 + k * 0.01          # a distinct winner per table
        tables.append((t(bat), t(flags)))
        wants.append(ob.schedule_decomposed(bat, flags, mb, off, ids))
    best = torch.zeros(2, dtype=torch.int64, device="cuda")
    outs = [(torch.empty(r, dtype=torch.int32, device="cuda"), torch.empty(r, dtype=torch.float64, device="cuda"))
            for _ in range(6)]
    for i, (oi, os_) in enumerate(outs):
        d_bat, d_flags = tables[i & 1]
        dev.node_best(d_bat, d_flags, best)
        dev.schedule(d_bat, d_flags, best, d_mb, d_off, d_ids, oi, os_, after_best=True)
    torch.cuda.synchronize()
    for i, (oi, os_) in enumerate(outs):
        _eq((oi.cpu().numpy(), os_.cpu().numpy()), wants[i & 1], f"launch {i}")


@pytest.mark.parametrize("n,r", [(1, 1), (7, 0), (1000, 5), (4097, 70_000), (100_000, 300_000),
                                 (1_000_000, 1_000_000), (2_000_001, 1000)])
def test_schedule_fused_equals_k1_k2(dev, pkg, n, r):
    """klm_dev_schedule_fused (K1 and K2 in one cooperative launch) against the oracle and
    against `best` as klm_dev_node_best writes it. Two tables with different winners alternate
    back to back on the same scratch without synchronisation: the self-resetting counters must be
    ready for every next launch, and no CTA may decide with the previous launch's winner."""
    import torch
    rng = np.random.default_rng(3 * n + r)
    mb, off, ids = pkg.synth.requests(6, r, n)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    d_mb = t(mb if r else np.zeros(1))[:r]
    d_off, d_ids = t(off), t(ids if ids.size else np.zeros(1, np.int32))
    tables, wants, bests = [], [], []
    for k in range(2):
        bat, flags = pkg.synth.node_table(8 + k, n)
        bat = bat.copy()
        bat[rng.integers(0, n)] = 99.95 + k * 0.01
        tables.append((t(bat), t(flags)))
        wants.append(ob.schedule_decomposed(bat, flags, mb, off, ids))
        b = torch.zeros(2, dtype=torch.int64, device="cuda")
        dev.node_best(tables[k][0], tables[k][1], b)
        bests.append(b.cpu().numpy().copy())
    outs = [(torch.full((max(r, 1),), -7, dtype=torch.int32, device="cuda")[:r],
             torch.empty(max(r, 1), dtype=torch.float64, device="cuda")[:r],
             torch.zeros(2, dtype=torch.int64, device="cuda")) for _ in range(6)]
    scratch = torch.zeros(int(dev.lib.klm_dev_node_best_scratch_bytes()), dtype=torch.uint8, device="cuda")
    for i, (oi, os_, b) in enumerate(outs):
        d_bat, d_flags = tables[i & 1]
        dev.schedule_fused(d_bat, d_flags, d_mb, d_off, d_ids, oi, os_, best=b if i != 3 else None,
                           scratch=scratch)
    torch.cuda.synchronize()
    assert not scratch[:8].any(), "the barrier counters must be back at zero"
    for i, (oi, os_, b) in enumerate(outs):
        _eq((oi.cpu().numpy(), os_.cpu().numpy()), wants[i & 1], f"launch {i}")
        if i != 3:
            assert np.array_equal(b.cpu().numpy(), bests[i & 1]), i


def test_schedule_fused_unaligned_and_based(dev, pkg):
    """Columns that are not 16-byte aligned (scalar scan) and a node_base: indices are global."""
    import torch
    n, r, base = 5001, 3000, 4096
    bat, flags = pkg.synth.node_table(11, n + 3)
    mb, off, ids = pkg.synth.requests(11, r, n)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    d_bat, d_flags = t(bat)[1:1 + n], t(flags)[3:3 + n]
    want = ob.schedule_decomposed(bat[1:1 + n], flags[3:3 + n], mb, off, ids)
    oi = torch.empty(r, dtype=torch.int32, device="cuda")
    os_ = torch.empty(r, dtype=torch.float64, device="cuda")
    dev.schedule_fused(d_bat, d_flags, t(mb), t(off), t(ids), oi, os_)
    torch.cuda.synchronize()
    _eq((oi.cpu().numpy(), os_.cpu().numpy()), want, "unaligned")
    # the same table as the shard [base, base + n) of a larger one: preferred ids are global
    ids_g = (ids.astype(np.int64) + base).astype(np.int32)
    dev.schedule_fused(d_bat, d_flags, t(mb), t(off), t(ids_g), oi, os_, node_base=base)
    torch.cuda.synchronize()
    gi = oi.cpu().numpy()
    assert np.array_equal(np.where(gi >= 0, gi - base, gi), want[0])
    assert np.array_equal(os_.cpu().numpy(), want[1])


def test_bruteforce_many_prefs(handle, pkg):
    """More preferred ids than the brute-force kernel keeps in registers."""
    rng = np.random.default_rng(5)
    bat, flags, mb, off, ids = helpers.random_problem(rng, 3000, 500, kmax=12, grid=1.0)
    want = ob.schedule_literal(bat, flags, mb, off, ids)
    dec, bru = _both(handle, pkg, bat, flags, mb, off, ids)
    _eq(dec, want, "decomposed")
    _eq(bru, want, "bruteforce")


def test_handle_is_thread_safe(pkg):
    """The reference's callers are concurrent goroutines on arbitrary OS threads (reconcile
    loop, Collect loop, HTTP handlers - SURVEY.md §8b): many threads share one handle, each
    call must see a consistent snapshot and return its own results."""
    import threading
    rng = np.random.default_rng(3)
    problems = [helpers.random_problem(rng, 2000 + 37 * i, 500, kmax=3, grid=1.0) for i in range(6)]
    wants = [ob.schedule_literal(*p) for p in problems]
    snap = pkg.synth.cluster_snapshot(2, 1500, 4000, gpus_per_node=1)
    want_cluster = ob.cluster_aggregate(snap)
    errors = []
    with pkg.Handle(0) as h:
        lock = threading.Lock()   # nodes_set + schedule_batch form one reconcile: keep them together

        def scheduler(i):
            try:
                for _ in range(15):
                    bat, flags, mb, off, ids = problems[i]
                    with lock:
                        h.nodes_set(bat, flags)
                        got = h.schedule_batch(mb, off, ids)
                    assert np.array_equal(got[0], wants[i][0]) and np.array_equal(got[1], wants[i][1])
            except Exception as e:  # noqa: BLE001
                errors.append(repr(e))

        def collector():
            try:
                for _ in range(40):
                    got = h.cluster_aggregate(*(snap[k] for k in ("cpu_capacity", "cpu_usage", "memory_capacity",
                                                                  "memory_usage", "gpu_count", "healthy",
                                                                  "gpu_usage", "pod_running")))
                    assert got == want_cluster
            except Exception as e:  # noqa: BLE001
                errors.append(repr(e))

        threads = [threading.Thread(target=scheduler, args=(i,)) for i in range(6)] + \
                  [threading.Thread(target=collector) for _ in range(3)]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
    assert not errors, errors[:3]


def test_offsets_that_decrease_are_refused(handle, pkg):
    """pref_off / ctr_off index device arrays: offsets that decrease (or leave [0, off[R]]) would
    make a kernel read out of bounds, so the host-buffer calls check them (small staged batch,
    plain batch, chunked batch, klm_reconcile, klm_pod_rates)."""
    b = pkg.binding
    rng = np.random.default_rng(12)
    for r in (50, 40_000, 300_000):
        bat, flags, mb, off, ids = helpers.random_problem(rng, 500, r, kmax=2, grid=1.0)
        ids = np.clip(ids, 0, 499).astype(np.int32)
        handle.nodes_set(bat, flags)
        bad = off.copy()
        j = r // 2
        bad[j] = bad[j + 1] + 3 if bad[j + 1] + 3 <= bad[-1] else bad[j - 1] - 1   # one inner offset out of order
        assert not np.all(np.diff(bad) >= 0)
        with pytest.raises(b.KlmError) as e:
            handle.schedule_batch(mb, bad, ids)
        assert e.value.code == b.KLM_EINVAL
        with pytest.raises(b.KlmError):
            handle.reconcile(bat, flags, mb, bad, ids)
        got = handle.schedule_batch(mb, off, ids)                     # the handle is still usable
        want = ob.schedule_decomposed(bat, flags, mb, off, ids)
        assert np.array_equal(got[0], want[0])
