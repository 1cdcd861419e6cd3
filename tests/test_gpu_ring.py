"""GPU parity: ring ingest and window statistics vs the oracle. min/max/p95 bit-exact
(they are input elements); mean within rtol 1e-5 (north-star tolerance; the f64 sums differ
only in association order, ~1e-15)."""
import numpy as np
import pytest

import oracle_binding as ob

pytestmark = pytest.mark.gpu
RTOL = 1e-5


def check_stats(got, want, what=""):
    for k in ("min", "max", "p95"):
        # bit patterns, not values: a zero result is +0.0 in every implementation
        bad = np.flatnonzero(got[k].view(np.uint32) != want[k].view(np.uint32))
        assert bad.size == 0, f"{what} {k}: {bad.size} mismatches, first at {bad[:5]}: {got[k][bad[:5]]} vs {want[k][bad[:5]]}"
    assert np.allclose(got["mean"], want["mean"], rtol=RTOL, atol=1e-300), f"{what} mean"


def distributions(rng, s, t):
    rows = np.empty((s, t), np.float32)
    kinds = 20
    for i in range(s):
        k = i % kinds
        if k == 0: v = rng.random(t) * 100
        elif k == 1: v = rng.standard_normal(t) * 5
        elif k == 2: v = np.round(rng.random(t) * 10)              # heavy ties
        elif k == 3: v = np.full(t, 7.25)                          # constant
        elif k == 4: v = np.sort(rng.random(t) * 1e6)              # ascending
        elif k == 5: v = np.sort(rng.random(t))[::-1] - 0.5        # descending, mixed sign
        elif k == 6: v = np.zeros(t)                               # all zero
        elif k == 7: v = 39.9042 + 0.001 * np.cos(0.1 * np.arange(t))  # tiny relative variance
        elif k == 8: v = np.where(rng.random(t) < 0.97, 0.0, rng.random(t) * 1e9)  # sparse spikes
        elif k == 9: v = -np.abs(rng.standard_normal(t)) * 1e-30   # tiny negatives
        elif k == 10: v = np.where(rng.random(t) < 0.5, -0.0, 0.0)  # signed zeros
        elif k == 11: v = np.maximum(100 - 0.1 * np.arange(t), 0)  # battery drain with a zero plateau
        elif k == 12: v = np.floor(8 + 5 * rng.random(t))          # five integer values (satellite count)
        elif k == 13: v = np.floor(rng.random(t) * 1.05)           # ~5 % ones: p95 sits on the 0/1 edge
        elif k == 14: v = np.floor(rng.random(t) * 1.02)           # ~2 % ones
        elif k == 15: v = np.round(rng.random(t) * 100)            # ~100 distinct values
        elif k == 16: v = np.round(rng.random(t) * 1000) / 10      # 0.1 grid
        elif k == 17: v = np.where(rng.random(t) < 0.9, 5.0, 5.0 + rng.random(t))  # 90 % tied at the min
        elif k == 18: v = np.where(rng.random(t) < 0.9, 5.0 - rng.random(t), 5.0)  # 10 % tied at the max
        else: v = np.exp(rng.standard_normal(t) * 3)               # heavy right tail (lognormal)
        rows[i] = v.astype(np.float32)
    return rows


# The warp kernel has one instantiation per size class (3, 5, 8, 12, 15, 16, 20, 24, 29, 32 rows of
# 128 samples; from 20 rows on split between registers and shared memory): sizes around each
# class boundary, plus sizes that fall in the middle of a class (generic pass)
@pytest.mark.parametrize("t", [1, 3, 4, 5, 20, 127, 128, 129, 300, 384, 385, 600, 640, 641, 900, 1000, 1024, 1025,
                               1200, 1536, 1537, 1800, 1920, 1921, 2048, 2049, 2431, 2432, 2433, 2500, 2560,
                               2561, 3000, 3072, 3073, 3583, 3584, 3585, 3600, 3711, 3712, 3713, 4096, 5000,
                               20000])
def test_window_stats_distributions(handle, t):
    rng = np.random.default_rng(t)
    s = 80
    rows = distributions(rng, s, t)
    handle.ring_create(s, t)
    handle.ring_load(0, rows)
    got = handle.ring_stats()
    pitch = (t + 3) & ~3
    ring = np.zeros((s, pitch), np.float32)
    ring[:, :t] = rows
    want = ob.window_stats(ring, np.full(s, t, np.uint32), t, pitch)
    check_stats(got, want, f"t={t}")


def test_partial_windows_and_ranges(handle):
    """seen < window (ring still filling), seen == 0, and sub-ranges of series."""
    rng = np.random.default_rng(9)
    s, t = 200, 3600
    ring = np.zeros((s, t), np.float32)
    seen = np.zeros(s, np.uint32)
    handle.ring_create(s, t)
    counts = rng.integers(0, t + 1, size=s)
    counts[:4] = [0, 1, 2, t]
    counts[4:12] = [2431, 2432, 2433, 2436, 2559, 2560, 2561, 3599]   # around the register / shared split
    for i in range(s):
        c = int(counts[i])
        if c:
            data = (rng.random((1, c)) * 50 - 10).astype(np.float32)
            handle.ring_load(i, data)
            ring[i, :c] = data[0]
            seen[i] = c
    want = ob.window_stats(ring, seen, t)
    check_stats(handle.ring_stats(), want, "partial")
    sub = handle.ring_stats(17, 93)
    for k in want:
        assert np.array_equal(sub[k], want[k][17:93]) or k == "mean"
    nop = handle.ring_stats(5, 5)
    assert nop["min"].size == 0
    assert handle.ring_stats(p95=False)["p95"] is None


def test_ingest_matches_oracle_and_wraps(handle):
    """Tick ingestion and scattered ingestion, past the wrap-around point."""
    rng = np.random.default_rng(11)
    s, t = 300, 37
    ring = np.zeros((s, (t + 3) & ~3), np.float32)
    seen = np.zeros(s, np.uint32)
    handle.ring_create(s, t)
    for tick in range(2 * t + 5):
        v = (rng.random(s) * 100).astype(np.float32)
        handle.ring_ingest_tick(v)
        ob.ring_ingest(ring, seen, t, v)
        if tick % 9 == 0:
            ids = rng.choice(s, size=40, replace=False).astype(np.int64)
            vv = rng.standard_normal(40).astype(np.float32)
            handle.ring_ingest(ids, vv)
            ob.ring_ingest(ring, seen, t, vv, series_ids=ids)
        if tick in (0, 5, t - 1, t, 2 * t + 4):
            check_stats(handle.ring_stats(), ob.window_stats(ring, seen, t, ring.shape[1]), f"tick {tick}")


def test_c2_shape_sample(handle, pkg):
    """C2 series (pod cpu/memory, node rates) at the real window; a 2k-series slice."""
    s, t = 2048, 3600
    scales = pkg.synth.c2_series_scales(1000, 10_000)
    rows = pkg.synth.c2_ring_rows(2, scales, 19_000, 19_000 + s, t)   # spans pod and node series
    handle.ring_create(s, t)
    handle.ring_load(0, rows)
    check_stats(handle.ring_stats(), ob.window_stats(rows, np.full(s, t, np.uint32), t), "C2")


def test_c3_shape_sample(handle, pkg):
    """C3 UAV channels (simulator waveforms, constants, zero channels) for 24 nodes x 64."""
    t = 3600
    rows = pkg.synth.uav_channel_rows(3, 100, 124, t)
    s = rows.shape[0]
    handle.ring_create(s, t)
    handle.ring_load(0, rows)
    check_stats(handle.ring_stats(), ob.window_stats(rows, np.full(s, t, np.uint32), t), "C3")


def test_full_c2_properties(dev, pkg):
    """Full C2 ring (26k series x 3600) generated on the device: size-independent
    properties (min <= mean <= max, min <= p95 <= max, exactly ceil(.95 n) elements <= p95
    and at least n - rank elements >= p95) plus an oracle check of a strided sample."""
    import torch
    s, t = 26_000, 3600
    scales = pkg.synth.c2_series_scales(1000, 10_000)
    ring = pkg.synth.c2_ring_rows_torch(2, scales, 0, s, t)
    seen = torch.full((s,), t, dtype=torch.int32, device="cuda")
    o = {k: torch.empty(s, dtype=torch.float64 if k == "mean" else torch.float32, device="cuda")
         for k in ("min", "max", "mean", "p95")}
    dev.window_stats(ring, seen, t, o["min"], o["max"], o["mean"], o["p95"])
    torch.cuda.synchronize()
    assert (o["min"].double() <= o["mean"]).all() and (o["mean"] <= o["max"].double()).all()
    assert (o["min"] <= o["p95"]).all() and (o["p95"] <= o["max"]).all()
    rank = (19 * t + 19) // 20 - 1
    le = (ring <= o["p95"][:, None]).sum(1)
    lt = (ring < o["p95"][:, None]).sum(1)
    assert (le >= rank + 1).all() and (lt <= rank).all()
    assert torch.equal(o["min"], ring.min(1).values) and torch.equal(o["max"], ring.max(1).values)
    pick = torch.arange(0, s, 53, device="cuda")
    rows = ring[pick].cpu().numpy()
    want = ob.window_stats(rows, np.full(rows.shape[0], t, np.uint32), t)
    got = {k: v[pick].cpu().numpy() for k, v in o.items()}
    check_stats(got, want, "full C2 sample")


@pytest.mark.parametrize("t", [100, 1000, 3600, 4096, 8192])
def test_k7_variants_agree(dev, t):
    """Every K7 variant (streaming, register-resident warp, shared-memory block)
    must match the oracle bit for bit."""
    import torch
    rng = np.random.default_rng(1000 + t)
    s = 400
    rows = distributions(rng, s, t)
    pitch = (t + 3) & ~3
    ring = torch.zeros((s, pitch), dtype=torch.float32, device="cuda")
    ring[:, :t] = torch.from_numpy(rows).cuda()
    seen_np = rng.integers(0, 2 * t, size=s)
    seen_np[:3] = [0, 1, t]
    seen = torch.from_numpy(seen_np.astype(np.int32)).cuda()
    want = ob.window_stats(ring.cpu().numpy(), seen_np.astype(np.uint32), t, pitch)
    for variant in ("stream", "warp", "block"):
        if t > 4096 and variant == "warp":
            continue
        o = {k: torch.full((s,), -123.0, dtype=torch.float64 if k == "mean" else torch.float32, device="cuda")
             for k in ("min", "max", "mean", "p95")}
        dev.window_stats(ring, seen, t, o["min"], o["max"], o["mean"], o["p95"], variant=variant)
        torch.cuda.synchronize()
        check_stats({k: v.cpu().numpy() for k, v in o.items()}, want, variant)


@pytest.mark.parametrize("n,p", [(333, 2111), (40_000, 90_001), (0, 17), (5, 0)],
                         ids=["staged", "direct", "no_nodes", "no_pods"])
def test_collect_cycle_matches_separate_calls(handle, pkg, n, p):
    """klm_collect (one Manager.Collect cycle in one call) == ingest + stats + rates +
    aggregate done separately, and == the oracle. Small snapshots cross PCIe as one staged
    block each way, large ones column by column; both run beside the ring branch."""
    rng = np.random.default_rng(77)
    s, t = 700, 1200
    rows = distributions(rng, s, t)
    snap = pkg.synth.cluster_snapshot(2, n, p, gpus_per_node=1)
    ring = np.ascontiguousarray(rows)
    seen = np.full(s, t - 1, np.uint32)
    handle.ring_create(s, t)
    handle.ring_load(0, rows[:, : t - 1])
    ring_o = np.zeros((s, t), np.float32)
    ring_o[:, : t - 1] = rows[:, : t - 1]
    for cycle in range(3):
        tick = (rng.random(s) * 100).astype(np.float32)
        got = handle.collect(tick, snap)
        ob.ring_ingest(ring_o, seen, t, tick)
        check_stats(got, ob.window_stats(ring_o, seen, t), f"collect cycle {cycle}")
        assert got["cluster"] == ob.cluster_aggregate(snap)
        want = ob.node_rates(snap)
        for k in want:
            assert np.array_equal(got[k], want[k]), k
    again = handle.collect(None, snap)      # no tick: statistics unchanged
    for k in ("min", "max", "p95", "mean"):
        assert np.array_equal(again[k], got[k])


def test_prepared_calls_match_wrappers(pkg):
    """collect_prepared / reconcile_prepared (structs built once, bare C calls per cycle, what
    bench.py's e2e leg times) give exactly what the convenience wrappers give."""
    rng = np.random.default_rng(5)
    s, t, n, p, r = 300, 500, 120, 900, 7
    rows = distributions(rng, s, t)
    snap = pkg.synth.cluster_snapshot(2, n, p, gpus_per_node=2)
    bat, flags = pkg.synth.node_table(2, n)
    mb, off, ids = pkg.synth.requests(2, r, n, k_fixed=2)
    with pkg.Handle(0) as a, pkg.Handle(0) as b:
        for h in (a, b):
            h.ring_create(s, t)
            h.ring_load(0, rows[:, : t - 1])
        out = {"cpu_rate": np.empty(n), "memory_rate": np.empty(n), "disk_usage": np.empty(n, np.int64),
               "disk_rate": np.empty(n), "under_pressure": np.empty(n, np.uint8), "min": np.empty(s, np.float32),
               "max": np.empty(s, np.float32), "mean": np.empty(s), "p95": np.empty(s, np.float32)}
        res_i, res_s = np.empty(r, np.int32), np.empty(r)
        collect = b.collect_prepared(snap, out)
        reconcile = b.reconcile_prepared(bat, flags, mb, off, ids, res_i, res_s)
        for cycle in range(3):
            tick = (rng.random(s) * 100).astype(np.float32)
            want = a.collect(tick, snap)
            cl = collect(tick)
            assert cl.as_dict() == want["cluster"]
            for k in out:
                assert np.array_equal(out[k], want[k]), (cycle, k)
            a.nodes_set(bat, flags)
            wi, ws = a.schedule_batch(mb, off, ids)
            reconcile()
            assert np.array_equal(res_i, wi) and np.array_equal(res_s, ws)
        with pytest.raises(ValueError):
            b.collect_prepared(snap, dict(out, min=np.empty(s, np.float64)))  # wrong dtype: refused
        # the packed result block (one download): same numbers, as views into one array
        pout = {k: v for k, v in out.items() if k not in ("min", "max", "mean", "p95")}
        pcollect = b.collect_prepared(snap, pout, packed=True)
        tick = (rng.random(s) * 100).astype(np.float32)
        want = a.collect(tick, snap)
        pcollect(tick)
        for k in ("min", "max", "mean", "p95", "cpu_rate", "under_pressure"):
            assert np.array_equal(pout[k], want[k]), k
        import ctypes as C
        bad = pkg.binding.CollectOut(None, *([None] * 5), pout["min"].ctypes.data, None, None, None,
                                     pout["min"].ctypes.data)
        cin = pkg.binding.CollectIn(None, None, 0, None, None, None, 0)
        assert b._lib.klm_collect(b._h, C.byref(cin), C.byref(bad)) == pkg.binding.KLM_EINVAL


def test_collect_pipeline_two_in_flight(pkg):
    """klm_collect_begin / klm_collect_end: with two cycles in flight every cycle still gives
    exactly what the synchronous klm_collect gives for the same tick and snapshot (the snapshot
    changes between cycles: begin stages it), and the state rules hold: a third begin, the
    synchronous ring calls while cycles are in flight, and an end with nothing in flight are
    refused with KLM_ESTATE."""
    rng = np.random.default_rng(21)
    s, t, n, p = 900, 3600, 150, 700
    rows = distributions(rng, s, t)
    with pkg.Handle(0) as a, pkg.Handle(0) as b:
        for h in (a, b):
            h.ring_create(s, t)
            h.ring_load(0, rows[:, : t - 1])
        snap = {k: np.ascontiguousarray(v) for k, v in pkg.synth.cluster_snapshot(2, n, p, gpus_per_node=1).items()}
        begin, end = b.collect_pipeline(snap)
        ticks = [(rng.random(s) * 100).astype(np.float32) for _ in range(7)]
        wants = []

        def sync_cycle(i):
            snap["cpu_usage"][:] = (snap["cpu_capacity"] * rng.random(n)).astype(np.int64)  # changes per cycle
            w = a.collect(ticks[i], snap)
            wants.append({k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in w.items()})

        def check(i, got):
            w = wants[i]
            assert got["cluster"] == w["cluster"], i
            for k in ("min", "max", "p95", "mean", "cpu_rate", "memory_rate", "disk_usage", "disk_rate",
                      "under_pressure"):
                assert np.array_equal(got[k], w[k]), (i, k)

        sync_cycle(0); begin(ticks[0])
        for i in range(1, 7):
            sync_cycle(i); begin(ticks[i])          # two in flight
            if i == 3:
                with pytest.raises(pkg.KlmError) as ei:
                    begin(ticks[i])                  # a third is refused ...
                assert ei.value.code == pkg.binding.KLM_ESTATE
                with pytest.raises(pkg.KlmError):
                    b.ring_stats()                   # ... and so is a synchronous ring call
            check(i - 1, end())
        check(6, end())
        with pytest.raises(pkg.KlmError):
            end()                                    # nothing in flight
        again = b.collect(None, snap)                # idle again: the synchronous call works
        assert np.array_equal(again["p95"], wants[6]["p95"])


def test_pushes_queue_behind_cycles_in_flight(pkg):
    """UpdateUAVReport (manager.go:391-449) is called from HTTP handlers at any time, also while
    the pipelined collector has cycles in flight: klm_ring_ingest / klm_ring_ingest_tick then queue
    and the next klm_collect_begin applies the queue, in arrival order, before its own tick. A
    test-side lock records the arrival order; a second handle replays it with synchronous calls
    (itself checked against the oracle elsewhere) and every cycle's statistics must be identical."""
    import threading
    rng = np.random.default_rng(33)
    s, t, cycles = 640, 1200, 12
    rows = distributions(rng, s, t)
    with pkg.Handle(0) as live, pkg.Handle(0) as replay:
        for h in (live, replay):
            h.ring_create(s, t)
            h.ring_load(0, rows[:, : t - 7])
        snap = {k: np.ascontiguousarray(v) for k, v in pkg.synth.cluster_snapshot(2, 10, 20).items()}
        begin, end = live.collect_pipeline(snap)
        order, lock, stop = [], threading.Lock(), threading.Event()
        errors = []

        def pusher(seed):
            r = np.random.default_rng(seed)
            try:
                while not stop.is_set():
                    if r.random() < 0.2:
                        v = (r.random(s) * 50).astype(np.float32)
                        with lock:
                            live.ring_ingest_tick(v)
                            order.append(("tick", v))
                    else:
                        ids = r.choice(s, size=int(r.integers(1, 40)), replace=False).astype(np.int64)
                        v = (r.random(ids.size) * 50).astype(np.float32)
                        with lock:
                            live.ring_ingest(ids, v)
                            order.append(("push", ids, v))
            except Exception as e:  # noqa: BLE001
                errors.append(repr(e))

        threads = [threading.Thread(target=pusher, args=(k,)) for k in range(3)]
        got = []
        ticks = [(rng.random(s) * 100).astype(np.float32) for _ in range(cycles)]
        with lock:
            begin(ticks[0]); order.append(("cycle", 0))
        for th in threads:
            th.start()
        for i in range(1, cycles):
            with lock:
                begin(ticks[i]); order.append(("cycle", i))
            res = end()
            got.append({k: res[k].copy() for k in ("min", "max", "mean", "p95")})
        stop.set()
        for th in threads:
            th.join()
        res = end()
        got.append({k: res[k].copy() for k in ("min", "max", "mean", "p95")})
        assert not errors, errors[:2]
        assert sum(1 for o in order if o[0] != "cycle") > 20, "the pushers never ran"
        final = live.ring_stats()            # idle again: applies what was pushed after the last begin
        # replay the recorded order synchronously
        want = []
        for o in order:
            if o[0] == "tick":
                replay.ring_ingest_tick(o[1])
            elif o[0] == "push":
                replay.ring_ingest(o[1], o[2])
            else:
                replay.ring_ingest_tick(ticks[o[1]])
                want.append(replay.ring_stats())
        for i, (g, w) in enumerate(zip(got, want)):
            for k in ("min", "max", "mean", "p95"):
                assert np.array_equal(g[k], w[k]), (i, k)
        want_final = replay.ring_stats()
        for k in ("min", "max", "mean", "p95"):
            assert np.array_equal(final[k], want_final[k]), k


@pytest.mark.parametrize("t", [900, 1000, 3600, 4096, 8192])
def test_hinted_window_stats_exact_over_ticks(dev, t):
    """klm_dev_window_stats_hinted: the bracket carried in `state` is only a pivot. Over a run of
    ticks (each appends one sample per series, wrapping the ring) every call must equal the oracle
    bit for bit -- from a NaN state, from the brackets the kernel left, and from garbage brackets
    (reversed, infinite, far off, another series')."""
    import torch
    rng = np.random.default_rng(7000 + t)
    s = 240
    rows = distributions(rng, s, t)
    pitch = (t + 3) & ~3
    ring = torch.zeros((s, pitch), dtype=torch.float32, device="cuda")
    ring[:, :t] = torch.from_numpy(rows).cuda()
    seen_np = np.full(s, t, np.int64)
    seen_np[:4] = [0, 1, t - 1, t // 2]                     # rings that are still filling
    seen = torch.from_numpy(seen_np.astype(np.int32)).cuda()
    state = torch.full((s, 2), float("nan"), dtype=torch.float32, device="cuda")
    o = {k: torch.empty(s, dtype=torch.float64 if k == "mean" else torch.float32, device="cuda")
         for k in ("min", "max", "mean", "p95")}
    used = 0
    for tick in range(12):
        if tick:
            vals = torch.from_numpy(distributions(rng, s, 1)[:, 0].copy()).cuda()
            dev.ring_ingest(ring, seen, t, vals)
            seen_np += 1
        if tick == 5:    # garbage brackets
            g = torch.from_numpy(rng.standard_normal((s, 2)).astype(np.float32) * 1e3).cuda()
            g[::7, 0] = float("inf"); g[1::7, 1] = float("-inf"); g[2::7] = 0.0; g[3::7, 0] = -1e30; g[3::7, 1] = 1e30
            state.copy_(g)
        if tick == 8:    # another series' brackets
            state.copy_(state.roll(17, 0))
        dev.window_stats(ring, seen, t, o["min"], o["max"], o["mean"], o["p95"], state=state)
        torch.cuda.synchronize()
        want = ob.window_stats(ring.cpu().numpy(), seen_np.astype(np.uint32), t, pitch)
        check_stats({k: v.cpu().numpy() for k, v in o.items()}, want, f"hinted t={t} tick={tick}")
        st = state.cpu().numpy()
        used += int(np.sum(st[:, 0] < st[:, 1]))
    assert used > 0 or t < 896, "no series ever produced a bracket"


@pytest.mark.parametrize("t", [1200, 3600, 4096, 6000])
def test_every_call_leaves_a_bracket(dev, t):
    """A series must not get stuck on the cold path: whatever way a call settles the p95 (sampled
    pivot, retry, counting fallback after a list overflow, a block of equal samples), it leaves a
    bracket or a tie value in `state`, and the next call uses it and stays exact."""
    import torch
    rng = np.random.default_rng(9100 + t)
    s = 1500
    rows = distributions(rng, s, t)
    pitch = (t + 3) & ~3
    ring = torch.zeros((s, pitch), dtype=torch.float32, device="cuda")
    ring[:, :t] = torch.from_numpy(rows).cuda()
    seen_np = np.full(s, t, np.int64)
    seen = torch.from_numpy(seen_np.astype(np.int32)).cuda()
    state = torch.full((s, 2), float("nan"), dtype=torch.float32, device="cuda")
    o = {k: torch.empty(s, dtype=torch.float64 if k == "mean" else torch.float32, device="cuda")
         for k in ("min", "max", "mean", "p95")}
    for tick in range(3):
        if tick:
            vals = torch.from_numpy(distributions(rng, s, 1)[:, 0].copy()).cuda()
            dev.ring_ingest(ring, seen, t, vals)
            seen_np += 1
        dev.window_stats(ring, seen, t, o["min"], o["max"], o["mean"], o["p95"], state=state)
        torch.cuda.synchronize()
        want = ob.window_stats(ring.cpu().numpy(), seen_np.astype(np.uint32), t, pitch)
        check_stats({k: v.cpu().numpy() for k, v in o.items()}, want, f"t={t} tick={tick}")
        st = state.cpu().numpy()
        finite_p95 = np.isfinite(want["p95"])
        left = ~np.isnan(st[:, 0]) & ~np.isnan(st[:, 1])
        # (a bracket whose floor has fewer than three samples between it and the answer is dropped on
        # purpose -- window_stats.cuh, kHintRefresh -- which right after a cold start is a few per cent)
        assert left[finite_p95].mean() > 0.88, (t, tick, float(left[finite_p95].mean()))


def test_full_c3_hinted_against_oracle(dev, pkg):
    """C3 at full size -- 5000 UAV nodes x 64 channels = 320 000 series x 3600 samples (4.6 GB),
    the BASELINE configuration -- through the product path (brackets carried between calls, the
    landing-ring kernel): every series of the last of three ticks is compared with the oracle, bit
    for bit (the threaded quickselect oracle, itself pinned to the sorting one on CPU)."""
    import os
    import torch
    t, nodes, step = 3600, 5000, 250
    s = nodes * 64
    ring = torch.empty((s, t), dtype=torch.float32, device="cuda")
    for a in range(0, nodes, step):
        rows = pkg.synth.uav_channel_rows(3, a, a + step, t)
        ring[a * 64:a * 64 + rows.shape[0]] = torch.from_numpy(rows).cuda()
    seen = torch.full((s,), t, dtype=torch.int32, device="cuda")
    state = torch.full((s, 2), float("nan"), dtype=torch.float32, device="cuda")
    o = {k: torch.empty(s, dtype=torch.float64 if k == "mean" else torch.float32, device="cuda")
         for k in ("min", "max", "mean", "p95")}
    for tick in range(3):
        if tick:   # a fresh sample per series: an earlier sample of the same channel
            vals = ring[:, 7 * tick].clone()
            dev.ring_ingest(ring, seen, t, vals)
        dev.window_stats(ring, seen, t, o["min"], o["max"], o["mean"], o["p95"], state=state)
    torch.cuda.synchronize()
    st = state.cpu().numpy()
    assert np.mean(~np.isnan(st[:, 0])) > 0.99, "the calls left no brackets"
    got = {k: v.cpu().numpy() for k, v in o.items()}
    threads = os.cpu_count() or 1
    for a in range(0, s, 40_000):
        b = min(s, a + 40_000)
        rows = ring[a:b].cpu().numpy()
        want = ob.window_stats(rows, np.full(b - a, t + 2, np.uint32), t, fast=True, threads=threads)
        check_stats({k: v[a:b] for k, v in got.items()}, want, f"full C3 series {a}..{b}")


@pytest.mark.parametrize("t", [7, 130, 900, 1000, 3600, 4096, 8192])
def test_hinted_over_many_ticks_everywhere_in_the_row(dev, t):
    """Ten ticks of ingest + statistics with carried brackets on rings that are empty, filling,
    exactly full, wrapped and saturated, the new sample landing in every part of a row, with cold,
    carried and garbage brackets: every result against the oracle. (The t = 130 case is the one that
    exposed the -0.0 / +0.0 pivot fault: a p95 falling from a spike back onto a plateau of zeros.)"""
    import torch
    rng = np.random.default_rng(4400 + t)
    s = 320
    rows = distributions(rng, s, t)
    pitch = (t + 3) & ~3
    ring = torch.zeros((s, pitch), dtype=torch.float32, device="cuda")
    ring[:, :t] = torch.from_numpy(rows).cuda()
    seen_np = (t + rng.integers(0, t, s)).astype(np.int64)
    seen_np[:8] = [0, 1, 2, t - 2, t - 1, t, max(t // 2, 1), 2 * t - 1]
    seen_np[8] = 0xffffffff           # saturated counter: stays, the slot is (2^32 - 1) % t
    seen = torch.from_numpy(seen_np.astype(np.uint32).view(np.int32)).cuda()
    state = torch.full((s, 2), float("nan"), dtype=torch.float32, device="cuda")
    o = {k: torch.empty(s, dtype=torch.float64 if k == "mean" else torch.float32, device="cuda")
         for k in ("min", "max", "mean", "p95")}
    for tick in range(10):
        vals = torch.from_numpy(distributions(rng, s, 1)[:, 0].copy()).cuda()
        if tick == 6:
            state.copy_(torch.from_numpy(rng.standard_normal((s, 2)).astype(np.float32) * 100).cuda())
        dev.ring_ingest(ring, seen, t, vals)
        dev.window_stats(ring, seen, t, o["min"], o["max"], o["mean"], o["p95"], state=state)
        torch.cuda.synchronize()
        sn = seen.cpu().numpy().view(np.uint32)
        want = ob.window_stats(ring.cpu().numpy(), sn, t, pitch)
        check_stats({k: v.cpu().numpy() for k, v in o.items()}, want, f"t={t} tick {tick}")


@pytest.mark.parametrize("t", [130, 1000, 3600])
def test_p95_on_and_off_a_plateau_of_zeros(dev, t):
    """The p95 of a series that is a plateau of zeros (both signs) next to about 5 % spikes moves on
    and off the plateau from tick to tick. -0.0 and +0.0 are neighbouring keys but one value: a
    pivot that steps "just below zero" must still list every zero. All variants, stateless and
    with carried brackets, against the oracle."""
    import torch
    rng = np.random.default_rng(7700 + t)
    s = 256
    k = t - ((19 * t + 19) // 20 - 1)                       # samples at or above the p95
    rows = np.where(rng.random((s, t)) < 0.5, -0.0, 0.0).astype(np.float32)
    for i in range(s):
        m = max(k + int(rng.integers(-3, 4)), 0)            # spikes: around k of them
        idx = rng.choice(t, size=min(m, t), replace=False)
        sign = 1.0 if i % 3 else -1.0                       # every third series: negative dips instead
        rows[i, idx] = (sign * (rng.random(idx.size) * 50 + 1e-3)).astype(np.float32)
    pitch = (t + 3) & ~3
    ring = torch.zeros((s, pitch), dtype=torch.float32, device="cuda")
    ring[:, :t] = torch.from_numpy(rows).cuda()
    seen_np = np.full(s, t, np.int64)
    seen = torch.from_numpy(seen_np.astype(np.int32)).cuda()
    o = {kk: torch.empty(s, dtype=torch.float64 if kk == "mean" else torch.float32, device="cuda")
         for kk in ("min", "max", "mean", "p95")}
    want = ob.window_stats(ring.cpu().numpy(), seen_np.astype(np.uint32), t, pitch)
    for variant in (None, "warp", "stream", "block"):
        dev.window_stats(ring, seen, t, o["min"], o["max"], o["mean"], o["p95"], variant=variant)
        torch.cuda.synchronize()
        check_stats({kk: v.cpu().numpy() for kk, v in o.items()}, want, f"plateau t={t} {variant}")
    for rep in range(2):
        state = torch.full((s, 2), float("nan"), dtype=torch.float32, device="cuda")
        for tick in range(14):
            spike = rng.random(s) < 0.5
            vals_np = np.where(spike, rng.random(s) * 50 + 1e-3, np.where(rng.random(s) < 0.5, -0.0, 0.0))
            vals_np[::3] = -np.abs(vals_np[::3])
            vals = torch.from_numpy(vals_np.astype(np.float32)).cuda()
            dev.ring_ingest(ring, seen, t, vals)
            dev.window_stats(ring, seen, t, o["min"], o["max"], o["mean"], o["p95"], state=state)
            seen_np += 1
            torch.cuda.synchronize()
            want = ob.window_stats(ring.cpu().numpy(), seen_np.astype(np.uint32), t, pitch)
            check_stats({kk: v.cpu().numpy() for kk, v in o.items()}, want, f"plateau t={t} rep={rep} tick={tick}")


def test_windows_holding_neg_inf_samples(dev):
    """A case the fuzzer (tools/k7_fuzz.py) found: 33-sample windows that contain -inf samples. Such
    a window lists every sample above -inf and selects with a = -inf, and the -inf PADDING of the lane
    lists then passed the `>= a` test of the gather and overwrote real candidates (p95 several ranks
    too low; in the tree since round 1). Fixture: the failing rings as generated, 184 series."""
    import os
    import torch
    d = np.load(os.path.join(os.path.dirname(__file__), "golden", "k7_case_neg_inf_small_window.npz"))
    rows, seen_np, t = d["rows"], d["seen"].astype(np.uint32), int(d["t"])
    pitch = rows.shape[1]
    want = ob.window_stats(rows, seen_np, t, pitch)
    ring = torch.from_numpy(rows).cuda()
    seen = torch.from_numpy(seen_np.view(np.int32).copy()).cuda()
    s = rows.shape[0]
    o = {k: torch.empty(s, dtype=torch.float64 if k == "mean" else torch.float32, device="cuda")
         for k in ("min", "max", "mean", "p95")}
    state = torch.full((s, 2), float("nan"), dtype=torch.float32, device="cuda")
    for variant in (None, "warp", "stream", "block", "hinted", "hinted"):
        if variant == "hinted":
            dev.window_stats(ring, seen, t, o["min"], o["max"], o["mean"], o["p95"], state=state)
        else:
            dev.window_stats(ring, seen, t, o["min"], o["max"], o["mean"], o["p95"], variant=variant)
        torch.cuda.synchronize()
        got = {k: v.cpu().numpy() for k, v in o.items()}
        for k in ("min", "max", "p95"):
            assert np.array_equal(got[k].view(np.uint32), want[k].view(np.uint32)), (variant, k)
