"""GPU parity of node-range / request sharding without an all-gather (klm.h, "node-range
sharding"): K1 per shard publishes into every rank's exchange slots, the sharded K2 of each rank
decides its slice of the requests and probes preferred nodes where they live. Checked bit for
bit against the oracle (controller.go:174-221 + :160-161 on the WHOLE table).

With one GPU the ranks are emulated on it: all K1s are enqueued before any K2 on one stream, so
no kernel ever waits for a kernel that is not already complete (two kernels that wait on each
other must never share a GPU). With two or more GPUs the multi-device handle is tested through
klm_* (skipped, not passed, on a one-GPU box)."""
import ctypes as C

import numpy as np
import pytest

import helpers
import oracle_binding as ob

pytestmark = pytest.mark.gpu


def _run_emulated(dev, pkg, bat, flags, mb, off, ids, g, replicated, epochs=(1,)):
    import torch
    n, r = bat.size, mb.size
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    d_mb, d_off = t(mb), t(off)
    d_ids = t(ids if ids.size else np.zeros(1, np.int32))
    ranges = [pkg.binding.shard_range(n, g, k) for k in range(g)]
    spread = replicated == "spread"
    if spread:       # every rank has a replica; only its own shard is filled in, K1 pushes it to the others
        reps = [(torch.full((max(n, 1),), -1.0, dtype=torch.float64, device="cuda"),
                 torch.zeros(max(n, 1), dtype=torch.uint8, device="cuda")) for _ in range(g)]
        for k, (lo, hi) in enumerate(ranges):
            reps[k][0][lo:hi] = t(bat[lo:hi])
            reps[k][1][lo:hi] = t(flags[lo:hi])
        parts = [(reps[k][0][lo:hi], reps[k][1][lo:hi]) for k, (lo, hi) in enumerate(ranges)]
    elif replicated:   # request-shard layout: one table, the parts are slices of it
        fb, ff = t(bat), t(flags)
        parts = [(fb[lo:hi], ff[lo:hi]) for lo, hi in ranges]
    else:            # node-shard layout: every rank's shard is its own allocation
        parts = [(t(bat[lo:hi]) if hi > lo else torch.empty(0, dtype=torch.float64, device="cuda"),
                  t(flags[lo:hi]) if hi > lo else torch.empty(0, dtype=torch.uint8, device="cuda"))
                 for lo, hi in ranges]
    slots = [dev.shard_slots() for _ in range(g)]
    scratch = [torch.zeros_like(dev.best_scratch) for _ in range(g)]
    slot_ptrs = [s.data_ptr() for s in slots]
    bp = [p[0].data_ptr() if p[0].numel() else 0 for p in parts]
    fp = [p[1].data_ptr() if p[1].numel() else 0 for p in parts]
    out_idx = torch.full((r,), -7, dtype=torch.int32, device="cuda")
    out_score = torch.full((r,), -7.0, dtype=torch.float64, device="cuda")
    for e in epochs:
        for k, (lo, hi) in enumerate(ranges):      # every rank's K1 first ...
            if spread:
                dev.node_best_spread(parts[k][0], parts[k][1], lo, g, k, slot_ptrs, [x[0].data_ptr() for x in reps],
                                     [x[1].data_ptr() for x in reps], e, scratch=scratch[k])
            else:
                dev.node_best_publish(parts[k][0], parts[k][1], lo, g, k, slot_ptrs, e, scratch=scratch[k])
        for k in range(g):                         # ... then every rank's K2 on its request slice
            r0, r1 = r * k // g, r * (k + 1) // g
            if r1 == r0:
                continue
            if spread:   # rank k probes its own replica, which the peers have completed
                bp = [reps[k][0][lo:].data_ptr() if hi > lo else 0 for lo, hi in ranges]
                fp = [reps[k][1][lo:].data_ptr() if hi > lo else 0 for lo, hi in ranges]
            # offsets stay absolute: the slice of pref_off indexes the whole pref_ids array
            dev.schedule_sharded(n, bp, fp, slots[k], e, d_mb[r0:r1], d_off[r0:r1 + 1], d_ids,
                                 out_idx[r0:r1], out_score[r0:r1], table_from_k1=spread)
    torch.cuda.synchronize()
    for s in slots:
        assert dev.shard_error(s) == 0
    return out_idx.cpu().numpy(), out_score.cpu().numpy()


@pytest.mark.parametrize("g", [1, 2, 3, 8, 16])
@pytest.mark.parametrize("replicated", [False, True, "spread"])
def test_sharded_equals_oracle_tie_heavy(dev, pkg, g, replicated):
    rng = np.random.default_rng(100 + g)
    for n, r, kmax in ((1, 5, 2), (17, 40, 3), (1000, 3000, 3), (40_003, 20_001, 6)):
        bat, flags, mb, off, ids = helpers.random_problem(rng, n, r, kmax=kmax, grid=1.0)
        want = ob.schedule_literal(bat, flags, mb, off, ids) if n * r <= 4e7 else \
            ob.schedule_decomposed(bat, flags, mb, off, ids)
        got = _run_emulated(dev, pkg, bat, flags, mb, off, ids, g, replicated, epochs=(1, 2, 3))
        assert np.array_equal(got[0], want[0]), (n, r, g)
        assert np.array_equal(got[1], want[1]), (n, r, g)


def test_sharded_no_eligible_node_and_empty_shards(dev, pkg):
    # 5 nodes over 8 ranks: shards of 16 nodes, ranks 1..7 own nothing; then nobody is eligible
    bat = np.array([50.0, 60.0, 60.0, 10.0, 99.0])
    flags = np.array([3, 3, 3, 3, 1], np.uint8)
    mb = np.array([0.0, 55.0, 61.0, 0.0])
    off = np.array([0, 0, 1, 2, 4], np.int32)
    ids = np.array([3, 4, 0, 2], np.int32)
    want = ob.schedule_literal(bat, flags, mb, off, ids)
    got = _run_emulated(dev, pkg, bat, flags, mb, off, ids, 8, False)
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
    flags[:] = 2
    want = ob.schedule_literal(bat, flags, mb, off, ids)
    got = _run_emulated(dev, pkg, bat, flags, mb, off, ids, 8, False)
    assert (got[0] == -1).all() and np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])


@pytest.mark.parametrize("g", [2, 4, 8])
def test_c5_full_size_sharded(dev, pkg, g):
    """C5 at full size: 1M nodes in g shards, 1M requests in g slices."""
    n, r = 1_000_000, 1_000_000
    bat, flags = pkg.synth.node_table(5, n)
    mb, off, ids = pkg.synth.requests(5, r, n)
    want = ob.schedule_decomposed(bat, flags, mb, off, ids)
    got = _run_emulated(dev, pkg, bat, flags, mb, off, ids, g, False)
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
    got = _run_emulated(dev, pkg, bat, flags, mb, off, ids, g, "spread")   # K1 fused with the all-gather
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])


def test_publish_argument_checks(dev, pkg):
    import torch
    lib = pkg.binding.load()
    bat = torch.zeros(16, dtype=torch.float64, device="cuda")
    fl = torch.zeros(16, dtype=torch.uint8, device="cuda")
    slots = dev.shard_slots()
    arr = (C.c_void_p * 1)(slots.data_ptr())
    sc = dev.best_scratch.data_ptr()
    b = pkg.binding
    # epoch 0 is the "never published" value of zero-filled slots
    assert lib.klm_dev_node_best_publish(bat.data_ptr(), fl.data_ptr(), 16, 0, 1, 0, arr, 0, None, sc, None) == b.KLM_EINVAL
    assert lib.klm_dev_node_best_publish(bat.data_ptr(), fl.data_ptr(), 16, 0, 17, 0, arr, 1, None, sc, None) == b.KLM_ELIMIT
    assert lib.klm_dev_node_best_publish(bat.data_ptr(), fl.data_ptr(), 16, 0, 1, 1, arr, 1, None, sc, None) == b.KLM_EINVAL


# ------------------------------------------------------------------ several GPUs, through klm_*

def _gpus():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("mode_name", ["SHARD_REQUESTS", "SHARD_NODES"])
def test_multi_device_handle_matches_oracle(pkg, mode_name):
    g = _gpus()
    if g < 2:
        pytest.skip("needs at least two GPUs (one handle over several devices)")
    b = pkg.binding
    mode = getattr(b, mode_name)
    rng = np.random.default_rng(7)
    with pkg.Handle(list(range(g))) as h:
        for n, r, kmax in ((3, 3, 2), (1000, 3000, 3), (100_000, 300_000, 3), (1_000_000, 1_000_000, 2)):
            if n >= 100_000:
                bat, flags = pkg.synth.node_table(5, n)
                mb, off, ids = pkg.synth.requests(5, r, n)
            else:
                bat, flags, mb, off, ids = helpers.random_problem(rng, n, r, kmax=kmax, grid=1.0)
            want = ob.schedule_decomposed(bat, flags, mb, off, ids)
            got = h.reconcile(bat, flags, mb, off, ids, algo=b.ALGO_DECOMPOSED | mode)
            assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1]), (n, r, "reconcile")
            got = h.schedule_batch(mb, off, ids, algo=b.ALGO_DECOMPOSED | mode)   # the snapshot stays
            assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1]), (n, r, "batch")
            h.nodes_set(bat, flags)
            for m2 in (mode, b.SHARD_REQUESTS, b.SHARD_NODES):   # both layouts from one snapshot
                got = h.schedule_batch(mb, off, ids, algo=b.ALGO_DECOMPOSED | m2)
                assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1]), (n, r, m2)


def test_multi_device_handle_kats_and_verifier(pkg):
    g = _gpus()
    if g < 2:
        pytest.skip("needs at least two GPUs (one handle over several devices)")
    b = pkg.binding
    rng = np.random.default_rng(9)
    with pkg.Handle([0, 1]) as h:
        for n, r in ((3, 4), (200, 300), (3000, 500)):
            bat, flags, mb, off, ids = helpers.random_problem(rng, n, r, kmax=5, grid=1.0)
            want = ob.schedule_literal(bat, flags, mb, off, ids)
            for mode in (b.SHARD_REQUESTS, b.SHARD_NODES, b.SHARD_REQUESTS | b.ALGO_BRUTEFORCE):
                got = h.reconcile(bat, flags, mb, off, ids, algo=mode)
                assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1]), (n, r, mode)
        with pytest.raises(b.KlmError):   # the verifier scans whole tables
            h.reconcile(bat, flags, mb, off, ids, algo=b.SHARD_NODES | b.ALGO_BRUTEFORCE)
        with pytest.raises(b.KlmError):
            h.reconcile(bat, flags, mb, off, ids, algo=0x200)
    with pytest.raises(b.KlmError):       # a device listed twice
        pkg.Handle([0, 0])
