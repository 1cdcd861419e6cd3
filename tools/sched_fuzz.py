#!/usr/bin/env python
"""Fuzz the scheduling kernels against the oracle: K1 + K2, K2 launched after K1 (dependent
launch), the fused launch, the brute-force scan and node-range shards (2..8 ranks emulated on one
GPU, probe and spread layouts), on adversarial fleets -- ties everywhere, signed zeros, huge and
tiny and negative batteries, garbage in the unused flag bits, nobody eligible, preferred lists of
0..12 ids with duplicates and ids outside the table, thresholds that are negative, zero, huge.
   python tools/sched_fuzz.py [seconds] [seed]   -> one JSON line; exit 1 on the first mismatch
The oracle is test infrastructure: this tool checks the product, it is not part of it."""
import ctypes
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as ge  # noqa: E402
import oracle_binding as ob  # noqa: E402

BATS = np.array([0.0, -0.0, 100.0, 99.95, 50.0, 50.0, 20.0, 1e-300, -1e-300, 1e300, -5.0, 89.99999999999999, 90.0])
MINS = np.array([0.0, -0.0, -5.0, 20.0, 50.0, 90.0, 100.0, 100.00000000000001, 1e300, 1e-300])


def fleet(rng, n):
    kind = int(rng.integers(0, 6))
    if kind == 0:
        bat = rng.random(n) * 100
    elif kind == 1:
        bat = rng.choice(BATS, n)
    elif kind == 2:
        bat = np.round(rng.random(n) * 4) * 25.0          # five values: ties everywhere
    elif kind == 3:
        bat = np.full(n, float(rng.choice(BATS)))
    elif kind == 4:
        bat = np.sort(rng.random(n) * 100)[::-1].copy()
    else:
        bat = np.where(rng.random(n) < 0.5, rng.choice(BATS, n), rng.random(n) * 100)
    p_ok = float(rng.choice([0.0, 0.02, 0.5, 0.95, 1.0]))
    flags = np.where(rng.random(n) < p_ok, 3, rng.integers(0, 3, n)).astype(np.uint8)
    flags |= (rng.integers(0, 64, n).astype(np.uint8) << 2)   # bits the kernels must ignore
    return bat.astype(np.float64), flags


def requests(rng, r, n):
    mb = np.where(rng.random(r) < 0.5, rng.choice(MINS, r), rng.random(r) * 100).astype(np.float64)
    kmax = int(rng.choice([0, 1, 2, 3, 12]))
    k = rng.integers(0, kmax + 1, r)
    off = np.zeros(r + 1, np.int64)
    np.cumsum(k, out=off[1:])
    tot = int(off[-1])
    ids = rng.integers(-2, n + 2, tot)
    hot = rng.random(tot) < 0.3                       # a few popular nodes: duplicates, and often the winner
    ids[hot] = rng.integers(0, min(n, 4), int(hot.sum()))
    return mb, off.astype(np.int32), ids.astype(np.int32)


def main():
    budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
    seed = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    rng = np.random.default_rng(seed)
    pkg = ge.load_package()
    dev = pkg.Dev()
    cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    t_end = time.time() + budget
    rounds = checks = 0

    def fail(what, got, want, extra):
        bad = np.flatnonzero((got[0] != want[0]) | (got[1].view(np.uint64) != want[1].view(np.uint64)))[:4]
        raise SystemExit(json.dumps({"fail": what, "requests": bad.tolist(), "got_idx": got[0][bad].tolist(),
                                     "want_idx": want[0][bad].tolist(), "got_score": got[1][bad].tolist(),
                                     "want_score": want[1][bad].tolist(), **extra}))

    def same(got, want):
        # scores bit for bit except the sign of a zero (Go compares values; -0.0 + 10 is 10 either way)
        return np.array_equal(got[0], want[0]) and np.array_equal(got[1] + 0.0, want[1] + 0.0)

    while time.time() < t_end:
        n = int(rng.choice([1, 2, 3, 17, 255, 256, 257, 4095, 4096, 4097, 5000])) if rng.random() < 0.8 else int(rng.integers(1, 70000))
        r = int(rng.choice([0, 1, 2, 255, 256, 257, 1023, 1025, 3000])) if rng.random() < 0.8 else int(rng.integers(0, 20000))
        bat, flags = fleet(rng, n)
        mb, off, ids = requests(rng, r, n)
        want = ob.schedule_decomposed(bat, flags, mb, off, ids)
        extra = {"seed": seed, "round": rounds, "n": n, "r": r}
        if n * max(r, 1) <= 3_000_000 and r:
            lit = ob.schedule_literal(bat, flags, mb, off, ids)
            if not same(lit, want):
                fail("oracle: literal != decomposed", want, lit, extra)
        d_bat, d_flags = cu(bat), cu(flags)
        d_mb = cu(mb if r else np.zeros(1))[:r]
        d_off, d_ids = cu(off), cu(ids if ids.size else np.zeros(1, np.int32))
        oi = torch.full((max(r, 1),), -9, dtype=torch.int32, device="cuda")[:r]
        os_ = torch.full((max(r, 1),), -9.0, dtype=torch.float64, device="cuda")[:r]
        best = torch.zeros(2, dtype=torch.int64, device="cuda")

        def got():
            torch.cuda.synchronize()
            return oi.cpu().numpy().copy(), os_.cpu().numpy().copy()

        dev.node_best(d_bat, d_flags, best)
        dev.schedule(d_bat, d_flags, best, d_mb, d_off, d_ids, oi, os_)
        if not same(got(), want):
            fail("K1 + K2", got(), want, extra)
        oi.fill_(-9); os_.fill_(-9)
        dev.node_best(d_bat, d_flags, best)
        dev.schedule(d_bat, d_flags, best, d_mb, d_off, d_ids, oi, os_, after_best=True)
        if not same(got(), want):
            fail("K1 + K2 (dependent launch)", got(), want, extra)
        oi.fill_(-9); os_.fill_(-9)
        dev.schedule_fused(d_bat, d_flags, d_mb, d_off, d_ids, oi, os_)
        if not same(got(), want):
            fail("fused", got(), want, extra)
        checks += 3
        if n * max(r, 1) <= 20_000_000 and r:
            oi.fill_(-9); os_.fill_(-9)
            dev.schedule_bruteforce(d_bat, d_flags, d_mb, d_off, d_ids, oi, os_)
            if not same(got(), want):
                fail("brute force", got(), want, extra)
            checks += 1
        # node-range shards, ranks emulated on this GPU: every K1 is enqueued before any K2
        g_parts = int(rng.choice([2, 3, 4, 8]))
        ranges = [pkg.binding.shard_range(n, g_parts, q) for q in range(g_parts)]
        for spread in (False, True):
            slots = [dev.shard_slots() for _ in range(g_parts)]
            slot_ptrs = [s_.data_ptr() for s_ in slots]
            epoch = int(rng.integers(1, 1000))
            if spread:
                reps = [(torch.full((n,), -1.0, dtype=torch.float64, device="cuda"),
                         torch.zeros(n, dtype=torch.uint8, device="cuda")) for _ in range(g_parts)]
                for q, (a, b) in enumerate(ranges):
                    reps[q][0][a:b] = d_bat[a:b]
                    reps[q][1][a:b] = d_flags[a:b]
                for q, (a, b) in enumerate(ranges):
                    dev.node_best_spread(reps[q][0][a:b], reps[q][1][a:b], a, g_parts, q, slot_ptrs,
                                         [x[0].data_ptr() for x in reps], [x[1].data_ptr() for x in reps], epoch)
            else:
                for q, (a, b) in enumerate(ranges):
                    dev.node_best_publish(d_bat[a:b], d_flags[a:b], a, g_parts, q, slot_ptrs, epoch)
            outs = []
            for q in range(g_parts):
                r0, r1 = r * q // g_parts, r * (q + 1) // g_parts
                if r1 == r0:
                    outs.append((np.zeros(0, np.int32), np.zeros(0)))
                    continue
                sub_off = (off[r0:r1 + 1] - off[r0]).astype(np.int32)
                sub_ids = ids[off[r0]:off[r1]]
                qi = torch.full((r1 - r0,), -9, dtype=torch.int32, device="cuda")
                qs = torch.full((r1 - r0,), -9.0, dtype=torch.float64, device="cuda")
                if spread:
                    bp = [reps[q][0][a:].data_ptr() if b > a else 0 for a, b in ranges]
                    fp = [reps[q][1][a:].data_ptr() if b > a else 0 for a, b in ranges]
                else:
                    bp = [d_bat[a:b].data_ptr() if b > a else 0 for a, b in ranges]
                    fp = [d_flags[a:b].data_ptr() if b > a else 0 for a, b in ranges]
                dev.schedule_sharded(n, bp, fp, slots[q], epoch, cu(mb[r0:r1]), cu(sub_off),
                                     cu(sub_ids if sub_ids.size else np.zeros(1, np.int32)), qi, qs,
                                     table_from_k1=spread)
                torch.cuda.synchronize()
                outs.append((qi.cpu().numpy(), qs.cpu().numpy()))
            gs = (np.concatenate([x[0] for x in outs]) if r else np.zeros(0, np.int32),
                  np.concatenate([x[1] for x in outs]) if r else np.zeros(0))
            if not same(gs, want):
                fail(f"sharded G={g_parts} spread={spread}", gs, want, extra)
            if any(dev.shard_error(s_) for s_ in slots):
                raise SystemExit(json.dumps({"fail": "shard error word set", **extra}))
            checks += 1
        rounds += 1
    print(json.dumps({"ok": True, "seed": seed, "rounds": rounds, "checks": checks, "seconds": budget}))


if __name__ == "__main__":
    main()
