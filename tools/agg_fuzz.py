#!/usr/bin/env python
"""Fuzz the aggregation side against the oracle through the host-buffer ABI: cluster aggregate,
node / pod rates, fleet scan, the analysis predicates and the top-k batch, on adversarial columns
(zero and negative capacities, usage above capacity, values near 2^62, empty CSR segments, every
count from 0 up across the kernels' tile boundaries).
   python tools/agg_fuzz.py [seconds] [seed]   -> one JSON line; exit 1 on the first mismatch"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as ge  # noqa: E402
import oracle_binding as ob  # noqa: E402

BIG = np.array([0, 1, -1, 7, 1000, 4000, 1 << 30, 1 << 40, (1 << 62) - 1, -(1 << 40), 90, 100], np.int64)


def ints(rng, n, signed=True):
    kind = int(rng.integers(0, 4))
    if kind == 0:
        v = rng.integers(0, 64000, n)
    elif kind == 1:
        v = rng.choice(BIG, n)
    elif kind == 2:
        v = rng.integers(0, 1 << 40, n)
    else:
        v = np.where(rng.random(n) < 0.2, rng.choice(BIG, n), rng.integers(0, 100000, n))
    v = v.astype(np.int64)
    return v if signed else np.abs(v)


def fail(what, key, got, want, extra):
    got, want = np.asarray(got), np.asarray(want)
    bad = np.flatnonzero(~((got == want) | ((got != got) & (want != want))).reshape(-1))[:4]
    raise SystemExit(json.dumps({"fail": what, "key": key, "at": bad.tolist(), "got": got.reshape(-1)[bad].tolist(),
                                 "want": want.reshape(-1)[bad].tolist(), **extra}))


def eq(a, b):
    a, b = np.asarray(a), np.asarray(b)
    if a.dtype.kind == "f":
        return a.shape == b.shape and bool(np.all((a == b) | ((a != a) & (b != b))))
    return np.array_equal(a, b)


def main():
    budget = float(sys.argv[1]) if len(sys.argv) > 1 else 40.0
    seed = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    rng = np.random.default_rng(seed)
    pkg = ge.load_package()
    sizes = [0, 1, 2, 3, 31, 32, 33, 255, 256, 257, 1023, 1024, 1025, 4097, 70001]
    t_end = time.time() + budget
    rounds = checks = 0
    with pkg.Handle(0) as h:
        while time.time() < t_end:
            n = int(rng.choice(sizes)) if rng.random() < 0.8 else int(rng.integers(0, 200000))
            p = int(rng.choice(sizes)) if rng.random() < 0.8 else int(rng.integers(0, 300000))
            extra = {"seed": seed, "round": rounds, "n": n, "p": p}
            # --- cluster aggregate
            g = int(rng.integers(0, 3))
            snap = {"cpu_capacity": ints(rng, n), "cpu_usage": ints(rng, n), "memory_capacity": ints(rng, n),
                    "memory_usage": ints(rng, n), "gpu_count": np.full(n, g, np.int32),
                    "healthy": (rng.random(n) < rng.random()).astype(np.uint8),
                    "gpu_usage": np.where(rng.random(n * g) < 0.3, rng.choice([0.0, 50.0, 49.999999, 100.0], n * g),
                                          rng.random(n * g) * 100),
                    "pod_running": (rng.random(p) < rng.random()).astype(np.uint8)}
            # keep the int64 sums inside int64 (Go would wrap too, but wrapping sums are not a use case)
            for k in ("cpu_capacity", "cpu_usage", "memory_capacity", "memory_usage"):
                if n:
                    snap[k] = np.clip(snap[k], -(1 << 62) // max(n, 1), (1 << 62) // max(n, 1))
            got = h.cluster_aggregate(*(snap[k] for k in ("cpu_capacity", "cpu_usage", "memory_capacity", "memory_usage",
                                                          "gpu_count", "healthy", "gpu_usage", "pod_running")))
            want = ob.cluster_aggregate(snap)
            for k in want:
                if not eq(got[k], want[k]):
                    fail("cluster_aggregate", k, [got[k]], [want[k]], extra)
            # --- node rates
            snap["disk_capacity"], snap["disk_allocatable"] = ints(rng, n), ints(rng, n)
            with_alloc = bool(rng.random() < 0.7)
            got = h.node_rates(snap["cpu_capacity"], snap["cpu_usage"], snap["memory_capacity"], snap["memory_usage"],
                               snap["disk_capacity"], snap["disk_allocatable"] if with_alloc else None)
            want = ob.node_rates(snap, with_alloc=with_alloc)
            for k in want:
                if not eq(got[k], want[k]):
                    fail("node_rates", k, got[k], want[k], extra)
            # --- available resources / network quality
            du = ints(rng, n)
            got = h.node_available(snap["cpu_capacity"], snap["cpu_usage"], snap["memory_capacity"], snap["memory_usage"],
                                   snap["disk_capacity"], du)
            want = ob.node_available(snap["cpu_capacity"], snap["cpu_usage"], snap["memory_capacity"],
                                     snap["memory_usage"], snap["disk_capacity"], du)
            for i in range(3):
                if not eq(got[i], want[i]):
                    fail("node_available", i, got[i], want[i], extra)
            conn = (rng.random(n) < 0.8).astype(np.uint8)
            rtt = np.where(rng.random(n) < 0.4, rng.choice([0.0, 10.0, 9.999999, 50.0, 100.0, 100.000001, -1.0, 1e9], n),
                           rng.random(n) * 200)
            if not eq(h.network_quality(conn, rtt), ob.network_quality(conn, rtt)):
                fail("network_quality", "q", h.network_quality(conn, rtt), ob.network_quality(conn, rtt), extra)
            # --- pod rates / restarts / node health: CSR segments, many of them empty
            kc = rng.integers(0, int(rng.choice([1, 2, 4, 9])), p)
            off = np.zeros(p + 1, np.int64)
            np.cumsum(kc, out=off[1:])
            tot = int(off[-1])
            c = {"ctr_off": off.astype(np.int32)}
            for k in ("cpu_usage", "memory_usage", "cpu_request", "cpu_limit", "memory_request", "memory_limit"):
                c[k] = np.clip(ints(rng, tot), -(1 << 58), 1 << 58)
            got = h.pod_rates(*(c[k] for k in ("ctr_off", "cpu_usage", "memory_usage", "cpu_request", "cpu_limit",
                                               "memory_request", "memory_limit")))
            want = ob.pod_rates(c)
            for k in want:
                if not eq(got[k], want[k]):
                    fail("pod_rates", k, got[k], want[k], extra)
            rc = rng.integers(0, 1 << 20, tot).astype(np.int32)
            if not eq(h.pod_restarts(c["ctr_off"], rc), ob.pod_restarts(c["ctr_off"], rc)):
                fail("pod_restarts", "sum", h.pod_restarts(c["ctr_off"], rc), ob.pod_restarts(c["ctr_off"], rc), extra)
            ct, cs = rng.integers(0, 8, tot).astype(np.uint8), rng.integers(0, 4, tot).astype(np.uint8)
            gh, wh = h.node_health(c["ctr_off"], ct, cs), ob.node_health(c["ctr_off"], ct, cs)
            for i in range(2):
                if not eq(gh[i], wh[i]):
                    fail("node_health", i, gh[i], wh[i], extra)
            # --- fleet scan
            ok = (rng.random(n) < rng.random()).astype(np.uint8)
            pct = np.where(rng.random(n) < 0.3, rng.choice([0.0, 20.0, 19.999999, 100.0, -1.0], n), rng.random(n) * 100)
            thr = float(rng.choice([20.0, 0.0, 101.0, 19.999999, -5.0]))
            gf, wf = h.fleet_scan(ok, pct, thr), ob.fleet_scan(ok, pct, thr)
            if gf[0] != wf[0] or not eq(gf[1], wf[1]):
                fail("fleet_scan", "low", gf[1], wf[1], {**extra, "healthy": [int(gf[0]), int(wf[0])]})
            # --- top-k batch
            if n:
                bat = np.where(rng.random(n) < 0.5, np.round(rng.random(n) * 4) * 25.0, rng.random(n) * 100)
                flags = np.where(rng.random(n) < float(rng.choice([0.0, 0.5, 1.0])), 3, rng.integers(0, 3, n)).astype(np.uint8)
                r = int(rng.choice([0, 1, 33, 257, 2000]))
                mb = rng.choice([0.0, 20.0, 50.0, 100.0, 100.5], r)
                kk = rng.integers(0, 5, r)
                o2 = np.zeros(r + 1, np.int64)
                np.cumsum(kk, out=o2[1:])
                ids = rng.integers(0, n, int(o2[-1])).astype(np.int32)
                k = int(rng.integers(1, 9))   # klm_topk_max() == 8
                h.nodes_set(bat, flags)
                gi, gs = h.schedule_batch_topk(mb, o2.astype(np.int32), ids, k)
                wi, ws = ob.schedule_topk(bat, flags, mb, o2.astype(np.int32), ids, k)
                if not eq(gi, wi):
                    fail("topk", "idx", gi, wi, {**extra, "k": k, "r": r})
                if not eq(gs, ws):
                    fail("topk", "score", gs, ws, {**extra, "k": k, "r": r})
                checks += 1
            checks += 9
            rounds += 1
    print(json.dumps({"ok": True, "seed": seed, "rounds": rounds, "checks": checks, "seconds": budget}))


if __name__ == "__main__":
    main()
