#!/usr/bin/env python
"""Per-kernel timing table (not the driver's bench: that is bench.py).

Times every kernel of libklm.so alone at the sizes of the BASELINE.json configs, with CUDA
events on the launching stream, a 256 MiB L2 flush between iterations when the inputs fit in
L2, and reports algorithmic bytes / time against the measured HBM copy peak. One JSON line
per kernel; `python bench_kernels.py > profiles/rNN_kernel_table.jsonl`.
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402


def main():
    pkg = ge.load_package()
    dev = pkg.Dev()
    syn = pkg.synth
    peak = 6539.9
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peak = json.load(open(p))["hbm_gbs"]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()

    def timeit(name, fn, alg_bytes, note="", iters=20, do_flush=True, units=None):
        # A train of launches keeps the GPU busy, so the idle-GPU launch latency (~8 us) that a
        # single event-bracketed launch would include is not counted. With a flush between
        # launches the flush-only train is timed too and subtracted.
        def train(with_fn):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            for _ in range(iters):
                if do_flush:
                    flush.zero_()
                if with_fn:
                    fn()
            e1.record()
            e1.synchronize()
            return e0.elapsed_time(e1) / iters

        if os.environ.get("KLM_BK_ONCE"):  # under ncu: one warm-up and one launch per kernel
            fn()
            flush.zero_()
            fn()
            torch.cuda.synchronize()
            print(json.dumps({"kernel": name, "algorithmic_bytes": int(alg_bytes)}), flush=True)
            return
        for _ in range(3):
            fn()
        reps = [train(True) - (train(False) if do_flush else 0.0) for _ in range(5)]
        t = float(np.median(reps))
        gbs = alg_bytes / t / 1e6
        line = {"kernel": name, "us": round(t * 1e3, 2), "algorithmic_bytes": int(alg_bytes),
                "GB/s": round(gbs, 1), "frac_of_measured_hbm_peak": round(gbs / peak, 4), "note": note}
        if units:
            line[units[0]] = round(units[1] / (t * 1e-3), 1)
        print(json.dumps(line), flush=True)

    # ---- K1 / K2 / K3 / K2b
    for n, r, cfg in ((100_000, 1_000_000, 4), (1_000_000, 1_000_000, 5)):
        bat, flags = syn.node_table(cfg, n)
        mb, off, ids = syn.requests(cfg, r, n)
        d_bat, d_flags, d_mb, d_off, d_ids = cu(bat), cu(flags), cu(mb), cu(off), cu(ids)
        best = torch.empty(2, dtype=torch.int64, device="cuda")
        oi = torch.empty(r, dtype=torch.int32, device="cuda")
        os_ = torch.empty(r, dtype=torch.float64, device="cuda")
        k1 = dev.prebind("node_best", d_bat, d_flags, best)
        k2 = dev.prebind("schedule", d_bat, d_flags, best, d_mb, d_off, d_ids, oi, os_)
        timeit(f"K1 node_best N={n}", k1, 9 * n, "latency-bound below ~1 MB")
        timeit(f"K2 schedule R={r} N={n}", k2, 24 * r + 13 * ids.size, units=("decisions/s", r))
        kf = dev.prebind("schedule_fused", d_bat, d_flags, d_mb, d_off, d_ids, oi, os_)
        timeit(f"K1+K2 schedule_fused R={r} N={n}", kf, 9 * n + 24 * r + 13 * ids.size,
               "one cooperative launch", units=("decisions/s", r))
    g = 8
    pi = torch.randint(-1, 1000, (g, r), dtype=torch.int32, device="cuda")
    ps = torch.rand((g, r), dtype=torch.float64, device="cuda")
    k3 = dev.prebind("merge_partials", pi, ps, oi, os_)
    timeit(f"K3 merge_partials G={g} R={r}", k3, r * (12 * g + 12), units=("decisions/s", r))
    rb, nb = 20_000, 100_000
    bat, flags = syn.node_table(4, nb)
    mb, off, ids = syn.requests(4, rb, nb)
    d_bat, d_flags, d_mb, d_off, d_ids = cu(bat), cu(flags), cu(mb), cu(off), cu(ids)
    k2b = dev.prebind("schedule_bruteforce", d_bat, d_flags, d_mb, d_off, d_ids, oi[:rb], os_[:rb])
    timeit(f"K2b bruteforce R={rb} N={nb}", k2b, 9 * nb + 24 * rb + 13 * ids.size,
           "on-chip bound by construction: smem_read_GB/s = R*N*9 B of shared-memory reads / time", iters=5,
           units=("smem_read_GB/s", rb * nb * 9 / 1e9))
    # C5 with a single request: what one decision costs when the nodes are sharded
    n1, g1 = 1_000_000, 8
    bat, flags = syn.node_table(5, n1)
    mb, off, ids = syn.requests(5, 1, n1, k_fixed=2)
    d_bat, d_flags, d_mb, d_off, d_ids = cu(bat), cu(flags), cu(mb), cu(off), cu(ids)
    k1 = dev.prebind("node_best", d_bat[: n1 // g1], d_flags[: n1 // g1], best)
    k2 = dev.prebind("schedule", d_bat[: n1 // g1], d_flags[: n1 // g1], best, d_mb, d_off, d_ids, oi[:1], os_[:1],
                     after_best=True)
    k3 = dev.prebind("merge_partials", pi[:, :1].contiguous(), ps[:, :1].contiguous(), oi[:1], os_[:1])

    def one_decision():
        k1(); k2(); k3()
    timeit(f"C5 R=1: K1+K2 on one of {g1} node shards + K3", one_decision, 9 * (n1 // g1) + 50 + 16 * g1,
           "latency of one sharded decision, exchange excluded")

    # ---- K4 / K5 / fleet scan
    n, pods = 100_000, 1_000_000
    snap = syn.cluster_snapshot(4, n, pods, gpus_per_node=2)
    cols = {k: cu(snap[k]) for k in ("cpu_capacity", "cpu_usage", "memory_capacity", "memory_usage",
                                     "gpu_count", "healthy", "gpu_usage")}
    d_pods = cu(snap["pod_running"])
    out = torch.zeros(104, dtype=torch.uint8, device="cuda")
    k4 = dev.prebind("cluster_aggregate", cols, d_pods, out)
    timeit(f"K4 cluster_aggregate N={n} P={pods}", k4, 37 * n + 8 * 2 * n + pods)
    n = 1_000_000
    snap = syn.cluster_snapshot(5, n, 1)
    cols = {k: cu(snap[k]) for k in ("cpu_capacity", "cpu_usage", "memory_capacity", "memory_usage")}
    rates = {"cpu_rate": torch.empty(n, dtype=torch.float64, device="cuda"),
             "memory_rate": torch.empty(n, dtype=torch.float64, device="cuda"),
             "disk_usage": torch.empty(n, dtype=torch.int64, device="cuda"),
             "disk_rate": torch.empty(n, dtype=torch.float64, device="cuda"),
             "under_pressure": torch.empty(n, dtype=torch.uint8, device="cuda")}
    k5 = dev.prebind("node_rates", cols, cu(snap["disk_capacity"]), cu(snap["disk_allocatable"]), rates)
    timeit(f"K5 node_rates N={n}", k5, (48 + 33) * n)

    # ---- K6 / K7
    s, t = 26_000, int(os.environ.get("KLM_BK_WINDOW", "3600"))
    ring = (torch.rand((s, t), device="cuda") * 100).float()
    seen = torch.full((s,), t, dtype=torch.int32, device="cuda")
    tick = torch.rand(s, device="cuda")
    k6 = dev.prebind("ring_ingest", ring, seen, t, tick)
    timeit(f"K6 ring_ingest S={s}", k6, 8 * s, "one 32 B sector per sample in DRAM", do_flush=False)
    if t == 3600:  # C3's ingest tick: 5000 nodes x 64 channels (rows of a larger ring; same scatter)
        s3 = 320_000
        ring3 = torch.empty((s3, t), dtype=torch.float32, device="cuda")
        seen3 = torch.full((s3,), t, dtype=torch.int32, device="cuda")
        k63 = dev.prebind("ring_ingest", ring3, seen3, t, torch.rand(s3, device="cuda"))
        timeit(f"K6 ring_ingest S={s3} (C3 tick)", k63, 8 * s3, "sector traffic 36 B/sample (4 read + 32 written)",
               do_flush=False, units=("samples/s", s3))
        del ring3, seen3
    o = [torch.empty(s, dtype=torch.float32, device="cuda") for _ in range(3)]
    om = torch.empty(s, dtype=torch.float64, device="cuda")
    for variant in ("warp", "stream", "block"):
        k7 = dev.prebind("window_stats", ring, seen, t, o[0], o[1], om, o[2], variant=variant)
        timeit(f"K7 window_stats[{variant}] S={s} T={t}", k7, s * (4 * t + 20), do_flush=False,
               units=("samples/s", s * t))
    if t <= 8192 and not os.environ.get("KLM_BK_ONCE"):
        # the streaming product path: one fresh sample per series between launches, brackets carried
        # in `state`; a train of (tick, K7) pairs minus the train of ticks alone
        state = torch.full((s, 2), float("nan"), dtype=torch.float32, device="cuda")
        ticks = (torch.rand((256, s), device="cuda") * 100).float()
        k6s = [dev.prebind("ring_ingest", ring, seen, t, ticks[i]) for i in range(256)]
        kh = dev.prebind("window_stats", ring, seen, t, o[0], o[1], om, o[2], state=state)
        cur = [0]

        def pair():
            k6s[cur[0] % 256]()
            cur[0] += 1
            kh()

        def tick_only():
            k6s[cur[0] % 256]()
            cur[0] += 1
        for _ in range(6):
            pair()

        def train(fn, iters=20):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            for _ in range(iters):
                fn()
            e1.record()
            e1.synchronize()
            return e0.elapsed_time(e1) / iters
        tt = float(np.median([train(pair) for _ in range(7)])) - float(np.median([train(tick_only) for _ in range(7)]))
        alg = s * (4 * t + 36)
        print(json.dumps({"kernel": f"K7 window_stats[hinted: brackets carried, landing ring] S={s} T={t}",
                          "us": round(tt * 1e3, 2), "algorithmic_bytes": int(alg), "GB/s": round(alg / tt / 1e6, 1),
                          "frac_of_measured_hbm_peak": round(alg / tt / 1e6 / peak, 4),
                          "note": "train of (tick, K7) pairs minus the train of ticks", "samples/s": round(s * t / (tt * 1e-3), 1)}),
              flush=True)


if __name__ == "__main__":
    main()
