#!/usr/bin/env python
"""bench.py — the hot path of k8s-llm-monitor on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c3|c4|c5]
                  [--impl ours|reference]

One "step" is one pass of the hot path over one batch of synthetic input:
  c2 (default, BASELINE.json configs[1]): 1k nodes x 10k pods -> 26 000 metric series x
      3600-sample windows; tick ingest (K6) + window min/max/mean/p95 (K7) + node rates (K5)
      + cluster aggregate (K4) + single-request scoring (K1 + K2).  metric: samples/s.
  c3: 5k UAV nodes x 64 channels x 3600 + 5k requests (K=2).        metric: samples/s.
  c4: 100k nodes x 1M requests, request-sharded across ranks.        metric: decisions/s.
  c5: 1M nodes node-range-sharded, NCCL all-gather + K3 merge.       metric: decisions/s.
`value` is timed with CUDA events on the launching stream with all inputs resident in
HBM; `e2e` goes through the host-buffer C ABI (klm_* calls, pinned host arrays in and out).
--impl reference times the CPU implementation of the same path (the reference is Go and
cannot be built here, so this is the C restatement under oracle/, all host threads).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = {"c2": ("metric samples aggregated/sec", "samples/s"), "c3": ("metric samples aggregated/sec", "samples/s"),
          "c4": ("scheduling decisions/sec", "decisions/s"), "c5": ("scheduling decisions/sec", "decisions/s")}
SHAPES = {"c2": dict(nodes=1000, pods=10_000, series=26_000, window=3600, requests=1),
          "c3": dict(nodes=5000, pods=0, series=320_000, window=3600, requests=5000),
          "c4": dict(nodes=100_000, requests=1_000_000),
          "c5": dict(nodes=1_000_000, requests=1_000_000)}
CFG_ID = {"c2": 2, "c3": 3, "c4": 4, "c5": 5}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)), "measured (MEASURED_PEAKS.json)"
    return {"hbm_gbs": 6650.0}, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """Polls NVML for SM clock and throttle reasons while the timed region runs."""

    def __init__(self, index: int):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._t = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown",
                 0x4: "sw_power_cap", 0x80: "hw_power_brake_slowdown"}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, nm in names.items():
                    if r & bit:
                        self.reasons.add(nm)
            except Exception:
                pass
            time.sleep(0.002)

    def start(self):
        if self.nv:
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()

    def stop(self):
        self._stop.set()
        if self._t:
            self._t.join()
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "n_samples": len(self.samples)}


# ------------------------------------------------------------------ our arm

def run_ours(args):
    import torch
    import torch.distributed as dist
    import __graft_entry__ as ge

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = ge.load_package()
    dev = pkg.Dev()
    syn = pkg.synth
    wl = args.workload
    cid = CFG_ID[wl]
    shp = SHAPES[wl]
    cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    launches = [0]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    k7_events = []
    extra = {}

    if wl in ("c2", "c3"):
        S, T, N, P, R = shp["series"], shp["window"], shp["nodes"], shp["pods"], shp["requests"]
        s_base = rank * S  # weak scaling: every rank owns its own S series (disjoint seeds)
        if wl == "c2":
            scales = np.tile(syn.c2_series_scales(N, P), 1)
            chunks = []
            for a in range(0, S, 2000):
                b = min(S, a + 2000)
                u = syn.uniform_torch(syn.seed_for(cid, 30), (b - a) * T, (s_base + a) * T)
                sc = torch.as_tensor(scales[a:b], dtype=torch.float64, device="cuda")[:, None]
                chunks.append((u.reshape(b - a, T) * sc).to(torch.float32))
            ring = torch.cat(chunks)
            del chunks
        else:
            ring = torch.empty((S, T), dtype=torch.float32, device="cuda")
            step_nodes = 250
            for a in range(0, N, step_nodes):
                rows = syn.uav_channel_rows(cid, rank * N + a, rank * N + min(N, a + step_nodes), T)
                ring[a * 64:a * 64 + rows.shape[0]] = torch.from_numpy(rows).cuda()
        assert ring.shape[1] % 4 == 0
        seen = torch.full((S,), T, dtype=torch.int32, device="cuda")
        # fresh samples for every tick (a pool of 64 vectors, cycled): re-ingesting one
        # vector would pile up identical values in every window and distort the p95 work
        n_ticks = 64
        if wl == "c2":
            u = syn.uniform_torch(syn.seed_for(cid, 31), n_ticks * S, s_base * n_ticks).reshape(n_ticks, S)
            ticks = (u * torch.as_tensor(scales, dtype=torch.float64, device="cuda")[None, :]).to(torch.float32)
        else:
            ticks = ring[:, torch.arange(0, T, T // n_ticks, device="cuda")[:n_ticks]].t().contiguous()
        tick = ticks[0]
        o_min = torch.empty(S, dtype=torch.float32, device="cuda")
        o_max = torch.empty_like(o_min)
        o_p95 = torch.empty_like(o_min)
        o_mean = torch.empty(S, dtype=torch.float64, device="cuda")
        bat, flags = syn.node_table(cid, N)
        mb, off, ids = syn.requests(cid, R, N, k_fixed=2 if wl == "c3" else None)
        snap = syn.cluster_snapshot(cid, N, max(P, 1))
        d_bat, d_flags, d_mb, d_off = cu(bat), cu(flags), cu(mb), cu(off)
        d_ids = cu(ids if ids.size else np.zeros(1, np.int32))
        cols = {k: cu(snap[k]) for k in ("cpu_capacity", "cpu_usage", "memory_capacity", "memory_usage",
                                         "gpu_count", "healthy")}
        cols["gpu_usage"] = None
        d_pods = cu(snap["pod_running"])
        d_dcap, d_dalloc = cu(snap["disk_capacity"]), cu(snap["disk_allocatable"])
        rates = {"cpu_rate": torch.empty(N, dtype=torch.float64, device="cuda"),
                 "memory_rate": torch.empty(N, dtype=torch.float64, device="cuda"),
                 "disk_usage": torch.empty(N, dtype=torch.int64, device="cuda"),
                 "disk_rate": torch.empty(N, dtype=torch.float64, device="cuda"),
                 "under_pressure": torch.empty(N, dtype=torch.uint8, device="cuda")}
        d_cluster = torch.zeros(104, dtype=torch.uint8, device="cuda")
        best = torch.empty(2, dtype=torch.int64, device="cuda")
        o_idx = torch.empty(R, dtype=torch.int32, device="cuda")
        o_score = torch.empty(R, dtype=torch.float64, device="cuda")
        units = S * T
        # arguments are validated once; a launch in the loop is then a single ctypes call
        k6s = [dev.prebind("ring_ingest", ring, seen, T, ticks[i]) for i in range(n_ticks)]   # K6
        tick_no = [0]
        k7 = dev.prebind("window_stats", ring, seen, T, o_min, o_max, o_mean, o_p95)          # K7
        # the snapshot kernels do not touch the ring: they run on a second stream, beside K7
        side = torch.cuda.Stream()
        with torch.cuda.stream(side):
            k5 = dev.prebind("node_rates", cols, d_dcap, d_dalloc, rates)                         # K5
            k4 = dev.prebind("cluster_aggregate", cols, d_pods, d_cluster)                        # K4
            k1 = dev.prebind("node_best", d_bat, d_flags, best)                                   # K1
            k2 = dev.prebind("schedule", d_bat, d_flags, best, d_mb, d_off, d_ids, o_idx, o_score,
                             after_best=True)                                                # K2, overlapping K1
        side.wait_stream(torch.cuda.current_stream())
        extra["streams"] = "K6+K7 on the main stream, K5+K4+K1+K2 on a second stream; joined before the end event"
        ev_pool = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
                   for _ in range(args.steps)]

        def step(timed):
            k6s[tick_no[0] % n_ticks]()
            tick_no[0] += 1
            if timed:
                e0, e1 = ev_pool[len(k7_events)]
                e0.record()
            k7()
            if timed:
                e1.record()
                k7_events.append((e0, e1))
            k5(); k4(); k1(); k2()
            launches[0] += 6

        k7_bytes = S * (4 * T + 20)
        extra["decisions_per_step"] = R
    else:
        N, R = shp["nodes"], shp["requests"]
        bat, flags = syn.node_table(cid, N)
        if wl == "c4":  # request-sharded: every rank holds the whole node table, R/world requests
            r0, r1 = R * rank // world, R * (rank + 1) // world
            mb, off, ids = syn.requests(cid, r1 - r0, N, start=r0)
            n0, n1 = 0, N
        else:           # node-range-sharded: every rank holds N/world nodes and sees all requests
            n0, n1 = N * rank // world, N * (rank + 1) // world
            mb, off, ids = syn.requests(cid, R, N)
        d_bat, d_flags = cu(bat[n0:n1]), cu(flags[n0:n1])
        d_mb, d_off, d_ids = cu(mb), cu(off), cu(ids)
        rl = mb.size
        best = torch.empty(2, dtype=torch.int64, device="cuda")
        o_idx = torch.empty(rl, dtype=torch.int32, device="cuda")
        o_score = torch.empty(rl, dtype=torch.float64, device="cuda")
        peer = None
        if wl == "c5" and world > 1:
            g_idx = torch.empty(world * rl, dtype=torch.int32, device="cuda")
            g_score = torch.empty(world * rl, dtype=torch.float64, device="cuda")
            m_idx, m_score = torch.empty_like(o_idx), torch.empty_like(o_score)
            if not args.nccl_merge:
                # partial winners live in symmetric memory: K3 reads the peers' arrays in place
                # over NVLink and merges only this rank's slice of the requests
                try:
                    import torch.distributed._symmetric_memory as symm_mem
                    o_idx = symm_mem.empty(rl, dtype=torch.int32, device="cuda")
                    o_score = symm_mem.empty(rl, dtype=torch.float64, device="cuda")
                    h_i = symm_mem.rendezvous(o_idx, dist.group.WORLD)
                    h_s = symm_mem.rendezvous(o_score, dist.group.WORLD)
                    q0, q1 = R * rank // world, R * (rank + 1) // world
                    p_idx = torch.empty(q1 - q0, dtype=torch.int32, device="cuda")
                    p_score = torch.empty(q1 - q0, dtype=torch.float64, device="cuda")
                    k3p = dev.prebind("merge_partials_peer", [int(x) for x in h_i.buffer_ptrs],
                                      [int(x) for x in h_s.buffer_ptrs], q0, q1, p_idx, p_score)
                    peer = (h_i, k3p)
                except Exception as e:  # fall back to the NCCL all-gather + K3 path, and say so
                    extra["peer_merge_unavailable"] = str(e)[:200]
            extra["merge"] = ("K3 over symmetric (peer) memory, result sharded by request" if peer
                              else "NCCL all-gather + K3, result replicated")
        units = R  # whole job
        k1 = dev.prebind("node_best", d_bat, d_flags, best, node_base=n0)                     # K1
        # K2 is launched programmatically dependent on K1 (klm_dev_schedule_after_best): its loads
        # overlap K1, which alone is ~9 us of latency on a handful of SMs
        k2 = dev.prebind("schedule", d_bat, d_flags, best, d_mb, d_off, d_ids, o_idx, o_score, node_base=n0,
                         after_best=True)                                                 # K2
        ev_pool = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
                   for _ in range(args.steps)]

        def step(timed):
            if timed:
                e0, e1 = ev_pool[len(k7_events)]
                e0.record()
            k1()
            k2()
            if timed:
                e1.record()
                k7_events.append((e0, e1))
            launches[0] += 2
            if peer is not None:
                peer[0].barrier()   # every rank's partials are complete
                peer[1]()           # K3 over peer memory
                peer[0].barrier()   # nobody still reads the buffers the next step overwrites
                launches[0] += 1
            elif wl == "c5" and world > 1:
                dist.all_gather_into_tensor(g_idx, o_idx)
                dist.all_gather_into_tensor(g_score, o_score)
                dev.merge_partials(g_idx.view(world, rl), g_score.view(world, rl), m_idx, m_score)  # K3
                launches[0] += 1

        k7_bytes = 9 * (n1 - n0) + int(24 * rl + 13 * ids.size)

    if wl == "c5" and world > 1:
        # self-check (untimed): this rank's merged slice == the unsharded kernels on the whole table
        step(False)
        torch.cuda.synchronize()
        fb, ff = cu(bat), cu(flags)
        fbest = torch.empty(2, dtype=torch.int64, device="cuda")
        fi, fs = torch.empty(rl, dtype=torch.int32, device="cuda"), torch.empty(rl, dtype=torch.float64, device="cuda")
        dev.node_best(fb, ff, fbest)
        dev.schedule(fb, ff, fbest, d_mb, d_off, d_ids, fi, fs)
        torch.cuda.synchronize()
        if peer is not None:
            ok = bool(torch.equal(p_idx, fi[q0:q1]) and torch.equal(p_score, fs[q0:q1]))
        else:
            ok = bool(torch.equal(m_idx, fi) and torch.equal(m_score, fs))
        extra["sharded_equals_unsharded"] = ok
        if not ok:
            raise SystemExit("c5: node-sharded result differs from the unsharded one")
        del fb, ff, fi, fs

    # L2 hygiene: c2/c3 inputs (374 MB / 4.6 GB) exceed the 126 MB L2; c4/c5 inputs (~45 MB)
    # do not, so those workloads flush L2 between steps.
    flush = None
    if wl in ("c4", "c5"):
        flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    for _ in range(max(args.warmup, 3)):
        step(False)
        if flush is not None:
            flush.zero_()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    launches[0] = 0
    if flush is None:
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        side.wait_stream(torch.cuda.current_stream())   # the side stream starts at t0
        for _ in range(args.steps):
            step(True)
        torch.cuda.current_stream().wait_stream(side)   # ... and is joined before t1
        t1.record()
        barrier()
        total_ms = t0.elapsed_time(t1)
    else:
        total_ms = 0.0
        for _ in range(args.steps):
            flush.zero_()
            t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0.record()
            step(True)
            t1.record()
            t1.synchronize()
            total_ms += t0.elapsed_time(t1)
        barrier()
    clocks = sampler.stop()
    n_launch = launches[0]
    kern_ms = float(np.mean([a.elapsed_time(b) for a, b in k7_events]))
    if world > 1:
        tt = torch.tensor([total_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        total_ms = float(tt.item())
    ms_per_step = total_ms / args.steps
    weak = wl in ("c2", "c3")
    job_units = units * world if weak else units
    value = job_units / (ms_per_step * 1e-3)

    # ---- e2e through the host-buffer C ABI (rank-local handle, pinned host arrays)
    e2e = run_e2e(pkg, wl, cid, shp, rank, world, args, locals())
    if world > 1:
        te = torch.tensor([e2e["ms_per_step"]], dtype=torch.float64, device="cuda")
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
        e2e["ms_per_step"] = float(te.item())
    e2e_value = job_units / (e2e["ms_per_step"] * 1e-3)

    pk, pk_src = peaks()
    achieved = k7_bytes / (kern_ms * 1e-3) / 1e9
    line = {
        "metric": METRIC[wl][0], "value": value, "unit": METRIC[wl][1], "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": "weak" if weak else "strong", "vs_baseline": None,
        "dtype": "f32 samples, f64 sums and scores, int64 aggregates", "data": "synthetic",
        "config": {"workload": f"{wl}: " + ", ".join(f"{k}={v}" for k, v in shp.items()),
                   "per_rank": "own series shard" if weak else ("request shard" if wl == "c4" else "node-range shard"),
                   "l2": f"inputs exceed L2 ({k7_bytes / 1e6:.0f} MB of ring per step)" if flush is None else "256 MiB L2 flush between steps",
                   **extra},
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": METRIC[wl][1], "h2d_bytes_per_step": e2e["h2d"],
                "d2h_bytes_per_step": e2e["d2h"], "timing": "host wall clock around the klm_* calls of K steps, priming and draining included",
                "calls_per_step": "klm_collect_begin + klm_collect_end (two cycles in flight, one handle) + klm_reconcile (scheduler handle)" if weak else "klm_nodes_set + klm_schedule_batch"},
        "gpu_launches": n_launch,
        "roofline": {"bound": "hbm", "kernel": "window_stats" if weak else "node_best+schedule",
                     "achieved": achieved, "peak": pk["hbm_gbs"], "unit": "GB/s",
                     "frac": achieved / pk["hbm_gbs"], "frac_of_nominal_8TBs": achieved / 8000.0,
                     "peak_source": pk_src, "algorithmic_bytes_per_launch": k7_bytes,
                     "avg_launch_ms": kern_ms,
                     # dram__bytes_read.sum + dram__bytes_write.sum of one K7 launch at c2, from the
                     # ncu --set full capture summarised in profiles/r01_k7_warp_v3_split_ncu_summary.txt
                     "traffic": 379.089e6 if (wl == "c2" and world >= 1) else None},
    }
    if rank == 0:
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline(wl, cid, shp, budget_s=12.0)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def run_e2e(pkg, wl, cid, shp, rank, world, args, env):
    """Same step through klm_* with host buffers; host<->device copies inside the timed
    region (tick samples / node table / requests in; statistics / decisions out)."""
    import torch
    syn = pkg.synth
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()
    steps = max(3, min(args.steps, 50))
    finish, sink = None, [0.0]
    with pkg.Handle(torch.cuda.current_device()) as h:
        if wl in ("c2", "c3"):
            S, T, N = shp["series"], shp["window"], shp["nodes"]
            ring = env["ring"]
            h.ring_create(S, T)
            # history is state, loaded once (device->host->handle); a step ingests one tick
            for a in range(0, S, 4096):
                h.ring_load(a, ring[a:a + 4096, :T - 1].cpu().numpy())
            ticks_h = pin(env["ticks"].cpu().numpy())
            tick = ticks_h[0]
            e2e_tick = [0]
            bat, flags, mb, off = pin(env["bat"]), pin(env["flags"]), pin(env["mb"]), pin(env["off"])
            ids = pin(env["ids"] if env["ids"].size else np.zeros(1, np.int32))
            snap = {k: pin(v) for k, v in env["snap"].items()}
            snap_h = dict(snap)
            snap_h["gpu_usage"] = np.zeros(0)   # the C2 / C3 nodes carry no GPUs
            outs = {"cpu_rate": pin(np.empty(N)), "memory_rate": pin(np.empty(N)),
                    "disk_usage": pin(np.empty(N, np.int64)), "disk_rate": pin(np.empty(N)),
                    "under_pressure": pin(np.empty(N, np.uint8)), "min": pin(np.empty(S, np.float32)),
                    "max": pin(np.empty(S, np.float32)), "mean": pin(np.empty(S)),
                    "p95": pin(np.empty(S, np.float32))}
            col_keys = ("cpu_capacity", "cpu_usage", "memory_capacity", "memory_usage", "gpu_count", "healthy",
                        "pod_running", "disk_capacity", "disk_allocatable")
            h2d = tick.nbytes + bat.nbytes + flags.nbytes + mb.nbytes + off.nbytes + env["ids"].nbytes + \
                sum(snap[k].nbytes for k in col_keys)
            d2h = S * 20 + 104 + N * 33 + mb.size * 12

            # argument structs built once: the timed region is the C ABI calls themselves, as a
            # compiled (cgo) caller issues them, not the Python wrapper's array coercions
            res_i, res_s = pin(np.empty(mb.size, np.int32)), pin(np.empty(mb.size))
            # The metrics manager and the scheduler are separate components (separate goroutines
            # in the reference): each owns a handle. The collector keeps two cycles in flight
            # (klm_collect_begin / klm_collect_end): the tick upload and the result download of one
            # cycle overlap the window statistics of the next; every cycle's results are read.
            snap_x = {k: np.ascontiguousarray(v) for k, v in snap_h.items()}
            begin, end = h.collect_pipeline(snap_x)
            h_sched = pkg.Handle(torch.cuda.current_device())
            reconcile = h_sched.reconcile_prepared(bat, flags, mb, off, ids, res_i, res_s)
            in_flight = [0]

            def next_tick():
                t = ticks_h[e2e_tick[0] % ticks_h.shape[0]]
                e2e_tick[0] += 1
                return t

            def step():
                if not in_flight[0]:          # prime the pipeline (inside the timed region)
                    begin(next_tick())
                    in_flight[0] = 1
                begin(next_tick())            # cycle i + 1 goes in ...
                res = end()                   # ... cycle i comes out: statistics, rates, cluster metrics
                sink[0] = float(res["p95"][0]) + res["cluster"]["total_nodes"]
                reconcile()                   # the scheduler's reconcile: snapshot + batch, one call

            def finish():                     # drain: the last cycle begun is completed and read
                while in_flight[0]:
                    end()
                    in_flight[0] = 0
                h_sched.close()
        else:
            n0, n1 = env["n0"], env["n1"]
            bat, flags = pin(env["bat"][n0:n1]), pin(env["flags"][n0:n1])
            mb, off, ids = pin(env["mb"]), pin(env["off"]), pin(env["ids"])
            h2d = bat.nbytes + flags.nbytes + mb.nbytes + off.nbytes + ids.nbytes
            d2h = mb.size * 12
            if wl == "c5" and world > 1 and env.get("peer") is not None:
                dev = env["dev"]
                hb = [torch.from_numpy(x) for x in (bat, flags, mb, off, ids)]
                db = [torch.empty_like(x, device="cuda") for x in hb]
                res_i = torch.empty(env["p_idx"].numel(), dtype=torch.int32).pin_memory()
                res_s = torch.empty(env["p_idx"].numel(), dtype=torch.float64).pin_memory()
                d2h = res_i.numel() * 12

                def step():
                    for d, s_ in zip(db, hb):
                        d.copy_(s_, non_blocking=True)
                    dev.node_best(db[0], db[1], env["best"], node_base=n0)
                    dev.schedule(db[0], db[1], env["best"], db[2], db[3], db[4], env["o_idx"], env["o_score"],
                                 node_base=n0)
                    env["peer"][0].barrier()
                    env["peer"][1]()
                    env["peer"][0].barrier()
                    res_i.copy_(env["p_idx"], non_blocking=True)
                    res_s.copy_(env["p_score"], non_blocking=True)
                    torch.cuda.synchronize()
            elif wl == "c5" and world > 1:
                # node-range sharding needs the device-pointer layer + NCCL; its e2e adds
                # the H2D of the rank's inputs and the D2H of the merged result
                import torch.distributed as dist
                dev = env["dev"]
                hb = [torch.from_numpy(x) for x in (bat, flags, mb, off, ids)]
                db = [torch.empty_like(x, device="cuda") for x in hb]
                res_i = torch.empty(mb.size, dtype=torch.int32).pin_memory()
                res_s = torch.empty(mb.size, dtype=torch.float64).pin_memory()

                def step():
                    for d, s in zip(db, hb):
                        d.copy_(s, non_blocking=True)
                    dev.node_best(db[0], db[1], env["best"], node_base=n0)
                    dev.schedule(db[0], db[1], env["best"], db[2], db[3], db[4], env["o_idx"], env["o_score"],
                                 node_base=n0)
                    dist.all_gather_into_tensor(env["g_idx"], env["o_idx"])
                    dist.all_gather_into_tensor(env["g_score"], env["o_score"])
                    dev.merge_partials(env["g_idx"].view(world, -1), env["g_score"].view(world, -1),
                                       env["m_idx"], env["m_score"])
                    res_i.copy_(env["m_idx"], non_blocking=True)
                    res_s.copy_(env["m_score"], non_blocking=True)
                    torch.cuda.synchronize()
            else:
                # results land in pinned arrays, so a large batch pipelines over PCIe
                res_i, res_s = pin(np.empty(mb.size, np.int32)), pin(np.empty(mb.size))
                step = h.reconcile_prepared(bat, flags, mb, off, ids, res_i, res_s)
        for _ in range(3):
            step()
        env["barrier"]()
        t0 = time.perf_counter()
        for _ in range(steps):
            step()
        if finish is not None:
            finish()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
    return {"ms_per_step": dt / steps * 1e3, "h2d": int(h2d), "d2h": int(d2h)}


# ------------------------------------------------------------------ CPU legs (oracle; checker timed as baseline)

def cpu_step_factory(wl, cid, shp, sample_units):
    """Returns (step_fn, units_per_step, description) running the oracle on a bounded
    sample of the workload with every host thread."""
    import oracle_binding as ob
    import __graft_entry__ as ge
    syn = ge.load_package().synth
    threads = os.cpu_count() or 1
    if wl in ("c2", "c3"):
        T, N = shp["window"], shp["nodes"]
        s = max(threads, min(shp["series"], sample_units // T))
        if wl == "c2":
            scales = syn.c2_series_scales(N, shp["pods"])
            rows = syn.c2_ring_rows(cid, scales, 0, s, T)
        else:
            rows = syn.uav_channel_rows(cid, 0, (s + 63) // 64, T)[:s]
        rows = np.ascontiguousarray(rows)
        seen = np.full(s, T, np.uint32)
        bat, flags = syn.node_table(cid, N)
        mb, off, ids = syn.requests(cid, shp["requests"], N, k_fixed=2 if wl == "c3" else None)
        snap = syn.cluster_snapshot(cid, N, max(shp["pods"], 1))
        frac = s / shp["series"]
        r_s = max(1, int(round(shp["requests"] * frac)))
        sub_off = (off[: r_s + 1] - off[0]).astype(np.int32)

        def step():
            ob.window_stats(rows, seen, T, fast=True, threads=threads)
            ob.node_rates(snap)
            ob.cluster_aggregate(snap)
            ob.schedule_scan(bat, flags, mb[:r_s], sub_off, ids[: off[r_s]], threads=threads)
        return step, s * T, f"{s} of {shp['series']} series x {T} samples (+ aggregate, rates, {r_s} request(s)), {threads} threads", threads
    N, R = shp["nodes"], shp["requests"]
    r_s = max(threads, min(R, sample_units))
    bat, flags = syn.node_table(cid, N)
    mb, off, ids = syn.requests(cid, r_s, N)

    def step():
        ob.schedule_scan(bat, flags, mb, off, ids, threads=threads)
    return step, r_s, f"{r_s} of {R} requests, each scanning all {N} nodes (controller.go:181-218), {threads} threads", threads


def cpu_baseline(wl, cid, shp, budget_s):
    # calibrate on a small sample, then size the bounded sample to ~budget_s of CPU work
    probe_units = 3600 * 64 if wl in ("c2", "c3") else 64
    step, u, _, threads = cpu_step_factory(wl, cid, shp, probe_units)
    t0 = time.perf_counter()
    step()
    rate = u / max(time.perf_counter() - t0, 1e-6)
    step, u, desc, threads = cpu_step_factory(wl, cid, shp, int(rate * budget_s))
    step()  # warm the caches and the thread pool
    reps, t0 = 0, time.perf_counter()
    while reps < 3 or time.perf_counter() - t0 < min(budget_s, 8.0):
        step()
        reps += 1
        if time.perf_counter() - t0 > 2 * budget_s:
            break
    dt = (time.perf_counter() - t0) / reps
    return {"value": u / dt, "unit": METRIC[wl][1], "cores": threads, "kind": "port",
            "sample": desc + f"; mean of {reps} passes, {dt:.3f} s each",
            "note": "C restatement of the Go reference (Go toolchain unavailable); window stats have no "
                    "reference code (quickselect baseline)"}


def run_reference(args):
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    wl = args.workload
    cid, shp = CFG_ID[wl], SHAPES[wl]
    probe_units = 3600 * 64 if wl in ("c2", "c3") else 64
    step, u, _, threads = cpu_step_factory(wl, cid, shp, probe_units)
    t0 = time.perf_counter()
    step()
    rate = u / max(time.perf_counter() - t0, 1e-6)
    total_steps = args.steps + max(args.warmup, 1)
    per_step_s = min(10.0, 120.0 / total_steps)
    step, u, desc, threads = cpu_step_factory(wl, cid, shp, max(int(rate * per_step_s), probe_units))
    for _ in range(max(args.warmup, 1)):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = (time.perf_counter() - t0) / args.steps
    v = u / dt
    weak = wl in ("c2", "c3")
    line = {"impl": "reference", "metric": METRIC[wl][0], "value": v, "unit": METRIC[wl][1],
            "n_gpus": args.gpus, "steps": args.steps, "warmup": max(args.warmup, 1), "ms_per_step": dt * 1e3,
            "higher_is_better": True, "scaling": "weak" if weak else "strong", "vs_baseline": None,
            "dtype": "f32 samples, f64 sums and scores, int64 aggregates", "data": "synthetic",
            "config": {"workload": f"{wl}: " + ", ".join(f"{k}={v}" for k, v in shp.items()),
                       "note": "host CPU only; one process regardless of --gpus"},
            "cpu_baseline": {"value": v, "unit": METRIC[wl][1], "cores": threads, "kind": "port",
                             "sample": desc + " per step"},
            "e2e": {"value": v, "unit": METRIC[wl][1], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=None)
    ap.add_argument("--workload", choices=sorted(SHAPES), default="c2")
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--nccl-merge", action="store_true",
                    help="c5, N > 1: merge through an NCCL all-gather instead of peer memory")
    args = ap.parse_args()
    if args.impl == "reference":
        args.steps = args.steps or 5
        args.warmup = 1 if args.warmup is None else args.warmup
        run_reference(args)
    else:
        args.steps = args.steps or (200 if args.workload == "c2" else 50)
        args.warmup = 20 if args.warmup is None else args.warmup
        run_ours(args)


if __name__ == "__main__":
    main()
