#!/usr/bin/env python
"""Which way do the series of C2 / C3 take through the hinted K7, tick after tick? Needs the
tuning build with path counters:  bash tools/build_variant.sh stats "-DKLM_WSS_STATS"
   KLM_LIB_NAME=libklm_stats.so python tools/k7_paths.py"""
import ctypes as C
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

pkg = ge.load_package()
dev, syn = pkg.Dev(), pkg.synth
names = ["tie_held", "bracket_held", "bracket_missed", "sampled", "narrow", "short", "retried", "stepped_below"]


def run(label, ring, ticks):
    s, t = ring.shape
    seen = torch.full((s,), t, dtype=torch.int32, device="cuda")
    state = torch.full((s, 2), float("nan"), dtype=torch.float32, device="cuda")
    o = [torch.empty(s, dtype=torch.float32, device="cuda") for _ in range(3)]
    om = torch.empty(s, dtype=torch.float64, device="cuda")
    buf = (C.c_ulonglong * 8)()
    for i in range(ticks.shape[0]):
        dev.ring_ingest(ring, seen, t, ticks[i].contiguous())
        dev.lib.klm_dbg_wss_paths(None, 1)
        dev.window_stats(ring, seen, t, o[0], o[1], om, o[2], state=state)
        torch.cuda.synchronize()
        dev.lib.klm_dbg_wss_paths(buf, 0)
        if i in (0, 1, 2, 5, 11, 100, 400, 800, 1200, ticks.shape[0] - 1):
            print(json.dumps({"workload": label, "tick": i, **{n: int(buf[k]) for k, n in enumerate(names)}}), flush=True)


dev.lib.klm_dbg_wss_paths.argtypes = [C.c_void_p, C.c_int]
T = 3600
# C3: 400 UAV nodes x 64 channels (a slice of the 5000)
rows = syn.uav_channel_rows(3, 0, 400, T)
ring = torch.from_numpy(rows).cuda()
ticks = ring[:, torch.arange(0, T, T // 12, device="cuda")[:12]].t().contiguous()
run("c3 (400 nodes)", ring, ticks)
# C2
scales = syn.c2_series_scales(1000, 10_000)
ring = syn.c2_ring_rows_torch(2, scales, 0, 26000, T)
sc = torch.as_tensor(scales, dtype=torch.float64, device="cuda")[None, :]
NT = 1600
ticks = (syn.uniform_torch(syn.seed_for(2, 31), NT * 26000, 0).reshape(NT, 26000) * sc).to(torch.float32)
run("c2", ring, ticks)
