#!/usr/bin/env python
"""Host wall clock of klm_reconcile on one handle over 1..G GPUs, for a few (nodes, requests)
shapes and both sharding modes: where the end-to-end time of the multi-device path goes."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

pkg = ge.load_package()
syn, b = pkg.synth, pkg.binding
pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()
G = torch.cuda.device_count()
for n, r in ((1_000_000, 1000), (1000, 1_000_000), (100_000, 1_000_000), (1_000_000, 1_000_000)):
    bat, flags = syn.node_table(5, n)
    mb, off, ids = syn.requests(5, r, n)
    a = [pin(x) for x in (bat, flags, mb, off, ids)]
    ri, rs = pin(np.empty(r, np.int32)), pin(np.empty(r))
    for devs in ([0], list(range(G))):
        for mode, name in ((b.SHARD_REQUESTS, "requests"), (b.SHARD_NODES, "nodes")):
            if len(devs) == 1 and mode:
                continue
            with pkg.Handle(devs) as h:
                call = h.reconcile_prepared(*a, ri, rs, algo=mode)
                for _ in range(3):
                    call()
                t0 = time.perf_counter()
                for _ in range(10):
                    call()
                dt = (time.perf_counter() - t0) / 10
            print(json.dumps({"nodes": n, "requests": r, "gpus": len(devs), "mode": name, "ms": round(dt * 1e3, 4),
                              "MB_in": round(sum(x.nbytes for x in a) / 1e6, 1), "MB_out": round((ri.nbytes + rs.nbytes) / 1e6, 1)}),
                  flush=True)
