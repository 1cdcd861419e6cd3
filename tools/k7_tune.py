#!/usr/bin/env python
"""K7 tuning harness: time klm_dev_window_stats of several builds of the library on the C2 ring.

  python tools/k7_tune.py libklm.so libklm_foo.so ...      (parent: one subprocess per build)

Each build runs in its own process (the library is loaded once per process), on the same seeded
ring: 26 000 series x 3600 samples, three distributions (uniform, the C2 generator's scaled
uniform, quantised with heavy ties). Prints one JSON line per build with the median time of
trains of 20 launches, and a digest of the outputs; the parent flags digests that differ from
the first build's. Experiments only; the driver's bench is bench.py.
"""
from __future__ import annotations

import hashlib
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def worker():
    import numpy as np
    import torch
    sys.path.insert(0, ROOT)
    import __graft_entry__ as ge
    pkg = ge.load_package()
    dev = pkg.Dev()
    s, t = int(os.environ.get("K7_SERIES", 26000)), int(os.environ.get("K7_WINDOW", 3600))
    pitch = (t + 3) & ~3
    g = torch.Generator(device="cuda").manual_seed(1234)
    out = {"lib": os.environ.get("KLM_LIB_NAME", "libklm.so")}
    digest = hashlib.sha256()
    for dist in ("uniform", "quantised"):
        ring = torch.rand((s, pitch), device="cuda", generator=g) * 100
        if dist == "quantised":
            ring = torch.round(ring / 5) * 5
        ring = ring.float().contiguous()
        seen = torch.full((s,), t, dtype=torch.int32, device="cuda")
        o = [torch.empty(s, dtype=torch.float32, device="cuda") for _ in range(3)]
        om = torch.empty(s, dtype=torch.float64, device="cuda")
        k7 = dev.prebind("window_stats", ring, seen, t, o[0], o[1], om, o[2])
        for _ in range(5):
            k7()
        torch.cuda.synchronize()
        reps = []
        for _ in range(7):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(20):
                k7()
            e1.record()
            e1.synchronize()
            reps.append(e0.elapsed_time(e1) / 20 * 1e3)
        out[dist + "_us"] = round(float(np.median(reps)), 2)
        if hasattr(dev.lib, "klm_dev_window_stats_hinted") and t <= 8192:
            # streaming: one fresh sample per series between calls, the bracket carried in `state`
            state = torch.full((s, 2), float("nan"), dtype=torch.float32, device="cuda")
            ticks = torch.rand((64, s), device="cuda", generator=g) * 100
            if dist == "quantised":
                ticks = torch.round(ticks / 5) * 5
            k6 = [dev.prebind("ring_ingest", ring, seen, t, ticks[i].contiguous()) for i in range(64)]
            kh = dev.prebind("window_stats", ring, seen, t, o[0], o[1], om, o[2], state=state)
            kc = dev.prebind("window_stats", ring, seen, t, o[0], o[1], om, o[2], variant="stream")
            def train(fns, reps=20):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for i in range(reps):
                    for f in fns:
                        (f[(7 + i) % 64] if isinstance(f, list) else f)()
                e1.record()
                e1.synchronize()
                return e0.elapsed_time(e1) / reps * 1e3
            for name, fn in (("hinted", kh), ("stream_cold", kc)):
                for i in range(6):
                    k6[i](); fn()
                # one tick + one K7 per trip, minus the ticks alone (single launches are below the
                # ~2 us resolution of CUDA events)
                both = float(np.median([train([k6, fn]) for _ in range(7)]))
                alone = float(np.median([train([k6]) for _ in range(7)]))
                out[f"{dist}_{name}_us"] = round(both - alone, 2)
            st = state.cpu().numpy()
            out[dist + "_bracketed_frac"] = round(float(np.mean(st[:, 0] < st[:, 1])), 3)
            out[dist + "_tie_frac"] = round(float(np.mean(st[:, 0] == st[:, 1])), 3)
            digest.update(o[2].cpu().numpy().tobytes())
        out[dist + "_min_us"] = round(float(min(reps)), 2)
        for x in (o[0], o[1], o[2]):
            digest.update(x.cpu().numpy().tobytes())
        digest.update(np.round(om.cpu().numpy(), 6).tobytes())
        del ring
    out["digest"] = digest.hexdigest()[:16]
    print(json.dumps(out), flush=True)


def main():
    if os.environ.get("K7_WORKER"):
        return worker()
    libs = sys.argv[1:] or ["libklm.so"]
    first = None
    for lib in libs:
        env = dict(os.environ, K7_WORKER="1", KLM_LIB_NAME=lib)
        r = subprocess.run([sys.executable, os.path.abspath(__file__)], env=env, stdout=subprocess.PIPE,
                           stderr=subprocess.STDOUT, text=True)
        line = r.stdout.strip().splitlines()[-1] if r.stdout.strip() else ""
        try:
            d = json.loads(line)
            first = first or d["digest"]
            d["same_as_first"] = d["digest"] == first
            print(json.dumps(d), flush=True)
        except Exception:
            print(json.dumps({"lib": lib, "error": r.stdout[-600:]}), flush=True)


if __name__ == "__main__":
    main()
