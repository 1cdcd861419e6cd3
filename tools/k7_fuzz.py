#!/usr/bin/env python
"""Fuzz the window statistics against the oracle: random window sizes, fill levels and
adversarial value mixes (few distinct values, plateaus next to spikes, signed zeros, infinities,
denormals, huge / tiny magnitudes, sorted runs, periodic patterns that pile one lane up), every K7
variant stateless, then runs of ticks with carried brackets (cold, carried, garbage, tie values).
   python tools/k7_fuzz.py [seconds] [seed]      -> one JSON line; exit 1 on the first mismatch
The oracle is test infrastructure: this tool checks the product, it is not part of it."""
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

SPECIAL = np.array([0.0, -0.0, 1.0, -1.0, np.inf, -np.inf, 1e-45, -1e-45, 1.17549435e-38, 3.4028235e38,
                    -3.4028235e38, 5.0, 5.0000005, 100.0, 1e-30], np.float32)


def series(rng, t):
    kind = int(rng.integers(0, 14))
    if kind == 0:
        v = rng.random(t) * 10.0 ** rng.integers(-3, 6)
    elif kind == 1:   # few distinct values
        v = rng.choice(rng.standard_normal(int(rng.integers(1, 6))) * 10, t)
    elif kind == 2:   # plateau + about 5 % spikes (count near the p95 rank)
        base = float(rng.choice(SPECIAL[:4]))
        frac = 0.05 + rng.standard_normal() * 0.01
        v = np.where(rng.random(t) < frac, base + rng.random(t) * 50 + 1e-3, base)
    elif kind == 3:   # specials only
        v = rng.choice(SPECIAL, t)
    elif kind == 4:   # specials mixed into noise
        v = np.where(rng.random(t) < 0.1, rng.choice(SPECIAL, t), rng.standard_normal(t))
    elif kind == 5:   # sorted runs
        v = np.sort(rng.standard_normal(t))
        if rng.random() < 0.5:
            v = v[::-1]
    elif kind == 6:   # period-128 pattern: the large values all fall to one lane
        v = np.where((np.arange(t) % 128) < int(rng.integers(1, 9)), 100 + rng.random(t), rng.random(t))
    elif kind == 7:   # dips below a plateau (the p95 is the plateau, the minimum a dip)
        v = np.where(rng.random(t) < 0.9, 7.0, 7.0 - rng.random(t) * 3)
    elif kind == 8:   # denormals
        v = (rng.integers(-50, 50, t) * 1.4e-45)
    elif kind == 9:   # two clusters far apart
        v = np.where(rng.random(t) < 0.06, 1e30 * (1 + rng.random(t)), rng.random(t) * 1e-30)
    elif kind == 10:  # integers 0..3
        v = rng.integers(0, 4, t).astype(np.float64)
    elif kind == 11:  # constant
        v = np.full(t, float(rng.choice(SPECIAL)))
    elif kind == 12:  # monotone ramp with a flat end
        v = np.minimum(np.arange(t) * 0.01, t * 0.0095)
    else:             # lognormal
        v = np.exp(rng.standard_normal(t) * 4)
    with np.errstate(over="ignore", invalid="ignore"):
        v = v.astype(np.float32)
    v[np.isnan(v)] = 0.0
    return v


def same(a, b):
    return np.array_equal(a.view(np.uint32), b.view(np.uint32))


def check(got, want, what, rows=None, seen=None, t=None):
    for k in ("min", "max", "p95"):
        if not same(got[k], want[k]):
            bad = np.flatnonzero(got[k].view(np.uint32) != want[k].view(np.uint32))[:4]
            info = {"fail": what, "stat": k, "series": bad.tolist(),
                    "got": [float(x) for x in got[k][bad]], "want": [float(x) for x in want[k][bad]]}
            if rows is not None:
                i = int(bad[0])
                info["seen"] = int(seen[i])
                info["row"] = [float(x) for x in rows[i, :min(t, 140)]]
                os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
                np.savez(os.path.join(ROOT, "gpurun_out", "k7_fuzz_fail.npz"), rows=rows, seen=np.asarray(seen), t=t)
            raise SystemExit(json.dumps(info))
    # mean: an f64 sum whose association order is the implementation's. It is exact (and then
    # bit-equal) unless the window mixes magnitudes more than 2^29 apart, e.g. +-3.4e38 next to 100;
    # the tolerance is therefore relative to the size of the terms, not of the (cancelled) result.
    if rows is not None:
        n = np.minimum(seen.astype(np.int64), t)
        with np.errstate(invalid="ignore", over="ignore"):
            mag = np.array([np.abs(rows[i, :n[i]].astype(np.float64)).sum() / max(n[i], 1) for i in range(len(n))])
    else:
        mag = np.abs(want["mean"])
    with np.errstate(invalid="ignore"):
        g, w = got["mean"], want["mean"]
        ok_each = (g == w) | (np.isnan(g) & np.isnan(w)) | (np.abs(g - w) <= 1e-5 * mag + 1e-300) | ~np.isfinite(mag)
    if not ok_each.all():
        bad = np.flatnonzero(~ok_each)[:4]
        info = {"fail": what, "stat": "mean", "series": bad.tolist(), "got": [float(x) for x in g[bad]],
                "want": [float(x) for x in w[bad]]}
        if rows is not None:
            i = int(bad[0])
            info["seen"] = int(seen[i])
            info["row"] = [float(x) for x in rows[i, :min(t, 140)]]
        raise SystemExit(json.dumps(info))


def main():
    budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
    seed = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    rng = np.random.default_rng(seed)
    dev = ge.load_package().Dev()
    t_end = time.time() + budget
    rounds = calls = 0
    sizes = [1, 2, 3, 4, 5, 7, 31, 32, 33, 127, 128, 129, 130, 255, 256, 895, 896, 897, 1000, 2047, 2048, 3600, 4095,
             4096, 4097, 6000, 8191, 8192, 8193, 12000]
    while time.time() < t_end:
        t = int(rng.choice(sizes)) if rng.random() < 0.7 else int(rng.integers(1, 9000))
        s = int(rng.integers(33, 200))
        pitch = (t + 3) & ~3
        rows = np.zeros((s, pitch), np.float32)
        for i in range(s):
            rows[i, :t] = series(rng, t)
        mode = rng.random(s)
        seen_np = np.where(mode < 0.6, t, np.where(mode < 0.8, rng.integers(0, t + 1, s), t + rng.integers(0, 3 * t + 1, s)))
        seen_np = seen_np.astype(np.int64)
        ring = torch.from_numpy(rows).cuda()
        seen = torch.from_numpy(seen_np.astype(np.uint32).view(np.int32)).cuda()
        o = {k: torch.empty(s, dtype=torch.float64 if k == "mean" else torch.float32, device="cuda")
             for k in ("min", "max", "mean", "p95")}
        want = ob.window_stats(rows, seen_np.astype(np.uint32), t, pitch)
        variants = [None, "block"] + (["warp"] if t <= 4096 else []) + (["stream"] if t <= 8192 else [])
        for v in variants:
            for kk in o.values():
                kk.fill_(-123.0)
            dev.window_stats(ring, seen, t, o["min"], o["max"], o["mean"], o["p95"], variant=v)
            torch.cuda.synchronize()
            check({k: x.cpu().numpy() for k, x in o.items()}, want, f"seed={seed} round={rounds} t={t} variant={v}",
                  rows, seen_np, t)
            calls += 1
        if t <= 8192:
            state = torch.full((s, 2), float("nan"), dtype=torch.float32, device="cuda")
            for tick in range(int(rng.integers(3, 12))):
                r = rng.random()
                if r < 0.15:     # garbage brackets
                    state.copy_(torch.from_numpy((rng.standard_normal((s, 2)) * 10.0 ** rng.integers(-2, 4)).astype(np.float32)).cuda())
                elif r < 0.25:   # tie values out of nowhere
                    g = torch.from_numpy(rng.choice(SPECIAL, s)).cuda()
                    state.copy_(torch.stack([g, g], 1))
                vals_np = np.array([series(rng, 1)[0] if rng.random() < 0.5 else rows[i, int(rng.integers(0, t))]
                                    for i in range(s)], np.float32)
                dev.ring_ingest(ring, seen, t, torch.from_numpy(vals_np).cuda())
                dev.window_stats(ring, seen, t, o["min"], o["max"], o["mean"], o["p95"], state=state)
                torch.cuda.synchronize()
                sn = seen.cpu().numpy().view(np.uint32)
                want = ob.window_stats(ring.cpu().numpy(), sn, t, pitch)
                check({k: x.cpu().numpy() for k, x in o.items()}, want,
                      f"seed={seed} round={rounds} t={t} hinted tick={tick}", ring.cpu().numpy(), sn, t)
                calls += 1
        rounds += 1
    print(json.dumps({"ok": True, "seed": seed, "rounds": rounds, "kernel_calls_checked": calls, "seconds": budget}))


if __name__ == "__main__":
    main()
