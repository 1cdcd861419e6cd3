"""Test-side helpers: flatten the reference's wire format into the C ABI's SoA form."""
from __future__ import annotations

import json
import os

import numpy as np

import controller_oracle as co

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_kats():
    with open(os.path.join(GOLDEN, "scheduler_kat.json"), encoding="utf-8") as f:
        return json.load(f)["cases"]


def flatten_items(items):
    """UAVMetric items -> (battery f64[N], flags u8[N], lower-cased names)."""
    n = len(items)
    battery = np.zeros(n, np.float64)
    flags = np.zeros(n, np.uint8)
    names = []
    for i, it in enumerate(items):
        spec = it.get("spec", {})
        name = spec.get("node_name") if isinstance(spec.get("node_name"), str) else ""
        battery[i] = co.read_float(spec, "battery", "remaining_percent")
        status = co.read_string(it.get("status", {}), "collection_status").lower()
        if name != "":
            flags[i] |= 1
        if status in ("", "active"):
            flags[i] |= 2
        names.append(name.lower())
    return battery, flags, names


def flatten_specs(specs, names):
    """SchedulingRequestSpec list -> (min_battery, pref_off, pref_ids): every index whose
    lower-cased name is preferred (duplicate names are distinct candidates)."""
    by_name = {}
    for i, nm in enumerate(names):
        by_name.setdefault(nm, []).append(i)
    mb, off, ids = [], [0], []
    for s in specs:
        mb.append(s.min_battery_percent)
        for p in dict.fromkeys(x.lower() for x in s.preferred_nodes):
            ids.extend(by_name.get(p, []))
        off.append(len(ids))
    return np.array(mb, np.float64), np.array(off, np.int32), np.array(ids, np.int32)


def random_problem(rng, n, r, kmax=3, grid=10.0, p_bad=0.1):
    """Tie-heavy random instance for the equivalence tests."""
    battery = np.round(rng.random(n) * 100 * grid) / grid
    flags = np.full(n, 3, np.uint8)
    bad = rng.random(n)
    flags[bad < p_bad / 2] = 2
    flags[(bad >= p_bad / 2) & (bad < p_bad)] = 1
    flags[(bad >= p_bad) & (bad < p_bad * 1.1)] = 0
    mb = rng.choice([-1.0, 0.0, 20.0, 50.0, 80.0, 99.0, 101.0], size=r)
    k = rng.integers(0, kmax + 1, size=r)
    off = np.zeros(r + 1, np.int32)
    off[1:] = np.cumsum(k)
    ids = rng.integers(-1, n + 1, size=int(off[-1])).astype(np.int32)  # includes out-of-range ids
    return battery, flags, mb, off, ids
