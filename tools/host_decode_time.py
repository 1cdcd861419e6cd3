"""CPU-only timing of the wire decoders at the C3 shape (no GPU needed): JSON lists -> decoded
items (tools/host_bench.py times the whole host layer, reconcile included, on a GPU box)."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

pkg = ge.load_package()
rng = np.random.default_rng(3)
n_nodes = n_req = 5000
items = [{"spec": {"uav_id": f"uav-{i}", "node_name": f"Node-{i:05d}",
                   "battery": {"remaining_percent": round(float(rng.random() * 100), 1)}},
          "status": {"last_update": "2025-10-17T00:00:00Z", "collection_status": "active"}} for i in range(n_nodes)]
reqs = [{"spec": {"workload": {"name": f"job-{r}", "namespace": "default"},
                  "minBatteryPercent": float(rng.choice([0, 20, 50, 80])),
                  "preferredNodes": [f"node-{int(rng.integers(0, n_nodes)):05d}" for _ in range(int(rng.integers(0, 3)))]}}
        for r in range(n_req)]
uj, rj = json.dumps({"items": items}), json.dumps({"items": reqs})
s = pkg.hostbinding.HostSession(None)


def best(fn, reps=7):
    b = 1e9
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        b = min(b, time.perf_counter() - t0)
    return b


tu = best(lambda: (s.clear(), s.load_uavs_json(uj)))
tr = best(lambda: (s.clear(), s.load_requests_json(rj, apimachinery_ints=False)))
print(json.dumps({"uav_items_per_s": n_nodes / tu, "requests_per_s": n_req / tr,
                  "MB_per_s": (len(uj) + len(rj)) / (tu + tr) / 1e6, "uav_ms": tu * 1e3, "req_ms": tr * 1e3}))
