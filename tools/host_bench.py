"""Throughput of the rows either side of the hot path at the C3 shape:
JSON lists in (UAVMetric items, SchedulingRequests) -> C++ decode + name resolution ->
one GPU reconcile -> status maps out (controller.go:229-236 fields, same JSON).
Timing only; parity of this layer is tests/test_gpu_host.py (the oracle stays in tests/).
One JSON line on stdout."""
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402


def main(n_nodes=5000, n_req=5000):
    pkg = ge.load_package()
    rng = np.random.default_rng(3)
    items = []
    for i in range(n_nodes):
        st = "active" if rng.random() > 0.05 else "stale"
        it = {"spec": {"uav_id": f"uav-{i}", "node_name": f"Node-{i:05d}" if rng.random() > 0.02 else "",
                       "battery": {"remaining_percent": round(float(rng.random() * 100), 1)}},
              "status": {"last_update": "2025-10-17T00:00:00Z", "collection_status": st}}
        items.append(it)
    reqs = []
    for r in range(n_req):
        k = int(rng.integers(0, 3))
        pref = [f"node-{int(rng.integers(0, n_nodes)):05d}" for _ in range(k)]   # lower-cased on purpose
        reqs.append({"spec": {"workload": {"name": f"job-{r}", "namespace": "default"},
                              "minBatteryPercent": float(rng.choice([0, 20, 50, 80])),
                              "preferredNodes": pref}})
    uav_json = json.dumps({"items": items})
    req_json = json.dumps({"items": reqs})

    def timed(fn, reps=5):
        best = 1e9
        for _ in range(reps):
            t0 = time.perf_counter()
            out = fn()
            best = min(best, time.perf_counter() - t0)
        return best, out

    with pkg.Handle(0) as h:
        host = pkg.hostbinding.HostSession(h)
        host.clear()
        t_uav, n1 = timed(lambda: (host.clear(), host.load_uavs_json(uav_json))[1])
        def load_both():
            host.clear()
            host.load_uavs_json(uav_json)
            return host.load_requests_json(req_json, apimachinery_ints=False)
        t_both, n2 = timed(load_both)
        t_req = t_both - t_uav
        assert n1 == n_nodes and n2 == n_req
        lib, s = host.lib, host.s
        t_rec, n3 = timed(lambda: lib.klmh_reconcile(s), reps=20)
        assert n3 == n_req
        buf = C.create_string_buffer(4096)
        stamp = b"2025-10-17T00:00:00Z"

        def encode_all():
            tot = 0
            for i in range(n_req):
                lib.klmh_status_json(s, i, stamp, buf, 4096)
                tot += len(buf.value)
            return tot
        t_enc_py, nbytes = timed(encode_all, reps=3)   # one ctypes call per status (Python's overhead)
        lib.klmh_status_json_all.restype = C.c_int64
        lib.klmh_status_json_all.argtypes = [C.c_void_p, C.c_char_p]
        t_enc, nbytes2 = timed(lambda: lib.klmh_status_json_all(s, stamp), reps=5)
        assert nbytes2 == nbytes
        host.close()
    line = {"what": "host layer, C3 shape", "nodes": n_nodes, "requests": n_req,
            "uav_json_bytes": len(uav_json), "request_json_bytes": len(req_json),
            "decode_uav_items_per_s": n_nodes / t_uav, "decode_requests_per_s": n_req / t_req,
            "decode_MB_per_s": (len(uav_json) + len(req_json)) / (t_uav + t_req) / 1e6,
            "reconcile_ms": t_rec * 1e3, "reconcile_requests_per_s": n_req / t_rec,
            "status_encode_per_s": n_req / t_enc, "status_encode_per_s_one_ctypes_call_each": n_req / t_enc_py,
            "status_json_bytes": nbytes,
            "end_to_end_requests_per_s": n_req / (t_uav + t_req + t_rec + t_enc),
            "parity": "tests/test_gpu_host.py"}
    print(json.dumps(line))


if __name__ == "__main__":
    main(*(int(a) for a in sys.argv[1:3]))
