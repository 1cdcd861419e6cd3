"""ctypes binding of oracle/libklm_oracle.so — the CPU checker. TEST INFRASTRUCTURE:
only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs import this."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
LIB = os.path.join(ORACLE_DIR, "libklm_oracle.so")

_lib = None


class ClusterOut(C.Structure):
    _fields_ = [(n, C.c_int64) for n in ("total_nodes", "healthy_nodes", "total_pods", "running_pods",
                                         "total_cpu", "used_cpu", "total_memory", "used_memory",
                                         "total_gpus", "available_gpus")] + [
        ("cpu_usage_rate", C.c_double), ("memory_usage_rate", C.c_double),
        ("health_status", C.c_int32), ("issues", C.c_uint32)]


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        src = os.path.join(ORACLE_DIR, "klm_oracle.c")
        if not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(src):
            subprocess.run(["make", "-C", ORACLE_DIR], check=True, stdout=subprocess.DEVNULL)
        _lib = C.CDLL(LIB)
    return _lib


def _a(x, dt):
    return np.ascontiguousarray(x, dtype=dt)


def _p(x):
    return None if x is None else C.c_void_p(x.ctypes.data)


def _schedule(fn_name, battery, flags, min_battery, pref_off, pref_ids, threads=None):
    b, f = _a(battery, np.float64), _a(flags, np.uint8)
    mb, off, ids = _a(min_battery, np.float64), _a(pref_off, np.int32), _a(pref_ids, np.int32)
    if ids.size == 0:
        ids = np.zeros(1, np.int32)
    r = mb.size
    oi, osc = np.empty(r, np.int32), np.empty(r, np.float64)
    fn = getattr(lib(), fn_name)
    fn.restype = C.c_int
    args = [C.c_int64(b.size), _p(b), _p(f), C.c_int64(r), _p(mb), _p(off), _p(ids), _p(oi), _p(osc)]
    if threads is not None:
        args.append(C.c_int(threads))
    rc = fn(*args)
    assert rc == 0
    return oi, osc


def schedule_literal(battery, flags, min_battery, pref_off, pref_ids):
    """controller.go:174-221 + :160-161, candidate list + stable sort."""
    return _schedule("klmo_schedule_literal", battery, flags, min_battery, pref_off, pref_ids)


def schedule_scan(battery, flags, min_battery, pref_off, pref_ids, threads=None):
    if threads:
        return _schedule("klmo_schedule_scan_mt", battery, flags, min_battery, pref_off, pref_ids, threads)
    return _schedule("klmo_schedule_scan", battery, flags, min_battery, pref_off, pref_ids)


def schedule_decomposed(battery, flags, min_battery, pref_off, pref_ids):
    return _schedule("klmo_schedule_decomposed", battery, flags, min_battery, pref_off, pref_ids)


def cluster_aggregate(snap: dict) -> dict:
    cols = [_a(snap["cpu_capacity"], np.int64), _a(snap["cpu_usage"], np.int64),
            _a(snap["memory_capacity"], np.int64), _a(snap["memory_usage"], np.int64),
            _a(snap["gpu_count"], np.int32), _a(snap["healthy"], np.uint8)]
    gu, pr = _a(snap["gpu_usage"], np.float64), _a(snap["pod_running"], np.uint8)
    out = ClusterOut()
    f = lib().klmo_cluster_aggregate
    f.restype = None
    f(C.c_int64(cols[0].size), *[_p(c) for c in cols], _p(gu), C.c_int64(gu.size), _p(pr),
      C.c_int64(pr.size), C.byref(out))
    return {n: getattr(out, n) for n, _ in out._fields_}


def node_rates(snap: dict, with_alloc=True) -> dict:
    a = [_a(snap[k], np.int64) for k in ("cpu_capacity", "cpu_usage", "memory_capacity", "memory_usage",
                                         "disk_capacity")]
    da = _a(snap["disk_allocatable"], np.int64) if with_alloc else None
    n = a[0].size
    out = {"cpu_rate": np.empty(n), "memory_rate": np.empty(n), "disk_usage": np.empty(n, np.int64),
           "disk_rate": np.empty(n), "under_pressure": np.empty(n, np.uint8)}
    f = lib().klmo_node_rates
    f.restype = None
    f(C.c_int64(n), *[_p(x) for x in a], _p(da), *[_p(v) for v in out.values()])
    return out


def pod_rates(c: dict) -> dict:
    off = _a(c["ctr_off"], np.int32)
    cols = [_a(c[k], np.int64) for k in ("cpu_usage", "memory_usage", "cpu_request", "cpu_limit",
                                         "memory_request", "memory_limit")]
    p = off.size - 1
    out = {"pod_totals": np.empty((6, p), np.int64), "cpu_rate": np.empty(p), "memory_rate": np.empty(p),
           "cpu_util": np.empty(p), "memory_util": np.empty(p), "over_limit": np.empty(p, np.uint8)}
    f = lib().klmo_pod_rates
    f.restype = None
    f(C.c_int64(p), _p(off), *[_p(x) for x in cols], *[_p(v) for v in out.values()])
    return out


def fleet_scan(status_ok, pct, threshold):
    ok, pc = _a(status_ok, np.uint8), _a(pct, np.float64)
    low = np.empty(max(ok.size, 1), np.int32)
    h, nl = C.c_int64(0), C.c_int64(0)
    f = lib().klmo_fleet_scan
    f.restype = None
    f(C.c_int64(ok.size), _p(ok), _p(pc), C.c_double(threshold), C.byref(h), _p(low), C.byref(nl))
    return h.value, low[: nl.value].copy()


def ring_ingest(ring, seen, window, values, series_ids=None):
    """In-place on ring (f32 [S, pitch]) and seen (u32 [S])."""
    assert ring.dtype == np.float32 and seen.dtype == np.uint32 and ring.flags.c_contiguous
    v = _a(values, np.float32)
    ids = None if series_ids is None else _a(series_ids, np.int64)
    f = lib().klmo_ring_ingest
    f.restype = None
    f(_p(ring), _p(seen), C.c_int64(ring.shape[0]), C.c_int32(window), C.c_int32(ring.shape[1]),
      C.c_int64(v.size), _p(ids), _p(v))


def window_stats(ring, seen, window, pitch=None, s0=0, s1=None, fast=False, threads=1) -> dict:
    r = _a(ring, np.float32)
    sn = _a(seen, np.uint32)
    if pitch is None:
        pitch = r.shape[1]
    if s1 is None:
        s1 = sn.size
    n = s1 - s0
    out = {"min": np.empty(n, np.float32), "max": np.empty(n, np.float32), "mean": np.empty(n, np.float64),
           "p95": np.empty(n, np.float32)}
    args = [_p(r), _p(sn), C.c_int32(window), C.c_int32(pitch), C.c_int64(s0), C.c_int64(s1),
            _p(out["min"]), _p(out["max"]), _p(out["mean"]), _p(out["p95"])]
    if fast:
        f = lib().klmo_window_stats_fast
        args.append(C.c_int(threads))
    else:
        f = lib().klmo_window_stats
    f.restype = C.c_int
    assert f(*args) == 0
    return out


def node_available(cpu_cap, cpu_use, mem_cap, mem_use, disk_cap, disk_use):
    a = [_a(x, np.int64) for x in (cpu_cap, cpu_use, mem_cap, mem_use, disk_cap, disk_use)]
    n = a[0].size
    out = [np.empty(n), np.empty(n), np.empty(n)]
    f = lib().klmo_node_available
    f.restype = None
    f(C.c_int64(n), *[_p(x) for x in a], *[_p(v) for v in out])
    return out


def network_quality(connected, rtt):
    c, r = _a(connected, np.uint8), _a(rtt, np.float64)
    q = np.empty(c.size, np.uint8)
    f = lib().klmo_network_quality
    f.restype = None
    f(C.c_int64(c.size), _p(c), _p(r), _p(q))
    return q


def pod_restarts(st_off, restart_count):
    off, cnt = _a(st_off, np.int32), _a(restart_count, np.int32)
    out = np.empty(off.size - 1, np.int32)
    f = lib().klmo_pod_restarts
    f.restype = None
    f(C.c_int64(out.size), _p(off), _p(cnt), _p(out))
    return out


def node_health(cond_off, cond_type, cond_status):
    off, t, s = _a(cond_off, np.int32), _a(cond_type, np.uint8), _a(cond_status, np.uint8)
    n = off.size - 1
    healthy, failing = np.empty(n, np.uint8), np.empty(n, np.uint8)
    f = lib().klmo_node_health
    f.restype = None
    f(C.c_int64(n), _p(off), _p(t), _p(s), _p(healthy), _p(failing))
    return healthy, failing


def node_features(battery, flags, age, wmin, wmax, wp95, max_age_s, temp_limit, horizon_s, span_s):
    b, f = _a(battery, np.float64), _a(flags, np.uint8)
    a = _a(age, np.float64)
    mn, mx, p = _a(wmin, np.float32), _a(wmax, np.float32), _a(wp95, np.float32)
    bo, fo = np.empty_like(b), np.empty_like(f)
    fn = lib().klmo_node_features
    fn.restype = None
    fn(C.c_int64(b.size), _p(b), _p(f), _p(a), _p(mn), _p(mx), _p(p), C.c_double(max_age_s), C.c_double(temp_limit),
       C.c_double(horizon_s), C.c_double(span_s), _p(bo), _p(fo))
    return bo, fo


def schedule_topk(battery, flags, min_battery, pref_off, pref_ids, k):
    b, f = _a(battery, np.float64), _a(flags, np.uint8)
    mb, off, ids = _a(min_battery, np.float64), _a(pref_off, np.int32), _a(pref_ids, np.int32)
    r = mb.size
    oi, os_ = np.empty((r, k), np.int32), np.empty((r, k), np.float64)
    fn = lib().klmo_schedule_topk
    fn.restype = C.c_int
    rc = fn(C.c_int64(b.size), _p(b), _p(f), C.c_int64(r), _p(mb), _p(off), _p(ids if ids.size else np.zeros(1, np.int32)),
            C.c_int(k), _p(oi), _p(os_))
    assert rc == 0
    return oi, os_
