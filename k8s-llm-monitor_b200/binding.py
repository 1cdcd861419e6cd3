"""ctypes binding of libklm.so (include/klm.h).

The library is the product; this module only declares its C signatures and
gives numpy / torch callers a thin, typed way in. There is no CPU path here:
if libklm.so is missing the import fails loudly, and every call needs a GPU.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Tuple

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# KLM_LIB_NAME picks a sibling build (tools/build_variant.sh: ablation / tuning experiments)
LIB_PATH = os.path.join(_HERE, os.environ.get("KLM_LIB_NAME", "libklm.so"))

KLM_OK, KLM_EINVAL, KLM_ECUDA, KLM_ENODEV, KLM_ESTATE, KLM_ELIMIT = 0, -1, -2, -3, -4, -5
NODE_HAS_NAME, NODE_STATUS_OK, NODE_ELIGIBLE = 1, 2, 3
ALGO_DECOMPOSED, ALGO_BRUTEFORCE = 0, 1
SHARD_REQUESTS, SHARD_NODES = 0x000, 0x100  # or-ed into the `algo` / mode argument
HEALTH_NAMES = ("healthy", "warning", "critical")  # manager.go:558-564


class KlmError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"libklm error {code}: {msg}")
        self.code = code


class NodeCols(C.Structure):  # klm_node_cols
    _fields_ = [
        ("cpu_capacity", C.c_void_p), ("cpu_usage", C.c_void_p),
        ("memory_capacity", C.c_void_p), ("memory_usage", C.c_void_p),
        ("gpu_count", C.c_void_p), ("healthy", C.c_void_p),
        ("gpu_usage", C.c_void_p), ("n_gpu_usage", C.c_int64),
    ]


class ClusterOut(C.Structure):  # klm_cluster_out
    _fields_ = [
        ("total_nodes", C.c_int64), ("healthy_nodes", C.c_int64),
        ("total_pods", C.c_int64), ("running_pods", C.c_int64),
        ("total_cpu", C.c_int64), ("used_cpu", C.c_int64),
        ("total_memory", C.c_int64), ("used_memory", C.c_int64),
        ("total_gpus", C.c_int64), ("available_gpus", C.c_int64),
        ("cpu_usage_rate", C.c_double), ("memory_usage_rate", C.c_double),
        ("health_status", C.c_int32), ("issues", C.c_uint32),
    ]

    def as_dict(self) -> dict:
        return {name: getattr(self, name) for name, _ in self._fields_}


class FeatureParams(C.Structure):  # klm_feature_params
    _fields_ = [("max_age_s", C.c_double), ("temp_limit", C.c_double), ("horizon_s", C.c_double),
                ("span_s", C.c_double)]


class CollectIn(C.Structure):  # klm_collect_in
    _fields_ = [("tick_values", C.c_void_p), ("nodes", C.POINTER(NodeCols)), ("n_nodes", C.c_int64),
                ("disk_capacity", C.c_void_p), ("disk_allocatable", C.c_void_p),
                ("pod_running", C.c_void_p), ("n_pods", C.c_int64)]


class CollectOut(C.Structure):  # klm_collect_out
    _fields_ = [("cluster", C.POINTER(ClusterOut)), ("cpu_rate", C.c_void_p), ("memory_rate", C.c_void_p),
                ("disk_usage", C.c_void_p), ("disk_rate", C.c_void_p), ("under_pressure", C.c_void_p),
                ("win_min", C.c_void_p), ("win_max", C.c_void_p), ("win_mean", C.c_void_p),
                ("win_p95", C.c_void_p), ("win_packed", C.c_void_p)]


CLUSTER_OUT_DTYPE = np.dtype([
    ("total_nodes", "<i8"), ("healthy_nodes", "<i8"), ("total_pods", "<i8"), ("running_pods", "<i8"),
    ("total_cpu", "<i8"), ("used_cpu", "<i8"), ("total_memory", "<i8"), ("used_memory", "<i8"),
    ("total_gpus", "<i8"), ("available_gpus", "<i8"), ("cpu_usage_rate", "<f8"),
    ("memory_usage_rate", "<f8"), ("health_status", "<i4"), ("issues", "<u4")])
assert CLUSTER_OUT_DTYPE.itemsize == C.sizeof(ClusterOut) == 104

_P = C.c_void_p
_I64, _I32, _F64 = C.c_int64, C.c_int32, C.c_double

# name -> (restype, argtypes); every symbol include/klm.h declares
SIGNATURES = {
    "klm_create": (C.c_int, [C.POINTER(C.c_int), C.c_int, C.POINTER(_P)]),
    "klm_destroy": (None, [_P]),
    "klm_last_error": (C.c_char_p, []),
    "klm_abi_version": (C.c_int, []),
    "klm_collect_packed_bytes": (_I64, [_I64]),
    "klm_collect_begin": (C.c_int, [_P, C.POINTER(CollectIn), C.POINTER(CollectOut)]),
    "klm_collect_end": (C.c_int, [_P]),
    "klm_host_alloc": (C.c_int, [C.c_size_t, C.POINTER(_P)]),
    "klm_host_free": (C.c_int, [_P]),
    "klm_nodes_set": (C.c_int, [_P, _I64, _P, _P]),
    "klm_schedule_batch": (C.c_int, [_P, _I64, _P, _P, _P, _P, _P, C.c_int]),
    "klm_topk_max": (C.c_int, []),
    "klm_schedule_batch_topk": (C.c_int, [_P, _I64, _P, _P, _P, C.c_int, _P, _P]),
    "klm_dev_node_topk_scratch_bytes": (_I64, []),
    "klm_dev_node_topk": (C.c_int, [_P, _P, _I64, _I64, C.c_int, _P, _P, _P]),
    "klm_dev_schedule_topk": (C.c_int, [_P, _P, _I64, _I64, _P, C.c_int, _I64, _P, _P, _P, _P, _P, _P]),
    "klm_reconcile": (C.c_int, [_P, _I64, _P, _P, _I64, _P, _P, _P, _P, _P, C.c_int]),
    "klm_cluster_aggregate": (C.c_int, [_P, C.POINTER(NodeCols), _I64, _P, _I64, C.POINTER(ClusterOut)]),
    "klm_node_rates": (C.c_int, [_P, _I64] + [_P] * 11),
    "klm_pod_rates": (C.c_int, [_P, _I64] + [_P] * 13),
    "klm_fleet_scan": (C.c_int, [_P, _I64, _P, _P, _F64, C.POINTER(_I64), _P, C.POINTER(_I64)]),
    "klm_node_available": (C.c_int, [_P, _I64] + [_P] * 9),
    "klm_network_quality": (C.c_int, [_P, _I64, _P, _P, _P]),
    "klm_pod_restarts": (C.c_int, [_P, _I64, _P, _P, _P]),
    "klm_node_health": (C.c_int, [_P, _I64, _P, _P, _P, _P, _P]),
    "klm_dev_node_available": (C.c_int, [_I64] + [_P] * 10),
    "klm_dev_network_quality": (C.c_int, [_I64, _P, _P, _P, _P]),
    "klm_dev_pod_restarts": (C.c_int, [_I64, _P, _P, _P, _P]),
    "klm_dev_node_health": (C.c_int, [_I64, _P, _P, _P, _P, _P, _P]),
    "klm_dev_node_features": (C.c_int, [_I64, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "klm_ring_create": (C.c_int, [_P, _I64, _I32]),
    "klm_ring_ingest": (C.c_int, [_P, _I64, _P, _P]),
    "klm_ring_ingest_tick": (C.c_int, [_P, _P]),
    "klm_ring_load": (C.c_int, [_P, _I64, _I64, _I32, _P]),
    "klm_ring_stats": (C.c_int, [_P, _I64, _I64, _P, _P, _P, _P]),
    "klm_collect": (C.c_int, [_P, C.POINTER(CollectIn), C.POINTER(CollectOut)]),
    "klm_dev_node_best_scratch_bytes": (_I64, []),
    "klm_dev_node_best": (C.c_int, [_P, _P, _I64, _I64, _P, _P, _P]),
    "klm_dev_node_best_static_max_nodes": (_I64, []),
    "klm_dev_node_best_static": (C.c_int, [_P, _P, _I64, _I64, _P, _P]),
    "klm_dev_schedule": (C.c_int, [_P, _P, _I64, _I64, _P, _I64, _P, _P, _P, _P, _P, _P]),
    "klm_dev_schedule_after_best": (C.c_int, [_P, _P, _I64, _I64, _P, _I64, _P, _P, _P, _P, _P, _P]),
    "klm_dev_schedule_fused": (C.c_int, [_P, _P, _I64, _I64, _P, _I64, _P, _P, _P, _P, _P, _P, _P]),
    "klm_dev_schedule_bruteforce": (C.c_int, [_P, _P, _I64, _I64, _I64, _P, _P, _P, _P, _P, _P]),
    "klm_dev_merge_partials": (C.c_int, [C.c_int, _I64, _P, _P, _P, _P, _P]),
    "klm_dev_shard_slots_bytes": (_I64, []),
    "klm_shard_range": (None, [_I64, C.c_int, C.c_int, C.POINTER(_I64), C.POINTER(_I64)]),
    "klm_dev_node_best_publish": (C.c_int, [_P, _P, _I64, _I64, C.c_int, C.c_int, _P, C.c_uint64, _P, _P, _P]),
    "klm_dev_schedule_sharded": (C.c_int, [C.c_int, _I64, _P, _P, _P, C.c_uint64, _I64, _P, _P, _P, _P, _P, C.c_int, _P]),
    "klm_dev_shard_error": (C.c_int, [_P, C.POINTER(C.c_int), _P]),
    "klm_dev_node_best_spread": (C.c_int, [_P, _P, _I64, _I64, C.c_int, C.c_int, _P, _P, _P, C.c_uint64, _P, _P, _P]),
    "klm_dev_shard_rendezvous": (C.c_int, [C.c_int, C.c_int, _P, _P, C.c_uint64, _P]),
    "klm_dev_merge_partials_peer": (C.c_int, [C.c_int, _P, _P, _I64, _I64, _P, _P, _P]),
    "klm_dev_cluster_scratch_bytes": (_I64, []),
    "klm_dev_cluster_aggregate": (C.c_int, [C.POINTER(NodeCols), _I64, _P, _I64, _P, _P, _P]),
    "klm_dev_cluster_sums": (C.c_int, [C.POINTER(NodeCols), _I64, _P, _I64, _P, _P, _P]),
    "klm_dev_cluster_finish": (C.c_int, [_P, _P, _P]),
    "klm_dev_node_rates": (C.c_int, [_I64] + [_P] * 12),
    "klm_dev_pod_rates": (C.c_int, [_I64] + [_P] * 14),
    "klm_dev_fleet_scan_scratch_bytes": (_I64, [_I64]),
    "klm_dev_fleet_scan": (C.c_int, [_I64, _P, _P, _F64, _P, _P, _P, _P]),
    "klm_dev_ring_ingest": (C.c_int, [_P, _P, _I64, _I32, _I32, _I64, _P, _P, _P]),
    "klm_dev_window_stats": (C.c_int, [_P, _P, _I32, _I32, _I64, _I64, _P, _P, _P, _P, _P]),
    "klm_dev_window_stats_stream": (C.c_int, [_P, _P, _I32, _I32, _I64, _I64, _P, _P, _P, _P, _P]),
    "klm_dev_window_stats_hinted": (C.c_int, [_P, _P, _I32, _I32, _I64, _I64, _P, _P, _P, _P, _P, _P]),
    "klm_dev_window_stats_warp": (C.c_int, [_P, _P, _I32, _I32, _I64, _I64, _P, _P, _P, _P, _P]),
    "klm_dev_window_stats_block": (C.c_int, [_P, _P, _I32, _I32, _I64, _I64, _P, _P, _P, _P, _P]),
}

_lib: Optional[C.CDLL] = None


def load() -> C.CDLL:
    """Load libklm.so and declare every signature. Fails loudly if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(make -C k8s-llm-monitor_b200/csrc). libklm has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here = header/library mismatch
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def host_array(shape, dtype) -> np.ndarray:
    """numpy array on page-locked memory from klm_host_alloc (freed when the array dies)."""
    import weakref
    lib = load()
    dt = np.dtype(dtype)
    n = int(np.prod(shape))
    p = _P()
    rc = lib.klm_host_alloc(max(1, n * dt.itemsize), C.byref(p))
    if rc != KLM_OK:
        raise KlmError(rc, last_error())
    buf = (C.c_char * max(1, n * dt.itemsize)).from_address(p.value)
    weakref.finalize(buf, lib.klm_host_free, p.value)
    return np.frombuffer(buf, dtype=dt, count=n).reshape(shape)


def last_error() -> str:
    return load().klm_last_error().decode("utf-8", "replace")


def check(rc: int) -> None:
    if rc != KLM_OK:
        raise KlmError(rc, last_error())


def _np(a, dtype) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=dtype)


def _ptr(a: Optional[np.ndarray]) -> Optional[int]:
    return None if a is None else a.ctypes.data


def shard_range(n_nodes: int, n_parts: int, part: int) -> Tuple[int, int]:
    """klm_shard_range: the node range [lo, hi) that rank `part` of `n_parts` owns."""
    lo, hi = _I64(0), _I64(0)
    load().klm_shard_range(n_nodes, n_parts, part, C.byref(lo), C.byref(hi))
    return lo.value, hi.value


class Handle:
    """Host-buffer API: numpy in, numpy out, everything computed on the GPU."""

    def __init__(self, device=0):
        """device: one ordinal, or a sequence of ordinals (one handle over several GPUs: the
        scheduling calls then spread every batch over all of them, see SHARD_*)."""
        self._lib = load()
        self._h = _P()
        ids = [int(device)] if isinstance(device, (int, np.integer)) else [int(d) for d in device]
        self.devices = ids
        dev = (C.c_int * len(ids))(*ids)
        check(self._lib.klm_create(dev, len(ids), C.byref(self._h)))

    def close(self) -> None:
        if getattr(self, "_h", None) is not None and self._h:
            self._lib.klm_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # -- scheduling (controller.go:152-161, :174-221)
    def nodes_set(self, battery, flags) -> None:
        b, f = _np(battery, np.float64), _np(flags, np.uint8)
        if b.shape != f.shape or b.ndim != 1:
            raise ValueError("battery and flags must be 1-D and the same length")
        check(self._lib.klm_nodes_set(self._h, b.size, _ptr(b), _ptr(f)))

    def schedule_batch(self, min_battery, pref_off, pref_ids, algo: int = ALGO_DECOMPOSED
                       ) -> Tuple[np.ndarray, np.ndarray]:
        mb, off, ids = _np(min_battery, np.float64), _np(pref_off, np.int32), _np(pref_ids, np.int32)
        r = mb.size
        if off.size != r + 1:
            raise ValueError("pref_off must have n_requests + 1 entries")
        out_idx = np.empty(r, np.int32)
        out_score = np.empty(r, np.float64)
        if ids.size == 0:
            ids = np.zeros(1, np.int32)
        check(self._lib.klm_schedule_batch(self._h, r, _ptr(mb), _ptr(off), _ptr(ids),
                                           _ptr(out_idx), _ptr(out_score), algo))
        return out_idx, out_score

    def schedule_batch_topk(self, min_battery, pref_off, pref_ids, k: int) -> Tuple[np.ndarray, np.ndarray]:
        """klm_schedule_batch_topk: ([R, k] indices, [R, k] scores), -1 / 0 padded."""
        mb, off, ids = _np(min_battery, np.float64), _np(pref_off, np.int32), _np(pref_ids, np.int32)
        r = mb.size
        out_idx, out_score = np.empty((r, k), np.int32), np.empty((r, k), np.float64)
        if ids.size == 0:
            ids = np.zeros(1, np.int32)
        check(self._lib.klm_schedule_batch_topk(self._h, r, _ptr(mb), _ptr(off), _ptr(ids), k, _ptr(out_idx),
                                                _ptr(out_score)))
        return out_idx, out_score

    def reconcile(self, battery, flags, min_battery, pref_off, pref_ids, algo: int = ALGO_DECOMPOSED
                  ) -> Tuple[np.ndarray, np.ndarray]:
        """klm_reconcile: snapshot + batch in one call (== nodes_set then schedule_batch)."""
        b, f = _np(battery, np.float64), _np(flags, np.uint8)
        mb, off, ids = _np(min_battery, np.float64), _np(pref_off, np.int32), _np(pref_ids, np.int32)
        r = mb.size
        if b.shape != f.shape or b.ndim != 1 or off.size != r + 1:
            raise ValueError("reconcile: array lengths disagree")
        out_idx, out_score = np.empty(r, np.int32), np.empty(r, np.float64)
        if ids.size == 0:
            ids = np.zeros(1, np.int32)
        check(self._lib.klm_reconcile(self._h, b.size, _ptr(b), _ptr(f), r, _ptr(mb), _ptr(off), _ptr(ids),
                                      _ptr(out_idx), _ptr(out_score), algo))
        return out_idx, out_score

    # -- element-wise leftovers (types.go:151-156, :187-199; pod_metrics.go:174-177; node_metrics.go:142-164)
    def node_available(self, cpu_capacity, cpu_usage, memory_capacity, memory_usage, disk_capacity, disk_usage):
        a = [_np(x, np.int64) for x in (cpu_capacity, cpu_usage, memory_capacity, memory_usage, disk_capacity,
                                        disk_usage)]
        n = a[0].size
        out = [np.empty(n), np.empty(n), np.empty(n)]
        check(self._lib.klm_node_available(self._h, n, *[_ptr(x) for x in a], *[_ptr(o) for o in out]))
        return out

    def network_quality(self, connected, rtt_ms) -> np.ndarray:
        c, r = _np(connected, np.uint8), _np(rtt_ms, np.float64)
        q = np.empty(c.size, np.uint8)
        check(self._lib.klm_network_quality(self._h, c.size, _ptr(c), _ptr(r), _ptr(q)))
        return q

    def pod_restarts(self, st_off, restart_count) -> np.ndarray:
        off, cnt = _np(st_off, np.int32), _np(restart_count, np.int32)
        out = np.empty(off.size - 1, np.int32)
        if cnt.size == 0:
            cnt = np.zeros(1, np.int32)
        check(self._lib.klm_pod_restarts(self._h, out.size, _ptr(off), _ptr(cnt), _ptr(out)))
        return out

    def node_health(self, cond_off, cond_type, cond_status):
        off, t, s = _np(cond_off, np.int32), _np(cond_type, np.uint8), _np(cond_status, np.uint8)
        n = off.size - 1
        healthy, failing = np.empty(n, np.uint8), np.empty(n, np.uint8)
        if t.size == 0:
            t, s = np.zeros(1, np.uint8), np.zeros(1, np.uint8)
        check(self._lib.klm_node_health(self._h, n, _ptr(off), _ptr(t), _ptr(s), _ptr(healthy), _ptr(failing)))
        return healthy, failing

    # -- aggregation (manager.go:493-565)
    def cluster_aggregate(self, cpu_capacity, cpu_usage, memory_capacity, memory_usage, gpu_count,
                          healthy, gpu_usage, pod_running) -> dict:
        cols = [_np(cpu_capacity, np.int64), _np(cpu_usage, np.int64), _np(memory_capacity, np.int64),
                _np(memory_usage, np.int64), _np(gpu_count, np.int32), _np(healthy, np.uint8)]
        gu = _np(gpu_usage, np.float64)
        pr = _np(pod_running, np.uint8)
        n = cols[0].size
        nc = NodeCols(*[a.ctypes.data if a.size else None for a in cols],
                      gu.ctypes.data if gu.size else None, gu.size)
        out = ClusterOut()
        check(self._lib.klm_cluster_aggregate(self._h, C.byref(nc), n, _ptr(pr) if pr.size else None,
                                              pr.size, C.byref(out)))
        return out.as_dict()

    def node_rates(self, cpu_capacity, cpu_usage, memory_capacity, memory_usage, disk_capacity,
                   disk_allocatable=None) -> dict:
        a = [_np(x, np.int64) for x in (cpu_capacity, cpu_usage, memory_capacity, memory_usage, disk_capacity)]
        da = None if disk_allocatable is None else _np(disk_allocatable, np.int64)
        n = a[0].size
        out = {"cpu_rate": np.empty(n, np.float64), "memory_rate": np.empty(n, np.float64),
               "disk_usage": np.empty(n, np.int64), "disk_rate": np.empty(n, np.float64),
               "under_pressure": np.empty(n, np.uint8)}
        check(self._lib.klm_node_rates(self._h, n, *[_ptr(x) for x in a], _ptr(da),
                                       *[_ptr(v) for v in out.values()]))
        return out

    def pod_rates(self, ctr_off, cpu_usage, memory_usage, cpu_request, cpu_limit, memory_request,
                  memory_limit) -> dict:
        off = _np(ctr_off, np.int32)
        cols = [_np(x, np.int64) for x in (cpu_usage, memory_usage, cpu_request, cpu_limit,
                                           memory_request, memory_limit)]
        p = off.size - 1
        out = {"pod_totals": np.empty((6, p), np.int64), "cpu_rate": np.empty(p, np.float64),
               "memory_rate": np.empty(p, np.float64), "cpu_util": np.empty(p, np.float64),
               "memory_util": np.empty(p, np.float64), "over_limit": np.empty(p, np.uint8)}
        check(self._lib.klm_pod_rates(self._h, p, _ptr(off), *[_ptr(x) if x.size else None for x in cols],
                                      *[_ptr(v) for v in out.values()]))
        return out

    def fleet_scan(self, status_ok, remaining_percent, threshold: float) -> Tuple[int, np.ndarray]:
        ok, pct = _np(status_ok, np.uint8), _np(remaining_percent, np.float64)
        n = ok.size
        low = np.empty(max(n, 1), np.int32)
        healthy, n_low = _I64(0), _I64(0)
        check(self._lib.klm_fleet_scan(self._h, n, _ptr(ok), _ptr(pct), float(threshold),
                                       C.byref(healthy), _ptr(low), C.byref(n_low)))
        return healthy.value, low[: n_low.value].copy()

    # -- metric ring
    def ring_create(self, n_series: int, window: int) -> None:
        check(self._lib.klm_ring_create(self._h, n_series, window))
        self._ring_shape = (n_series, window)

    def ring_ingest(self, series_ids, values) -> None:
        ids, v = _np(series_ids, np.int64), _np(values, np.float32)
        check(self._lib.klm_ring_ingest(self._h, ids.size, _ptr(ids), _ptr(v)))

    def ring_ingest_tick(self, values) -> None:
        v = _np(values, np.float32)
        if v.size != self._ring_shape[0]:
            raise ValueError("one value per series expected")
        check(self._lib.klm_ring_ingest_tick(self._h, _ptr(v)))

    def ring_load(self, s0: int, data) -> None:
        d = _np(data, np.float32)
        if d.ndim != 2:
            raise ValueError("data must be [rows, count]")
        check(self._lib.klm_ring_load(self._h, s0, s0 + d.shape[0], d.shape[1], _ptr(d)))

    def collect(self, tick_values, snap: dict, out: Optional[dict] = None) -> dict:
        """One Manager.Collect cycle (klm_collect): tick ingest + window stats + node rates +
        cluster aggregate, one synchronisation. `snap` holds the node / pod columns
        (cluster_snapshot layout); `out` may carry preallocated (e.g. pinned) arrays."""
        cols = [_np(snap[k], dt) for k, dt in (("cpu_capacity", np.int64), ("cpu_usage", np.int64),
                                               ("memory_capacity", np.int64), ("memory_usage", np.int64),
                                               ("gpu_count", np.int32), ("healthy", np.uint8))]
        gu = _np(snap.get("gpu_usage", np.zeros(0)), np.float64)
        pr = _np(snap["pod_running"], np.uint8)
        dc = None if snap.get("disk_capacity") is None else _np(snap["disk_capacity"], np.int64)
        da = None if snap.get("disk_allocatable") is None else _np(snap["disk_allocatable"], np.int64)
        tick = None if tick_values is None else _np(tick_values, np.float32)
        n, s = cols[0].size, self._ring_shape[0]
        nc = NodeCols(*[a.ctypes.data if a.size else None for a in cols], gu.ctypes.data if gu.size else None,
                      gu.size)
        o = out if out is not None else {}
        o.setdefault("cpu_rate", np.empty(n, np.float64)); o.setdefault("memory_rate", np.empty(n, np.float64))
        o.setdefault("disk_usage", np.empty(n, np.int64)); o.setdefault("disk_rate", np.empty(n, np.float64))
        o.setdefault("under_pressure", np.empty(n, np.uint8))
        o.setdefault("min", np.empty(s, np.float32)); o.setdefault("max", np.empty(s, np.float32))
        o.setdefault("mean", np.empty(s, np.float64)); o.setdefault("p95", np.empty(s, np.float32))
        cl = ClusterOut()
        cin = CollectIn(_ptr(tick), C.pointer(nc), n, _ptr(dc), _ptr(da), _ptr(pr) if pr.size else None, pr.size)
        cout = CollectOut(C.pointer(cl), *[_ptr(o[k]) for k in ("cpu_rate", "memory_rate", "disk_usage",
                                                                  "disk_rate", "under_pressure", "min", "max",
                                                                  "mean", "p95")])
        check(self._lib.klm_collect(self._h, C.byref(cin), C.byref(cout)))
        o["cluster"] = cl.as_dict()
        return o

    # -- prepared calls: argument structs built once, then only the C call per cycle. This is
    #    what a compiled caller (the cgo shim) pays; the convenience wrappers above spend
    #    ~30 us per call on array coercion and struct building. Arrays are used in place, so
    #    they must already be contiguous, of the exact dtype, and stay alive and unmoved.
    @staticmethod
    def _exact(a, dtype, name):
        if not isinstance(a, np.ndarray) or a.dtype != dtype or not a.flags["C_CONTIGUOUS"]:
            raise ValueError(f"{name}: prepared calls need a C-contiguous numpy array of {np.dtype(dtype)}")
        return a.ctypes.data if a.size else None

    def collect_prepared(self, snap: dict, out: dict, packed: bool = False):
        """Returns call(tick) -> ClusterOut dict; fills `out` (keys as collect()) in place.
        packed=True: the four window statistics come back as one page-locked block
        (klm_collect_out.win_packed, a single download); out["min"/"max"/"p95"/"mean"] are
        then replaced by views into it."""
        ex = self._exact
        cols = [ex(snap[k], dt, k) for k, dt in (("cpu_capacity", np.int64), ("cpu_usage", np.int64),
                                                 ("memory_capacity", np.int64), ("memory_usage", np.int64),
                                                 ("gpu_count", np.int32), ("healthy", np.uint8))]
        gu = snap.get("gpu_usage")
        n_gu = 0 if gu is None else gu.size
        nc = NodeCols(*cols, ex(gu, np.float64, "gpu_usage") if n_gu else None, n_gu)
        n, p = snap["cpu_capacity"].size, snap["pod_running"].size
        cl = ClusterOut()
        cin = CollectIn(None, C.pointer(nc), n, ex(snap["disk_capacity"], np.int64, "disk_capacity"),
                        ex(snap["disk_allocatable"], np.int64, "disk_allocatable"),
                        ex(snap["pod_running"], np.uint8, "pod_running"), p)
        dts = (("cpu_rate", np.float64), ("memory_rate", np.float64), ("disk_usage", np.int64),
               ("disk_rate", np.float64), ("under_pressure", np.uint8), ("min", np.float32),
               ("max", np.float32), ("mean", np.float64), ("p95", np.float32))
        s = self._ring_shape[0]
        block = None
        if packed:
            nbytes = self._lib.klm_collect_packed_bytes(s)
            block = host_array(nbytes, np.uint8)
            c4 = (s * 4 + 255) & ~255
            out["min"] = block[0:s * 4].view(np.float32)
            out["max"] = block[c4:c4 + s * 4].view(np.float32)
            out["p95"] = block[2 * c4:2 * c4 + s * 4].view(np.float32)
            out["mean"] = block[3 * c4:3 * c4 + s * 8].view(np.float64)
            ptrs = [ex(out[k], dt, k) for k, dt in dts[:5]] + [None] * 4 + [block.ctypes.data]
        else:
            ptrs = [ex(out[k], dt, k) for k, dt in dts]
        cout = CollectOut(C.pointer(cl), *ptrs)
        fn, h, pin_, pout = self._lib.klm_collect, self._h, C.byref(cin), C.byref(cout)
        keep = (nc, cl, cin, cout, snap, out, block)

        def call(tick):
            if tick is not None:
                if tick.dtype != np.float32 or tick.size != s or not tick.flags["C_CONTIGUOUS"]:
                    raise ValueError("tick: need float32[n_series], C-contiguous")
                cin.tick_values = tick.ctypes.data
            else:
                cin.tick_values = None
            rc = fn(h, pin_, pout)
            if rc != KLM_OK:
                raise KlmError(rc, last_error())
            return cl
        call.keep = keep
        return call

    def collect_pipeline(self, snap: dict):
        """klm_collect_begin / klm_collect_end with two cycles in flight. Returns
        (begin(tick), end()): end() completes the oldest cycle begun and returns its result dict
        (keys as collect(): cluster, cpu_rate, ..., min, max, mean, p95 -- views into that cycle's
        page-locked block, valid until the cycle after next is begun). `snap` arrays are staged
        by begin, so they may change between cycles; they must be exact-dtype C-contiguous."""
        ex = self._exact
        s = self._ring_shape[0]
        n, p = snap["cpu_capacity"].size, snap["pod_running"].size
        gu = snap.get("gpu_usage")
        n_gu = 0 if gu is None else gu.size
        cols = [ex(snap[k], dt, k) for k, dt in (("cpu_capacity", np.int64), ("cpu_usage", np.int64),
                                                 ("memory_capacity", np.int64), ("memory_usage", np.int64),
                                                 ("gpu_count", np.int32), ("healthy", np.uint8))]
        nc = NodeCols(*cols, ex(gu, np.float64, "gpu_usage") if n_gu else None, n_gu)
        cin = CollectIn(None, C.pointer(nc), n, ex(snap["disk_capacity"], np.int64, "disk_capacity"),
                        ex(snap["disk_allocatable"], np.int64, "disk_allocatable"),
                        ex(snap["pod_running"], np.uint8, "pod_running"), p)
        nbytes = self._lib.klm_collect_packed_bytes(s)
        c4 = (s * 4 + 255) & ~255
        sets = []
        for _ in range(2):
            block = host_array(nbytes, np.uint8)
            o = {"cpu_rate": np.empty(n), "memory_rate": np.empty(n), "disk_usage": np.empty(n, np.int64),
                 "disk_rate": np.empty(n), "under_pressure": np.empty(n, np.uint8),
                 "min": block[0:s * 4].view(np.float32), "max": block[c4:c4 + s * 4].view(np.float32),
                 "p95": block[2 * c4:2 * c4 + s * 4].view(np.float32),
                 "mean": block[3 * c4:3 * c4 + s * 8].view(np.float64)}
            cl = ClusterOut()
            cout = CollectOut(C.pointer(cl), *[o[k].ctypes.data if o[k].size else None
                                               for k in ("cpu_rate", "memory_rate", "disk_usage", "disk_rate",
                                                         "under_pressure")],
                              None, None, None, None, block.ctypes.data)
            sets.append((o, cl, cout, block))
        b_fn, e_fn, h, pin_ = self._lib.klm_collect_begin, self._lib.klm_collect_end, self._h, C.byref(cin)
        seq = {"begun": 0, "ended": 0}

        def begin(tick):
            if tick is not None and (tick.dtype != np.float32 or tick.size != s or not tick.flags["C_CONTIGUOUS"]):
                raise ValueError("tick: need float32[n_series], C-contiguous")
            cin.tick_values = None if tick is None else tick.ctypes.data
            rc = b_fn(h, pin_, C.byref(sets[seq["begun"] & 1][2]))
            if rc != KLM_OK:
                raise KlmError(rc, last_error())
            seq["begun"] += 1

        def end():
            rc = e_fn(h)
            if rc != KLM_OK:
                raise KlmError(rc, last_error())
            o, cl, _, _ = sets[seq["ended"] & 1]
            seq["ended"] += 1
            o["cluster"] = cl.as_dict()
            return o
        begin.keep = (nc, cin, sets, snap)
        return begin, end

    def reconcile_prepared(self, battery, flags, min_battery, pref_off, pref_ids, out_idx, out_score,
                           algo: int = ALGO_DECOMPOSED):
        """Returns call(): klm_reconcile on the given arrays, results in out_idx / out_score
        (one reconcile: snapshot, then every pending request)."""
        ex = self._exact
        n, r = battery.size, min_battery.size
        if flags.size != n or pref_off.size != r + 1 or out_idx.size != r or out_score.size != r:
            raise ValueError("reconcile_prepared: array lengths disagree")
        a = [ex(battery, np.float64, "battery"), ex(flags, np.uint8, "flags"),
             ex(min_battery, np.float64, "min_battery"), ex(pref_off, np.int32, "pref_off"),
             ex(pref_ids, np.int32, "pref_ids"), ex(out_idx, np.int32, "out_idx"),
             ex(out_score, np.float64, "out_score")]
        fn, h = self._lib.klm_reconcile, self._h

        def call():
            rc = fn(h, n, a[0], a[1], r, a[2], a[3], a[4], a[5], a[6], algo)
            if rc != KLM_OK:
                raise KlmError(rc, last_error())
        call.keep = (battery, flags, min_battery, pref_off, pref_ids, out_idx, out_score)
        return call

    def ring_stats(self, s0: int = 0, s1: Optional[int] = None, p95: bool = True) -> dict:
        if s1 is None:
            s1 = self._ring_shape[0]
        n = s1 - s0
        out = {"min": np.empty(n, np.float32), "max": np.empty(n, np.float32),
               "mean": np.empty(n, np.float64), "p95": np.empty(n, np.float32) if p95 else None}
        check(self._lib.klm_ring_stats(self._h, s0, s1, _ptr(out["min"]), _ptr(out["max"]),
                                       _ptr(out["mean"]), _ptr(out["p95"])))
        return out


# --------------------------------------------------------------------------
# Device-pointer layer for callers that keep their data in HBM (torch tensors).
# torch is plumbing here: it owns the memory and the stream, nothing else.
# --------------------------------------------------------------------------

def _dp(t) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _stream_ptr(stream=None) -> int:
    import torch
    s = stream if stream is not None else torch.cuda.current_stream()
    return s.cuda_stream


def _need(t, dtype, name):
    if t is None:
        return
    if not t.is_cuda or t.dtype != dtype or not t.is_contiguous():
        raise ValueError(f"{name}: need a contiguous CUDA tensor of {dtype}, got {t.dtype} on {t.device}")


class _Recorder:
    """Stands in for the library while a Dev call is being recorded by Dev.prebind()."""

    def __init__(self, lib):
        self._lib = lib
        self.calls = []

    def __getattr__(self, name):
        fn = getattr(self._lib, name)

        def rec(*args):
            self.calls.append((fn, args))
            return KLM_OK
        return rec


class Dev:
    """klm_dev_* calls on torch CUDA tensors; only enqueues work on the current stream."""

    def __init__(self, device=None):
        import torch
        self.torch = torch
        self.lib = load()
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        with torch.cuda.device(self.device):
            # scratch areas carry tickets that must start at zero
            self.best_scratch = torch.zeros(self.lib.klm_dev_node_best_scratch_bytes(), dtype=torch.uint8,
                                            device=self.device)
            self.cluster_scratch = torch.zeros(self.lib.klm_dev_cluster_scratch_bytes(), dtype=torch.uint8,
                                               device=self.device)

    def prebind(self, method: str, *args, **kwargs):
        """Validate the arguments of one Dev call now and return a zero-argument callable
        that just re-issues the C call with the same pointers (hot loops: a launch is then
        one ctypes call, ~1 us of host time). The tensors must stay alive and in place."""
        rec = _Recorder(self.lib)
        real, self.lib = self.lib, rec
        try:
            getattr(self, method)(*args, **kwargs)
        finally:
            self.lib = real
        calls = [(fn, tuple(a)) for fn, a in rec.calls if fn.restype is C.c_int]
        if not calls:
            raise ValueError(f"{method} issued no launch")

        def issue():
            for fn, a in calls:
                rc = fn(*a)
                if rc != KLM_OK:
                    raise KlmError(rc, last_error())
        return issue

    def node_best(self, battery, flags, best, node_base: int = 0, stream=None, scratch=None):
        t = self.torch
        _need(battery, t.float64, "battery"); _need(flags, t.uint8, "flags"); _need(best, t.int64, "best")
        sc = self.best_scratch if scratch is None else scratch
        check(self.lib.klm_dev_node_best(_dp(battery), _dp(flags), battery.numel(), node_base, _dp(best),
                                         _dp(sc), _stream_ptr(stream)))

    def schedule(self, battery, flags, best, min_battery, pref_off, pref_ids, out_idx, out_score,
                 node_base: int = 0, stream=None, after_best: bool = False):
        """after_best=True: klm_dev_schedule_after_best (overlaps the node_best launched just
        before it on the same stream; see the contract in klm.h)."""
        t = self.torch
        _need(min_battery, t.float64, "min_battery"); _need(pref_off, t.int32, "pref_off")
        _need(pref_ids, t.int32, "pref_ids"); _need(out_idx, t.int32, "out_idx")
        _need(out_score, t.float64, "out_score")
        fn = self.lib.klm_dev_schedule_after_best if after_best else self.lib.klm_dev_schedule
        check(fn(_dp(battery), _dp(flags), battery.numel(), node_base, _dp(best),
                                        min_battery.numel(), _dp(min_battery), _dp(pref_off), _dp(pref_ids),
                                        _dp(out_idx), _dp(out_score), _stream_ptr(stream)))

    def schedule_fused(self, battery, flags, min_battery, pref_off, pref_ids, out_idx, out_score,
                       best=None, node_base: int = 0, stream=None, scratch=None):
        """klm_dev_schedule_fused: K1 and K2 in one cooperative launch (best: optional int64[2])."""
        t = self.torch
        _need(battery, t.float64, "battery"); _need(flags, t.uint8, "flags")
        _need(min_battery, t.float64, "min_battery"); _need(pref_off, t.int32, "pref_off")
        _need(pref_ids, t.int32, "pref_ids"); _need(out_idx, t.int32, "out_idx")
        _need(out_score, t.float64, "out_score")
        if best is not None:
            _need(best, t.int64, "best")
        sc = self.best_scratch if scratch is None else scratch
        check(self.lib.klm_dev_schedule_fused(_dp(battery), _dp(flags), battery.numel(), node_base,
                                              _dp(best) if best is not None else None, min_battery.numel(),
                                              _dp(min_battery), _dp(pref_off), _dp(pref_ids), _dp(out_idx),
                                              _dp(out_score), _dp(sc), _stream_ptr(stream)))

    def schedule_bruteforce(self, battery, flags, min_battery, pref_off, pref_ids, out_idx, out_score,
                            node_base: int = 0, stream=None):
        check(self.lib.klm_dev_schedule_bruteforce(_dp(battery), _dp(flags), battery.numel(), node_base,
                                                   min_battery.numel(), _dp(min_battery), _dp(pref_off),
                                                   _dp(pref_ids), _dp(out_idx), _dp(out_score),
                                                   _stream_ptr(stream)))

    def merge_partials(self, part_idx, part_score, out_idx, out_score, stream=None):
        t = self.torch
        _need(part_idx, t.int32, "part_idx"); _need(part_score, t.float64, "part_score")
        n_parts, r = part_idx.shape
        check(self.lib.klm_dev_merge_partials(n_parts, r, _dp(part_idx), _dp(part_score), _dp(out_idx),
                                              _dp(out_score), _stream_ptr(stream)))

    # -- node-range sharding without an all-gather (klm.h, "node-range sharding")
    def shard_slots(self, device=None):
        """A zero-filled exchange-slot array (uint8 tensor) for one rank."""
        return self.torch.zeros(int(self.lib.klm_dev_shard_slots_bytes()), dtype=self.torch.uint8,
                                device="cuda" if device is None else device)

    def node_best_publish(self, battery, flags, node_base: int, n_parts: int, my_part: int, slot_ptrs, epoch,
                          best=None, scratch=None, stream=None):
        """slot_ptrs: device addresses (ints) of every rank's slot array, as seen from this device.
        epoch: int, or a ctypes.c_uint64 whose value a prebound call re-reads at every issue."""
        t = self.torch
        _need(battery, t.float64, "battery"); _need(flags, t.uint8, "flags")
        arr = (C.c_void_p * n_parts)(*[int(x) for x in slot_ptrs])
        sc = self.best_scratch if scratch is None else scratch
        check(self.lib.klm_dev_node_best_publish(_dp(battery), _dp(flags), battery.numel(), node_base, n_parts,
                                                 my_part, arr, epoch, _dp(best) if best is not None else None,
                                                 _dp(sc), _stream_ptr(stream)))

    def node_best_spread(self, battery, flags, node_base: int, n_parts: int, my_part: int, slot_ptrs, rep_bat_ptrs,
                         rep_flg_ptrs, epoch, best=None, scratch=None, stream=None):
        """node_best_publish that also pushes the shard into every peer's replica (rep_*_ptrs: device
        addresses of the replica BASES of all ranks, as seen from this device)."""
        t = self.torch
        _need(battery, t.float64, "battery"); _need(flags, t.uint8, "flags")
        arr = (C.c_void_p * n_parts)(*[int(x) for x in slot_ptrs])
        ab = (C.c_void_p * n_parts)(*[int(x) for x in rep_bat_ptrs])
        af = (C.c_void_p * n_parts)(*[int(x) for x in rep_flg_ptrs])
        sc = self.best_scratch if scratch is None else scratch
        check(self.lib.klm_dev_node_best_spread(_dp(battery), _dp(flags), battery.numel(), node_base, n_parts,
                                                my_part, arr, ab, af, epoch, _dp(best) if best is not None else None,
                                                _dp(sc), _stream_ptr(stream)))

    def schedule_sharded(self, n_total: int, bat_ptrs, flg_ptrs, my_slots, epoch, min_battery, pref_off, pref_ids,
                         out_idx, out_score, stream=None, table_from_k1: bool = False):
        """bat_ptrs / flg_ptrs: device addresses (ints) of every rank's shard columns.
        table_from_k1: the columns are filled by this epoch's node_best_spread calls."""
        t = self.torch
        _need(min_battery, t.float64, "min_battery"); _need(pref_off, t.int32, "pref_off")
        _need(pref_ids, t.int32, "pref_ids"); _need(out_idx, t.int32, "out_idx")
        _need(out_score, t.float64, "out_score")
        n = len(bat_ptrs)
        ab = (C.c_void_p * n)(*[int(x) for x in bat_ptrs])
        af = (C.c_void_p * n)(*[int(x) for x in flg_ptrs])
        check(self.lib.klm_dev_schedule_sharded(n, n_total, ab, af, _dp(my_slots), epoch, min_battery.numel(),
                                                _dp(min_battery), _dp(pref_off), _dp(pref_ids), _dp(out_idx),
                                                _dp(out_score), int(table_from_k1), _stream_ptr(stream)))

    def node_features(self, battery, flags, heartbeat_age, win_min, win_max, win_p95, params: dict, battery_out,
                      flags_out, stream=None):
        """klm_dev_node_features: window statistics + heartbeat age -> the node table the scheduler sees."""
        prm = FeatureParams(**params)
        check(self.lib.klm_dev_node_features(battery.numel(), _dp(battery), _dp(flags), _dp(heartbeat_age), _dp(win_min),
                                             _dp(win_max), _dp(win_p95), C.byref(prm), _dp(battery_out),
                                             _dp(flags_out), _stream_ptr(stream)))

    def shard_rendezvous(self, n_parts: int, my_part: int, slot_ptrs, my_slots, epoch, stream=None):
        """Device-side rendezvous of the ranks (slot arrays of its own; epoch as in node_best_publish)."""
        arr = (C.c_void_p * n_parts)(*[int(x) for x in slot_ptrs])
        check(self.lib.klm_dev_shard_rendezvous(n_parts, my_part, arr, _dp(my_slots), epoch, _stream_ptr(stream)))

    def shard_error(self, my_slots, stream=None) -> int:
        out = C.c_int(0)
        check(self.lib.klm_dev_shard_error(_dp(my_slots), C.byref(out), _stream_ptr(stream)))
        return out.value

    def merge_partials_peer(self, idx_ptrs, score_ptrs, r0: int, r1: int, out_idx, out_score, stream=None):
        """idx_ptrs / score_ptrs: device addresses (ints) of every rank's partial arrays."""
        n = len(idx_ptrs)
        ai = (C.c_void_p * n)(*idx_ptrs)
        asc = (C.c_void_p * n)(*score_ptrs)
        check(self.lib.klm_dev_merge_partials_peer(n, ai, asc, r0, r1, _dp(out_idx), _dp(out_score),
                                                   _stream_ptr(stream)))

    def _cols(self, cols: dict) -> NodeCols:
        gu = cols.get("gpu_usage")
        return NodeCols(_dp(cols["cpu_capacity"]), _dp(cols["cpu_usage"]), _dp(cols["memory_capacity"]),
                        _dp(cols["memory_usage"]), _dp(cols["gpu_count"]), _dp(cols["healthy"]),
                        _dp(gu) if gu is not None and gu.numel() else None,
                        0 if gu is None else gu.numel())

    def cluster_aggregate(self, cols: dict, pod_running, out, stream=None):
        """out: uint8 tensor of 104 bytes (klm_cluster_out)."""
        nc = self._cols(cols)
        check(self.lib.klm_dev_cluster_aggregate(C.byref(nc), cols["cpu_capacity"].numel(), _dp(pod_running),
                                                 pod_running.numel(), _dp(out), _dp(self.cluster_scratch),
                                                 _stream_ptr(stream)))

    def cluster_sums(self, cols: dict, pod_running, sums, stream=None):
        nc = self._cols(cols)
        check(self.lib.klm_dev_cluster_sums(C.byref(nc), cols["cpu_capacity"].numel(), _dp(pod_running),
                                            pod_running.numel(), _dp(sums), _dp(self.cluster_scratch),
                                            _stream_ptr(stream)))

    def cluster_finish(self, sums, out, stream=None):
        check(self.lib.klm_dev_cluster_finish(_dp(sums), _dp(out), _stream_ptr(stream)))

    def node_rates(self, cols: dict, disk_capacity, disk_allocatable, out: dict, stream=None):
        check(self.lib.klm_dev_node_rates(cols["cpu_capacity"].numel(), _dp(cols["cpu_capacity"]),
                                          _dp(cols["cpu_usage"]), _dp(cols["memory_capacity"]),
                                          _dp(cols["memory_usage"]), _dp(disk_capacity), _dp(disk_allocatable),
                                          _dp(out.get("cpu_rate")), _dp(out.get("memory_rate")),
                                          _dp(out.get("disk_usage")), _dp(out.get("disk_rate")),
                                          _dp(out.get("under_pressure")), _stream_ptr(stream)))

    def ring_ingest(self, ring, seen, window: int, values, series_ids=None, stream=None):
        t = self.torch
        _need(ring, t.float32, "ring"); _need(seen, t.int32, "seen"); _need(values, t.float32, "values")
        n_series, pitch = ring.shape
        check(self.lib.klm_dev_ring_ingest(_dp(ring), _dp(seen), n_series, window, pitch, values.numel(),
                                           _dp(series_ids), _dp(values), _stream_ptr(stream)))

    def window_stats(self, ring, seen, window: int, out_min, out_max, out_mean, out_p95,
                     s0: int = 0, s1: Optional[int] = None, stream=None, block_variant: bool = False,
                     variant: Optional[str] = None, state=None):
        """state: float32 [n_series, 2] tensor (NaN-filled at first) -> klm_dev_window_stats_hinted."""
        t = self.torch
        _need(ring, t.float32, "ring"); _need(seen, t.int32, "seen")
        _need(out_min, t.float32, "out_min"); _need(out_max, t.float32, "out_max")
        _need(out_mean, t.float64, "out_mean"); _need(out_p95, t.float32, "out_p95")
        n_series, pitch = ring.shape
        if s1 is None:
            s1 = n_series
        if block_variant:
            variant = "block"
        fn = {None: self.lib.klm_dev_window_stats,
              "warp": self.lib.klm_dev_window_stats_warp, "stream": self.lib.klm_dev_window_stats_stream,
              "block": self.lib.klm_dev_window_stats_block}[variant]
        if state is not None:
            _need(state, t.float32, "state")
            check(self.lib.klm_dev_window_stats_hinted(_dp(ring), _dp(seen), window, pitch, s0, s1, _dp(out_min),
                                                       _dp(out_max), _dp(out_mean), _dp(out_p95), _dp(state),
                                                       _stream_ptr(stream)))
            return
        check(fn(_dp(ring), _dp(seen), window, pitch, s0, s1, _dp(out_min), _dp(out_max), _dp(out_mean),
                 _dp(out_p95), _stream_ptr(stream)))
