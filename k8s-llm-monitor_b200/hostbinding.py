"""ctypes view of libklm_host.so, the C++ host layer (host/klm_host.h) that mirrors the
reference's scheduler Controller and metrics Manager around the GPU path. Used by tests."""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional

from . import binding

LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libklm_host.so")
_lib = None
_CAP = 512


def load() -> C.CDLL:
    global _lib
    if _lib is None:
        binding.load()  # libklm.so first (the host library links against it)
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} not found: run __graft_entry__.build()")
        lib = C.CDLL(LIB_PATH)
        lib.klmh_open.restype = C.c_void_p
        lib.klmh_open.argtypes = [C.c_void_p]
        lib.klmh_error.restype = C.c_char_p
        lib.klmh_error.argtypes = [C.c_void_p]
        lib.klmh_healthy_uav_count.restype = C.c_int64
        lib.klmh_update_uav_report.restype = C.c_int64
        for name in ("klmh_close", "klmh_clear", "klmh_add_uav", "klmh_add_request", "klmh_reconcile", "klmh_reconcile_capacity", "klmh_to_lower", "klmh_write_statuses", "klmh_add_status",
                     "klmh_status", "klmh_add_node", "klmh_add_pod", "klmh_calculate_cluster_metrics",
                     "klmh_issue", "klmh_add_fleet_uav", "klmh_healthy_uav_count", "klmh_low_battery_uavs",
                     "klmh_name", "klmh_load_uavs_json", "klmh_load_requests_json", "klmh_request_info",
                     "klmh_uav_info", "klmh_status_json", "klmh_format_float", "klmh_report_decode",
                     "klmh_enable_uav_ring", "klmh_update_uav_report", "klmh_uav_entry"):
            getattr(lib, name)
        _lib = lib
    return _lib


def _b(s: Optional[str]):
    return None if s is None else s.encode("utf-8")


def format_json_float(v: float) -> str:
    """A float64 as Go's encoding/json writes it (host/klm_wire.h FormatJSONFloat)."""
    buf = C.create_string_buffer(64)
    load().klmh_format_float(C.c_double(v), buf, 64)
    return buf.value.decode()


def to_lower(text: str) -> str:
    """strings.ToLower as the host layer applies it (host/klm_host.h ToLower)."""
    raw = text.encode("utf-8", "surrogateescape") if isinstance(text, str) else text
    buf = C.create_string_buffer(4 * len(raw) + 8)
    load().klmh_to_lower(raw, buf, len(buf))
    return buf.value.decode("utf-8", "replace")


UPDATE_CB = C.CFUNCTYPE(C.c_int, C.c_char_p, C.c_char_p, C.c_char_p)


class HostSession:
    def __init__(self, handle: Optional["binding.Handle"]):
        """handle=None gives a decode / encode only session (usable without a GPU)."""
        self.lib = load()
        self.s = C.c_void_p(self.lib.klmh_open(handle._h if handle is not None else None))
        self._keep = handle

    def close(self):
        if self.s:
            self.lib.klmh_close(self.s)
            self.s = None

    def clear(self):
        self.lib.klmh_clear(self.s)

    def _err(self):
        raise RuntimeError(self.lib.klmh_error(self.s).decode("utf-8", "replace"))

    # --- Controller
    def add_uav(self, node_name: Optional[str], uav_id: str, battery: Optional[float], status: Optional[str]):
        self.lib.klmh_add_uav(self.s, _b(node_name), _b(uav_id), C.c_int(0 if battery is None else 1),
                              C.c_double(0.0 if battery is None else float(battery)), _b(status))

    def add_request(self, phase: Optional[str], wl_name: str, wl_ns: str, min_battery: float, preferred: List[str]):
        arr = (C.c_char_p * max(len(preferred), 1))(*[p.encode("utf-8") for p in preferred])
        self.lib.klmh_add_request(self.s, _b(phase), _b(wl_name), _b(wl_ns), C.c_double(min_battery),
                                  C.c_int(len(preferred)), arr)

    # --- wire format (host/klm_wire.h)
    def load_uavs_json(self, text: str) -> int:
        n = self.lib.klmh_load_uavs_json(self.s, text.encode("utf-8"))
        if n < 0:
            self._err()
        return n

    def load_requests_json(self, text: str, apimachinery_ints: bool = True) -> int:
        n = self.lib.klmh_load_requests_json(self.s, text.encode("utf-8"), C.c_int(int(apimachinery_ints)))
        if n < 0:
            self._err()
        return n

    def request_info(self, i: int) -> dict:
        mb, npref = C.c_double(0), C.c_int(0)
        name, phase = C.create_string_buffer(_CAP), C.create_string_buffer(_CAP)
        hp = self.lib.klmh_request_info(self.s, i, C.byref(mb), C.byref(npref), name, phase, _CAP)
        return {"min_battery": mb.value, "n_pref": npref.value, "workload": name.value.decode(),
                "has_phase": hp == 1, "phase": phase.value.decode()}

    def uav_info(self, i: int) -> dict:
        bat, hb = C.c_double(0), C.c_int(0)
        name, status = C.create_string_buffer(_CAP), C.create_string_buffer(_CAP)
        f = self.lib.klmh_uav_info(self.s, i, C.byref(bat), C.byref(hb), name, status, _CAP)
        return {"battery": bat.value, "has_battery": bool(hb.value), "name": name.value.decode(),
                "has_name": bool(f & 1), "status": status.value.decode(), "has_status": bool(f & 2)}

    def status_json(self, i: int, last_updated: str) -> str:
        buf = C.create_string_buffer(2048)
        self.lib.klmh_status_json(self.s, i, last_updated.encode(), buf, 2048)
        return buf.value.decode("utf-8")

    # --- UAV push ingestion (manager.go:391-449; host/klm_wire.h UAVReport)
    def report_decode(self, text: str) -> dict:
        import numpy as np
        ch = np.zeros(64, np.float32)
        bufs = [C.create_string_buffer(_CAP) for _ in range(3)]
        pct = C.c_double(0)
        rc = self.lib.klmh_report_decode(self.s, text.encode("utf-8"), ch.ctypes.data_as(C.c_void_p), bufs[0], bufs[1],
                                         bufs[2], C.byref(pct), _CAP)
        if rc < 0:
            self._err()
        return {"has_state": rc == 1, "channels": ch, "node_name": bufs[0].value.decode(),
                "status": bufs[1].value.decode(), "source": bufs[2].value.decode(), "remaining_percent": pct.value}

    def enable_uav_ring(self, max_nodes: int, window: int):
        if self.lib.klmh_enable_uav_ring(self.s, C.c_int64(max_nodes), C.c_int(window)) < 0:
            self._err()

    def update_uav_report(self, text: str, now: str = "2026-01-01T00:00:00Z") -> int:
        rc = self.lib.klmh_update_uav_report(self.s, text.encode("utf-8"), now.encode())
        if rc == -2:
            self._err()
        return rc

    def uav_entry(self, i: int) -> dict:
        pct = C.c_double(0)
        bufs = [C.create_string_buffer(_CAP) for _ in range(3)]
        rc = self.lib.klmh_uav_entry(self.s, C.c_int64(i), C.byref(pct), bufs[0], bufs[1], bufs[2], _CAP)
        return {"has_state": rc == 1, "remaining_percent": pct.value, "node_name": bufs[0].value.decode(),
                "status": bufs[1].value.decode(), "last_heartbeat": bufs[2].value.decode()}

    def add_status(self, updated: bool, phase: str, node: str, uav: str, score: float, message: str):
        self.lib.klmh_add_status(self.s, C.c_int(int(updated)), _b(phase), _b(node), _b(uav), C.c_double(score),
                                 _b(message))

    def write_statuses(self, now: str, workers: int, callback) -> int:
        """WriteStatuses over the session's statuses: callback(ns, name, status_json) -> bool (True = ok),
        called from `workers` threads. Returns the number of failed updates."""
        cb = UPDATE_CB(lambda ns, name, js: 0 if callback(ns.decode(), name.decode(), js.decode("utf-8")) else 1)
        n = self.lib.klmh_write_statuses(self.s, now.encode(), C.c_int(workers), cb)
        if n < 0:
            self._err()
        return n

    def reconcile_capacity(self, capacity):
        """Controller::ReconcileWithCapacity: requests in order, node i assignable capacity[i] times."""
        cap = (C.c_int32 * max(len(capacity), 1))(*[int(c) for c in capacity])
        n = self.lib.klmh_reconcile_capacity(self.s, cap, C.c_int(len(capacity)))
        return self._statuses(n)

    def reconcile(self):
        n = self.lib.klmh_reconcile(self.s)
        return self._statuses(n)

    def _statuses(self, n):
        if n < 0:
            self._err()
        out = []
        for i in range(n):
            bufs = [C.create_string_buffer(_CAP) for _ in range(4)]
            score = C.c_double(0)
            upd = self.lib.klmh_status(self.s, i, bufs[0], bufs[1], bufs[2], C.byref(score), bufs[3], _CAP)
            out.append({"updated": bool(upd), "phase": bufs[0].value.decode(), "node": bufs[1].value.decode(),
                        "uav": bufs[2].value.decode(), "score": score.value,
                        "message": bufs[3].value.decode("utf-8")})
        return out

    # --- Manager
    def add_node(self, cpu_cap, cpu_use, mem_cap, mem_use, healthy, gpu_usage=()):
        g = (C.c_double * max(len(gpu_usage), 1))(*gpu_usage)
        self.lib.klmh_add_node(self.s, C.c_int64(cpu_cap), C.c_int64(cpu_use), C.c_int64(mem_cap),
                               C.c_int64(mem_use), C.c_int(int(healthy)), C.c_int(len(gpu_usage)),
                               C.c_int(len(gpu_usage)), g)

    def add_pod(self, phase: str):
        self.lib.klmh_add_pod(self.s, _b(phase))

    def calculate_cluster_metrics(self):
        out = binding.ClusterOut()
        health = C.create_string_buffer(_CAP)
        n = self.lib.klmh_calculate_cluster_metrics(self.s, C.byref(out), health, _CAP)
        if n < 0:
            self._err()
        issues = []
        for i in range(n):
            b = C.create_string_buffer(_CAP)
            self.lib.klmh_issue(self.s, i, b, _CAP)
            issues.append(b.value.decode())
        d = out.as_dict()
        d.pop("health_status"); d.pop("issues")
        d["HealthStatus"] = health.value.decode()
        d["Issues"] = issues
        return d

    def add_fleet_uav(self, node_name: str, system_status: str, pct: float):
        self.lib.klmh_add_fleet_uav(self.s, _b(node_name), _b(system_status), C.c_double(pct))

    def healthy_uav_count(self) -> int:
        n = self.lib.klmh_healthy_uav_count(self.s)
        if n < 0:
            self._err()
        return n

    def low_battery_uavs(self, threshold: float) -> List[str]:
        n = self.lib.klmh_low_battery_uavs(self.s, C.c_double(threshold))
        if n < 0:
            self._err()
        out = []
        for i in range(n):
            b = C.create_string_buffer(_CAP)
            self.lib.klmh_name(self.s, i, b, _CAP)
            out.append(b.value.decode())
        return out
