"""Where does the e2e step time go (host wall clock per klm_* call, C2 sizes)."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge
pkg = ge.load_package(); syn = pkg.synth
pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()
S, T, N, P = 26000, 3600, 1000, 10000
h = pkg.Handle(0)
h.ring_create(S, T)
rows = (np.random.default_rng(1).random((4096, T - 1)) * 100).astype(np.float32)
for a in range(0, S, 4096):
    h.ring_load(a, rows[: min(4096, S - a)])
snap = {k: pin(v) for k, v in syn.cluster_snapshot(2, N, P).items()}
snap["gpu_usage"] = np.zeros(0)
bat, flags = map(pin, syn.node_table(2, N))
mb, off, ids = syn.requests(2, 1, N)
ids = ids if ids.size else np.zeros(1, np.int32)
mb, off, ids = pin(mb), pin(off), pin(ids)
ticks = pin((np.random.default_rng(2).random((64, S)) * 100).astype(np.float32))
outs = {"cpu_rate": pin(np.empty(N)), "memory_rate": pin(np.empty(N)), "disk_usage": pin(np.empty(N, np.int64)),
        "disk_rate": pin(np.empty(N)), "under_pressure": pin(np.empty(N, np.uint8)),
        "min": pin(np.empty(S, np.float32)), "max": pin(np.empty(S, np.float32)), "mean": pin(np.empty(S)),
        "p95": pin(np.empty(S, np.float32))}
empty = {k: v[:0] for k, v in snap.items()}

def t(name, fn, n=200):
    for _ in range(10): fn()
    t0 = time.perf_counter()
    for _ in range(n): fn()
    print(f"{name:44s} {(time.perf_counter() - t0) / n * 1e6:8.1f} us")

i = [0]
def collect_full():
    h.collect(ticks[i[0] % 64], snap, outs); i[0] += 1
t("collect (tick + stats + snapshot)", collect_full)
t("collect (no tick; stats + snapshot)", lambda: h.collect(None, snap, outs))
t("nodes_set (1000 nodes)", lambda: h.nodes_set(bat, flags))
t("schedule_batch (1 request)", lambda: h.schedule_batch(mb, off, ids))
def sched():
    h.nodes_set(bat, flags); h.schedule_batch(mb, off, ids)
t("nodes_set + schedule_batch", sched)
t("ring_stats only (K7 + 4 D2H)", lambda: h.ring_stats())
import ctypes
t("python ctypes floor (klm_abi_version)", lambda: h._lib.klm_abi_version(), 2000)

# raw C call with the argument structs built once (what a cgo caller pays); one input struct
# per tick vector, because re-ingesting one vector fills the windows with ties
b = pkg.binding
C = ctypes
cols = [snap[k] for k in ("cpu_capacity", "cpu_usage", "memory_capacity", "memory_usage", "gpu_count", "healthy")]
nc = b.NodeCols(*[a.ctypes.data for a in cols], None, 0)
cl = b.ClusterOut()
fresh = pin((np.random.default_rng(3).random((3 * 210, S)) * 100).astype(np.float32))
fn, hh = h._lib.klm_collect, h._h


def raw(name, full, keys, base):
    cins = [b.CollectIn(fresh[base + k].ctypes.data, C.pointer(nc) if full else None, N if full else 0,
                        snap["disk_capacity"].ctypes.data if full else None,
                        snap["disk_allocatable"].ctypes.data if full else None,
                        snap["pod_running"].ctypes.data if full else None, P if full else 0) for k in range(210)]
    small = [outs[k].ctypes.data if full else None for k in ("cpu_rate", "memory_rate", "disk_usage", "disk_rate",
                                                               "under_pressure")]
    win = [outs[k].ctypes.data if k in keys else None for k in ("min", "max", "mean", "p95")]
    cout = b.CollectOut(C.pointer(cl), *small, *win)
    po, k = C.byref(cout), [0]

    def call():
        rc = fn(hh, C.byref(cins[k[0]]), po)
        k[0] += 1
        assert rc == 0
    t(name, call)


raw("raw klm_collect (prebuilt structs)", True, ("min", "max", "mean", "p95"), 0)
raw("raw klm_collect, ring branch only", False, ("min", "max", "mean", "p95"), 210)
raw("raw klm_collect, ring branch, p95 only out", False, ("p95",), 420)
