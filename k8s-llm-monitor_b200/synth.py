"""Seeded synthetic cluster snapshots for the five BASELINE.json configs.

Counter-based splitmix64 (stream = seed, element = index), so any slice of any
array can be produced independently and identically by numpy on the host and by
torch on the GPU. The same bytes are fed to the CPU oracle and to the CUDA path.
Distributions follow SURVEY.md §8(d); the waveforms of the UAV channels follow
the reference simulator's formulas (pkg/uav/mavlink_simulator.go:265-322) with
math/rand replaced by the seeded generator.
"""
from __future__ import annotations

import numpy as np

SEED_BASE = 0xB200
_GOLDEN = 0x9E3779B97F4A7C15
_M1 = 0xBF58476D1CE4E5B9
_M2 = 0x94D049BB133111EB


def seed_for(config_id: int, stream: int = 0) -> int:
    return (SEED_BASE + config_id) * 1000003 + stream


# ------------------------------------------------------------------ numpy

def splitmix64(seed: int, n: int, start: int = 0) -> np.ndarray:
    """n 64-bit words: element i is mix(seed + (start + i + 1) * golden)."""
    with np.errstate(over="ignore"):
        i = np.arange(start + 1, start + n + 1, dtype=np.uint64)
        z = np.uint64(seed & 0xFFFFFFFFFFFFFFFF) + i * np.uint64(_GOLDEN)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(_M1)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(_M2)
        return z ^ (z >> np.uint64(31))


def uniform(seed: int, n: int, start: int = 0) -> np.ndarray:
    """float64 in [0, 1): top 53 bits."""
    return (splitmix64(seed, n, start) >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


def node_table(config_id: int, n: int):
    """battery f64[n] on a 0.1 % grid (many exact ties), flags u8[n]:
    2 % empty name, 5 % non-active status, 1 % battery missing (= 0)."""
    u = uniform(seed_for(config_id, 1), n)
    battery = np.round(u * 100.0 * 10.0) / 10.0
    battery = np.minimum(battery, 99.9)
    r = uniform(seed_for(config_id, 2), n)
    flags = np.full(n, 3, np.uint8)
    flags[r < 0.02] &= np.uint8(~1 & 0xFF)          # empty node name
    flags[(r >= 0.02) & (r < 0.07)] &= np.uint8(~2 & 0xFF)  # status not active
    battery[(r >= 0.07) & (r < 0.08)] = 0.0          # battery field missing -> readFloat gives 0
    return battery, flags


_MIN_CHOICES = np.array([20.0, 50.0, 60.0, 80.0, 95.0])


def requests(config_id: int, r: int, n_nodes: int, k_fixed: int | None = None, start: int = 0):
    """min_battery f64[r] (10 % zero, else uniform over {20,50,60,80,95});
    K preferred ids per request: k_fixed, or {0: 30 %, 1: 40 %, 2: 30 %};
    ids uniform over the nodes, ineligible ones included."""
    u = uniform(seed_for(config_id, 3), r, start)
    v = uniform(seed_for(config_id, 4), r, start)
    min_battery = np.where(u < 0.10, 0.0, _MIN_CHOICES[np.minimum((v * 5).astype(np.int64), 4)])
    if k_fixed is None:
        w = uniform(seed_for(config_id, 5), r, start)
        k = np.where(w < 0.30, 0, np.where(w < 0.70, 1, 2)).astype(np.int64)
    else:
        k = np.full(r, k_fixed, np.int64)
    off = np.zeros(r + 1, np.int64)
    np.cumsum(k, out=off[1:])
    total = int(off[-1])
    # id j of request q is element (start + q) * 4 + j of its own stream
    slot = np.repeat(np.arange(r, dtype=np.int64), k)
    j = np.arange(total, dtype=np.int64) - off[:-1][slot]
    words = splitmix64(seed_for(config_id, 6), (r + 1) * 4, start * 4)
    ids = (words[slot * 4 + j] % np.uint64(max(n_nodes, 1))).astype(np.int32)
    return min_battery, off.astype(np.int32), ids


def cluster_snapshot(config_id: int, n: int, p: int, gpus_per_node: int = 0):
    """Node/pod columns for calculateClusterMetrics (manager.go:493-565)."""
    cpu_choices = np.array([4000, 8000, 16000, 64000], np.int64)
    mem_choices = np.array([8, 16, 64, 256], np.int64) * (1 << 30)
    a = uniform(seed_for(config_id, 10), n)
    b = uniform(seed_for(config_id, 11), n)
    cpu_cap = cpu_choices[np.minimum((a * 4).astype(np.int64), 3)]
    mem_cap = mem_choices[np.minimum((b * 4).astype(np.int64), 3)]
    cpu_use = (cpu_cap.astype(np.float64) * uniform(seed_for(config_id, 12), n)).astype(np.int64)
    mem_use = (mem_cap.astype(np.float64) * uniform(seed_for(config_id, 13), n)).astype(np.int64)
    healthy = (uniform(seed_for(config_id, 14), n) >= 0.03).astype(np.uint8)
    gpu_count = np.full(n, gpus_per_node, np.int32)
    gpu_usage = uniform(seed_for(config_id, 15), n * gpus_per_node) * 100.0
    pod_running = (uniform(seed_for(config_id, 16), p) >= 0.08).astype(np.uint8)
    disk_cap = np.full(n, 512 << 30, np.int64)
    disk_alloc = (disk_cap.astype(np.float64) * (0.5 + 0.5 * uniform(seed_for(config_id, 17), n))).astype(np.int64)
    return {"cpu_capacity": cpu_cap, "cpu_usage": cpu_use, "memory_capacity": mem_cap,
            "memory_usage": mem_use, "gpu_count": gpu_count, "healthy": healthy,
            "gpu_usage": gpu_usage, "pod_running": pod_running,
            "disk_capacity": disk_cap, "disk_allocatable": disk_alloc}


def pod_containers(config_id: int, p: int):
    """CSR of containers per pod (1-3) with requests/limits/usage columns."""
    k = 1 + np.minimum((uniform(seed_for(config_id, 20), p) * 3).astype(np.int64), 2)
    off = np.zeros(p + 1, np.int64)
    np.cumsum(k, out=off[1:])
    c = int(off[-1])
    u = [uniform(seed_for(config_id, 21 + i), c) for i in range(6)]
    cpu_req = (u[0] * 2000).astype(np.int64)
    cpu_lim = np.where(u[1] < 0.2, 0, cpu_req + (u[1] * 2000).astype(np.int64))
    mem_req = (u[2] * (4 << 30)).astype(np.int64)
    mem_lim = np.where(u[3] < 0.2, 0, mem_req + (u[3] * (4 << 30)).astype(np.int64))
    cpu_use = (u[4] * 2500).astype(np.int64)
    mem_use = (u[5] * (5 << 30)).astype(np.int64)
    return {"ctr_off": off.astype(np.int32), "cpu_usage": cpu_use, "memory_usage": mem_use,
            "cpu_request": cpu_req, "cpu_limit": cpu_lim, "memory_request": mem_req,
            "memory_limit": mem_lim}


# C2 series layout: 2 per pod (cpu millicores, memory bytes) then 6 per node
# (cpu, memory, cpu rate, memory rate, disk rate, network latency ms).
def c2_series_scales(n_nodes: int, n_pods: int) -> np.ndarray:
    pod = np.tile(np.array([4000.0, float(1 << 34)]), n_pods)
    node = np.tile(np.array([4000.0, float(1 << 34), 100.0, 100.0, 100.0, 100.0]), n_nodes)
    return np.concatenate([pod, node])


def c2_ring_rows(config_id: int, scales: np.ndarray, s0: int, s1: int, t: int) -> np.ndarray:
    """f32 [s1 - s0, t]: U[0, scale_s), element (s, k) is word s * t + k of the stream."""
    u = uniform(seed_for(config_id, 30), (s1 - s0) * t, s0 * t).reshape(s1 - s0, t)
    return (u * scales[s0:s1, None]).astype(np.float32)


# ---- C3: UAV channels (SURVEY.md §8a A11 channel map; simulator formulas)
N_UAV_CHANNELS = 64


def uav_channel_rows(config_id: int, node0: int, node1: int, t: int, dt: float = 1.0) -> np.ndarray:
    """f32 [(node1 - node0) * 64, t]: the 64-channel history of each UAV node.
    Armed AUTO flight on a circle (mavlink_simulator.go:272-322), per-node phase
    offset, battery draining from a per-node start level; rand.Float64 noise
    terms come from the seeded stream. Channels 43-63 are reserved (zero)."""
    nn = node1 - node0
    ch = N_UAV_CHANNELS
    out = np.zeros((nn, ch, t), np.float32)
    phase = uniform(seed_for(config_id, 40), nn, node0)[:, None] * 600.0
    start_pct = 40.0 + 60.0 * uniform(seed_for(config_id, 41), nn, node0)[:, None]
    el = np.arange(t, dtype=np.float64)[None, :] * dt + phase  # elapsed seconds

    def noise(k):  # one U[0,1) per (node, sample) for noise slot k
        return uniform(seed_for(config_id, 50 + k), nn * t, node0 * t).reshape(nn, t)

    omega, radius = 0.1, 0.001
    gs = 5.0 + noise(0) * 0.5
    cog = np.mod(omega * el * 180 / np.pi, 360)
    # 0.1 % per second while armed, clamped at 0 (mavlink_simulator.go:309-314)
    pct = np.maximum(start_pct - 0.1 * dt * np.arange(t, dtype=np.float64)[None, :], 0.0)
    throttle = 50.0 + 20.0 * np.sin(0.1 * el)
    total_cap = 5000.0
    cur = 10.0 + throttle * 0.2
    rem_cap = total_cap * pct / 100.0
    vals = {
        0: 39.9042 + radius * np.cos(omega * el), 1: 116.4074 + radius * np.sin(omega * el),
        2: 100.0 + 50.0 + 10.0 * np.sin(0.05 * el), 3: 50.0 + 10.0 * np.sin(0.05 * el),
        4: 0.8 + 0.4 * noise(1), 5: np.floor(8 + 5 * noise(2)), 6: np.full((nn, t), 3.0),
        7: gs, 8: cog,
        9: 5.0 * np.sin(0.5 * el) + noise(3) * 0.5, 10: 3.0 * np.cos(0.3 * el) + noise(4) * 0.3,
        11: np.mod(cog, 360), 12: noise(5) * 2.0 - 1.0, 13: noise(6) * 2.0 - 1.0, 14: noise(7) * 5.0 - 2.5,
        15: gs + noise(8) * 0.5, 16: gs, 17: np.cos(0.05 * el) * 2.0, 18: throttle,
        19: 22.2 - (100.0 - pct) * 0.04, 20: cur, 21: pct, 22: rem_cap, 23: np.full((nn, t), total_cap),
        24: 25.0 + (100.0 - pct) * 0.3, 25: np.full((nn, t), 6.0), 26: np.floor(rem_cap / cur * 3600.0 / 1000.0),
        27: np.floor(np.mod(el / 60.0, 10)), 28: np.full((nn, t), 10.0), 29: 100.0 * noise(9),
        30: np.floor(20 * noise(10)), 31: np.floor(noise(11) * 1.02), 32: np.floor(noise(12) * 1.05),
        33: np.ones((nn, t)),
    }
    for c in range(34, 40):
        vals[c] = np.ones((nn, t))
    vals[40] = np.full((nn, t), 3.0)  # mode AUTO
    vals[41] = np.ones((nn, t))       # mission ACTIVE
    vals[42] = np.zeros((nn, t))      # system status OK
    for c, v in vals.items():
        out[:, c, :] = np.broadcast_to(v, (nn, t)).astype(np.float32)
    return out.reshape(nn * ch, t)


# ------------------------------------------------------------------ torch (device-side generation)

def _to_i64(x: int) -> int:
    x &= 0xFFFFFFFFFFFFFFFF
    return x - (1 << 64) if x >= (1 << 63) else x


def splitmix64_torch(seed: int, n: int, start: int = 0, device="cuda"):
    """Same words as splitmix64(), as int64 bit patterns, computed on `device`.
    int64 multiply wraps; logical right shifts are arithmetic shifts + a mask."""
    import torch
    i = torch.arange(start + 1, start + n + 1, dtype=torch.int64, device=device)
    z = i * _to_i64(_GOLDEN) + _to_i64(seed)

    def lsr(v, k):
        return (v >> k) & ((1 << (64 - k)) - 1)

    z = (z ^ lsr(z, 30)) * _to_i64(_M1)
    z = (z ^ lsr(z, 27)) * _to_i64(_M2)
    return z ^ lsr(z, 31)


def uniform_torch(seed: int, n: int, start: int = 0, device="cuda"):
    import torch
    w = splitmix64_torch(seed, n, start, device)
    return ((w >> 11) & ((1 << 53) - 1)).to(torch.float64) * (1.0 / 9007199254740992.0)


def c2_ring_rows_torch(config_id: int, scales, s0: int, s1: int, t: int, device="cuda"):
    """Bit-identical to c2_ring_rows (integer mixing + exact IEEE multiplies)."""
    import torch
    u = uniform_torch(seed_for(config_id, 30), (s1 - s0) * t, s0 * t, device).reshape(s1 - s0, t)
    sc = torch.as_tensor(scales[s0:s1], dtype=torch.float64, device=device)[:, None]
    return (u * sc).to(torch.float32)
