"""Host-side multi-GPU plumbing: one process per GPU, torch.distributed for the exchange.

Two ways to split a scheduling batch:
  * by request  — every rank holds the whole node table and a contiguous slice of the
    requests; results are disjoint slices, no exchange on the data path;
  * by node range — rank r holds nodes [r N / G, (r + 1) N / G) and sees every request; each
    rank produces its partial winner per request with GLOBAL node indices, the partials are
    all-gathered (NCCL over NVLink on GPUs, gloo in the CPU tests) and folded by
    (score desc, global index asc), which is exactly the reference's "stable sort, take [0]"
    (internal/scheduler/controller.go:160-161) because the ranges are contiguous.

The compute steps are injected (`partial_fn`, `merge_fn`): on a GPU they are the klm_dev_*
launchers (see `gpu_ops`); the CPU tests pass stand-ins so the partition / gather / merge
logic can run under gloo without a device.
"""
from __future__ import annotations

from typing import Callable, Tuple


def split_range(n: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous balanced split of range(n): rank r gets [n r / G, n (r + 1) / G)."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    return n * rank // world, n * (rank + 1) // world


class NodeShardedScheduler:
    """Node-range sharding of one scheduling batch across the ranks of a process group."""

    def __init__(self, n_nodes_total: int, partial_fn: Callable, merge_fn: Callable, group=None):
        import torch.distributed as dist
        self.dist = dist
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.n_total = n_nodes_total
        self.lo, self.hi = split_range(n_nodes_total, self.world, self.rank)
        self.partial_fn = partial_fn
        self.merge_fn = merge_fn

    def schedule(self, local_battery, local_flags, min_battery, pref_off, pref_ids):
        """local_* cover nodes [lo, hi); the request arrays are identical on every rank.
        Returns (idx, score) for every request, identical on every rank."""
        import torch
        if local_battery.shape[0] != self.hi - self.lo:
            raise ValueError("local node arrays do not match this rank's range")
        r = min_battery.shape[0]
        p_idx, p_score = self.partial_fn(local_battery, local_flags, self.lo, min_battery, pref_off, pref_ids)
        # flat [world * r] outputs (rank-major) are accepted by both NCCL and gloo
        g_idx = torch.empty(self.world * r, dtype=torch.int32, device=p_idx.device)
        g_score = torch.empty(self.world * r, dtype=torch.float64, device=p_idx.device)
        self.dist.all_gather_into_tensor(g_idx, p_idx.contiguous(), group=self.group)
        self.dist.all_gather_into_tensor(g_score, p_score.contiguous(), group=self.group)
        return self.merge_fn(g_idx.view(self.world, r), g_score.view(self.world, r))


class FusedShardedScheduler:
    """The round-2 form of node-range sharding (klm.h, "node-range sharding without an all-gather"):
    every rank owns one shard of the node table AND one slice of the requests.

      1. best_fn(local_battery, local_flags, node_base) -> (battery, global index) of the shard's
         best eligible node (K1). On GPUs the 16-byte result is stored into every rank's exchange
         slots over NVLink by the kernel itself; here an all_gather of one (f64, i64) record per
         rank stands in for those stores.
      2. The G records are folded: largest battery, lowest global index on ties
         (controller.go:160 across shards).
      3. decide_fn(table_view, g_idx, g_bat, request slice) -> the slice's decisions (sharded K2).
         table_view(ids) returns (flags, battery) of arbitrary global node ids: on GPUs a peer
         pointer read, or a read of the local replica that klm_dev_node_best_spread filled; here
         the shards are all-gathered once per snapshot to play that role.
    Results are disjoint request slices; no rank ever holds partial winners of other ranks'
    requests, and nothing is merged afterwards."""

    def __init__(self, n_nodes_total: int, best_fn: Callable, decide_fn: Callable, group=None):
        import torch.distributed as dist
        self.dist, self.group = dist, group
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.n_total = n_nodes_total
        self.best_fn, self.decide_fn = best_fn, decide_fn
        self.lo, self.hi = shard_range(n_nodes_total, self.world, self.rank)

    def schedule(self, local_battery, local_flags, min_battery, pref_off, pref_ids):
        """local_* cover nodes [lo, hi); the request arrays are this rank's SLICE (offsets local to
        the slice). Returns (idx, score) for the slice."""
        import torch
        if local_battery.shape[0] != self.hi - self.lo:
            raise ValueError("local node arrays do not match this rank's shard")
        bat, idx = self.best_fn(local_battery, local_flags, self.lo)
        rec_b = torch.tensor([float(bat)], dtype=torch.float64)
        rec_i = torch.tensor([int(idx)], dtype=torch.int64)
        all_b = [torch.zeros(1, dtype=torch.float64) for _ in range(self.world)]
        all_i = [torch.zeros(1, dtype=torch.int64) for _ in range(self.world)]
        self.dist.all_gather(all_b, rec_b, group=self.group)     # the slot exchange
        self.dist.all_gather(all_i, rec_i, group=self.group)
        g_idx, g_bat = -1, 0.0
        for b, i in zip(all_b, all_i):                           # the fold of the sharded K2
            b, i = float(b), int(i)
            if i >= 0 and (g_idx < 0 or b > g_bat or (b == g_bat and i < g_idx)):
                g_idx, g_bat = i, b
        shards_b, shards_f = [None] * self.world, [None] * self.world
        self.dist.all_gather_object(shards_b, local_battery, group=self.group)   # the view of the peers' shards
        self.dist.all_gather_object(shards_f, local_flags, group=self.group)
        ranges = [shard_range(self.n_total, self.world, k) for k in range(self.world)]

        def view(node_id: int):
            if node_id < 0 or node_id >= self.n_total:
                return None
            L = ranges[0][1] - ranges[0][0]
            k = min(node_id // L, self.world - 1) if L > 0 else 0
            lo = ranges[k][0]
            return int(shards_f[k][node_id - lo]), float(shards_b[k][node_id - lo])

        return self.decide_fn(view, g_idx, g_bat, min_battery, pref_off, pref_ids)


def shard_range(n_nodes: int, world: int, rank: int) -> Tuple[int, int]:
    """klm_shard_range: shards of ceil(n / world) nodes rounded up to a multiple of 16."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    length = (-(-max(n_nodes, 0) // world) + 15) // 16 * 16
    lo, hi = min(rank * length, n_nodes), min((rank + 1) * length, n_nodes)
    return lo, hi


def request_shard(n_requests: int, world: int, rank: int) -> Tuple[int, int]:
    """Request sharding needs no exchange: rank r schedules requests [lo, hi) and owns
    out[lo:hi]."""
    return split_range(n_requests, world, rank)


def gpu_ops(dev):
    """(partial_fn, merge_fn) backed by the CUDA kernels (K1 + K2 with node_base, K3)."""
    import torch

    def partial_fn(battery, flags, node_base, min_battery, pref_off, pref_ids):
        r = min_battery.shape[0]
        best = torch.empty(2, dtype=torch.int64, device=battery.device)
        idx = torch.empty(r, dtype=torch.int32, device=battery.device)
        score = torch.empty(r, dtype=torch.float64, device=battery.device)
        dev.node_best(battery, flags, best, node_base=node_base)
        dev.schedule(battery, flags, best, min_battery, pref_off, pref_ids, idx, score, node_base=node_base)
        return idx, score

    def merge_fn(g_idx, g_score):
        r = g_idx.shape[1]
        idx = torch.empty(r, dtype=torch.int32, device=g_idx.device)
        score = torch.empty(r, dtype=torch.float64, device=g_idx.device)
        dev.merge_partials(g_idx, g_score, idx, score)
        return idx, score

    return partial_fn, merge_fn
