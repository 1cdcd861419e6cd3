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
