#!/usr/bin/env python
"""Where a sharded C5 step goes (run under torchrun, one rank per GPU): per rank, with an L2 flush
and a device-side rendezvous in front of every measurement, CUDA-event times of
  k1_publish   K1 over the rank's shard + 16-byte publication
  k1_spread    the same + the push of the shard into every peer's replica
  k2_local     sharded K2 on the rank's request slice, slots already published, local replica
  k2_probe     the same, probing the peers' shards through peer pointers
  step_spread  K1 spread + K2 (table_from_k1)      step_probe  K1 publish + K2 probing peers
Prints one JSON line per rank."""
import ctypes
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
pkg = ge.load_package()
dev, syn, b = pkg.Dev(), pkg.synth, pkg.binding
N = int(os.environ.get("C5_NODES", 1_000_000))
R = int(os.environ.get("C5_REQUESTS", 1_000_000))
cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
bat, flags = syn.node_table(5, N)
r0, r1 = R * rank // world, R * (rank + 1) // world
mb, off, ids = syn.requests(5, r1 - r0, N, start=r0)
d_mb, d_off, d_ids = cu(mb), cu(off), cu(ids)
lo, hi = b.shard_range(N, world, rank)
ranges = [b.shard_range(N, world, k) for k in range(world)]
sb, sf = symm_mem.empty(N, dtype=torch.float64, device="cuda"), symm_mem.empty(N, dtype=torch.uint8, device="cuda")
hb, hf = symm_mem.rendezvous(sb, dist.group.WORLD), symm_mem.rendezvous(sf, dist.group.WORLD)
sb.copy_(cu(bat)); sf.copy_(cu(flags))           # complete replicas: every variant gives the same answers
nb = int(dev.lib.klm_dev_shard_slots_bytes())
slots, sync = symm_mem.empty(nb, dtype=torch.uint8, device="cuda"), symm_mem.empty(nb, dtype=torch.uint8, device="cuda")
slots.zero_(); sync.zero_()
hs, hy = symm_mem.rendezvous(slots, dist.group.WORLD), symm_mem.rendezvous(sync, dist.group.WORLD)
slot_ptrs, sync_ptrs = [int(x) for x in hs.buffer_ptrs], [int(x) for x in hy.buffer_ptrs]
rep_b, rep_f = [int(x) for x in hb.buffer_ptrs], [int(x) for x in hf.buffer_ptrs]
loc_b = [sb[a:].data_ptr() if e > a else 0 for a, e in ranges]
loc_f = [sf[a:].data_ptr() if e > a else 0 for a, e in ranges]
peer_b = [rep_b[k] + 8 * ranges[k][0] for k in range(world)]     # rank k's shard inside ITS replica
peer_f = [rep_f[k] + ranges[k][0] for k in range(world)]
o_idx = torch.empty(r1 - r0, dtype=torch.int32, device="cuda")
o_score = torch.empty(r1 - r0, dtype=torch.float64, device="cuda")
epoch, sepoch = ctypes.c_uint64(0), ctypes.c_uint64(0)
k1p = dev.prebind("node_best_publish", sb[lo:hi], sf[lo:hi], lo, world, rank, slot_ptrs, epoch)
k1s = dev.prebind("node_best_spread", sb[lo:hi], sf[lo:hi], lo, world, rank, slot_ptrs, rep_b, rep_f, epoch)
k2l = dev.prebind("schedule_sharded", N, loc_b, loc_f, slots, epoch, d_mb, d_off, d_ids, o_idx, o_score)
k2w = dev.prebind("schedule_sharded", N, loc_b, loc_f, slots, epoch, d_mb, d_off, d_ids, o_idx, o_score, table_from_k1=True)
k2p = dev.prebind("schedule_sharded", N, peer_b, peer_f, slots, epoch, d_mb, d_off, d_ids, o_idx, o_score)
rdv = dev.prebind("shard_rendezvous", world, rank, sync_ptrs, sync, sepoch)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
dist.barrier(); torch.cuda.synchronize()


def measure(fns, bump, reps=60):
    ts = []
    for i in range(reps + 5):
        flush.zero_()
        sepoch.value += 1
        rdv()
        if bump:
            epoch.value += 1
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for f in fns:
            f()
        e1.record()
        e1.synchronize()
        if i >= 5:
            ts.append(e0.elapsed_time(e1) * 1e3)
    return round(float(np.median(ts)), 2)


out = {"rank": rank, "world": world, "nodes": N, "requests": R}
out["k1_publish_us"] = measure([k1p], True)
out["k1_spread_us"] = measure([k1s], True)
out["k2_local_us"] = measure([k2l], False)     # epoch unchanged: the slots already carry it
out["k2_probe_us"] = measure([k2p], False)
out["step_spread_us"] = measure([k1s, k2w], True)
out["step_probe_us"] = measure([k1p, k2p], True)
out["step_publish_local_us"] = measure([k1p, k2l], True)
res = [None] * world
dist.all_gather_object(res, out)
if rank == 0:
    for r in res:
        print(json.dumps(r), flush=True)
dist.barrier()
dist.destroy_process_group()
