"""CPU, world_size 2 over gloo: the partition / all-gather / merge logic of node-range and
request sharding (k8s-llm-monitor_b200/sharding.py). The CUDA kernels cannot run here, so
the two compute steps are stand-ins built on the oracle; what is under test is the host
logic around them (ranges, global indices, gather layout, tie-breaking across ranks)."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, n, r, seed, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import __graft_entry__ as ge
    import helpers
    import oracle_binding as ob
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    pkg = ge.load_package()
    sh = pkg.sharding
    rng = np.random.default_rng(seed)                      # same problem on every rank
    bat, flags, mb, off, ids = helpers.random_problem(rng, n, r, kmax=4, grid=1.0)
    ids = np.clip(ids, 0, n - 1).astype(np.int32)

    def partial_fn(lb, lf, base, mbt, offt, idst):
        # stand-in for K1 + K2: oracle on the local range, global ids shifted to local ones
        local = idst.numpy().astype(np.int64) - base
        local = np.where((local >= 0) & (local < lb.shape[0]), local, -1).astype(np.int32)
        i, s = ob.schedule_decomposed(lb.numpy(), lf.numpy(), mbt.numpy(), offt.numpy(), local)
        gi = np.where(i >= 0, i + base, -1).astype(np.int32)
        return torch.from_numpy(gi), torch.from_numpy(s)

    def merge_fn(g_idx, g_score):
        # stand-in for K3: (score desc, global index asc) over the ranks
        gi, gs = g_idx.numpy(), g_score.numpy()
        out_i = np.full(gi.shape[1], -1, np.int32)
        out_s = np.zeros(gi.shape[1])
        for g in range(gi.shape[0]):
            take = (gi[g] >= 0) & ((out_i < 0) | (gs[g] > out_s) | ((gs[g] == out_s) & (gi[g] < out_i)))
            out_i = np.where(take, gi[g], out_i)
            out_s = np.where(take, gs[g], out_s)
        return torch.from_numpy(out_i), torch.from_numpy(out_s)

    ns = sh.NodeShardedScheduler(n, partial_fn, merge_fn)
    lo, hi = ns.lo, ns.hi
    assert (lo, hi) == sh.split_range(n, world, rank)
    t = torch.from_numpy
    idx, score = ns.schedule(t(bat[lo:hi].copy()), t(flags[lo:hi].copy()), t(mb), t(off), t(ids))
    want_i, want_s = ob.schedule_literal(bat, flags, mb, off, ids)
    ok_nodes = np.array_equal(idx.numpy(), want_i) and np.array_equal(score.numpy(), want_s)

    # request sharding: each rank schedules its slice; gathered slices == whole batch
    r0, r1 = sh.request_shard(r, world, rank)
    sub_off = (off[r0:r1 + 1] - off[r0]).astype(np.int32)
    si, ss = ob.schedule_decomposed(bat, flags, mb[r0:r1], sub_off, ids[off[r0]:off[r1]])
    parts = [None] * world
    dist.all_gather_object(parts, (r0, r1, si, ss))
    full_i = np.concatenate([p[2] for p in sorted(parts, key=lambda p: p[0])])
    full_s = np.concatenate([p[3] for p in sorted(parts, key=lambda p: p[0])])
    ok_req = np.array_equal(full_i, want_i) and np.array_equal(full_s, want_s)
    covered = sorted((p[0], p[1]) for p in parts)
    ok_cover = covered[0][0] == 0 and covered[-1][1] == r and all(a[1] == b[0] for a, b in zip(covered, covered[1:]))
    # the fused form (round 2): every rank owns a node shard and a request slice; the 16-byte
    # exchange + fold + probes of the peers' shards give the slice's decisions directly
    def best_fn(lb, lf, base):
        bb, bi = 0.0, -1
        for i in range(lb.shape[0]):                       # controller.go:190, :198, :160-161
            if (int(lf[i]) & 3) == 3 and (bi < 0 or float(lb[i]) > bb):
                bb, bi = float(lb[i]), i
        return (bb, bi + base) if bi >= 0 else (0.0, -1)

    def decide_fn(view, g, gb, mbt, offt, idst):
        oi = np.full(mbt.shape[0], -1, np.int32)
        os_ = np.zeros(mbt.shape[0])
        for q in range(mbt.shape[0]):
            m = float(mbt[q])
            if g < 0 or (m > 0 and gb < m):
                continue
            bi, bs = g, gb
            for k in range(int(offt[q]), int(offt[q + 1])):
                got = view(int(idst[k]))
                if got is None or (got[0] & 3) != 3 or (m > 0 and got[1] < m):
                    continue
                sc = got[1] + 10
                if sc > bs or (sc == bs and int(idst[k]) < bi):
                    bi, bs = int(idst[k]), sc
            oi[q], os_[q] = bi, bs
        return oi, os_

    fs = sh.FusedShardedScheduler(n, best_fn, decide_fn)
    assert (fs.lo, fs.hi) == pkg.binding.shard_range(n, world, rank)     # the library's split
    q0, q1 = sh.request_shard(r, world, rank)
    f_off = (off[q0:q1 + 1] - off[q0]).astype(np.int32)
    fi, fsc = fs.schedule(bat[fs.lo:fs.hi].copy(), flags[fs.lo:fs.hi].copy(), mb[q0:q1], f_off, ids[off[q0]:off[q1]])
    ok_fused = np.array_equal(fi, want_i[q0:q1]) and np.array_equal(fsc, want_s[q0:q1])
    with open(os.path.join(out_dir, f"rank{rank}.txt"), "w") as f:
        f.write(f"{int(ok_nodes)} {int(ok_req)} {int(ok_cover)} {int(ok_fused)}")
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("n,r", [(1001, 257), (2, 50), (40, 33)])
def test_sharding_two_ranks_gloo(tmp_path, n, r):
    world = 2
    port = 29500 + (os.getpid() % 2000) + n % 7
    mp.spawn(_worker, args=(world, port, n, r, 11, str(tmp_path)), nprocs=world, join=True)
    for rank in range(world):
        assert open(tmp_path / f"rank{rank}.txt").read() == "1 1 1 1"


def test_split_range_properties(pkg):
    sh = pkg.sharding
    for n in (0, 1, 7, 100_000, 1_000_003):
        for g in (1, 2, 4, 8):
            parts = [sh.split_range(n, g, k) for k in range(g)]
            assert parts[0][0] == 0 and parts[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(parts, parts[1:]))
            sizes = [b - a for a, b in parts]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sh.split_range(10, 2, 2)
