"""GPU: the opt-in chain that connects the window statistics to the scheduler (SURVEY.md 8f row 4;
no reference semantics, parity unpinned): UAV ring -> K7 -> klm_dev_node_features (stale
heartbeats, hot batteries, battery projected at the window's drain) -> K1 + K2. Every stage is
checked against the oracle's restatement of the same definition, and the decisions against the
oracle scheduler run on the oracle's derived table."""
import numpy as np
import pytest

import oracle_binding as ob

pytestmark = pytest.mark.gpu


def test_window_features_feed_the_scheduler(dev, pkg):
    import torch
    nodes, t = 300, 600
    rows = pkg.synth.uav_channel_rows(3, 0, nodes, t)                      # [nodes * 64, t]
    rng = np.random.default_rng(4)
    rows = rows.reshape(nodes, 64, t).copy()
    rows[::9, 24, -40:] = 95.0                                             # some batteries run hot lately
    rows = rows.reshape(nodes * 64, t)
    ring = torch.from_numpy(rows).cuda()
    seen = torch.full((nodes * 64,), t, dtype=torch.int32, device="cuda")
    o = {k: torch.empty(nodes * 64, dtype=torch.float64 if k == "mean" else torch.float32, device="cuda")
         for k in ("min", "max", "mean", "p95")}
    state = torch.full((nodes * 64, 2), float("nan"), dtype=torch.float32, device="cuda")
    dev.window_stats(ring, seen, t, o["min"], o["max"], o["mean"], o["p95"], state=state)
    battery = rows.reshape(nodes, 64, t)[:, 21, -1].astype(np.float64)     # channel 21's latest sample
    flags = np.full(nodes, 3, np.uint8)
    flags[rng.random(nodes) < 0.05] = 1
    age = rng.choice([1.0, 5.0, 29.9, 30.0, 30.1, 400.0], size=nodes)
    prm = dict(max_age_s=30.0, temp_limit=60.0, horizon_s=120.0, span_s=float(t))
    d_b, d_f = torch.from_numpy(battery).cuda(), torch.from_numpy(flags).cuda()
    b_out, f_out = torch.empty_like(d_b), torch.empty_like(d_f)
    dev.node_features(d_b, d_f, torch.from_numpy(age).cuda(), o["min"], o["max"], o["p95"], prm, b_out, f_out)
    torch.cuda.synchronize()
    want_stats = ob.window_stats(rows, np.full(nodes * 64, t, np.uint32), t, t)
    wb, wf = ob.node_features(battery, flags, age, want_stats["min"], want_stats["max"], want_stats["p95"], **prm)
    assert np.array_equal(b_out.cpu().numpy().view(np.uint64), wb.view(np.uint64))
    assert np.array_equal(f_out.cpu().numpy(), wf)
    assert 0 < int((wf != flags).sum()) < nodes and (wb < battery).any() and (wb >= 0).all()
    # the scheduler on the derived table
    mb, off, ids = pkg.synth.requests(3, 2000, nodes, k_fixed=2)
    best = torch.empty(2, dtype=torch.int64, device="cuda")
    oi = torch.empty(mb.size, dtype=torch.int32, device="cuda")
    os_ = torch.empty(mb.size, dtype=torch.float64, device="cuda")
    t_ = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    dev.node_best(b_out, f_out, best)
    dev.schedule(b_out, f_out, best, t_(mb), t_(off), t_(ids), oi, os_, after_best=True)
    torch.cuda.synchronize()
    wi, ws = ob.schedule_literal(wb, wf, mb, off, ids)
    assert np.array_equal(oi.cpu().numpy(), wi) and np.array_equal(os_.cpu().numpy(), ws)
    # every rule can be switched off
    off_prm = dict(max_age_s=0.0, temp_limit=0.0, horizon_s=0.0, span_s=0.0)
    dev.node_features(d_b, d_f, None, None, None, None, off_prm, b_out, f_out)
    torch.cuda.synchronize()
    assert np.array_equal(b_out.cpu().numpy(), battery) and np.array_equal(f_out.cpu().numpy(), flags)
