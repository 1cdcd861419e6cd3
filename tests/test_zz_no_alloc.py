"""GPU: klm_dev_node_best_static — the node_best_reduce entry point for hosts that may not
allocate device memory (their HBM is fully reserved, or cudaMalloc is banned on the
reconcile path because it synchronises the device). Its scratch is a fixed array inside
libklm.so instead of a caller buffer.

Called through the C ABI and compared with klm_dev_node_best (caller scratch) on two calls
in a row: the shared array keeps the first call's contents, so the second call proves the
entry point clears what has to start at zero. Results are integers and f64 bit patterns:
the comparison is exact, like every other scheduling test here.

Collected last (file name), so a -x run has already exercised everything else.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _both(pkg, battery, flags, node_base):
    import torch
    lib = pkg.load()
    dev = pkg.Dev()
    stream = torch.cuda.current_stream().cuda_stream
    d_bat = torch.from_numpy(battery).cuda()
    d_flags = torch.from_numpy(flags).cuda()
    ref = torch.full((2,), -7, dtype=torch.int64, device="cuda")
    got = torch.full((2,), -9, dtype=torch.int64, device="cuda")
    dev.node_best(d_bat, d_flags, ref, node_base=node_base)          # existing entry point
    rc = lib.klm_dev_node_best_static(d_bat.data_ptr(), d_flags.data_ptr(), battery.size, node_base,
                                      got.data_ptr(), stream)        # no scratch argument
    assert rc == 0, pkg.binding.last_error()
    torch.cuda.synchronize()
    return ref.cpu().numpy(), got.cpu().numpy()


def test_static_scratch_matches_on_two_calls_in_a_row(pkg):
    import torch
    assert torch.cuda.is_available()
    torch.cuda.set_device(0)
    # call 1: many blocks, heavy ties (0.1 grid), ineligible nodes mixed in
    bat1, flags1 = pkg.synth.node_table(5, 1_000_003)
    ref, got = _both(pkg, bat1, flags1, node_base=0)
    assert np.array_equal(ref, got), (ref, got)
    assert ref[0] >= 0
    # call 2, straight after: a different, smaller table with an index offset; then a table
    # where nobody is eligible and an empty one
    bat2, flags2 = pkg.synth.node_table(3, 5000)
    ref, got = _both(pkg, bat2, flags2, node_base=123_456)
    assert np.array_equal(ref, got), (ref, got)
    ref, got = _both(pkg, bat2, np.ones_like(flags2), node_base=0)
    assert np.array_equal(ref, got) and got[0] == -1
    ref, got = _both(pkg, np.zeros(0), np.zeros(0, np.uint8), node_base=0)
    assert np.array_equal(ref, got) and got[0] == -1


def test_static_scratch_rejects_oversize(pkg):
    import torch
    torch.cuda.set_device(0)
    lib = pkg.load()
    limit = lib.klm_dev_node_best_static_max_nodes()
    assert limit == 4_849_664
    best = torch.zeros(2, dtype=torch.int64, device="cuda")
    one = torch.zeros(1, dtype=torch.float64, device="cuda")
    flag = torch.zeros(1, dtype=torch.uint8, device="cuda")
    # the size check comes before any memory is touched, so tiny buffers are safe here
    rc = lib.klm_dev_node_best_static(one.data_ptr(), flag.data_ptr(), limit + 1, 0, best.data_ptr(),
                                      torch.cuda.current_stream().cuda_stream)
    assert rc == pkg.binding.KLM_ELIMIT
    assert "exceed" in pkg.binding.last_error()


def test_static_k1_feeds_the_dependent_k2(pkg):
    """The pairing the header names: klm_dev_node_best_static followed by klm_dev_schedule_after_best
    on the same stream must decide like the allocating pair, on two rounds in a row."""
    import torch
    import oracle_binding as ob
    torch.cuda.set_device(0)
    lib, dev = pkg.load(), pkg.Dev()
    stream = torch.cuda.current_stream().cuda_stream
    for seed, n, r in ((4, 120_000, 50_000), (5, 3000, 800)):
        bat, flags = pkg.synth.node_table(seed, n)
        mb, off, ids = pkg.synth.requests(seed, r, n)
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
        d_bat, d_flags, d_mb, d_off, d_ids = t(bat), t(flags), t(mb), t(off), t(ids)
        best = torch.full((2,), -9, dtype=torch.int64, device="cuda")
        oi = torch.empty(r, dtype=torch.int32, device="cuda")
        os_ = torch.empty(r, dtype=torch.float64, device="cuda")
        rc = lib.klm_dev_node_best_static(d_bat.data_ptr(), d_flags.data_ptr(), n, 0, best.data_ptr(), stream)
        assert rc == 0, pkg.binding.last_error()
        dev.schedule(d_bat, d_flags, best, d_mb, d_off, d_ids, oi, os_, after_best=True)
        torch.cuda.synchronize()
        want = ob.schedule_decomposed(bat, flags, mb, off, ids)
        assert np.array_equal(oi.cpu().numpy(), want[0]) and np.array_equal(os_.cpu().numpy(), want[1])


def test_static_scratch_on_a_second_device(pkg):
    """Every GPU that loads the library has its own copy of the array; its address is looked up
    with that device current."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    lib = pkg.load()
    bat, flags = pkg.synth.node_table(6, 200_000)
    want = None
    for d in (1, 0, 1):
        with torch.cuda.device(d):
            d_bat, d_flags = torch.from_numpy(bat).cuda(), torch.from_numpy(flags).cuda()
            got = torch.full((2,), -9, dtype=torch.int64, device="cuda")
            rc = lib.klm_dev_node_best_static(d_bat.data_ptr(), d_flags.data_ptr(), bat.size, 0, got.data_ptr(),
                                              torch.cuda.current_stream().cuda_stream)
            assert rc == 0, pkg.binding.last_error()
            torch.cuda.synchronize()
            g = got.cpu().numpy()
            want = g if want is None else want
            assert np.array_equal(g, want) and g[0] >= 0
    torch.cuda.set_device(0)
