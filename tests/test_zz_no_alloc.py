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
