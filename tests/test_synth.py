"""CPU: the seeded generators are deterministic, sliceable, and the torch (device-side)
variants produce the same bits as numpy."""
import numpy as np
import torch


def test_splitmix_known_answer(pkg):
    # splitmix64 with seed 0: first outputs of the published reference implementation
    w = pkg.synth.splitmix64(0, 3)
    assert [int(x) for x in w] == [0xE220A8397B1DCDAF, 0x6E789E6AA1B965F4, 0x06C45D188009454F]


def test_slices_are_independent(pkg):
    s = pkg.synth
    full = s.uniform(s.seed_for(2, 30), 1000)
    assert np.array_equal(full[100:300], s.uniform(s.seed_for(2, 30), 200, 100))
    sc = s.c2_series_scales(4, 10)
    a = s.c2_ring_rows(2, sc, 0, sc.size, 50)
    b = s.c2_ring_rows(2, sc, 7, 19, 50)
    assert np.array_equal(a[7:19], b)
    mb, off, ids = s.requests(4, 1000, 500)
    mb2, off2, ids2 = s.requests(4, 300, 500, start=200)
    assert np.array_equal(mb[200:500], mb2)
    assert np.array_equal(ids[off[200]:off[500]], ids2)


def test_torch_generators_bit_equal(pkg):
    s = pkg.synth
    w = s.splitmix64(s.seed_for(2, 30), 4096, 17)
    wt = s.splitmix64_torch(s.seed_for(2, 30), 4096, 17, device="cpu").numpy().view(np.uint64)
    assert np.array_equal(w, wt)
    sc = s.c2_series_scales(3, 5)
    a = s.c2_ring_rows(2, sc, 2, 20, 64)
    b = s.c2_ring_rows_torch(2, sc, 2, 20, 64, device="cpu").numpy()
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))


def test_node_table_shape(pkg):
    bat, flags = pkg.synth.node_table(2, 20000)
    assert bat.min() >= 0 and bat.max() <= 99.9
    assert np.allclose(bat * 10, np.round(bat * 10))  # 0.1 grid -> ties
    assert 0.01 < np.mean((flags & 1) == 0) < 0.03 and 0.04 < np.mean((flags & 2) == 0) < 0.06
    rows = pkg.synth.uav_channel_rows(3, 0, 2, 600)
    assert rows.shape == (128, 600) and np.isfinite(rows).all()
    assert (rows[43:64] == 0).all()
