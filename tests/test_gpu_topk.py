"""GPU parity of top-k scheduling against the literal oracle: the candidate
list of controller.go:174-221, stable descending sort (:160), cut at k. k = 1 must reproduce
klm_schedule_batch bit for bit; beyond [0] the reference has no semantics (parity unpinned)."""
import numpy as np
import pytest

import helpers
import oracle_binding as ob

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("k", [1, 2, 3, 8])
def test_topk_equals_literal_sort(handle, pkg, k):
    rng = np.random.default_rng(50 + k)
    for n, r, kmax, grid in ((1, 4, 2, 1.0), (5, 30, 3, 1.0), (300, 400, 4, 0.2), (5000, 1500, 12, 1.0),
                             (70_000, 300, 3, 10.0)):
        bat, flags, mb, off, ids = helpers.random_problem(rng, n, r, kmax=kmax, grid=grid)
        handle.nodes_set(bat, flags)
        got = handle.schedule_batch_topk(mb, off, ids, k)
        want = ob.schedule_topk(bat, flags, mb, off, ids, k)
        assert np.array_equal(got[0], want[0]), (n, r, k)
        assert np.array_equal(got[1], want[1]), (n, r, k)
        one = handle.schedule_batch(mb, off, ids)
        assert np.array_equal(got[0][:, 0], one[0]) and np.array_equal(got[1][:, 0], one[1])


def test_topk_fixture_and_limits(handle, pkg):
    # examples/multi-pod-request.yaml on the fleet of deployments/uav-simulator.yaml (85, 75, 90;
    # min 60, agent-0 and agent-1 preferred): scores 85, 85, 100 -> agent-1, then the 85/85 tie by index
    handle.nodes_set(np.array([85.0, 75.0, 90.0]), np.array([3, 3, 3], np.uint8))
    idx, score = handle.schedule_batch_topk(np.array([60.0]), np.array([0, 2], np.int32), np.array([1, 2], np.int32), 3)
    assert idx.tolist() == [[2, 0, 1]] and score.tolist() == [[100.0, 85.0, 85.0]]
    idx, score = handle.schedule_batch_topk(np.array([80.0]), np.array([0, 0], np.int32), np.zeros(0, np.int32), 4)
    assert idx.tolist() == [[2, 0, -1, -1]] and score.tolist() == [[90.0, 85.0, 0.0, 0.0]]
    b = pkg.binding
    with pytest.raises(b.KlmError) as e:
        handle.schedule_batch_topk(np.array([0.0]), np.array([0, 0], np.int32), np.zeros(0, np.int32), 9)
    assert e.value.code == b.KLM_ELIMIT
