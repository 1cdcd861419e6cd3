"""A short run of tools/k7_fuzz.py under pytest: adversarial rings (signed zeros, infinities,
denormals, plateaus next to spikes, one-lane pile-ups, random fill levels) through every K7
variant and through runs of ticks with carried / garbage brackets, against the oracle. The long
runs (minutes, several seeds) are done by hand with the same tool; this keeps the path alive."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("seed", [11, 12])
def test_k7_fuzz_short(seed):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "k7_fuzz.py"), "8", str(seed)],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, (r.stdout[-800:], r.stderr[-800:])
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["ok"] and line["kernel_calls_checked"] > 50


@pytest.mark.parametrize("seed", [21])
def test_sched_fuzz_short(seed):
    """tools/sched_fuzz.py: K1 + K2, the dependent launch, the fused launch, the brute-force scan and
    node-range shards (probe and spread, 2..8 emulated ranks) on adversarial fleets and requests."""
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "sched_fuzz.py"), "8", str(seed)],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, (r.stdout[-800:], r.stderr[-800:])
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["ok"] and line["checks"] > 100


def test_agg_fuzz_short():
    """tools/agg_fuzz.py: cluster aggregate, node / pod rates, predicates, fleet scan and the top-k
    batch on adversarial integer columns and CSR shapes, through the host-buffer ABI."""
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "agg_fuzz.py"), "8", "31"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, (r.stdout[-800:], r.stderr[-800:])
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["ok"] and line["checks"] > 50
