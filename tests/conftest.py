import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session")
def pkg():
    import __graft_entry__ as ge
    lib = os.path.join(ge.PKG_DIR, "libklm.so")
    if not os.path.exists(lib):
        ge.build()
    return ge.load_package()


@pytest.fixture(scope="session")
def handle(pkg):
    # GPU tests only: fails loudly (no fallback) when there is no device
    h = pkg.Handle(0)
    yield h
    h.close()


@pytest.fixture(scope="session")
def dev(pkg):
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    torch.cuda.set_device(0)
    return pkg.Dev()
