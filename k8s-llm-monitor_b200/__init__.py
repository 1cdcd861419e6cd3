"""klm_b200 — B200-native data path for k8s-llm-monitor's scheduler and metrics manager.

The product is libklm.so (hand-written sm_100a kernels behind the C ABI of
include/klm.h). This package holds its sources (csrc/), the C++ host layer that
mirrors the reference's controller/manager interfaces (host/), a ctypes binding
and the seeded synthetic-cluster generators used by the tests and bench.py.

The directory name has a hyphen, so import it through
`__graft_entry__.load_package()` (registers it as `klm_b200`).
"""
from . import binding, hostbinding, sharding, synth  # noqa: F401
from .binding import Dev, Handle, KlmError, load  # noqa: F401
