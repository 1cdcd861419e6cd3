"""CPU: libklm.so loads and exports exactly what include/klm.h declares (no compute calls)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "klm.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(klm_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_symbols():
    syms = header_symbols()
    assert "klm_schedule_batch" in syms and "klm_dev_window_stats" in syms and len(syms) >= 30


def test_library_exports_every_declared_symbol(pkg):
    lib = ctypes.CDLL(pkg.binding.LIB_PATH)
    missing = [s for s in header_symbols() if not hasattr(lib, s)]
    assert not missing, f"declared in klm.h but not exported: {missing}"


def test_binding_covers_header(pkg):
    assert sorted(pkg.binding.SIGNATURES) == header_symbols()
    pkg.load()


def test_no_cpu_fallback_without_device(pkg):
    """On a machine without a GPU the handle must fail loudly, not fall back."""
    import torch
    if torch.cuda.is_available():
        return
    try:
        pkg.Handle(0)
    except pkg.KlmError as e:
        assert e.code == pkg.binding.KLM_ENODEV and "no CPU fallback" in str(e)
    else:
        raise AssertionError("klm_create succeeded without a CUDA device")


def test_product_does_not_touch_oracle():
    """Nothing under the package may reference the oracle (it is test infrastructure)."""
    pkg_dir = os.path.join(ROOT, "k8s-llm-monitor_b200")
    bad = []
    for dp, _, files in os.walk(pkg_dir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".hpp", "Makefile")):
                text = open(os.path.join(dp, f), errors="replace").read()
                # mentions in comments are fine; linking, loading or calling is not
                if re.search(r"klmo_\w+\s*\(|libklm_oracle|oracle_binding|controller_oracle|-lklm_oracle", text):
                    bad.append(os.path.join(dp, f))
    assert not bad, bad


def test_host_layer_loads(pkg):
    """libklm_host.so (C++ Controller / Manager mirror) loads and exports its C view."""
    lib = pkg.hostbinding.load()
    for name in ("klmh_open", "klmh_reconcile", "klmh_status", "klmh_calculate_cluster_metrics",
                 "klmh_low_battery_uavs"):
        assert hasattr(lib, name)


def test_shard_range_partitions_the_nodes(pkg):
    """klm_shard_range (host arithmetic, no GPU): contiguous, ordered, covering, shards of a
    multiple of 16 nodes so that every shard's columns keep 16-byte alignment."""
    sr = pkg.binding.shard_range
    for n in (0, 1, 3, 16, 17, 1000, 99_999, 100_000, 1_000_000, 2**31 - 1):
        for g in (1, 2, 3, 4, 8, 16):
            prev = 0
            for k in range(g):
                lo, hi = sr(n, g, k)
                assert lo == prev and lo <= hi <= n
                assert lo % 16 == 0 or lo == n
                prev = hi
            assert prev == n


def test_go_shim_names_every_host_buffer_symbol():
    """go/klm/klm.go (uncompiled here: no Go toolchain) must at least bind every klm_* host-buffer
    entry point the header declares; a regex over the source is what can be checked without a
    compiler. It must also fetch klm_last_error on a locked OS thread (the error text is
    thread-local and goroutines migrate)."""
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hdr = open(os.path.join(root, "include", "klm.h")).read()
    go = open(os.path.join(root, "go", "klm", "klm.go")).read()
    names = sorted(set(re.findall(r"\b(klm_(?!dev_)\w+)\s*\(", hdr)))
    assert len(names) > 25
    missing = [n for n in names if f"C.{n}(" not in go]
    assert not missing, missing
    assert "runtime.LockOSThread()" in go and go.count("C.klm_last_error()") == 1


def test_go_shim_calls_match_the_header_arity():
    """No Go toolchain here or on the GPU box, so the cgo shim has never been compiled; at least every
    C.klm_*(...) call in it must pass as many arguments as the prototype in include/klm.h takes
    (a mismatch would be a compile error on the maintainer's side)."""
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    h = re.sub(r"/\*.*?\*/", "", open(os.path.join(root, "include", "klm.h")).read(), flags=re.S)
    protos = {}
    for m in re.finditer(r"\b(?:int|int64_t|void|const char\*|void\*)\s+\*?(klm_[a-z0-9_]+)\s*\(([^;{]*?)\)\s*;", h, flags=re.S):
        args = m.group(2).strip()
        protos[m.group(1)] = 0 if args in ("", "void") else len(re.split(r",(?![^()]*\))", args))
    assert len(protos) >= 60
    go = open(os.path.join(root, "go", "klm", "klm.go")).read()

    def top_level_args(s):
        depth, cur, out = 0, "", []
        for ch in s:
            depth += ch in "([{"
            depth -= ch in ")]}"
            if ch == "," and depth == 0:
                out.append(cur)
                cur = ""
            else:
                cur += ch
        return out + ([cur] if cur.strip() else [])

    bad, seen = [], 0
    for m in re.finditer(r"C\.(klm_[a-z0-9_]+)\(", go):
        i = j = m.end()
        depth = 1
        while depth:
            depth += go[j] == "("
            depth -= go[j] == ")"
            j += 1
        n = len(top_level_args(go[i:j - 1]))
        seen += 1
        if protos.get(m.group(1)) != n:
            bad.append((m.group(1), protos.get(m.group(1)), n))
    assert seen >= 30 and not bad, bad
