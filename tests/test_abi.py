"""The C-ABI library loads and exports exactly what include/b200p.h declares (no compute without a GPU)."""
import ctypes
import os
import re

import pytest

from mlx_parallm_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    src = open(os.path.join(ROOT, "include", "b200p.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200p_[a-z0-9_]+)\s*\(", src)))


def test_library_is_built_in_tree():
    assert os.path.exists(_lib.LIB_PATH), "run `make -C mlx_parallm_b200/csrc` (or __graft_entry__.build())"


def test_every_declared_symbol_is_exported():
    L = ctypes.CDLL(_lib.LIB_PATH)
    fns = header_functions()
    assert len(fns) >= 15
    missing = [f for f in fns if not hasattr(L, f)]
    assert not missing, missing


def test_python_binding_covers_the_header():
    assert sorted(_lib.SIGNATURES) == header_functions()


def test_version_and_error_plumbing():
    L = _lib.lib()
    assert b"sm_100a" in L.b200p_version()
    assert L.b200p_qmm_set_path(7) == -1                 # B200P_ERR_ARG, no device needed
    assert "path" in _lib.last_error()
    L.b200p_qmm_set_path(0)


def test_no_cpu_fallback():
    """Product ops refuse CPU tensors instead of computing somewhere else."""
    import torch
    from mlx_parallm_b200 import ops
    w = torch.zeros((2, 8), dtype=torch.int32)
    s = torch.zeros((2, 1), dtype=torch.float16)
    with pytest.raises(_lib.B200PError):
        ops.dequantize(w, s, s)


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "mlx_parallm_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dp, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), os.path.join(dp, f)


def test_environment_switches_are_documented_and_exist():
    """INTEGRATION.md's table of B200P_* environment switches against the sources: everything the library, the package, the
    bench or the tests read is documented, and nothing documented has gone away."""
    import glob
    import re
    root = ROOT if "ROOT" in globals() else os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    read = set()
    for f in glob.glob(os.path.join(root, "mlx_parallm_b200", "csrc", "*.cu*")):
        read |= set(re.findall(r'(?:getenv|env_int)\("(B200P_[A-Z0-9_]+)"', open(f).read()))
    py = glob.glob(os.path.join(root, "mlx_parallm_b200", "**", "*.py"), recursive=True) + [os.path.join(root, "bench.py")] + \
        glob.glob(os.path.join(root, "tests", "*.py"))
    for f in py:
        read |= set(re.findall(r'environ(?:\.get)?[\(\[]\s*"(B200P_[A-Z0-9_]+)"', open(f).read()))
    table = open(os.path.join(root, "INTEGRATION.md")).read().split("## Environment switches", 1)[1]
    documented = set(re.findall(r"`(B200P_[A-Z0-9_]+)`", "\n".join(l.split("|")[1] for l in table.splitlines() if l.startswith("| `B200P_"))))
    assert read - documented == set(), f"read but not documented in INTEGRATION.md: {sorted(read - documented)}"
    assert documented - read == set(), f"documented but never read: {sorted(documented - read)}"
