"""The C-ABI library loads and exports every symbol include/cfd_b200.h declares (no compute without a GPU)."""
import ctypes as C
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    txt = open(os.path.join(ROOT, "include", "cfd_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(cfd_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol(cfd):
    from cfd_b200 import _lib
    lib = cfd.load()
    names = declared_functions()
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/cfd_b200.h but not exported"
        assert n in _lib.SYMBOLS, f"{n} has no ctypes prototype in cfd_b200/_lib.py"
    assert set(_lib.SYMBOLS) <= set(names)
    assert b"sm_100a" in lib.cfd_version()


def test_struct_layout_matches_header(cfd):
    from cfd_b200 import _lib
    # 5 ints + 8 ints + (pad) 8 doubles + 8 ints + double = natural C layout
    assert C.sizeof(_lib.CfdConfig) == 5 * 4 + 8 * 4 + 4 + 8 * 8 + 8 * 4 + 8 + 2 * 4 + 3 * 4 + 4 + 3 * 8 + 4 * 4 + 8 + 8
    assert C.sizeof(_lib.CfdHeatConfig) == 8 + 14 * 8 + 5 * 4 + 4


def test_no_cpu_fallback(cfd):
    """Without a CUDA device context creation must fail loudly, never compute on the host."""
    import torch
    if torch.cuda.is_available():
        return
    import pytest
    with pytest.raises(cfd.CfdError, match="no CUDA device"):
        cfd.CavityFVM(10, 10)
    with pytest.raises(cfd.CfdError, match="no CUDA device"):
        cfd.Heat2D(8, 8)


def test_product_never_imports_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "cfd_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "from oracle" not in src and "liboracle" not in src, f


def test_host_setgrid_matches_oracle_bitwise(cfd, oracle):
    import numpy as np
    g = cfd.setgrid(37, 11, 0.0, 1.0, -0.5, 2.0)
    o = oracle.Cavity(37, 11, 100.0, xlim=(0.0, 1.0), ylim=(-0.5, 2.0))
    for k, v in g.items():
        assert np.array_equal(v, getattr(o, k)), k
