"""CPU suite: the C-ABI library builds for sm_100a, loads, and exports every symbol include/rx_cuda.h declares.
No compute call is made here (no GPU in this container); on a box without CUDA devices rx_create must fail loudly."""
import ctypes as C
import os
import re

import pytest

from rxcase import ROOT, load_package


def declared_symbols():
    txt = open(os.path.join(ROOT, "include", "rx_cuda.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(rx_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    pk = load_package()
    pk.build_library()
    lib = pk.load_library()
    syms = declared_symbols()
    assert len(syms) >= 25
    for s in syms:
        assert hasattr(lib, s), "librx_cuda.so does not export %s" % s


def test_python_mirror_lists_the_same_symbols():
    pk = load_package()
    from importlib import import_module
    rs = import_module(pk.__name__ + ".rx_solver")
    assert sorted(rs.EXPORTS) == declared_symbols()


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    pk = load_package()
    lib = pk.load_library()
    assert [lib.rx_supported_species(n) for n in (9, 5, 3, 6, 1)] == [1, 1, 1, 0, 0]   # the mixtures the reference ships
    from rxcase import Case, GOLDEN, gpu_solver
    case = Case(os.path.join(GOLDEN, "channel_lam_gg.npz"))
    with pytest.raises(pk.RxError) as e:
        gpu_solver(case)
    assert "no CUDA device" in str(e.value)
