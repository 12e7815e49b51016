"""The device's exp / pow / log (csrc/rx_libm.cuh: glibc 2.39's __exp_fma / __pow_fma / __log_fma restated operation by operation
over the tables of the image's libm.so.6) must return the SAME BITS as the host's libm — the functions the CPU reference calls.
Checked over 10^6 arguments each, on the ranges the path uses and beyond.  Bit-exact: integer comparison of the patterns."""
import ctypes as C
import ctypes.util

import numpy as np
import pytest

from rxcase import load_package

pytestmark = pytest.mark.gpu
N = 1_000_000


def _host(fn, *args):
    """glibc's own routine, one call per element (numpy's exp/log/power are its own SIMD code, not libm)."""
    lib = C.CDLL(ctypes.util.find_library("m") or "libm.so.6")
    f = getattr(lib, fn)
    f.restype = C.c_double
    f.argtypes = [C.c_double] * len(args)
    return np.array([f(*a) for a in zip(*[x.tolist() for x in args])])


def _same_bits(a, b):
    return a.view(np.uint64) == b.view(np.uint64)


def test_exp_bits():
    pk = load_package()
    rng = np.random.default_rng(1)
    x = np.concatenate([-rng.uniform(2.0, 110.0, N // 2),          # -Ta/T of the Arrhenius rates (T 200..6000 K)
                        rng.uniform(-700.0, 700.0, N // 4), rng.normal(0.0, 1.0, N // 4), [0.0, -0.0, 1e-300, -1e-20, 708.0, -708.0, 1e-17]])
    got = pk.libm_eval("exp", x); want = _host("exp", x)
    bad = ~_same_bits(got, want)
    assert not bad.any(), (int(bad.sum()), x[bad][:5], got[bad][:5], want[bad][:5])


def test_pow_bits():
    pk = load_package()
    rng = np.random.default_rng(2)
    x = np.concatenate([rng.uniform(200.0, 6000.0, N // 4),                      # T^1.75 of the diffusion coefficients
                        10.0 ** rng.uniform(-30.0, 4.0, N // 4),                 # concentrations^0.5
                        rng.uniform(0.05, 30.0, N // 4),                         # (1/rho)^gamma, (P/s)^(1/gamma) of the outlet state
                        10.0 ** rng.uniform(-200.0, 200.0, N // 4)])
    y = np.concatenate([np.full(N // 4, 1.75), np.full(N // 4, 0.5), rng.uniform(0.6, 1.7, N // 4), rng.uniform(-1.5, 1.5, N // 4)])
    got = pk.libm_eval("pow", x, y); want = _host("pow", x, y)
    bad = ~_same_bits(got, want)
    assert not bad.any(), (int(bad.sum()), x[bad][:5], y[bad][:5], got[bad][:5], want[bad][:5])


def test_log_bits():
    pk = load_package()
    rng = np.random.default_rng(3)
    x = np.concatenate([rng.uniform(1e-3, 1.0, N // 4), rng.uniform(0.93, 1.07, N // 4), 10.0 ** rng.uniform(-300.0, 300.0, N // 4),
                        rng.uniform(1.0, 1e4, N // 4), [1.0, 0.9375, 1.0 + 2.0 ** -52]])
    got = pk.libm_eval("log", x); want = _host("log", x)
    bad = ~_same_bits(got, want)
    assert not bad.any(), (int(bad.sum()), x[bad][:5], got[bad][:5], want[bad][:5])
