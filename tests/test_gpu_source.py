"""Stand-alone Arrhenius (+ PaSR) source over independent points (rx_source_points / rx_source_points_dev)
against the CPU oracle (rxo_source_points, the restatement of CSourceReactive::ComputeChemistry,
numerics_direct_reactive.cpp:1728-1825 and ReactingModelLibrary :99-227, 803-920) on the seeded states of the source-term
microbenchmark: pure-oxidiser / pure-fuel points exercise the 1e-30 / 1e-15 / 1e-10 thresholds.  Tolerance 1e-12 of the
largest production rate of a species (exp / pow are libm on one side, CUDA on the other)."""
import ctypes as C

import numpy as np
import pytest

import rxcase
from rxcase import load_package

pytestmark = pytest.mark.gpu


def _oracle_source(chem, rans, cMu, lb, T, rho, Y, omT, vol):
    lib = C.CDLL(rxcase.build_oracle())
    keep = []
    k = lambda a: (keep.append(np.ascontiguousarray(a, dtype=np.float64)), keep[-1])[1]
    Ns, Nr = chem["Ns"], chem["Nr"]
    ch = rxcase._Chem(Ns, Nr, chem["nKnots"], *[rxcase._p(k(np.asarray(chem[n]).reshape(-1))) for n in
                     ["mMasses", "diffVol", "x", "y", "y2", "As", "Betas", "Ta", "reversible", "availBack", "AsB", "BetasB", "TaB",
                      "nuR", "nuP", "expR", "expP"]])
    res = np.empty((len(T), Ns))
    lib.rxo_source_points.restype = None
    lib.rxo_source_points(C.byref(ch), int(rans), C.c_double(cMu), C.c_double(lb), C.c_int64(len(T)), rxcase._p(k(T)), rxcase._p(k(rho)),
                          rxcase._p(k(Y)), rxcase._p(k(omT)), rxcase._p(k(vol)), rxcase._p(res))
    return res


@pytest.mark.parametrize("rans", [False, True])
def test_source_points_match_the_reference_library(rans):
    """Against the reference's own Framework::ReactingModelLibrary driven directly (oracle/ref_source_harness.cpp, fixture
    tests/golden/source_points.npz by oracle/make_source_golden.py): SetSourceTerm -> Set_DfrDrhos -> AssemblePaSRConstant ->
    GetMassProductionTerm (reacting_model_library.cpp:99-227) on the microbenchmark's own states (T up to 2500 K, 30 % pure
    oxidiser, 10 % pure fuel).  1e-12 of the largest production rate of a species, and 1e-11 point-wise with the 1e-12 floor."""
    pk = load_package()
    chem = pk.load_mechanism()
    g = np.load(rxcase.GOLDEN + "/source_points.npz")
    want = g["res_pasr" if rans else "res_laminar"]
    got = pk.source_points(chem, g["T"], g["rho"], g["Y"], g["vol"], omegaTurb=g["omegaTurb"], rans=rans, cMu=float(g["cMu"]), pasrLb=float(g["pasrLb"]))
    scale = np.abs(want).max(axis=0)
    err = np.abs(got - want).max(axis=0) / np.where(scale > 0, scale, 1.0)
    pw = rxcase.pointwise_rel_err(got, want)
    print("source vs reference library (rans=%s): column-relative %.3e, point-wise %.3e" % (rans, err.max(), pw.max()))
    assert err.max() < 1e-12, err
    assert pw.max() < 1e-11, pw
    assert np.array_equal(got == 0.0, want == 0.0)   # the thresholds (Y < 1e-15, single species) cut at the same points


@pytest.mark.parametrize("rans", [False, True])
@pytest.mark.parametrize("n", [1, 4097])
def test_source_points_host_form_matches_oracle(rans, n):
    pk = load_package()
    chem = pk.load_mechanism()
    T, rho, Y, omT, vol = pk.synthetic_points(chem, n, seed=7 + n)
    got = pk.source_points(chem, T, rho, Y, vol, omegaTurb=omT, rans=rans)
    want = _oracle_source(chem, rans, 0.09, 0.2, T, rho, Y, omT, vol)
    assert np.all(np.isfinite(got))
    scale = np.abs(want).max(axis=0)
    err = np.abs(got - want).max(axis=0) / np.where(scale > 0, scale, 1.0)
    assert err.max() < 1e-12, err
    if n > 1:
        assert np.abs(want).max() > 0       # the sample does react
        pure = (Y.max(axis=1) == 1.0)
        assert pure.any() and np.all(got[pure] == 0.0)   # a single species cannot react: exact zeros on both sides


def test_source_points_device_form_equals_host_form():
    import torch
    pk = load_package()
    chem = pk.load_mechanism()
    n = 10000
    T, rho, Y, omT, vol = pk.synthetic_points(chem, n, seed=3)
    opts = pk.default_opts(rans=True)
    host = pk.source_points(chem, T, rho, Y, vol, omegaTurb=omT, rans=True, cMu=opts["cMu"], pasrLb=opts["pasrLb"])
    s = pk.CReactiveNSSolverGPU(pk.channel_mesh(5, 4), chem, opts)
    dev = torch.device("cuda", 0)
    t = lambda a: torch.as_tensor(np.ascontiguousarray(a), device=dev)
    dT, dr, dY, dO, dV = t(T), t(rho), t(Y.T), t(omT), t(vol)      # Y as SoA [Ns][n]
    dR = torch.empty((chem["Ns"], n), dtype=torch.float64, device=dev)
    s.SourcePointsDev(n, dT.data_ptr(), dr.data_ptr(), dY.data_ptr(), dO.data_ptr(), dV.data_ptr(), dR.data_ptr())
    s.Synchronize()
    np.testing.assert_array_equal(dR.cpu().numpy().T, host)
    s.close()
