"""The kernels' per-point / per-edge arithmetic WITHOUT a GPU: csrc/rx_device.cuh and csrc/rx_flux.cuh define their functions with RX_HD (plain
`inline` outside nvcc), so tests/host_emu.cpp compiles the very source the kernels run — cons2prim (temperature solve, bisection fallback,
non-physical retry), the Wilke transport rules, the AUSM+-up flux, the Arrhenius + PaSR production term — for the host, and this file checks it
against the reference's golden dumps bit for bit.  On the host exp / pow / log are glibc's (the bits rx_libm.cuh reproduces on the device) and
divisions are IEEE (what the device's fast paths equal), so equality here plus the device-only tests of those two pieces (tests/test_gpu_libm.py,
the stage parity of tests/test_gpu_parity.py) is the chain that pins the CUDA path.  What this file cannot see: the kernels' own layout, the
two-lanes-per-edge viscous kernel and the batched transport kernel (device-only forms of the same arithmetic) — those are `-m gpu` tests."""
import ctypes as C
import os
import shutil
import subprocess

import numpy as np
import pytest

import rxcase
from rxcase import Case, GOLDEN, ROOT, load_package, chem_arrays, opts_dict

PKG = os.path.join(ROOT, "development-of-a-turbulent-numerical-solver-for-reactive-flows-in-su2_b200")


@pytest.fixture(scope="module")
def emu(tmp_path_factory):
    if shutil.which("g++") is None:
        pytest.skip("no g++")
    so = str(tmp_path_factory.mktemp("host_emu") / "libhost_emu.so")
    # RX_EMU_UBSAN=1 (set by test_kernel_arithmetic_is_free_of_undefined_behaviour, which runs this file once more in a child process with libubsan
    # preloaded): the same source under UndefinedBehaviorSanitizer, any report aborts
    san = ["-O1", "-g", "-fsanitize=undefined", "-fno-sanitize-recover=undefined"] if os.environ.get("RX_EMU_UBSAN") == "1" else ["-O2"]
    # -ffp-contract=off: no fused multiply-add, as the CUDA build (-fmad=false) and the reference's own x86-64 build
    cc = subprocess.run(["g++", "-std=c++17"] + san + ["-x", "c++", "-shared", "-fPIC", "-ffp-contract=off", "-I", os.path.join(ROOT, "include"),
                         "-I", os.path.join(PKG, "csrc"), os.path.join(ROOT, "tests", "host_emu.cpp"), "-o", so],
                        stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert cc.returncode == 0, cc.stdout[-3000:]
    return C.CDLL(so)


def _structs(pk, case, keep):
    import importlib
    rs = importlib.import_module(pk.__name__ + ".rx_solver")
    ch = rs._chem_struct(chem_arrays(case), keep)
    od = dict(secondOrder=0, limiter=0, limCoeff=0.3, refElemLength=0.1, timeStepping=0, unstCfl=0.0, deltaUnstTime=0.0, dualTime=0, deltaUnstTimeND=0.0,
              viscLimiter=0, device=0)
    od.update(opts_dict(case))
    return ch, rs._Opts(**{k: od[k] for k, _ in rs._Opts._fields_})


_d = lambda a: a.ctypes.data_as(C.POINTER(C.c_double))


@pytest.mark.parametrize("name", ["channel_lam_rk", "channel_rans", "channel_euler", "channel_lam_edge", "channel_lam_clip", "channel_rans_ign",
                                  "channel_air3_lam", "channel_mix5_lam", "jet_rans_perturbed"])
def test_cons2prim_and_transport_source_equal_the_reference_dumps(emu, name):
    """cons2prim<NS> exactly as k_primitive calls it, on the conserved state, secant start value and k of every dumped iteration, against the
    reference's primitive dump (T, u, v, P, rho, h, a, Y_s) — nine-, five- and three-species mixtures, the bisection fallback and the
    Solution_Old retry of the edge-case fixtures, the ignition overwrite; then transport<NS> against the dumped viscosity and conductivity."""
    pk = load_package()
    case = Case("%s/%s.npz" % (GOLDEN, name))
    keep = []
    ch, op = _structs(pk, case, keep)
    Ns = case.Ns
    its = sorted({int(k[2:6]) for k in case.arr if k.startswith("it")})
    checked = 0
    for it in its:
        U = np.ascontiguousarray(case.stage(it, 0, "U0")); T0 = np.ascontiguousarray(case.stage(it, 0, "prev_V")[:, 0])
        ke = np.ascontiguousarray(case.stage(it, 0, "turb")[:, 1])
        Uold = np.ascontiguousarray(case.stage(it, 0, "Uold")) if case.has("it%04d_rk0_Uold" % it) else None
        n = U.shape[0]
        V = np.empty((n, 7 + Ns)); bad = np.zeros(n, dtype=np.int32)
        fails = emu.emu_cons2prim(C.byref(ch), C.byref(op), C.c_int64(it), C.c_int64(n), _d(U), _d(Uold) if Uold is not None else None, _d(T0), _d(ke), _d(V),
                                  bad.ctypes.data_as(C.POINTER(C.c_int32)))
        want = case.stage(it, 0, "pre_V")
        assert fails == 0
        assert np.array_equal(V, want), "%s it %d: cons2prim differs at %d entries, worst %.3e" % (name, it, int((V != want).sum()), float(np.abs(V - want).max()))
        if case.meta["cfg"]["viscous"]:
            aux = case.stage(it, 0, "pre_aux")
            mu = np.empty(n); ka = np.empty(n)
            Tn = np.ascontiguousarray(V[:, 0]); Y = np.ascontiguousarray(V[:, 7:])
            if case.meta["cfg"]["ignition"]:   # the transport properties are those of the temperature SetPrimVar found, before the ignition overwrite
                op2 = type(op).from_buffer_copy(op); op2.ignition = 0   # (variable_direct_reactive.cpp:1193-1228 runs inside SetPrimVar; the overwrite comes after, :1013-1024)
                V2 = np.empty_like(V)
                emu.emu_cons2prim(C.byref(ch), C.byref(op2), C.c_int64(it), C.c_int64(n), _d(U), _d(Uold) if Uold is not None else None, _d(T0), _d(ke), _d(V2), None)
                Tn = np.ascontiguousarray(V2[:, 0])
            assert emu.emu_transport(C.byref(ch), C.c_int64(n), _d(Tn), _d(Y), _d(mu), _d(ka)) == 0
            assert np.array_equal(mu, aux[:, 2]) and np.array_equal(ka, aux[:, 3]), "%s it %d: transport differs" % (name, it)
        checked += 1
    assert checked >= 1


@pytest.mark.parametrize("name", ["channel_lam_rk", "channel_euler", "channel_lam_gg"])
def test_ausm_source_equals_the_reference_convective_residual(emu, name):
    """ausm<NS> on every interior edge of the dumped primitive state, accumulated at the two end points in edge order as the reference's edge loop
    does (+ at node 0, - at node 1), against the dumped residual after Upwind_Residual — first-order fixtures that keep the intermediate residuals."""
    case = Case("%s/%s.npz" % (GOLDEN, name))
    assert not case.meta["cfg"].get("second_order", 0)
    Ns = case.Ns
    en = case["geo_edge_nodes"].astype(np.int64); nrm = np.ascontiguousarray(case["geo_edge_normal"])
    its = [it for it in sorted({int(k[2:6]) for k in case.arr if k.startswith("it")}) if case.has("it%04d_rk0_res_conv" % it)]
    assert its
    for it in its:
        V = case.stage(it, 0, "pre_V")
        Vi = np.ascontiguousarray(V[en[:, 0]]); Vj = np.ascontiguousarray(V[en[:, 1]])
        F = np.empty((en.shape[0], 4 + Ns))
        assert emu.emu_ausm(Ns, C.c_double(case.meta["cfg"]["mach"]), C.c_int64(en.shape[0]), _d(Vi), _d(Vj), _d(nrm), _d(F)) == 0
        R = np.zeros((case.nPoint, 4 + Ns))
        for e in range(en.shape[0]):          # sequential, in edge order: the order of the sums is part of the result
            R[en[e, 0]] += F[e]; R[en[e, 1]] -= F[e]
        want = case.stage(it, 0, "res_conv")
        nD = case.nPointDomain
        assert np.array_equal(R[:nD], want[:nD]), "%s it %d: worst %.3e" % (name, it, float(np.abs(R[:nD] - want[:nD]).max()))


@pytest.mark.parametrize("name", ["channel_lam_rk", "channel_lam_gg", "channel_rans", "channel_air3_lam"])
def test_viscous_flux_source_equals_the_reference_viscous_residual(emu, name):
    """viscous_flux<NS, EDGE = true> — mean gradients with the edge correction, Stefan-Maxwell diffusion fluxes through the restatement of
    Eigen's BiCGSTAB, for RANS the SST closure through the restatement of its pivoted Householder QR — on every interior edge of the dumped
    state, with 1/D_st formed from the dumped T and P as k_transport forms it, subtracted at node 0 and added at node 1 in edge order on top of
    the dumped convective residual: the dumped residual after Viscous_Residual, bit for bit (nine species laminar and RANS, three species)."""
    pk = load_package()
    case = Case("%s/%s.npz" % (GOLDEN, name))
    keep = []
    ch, op = _structs(pk, case, keep)
    Ns = case.Ns
    en = np.ascontiguousarray(case["geo_edge_nodes"].astype(np.int64)); nrm = np.ascontiguousarray(case["geo_edge_normal"])
    xy = np.ascontiguousarray(case["geo_coord"])
    its = [it for it in sorted({int(k[2:6]) for k in case.arr if k.startswith("it")}) if case.has("it%04d_rk0_res_visc" % it) and case.has("it%04d_rk0_res_conv" % it)]
    assert its
    for it in its:
        V = np.ascontiguousarray(case.stage(it, 0, "pre_V")); g = np.ascontiguousarray(case.stage(it, 0, "pre_grad")); aux = case.stage(it, 0, "pre_aux")
        mu = np.ascontiguousarray(aux[:, 2]); ka = np.ascontiguousarray(aux[:, 3]); tb = np.ascontiguousarray(case.stage(it, 0, "turb"))
        F = np.empty((en.shape[0], 4 + Ns))
        rc = emu.emu_viscous(C.byref(ch), C.byref(op), C.c_int64(en.shape[0]), en.ctypes.data_as(C.POINTER(C.c_int64)), _d(nrm), C.c_int64(case.nPoint), _d(V), _d(g),
                             _d(mu), _d(ka), _d(xy), _d(tb), _d(F))
        assert rc == 0
        R = case.stage(it, 0, "res_conv").copy()
        for e in range(en.shape[0]):
            R[en[e, 0]] -= F[e]; R[en[e, 1]] += F[e]
        want = case.stage(it, 0, "res_visc")
        nD = case.nPointDomain
        assert np.array_equal(R[:nD], want[:nD]), "%s it %d: %d entries differ, worst %.3e" % (name, it, int((R[:nD] != want[:nD]).sum()), float(np.abs(R[:nD] - want[:nD]).max()))


@pytest.mark.parametrize("rans", [False, True])
def test_mass_production_source_equals_the_restatement_and_the_reference_library(emu, rans):
    """mass_production<9> scaled as k_source_points scales it, on the states of the source-term fixture: bit-equal to the CPU restatement
    (rxo_source_points) and within the GPU test's tolerance of the reference's own ReactingModelLibrary (tests/golden/source_points.npz)."""
    import importlib
    pk = load_package()
    rs = importlib.import_module(pk.__name__ + ".rx_solver")
    chem = pk.load_mechanism()
    g = np.load(GOLDEN + "/source_points.npz")
    keep = []
    ch = rs._chem_struct(chem, keep)
    T, rho, Y, vol, omT = [np.ascontiguousarray(g[k], dtype=np.float64) for k in ("T", "rho", "Y", "vol", "omegaTurb")]
    n = T.shape[0]
    res = np.empty((n, 9))
    assert emu.emu_source(C.byref(ch), int(rans), C.c_double(float(g["cMu"])), C.c_double(float(g["pasrLb"])), C.c_int64(n), _d(T), _d(rho), _d(Y), _d(omT), _d(vol), _d(res)) == 0
    from test_gpu_source import _oracle_source
    want = _oracle_source(chem, rans, float(g["cMu"]), float(g["pasrLb"]), T, rho, Y, omT, vol)
    assert np.array_equal(res, want), "worst %.3e" % float(np.abs(res - want).max())
    lib = g["res_pasr" if rans else "res_laminar"]
    scale = np.abs(lib).max(axis=0)
    assert (np.abs(res - lib).max(axis=0) / np.where(scale > 0, scale, 1.0)).max() < 1e-12
    assert np.array_equal(res == 0.0, lib == 0.0)


def test_pivoted_qr_source_equals_the_eigen_pinned_restatement(emu):
    """colpiv_qr_solve<9, 2, 1> as the serial device form calls it, on the systems of the QR hooks (the SST closure's operator for random
    compositions with trace species, dense matrices over several decades, rank-deficient ones, zero rows / columns, the zero matrix): every
    entry equal to the oracle's restatement, which oracle/_ref/eigen_qr_check pins to Eigen 3.3.7's ColPivHouseholderQR bit for bit."""
    from test_oracle_golden import qr_test_systems, oracle_qr
    A, b = qr_test_systems(600)
    want = oracle_qr(A, b)
    Acm = np.ascontiguousarray(np.transpose(A, (0, 2, 1))).reshape(A.shape[0], 81)   # column-major, as the device form takes it
    b = np.ascontiguousarray(b)
    x = np.empty_like(b)
    assert emu.emu_qr(C.c_int64(A.shape[0]), _d(Acm), _d(b), _d(x)) == 0
    same = (x == want) | (np.isnan(x) & np.isnan(want))
    assert same.all(), "%d entries differ" % int((~same).sum())


def test_kernel_arithmetic_is_free_of_undefined_behaviour():
    """The tests above once more in a child process, with tests/host_emu.cpp built under UndefinedBehaviorSanitizer (array bounds of the unrolled
    per-species arrays, shifts, integer overflow, misaligned or null accesses): no report on any fixture."""
    if os.environ.get("RX_EMU_UBSAN") == "1":
        pytest.skip("this is the child run")
    if shutil.which("g++") is None:
        pytest.skip("no g++")
    ub = subprocess.run(["g++", "-print-file-name=libubsan.so"], stdout=subprocess.PIPE, text=True).stdout.strip()
    if not os.path.isabs(ub) or not os.path.exists(ub):
        pytest.skip("this g++ has no UBSan runtime")
    import sys
    env = dict(os.environ, RX_EMU_UBSAN="1", LD_PRELOAD=ub, UBSAN_OPTIONS="print_stacktrace=1")
    run = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-x", "-q", "-p", "no:cacheprovider"], cwd=os.path.dirname(os.path.abspath(__file__)),
                         env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=1200)
    assert run.returncode == 0 and "runtime error" not in run.stdout and " passed" in run.stdout, run.stdout[-4000:]
