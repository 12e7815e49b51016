"""CPU suite: the oracle (oracle/rx_oracle.cpp) against golden dumps of the reference's own classes.

Golden vectors: tests/golden/*.npz, produced by oracle/make_golden.py from oracle/_ref/ref_harness (the reference's
CReactiveNSSolver / CUpwReactiveAUSM / CAvgGradReactive_* / CSourceReactive / ReactingModelLibrary objects driven in the
order of CMultiGridIntegration::MultiGrid_Cycle, integration_time.cpp:143-181).
Bar: BIT-IDENTICAL at every stage of every case (TOL = 0), RANS included: the exp/pow/log of the image's glibc are what the reference
itself calls, and every sum of the two Eigen solvers on the path (BiCGSTAB, pivoted Householder QR) is taken in Eigen 3.3.7's own
order (oracle/eigen_qr_check.cpp pins the QR against the real headers).
"""
import re

import numpy as np
import pytest

from rxcase import Case, Oracle, GOLDEN, vec_rel_err, stage_rel_err

CASES = ["jet_rans_shipped", "jet_rans_perturbed", "channel_lam_rk", "channel_lam_gg", "channel_rans", "channel_lam_walls", "channel_lam_muscl", "channel_lam_bj", "channel_euler", "channel_rans_gg",
         "channel_lam_edge", "channel_lam_clip", "channel_rans_ign",
         # species-count generality: the reference's other two mixtures (TURBOLENT_FLAT_PLATE): 3-species air laminar / RANS, 5 species MUSCL
         "channel_air3_lam", "channel_air3_rans", "channel_mix5_lam",
         # the other two kinds of BC_Inlet (INLET_TYPE= TOTAL_CONDITIONS / MASS_FLOW, solver_direct_reactive.cpp:3283-3465)
         "channel_lam_inlet_total", "channel_rans_inlet_massflow",
         # BC_Supersonic_Inlet / BC_Supersonic_Outlet (:2998-3214, :3681-3807)
         "channel_lam_supersonic",
         # UNSTEADY_SIMULATION= TIME_STEPPING (global minimum time step; the Euler solver's CFL swap and fixed UNST_TIMESTEP)
         "channel_lam_unsteady", "channel_euler_unsteady", "channel_euler_unsteady_fixed",
         # explicit dual time stepping (SetResidual_DualTime :2172-2215, the pseudo-time step clamp :5274-5297, the level shift)
         "channel_lam_dualtime2", "channel_euler_dualtime1", "channel_rans_dualtime2",
         # VISCOUS_LIMITER_FLOW= YES (limited mean gradients in the interior viscous flux, numerics_direct_reactive.cpp:1482-1505)
         "channel_lam_visclim"]
TOL = 0.0


def drive(case, solver, check):
    """Run `solver` (Oracle-like API) through every dumped iteration of `case`, calling check(stage_key, array)."""
    cf = case.meta["cfg"]
    its = sorted({int(re.match(r"it(\d+)_", k).group(1)) for k in case.arr if k.startswith("it")})
    nrk = cf["nRK"] if cf["time_scheme"] == 1 else 1
    dual, prevUn = cf.get("dual_time", 0), None
    for it in its:
        if dual:   # dual time stepping: the stored time levels; where the reference closed a physical time step, shift them as it does
            Un, Un1 = case.stage(it, 0, "Un"), case.stage(it, 0, "Un1")
            if prevUn is not None and not np.array_equal(Un, prevUn):
                solver.shift_time_levels()
                a, b = solver.time_levels()
                check("it%04d_rk0_Un" % it, a); check("it%04d_rk0_Un1" % it, b)
            solver.set_time_levels(Un, Un1); solver.set_inner_iteration(it % cf.get("dual_inner", 2))
            prevUn = Un
        if case.has("it%04d_rk0_Uold" % it):   # Solution_Old at the start of the iteration: the non-physical fallback reads it
            solver.set_solution(case.stage(it, 0, "Uold")); solver.set_old_solution()
        solver.set_solution(case.stage(it, 0, "U0"))
        solver.set_temperature(case.stage(it, 0, "prev_V")[:, 0])
        solver.set_turbulence(case.stage(it, 0, "turb"))
        for rk in range(nrk):
            key = lambda w: "it%04d_rk%d_%s" % (it, rk, w)
            solver.preprocess(it, False)
            V, aux, g = solver.primitive()
            check(key("pre_V"), V); check(key("pre_aux"), aux)
            if cf["viscous"]:
                check(key("pre_grad"), g)
            if cf.get("limiter", 0) and hasattr(solver, "limiter"):
                check(key("pre_lim"), solver.limiter())
            if rk == 0:
                solver.set_old_solution(); solver.time_step(); check(key("dt"), solver.dt())
            solver.upwind_residual(); check(key("res_conv"), solver.residual())
            solver.viscous_residual(); check(key("res_visc"), solver.residual())
            solver.source_residual(); check(key("res_src"), solver.residual())
            if dual:
                solver.dual_time_residual(); check(key("res_dual"), solver.residual())
            solver.boundary_conditions(); check(key("res_bc"), solver.residual())
            rms = solver.explicit_update(cf["rk_alpha"][rk] if nrk > 1 else 1.0)
            check(key("U1"), solver.solution()); check(key("rms"), rms.reshape(1, -1))
        solver.preprocess(it, True)
        V, aux, g = solver.primitive()
        check("it%04d_rk0_out_V" % it, V); check("it%04d_rk0_out_aux" % it, aux)


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_dumps(name):
    case = Case("%s/%s.npz" % (GOLDEN, name))
    seen = []

    def check(key, a):
        if not case.has(key):
            return
        ref = case[key].reshape(a.shape)
        err = stage_rel_err(key, a, ref, case.Ns).max()
        seen.append(key)
        assert err <= TOL, "%s: %s vec-rel err %.3e" % (name, key, err)

    drive(case, Oracle(case), check)
    assert len(seen) >= 7


def test_bicgstab_restatement_zero_rhs_and_identity():
    import ctypes as C
    from rxcase import build_oracle
    lib = C.CDLL(build_oracle())
    n = 9
    A = np.asfortranarray(np.eye(n) * 2.0)
    x = np.empty(n); it = C.c_int(-1)
    rhs = np.zeros(n)
    P = lambda a: a.ctypes.data_as(C.POINTER(C.c_double))
    lib.rxo_bicgstab(n, P(A), P(rhs), P(x), C.byref(it))
    assert it.value == 0 and np.all(x == 0)          # rhs == 0 => x = 0 (BiCGSTAB.h:45-49)
    rhs = np.arange(1.0, n + 1)
    lib.rxo_bicgstab(n, P(A), P(rhs), P(x), C.byref(it))
    assert it.value == 1 and np.allclose(x, rhs / 2.0, rtol=1e-15)


@pytest.mark.parametrize("name", ["channel_rans", "channel_rans_gg", "channel_rans_ign", "jet_rans_shipped", "jet_rans_perturbed",
                                  "channel_air3_rans", "channel_rans_inlet_massflow", "channel_rans_dualtime2"])
def test_oracle_sst_transport_step_matches_reference(name):
    """rxo_turb_iterate — the restatement of CTurbSSTSolver's SingleGrid_Iteration (residuals, boundary conditions, FGMRES + LU-SGS in the
    reference's own order of operations, clipped update, Postprocessing) — against the harness dumps of the reference's turbulence solver
    after every dumped flow iteration: tres (LinSysRes before the solve), tdx (LinSysSol), tsol ((k, omega) after the update), trms, tpost."""
    case = Case("%s/%s.npz" % (GOLDEN, name))
    cf = case.meta["cfg"]
    nrk = cf["nRK"] if cf["time_scheme"] == 1 else 1
    its = sorted({int(k[2:6]) for k in case.arr if k.startswith("it") and k.endswith("_tres")})
    assert its
    o = Oracle(case)
    worst = {}
    for it in its:
        tpost = case.stage(it, 0, "tpost")
        o.set_solution(case.stage(it, 0, "U0")); o.set_temperature(case.stage(it, 0, "prev_V")[:, 0]); o.set_turbulence(case.stage(it, 0, "turb"))
        dual = cf.get("dual_time", 0)
        if dual:   # the flow solver's and the SST solver's stored time levels (the latter's dual-time source term reads both, :762-780)
            o.set_time_levels(case.stage(it, 0, "Un"), case.stage(it, 0, "Un1")); o.set_inner_iteration(it % cf.get("dual_inner", 2))
            o.turb_set_time_levels(case.stage(it, 0, "tUn"), case.stage(it, 0, "tUn1"))
        for rk in range(nrk):
            o.preprocess(it, False)
            if rk == 0:
                o.set_old_solution(); o.time_step()
            o.upwind_residual(); o.viscous_residual(); o.source_residual()
            if dual:
                o.dual_time_residual()
            o.boundary_conditions()
            o.explicit_update(cf["rk_alpha"][rk] if nrk > 1 else 1.0)
        o.preprocess(it, True)
        r = o.turb_iterate(case.stage(it, 0, "turb")[:, 1:3], case.stage(it, 0, "tblend"), tpost[:, 2], float(tpost[0, 11]), cf["kine_inf"], cf["omega_inf"],
                           cf["cfl_red_turb"], cf["relax_turb"], cf["lin_iter"], cf["lin_tol"])
        nD = case.nPointDomain
        def col(a, ref):
            sc = np.abs(ref).max(axis=0)
            return float((np.abs(a - ref).max(axis=0) / np.where(sc > 0, sc, 1.0)).max())
        e = {"tres": col(r["res"][:nD], case.stage(it, 0, "tres")[:nD]), "tdx": col(r["dx"][:nD], case.stage(it, 0, "tdx")[:nD]),
             "tsol": col(r["post"][:nD, 0:2], case.stage(it, 0, "tsol")[:nD]), "trms": float(np.abs(r["rms"] / case.stage(it, 0, "trms") - 1.0).max()),
             "tpost": col(r["post"][:nD], tpost[:nD])}
        for k, v in e.items():
            worst[k] = max(worst.get(k, 0.0), v)
    print(name, worst)
    assert all(v == 0.0 for v in worst.values()), worst   # the restatement follows the reference's order of operations: identical bits


def test_qr_restatement_equals_eigen_bit_for_bit():
    """oracle/_ref/eigen_qr_check (built by oracle/build_ref.sh against the Eigen headers the reference vendors): the oracle's
    colPivQRSolve against Eigen::ColPivHouseholderQR<MatrixXd>::solve on the SST closure's operator for random compositions, dense
    random matrices of orders 1..9 and rank-deficient / zero matrices — every solution entry identical."""
    import os
    import subprocess
    from rxcase import ROOT
    exe = os.path.join(ROOT, "oracle", "_ref", "eigen_qr_check")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/eigen_qr_check is not in this snapshot")
    r = subprocess.run([exe, "40000"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0 and "mismatched_entries=0 " in r.stdout and "eigen=3.3.7" in r.stdout, r.stdout


def qr_test_systems(n, seed=7):
    """9 x 9 systems for the QR hooks: the closure's operator (numerics_direct_reactive.cpp:860-880) for random compositions with
    trace species, dense random matrices over several decades, rank-deficient ones, zero columns / rows, the zero matrix."""
    rng = np.random.default_rng(seed)
    Mm = np.array([2.016, 31.998, 1.008, 15.999, 17.007, 33.006, 34.014, 18.015, 28.014])
    A = np.empty((n, 9, 9)); b = (rng.random((n, 2, 9)) - 0.5) * 10.0 ** (6 * rng.random((n, 2, 9)) - 3)
    for q in range(n):
        kind = q % 4
        if kind < 2:
            Y = rng.random(9) * 10.0 ** (-12 * rng.random(9) * (rng.random(9) < 0.5)); Y /= Y.sum()
            X = Y / Mm; X /= X.sum()
            mt = Mm.sum(); sig = X.sum()
            M = mt * (Y[:, None] / Mm[:, None] - X[:, None] / Mm[None, :])
            M[np.arange(9), np.arange(9)] = mt / Mm * (Y - X + sig)
        elif kind == 2:
            M = (rng.random((9, 9)) - 0.5) * 10.0 ** (8 * rng.random((9, 9)) - 4)
        else:
            r = (q // 4) % 9
            M = (rng.random((9, max(r, 1))) - 0.5) @ (rng.random((max(r, 1), 9)) - 0.5) if r else np.zeros((9, 9))
            if q % 5 == 0:
                M[:, 8] = 0.0
            if q % 7 == 0:
                M[0, :] = 0.0
        A[q] = M
    return A, b


def oracle_qr(A, b):
    import ctypes as C
    from rxcase import build_oracle
    lib = C.CDLL(build_oracle())
    P = lambda a: a.ctypes.data_as(C.POINTER(C.c_double))
    x = np.empty_like(b)
    for q in range(A.shape[0]):
        Af = np.ascontiguousarray(A[q].T)   # column-major
        for d in range(2):
            bb = np.ascontiguousarray(b[q, d]); xx = np.empty(9)
            lib.rxo_colpiv_qr_solve(9, P(Af), P(bb), P(xx))
            x[q, d] = xx
    return x


def test_qr_restatement_solves_the_systems():
    A, b = qr_test_systems(400)
    x = oracle_qr(A, b)
    for q in range(0, 400, 4):   # the closure's operator: well conditioned enough for a residual check
        for d in range(2):
            assert np.abs(A[q] @ x[q, d] - b[q, d]).max() <= 1e-9 * max(1.0, np.abs(b[q, d]).max()) * np.linalg.cond(A[q])


TPOST_COLS = ["k", "omega", "dist", "F1", "F2", "CDkw", "muT", "dk/dx", "dk/dy", "domega/dx", "domega/dy", "sigma_k"]


@pytest.mark.parametrize("name", ["channel_rans", "jet_rans_shipped", "jet_rans_perturbed", "channel_rans_gg"])
def test_oracle_turbulence_postprocessing_matches_reference(name):
    """SURVEY.md 8(f) rank 1, first slice: CTurbSSTSolver::Postprocessing (gradient of k and omega, blending functions, eddy
    viscosity) restated in rxo_turb_postprocess against the reference's own turbulence solver state after its iteration
    (`tpost` dump; flow inputs = the reference's rho, laminar viscosity and strain magnitude, `tpost_flow`)."""
    case = Case("%s/%s.npz" % (GOLDEN, name))
    key = sorted(k for k in case.arr if k.endswith("_tpost"))[-1]
    ref = case[key]
    o = Oracle(case)
    out = o.turb_postprocess(ref[:, 0:2], ref[:, 2], case[key + "_flow"], sigma_k=float(ref[0, 11]))
    nD = case.nPointDomain
    for c, nm in enumerate(TPOST_COLS):
        scale = np.abs(ref[:nD, c]).max()
        err = np.abs(out[:nD, c] - ref[:nD, c]).max() / (scale if scale > 0 else 1.0)
        assert err <= TOL, "%s column %s: %.3e" % (name, nm, err)


@pytest.mark.parametrize("rans", [False, True])
def test_oracle_source_points_match_the_reference_library(rans):
    """rxo_source_points against the reference's own Framework::ReactingModelLibrary driven directly over independent points
    (oracle/ref_source_harness.cpp -> tests/golden/source_points.npz): pins the restatement — and through
    it the stand-alone source kernel — on the states of the source microbenchmark (T up to 2500 K, pure-oxidiser / pure-fuel points)."""
    import ctypes as C
    import rxcase
    pk = rxcase.load_package()
    chem = pk.load_mechanism()
    g = np.load(GOLDEN + "/source_points.npz")
    lib = C.CDLL(rxcase.build_oracle())
    keep = []
    k = lambda a: (keep.append(np.ascontiguousarray(a, dtype=np.float64)), keep[-1])[1]
    Ns, Nr = chem["Ns"], chem["Nr"]
    ch = rxcase._Chem(Ns, Nr, chem["nKnots"], *[rxcase._p(k(np.asarray(chem[n]).reshape(-1))) for n in
                     ["mMasses", "diffVol", "x", "y", "y2", "As", "Betas", "Ta", "reversible", "availBack", "AsB", "BetasB", "TaB",
                      "nuR", "nuP", "expR", "expP"]])
    n = len(g["T"])
    res = np.empty((n, Ns))
    lib.rxo_source_points.restype = None
    lib.rxo_source_points(C.byref(ch), int(rans), C.c_double(float(g["cMu"])), C.c_double(float(g["pasrLb"])), C.c_int64(n), rxcase._p(k(g["T"])),
                          rxcase._p(k(g["rho"])), rxcase._p(k(g["Y"])), rxcase._p(k(g["omegaTurb"])), rxcase._p(k(g["vol"])), rxcase._p(res))
    want = g["res_pasr" if rans else "res_laminar"]
    assert np.abs(want).max() > 0
    pw = rxcase.pointwise_rel_err(res, want)
    assert pw.max() <= TOL, pw
