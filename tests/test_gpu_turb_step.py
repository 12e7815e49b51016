"""The SST transport step on the device (rx_turb_iterate; csrc/rx_turb_step.cuh) against the reference's own CTurbSSTSolver, stage by
stage: after the flow iteration of every dumped iteration of the RANS golden cases the harness ran the turbulence solver's
Preprocessing / Space_Integration / Time_Integration / Postprocessing (integration_time.cpp:777-834) and dumped
  tres  LinSysRes after Upwind + Viscous + Source residual and the boundary conditions,
  tdx   the solution of the implicit system (FGMRES, 5 iterations, LU-SGS),
  tsol  (k, omega) after the clipped conservative update,  trms  its residual norms,  tpost  the state after Postprocessing.
Tolerances: tres 1e-12 of the column maximum (same sums in the same order); tdx / tsol / tpost 1e-10 — the Krylov solve stops after five
iterations, so its result moves with the rounding of the inner products (tree sums here, sequential sums in CSysVector::dotProd)."""
import numpy as np
import pytest

from rxcase import Case, GOLDEN, gpu_solver

pytestmark = pytest.mark.gpu
CASES = ["channel_rans", "channel_rans_gg", "channel_rans_ign", "jet_rans_shipped", "jet_rans_perturbed", "channel_air3_rans", "channel_rans_inlet_massflow",
         "channel_rans_dualtime2"]   # dual time stepping: CTurbSolver::SetResidual_DualTime (:730-815) inside the step


def colerr(a, ref):
    scale = np.abs(ref).max(axis=0)
    return (np.abs(a - ref).max(axis=0) / np.where(scale > 0, scale, 1.0)).max()


@pytest.mark.parametrize("name", CASES)
def test_sst_step_matches_reference(name):
    case = Case("%s/%s.npz" % (GOLDEN, name))
    cf = case.meta["cfg"]
    assert cf["spatial_order_turb"] in (0, 1)   # FIRST_ORDER (option_structure.hpp): the device step is first order
    alphas = cf["rk_alpha"] if cf["time_scheme"] == 1 else [1.0]
    its = sorted({int(k[2:6]) for k in case.arr if k.startswith("it") and k.endswith("_tres")})
    assert its
    s = gpu_solver(case)
    worst = {}
    for it in its:
        tpost = case.stage(it, 0, "tpost")
        s.set_solution(case.stage(it, 0, "U0")); s.set_temperature(case.stage(it, 0, "prev_V")[:, 0]); s.set_turbulence(case.stage(it, 0, "turb"))
        if cf.get("dual_time", 0):   # the stored time levels of both solvers; the pseudo-time iteration index of the flow step
            s.set_time_levels(case.stage(it, 0, "Un"), case.stage(it, 0, "Un1")); s.set_inner_iteration(it % cf.get("dual_inner", 2))
        s.Iterate(it, alphas)
        s.SetWall_Distance(tpost[:, 2])
        s.TurbSetState(case.stage(it, 0, "turb")[:, 1:3], case.stage(it, 0, "tblend"))
        if cf.get("dual_time", 0):
            s.TurbSetTimeLevels(case.stage(it, 0, "tUn"), case.stage(it, 0, "tUn1"))
        rms, nlin = s.TurbIterate(cf["kine_inf"], cf["omega_inf"], cf["cfl_red_turb"], cf["relax_turb"], cf["lin_iter"], cf["lin_tol"])
        res, dx = s.TurbGetStep()
        post = s.GetTurbulence()
        nD = case.nPointDomain
        e = {"tres": colerr(res[:nD], case.stage(it, 0, "tres")[:nD]), "tdx": colerr(dx[:nD], case.stage(it, 0, "tdx")[:nD]),
             "tsol": colerr(post[:nD, 0:2], case.stage(it, 0, "tsol")[:nD]), "trms": float(np.abs(rms / case.stage(it, 0, "trms") - 1.0).max()),
             "tpost": colerr(post[:nD], tpost[:nD])}
        for k, v in e.items():
            worst[k] = max(worst.get(k, 0.0), float(v))
        assert nlin == cf["lin_iter"] or nlin < cf["lin_iter"]
    s.close()
    print(name, {k: float("%.2e" % v) for k, v in worst.items()})
    assert worst["tres"] <= 1e-12 and worst["trms"] <= 1e-12, worst
    assert worst["tdx"] <= 1e-10 and worst["tsol"] <= 1e-10 and worst["tpost"] <= 1e-10, worst


def test_sst_step_on_a_larger_mesh_equals_the_oracle_and_does_not_depend_on_the_sweep_block(monkeypatch):
    """A 160 x 90 channel (14 400 points, levels up to 45 points wide): the device step against the CPU restatement (rxo_turb_iterate,
    pinned bit for bit to the reference on the fixtures), and once more with 32-thread sweep blocks (RX_SWEEP_THREADS), which sends every
    level wider than the block through the loop the 16 M-point meshes use for their widest levels — same bits as the default block."""
    from rxcase import Oracle, load_package
    pk = load_package()
    chem = pk.load_mechanism()
    opts = pk.default_opts(rans=True, cfl=0.1)
    mesh = pk.channel_mesh(160, 90)
    U, T, tb = pk.synthetic_state(mesh, chem, seed=21, rans=True)
    alphas = [0.66667, 0.66667, 1.0]

    def gpu(threads):
        if threads:
            monkeypatch.setenv("RX_SWEEP_THREADS", str(threads))
        else:
            monkeypatch.delenv("RX_SWEEP_THREADS", raising=False)
        s = pk.CReactiveNSSolverGPU(mesh, chem, opts)
        s.SetSolution(U); s.SetTemperature(T); s.SetTurbulence(tb)
        s.ComputeWall_Distance()
        s.TurbPostprocessing(np.ascontiguousarray(tb[:, 1:3]))
        t0 = s.GetTurbulence()
        s.Iterate(1, alphas)
        rms, nlin = s.TurbIterate(0.135, 2.0e3)
        res, dx = s.TurbGetStep()
        out = (t0, res, dx, s.GetTurbulence(), rms, nlin)
        s.close()
        return out

    t0, res, dx, post, rms, nlin = gpu(0)
    _t, res32, dx32, post32, rms32, nlin32 = gpu(32)
    assert np.array_equal(dx, dx32) and np.array_equal(post, post32) and np.array_equal(res, res32) and nlin == nlin32
    o = Oracle.from_mesh(mesh, chem, opts)
    # the flow iteration on the device read the eddy viscosity, grad k and sigma_k its own Postprocessing formed from (k, omega)
    turb6 = np.ascontiguousarray(np.stack([t0[:, 6], t0[:, 0], t0[:, 1], t0[:, 7], t0[:, 8], t0[:, 11]], axis=1))
    o.set_solution(U); o.set_temperature(T); o.set_turbulence(turb6)
    for rk in range(3):
        o.preprocess(1, False)
        if rk == 0:
            o.set_old_solution(); o.time_step()
        o.upwind_residual(); o.viscous_residual(); o.source_residual(); o.boundary_conditions()
        o.explicit_update(alphas[rk])
    o.preprocess(1, True)
    r = o.turb_iterate(tb[:, 1:3], t0[:, 3:6], t0[:, 2], 0.85, 0.135, 2.0e3, 1.0, 1.0, 5, 1e-6)
    assert colerr(res, r["res"]) <= 1e-12 and colerr(dx, r["dx"]) <= 1e-10 and colerr(post, r["post"]) <= 1e-10 and nlin == r["iters"]
