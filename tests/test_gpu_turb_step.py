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
CASES = ["channel_rans", "channel_rans_gg", "channel_rans_ign", "jet_rans_shipped", "jet_rans_perturbed", "channel_air3_rans", "channel_rans_inlet_massflow"]


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
        s.Iterate(it, alphas)
        s.SetWall_Distance(tpost[:, 2])
        s.TurbSetState(case.stage(it, 0, "turb")[:, 1:3], case.stage(it, 0, "tblend"))
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
