"""SST Postprocessing on the device (first slice; csrc/rx_turb.cuh, rx_turb_postprocess) against the
reference's own turbulence solver state after its iteration (`tpost` dumps of oracle/ref_harness.cpp), and partitioned ==
single.  Tolerance 1e-12 of the column maximum: the flow inputs (rho, laminar viscosity, strain magnitude) are the device's own,
themselves within 1e-13 of the reference's, and pow / tanh are CUDA's (<= 2 ulp from glibc's)."""
import numpy as np
import pytest

from rxcase import Case, GOLDEN, gpu_solver, load_package
from test_oracle_golden import TPOST_COLS, drive
from test_gpu_partition import LocalExchanger

pytestmark = pytest.mark.gpu
TOL = 1e-12
WHAT_TURB_GRADIENT = 3


@pytest.mark.parametrize("name", ["channel_rans", "jet_rans_shipped", "jet_rans_perturbed", "channel_rans_gg"])
def test_turbulence_postprocessing_matches_reference(name):
    case = Case("%s/%s.npz" % (GOLDEN, name))
    key = sorted(k for k in case.arr if k.endswith("_tpost"))[-1]
    ref = case[key]
    s = gpu_solver(case)
    drive(case, s, lambda k, a: None)   # leaves the flow state where the reference's turbulence step found it (after the out_ pass)
    s.SetWall_Distance(ref[:, 2])
    s.TurbPostprocessing(ref[:, 0:2], sigma_k=float(ref[0, 11]))
    out = s.GetTurbulence()
    s.close()
    nD = case.nPointDomain
    worst = {}
    for c, nm in enumerate(TPOST_COLS):
        scale = np.abs(ref[:nD, c]).max()
        worst[nm] = np.abs(out[:nD, c] - ref[:nD, c]).max() / (scale if scale > 0 else 1.0)
    print(name, {k: float("%.2e" % v) for k, v in worst.items()})
    bad = {k: v for k, v in worst.items() if not (v <= TOL)}
    assert not bad, "%s: columns over tolerance: %s" % (name, bad)


def test_turbulence_postprocessing_needs_rans_context_and_wall_distance():
    pk = load_package()
    chem = pk.load_mechanism()
    mesh = pk.channel_mesh(12, 8)
    lam = pk.CReactiveNSSolverGPU(mesh, chem, pk.default_opts(rans=False, cfl=0.1))
    with pytest.raises(pk.RxError):
        lam.SetWall_Distance(np.ones(lam.nPoint))
    rans = pk.CReactiveNSSolverGPU(mesh, chem, pk.default_opts(rans=True, cfl=0.1))
    with pytest.raises(pk.RxError):
        rans.TurbPostprocessing(np.ones((rans.nPoint, 2)))   # no wall distance yet


def test_partitioned_turbulence_postprocessing_equals_single():
    """Three strips: gradients (stage 1), exchange of the halo gradients (what = 3), blending + eddy viscosity (stage 2)."""
    pk = load_package()
    nxg, ny, nranks = 48, 20, 3
    chem = pk.load_mechanism()
    opts = pk.default_opts(rans=True, cfl=0.1)

    def fields(xy):   # smooth positive k, omega and a wall distance, functions of the coordinates
        x, y = xy[:, 0], xy[:, 1]
        k = 1.0 + 0.5 * np.sin(40.0 * x) * np.cos(90.0 * y)
        w = 2.0e3 * (1.0 + 0.3 * np.cos(25.0 * x + 60.0 * y))
        return np.stack([k, w], axis=1), np.minimum(y, 0.025 - y) + 1e-5

    g = pk.channel_mesh(nxg, ny)
    Ug, Tg, tg = pk.synthetic_state(g, chem, seed=5, rans=True)
    ref = pk.CReactiveNSSolverGPU(g, chem, opts)
    ref.SetSolution(Ug); ref.SetTemperature(Tg); ref.SetTurbulence(tg)
    ref.Preprocessing(1, True, count_nonphysical=False)
    kw, dist = fields(np.asarray(g["coord"]))
    ref.SetWall_Distance(dist); ref.TurbPostprocessing(kw)
    want = ref.GetTurbulence()
    solvers, plans, globs = [], [], []
    for r in range(nranks):
        m, plan = pk.channel_strip_mesh(nxg, ny, r, nranks)
        U, T, tb = pk.synthetic_state(m, chem, seed=5, rans=True, glob_xy=plan["bbox"])
        s = pk.CReactiveNSSolverGPU(m, chem, opts)
        s.SetSolution(U); s.SetTemperature(T); s.SetTurbulence(tb)
        s.Preprocessing(1, True, count_nonphysical=False)
        solvers.append(s); plans.append(plan); globs.append(plan["glob"][1] * nxg + plan["glob"][0])
    ex = LocalExchanger(solvers, plans)
    ex.exchange(1)   # primitive gradients of the halo points, as after every Preprocessing (strain magnitude is local to the point)
    for s, glob in zip(solvers, globs):
        kwl, dl = fields(np.asarray(g["coord"])[glob])
        s.SetWall_Distance(dl); s.TurbPostprocessing(kwl, stage=1)
    ex.exchange(WHAT_TURB_GRADIENT)
    for s, glob in zip(solvers, globs):
        s.TurbPostprocessing(stage=2)
        got = s.GetTurbulence()
        n = s.nPointDomain
        for c, nm in enumerate(TPOST_COLS):
            scale = np.abs(want[:, c]).max()
            # owned points: everything
            err = np.abs(got[:n, c] - want[glob[:n], c]).max() / (scale if scale > 0 else 1.0)
            assert err <= TOL, (nm, err)
        for c in (6, 7, 8, 9, 10):   # halo layer: eddy viscosity (strain magnitude re-formed from the received gradients), exchanged gradients
            scale = np.abs(want[:, c]).max()
            err = np.abs(got[n:, c] - want[glob[n:], c]).max() / scale
            assert err <= TOL, (TPOST_COLS[c], "halo", err)


@pytest.mark.parametrize("name", ["channel_rans", "jet_rans_shipped"])
def test_wall_distance_from_the_context_matches_reference(name):
    """rx_compute_wall_distance (nearest node of the context's own isothermal / heat-flux markers, on the device) against the
    reference's GetWall_Distance (column 2 of `tpost`)."""
    case = Case("%s/%s.npz" % (GOLDEN, name))
    ref = case[sorted(k for k in case.arr if k.endswith("_tpost"))[-1]]
    s = gpu_solver(case)
    s.ComputeWall_Distance()
    drive(case, s, lambda k, a: None)
    s.TurbPostprocessing(ref[:, 0:2], sigma_k=float(ref[0, 11]))
    out = s.GetTurbulence()
    s.close()
    np.testing.assert_allclose(out[:, 2], ref[:, 2], rtol=1e-15, atol=0)
    assert np.abs(out[:, 6] - ref[:, 6]).max() <= TOL * np.abs(ref[:, 6]).max()   # eddy viscosity with the device's own distances
