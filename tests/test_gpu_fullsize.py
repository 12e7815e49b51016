"""The CUDA path at the FULL size of the headline configuration (BASELINE.json configs[1]: the 1415 x 1415-point, 4 M-cell triangulated
channel of bench.py, laminar reactive Navier-Stokes, 3-stage RK), where neither the compiled reference nor the CPU restatement finishes in
test time: size-independent properties of the scheme instead of stage-by-stage dumps.
  1. telescoping — every interior edge adds its flux to one end point and subtracts the same number from the other, so the convective +
     viscous residual summed over ALL points vanishes to rounding (1e-17 of the sum of magnitudes on the CPU restatement at small size);
  2. free-stream preservation — a uniform state leaves no convective or viscous residual at interior points (closed dual volumes);
  3. the fused iteration (rx_iterate) equals the staged calls bit for bit, and the same iteration from the same state run twice gives the
     same bits (no atomics, no order-dependent sums anywhere on the path)."""
import math

import numpy as np
import pytest

from rxcase import load_package

pytestmark = pytest.mark.gpu
NX = NY = 1415
RK = [0.66667, 0.66667, 1.0]


@pytest.fixture(scope="module")
def full():
    pk = load_package()
    chem = pk.load_mechanism()
    mesh = pk.channel_mesh(NX, NY)
    U, T, tb = pk.synthetic_state(mesh, chem, seed=31, rans=False)
    s = pk.CReactiveNSSolverGPU(mesh, chem, pk.default_opts(rans=False, cfl=0.1))   # one context for the three tests: each sets its own state
    yield pk, mesh, s, U, T, tb
    s.close()


def test_interior_fluxes_telescope_over_the_whole_mesh(full):
    pk, mesh, s, U, T, tb = full
    s.SetSolution(U); s.SetTemperature(T); s.SetTurbulence(tb)
    s.Preprocessing(1, False, count_nonphysical=False)
    s.Set_OldSolution(); s.SetTime_Step()
    s.Upwind_Residual(); s.Viscous_Residual()
    R = s.GetLinSysRes()
    assert R.shape == (NX * NY, 13) and np.isfinite(R).all()
    worst = 0.0
    for v in range(R.shape[1]):
        mag = float(np.abs(R[:, v]).sum())
        if mag == 0.0:
            continue
        worst = max(worst, abs(float(R[:, v].astype(np.longdouble).sum())) / mag)   # pairwise sum in extended precision
    print("telescoping: worst |sum R| / sum |R| = %.2e" % worst)
    assert worst <= 1e-10
    assert np.abs(R).max() > 0


def test_uniform_state_is_preserved_at_interior_points(full):
    pk, mesh, s, U, T, tb = full
    U = np.repeat(U[:1], U.shape[0], axis=0); T = np.repeat(T[:1], T.shape[0], axis=0)
    s.SetSolution(U); s.SetTemperature(T); s.SetTurbulence(tb)
    s.Preprocessing(1, False, count_nonphysical=False)
    s.Set_OldSolution(); s.SetTime_Step()
    s.Upwind_Residual(); s.Viscous_Residual()
    R = s.GetLinSysRes()
    V, _, _ = s.GetPrimitive()
    boundary = np.zeros(mesh["nPoint"], dtype=bool)
    for m in mesh["markers"]:
        boundary[m["node"]] = True
    # scale of one face flux: the largest flux density of the uniform state times the largest face
    rho, u, v, P, H = V[0, 4], V[0, 1], V[0, 2], V[0, 3], V[0, 5]
    speed = math.hypot(u, v) + V[0, 6]
    scale = max(rho * speed, rho * speed * speed + P, rho * speed * H) * float(np.sqrt((np.asarray(mesh["edgeNormal"]) ** 2).sum(1)).max())
    err = float(np.abs(R[~boundary]).max() / scale)
    print("free stream: max interior |R| / face flux = %.2e" % err)
    assert (~boundary).sum() > 1.9e6
    assert err <= 1e-11
    assert np.abs(R[boundary]).max() > 0   # open boundary faces carry flux until the boundary conditions close them


def test_fused_iteration_equals_staged_and_repeats_bit_for_bit(full):
    pk, mesh, s, U, T, tb = full
    s.SetSolution(U); s.SetTemperature(T); s.SetTurbulence(tb)
    s.SaveState()
    out = []
    for mode in ("fused", "fused", "staged"):
        s.RestoreState()
        if mode == "fused":
            rms = s.Iterate(1, RK)
        else:
            for rk, a in enumerate(RK):
                s.Preprocessing(1, False, count_nonphysical=False)
                if rk == 0:
                    s.Set_OldSolution(); s.SetTime_Step()
                s.Upwind_Residual(); s.Viscous_Residual(); s.Source_Residual(); s.BC_All()
                rms = s.ExplicitRK_Iteration(a)
            s.Preprocessing(1, True, count_nonphysical=False)
        out.append((s.GetSolution(), np.array(rms)))
    assert np.isfinite(out[0][0]).all() and not np.array_equal(out[0][0], U)
    assert np.array_equal(out[0][0], out[1][0]) and np.array_equal(out[0][1], out[1][1])   # the same iteration twice
    assert np.array_equal(out[0][0], out[2][0]) and np.array_equal(out[0][1], out[2][1])   # fused == staged
