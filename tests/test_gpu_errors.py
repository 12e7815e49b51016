"""Error behaviour of the entry points added in round 2: every misuse returns a negative rx_status with a message (RxError in the
Python mirror) instead of a wrong result or a crash — the reference's own habit on this path is to throw
(solver_direct_reactive.cpp:2772, 2861, 5384)."""
import numpy as np
import pytest

from rxcase import load_package

pytestmark = pytest.mark.gpu


def _ctx(pk, rans, nx=12, ny=8):
    chem = pk.load_mechanism()
    mesh = pk.channel_mesh(nx, ny)
    s = pk.CReactiveNSSolverGPU(mesh, chem, pk.default_opts(rans=rans, cfl=0.1))
    U, T, tb = pk.synthetic_state(mesh, chem, seed=5, rans=rans)
    s.SetSolution(U); s.SetTemperature(T); s.SetTurbulence(tb)
    return s, mesh


def test_halo_kinds_of_the_turbulence_fields_need_a_rans_context_and_a_communicator():
    pk = load_package()
    lam, _ = _ctx(pk, False)
    assert lam.HaloWidth(4) == 2 and lam.HaloWidth(3) == 4 and lam.HaloWidth(5) == 0
    for what in (0, 3, 4):
        with pytest.raises(pk.RxError):
            lam.HaloExchange(what)          # no communicator (rx_comm_init)
    lam.close()


def test_turbulence_step_argument_checks():
    pk = load_package()
    lam, _ = _ctx(pk, False)
    with pytest.raises(pk.RxError):
        lam.TurbIterate(0.135, 2.0e3)       # not a RANS context
    lam.close()
    rans, _ = _ctx(pk, True)
    with pytest.raises(pk.RxError):
        rans.TurbIterate(0.135, 2.0e3)      # no wall distance yet
    rans.ComputeWall_Distance()
    with pytest.raises(pk.RxError):
        rans.TurbIterate(0.135, 2.0e3, lin_iter=0)
    with pytest.raises(pk.RxError):
        rans.TurbIterate(0.135, 2.0e3, lin_iter=65)
    rans.close()


def test_partitioned_turbulence_step_without_communicator_is_refused():
    pk = load_package()
    chem = pk.load_mechanism()
    m, plan = pk.channel_strip_mesh(24, 9, 1, 3)
    s = pk.CReactiveNSSolverGPU(m, chem, pk.default_opts(rans=True, cfl=0.1))
    U, T, tb = pk.synthetic_state(m, chem, seed=5, rans=True, glob_xy=plan["bbox"])
    s.SetSolution(U); s.SetTemperature(T); s.SetTurbulence(tb)
    s.SetWall_Distance(np.full(s.nPoint, 1e-3))
    s.Preprocessing(1, False, count_nonphysical=False)
    with pytest.raises(pk.RxError, match="communicator"):
        s.TurbIterate(0.135, 2.0e3)
    with pytest.raises(pk.RxError):
        s.ComputeWall_Distance()            # a partition sees only its own wall nodes
    with pytest.raises(pk.RxError):
        s.WriteTecplot("/tmp/never.dat", np.zeros((1, 3), dtype=np.int64), kOmega=np.zeros((s.nPoint, 2)))   # one partition is part of the zone only
    s.close()


def test_output_writers_argument_checks(tmp_path):
    pk = load_package()
    rans, mesh = _ctx(pk, True)
    rans.Preprocessing(1, True, count_nonphysical=False)
    with pytest.raises(pk.RxError):
        rans.WriteRestart(str(tmp_path / "r.dat"))                       # a RANS restart needs (k, omega)
    with pytest.raises(pk.RxError):
        rans.WriteTecplot(str(tmp_path / "t.dat"), np.zeros((0, 3)))     # likewise
    with pytest.raises(pk.RxError):
        rans.WriteTecplot(str(tmp_path / "no_such_dir" / "t.dat"), np.zeros((0, 3)), kOmega=np.ones((rans.nPoint, 2)))
    rans.WriteTecplot(str(tmp_path / "t.dat"), np.array([[0, 1, 13]]), kOmega=np.ones((rans.nPoint, 2)), quads=np.array([[0, 1, 13, 12]]))
    lines = open(str(tmp_path / "t.dat")).read().split("\n")
    assert lines[2] == "ZONE NODES= %d, ELEMENTS= 2, DATAPACKING=POINT, ZONETYPE=FEQUADRILATERAL" % rans.nPoint
    assert lines[-3:] == ["1\t2\t14\t14", "1\t2\t14\t13", ""]
    rans.close()


def test_unsupported_species_count_is_refused():
    pk = load_package()
    chem = dict(pk.load_mechanism())
    Ns = 4
    chem["Ns"] = Ns
    for k in ("mMasses", "diffVol"):
        chem[k] = np.asarray(chem[k])[:Ns].copy()
    with pytest.raises(pk.RxError, match="Ns"):
        pk.CReactiveNSSolverGPU(pk.channel_mesh(6, 5), chem, pk.default_opts(rans=False, cfl=0.1))


def test_qr_hook_handles_empty_input():
    pk = load_package()
    xs, xp = pk.qr_eval(np.zeros((0, 81)), np.zeros((0, 2, 9)))
    assert xs.shape == (0, 2, 9) and xp.shape == (0, 2, 9)


def test_cfl_can_change_between_iterations_as_cfl_adapt_does():
    """rx_set_cfl (config->SetCFL by COutput::SetCFL_Number, driver_structure.cpp:2804): the time step of the next SetTime_Step equals
    the one of a context created with that CFL, bit for bit; a negative value is refused."""
    pk = load_package()
    a, mesh = _ctx(pk, False)
    chem = pk.load_mechanism()
    b = pk.CReactiveNSSolverGPU(mesh, chem, pk.default_opts(rans=False, cfl=0.37))
    U, T, tb = pk.synthetic_state(mesh, chem, seed=5, rans=False)
    b.SetSolution(U); b.SetTemperature(T); b.SetTurbulence(tb)
    for s in (a, b):
        s.Preprocessing(1, False, count_nonphysical=False)
    a.SetTime_Step(); dt0 = a.dt()[:, 0].copy()
    a.SetCFL(0.37)
    a.SetTime_Step(); b.SetTime_Step()
    assert np.array_equal(a.dt(), b.dt()) and not np.array_equal(a.dt()[:, 0], dt0)
    with pytest.raises(pk.RxError):
        a.SetCFL(-1.0)
    a.close(); b.close()


def test_dual_time_entry_points_need_a_dual_time_context():
    pk = load_package()
    lam, mesh = _ctx(pk, False)
    z = np.zeros((lam.nPoint, lam.nVar))
    for call in (lambda: lam.SetTimeLevels(z, z), lam.GetTimeLevels, lam.SetDualTime_Solver, lam.SetResidual_DualTime, lam.GetDelta_UnstTimeND):
        with pytest.raises(pk.RxError, match="dual time"):
            call()
    lam.close()
    chem = pk.load_mechanism()
    with pytest.raises(pk.RxError):   # neither a physical time step nor an unsteady CFL number
        pk.CReactiveNSSolverGPU(mesh, chem, dict(pk.default_opts(rans=False, cfl=0.1), dualTime=2))
    with pytest.raises(pk.RxError):   # excludes TIME_STEPPING
        pk.CReactiveNSSolverGPU(mesh, chem, dict(pk.default_opts(rans=False, cfl=0.1), dualTime=1, deltaUnstTimeND=1e-7, timeStepping=1))
    s = pk.CReactiveNSSolverGPU(mesh, chem, dict(pk.default_opts(rans=False, cfl=0.1), dualTime=2, deltaUnstTimeND=1e-7))
    U, T, tb = pk.synthetic_state(mesh, chem, seed=5, rans=False)
    s.SetSolution(U); s.SetTemperature(T); s.SetTurbulence(tb)
    a, b = s.GetTimeLevels()          # before any set / shift the levels equal the current solution (SetInitialCondition)
    assert np.array_equal(a, U) and np.array_equal(b, U)
    with pytest.raises(pk.RxError):
        s.SetInnerIteration(-1)
    with pytest.raises(pk.RxError, match="dual time"):   # the SST solver's levels exist in dual-time RANS contexts only
        s.TurbGetTimeLevels()
    s.close()
    rans = pk.CReactiveNSSolverGPU(mesh, chem, dict(pk.default_opts(rans=True, cfl=0.1), dualTime=2, deltaUnstTimeND=1e-7))
    U, T, tb = pk.synthetic_state(mesh, chem, seed=5, rans=True)
    rans.SetSolution(U); rans.SetTemperature(T); rans.SetTurbulence(tb)
    rans.TurbSetState(np.ascontiguousarray(tb[:, 1:3]), None)
    a, b = rans.TurbGetTimeLevels()   # before any set / shift: the current (k, omega)
    assert np.array_equal(a, tb[:, 1:3]) and np.array_equal(b, tb[:, 1:3])
    rans.close()


def test_viscous_limiter_needs_the_slope_limiter():
    """VISCOUS_LIMITER_FLOW multiplies gradients by the limiters of SPATIAL_ORDER_FLOW= 2ND_ORDER_LIMITER; without them the reference reads an
    unset array (solver_direct_reactive.cpp:5326) — rx_create refuses instead."""
    pk = load_package()
    chem = pk.load_mechanism()
    mesh = pk.channel_mesh(12, 8)
    with pytest.raises(pk.RxError, match="viscLimiter"):
        pk.CReactiveNSSolverGPU(mesh, chem, dict(pk.default_opts(rans=False, cfl=0.1), viscLimiter=1))
    with pytest.raises(pk.RxError, match="viscLimiter"):
        pk.CReactiveNSSolverGPU(mesh, chem, dict(pk.default_opts(rans=False, cfl=0.1), viscLimiter=1, secondOrder=1, limiter=0))
    s = pk.CReactiveNSSolverGPU(mesh, chem, dict(pk.default_opts(rans=False, cfl=0.1), viscLimiter=1, secondOrder=1, limiter=1))
    s.close()
