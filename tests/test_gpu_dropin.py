"""Drop-in boundary, end to end: the reference's own driver, geometry, config, integration and SST classes (compiled by
oracle/build_ref.sh) run with the flow solver object replaced by the package's C++ subclass CReactiveNSSolverGPU
(csrc/host), which forwards every CSolver virtual to librx_cuda.so.  The harness dumps the same stage arrays it dumped
for the CPU reference; they must agree with the golden fixtures to 1e-12 (grouped metric of rxcase.stage_rel_err)."""
import os
import subprocess

import numpy as np
import pytest

from rxcase import GOLDEN, ROOT, Case, stage_rel_err

pytestmark = pytest.mark.gpu

HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_harness_gpu")
HARNESS_CPU = os.path.join(ROOT, "oracle", "_ref", "ref_harness")
CASES = os.path.join(ROOT, "oracle", "_ref", "cases")
# name -> (iterations, seed)   (oracle/make_golden.py SPECS)
RUNS = {"channel_lam_rk": (2, 11), "channel_lam_gg": (1, 13), "channel_rans": (2, 17), "jet_rans_perturbed": (2, 7),
        "channel_lam_walls": (2, 19), "channel_lam_muscl": (2, 23), "channel_lam_bj": (2, 29), "channel_rans_gg": (2, 37),
        "channel_air3_rans": (2, 59), "channel_mix5_lam": (2, 61),   # the reference's 3- and 5-species mixtures
        "channel_lam_inlet_total": (2, 67), "channel_rans_inlet_massflow": (2, 71),   # INLET_TYPE= TOTAL_CONDITIONS / MASS_FLOW
        "channel_lam_supersonic": (2, 73),   # MARKER_SUPERSONIC_INLET / _OUTLET
        "channel_lam_unsteady": (3, 79), "channel_euler_unsteady": (3, 83), "channel_euler_unsteady_fixed": (2, 89),   # TIME_STEPPING
        "channel_lam_dualtime2": (3, 97), "channel_euler_dualtime1": (4, 101),
        "channel_rans_dualtime2": (4, 103),
        "channel_lam_visclim": (2, 107)}   # VISCOUS_LIMITER_FLOW= YES   # DUAL_TIME_STEPPING-2ND_ORDER / -1ST_ORDER, explicit pseudo-time iterations
CASE_ENV = {"channel_lam_inlet_total": {"RX_PERTURB_TAMP": "0", "RX_PERTURB_NOISE": "0"},   # oracle/make_golden.py ENVS
            # the time levels shifted on the device (CReactiveNSSolverGPU::ShiftTimeLevels next to CIntegration::SetDualTime_Solver); the Euler
            # case takes the default route, where the next physical step's first SetTime_Step uploads the levels the reference shifted in node[]
            "channel_lam_dualtime2": {"RX_DUAL_SHIFT_DEVICE": "1"}}
TOL_FIRST = 1e-12     # first iteration: identical inputs on both sides
TOL_CHAINED = 1e-12   # later iterations too: with the glibc-exact exp / pow (csrc/rx_libm.cuh) the laminar chains are bit-identical
TOL_FIRST_BY_CASE = {}


def _run(binary, d, out, nit, seed, env):
    log = subprocess.run([binary, "case.cfg", str(nit), out, "0", str(seed)], cwd=d, env=env,
                         stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
    assert log.returncode == 0, log.stdout[-3000:]
    return log.stdout


# RANS cases once more with SU2_RX_TURB_POST=1: muT, grad k, sigma_k formed on the device by rx_turb_postprocess from (k, omega)
# instead of copied from the host-side CTurbSSTSolver::Postprocessing
# and with SU2_RX_TURB=device: the SST solver object replaced by CTurbSSTSolverGPU as well — its whole iteration (residuals, boundary
# conditions, FGMRES + LU-SGS, update, Postprocessing) runs on the device (rx_turb_iterate); the reference's driver and integration
# classes still drive both solvers
PARAMS = [(n, False, False) for n in RUNS] + [("channel_rans", True, False), ("jet_rans_perturbed", True, False), ("channel_rans_gg", True, False)] + \
         [("channel_rans", False, True), ("jet_rans_perturbed", False, True), ("channel_rans_gg", False, True), ("channel_air3_rans", False, True),
          ("channel_rans_dualtime2", False, True)]   # dual time stepping with both solvers on the device


@pytest.mark.parametrize("name,turb_post,turb_dev", PARAMS)
def test_reference_driver_with_cuda_solver(name, turb_post, turb_dev, tmp_path):
    d = os.path.join(CASES, name)
    if not (os.path.exists(HARNESS) and os.path.exists(HARNESS_CPU) and os.path.exists(os.path.join(d, "case.cfg"))):
        pytest.skip("oracle/_ref harnesses or the case directory are not in this snapshot (built by oracle/build_ref.sh where the reference tree is present)")
    nit, seed = RUNS[name]
    out_g, out_c = str(tmp_path / "gpu"), str(tmp_path / "cpu")
    txt = _run(HARNESS, d, out_g, nit, seed, dict(os.environ, **CASE_ENV.get(name, {}), SU2_RX_BACKEND="cuda", SU2_RX_MIRROR="2", SU2_RX_TURB_POST="1" if turb_post else "0", RX_RESTART_OUT="1", RX_TECPLOT_OUT="1",
                                                     SU2_RX_TURB="device" if turb_dev else "host"))
    assert "CReactiveNSSolverGPU" in txt and (("CTurbSSTSolverGPU" in txt) == turb_dev)
    # restart file: the reference's COutput (fed by the subclass's host mirrors) and the subclass's own device writer, same run
    def _pair(stem):   # the reference's file (an unsteady run appends the time iteration: restart_flow_00002.dat) and the subclass's
        import glob
        ref = [f for f in glob.glob(os.path.join(d, stem + "*.dat")) if not f.endswith("_gpu.dat")]
        assert len(ref) == 1, ref
        a = open(ref[0], "rb").read(); b = open(os.path.join(d, stem + "_gpu.dat"), "rb").read()
        os.remove(ref[0]); os.remove(os.path.join(d, stem + "_gpu.dat"))
        return a, b
    if "rx_restart_write" in txt:
        a, b = _pair("restart_flow")
        assert a == b, "rx_restart_write differs from COutput::SetRestart on the same state"
    if "rx_tecplot_write" in txt:   # likewise the Tecplot volume file: COutput::SetTecplotASCII and the subclass's WriteTecplot
        a, b = _pair("flow")
        assert a == b, "rx_tecplot_write differs from COutput::SetTecplotASCII on the same state"
    _run(HARNESS_CPU, d, out_c, nit, seed, dict(os.environ, **CASE_ENV.get(name, {})))
    case = Case(os.path.join(GOLDEN, name + ".npz"))
    checked, worst = 0, (0.0, "")
    for f in sorted(os.listdir(out_c)):
        k = f[:-4]
        if not f.endswith(".npy") or not k.startswith("it"):
            continue
        what = k.split("_", 2)[-1]
        if what in ("U0", "turb", "tblend") or what.startswith("prev_"):
            continue   # inputs of the iteration (checked through what they produce)
        if turb_dev and what == "tres":
            continue   # the device assembles and solves in one call: LinSysRes is not mirrored before the solve (tests/test_gpu_turb_step.py compares it)
        a = np.load(os.path.join(out_g, f)); ref = np.load(os.path.join(out_c, f))
        if case.has(k) and case[k].shape == ref.shape:
            assert np.array_equal(ref, case[k]), "CPU reference run differs from the committed golden fixture: " + k
        if what in ("rms", "trms"):
            a = a.reshape(1, -1); ref = ref.reshape(1, -1)
        if what == "dt":
            err = np.abs(a - ref).max(axis=0) / np.maximum(np.abs(ref).max(axis=0), 1e-300)
        else:
            err = stage_rel_err(k, a, ref, case.Ns)
        e = float(np.max(err))
        tol = TOL_FIRST_BY_CASE.get(name, TOL_FIRST) if k.startswith("it0000") else TOL_CHAINED
        assert e < tol, "stage %s differs from the reference run by %.3e" % (k, e)
        if e > worst[0]:
            worst = (e, k)
        checked += 1
    assert checked >= 10, "no stage arrays were compared"
    print("worst", worst)
