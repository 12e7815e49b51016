"""Converged-field / residual-history parity over long runs (north_star: "the converged density/species/temperature fields and
residual history within 1e-10").

Reference: the compiled reference classes (oracle/_ref/ref_harness) run for 2000 iterations of the laminar 3-stage-RK channel
and 500 iterations of the shipped jet-reactor RANS + PaSR case (RESTART_SOL= NO), checkpoints and the per-iteration RMS of every
variable committed as tests/golden/long_*.npz (oracle/make_golden.py: LONG).  Loop followed: driver_structure.cpp:2654-2723,
monitor values integration_time.cpp:756-763 (CSolver::GetRes_RMS).

GPU side: the laminar case runs rx_iterate through the Python mirror; the jet case runs the reference's own driver + SST solver
with the flow solver replaced by CReactiveNSSolverGPU (oracle/_ref/ref_harness_gpu) — the drop-in itself.

Metrics (written to gpurun_out/longrun_drift.json; committed copy profiles/r2_longrun_drift.json):
  field drift  = max_p |a - ref| / max(|ref|, floor)  for rho, T and every rho_s at each checkpoint; floor for a species density =
                 1e-10 x the largest mixture density, i.e. a mass fraction of 1e-10 — the reference's own significance threshold
                 for a mass fraction (reacting_model_library.cpp:131, 310) — so that products present at mass fractions of
                 1e-18 ... 1e-48 (cold Arrhenius rates; they carry no digits the flow can see) are judged on the scale of the
                 mixture; the raw entry-relative figure of every species is reported next to it ("raw").
  history      = max_v |log10 RMS_gpu - log10 RMS_ref| per iteration over the variables whose RMS is above 1e-10 of the density
                 residual's (same reasoning); the raw per-variable maxima are reported as well.
"""
import json
import os
import subprocess

import numpy as np
import pytest

from rxcase import GOLDEN, ROOT, Case, gpu_solver

pytestmark = pytest.mark.gpu
TOL_FIELD = 1e-10
TOL_HIST = 1e-10
HARNESS_GPU = os.path.join(ROOT, "oracle", "_ref", "ref_harness_gpu")
CASES = os.path.join(ROOT, "oracle", "_ref", "cases")


def field_drift(a, ref, floor_abs=None, floor=1e-6):
    """max_p |a - ref| / max(|ref|, floor_abs)  per column; floor_abs defaults to floor x the column maximum."""
    a = np.asarray(a, dtype=np.float64).reshape(ref.shape[0], -1)
    r = np.asarray(ref, dtype=np.float64).reshape(ref.shape[0], -1)
    fl = floor * np.abs(r).max(axis=0)[None, :] if floor_abs is None else floor_abs
    den = np.maximum(np.abs(r), fl)
    return (np.abs(a - r) / np.where(den > 0, den, 1.0)).max(axis=0)


def species_drift(U, Uref, nVar):
    """(drift of rho, largest drift of a rho_s on the mass-fraction-1e-10 scale, largest raw entry-relative drift of a rho_s)"""
    cols = list(range(4, nVar))
    rho = field_drift(U[:, [0]], Uref[:, [0]])[0]
    ys = field_drift(U[:, cols], Uref[:, cols], floor_abs=1e-10 * np.abs(Uref[:, 0]).max())
    raw = field_drift(U[:, cols], Uref[:, cols], floor_abs=1e-300)
    return float(rho), float(ys.max()), float(raw.max())


def history_drift(h, href):
    """(max over significant variables per iteration, raw max per variable)"""
    d = np.zeros_like(href)
    pos = href > 1e-30   # a variable on the EPS^2 floor (absent species) has no digits to compare
    d[pos] = np.abs(np.log10(h[pos]) - np.log10(href[pos]))
    live = href > 1e-10 * href[:, [0]]
    return np.where(live, d, 0.0).max(axis=1), d.max(axis=0)


def _report(name, rec):
    out = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out, exist_ok=True)
    path = os.path.join(out, "longrun_drift.json")
    allr = json.load(open(path)) if os.path.exists(path) else {}
    allr[name] = rec
    json.dump(allr, open(path, "w"), indent=1, sort_keys=True)


def test_laminar_channel_2000_iterations_fields_and_history():
    g = np.load(os.path.join(GOLDEN, "long_channel_lam_rk.npz"))
    meta = json.loads(bytes(g["meta_json"]).decode())
    case = Case(os.path.join(GOLDEN, meta["base"] + ".npz"))
    alphas = case.meta["cfg"]["rk_alpha"]
    s = gpu_solver(case, {"cfl": float(meta["cfg"]["CFL_NUMBER"])})
    s.set_solution(g["ckpt000000_U"]); s.set_temperature(g["ckpt000000_T"]); s.set_turbulence(np.zeros((case.nPoint, 6)))
    nit = meta["nIter"]
    hist = np.empty((nit, case.nVar))
    rec = {"iterations": nit, "checkpoints": {}}
    worst = 0.0
    for it in range(nit):
        hist[it] = s.Iterate(it, alphas)
        if it + 1 in meta["ckpt"]:
            U = s.solution(); T = s.primitive()[0][:, 0]
            dr, dy, draw = species_drift(U, g["ckpt%06d_U" % (it + 1)], case.nVar)
            dT = field_drift(T.reshape(-1, 1), g["ckpt%06d_T" % (it + 1)].reshape(-1, 1))
            rec["checkpoints"][str(it + 1)] = {"rho": dr, "rho_s": dy, "rho_s_raw": draw, "T": float(dT[0])}
            worst = max(worst, dr, dy, float(dT[0]))
    s.close()
    hd, hraw = history_drift(hist, g["history_rms"])
    rec["history_max_dlog10"] = float(hd.max()); rec["history_argmax_iter"] = int(hd.argmax())
    rec["history_dlog10_at"] = {str(k): float(hd[:k].max()) for k in (1, 10, 100, 500, 1000, 2000) if k <= nit}
    rec["history_raw_max_dlog10_per_variable"] = [float(x) for x in hraw]
    rec["log10_rms_rho_first_last"] = [float(np.log10(hist[0, 0])), float(np.log10(hist[-1, 0]))]
    _report("long_channel_lam_rk", rec)
    print(json.dumps(rec, indent=1))
    assert worst <= TOL_FIELD, rec
    assert hd.max() <= TOL_HIST, rec


@pytest.mark.parametrize("turb", ["host", "device"])
def test_shipped_jet_rans_500_iterations_through_the_dropin(turb, tmp_path):
    """turb = host: the reference's CTurbSSTSolver on the CPU next to the CUDA flow solver; device: CTurbSSTSolverGPU as well — both
    solvers' iterations on the device (rx_iterate-equivalent staged calls + rx_turb_iterate), the reference's driver / integration on top."""
    d = os.path.join(CASES, "jet_rans_shipped")
    if not (os.path.exists(HARNESS_GPU) and os.path.exists(os.path.join(d, "case_long.cfg"))):
        pytest.skip("oracle/_ref/ref_harness_gpu or the case directory is not in this snapshot")
    g = np.load(os.path.join(GOLDEN, "long_jet_rans_shipped.npz"))
    meta = json.loads(bytes(g["meta_json"]).decode())
    out = str(tmp_path / "gpu")
    env = dict(os.environ, SU2_RX_BACKEND="cuda", SU2_RX_MIRROR="1", SU2_RX_TURB=turb, RX_CKPT=",".join(map(str, meta["ckpt"])), RX_QUIET="1")
    log = subprocess.run([HARNESS_GPU, "case_long.cfg", str(meta["nIter"]), out, str(10 ** 9)], cwd=d, env=env,
                         stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=1500)
    assert log.returncode == 0, log.stdout[-3000:]
    assert "CReactiveNSSolverGPU" in log.stdout and (("CTurbSSTSolverGPU" in log.stdout) == (turb == "device"))
    rec = {"iterations": meta["nIter"], "checkpoints": {}, "turbulence_solver": turb}
    worst = 0.0
    nVar = g["ckpt000000_U"].shape[1]
    for n in meta["ckpt"]:
        if n == 0:
            continue
        U = np.load(os.path.join(out, "ckpt%06d_U.npy" % n)); T = np.load(os.path.join(out, "ckpt%06d_T.npy" % n))
        kw = np.load(os.path.join(out, "ckpt%06d_kw.npy" % n))
        dr, dy, draw = species_drift(U, g["ckpt%06d_U" % n], nVar); dT = field_drift(T.reshape(-1, 1), g["ckpt%06d_T" % n].reshape(-1, 1))
        dK = field_drift(kw, g["ckpt%06d_kw" % n])
        rec["checkpoints"][str(n)] = {"rho": dr, "rho_s": dy, "rho_s_raw": draw, "T": float(dT[0]), "k": float(dK[0]), "omega": float(dK[1]), "muT": float(dK[2])}
        worst = max(worst, dr, dy, float(dT[0]))
    hd, hraw = history_drift(np.load(os.path.join(out, "history_rms.npy")), g["history_rms"])
    rec["history_max_dlog10"] = float(hd.max()); rec["history_argmax_iter"] = int(hd.argmax())
    rec["history_dlog10_at"] = {str(k): float(hd[:k].max()) for k in (1, 10, 100, 250, 500) if k <= meta["nIter"]}
    rec["history_raw_max_dlog10_per_variable"] = [float(x) for x in hraw]
    _report("long_jet_rans_shipped_turb_" + turb, rec)
    print(json.dumps(rec, indent=1))
    assert worst <= TOL_FIELD, rec
    assert hd.max() <= TOL_HIST, rec
