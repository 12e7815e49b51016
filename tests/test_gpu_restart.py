"""SU2 native restart files and the Tecplot volume file from / into the device-resident state (rx_restart_write / rx_restart_read /
rx_tecplot_write).

Golden bytes: the file the reference's own COutput::MergeSolution + SetRestart (output_structure.cpp:1991-2130, 3858-4060) wrote after
the last iteration of the stage fixtures channel_lam_rk and channel_rans (tests/golden/restart_files.npz, oracle/make_golden.py).
 * the device path is bit-identical to the reference, laminar and RANS, so the files must be IDENTICAL BYTE FOR BYTE;
 * round trip: a file written from the device state is read back by the reference's own Load_Restart + turbulence reader
   (solver_direct_reactive.cpp:4390-4520, solver_direct_turbulent.cpp:1046-1064) through ref_harness and by rx_restart_read;
 * per-partition pieces merged by PointID equal the single-partition file."""
import os
import subprocess

import numpy as np
import pytest

from rxcase import GOLDEN, ROOT, Case, gpu_solver, load_package
from test_oracle_golden import drive

pytestmark = pytest.mark.gpu
HARNESS_CPU = os.path.join(ROOT, "oracle", "_ref", "ref_harness")
CASES = os.path.join(ROOT, "oracle", "_ref", "cases")


def _state_at_write(name):
    """GPU solver driven through the fixture's last iteration: the state the reference had when it wrote the file."""
    case = Case(os.path.join(GOLDEN, name + ".npz"))
    s = gpu_solver(case)
    drive(case, s, lambda k, a: None)
    kw = None
    if case.meta["cfg"]["rans"]:
        kw = case[sorted(k for k in case.arr if k.endswith("_tpost"))[-1]][:, 0:2]   # (k, omega) after the SST iteration
    it = max(int(k[2:6]) for k in case.arr if k.startswith("it"))
    return case, s, kw, it


def _numbers(text):
    rows = [l.split("\t") for l in text.split("\n")[1:] if l and l[0].isdigit()]
    return np.array([[float(x) for x in r if x] for r in rows])


@pytest.mark.parametrize("name", ["channel_lam_rk", "channel_rans"])
def test_restart_file_equals_the_reference_writers(name, tmp_path):
    want = bytes(np.load(os.path.join(GOLDEN, "restart_files.npz"))[name])
    case, s, kw, it = _state_at_write(name)
    path = str(tmp_path / "restart_flow.dat")
    s.WriteRestart(path, kOmega=kw, globalIndex=case["geo_global_index"], ext_iter=it + 1)
    got = open(path, "rb").read()
    s.close()
    gl, wl = got.decode().split("\n"), want.decode().split("\n")
    assert gl[0] == wl[0], "header line differs"
    assert len(gl) == len(wl) and gl[-6:] == wl[-6:], "line count or trailer differs"
    assert got == want, "restart file is not byte-identical to the reference's"
    a, b = _numbers(got.decode()), _numbers(want.decode())
    assert a.shape == b.shape
    scale = np.abs(b).max(axis=0)
    assert (np.abs(a - b).max(axis=0) <= 1e-12 * np.where(scale > 0, scale, 1.0)).all()
    assert all(len(x) == len(y) for x, y in zip(gl, wl)), "field widths differ"


def _channel_triangles(nx, ny):
    """element list of oracle/make_golden.py: write_channel_su2 (every quad split in two, alternating diagonal)"""
    idx = lambda i, j: j * nx + i
    tr = []
    for j in range(ny - 1):
        for i in range(nx - 1):
            a, b, c, d = idx(i, j), idx(i + 1, j), idx(i + 1, j + 1), idx(i, j + 1)
            tr += [(a, b, c), (a, c, d)] if (i + j) % 2 == 0 else [(a, b, d), (b, c, d)]
    return np.array(tr, dtype=np.int64)


@pytest.mark.parametrize("name", ["channel_lam_rk", "channel_rans"])
def test_tecplot_file_equals_the_reference_writers(name, tmp_path):
    """rx_tecplot_write against the bytes of COutput::MergeConnectivity + SetTecplotASCII (output_tecplot.cpp:46-420) after the last
    iteration of the stage fixture: every state of the path is bit-identical, so the whole file must be IDENTICAL BYTE FOR BYTE
    (RANS: (k, omega) and the eddy viscosity column come from the reference's turbulence dump, as in the restart test)."""
    want = bytes(np.load(os.path.join(GOLDEN, "restart_files.npz"))[name + "_tecplot"])
    case, s, kw, it = _state_at_write(name)
    path = str(tmp_path / "flow.dat")
    s.WriteTecplot(path, _channel_triangles(41, 21), kOmega=kw, globalIndex=case["geo_global_index"])
    got = open(path, "rb").read()
    s.close()
    gl, wl = got.decode().split("\n"), want.decode().split("\n")
    assert gl[:3] == wl[:3], "title / VARIABLES / ZONE lines differ"
    assert len(gl) == len(wl) and gl[-1601:] == wl[-1601:], "connectivity differs"
    assert got == want, "Tecplot file is not byte-identical to the reference's"


@pytest.mark.parametrize("name", ["channel_lam_rk", "channel_rans"])
def test_written_file_is_read_back_by_the_reference_and_by_the_library(name, tmp_path):
    case, s, kw, it = _state_at_write(name)
    gi = case["geo_global_index"]
    U = s.solution()
    d = os.path.join(CASES, name)
    # the library's own reader, on a fresh context
    path = str(tmp_path / "restart_flow.dat")
    s.WriteRestart(path, kOmega=kw, globalIndex=gi, ext_iter=it + 1)
    s2 = gpu_solver(case)
    kw2, info = s2.LoadRestart(path, globalIndex=gi, nGlobal=case.nPoint)
    U2 = s2.solution()
    s2.close()
    assert info["extIter"] == it + 1 and info["bcThrust"] == 4000.0
    den = np.maximum(np.abs(U), 1e-300)
    assert (np.abs(U2 - U) / den).max() <= 2e-15            # sixteen significant digits in the file
    if kw is not None:
        assert (np.abs(kw2 - kw) / np.maximum(np.abs(kw), 1e-300)).max() <= 2e-15
    # the reference's readers (driver construction with RESTART_SOL= YES), state dumped before any iteration
    if not (os.path.exists(HARNESS_CPU) and os.path.exists(os.path.join(d, "case_restart.cfg"))):
        s.close()
        pytest.skip("oracle/_ref/ref_harness or the case directory is not in this snapshot")
    run = tmp_path / "run"
    run.mkdir()
    for f in os.listdir(d):
        os.symlink(os.path.join(d, f), str(run / f))
    s.WriteRestart(str(run / "restart_flow.dat"), kOmega=kw, globalIndex=gi, ext_iter=it + 1)
    s.close()
    log = subprocess.run([HARNESS_CPU, "case_restart.cfg", "0", str(tmp_path / "dump"), str(10 ** 9)], cwd=str(run), env=dict(os.environ, RX_CKPT="0"),
                         stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert log.returncode == 0, log.stdout[-2000:]
    Ur = np.load(str(tmp_path / "dump" / "ckpt000000_U.npy"))
    assert (np.abs(Ur - U) / den).max() <= 2e-15
    if kw is not None:
        kr = np.load(str(tmp_path / "dump" / "ckpt000000_kw.npy"))[:, 0:2]
        assert (np.abs(kr - kw) / np.maximum(np.abs(kw), 1e-300)).max() <= 2e-15


def test_partition_pieces_merge_to_the_single_partition_file(tmp_path):
    pk = load_package()
    chem = pk.load_mechanism()
    opts = pk.default_opts(rans=False, cfl=0.1)
    nx, ny, nr = 37, 11, 3
    g = pk.channel_mesh(nx, ny)
    Ug, Tg, tg = pk.synthetic_state(g, chem, seed=9)
    ref = pk.CReactiveNSSolverGPU(g, chem, opts)
    ref.SetSolution(Ug); ref.SetTemperature(Tg); ref.SetTurbulence(tg); ref.Preprocessing(1, True, count_nonphysical=False)
    whole = str(tmp_path / "whole.dat")
    ref.WriteRestart(whole, ext_iter=7)
    pieces = []
    for r in range(nr):
        m, plan = pk.channel_strip_mesh(nx, ny, r, nr)
        U, T, tb = pk.synthetic_state(m, chem, seed=9, glob_xy=plan["bbox"])
        s = pk.CReactiveNSSolverGPU(m, chem, opts)
        s.SetSolution(U); s.SetTemperature(T); s.SetTurbulence(tb); s.Preprocessing(1, True, count_nonphysical=False)
        glob = plan["glob"][1] * nx + plan["glob"][0]
        pieces.append(str(tmp_path / ("piece%d.dat" % r)))
        s.WriteRestart(pieces[-1], globalIndex=glob, piece=True)
        s.close()
    merged = str(tmp_path / "merged.dat")
    pk.merge_restart_pieces(pieces, merged, ref, ext_iter=7)
    ref.close()
    assert open(merged, "rb").read() == open(whole, "rb").read()
