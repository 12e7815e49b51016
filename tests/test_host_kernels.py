"""The CUDA kernels of the residual path run on the HOST, one simulated thread after another (tests/host_kernels.cpp over tests/cuda_host_shim.h),
in the launch order of rx_cuda.cu, against the reference's stage dumps — bit for bit, in the CPU test tier.  Where tests/test_host_emulation.py
checks the per-item arithmetic, this runs the kernels themselves: the structure-of-arrays indexing, the incident-edge lists of the owner-computes
accumulation, the neighbour loops of the gradient and limiter kernels, the marker blocks of the boundary kernels.  Every dumped iteration in full — all Runge-Kutta stages, each from the state the kernels themselves left: primitives, transport, gradients, limiters,
time step, the residual after the convective, viscous and source terms and after the boundary conditions, the clipped update, the Output pass.  Interior viscous edges go through the thread-per-edge kernel (on the device nine species use
the two-lanes-per-edge kernel of the same arithmetic, a `-m gpu` test).  TIME_STEPPING and dual time stepping included (their global minimum is a
plain loop here, a block reduction on the device)."""
import ctypes as C
import importlib
import os
import shutil
import subprocess

import numpy as np
import pytest

from rxcase import Case, GOLDEN, ROOT, load_package, chem_arrays, mesh_dict, opts_dict

PKG = os.path.join(ROOT, "development-of-a-turbulent-numerical-solver-for-reactive-flows-in-su2_b200")
_d = lambda a: a.ctypes.data_as(C.POINTER(C.c_double))


@pytest.fixture(scope="module")
def hk(tmp_path_factory):
    if shutil.which("g++") is None:
        pytest.skip("no g++")
    so = str(tmp_path_factory.mktemp("host_kernels") / "libhost_kernels.so")
    # RX_HK_SAN=1 (set by test_kernels_on_the_host_are_clean_under_sanitizers, which runs this file once more in a child process with the
    # sanitizer runtimes preloaded): every simulated thread under AddressSanitizer and UndefinedBehaviorSanitizer
    mode = os.environ.get("RX_HK_SAN")
    san = ["-O1", "-fsanitize=address,undefined", "-fno-sanitize-recover=undefined"] if mode == "1" else ["-O1", "-fsanitize=thread"] if mode == "tsan" else ["-O2"]   # no -g: it doubles the build time
    # -DRX_VP_NO_ASYNC_D: the pair kernel's 1/D prefetch through registers (the cp.async form is inline PTX)
    cc = subprocess.run(["g++", "-std=c++17", "-pthread", "-DRX_VP_NO_ASYNC_D"] + san + ["-x", "c++", "-shared", "-fPIC", "-ffp-contract=off", "-I", os.path.join(ROOT, "tests"),
                         "-I", os.path.join(ROOT, "include"), "-I", os.path.join(PKG, "csrc"), os.path.join(ROOT, "tests", "host_kernels.cpp"), "-o", so],
                        stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert cc.returncode == 0, cc.stdout[-3000:]
    L = C.CDLL(so)
    L.hk_create.restype = C.c_void_p
    return L


class HostKernels:
    def __init__(self, L, case, pair=False, mesh=None, chem=None, opts=None):
        """From a golden case, or (case = None) from a flat mesh dict, a mechanism dict and an options dict."""
        pk = load_package()
        rs = importlib.import_module(pk.__name__ + ".rx_solver")
        self.L, self.case = L, case
        self.keep = []
        k = lambda a, dt=np.float64: (self.keep.append(np.ascontiguousarray(a, dtype=dt)), self.keep[-1])[1]
        _i = lambda a: a.ctypes.data_as(C.POINTER(C.c_int64))
        mesh = mesh_dict(case) if case is not None else mesh
        chem = chem_arrays(case) if case is not None else chem
        Ns = int(chem["Ns"])
        nPoint, nDom = int(mesh["nPoint"]), int(mesh["nPointDomain"])
        ch = rs._chem_struct(chem, self.keep)
        mks = mesh["markers"]
        arr = (rs._Marker * max(len(mks), 1))()
        for q, m in enumerate(mks):
            node = k(m["node"], np.int64)
            arr[q] = rs._Marker(int(m["kind"]), len(node), _i(node), _i(k(m["normalNeighbor"], np.int64)), _d(k(m["normal"])), float(m.get("T", 0.0)),
                                float(m.get("Vmag", 0.0)), (C.c_double * 2)(*m.get("dir", [0.0, 0.0])), _d(k(m.get("Ys", np.zeros(Ns)))), float(m.get("P", 0.0)),
                                float(m.get("Twall", 0.0)), float(m.get("heatFlux", 0.0)))
        en = k(mesh["edgeNodes"], np.int64)
        ms = rs._Mesh(nPoint, nDom, int(en.shape[0]), _i(en), _d(k(mesh["edgeNormal"])), _d(k(mesh["coord"])), _d(k(mesh["volume"])),
                      _i(k(mesh["nbPtr"], np.int64)), _i(k(mesh["nbPoint"], np.int64)), len(mks), arr)
        od = dict(secondOrder=0, limiter=0, limCoeff=0.3, refElemLength=0.1, timeStepping=0, unstCfl=0.0, deltaUnstTime=0.0, dualTime=0, deltaUnstTimeND=0.0,
                  viscLimiter=0, device=0)
        od.update(opts_dict(case) if case is not None else opts)
        o = rs._Opts(**{f[0]: od[f[0]] for f in rs._Opts._fields_})
        self.keep += [arr, ms, ch, o]
        self.h = C.c_void_p(L.hk_create(C.byref(ms), C.byref(ch), C.byref(o)))
        assert self.h
        L.hk_use_pair(self.h, 1 if pair else 0)
        self.nP, self.nVar, self.nPrim, self.nGrad = nPoint, Ns + 4, Ns + 7, Ns + 4

    def close(self):
        self.L.hk_destroy(self.h)

    def get(self, what, shape):
        out = np.empty(shape)
        self.L.hk_get(self.h, what, _d(out))
        return out


NAMES = ["jet_rans_perturbed", "channel_lam_rk", "channel_lam_gg", "channel_rans", "channel_lam_walls", "channel_lam_muscl", "channel_lam_bj", "channel_euler",
         "channel_rans_gg", "channel_lam_edge", "channel_lam_clip", "channel_rans_ign", "channel_air3_lam", "channel_air3_rans", "channel_mix5_lam",
         "channel_lam_inlet_total", "channel_rans_inlet_massflow", "channel_lam_supersonic", "channel_lam_visclim",
         "channel_lam_unsteady", "channel_euler_unsteady", "channel_euler_unsteady_fixed", "channel_lam_dualtime2", "channel_euler_dualtime1", "channel_rans_dualtime2"]


PAIR_NAMES = ["channel_lam_rk", "channel_lam_gg", "channel_lam_walls", "channel_lam_visclim", "channel_lam_edge", "channel_lam_clip", "channel_lam_muscl",
              "channel_rans", "channel_rans_gg", "channel_rans_ign", "channel_rans_inlet_massflow", "channel_rans_dualtime2", "jet_rans_perturbed"]


@pytest.mark.parametrize("name", PAIR_NAMES)
def test_two_lanes_per_edge_viscous_kernel_on_the_host(hk, name):
    """k_edge_visc_pair — the nine-species viscous kernel of the device, two lanes per edge — with its two lanes run as two host threads in lockstep
    (shuffles and votes become exchanges between the two, tests/cuda_host_shim.h): the same stage dumps, bit for bit, through whole iterations.
    Laminar and RANS fixtures (the SST closure's pivoted QR worked by the pair on one shared copy of the operator).  Two host threads do not
    advance together the way the lanes of a warp do, so every hand-over between the lanes must go through a __syncwarp or a shuffle: the run is
    race-free under ThreadSanitizer, which is how the one missing __syncwarp of the pair QR (before lane 0 swaps the column norms the other lane
    has just scanned) was found."""
    _run_case(hk, name, pair=True)


@pytest.mark.parametrize("name", NAMES)
def test_kernels_on_the_host_reproduce_the_reference_stage_dumps(hk, name):
    _run_case(hk, name, pair=False)


def _run_case(hk, name, pair):
    case = Case("%s/%s.npz" % (GOLDEN, name))
    cf = case.meta["cfg"]
    s = HostKernels(hk, case, pair=pair)
    nD = case.nPointDomain
    its = sorted({int(k[2:6]) for k in case.arr if k.startswith("it")})
    seen = set()

    def check(it, rk, what, a):
        key = "it%04d_rk%d_%s" % (it, rk, what)
        if not case.has(key):
            return
        ref = case[key].reshape(a.shape)
        rows = slice(0, nD) if what.startswith("res") or what in ("dt", "U1", "tres", "tpost") else slice(None)
        assert np.array_equal(a[rows], ref[rows]), "%s %s: %d entries differ, worst %.3e" % (name, key, int((a[rows] != ref[rows]).sum()), float(np.abs(a[rows] - ref[rows]).max()))
        seen.add(what)

    nrk = cf["nRK"] if cf["time_scheme"] == 1 else 1
    for it in its:
        U = np.ascontiguousarray(case.stage(it, 0, "U0")); T = np.ascontiguousarray(case.stage(it, 0, "prev_V")[:, 0]); tb = np.ascontiguousarray(case.stage(it, 0, "turb"))
        if case.has("it%04d_rk0_Uold" % it):   # Solution_Old at the start of the iteration: the non-physical fallback reads it
            Uo = np.ascontiguousarray(case.stage(it, 0, "Uold"))
            hk.hk_set_state(s.h, _d(Uo), _d(T), _d(tb)); hk.hk_set_old(s.h)
        hk.hk_set_state(s.h, _d(U), _d(T), _d(tb))
        if cf.get("dual_time", 0):
            a = np.ascontiguousarray(case.stage(it, 0, "Un")); b = np.ascontiguousarray(case.stage(it, 0, "Un1"))
            hk.hk_set_time_levels(s.h, _d(a), _d(b)); hk.hk_set_inner_iteration(s.h, C.c_int64(it % cf.get("dual_inner", 2)))
        for rk in range(nrk):   # the later stages start from the state the kernels themselves left
            hk.hk_preprocess(s.h, C.c_int64(it), 0)
            check(it, rk, "pre_V", s.get(0, (s.nP, s.nPrim))); check(it, rk, "pre_aux", s.get(1, (s.nP, 6)))
            if cf["viscous"]:
                check(it, rk, "pre_grad", s.get(2, (s.nP, s.nGrad, 2)))
            if cf.get("limiter", 0):
                check(it, rk, "pre_lim", s.get(3, (s.nP, 4)))
            if rk == 0:
                hk.hk_set_old(s.h); hk.hk_time_step(s.h)
                check(it, rk, "dt", s.get(4, (s.nP, 3)))
            hk.hk_residual(s.h, 1); check(it, rk, "res_conv", s.get(5, (s.nP, s.nVar)))
            hk.hk_residual(s.h, 2); check(it, rk, "res_visc", s.get(5, (s.nP, s.nVar)))
            hk.hk_residual(s.h, 3); check(it, rk, "res_src", s.get(5, (s.nP, s.nVar)))
            if cf.get("dual_time", 0):
                hk.hk_dual_time(s.h); check(it, rk, "res_dual", s.get(5, (s.nP, s.nVar)))
            hk.hk_bc(s.h); check(it, rk, "res_bc", s.get(5, (s.nP, s.nVar)))
            hk.hk_update(s.h, C.c_double(cf["rk_alpha"][rk] if nrk > 1 else 1.0))
            check(it, rk, "U1", s.get(6, (s.nP, s.nVar)))
        hk.hk_preprocess(s.h, C.c_int64(it), 1)   # the Output pass
        check(it, 0, "out_V", s.get(0, (s.nP, s.nPrim))); check(it, 0, "out_aux", s.get(1, (s.nP, 6)))
        if cf["rans"] and case.has("it%04d_rk0_tres" % it):
            # the SST solver's step on the state the flow iteration left: its assembled residual (Upwind, Viscous, Source, dual-time term, boundary
            # conditions), then its Postprocessing on the (k, omega) the reference's solve produced
            tpost = case.stage(it, 0, "tpost")
            kw = np.ascontiguousarray(tb[:, 1:3]); bl = np.ascontiguousarray(case.stage(it, 0, "tblend")); dist = np.ascontiguousarray(tpost[:, 2])
            kn = np.ascontiguousarray(case.stage(it, 0, "tUn")) if cf.get("dual_time", 0) else None
            kn1 = np.ascontiguousarray(case.stage(it, 0, "tUn1")) if cf.get("dual_time", 0) else None
            res = np.empty((s.nP, 2))
            hk.hk_turb_residual(s.h, _d(kw), _d(bl), _d(dist), C.c_double(cf["kine_inf"]), C.c_double(cf["omega_inf"]), _d(kn) if kn is not None else None,
                                _d(kn1) if kn1 is not None else None, _d(res))
            check(it, 0, "tres", res)
            post = np.empty((s.nP, 12))
            ksol = np.ascontiguousarray(case.stage(it, 0, "tsol"))
            hk.hk_turb_post(s.h, _d(ksol), _d(dist), C.c_double(float(tpost[0, 11])), _d(post))
            check(it, 0, "tpost", post)
    s.close()
    print(name, "stages compared bit for bit:", sorted(seen))
    assert {"pre_V", "pre_aux", "dt", "res_bc", "U1"} <= seen, seen
    if cf["viscous"]:
        assert "pre_grad" in seen
    if cf.get("dual_time", 0):
        assert "res_dual" in seen
    if cf["rans"]:
        assert {"tres", "tpost"} <= seen, seen


@pytest.mark.parametrize("rans,pair", [(False, False), (False, True), (True, False), (True, True)])
def test_kernels_on_the_host_equal_the_restatement_on_seeded_states(hk, rans, pair):
    """Beyond the fixtures: three Runge-Kutta stages on a 23 x 13 channel from seeded shear-layer states with pure-fuel and pure-oxidiser regions,
    trace species and temperatures from 500 to 1300 K (the bench's generator), five seeds — the kernels on the host against the CPU restatement
    (pinned to the reference on the fixtures), every stage bit for bit."""
    from rxcase import Oracle
    pk = load_package()
    chem = pk.load_mechanism()
    mesh = pk.channel_mesh(23, 13)
    opts = pk.default_opts(rans=rans, cfl=0.1)
    alphas = [0.66667, 0.66667, 1.0]
    for seed in (1, 2, 3, 4, 5):
        U, T, tb = pk.synthetic_state(mesh, chem, seed=seed, rans=rans, Tpeak=500.0 + 60.0 * seed)
        o = Oracle.from_mesh(mesh, chem, opts)
        s = HostKernels(hk, None, pair=pair, mesh=mesh, chem=chem, opts=opts)
        o.set_solution(U); o.set_temperature(T); o.set_turbulence(tb)
        hk.hk_set_state(s.h, _d(np.ascontiguousarray(U)), _d(np.ascontiguousarray(T)), _d(np.ascontiguousarray(tb)))
        for rk, a in enumerate(alphas):
            o.preprocess(1, False); hk.hk_preprocess(s.h, C.c_int64(1), 0)
            V, aux, g = o.primitive()
            assert np.array_equal(s.get(0, V.shape), V) and np.array_equal(s.get(1, aux.shape), aux) and np.array_equal(s.get(2, g.shape), g), (seed, rk, "primitives")
            if rk == 0:
                o.set_old_solution(); o.time_step(); hk.hk_set_old(s.h); hk.hk_time_step(s.h)
                assert np.array_equal(s.get(4, (s.nP, 3)), o.dt()), (seed, "dt")
            for what, call in ((1, o.upwind_residual), (2, o.viscous_residual), (3, o.source_residual)):
                call(); hk.hk_residual(s.h, what)
                assert np.array_equal(s.get(5, (s.nP, s.nVar)), o.residual()), (seed, rk, "residual after stage %d" % what)
            o.boundary_conditions(); hk.hk_bc(s.h)
            assert np.array_equal(s.get(5, (s.nP, s.nVar)), o.residual()), (seed, rk, "boundary conditions")
            o.explicit_update(a); hk.hk_update(s.h, C.c_double(a))
            assert np.array_equal(s.get(6, (s.nP, s.nVar)), o.solution()), (seed, rk, "update")
        s.close()


def test_kernels_on_the_host_are_clean_under_sanitizers():
    """The same runs in a child process with tests/host_kernels.cpp built under AddressSanitizer and UndefinedBehaviorSanitizer: every access of
    every simulated thread — structure-of-arrays indexing, incident-edge and neighbour lists, marker blocks, the thread-private matrices in the
    stand-in for shared memory — stays inside its array, on all fixtures."""
    if os.environ.get("RX_HK_SAN"):
        pytest.skip("this is the child run")
    if shutil.which("g++") is None:
        pytest.skip("no g++")
    libs = [subprocess.run(["g++", "-print-file-name=" + n], stdout=subprocess.PIPE, text=True).stdout.strip() for n in ("libasan.so", "libubsan.so")]
    if not all(os.path.isabs(p) and os.path.exists(p) for p in libs):
        pytest.skip("this g++ has no sanitizer runtimes")
    import sys
    env = dict(os.environ, RX_HK_SAN="1", LD_PRELOAD=" ".join(libs), ASAN_OPTIONS="detect_leaks=0:verify_asan_link_order=0")
    run = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-x", "-q", "-p", "no:cacheprovider", "-k",
                          "not jet and (reproduce_the_reference or (two_lanes and (channel_lam_rk or channel_rans] or channel_lam_visclim)))"],   # the 9000-point jet case is slow under ASan
                         cwd=os.path.dirname(os.path.abspath(__file__)), env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=2400)
    if "Shadow memory range interleaves" in run.stdout or "ReserveShadowMemoryRange failed" in run.stdout:
        pytest.skip("the AddressSanitizer runtime cannot start in this process layout: " + run.stdout[-300:])
    assert run.returncode == 0 and "AddressSanitizer" not in run.stdout and "runtime error" not in run.stdout and " passed" in run.stdout, run.stdout[-4000:]


def test_two_lane_kernel_is_race_free_between_its_synchronisation_points():
    """k_edge_visc_pair on the host under ThreadSanitizer (child process; one laminar, one RANS and one dual-time RANS fixture): the two lanes are two
    threads that meet only at the kernel's shuffles, votes and __syncwarp calls, so any word one lane writes and the other reads without such a
    meeting in between is reported — on the device the lanes of a warp advance together and would hide it."""
    if os.environ.get("RX_HK_SAN"):
        pytest.skip("this is the child run")
    if shutil.which("g++") is None:
        pytest.skip("no g++")
    lib = subprocess.run(["g++", "-print-file-name=libtsan.so"], stdout=subprocess.PIPE, text=True).stdout.strip()
    if not (os.path.isabs(lib) and os.path.exists(lib)):
        pytest.skip("this g++ has no ThreadSanitizer runtime")
    import sys
    env = dict(os.environ, RX_HK_SAN="tsan", LD_PRELOAD=lib, TSAN_OPTIONS="halt_on_error=0 exitcode=0")
    run = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-x", "-q", "-p", "no:cacheprovider", "-k",
                          "two_lanes and (channel_lam_rk or channel_rans_dualtime2 or channel_rans])"],
                         cwd=os.path.dirname(os.path.abspath(__file__)), env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=2400)
    if "FATAL: ThreadSanitizer" in run.stdout or "unexpected memory mapping" in run.stdout:
        pytest.skip("the ThreadSanitizer runtime cannot start in this process layout: " + run.stdout[-300:])
    assert run.returncode == 0 and " passed" in run.stdout, run.stdout[-4000:]
    assert "ThreadSanitizer: data race" not in run.stdout, run.stdout[-6000:]
