"""Host-side mirror of the reference's solver interface for the reactive residual path, over the C ABI (ctypes).

Class and method names follow SU2_CFD/include/solver_reactive.hpp (CReactiveNSSolver: Preprocessing, SetTime_Step,
Upwind_Residual, Viscous_Residual, Source_Residual, BC_*, ExplicitEuler_Iteration, ExplicitRK_Iteration, Set_OldSolution);
each method forwards to exactly one entry point of include/rx_cuda.h.  Arrays are numpy, array-of-structures, in SU2's
point numbering — the layout of CVariable::Solution / CSysVector.

No CPU fallback: if librx_cuda.so is missing, cannot be loaded, or there is no CUDA device, RxError is raised.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "librx_cuda.so")
CSRC = os.path.join(HERE, "csrc")
HOST_PARTITION = os.path.join(CSRC, "host", "rx_partition.cpp")   # host-only graph tools of the same library
HOST_GEOMETRY = os.path.join(CSRC, "host", "rx_geometry.cpp")     # host-only dual-mesh builder
HOST_SU2_MESH = os.path.join(CSRC, "host", "rx_su2_mesh.cpp")     # host-only .su2 reader + geometry preprocessing

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-fmad=false", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared"]


class RxError(RuntimeError):
    pass


def build_library(force=False, verbose=False):
    """nvcc cross-compiles the CUDA extension for sm_100a (works without a GPU)."""
    srcs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh"))] + [HOST_PARTITION, HOST_GEOMETRY, HOST_SU2_MESH]
    srcs.append(os.path.join(os.path.dirname(HERE), "include", "rx_cuda.h"))
    if not force and os.path.exists(LIB_PATH) and all(os.path.getmtime(LIB_PATH) >= os.path.getmtime(s) for s in srcs):
        return LIB_PATH
    cmd = ["nvcc"] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB_PATH, os.path.join(CSRC, "rx_cuda.cu"), HOST_PARTITION, HOST_GEOMETRY, HOST_SU2_MESH]
    subprocess.check_call(cmd)
    return LIB_PATH


_P = C.POINTER(C.c_double)
_PI = C.POINTER(C.c_int64)


class _Chem(C.Structure):
    _fields_ = [("Ns", C.c_int), ("Nr", C.c_int), ("nKnots", C.c_int)] + [
        (n, _P) for n in ["mMasses", "diffVol", "x", "y", "y2", "As", "Betas", "Ta", "reversible", "availBack", "AsB",
                          "BetasB", "TaB", "nuR", "nuP", "expR", "expP"]]


class _Marker(C.Structure):
    _fields_ = [("kind", C.c_int), ("nVertex", C.c_int), ("node", _PI), ("normalNeighbor", _PI), ("normal", _P),
                ("T", C.c_double), ("Vmag", C.c_double), ("dir", C.c_double * 2), ("Ys", _P), ("P", C.c_double),
                ("Twall", C.c_double), ("heatFlux", C.c_double)]


class _Mesh(C.Structure):
    _fields_ = [("nPoint", C.c_int64), ("nPointDomain", C.c_int64), ("nEdge", C.c_int64), ("edgeNodes", _PI),
                ("edgeNormal", _P), ("coord", _P), ("volume", _P), ("nbPtr", _PI), ("nbPoint", _PI),
                ("nMarker", C.c_int), ("markers", C.POINTER(_Marker))]


class _Opts(C.Structure):
    _fields_ = [("viscous", C.c_int), ("rans", C.c_int), ("leastSquares", C.c_int),
                ("mInfty", C.c_double), ("cfl", C.c_double), ("maxDt", C.c_double), ("Tmin", C.c_double), ("Tmax", C.c_double),
                ("clipT", C.c_int), ("ignition", C.c_int), ("ignitionIter", C.c_int64), ("ignitionT", C.c_double),
                ("fuelIndex", C.c_int), ("oxidizerIndex", C.c_int),
                ("prandtlLam", C.c_double), ("prandtlTurb", C.c_double), ("lewisTurb", C.c_double), ("cMu", C.c_double),
                ("pasrLb", C.c_double), ("tkeInf", C.c_double),
                ("secondOrder", C.c_int), ("limiter", C.c_int), ("limCoeff", C.c_double), ("refElemLength", C.c_double),
                ("device", C.c_int), ("timeStepping", C.c_int), ("unstCfl", C.c_double), ("deltaUnstTime", C.c_double),
                ("dualTime", C.c_int), ("deltaUnstTimeND", C.c_double), ("viscLimiter", C.c_int)]


class _TurbOpts(C.Structure):
    _fields_ = [("kineInf", C.c_double), ("omegaInf", C.c_double), ("cflRed", C.c_double), ("relax", C.c_double), ("linIter", C.c_int), ("linTol", C.c_double)]


class _RestartInfo(C.Structure):
    _fields_ = [("aoa", C.c_double), ("sideslip", C.c_double), ("bcThrust", C.c_double), ("dCD_dCL", C.c_double), ("extIter", C.c_int64)]


def merge_restart_pieces(pieces, path, header_from, ext_iter=0, aoa=0.0, sideslip=0.0, bc_thrust=4000.0, dcd_dcl=0.0):
    """Per-partition pieces (WriteRestart(piece=True)) -> one restart file ordered by PointID, as COutput::MergeSolution + SetRestart
    leave it on the master rank.  header_from: a solver of the run (for the column names)."""
    lines = []
    for p in pieces:
        with open(p, "rb") as f:
            lines += [l for l in f.read().split(b"\n") if l]
    lines.sort(key=lambda l: int(l.split(b"\t", 1)[0]))
    nCons = header_from.nVar + (2 if header_from.opts["rans"] else 0)
    h = '"PointID"\t"x"\t"y"' + "".join('\t"Conservative_%d"' % (v + 1) for v in range(nCons)) + '\t"Pressure"\t"Temperature"\t"Mach"'
    if header_from.opts["viscous"]:
        h += '\t"Laminar_Viscosity"'
    if header_from.opts["rans"]:
        h += '\t"<greek>m</greek><sub>t</sub>"'
    with open(path, "wb") as f:
        f.write(h.encode() + b"\n" + b"\n".join(lines) + b"\n")
        f.write(("AOA= %.15e\nSIDESLIP_ANGLE= %.15e\nINITIAL_BCTHRUST= %.15e\nDCD_DCL_VALUE= %.15e\nEXT_ITER= %d\n" % (aoa, sideslip, bc_thrust, dcd_dcl, ext_iter)).encode())


EXPORTS = ["rx_supported_species", "rx_create", "rx_destroy", "rx_last_error", "rx_set_stream", "rx_synchronize",
           "rx_set_solution", "rx_set_temperature", "rx_set_turbulence", "rx_get_solution", "rx_get_primitive", "rx_get_dt",
           "rx_get_residual", "rx_preprocess", "rx_set_old_solution", "rx_time_step", "rx_residual_conv", "rx_residual_visc",
           "rx_residual_source", "rx_residual_all", "rx_kernel_time", "rx_bc", "rx_update_explicit", "rx_iterate", "rx_source_points", "rx_source_points_dev",
           "rx_halo_pack", "rx_halo_unpack", "rx_halo_width", "rx_launch_count", "rx_fp64_peak", "rx_bc_marker", "rx_get_bc_state", "rx_get_limiter",
           "rx_partition_kway", "rx_partition_backend", "rx_partition_mesh", "rx_part_free", "rx_rcm_order", "rx_get_residual_max", "rx_get_edge_order",
           "rx_set_wall_distance", "rx_turb_postprocess", "rx_get_turbulence", "rx_wall_distance", "rx_compute_wall_distance",
           "rx_residual_rms_static", "rx_get_status", "rx_libm_eval", "rx_qr_eval", "rx_state_snapshot", "rx_restart_write", "rx_restart_read", "rx_tecplot_write", "rx_set_cfl", "rx_dual_mesh_tri", "rx_dual_mesh_free", "rx_su2_mesh_read", "rx_su2_mesh_build", "rx_su2_mesh_free", "rx_su2_mesh_error",
           "rx_set_time_levels", "rx_get_time_levels", "rx_shift_time_levels", "rx_set_inner_iteration", "rx_get_delta_unst_time", "rx_residual_dual_time",
           "rx_turb_set_time_levels", "rx_turb_get_time_levels", "rx_turb_shift_time_levels",
           "rx_turb_set_state", "rx_turb_iterate", "rx_turb_get_step",
           "rx_comm_unique_id", "rx_comm_init", "rx_halo_exchange", "rx_comm_allreduce_rms"]

_lib = None


def load_library():
    """dlopen librx_cuda.so (built in-tree by build_library / __graft_entry__.build); raises RxError if absent."""
    global _lib
    if _lib is None:
        path = os.environ.get("RX_CUDA_LIB", LIB_PATH)   # tuning hook: a build variant of the same library (tools/ktime.py)
        if not os.path.exists(path):
            raise RxError("librx_cuda.so not built (run __graft_entry__.build()); there is no CPU fallback")
        try:
            _lib = C.CDLL(path)
        except OSError as e:
            raise RxError("cannot load %s: %s (no CPU fallback)" % (path, e))
        _lib.rx_last_error.restype = C.c_char_p
        _lib.rx_last_error.argtypes = [C.c_void_p]
        _lib.rx_launch_count.restype = C.c_int64
        _lib.rx_launch_count.argtypes = [C.c_void_p]
    return _lib


def _d(a):
    return a.ctypes.data_as(_P)


def _i(a):
    return a.ctypes.data_as(_PI)


def _c(a, dt=np.float64):
    return np.ascontiguousarray(a, dtype=dt)


def _chem_struct(chem, keep):
    """rx_chem over numpy arrays (kept alive in `keep`)."""
    k = lambda a, dt=np.float64: (keep.append(_c(a, dt)), keep[-1])[1]
    Ns, Nr = int(chem["Ns"]), int(chem["Nr"])
    pad = lambda a, n: k(np.asarray(a, dtype=np.float64).reshape(-1) if n else np.zeros(1))
    return _Chem(Ns, Nr, int(chem["nKnots"]), _d(k(chem["mMasses"])), _d(k(chem["diffVol"])), _d(k(chem["x"])),
                 _d(k(chem["y"])), _d(k(chem["y2"])),
                 *[_d(pad(chem[n], Nr)) for n in ["As", "Betas", "Ta", "reversible", "availBack", "AsB", "BetasB", "TaB"]],
                 *[_d(pad(chem[n], Nr * Ns)) for n in ["nuR", "nuP", "expR", "expP"]])


def source_points(chem, T, rho, Y, vol, omegaTurb=None, rans=False, cMu=0.09, pasrLb=0.2, device=0):
    """Stand-alone Arrhenius (+ PaSR when rans) source over independent points, host arrays in and out
    (CSourceReactive::ComputeChemistry, numerics_direct_reactive.cpp:1728-1825): returns res[n][Ns] = -omega_s * vol."""
    L = load_library()
    keep = []
    ch = _chem_struct(chem, keep)
    T = _c(T); rho = _c(rho); Y = _c(Y); vol = _c(vol)
    n = T.shape[0]
    res = np.empty((n, int(chem["Ns"])))
    om = _c(omegaTurb) if omegaTurb is not None else None
    rc = L.rx_source_points(C.byref(ch), int(bool(rans)), C.c_double(cMu), C.c_double(pasrLb), C.c_int64(n), _d(T), _d(rho), _d(Y),
                            _d(om) if om is not None else None, _d(vol), _d(res), int(device))
    if rc != 0:
        raise RxError("rx_source_points failed (%d): %s" % (rc, L.rx_last_error(None).decode()))
    return res


def libm_eval(which, x, y=None, device=0):
    """The device's glibc-exact exp ('exp'), pow ('pow', needs y) or log ('log') over host arrays (rx_libm_eval; csrc/rx_libm.cuh)."""
    L = load_library()
    w = {"exp": 0, "pow": 1, "log": 2}[which]
    x = _c(x); out = np.empty_like(x)
    yy = _c(y) if y is not None else None
    rc = L.rx_libm_eval(w, C.c_int64(x.size), _d(x), _d(yy) if yy is not None else None, _d(out), int(device))
    if rc != 0:
        raise RxError("rx_libm_eval failed (%d): %s" % (rc, L.rx_last_error(None).decode()))
    return out


def qr_eval(A, b, device=0):
    """Eigen's pivoted-QR solve of n 9 x 9 systems on the device (rx_qr_eval): A [n, 81] column-major, b [n, 2, 9] -> (xSerial, xPair)."""
    L = load_library()
    A = _c(A); b = _c(b)
    n = A.size // 81
    xs = np.empty_like(b); xp = np.empty_like(b)
    rc = L.rx_qr_eval(C.c_int64(n), _d(A), _d(b), _d(xs), _d(xp), int(device))
    if rc != 0:
        raise RxError("rx_qr_eval failed (%d): %s" % (rc, L.rx_last_error(None).decode()))
    return xs, xp


class _DualMesh(C.Structure):
    _fields_ = [("nEdge", C.c_int64), ("edgeNodes", C.POINTER(C.c_int64)), ("edgeNormal", C.POINTER(C.c_double)), ("volume", C.POINTER(C.c_double)),
                ("nbPtr", C.POINTER(C.c_int64)), ("nbPoint", C.POINTER(C.c_int64))]


def dual_mesh_tri(coord, tri):
    """Median-dual mesh of a triangulation by the library's host builder (rx_dual_mesh_tri): dict with nEdge, edgeNodes, edgeNormal,
    volume, nbPtr, nbPoint (numpy copies; the library's buffers are released)."""
    L = load_library()
    coord = _c(coord); tri = _c(tri, np.int64)
    nP, nT = coord.shape[0], tri.shape[0]
    dm = _DualMesh()
    L.rx_dual_mesh_free.restype = None
    rc = L.rx_dual_mesh_tri(C.c_int64(nP), _d(coord), C.c_int64(nT), _i(tri), C.byref(dm))
    if rc != 0:
        raise RxError("rx_dual_mesh_tri failed (%d)" % rc)
    try:
        nE = int(dm.nEdge)
        A = lambda p, n, shape: np.ctypeslib.as_array(p, shape=(n,)).reshape(shape).copy()
        return dict(nEdge=nE, edgeNodes=A(dm.edgeNodes, 2 * nE, (nE, 2)), edgeNormal=A(dm.edgeNormal, 2 * nE, (nE, 2)), volume=A(dm.volume, nP, (nP,)),
                    nbPtr=A(dm.nbPtr, nP + 1, (nP + 1,)), nbPoint=A(dm.nbPoint, 2 * nE, (2 * nE,)))
    finally:
        L.rx_dual_mesh_free(C.byref(dm))


def comm_unique_id():
    """ncclGetUniqueId through the library (rx_comm_unique_id): 128 bytes for rank 0 to hand to the other ranks."""
    L = load_library()
    buf = (C.c_char * 128)()
    rc = L.rx_comm_unique_id(buf)
    if rc != 0:
        raise RxError("rx_comm_unique_id failed (%d): %s" % (rc, L.rx_last_error(None).decode()))
    return bytes(buf)


STATIC_RMS_MAX_POINTS, STATIC_RMS_MAX_VARS = 4194304, 16   # include/rx_cuda.h: RX_STATIC_RMS_MAX_*


def residual_rms_static(R_dev_ptr, nPointDomain, ld, nVar, nPointGlobal, rms_dev_ptr, device=0, stream=0):
    """rx_residual_rms_static: residual norms of a device-resident SoA residual R_dev[nVar][ld] into rms_dev[nVar] without any
    allocation (static device scratch; no counterpart in the reference, see include/rx_cuda.h).  Device pointers as integers
    (e.g. torch tensor.data_ptr()); enqueued on `stream`, the host is not synchronised."""
    L = load_library()
    rc = L.rx_residual_rms_static(C.c_void_p(R_dev_ptr), C.c_int64(nPointDomain), C.c_int64(ld), int(nVar), C.c_int64(nPointGlobal),
                                  C.c_void_p(rms_dev_ptr), int(device), C.c_void_p(stream))
    if rc != 0:
        raise RxError("rx_residual_rms_static failed (%d): %s" % (rc, L.rx_last_error(None).decode()))


class CReactiveNSSolverGPU:
    """Mirror of CReactiveNSSolver (SU2_CFD/include/solver_reactive.hpp:398) whose hot virtuals run on the GPU.

    mesh  : dict(nPoint, nPointDomain, edgeNodes[nE,2], edgeNormal[nE,2], coord[nP,2], volume[nP], nbPtr, nbPoint,
                 markers=[dict(kind, node, normalNeighbor, normal, T, Vmag, dir, Ys, P, Twall)])   (flat CGeometry export)
    chem  : dict of ReactingModelLibrary tables (see include/rx_cuda.h: rx_chem)
    opts  : dict of CConfig-derived constants (rx_opts)
    """

    def __init__(self, mesh, chem, opts, device=0):
        L = self.lib = load_library()
        self._keep = []
        k = lambda a, dt=np.float64: (self._keep.append(_c(a, dt)), self._keep[-1])[1]
        Ns, Nr = int(chem["Ns"]), int(chem["Nr"])
        pad = lambda a, n: k(np.asarray(a, dtype=np.float64).reshape(-1) if n else np.zeros(1))
        ch = _Chem(Ns, Nr, int(chem["nKnots"]), _d(k(chem["mMasses"])), _d(k(chem["diffVol"])), _d(k(chem["x"])),
                   _d(k(chem["y"])), _d(k(chem["y2"])),
                   *[_d(pad(chem[n], Nr)) for n in ["As", "Betas", "Ta", "reversible", "availBack", "AsB", "BetasB", "TaB"]],
                   *[_d(pad(chem[n], Nr * Ns)) for n in ["nuR", "nuP", "expR", "expP"]])
        mks = mesh.get("markers", [])
        arr = (_Marker * max(len(mks), 1))()
        for q, m in enumerate(mks):
            node = k(m["node"], np.int64)
            arr[q] = _Marker(int(m["kind"]), len(node), _i(node), _i(k(m["normalNeighbor"], np.int64)), _d(k(m["normal"])),
                             float(m.get("T", 0.0)), float(m.get("Vmag", 0.0)), (C.c_double * 2)(*m.get("dir", [0.0, 0.0])),
                             _d(k(m.get("Ys", np.zeros(Ns)))), float(m.get("P", 0.0)), float(m.get("Twall", 0.0)), float(m.get("heatFlux", 0.0)))
        en = k(mesh["edgeNodes"], np.int64)
        self.nPoint, self.nPointDomain, self.nEdge = int(mesh["nPoint"]), int(mesh["nPointDomain"]), int(en.shape[0])
        ms = _Mesh(self.nPoint, self.nPointDomain, self.nEdge, _i(en), _d(k(mesh["edgeNormal"])), _d(k(mesh["coord"])),
                   _d(k(mesh["volume"])), _i(k(mesh["nbPtr"], np.int64)), _i(k(mesh["nbPoint"], np.int64)), len(mks), arr)
        od = dict(secondOrder=0, limiter=0, limCoeff=0.3, refElemLength=0.1, timeStepping=0, unstCfl=0.0, deltaUnstTime=0.0, dualTime=0, deltaUnstTimeND=0.0, viscLimiter=0)
        od.update(opts)
        od["device"] = device
        self.opts = od
        o = _Opts(**{f[0]: od[f[0]] for f in _Opts._fields_})
        self.Ns, self.nVar, self.nPrim, self.nGrad = Ns, Ns + 4, Ns + 7, Ns + 4
        ctx = C.c_void_p()
        rc = L.rx_create(C.byref(ctx), C.byref(ms), C.byref(ch), C.byref(o))
        if rc != 0:
            raise RxError("rx_create failed (%d): %s" % (rc, L.rx_last_error(None).decode()))
        self.ctx = ctx
        self.has_comm = False

    def _ck(self, rc):
        if rc != 0:
            raise RxError("rx call failed (%d): %s" % (rc, self.lib.rx_last_error(self.ctx).decode()))

    def close(self):
        if getattr(self, "ctx", None):
            self.lib.rx_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- mirrors of node[]->Solution / Primitive / LinSysRes
    def SetSolution(self, U): self._ck(self.lib.rx_set_solution(self.ctx, _d(_c(U))))

    def SetSolutionPtr(self, host_ptr):
        """Same, from a raw host address (e.g. a pinned torch tensor's data_ptr())."""
        self._ck(self.lib.rx_set_solution(self.ctx, C.c_void_p(host_ptr)))
    def SetTemperature(self, T): self._ck(self.lib.rx_set_temperature(self.ctx, _d(_c(T))))
    def SetTurbulence(self, t6): self._ck(self.lib.rx_set_turbulence(self.ctx, _d(_c(t6))))

    def GetSolution(self):
        U = np.empty((self.nPoint, self.nVar)); self._ck(self.lib.rx_get_solution(self.ctx, _d(U))); return U

    def GetPrimitive(self):
        V = np.empty((self.nPoint, self.nPrim)); aux = np.empty((self.nPoint, 6)); g = np.empty((self.nPoint, self.nGrad, 2))
        self._ck(self.lib.rx_get_primitive(self.ctx, _d(V), _d(aux), _d(g))); return V, aux, g

    def GetLimiter(self):
        L = np.empty((self.nPoint, 4)); self._ck(self.lib.rx_get_limiter(self.ctx, _d(L))); return L

    def GetDelta_Time(self):
        d = np.empty((self.nPoint, 3)); self._ck(self.lib.rx_get_dt(self.ctx, _d(d))); return d

    def GetLinSysRes(self):
        R = np.empty((self.nPoint, self.nVar)); self._ck(self.lib.rx_get_residual(self.ctx, _d(R))); return R

    # ---- CSolver virtuals (solver_reactive.hpp:127-363,441-554)
    def Preprocessing(self, ExtIter, Output=False, count_nonphysical=True):
        n = C.c_int64(0)
        self._ck(self.lib.rx_preprocess(self.ctx, C.c_int64(ExtIter), int(Output), C.byref(n) if count_nonphysical else None))
        return n.value

    def Set_OldSolution(self): self._ck(self.lib.rx_set_old_solution(self.ctx))
    def SetTime_Step(self): self._ck(self.lib.rx_time_step(self.ctx))
    def Upwind_Residual(self): self._ck(self.lib.rx_residual_conv(self.ctx))
    def Viscous_Residual(self): self._ck(self.lib.rx_residual_visc(self.ctx))
    def Source_Residual(self): self._ck(self.lib.rx_residual_source(self.ctx))
    def Residual_All(self): self._ck(self.lib.rx_residual_all(self.ctx))
    def BC_All(self): self._ck(self.lib.rx_bc(self.ctx))
    def BC_Marker(self, marker): self._ck(self.lib.rx_bc_marker(self.ctx, int(marker)))

    def KernelTime(self, which, reps=5):
        ms = C.c_double(0.0)
        self._ck(self.lib.rx_kernel_time(self.ctx, int(which), int(reps), C.byref(ms)))
        return ms.value

    def ExplicitRK_Iteration(self, alpha, want_rms=True):
        rms = np.empty(self.nVar)
        self._ck(self.lib.rx_update_explicit(self.ctx, C.c_double(alpha), _d(rms) if want_rms else None))
        return rms if want_rms else None

    def GetStatus(self):
        """[non-physical points of the last Preprocessing, NaN conv, NaN visc, NaN source, failed temperature solves]; raises RxError
        with the reference's message if any of the last four is set (rx_get_status)."""
        st = np.zeros(5, dtype=np.int64)
        self._ck(self.lib.rx_get_status(self.ctx, _i(st)))
        return st

    def GetRes_Max(self):
        """(max |LinSysRes| per variable, local index of the first point attaining it) of the last update with norms —
        CSolver::GetRes_Max / GetPoint_Max."""
        m = np.empty(self.nVar); ip = np.empty(self.nVar, dtype=np.int64)
        self._ck(self.lib.rx_get_residual_max(self.ctx, _d(m), _i(ip)))
        return m, ip

    def SetWall_Distance(self, dist):
        """geometry->node[]->GetWall_Distance() of every point, for TurbPostprocessing."""
        self._ck(self.lib.rx_set_wall_distance(self.ctx, _d(_c(dist))))

    def ComputeWall_Distance(self):
        """CPhysicalGeometry::ComputeWall_Distance on the device, from the context's own isothermal / heat-flux markers."""
        self._ck(self.lib.rx_compute_wall_distance(self.ctx))

    def TurbPostprocessing(self, kOmega=None, sigma_k=0.85, stage=0):
        """CTurbSSTSolver::Postprocessing on the device (rx_turb_postprocess): gradients of (k, omega), blending functions, eddy
        viscosity; fills the turbulence fields the flow path reads.  stage 1 / 2: the two halves around the halo exchange."""
        kw = _c(kOmega) if kOmega is not None else None
        self._ck(self.lib.rx_turb_postprocess(self.ctx, _d(kw) if kw is not None else None, C.c_double(sigma_k), int(stage)))

    def TurbSetState(self, kOmega=None, blend3=None):
        """(k, omega) of the turbulence solver and, optionally, F1, F2, CDkw of its previous Postprocessing (rx_turb_set_state)."""
        a = _c(kOmega) if kOmega is not None else None; b = _c(blend3) if blend3 is not None else None
        self._ck(self.lib.rx_turb_set_state(self.ctx, _d(a) if a is not None else None, _d(b) if b is not None else None))

    def TurbIterate(self, kine_inf, omega_inf, cfl_red=1.0, relax=1.0, lin_iter=5, lin_tol=1e-6):
        """One SingleGrid_Iteration of the SST solver on the device (rx_turb_iterate); returns (residual norms of k and omega, FGMRES iterations)."""
        o = _TurbOpts(kine_inf, omega_inf, cfl_red, relax, int(lin_iter), lin_tol)
        rms = np.empty(2); n = C.c_int(0)
        self._ck(self.lib.rx_turb_iterate(self.ctx, C.byref(o), _d(rms), C.byref(n)))
        return rms, n.value

    def TurbGetStep(self):
        """(LinSysRes before the solve, solution of the linear system) of the last TurbIterate, [nPoint][2] each."""
        r = np.empty((self.nPoint, 2)); x = np.empty((self.nPoint, 2))
        self._ck(self.lib.rx_turb_get_step(self.ctx, _d(r), _d(x)))
        return r, x

    def GetTurbulence(self):
        """[nPoint][12]: k, omega, dist, F1, F2, CDkw, muT, dk/dx, dk/dy, domega/dx, domega/dy, sigma_k."""
        out = np.empty((self.nPoint, 12))
        self._ck(self.lib.rx_get_turbulence(self.ctx, _d(out)))
        return out

    def EdgeOrder(self):
        """(order[nEdge] int32, key[nEdge] uint8): processing order of the viscous edge kernel and the per-edge keys it is
        built from (rx_get_edge_order; csrc/rx_order.cuh)."""
        order = np.empty(self.nEdge, dtype=np.int32); key = np.empty(self.nEdge, dtype=np.uint8)
        self._ck(self.lib.rx_get_edge_order(self.ctx, order.ctypes.data_as(C.POINTER(C.c_int32)), key.ctypes.data_as(C.POINTER(C.c_uint8))))
        return order, key

    def ExplicitEuler_Iteration(self, want_rms=True):
        return self.ExplicitRK_Iteration(1.0, want_rms)

    def Iterate(self, ExtIter, rk_alpha=(1.0,), want_rms=True):
        """One MultiGrid_Cycle + the Output preprocessing pass, all on the device (integration_time.cpp:42-181)."""
        a = _c(rk_alpha); rms = np.empty(self.nVar)
        self._ck(self.lib.rx_iterate(self.ctx, C.c_int64(ExtIter), len(a), _d(a), _d(rms) if want_rms else None))
        return rms if want_rms else None

    def CommInit(self, uid, rank, nranks, plan):
        """rx_comm_init: NCCL communicator + this partition's neighbour lists (plan["send"] / plan["recv"]: {neighbour rank: local point ids})."""
        nbs = sorted(set(int(q) for q in plan["send"]) | set(int(q) for q in plan["recv"]))
        sp, rp, spts, rpts = [0], [0], [], []
        for q in nbs:
            a = np.asarray(plan["send"].get(q, plan["send"].get(str(q), [])), dtype=np.int64); b = np.asarray(plan["recv"].get(q, plan["recv"].get(str(q), [])), dtype=np.int64)
            spts.append(a); rpts.append(b); sp.append(sp[-1] + a.size); rp.append(rp[-1] + b.size)
        cat = lambda l: _c(np.concatenate(l) if l else np.zeros(0), np.int64)
        nb = _c(np.array(nbs, dtype=np.int32), np.int32)
        spts, rpts, sp, rp = cat(spts), cat(rpts), _c(sp, np.int64), _c(rp, np.int64)
        self._ck(self.lib.rx_comm_init(self.ctx, C.c_char_p(uid), int(rank), int(nranks), len(nbs), nb.ctypes.data_as(C.POINTER(C.c_int)),
                                       _i(sp), _i(spts), _i(rp), _i(rpts)))
        self.has_comm = True

    def HaloExchange(self, what): self._ck(self.lib.rx_halo_exchange(self.ctx, int(what)))

    def CommAllreduceRms(self):
        rms = np.empty(self.nVar); self._ck(self.lib.rx_comm_allreduce_rms(self.ctx, _d(rms))); return rms

    def HaloWidth(self, what): return int(self.lib.rx_halo_width(self.ctx, int(what)))

    def HaloPack(self, what, n, ids_dev_ptr, buf_dev_ptr):
        self._ck(self.lib.rx_halo_pack(self.ctx, int(what), C.c_int64(n), C.c_void_p(ids_dev_ptr), C.c_void_p(buf_dev_ptr)))

    def HaloUnpack(self, what, n, ids_dev_ptr, buf_dev_ptr):
        self._ck(self.lib.rx_halo_unpack(self.ctx, int(what), C.c_int64(n), C.c_void_p(ids_dev_ptr), C.c_void_p(buf_dev_ptr)))

    def SourcePointsDev(self, n, T_ptr, rho_ptr, Y_ptr, omegaTurb_ptr, vol_ptr, res_ptr):
        """Device-pointer form of source_points (SoA Y[Ns][n], res[Ns][n]); constants (rans, C_mu, PaSR bound) are the context's."""
        v = lambda p: C.c_void_p(p) if p else None
        self._ck(self.lib.rx_source_points_dev(self.ctx, C.c_int64(n), v(T_ptr), v(rho_ptr), v(Y_ptr), v(omegaTurb_ptr), v(vol_ptr), v(res_ptr)))

    def WriteRestart(self, path, kOmega=None, globalIndex=None, ext_iter=0, piece=False, aoa=0.0, sideslip=0.0, bc_thrust=4000.0, dcd_dcl=0.0):
        """COutput::SetRestart from the device state (rx_restart_write): the SU2 native ASCII restart file."""
        info = _RestartInfo(aoa, sideslip, bc_thrust, dcd_dcl, int(ext_iter))
        kw = _c(kOmega) if kOmega is not None else None
        gi = _c(globalIndex, np.int64) if globalIndex is not None else None
        self._ck(self.lib.rx_restart_write(self.ctx, os.fsencode(path), _i(gi) if gi is not None else None, _d(kw) if kw is not None else None,
                                           C.byref(info), int(bool(piece))))

    def WriteTecplot(self, path, triangles, kOmega=None, globalIndex=None, quads=None, zone_prefix=None):
        """COutput::SetTecplotASCII (volume solution) from the device state (rx_tecplot_write); triangles [nTria][3], quads [nQuad][4], 0-based."""
        kw = _c(kOmega) if kOmega is not None else None
        gi = _c(globalIndex, np.int64) if globalIndex is not None else None
        tr = _c(triangles if triangles is not None else np.zeros((0, 3)), np.int64).reshape(-1, 3)
        qd = _c(quads if quads is not None else np.zeros((0, 4)), np.int64).reshape(-1, 4)
        self._ck(self.lib.rx_tecplot_write(self.ctx, os.fsencode(path), _i(gi) if gi is not None else None, _d(kw) if kw is not None else None,
                                           C.c_int64(tr.shape[0]), _i(tr) if tr.shape[0] else None, C.c_int64(qd.shape[0]), _i(qd) if qd.shape[0] else None,
                                           zone_prefix.encode() if zone_prefix else None))

    def LoadRestart(self, path, globalIndex=None, nGlobal=0):
        """CReactiveNSSolver::Load_Restart + the turbulence reader (rx_restart_read): sets the conserved state from the file; returns
        ((k, omega)[nPoint][2] or None, dict of the trailer values)."""
        info = _RestartInfo()
        kw = np.zeros((self.nPoint, 2)) if self.opts["rans"] else None
        gi = _c(globalIndex, np.int64) if globalIndex is not None else None
        self._ck(self.lib.rx_restart_read(self.ctx, os.fsencode(path), _i(gi) if gi is not None else None, C.c_int64(nGlobal),
                                          _d(kw) if kw is not None else None, C.byref(info)))
        return kw, dict(aoa=info.aoa, sideslip=info.sideslip, bcThrust=info.bcThrust, dCD_dCL=info.dCD_dCL, extIter=info.extIter)

    def SetCFL(self, cfl):
        """config->SetCFL between iterations (CFL_ADAPT): the value SetTime_Step reads from now on (rx_set_cfl)."""
        self._ck(self.lib.rx_set_cfl(self.ctx, C.c_double(cfl)))
        self.opts["cfl"] = float(cfl)

    # ---- dual time stepping (rx_opts.dualTime): stored time levels, the pseudo-time iteration index, the physical-time source term
    def SetTimeLevels(self, Un, Un1):
        Un = np.ascontiguousarray(Un, dtype=np.float64); Un1 = np.ascontiguousarray(Un1, dtype=np.float64)
        self._ck(self.lib.rx_set_time_levels(self.ctx, _d(Un), _d(Un1)))

    def GetTimeLevels(self):
        a = np.empty((self.nPoint, self.nVar)); b = np.empty((self.nPoint, self.nVar))
        self._ck(self.lib.rx_get_time_levels(self.ctx, _d(a), _d(b)))
        return a, b

    def SetDualTime_Solver(self):
        """CIntegration::SetDualTime_Solver (integration_structure.cpp:568-584): close a physical time step, U^(n-1) <- U^n <- U."""
        self._ck(self.lib.rx_shift_time_levels(self.ctx))

    def TurbSetTimeLevels(self, kwN, kwN1):
        a = np.ascontiguousarray(kwN, dtype=np.float64); b = np.ascontiguousarray(kwN1, dtype=np.float64)
        self._ck(self.lib.rx_turb_set_time_levels(self.ctx, _d(a), _d(b)))

    def TurbGetTimeLevels(self):
        a = np.empty((self.nPoint, 2)); b = np.empty((self.nPoint, 2))
        self._ck(self.lib.rx_turb_get_time_levels(self.ctx, _d(a), _d(b)))
        return a, b

    def TurbSetDualTime_Solver(self): self._ck(self.lib.rx_turb_shift_time_levels(self.ctx))

    def SetInnerIteration(self, it): self._ck(self.lib.rx_set_inner_iteration(self.ctx, C.c_int64(it)))

    def GetDelta_UnstTimeND(self):
        v = C.c_double(0.0); self._ck(self.lib.rx_get_delta_unst_time(self.ctx, C.byref(v))); return v.value

    def SetResidual_DualTime(self): self._ck(self.lib.rx_residual_dual_time(self.ctx))

    def SaveState(self): self._ck(self.lib.rx_state_snapshot(self.ctx, 0))
    def RestoreState(self): self._ck(self.lib.rx_state_snapshot(self.ctx, 1))

    def GetSolutionPtr(self, host_ptr):
        """rx_get_solution into a raw host address (e.g. a pinned torch tensor's data_ptr())."""
        self._ck(self.lib.rx_get_solution(self.ctx, C.c_void_p(host_ptr)))

    def Synchronize(self): self._ck(self.lib.rx_synchronize(self.ctx))
    def LaunchCount(self): return int(self.lib.rx_launch_count(self.ctx))

    def Fp64Peak(self):
        """Measured FP64 TFLOP/s of the device (DFMA loop), the compute roof of the viscous / source kernels."""
        v = C.c_double(0.0)
        self._ck(self.lib.rx_fp64_peak(self.ctx, C.byref(v)))
        return float(v.value)

    # ---- the stage API used by the parity tests (same names as tests/rxcase.Oracle)
    set_solution = SetSolution; set_temperature = SetTemperature; set_turbulence = SetTurbulence
    solution = GetSolution; primitive = GetPrimitive; dt = GetDelta_Time; residual = GetLinSysRes; limiter = GetLimiter
    def preprocess(self, it, output=False): return self.Preprocessing(it, output)
    set_old_solution = Set_OldSolution; time_step = SetTime_Step; upwind_residual = Upwind_Residual
    viscous_residual = Viscous_Residual; source_residual = Source_Residual; boundary_conditions = BC_All
    set_time_levels = SetTimeLevels; time_levels = GetTimeLevels; shift_time_levels = SetDualTime_Solver
    set_inner_iteration = SetInnerIteration; dual_time_residual = SetResidual_DualTime
    def explicit_update(self, alpha=1.0): return self.ExplicitRK_Iteration(alpha)
