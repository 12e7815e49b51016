"""Test helpers: load a golden case (dump of the reference's own classes, see oracle/ref_harness.cpp and
oracle/make_golden.py) and drive the CPU oracle (oracle/rx_oracle.cpp) through ctypes.

TEST INFRASTRUCTURE ONLY — nothing here is on the product path.
"""
import ctypes as C
import json
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
ORACLE_SRC = os.path.join(ROOT, "oracle", "rx_oracle.cpp")
ORACLE_SO = os.path.join(ROOT, "oracle", "_build", "librx_oracle.so")

KIND_INLET, KIND_OUTLET, KIND_ISOTHERMAL, KIND_HEATFLUX, KIND_EULER_WALL, KIND_INLET_TOTAL, KIND_INLET_MASSFLOW, KIND_SUPERSONIC_INLET, KIND_SUPERSONIC_OUTLET = 1, 2, 3, 4, 5, 6, 7, 8, 9
_BC_KIND = {"inlet": KIND_INLET, "outlet": KIND_OUTLET, "isothermal": KIND_ISOTHERMAL, "heatflux": KIND_HEATFLUX, "euler_wall": KIND_EULER_WALL,
            "supersonic_inlet": KIND_SUPERSONIC_INLET, "supersonic_outlet": KIND_SUPERSONIC_OUTLET}


def build_oracle(force=False):
    """Compile the C++ restatement (gcc only, no FMA contraction: x86-64 baseline ISA like the reference build)."""
    if force or not os.path.exists(ORACLE_SO) or os.path.getmtime(ORACLE_SO) < os.path.getmtime(ORACLE_SRC):
        os.makedirs(os.path.dirname(ORACLE_SO), exist_ok=True)
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-w",
                               "-o", ORACLE_SO, ORACLE_SRC])
    return ORACLE_SO


class Case:
    """A golden case: meta (json), geometry + chemistry tables, and per-iteration stage dumps."""

    def __init__(self, path):
        if os.path.isdir(path):
            self.meta = json.load(open(os.path.join(path, "meta.json")))
            self.arr = {f[:-4]: np.load(os.path.join(path, f)) for f in os.listdir(path) if f.endswith(".npy")}
        else:
            z = np.load(path, allow_pickle=False)
            self.meta = json.loads(bytes(z["meta_json"]).decode())
            self.arr = {k: z[k] for k in z.files if k != "meta_json"}
        m = self.meta
        self.nPoint, self.nPointDomain, self.nEdge = m["nPoint"], m["nPointDomain"], m["nEdge"]
        self.Ns, self.Nr = m["chem"]["Ns"], m["chem"]["Nr"]
        self.nVar, self.nPrim, self.nGrad = self.Ns + 4, self.Ns + 7, self.Ns + 4

    def __getitem__(self, k):
        return self.arr[k]

    def has(self, k):
        return k in self.arr

    def stage(self, it, rk, what):
        return self.arr["it%04d_rk%d_%s" % (it, rk, what)]


def _p(a, t=C.c_double):
    return a.ctypes.data_as(C.POINTER(t))


class _Chem(C.Structure):
    _fields_ = [("Ns", C.c_int), ("Nr", C.c_int), ("nKnots", C.c_int)] + [
        (n, C.POINTER(C.c_double)) for n in
        ["mMasses", "diffVol", "x", "y", "y2", "As", "Betas", "Ta", "reversible", "availBack", "AsB", "BetasB", "TaB",
         "nuR", "nuP", "expR", "expP"]]


class _Marker(C.Structure):
    _fields_ = [("kind", C.c_int), ("nVertex", C.c_int), ("node", C.POINTER(C.c_int64)),
                ("normalNeighbor", C.POINTER(C.c_int64)), ("normal", C.POINTER(C.c_double)),
                ("T", C.c_double), ("Vmag", C.c_double), ("dir", C.c_double * 2), ("Ys", C.POINTER(C.c_double)),
                ("P", C.c_double), ("Twall", C.c_double), ("heatFlux", C.c_double)]


class _Mesh(C.Structure):
    _fields_ = [("nPoint", C.c_int64), ("nPointDomain", C.c_int64), ("nEdge", C.c_int64),
                ("edgeNodes", C.POINTER(C.c_int64)), ("edgeNormal", C.POINTER(C.c_double)),
                ("coord", C.POINTER(C.c_double)), ("volume", C.POINTER(C.c_double)),
                ("nbPtr", C.POINTER(C.c_int64)), ("nbPoint", C.POINTER(C.c_int64)),
                ("nMarker", C.c_int), ("markers", C.POINTER(_Marker))]


class _Opts(C.Structure):
    _fields_ = [("viscous", C.c_int), ("rans", C.c_int), ("leastSquares", C.c_int),
                ("mInfty", C.c_double), ("cfl", C.c_double), ("maxDt", C.c_double), ("Tmin", C.c_double), ("Tmax", C.c_double),
                ("clipT", C.c_int), ("ignition", C.c_int), ("ignitionIter", C.c_int64), ("ignitionT", C.c_double),
                ("fuelIndex", C.c_int), ("oxidizerIndex", C.c_int),
                ("prandtlLam", C.c_double), ("prandtlTurb", C.c_double), ("lewisTurb", C.c_double), ("cMu", C.c_double),
                ("pasrLb", C.c_double), ("tkeInf", C.c_double),
                ("secondOrder", C.c_int), ("limiter", C.c_int), ("limCoeff", C.c_double), ("refElemLength", C.c_double),
                ("timeStepping", C.c_int), ("unstCfl", C.c_double), ("deltaUnstTime", C.c_double),
                ("dualTime", C.c_int), ("deltaUnstTimeND", C.c_double), ("viscLimiter", C.c_int)]


def chem_arrays(case):
    """Chemistry tables as contiguous float64 arrays keyed like rxo_chem / rx_chem."""
    ch = case.meta["chem"]
    Ns, Nr = ch["Ns"], ch["Nr"]
    f = lambda k, n: np.ascontiguousarray(np.array(ch.get(k, [0.0] * n), dtype=np.float64).reshape(-1)) if n else np.zeros(1)
    return dict(
        Ns=Ns, Nr=Nr, nKnots=int(case["chem_x"].shape[0]),
        mMasses=f("mMasses", Ns), diffVol=f("Diff_Volumes", Ns),
        x=np.ascontiguousarray(case["chem_x"]), y=np.ascontiguousarray(case["chem_y"]), y2=np.ascontiguousarray(case["chem_y2"]),
        As=f("As", Nr), Betas=f("Betas", Nr), Ta=f("Temps_Activation", Nr), reversible=f("Reversible", Nr),
        availBack=f("Available_Backward", Nr), AsB=f("As_back", Nr), BetasB=f("Betas_back", Nr), TaB=f("Temps_Activation_back", Nr),
        nuR=f("Stoich_Reactants", Nr * Ns), nuP=f("Stoich_Products", Nr * Ns),
        expR=f("Stoich_Reactants_Exp", Nr * Ns), expP=f("Stoich_Products_Exp", Nr * Ns))


def opts_dict(case):
    cf = case.meta["cfg"]
    return dict(viscous=cf["viscous"], rans=cf["rans"], leastSquares=1 if cf["grad_method"] == 2 else 0,
                mInfty=cf["mach"], cfl=cf["cfl"], maxDt=cf["max_dt"], Tmin=cf["Tmin"], Tmax=cf["Tmax"], clipT=cf["clip_T"],
                ignition=cf["ignition"], ignitionIter=cf["ignition_iter"], ignitionT=cf["ignition_T"],
                fuelIndex=cf["fuel_index"], oxidizerIndex=cf["oxidizer_index"], prandtlLam=cf["prandtl_lam"],
                prandtlTurb=cf["prandtl_turb"], lewisTurb=cf["lewis_turb"], cMu=cf["c_mu"], pasrLb=cf["pasr_lb"],
                tkeInf=cf["tke_inf"], secondOrder=cf.get("second_order", 0),
                limiter=(2 if cf.get("limiter_kind", 0) == 1 else 1) if cf.get("limiter", 0) else 0,   # 1 VENKATAKRISHNAN, 2 BARTH_JESPERSEN
                limCoeff=cf.get("limiter_coeff", 0.3), refElemLength=cf.get("ref_elem_length", 0.1),
                timeStepping=cf.get("time_stepping", 0), unstCfl=cf.get("unst_cfl", 0.0), deltaUnstTime=cf.get("unst_dt", 0.0),
                dualTime=cf.get("dual_time", 0), deltaUnstTimeND=cf.get("unst_dt_nd", 0.0), viscLimiter=cf.get("visc_limiter", 0))


def marker_list(case):
    out = []
    for k, mk in enumerate(case.meta["markers"]):
        kind = _BC_KIND.get(mk["bc"], 0)
        if mk["bc"] == "inlet":   # CConfig::GetKind_Inlet (option_structure.hpp:880-886): 1 total conditions, 2 mass flow, 4 temperature imposed
            kind = {1: KIND_INLET_TOTAL, 2: KIND_INLET_MASSFLOW}.get(int(case.meta["cfg"].get("kind_inlet", 4)), KIND_INLET)
        d = dict(kind=kind, tag=mk["tag"],
                 node=np.ascontiguousarray(case["geo_marker%d_node" % k], dtype=np.int64),
                 normalNeighbor=np.ascontiguousarray(case["geo_marker%d_normal_neighbor" % k], dtype=np.int64),
                 normal=np.ascontiguousarray(case["geo_marker%d_normal" % k], dtype=np.float64),
                 T=mk.get("T", 0.0), Vmag=mk.get("Vmag", 0.0), dir=mk.get("dir", [0.0, 0.0]),
                 Ys=np.array(mk.get("Ys", [0.0] * case.Ns), dtype=np.float64), P=mk.get("P", 0.0), Twall=mk.get("Twall", 0.0), heatFlux=mk.get("q", 0.0))
        out.append(d)
    return out


class Oracle:
    """ctypes driver of oracle/rx_oracle.cpp for one case."""

    def __init__(self, case, opts_override=None):
        self.case = case
        od = opts_dict(case)
        if opts_override:
            od.update(opts_override)
        self._create(mesh_dict(case), chem_arrays(case), od)

    @classmethod
    def from_mesh(cls, mesh, chem, opts):
        """Oracle on a flat mesh dict (the package's mesh builders) instead of a golden case."""
        self = cls.__new__(cls)
        self.case = None
        self._create(mesh, chem, dict(opts))
        return self

    def _create(self, mesh, ca, od):
        self.lib = C.CDLL(build_oracle())
        L = self.lib
        L.rxo_create.restype = C.c_void_p
        L.rxo_preprocess.restype = C.c_int64
        self._keep = []
        ca = {k: (np.ascontiguousarray(v, dtype=np.float64) if isinstance(v, np.ndarray) else v) for k, v in ca.items()}
        ch = _Chem(ca["Ns"], ca["Nr"], ca["nKnots"], *[_p(ca[n[0]]) for n in _Chem._fields_[3:]])
        self._keep.append(ca)
        self.chem_struct = ch
        Ns = ca["Ns"]
        mks = []
        for m in mesh["markers"]:
            mks.append(dict(kind=int(m["kind"]), node=np.ascontiguousarray(m["node"], dtype=np.int64),
                            normalNeighbor=np.ascontiguousarray(m["normalNeighbor"], dtype=np.int64),
                            normal=np.ascontiguousarray(m["normal"], dtype=np.float64).reshape(-1, 2),
                            T=float(m.get("T", 0.0)), Vmag=float(m.get("Vmag", 0.0)), dir=list(m.get("dir", [0.0, 0.0])),
                            Ys=np.ascontiguousarray(m.get("Ys", np.zeros(Ns)), dtype=np.float64), P=float(m.get("P", 0.0)),
                            Twall=float(m.get("Twall", 0.0)), heatFlux=float(m.get("heatFlux", 0.0))))
        arr = (_Marker * max(len(mks), 1))()
        for i, d in enumerate(mks):
            arr[i] = _Marker(d["kind"], len(d["node"]), _p(d["node"], C.c_int64), _p(d["normalNeighbor"], C.c_int64),
                             _p(d["normal"]), d["T"], d["Vmag"], (C.c_double * 2)(*d["dir"]), _p(d["Ys"]), d["P"], d["Twall"], d["heatFlux"])
        self._keep.append(mks)
        g = dict(en=np.ascontiguousarray(mesh["edgeNodes"], dtype=np.int64), nrm=np.ascontiguousarray(mesh["edgeNormal"], dtype=np.float64),
                 xy=np.ascontiguousarray(mesh["coord"], dtype=np.float64), vol=np.ascontiguousarray(mesh["volume"], dtype=np.float64),
                 ptr=np.ascontiguousarray(mesh["nbPtr"], dtype=np.int64), nb=np.ascontiguousarray(mesh["nbPoint"], dtype=np.int64))
        self._keep.append(g)
        nP, nE = int(mesh["nPoint"]), int(g["en"].shape[0])
        ms = _Mesh(nP, int(mesh["nPointDomain"]), nE, _p(g["en"], C.c_int64), _p(g["nrm"]), _p(g["xy"]), _p(g["vol"]),
                   _p(g["ptr"], C.c_int64), _p(g["nb"], C.c_int64), len(mks), arr)
        self.opts = od
        o = _Opts(**od)
        self.ctx = C.c_void_p(L.rxo_create(C.byref(ms), C.byref(ch), C.byref(o)))
        self.nP, self.nVar, self.nPrim, self.nGrad = nP, Ns + 4, Ns + 7, Ns + 4
        self._nE = nE

    def __del__(self):
        try:
            self.lib.rxo_destroy(self.ctx)
        except Exception:
            pass

    def set_solution(self, U):
        U = np.ascontiguousarray(U, dtype=np.float64); self.lib.rxo_set_solution(self.ctx, _p(U))

    def set_temperature(self, T):
        T = np.ascontiguousarray(T, dtype=np.float64); self.lib.rxo_set_temperature(self.ctx, _p(T))

    def turb_postprocess(self, kw, dist, flow3=None, sigma_k=0.85):
        """rxo_turb_postprocess: [nPoint][12] = k, omega, dist, F1, F2, CDkw, muT, grad k, grad omega, sigma_k."""
        kw = np.ascontiguousarray(kw, dtype=np.float64); dist = np.ascontiguousarray(dist, dtype=np.float64)
        out = np.empty((kw.shape[0], 12))
        f3 = np.ascontiguousarray(flow3, dtype=np.float64) if flow3 is not None else None
        self.lib.rxo_turb_postprocess(self.ctx, _p(kw), _p(dist), _p(f3) if f3 is not None else None, C.c_double(sigma_k), _p(out))
        return out

    def turb_iterate(self, kw, blend3, dist, sigma_k, kine_inf, omega_inf, cfl_red, relax, lin_iter, lin_tol):
        """rxo_turb_iterate: one SingleGrid_Iteration of the SST solver on the context's flow state -> dict(res, dx, rms, iters, post)."""
        kw = np.ascontiguousarray(kw, dtype=np.float64); bl = np.ascontiguousarray(blend3, dtype=np.float64); dist = np.ascontiguousarray(dist, dtype=np.float64)
        n = kw.shape[0]
        res = np.empty((n, 2)); dx = np.empty((n, 2)); rms = np.empty(2); post = np.empty((n, 12)); it = C.c_int(0)
        self.lib.rxo_turb_iterate.restype = C.c_int
        rc = self.lib.rxo_turb_iterate(self.ctx, _p(kw), _p(bl), _p(dist), C.c_double(sigma_k), C.c_double(kine_inf), C.c_double(omega_inf), C.c_double(cfl_red),
                                       C.c_double(relax), C.c_int(lin_iter), C.c_double(lin_tol), _p(res), _p(dx), _p(rms), C.byref(it), _p(post))
        assert rc == 0
        return dict(res=res, dx=dx, rms=rms, iters=it.value, post=post)

    def set_turbulence(self, t):
        t = np.ascontiguousarray(t, dtype=np.float64); self.lib.rxo_set_turbulence(self.ctx, _p(t))

    def solution(self):
        U = np.empty((self.nP, self.nVar)); self.lib.rxo_get_solution(self.ctx, _p(U)); return U

    def primitive(self):
        V = np.empty((self.nP, self.nPrim)); aux = np.empty((self.nP, 6)); g = np.empty((self.nP, self.nGrad, 2))
        self.lib.rxo_get_primitive(self.ctx, _p(V), _p(aux), _p(g)); return V, aux, g

    def dt(self):
        d = np.empty((self.nP, 3)); self.lib.rxo_get_dt(self.ctx, _p(d)); return d

    def residual(self):
        R = np.empty((self.nP, self.nVar)); self.lib.rxo_get_residual(self.ctx, _p(R)); return R

    def limiter(self):
        L = np.empty((self.nP, 4)); self.lib.rxo_get_limiter(self.ctx, _p(L)); return L

    def preprocess(self, ext_iter, output=False):
        return self.lib.rxo_preprocess(self.ctx, C.c_int64(ext_iter), int(output))

    def set_time_levels(self, Un, Un1):
        Un = np.ascontiguousarray(Un, dtype=np.float64); Un1 = np.ascontiguousarray(Un1, dtype=np.float64)
        self.lib.rxo_set_time_levels(self.ctx, _p(Un), _p(Un1))

    def time_levels(self):
        a = np.empty((self.nP, self.nVar)); b = np.empty((self.nP, self.nVar)); self.lib.rxo_get_time_levels(self.ctx, _p(a), _p(b)); return a, b

    def turb_set_time_levels(self, kwN, kwN1):
        a = np.ascontiguousarray(kwN, dtype=np.float64); b = np.ascontiguousarray(kwN1, dtype=np.float64)
        self.lib.rxo_turb_set_time_levels(self.ctx, _p(a), _p(b))

    def sm_iteration_counts(self):
        """BiCGSTAB iteration count of the Stefan-Maxwell solve of every interior edge, on the context's current primitive state."""
        it = np.zeros(self._nE, dtype=np.int32)
        self.lib.rxo_sm_iteration_counts(self.ctx, it.ctypes.data_as(C.POINTER(C.c_int32)))
        return it

    def set_inner_iteration(self, it): self.lib.rxo_set_inner_iteration(self.ctx, C.c_int64(it))
    def dual_time_residual(self): self.lib.rxo_dual_time_residual(self.ctx)
    def shift_time_levels(self): self.lib.rxo_shift_time_levels(self.ctx)
    def set_old_solution(self): self.lib.rxo_set_old_solution(self.ctx)
    def time_step(self): self.lib.rxo_time_step(self.ctx)
    def upwind_residual(self): self.lib.rxo_upwind_residual(self.ctx)
    def viscous_residual(self): self.lib.rxo_viscous_residual(self.ctx)
    def source_residual(self): self.lib.rxo_source_residual(self.ctx)
    def boundary_conditions(self): self.lib.rxo_boundary_conditions(self.ctx)

    def explicit_update(self, alpha=1.0):
        rms = np.empty(self.nVar); self.lib.rxo_explicit_update(self.ctx, C.c_double(alpha), _p(rms)); return rms


def group_rel_err(a, ref, groups):
    """Vector-relative error per group of columns:  max_{p, v in group}|a-ref| / max_{p, v in group}|ref|.
    Used for residual / conserved-variable arrays, where columns of one group share units (mass: rho and rho_s;
    momentum; energy) and a single species column can be numerically zero everywhere."""
    a = np.asarray(a, dtype=np.float64).reshape(ref.shape[0], -1)
    r = np.asarray(ref, dtype=np.float64).reshape(ref.shape[0], -1)
    out = []
    for g in groups:
        num = np.abs(a[:, g] - r[:, g]).max()
        den = np.abs(r[:, g]).max()
        out.append(num / den if den > 0 else num)
    return np.array(out)


def stage_rel_err(key, a, ref, Ns):
    """Error measure used by the parity tests for the stage array named `key` (see group_rel_err / vec_rel_err)."""
    what = key.split("_", 2)[-1]
    sp = list(range(4, 4 + Ns))
    if what in ("U0", "U1", "res_conv", "res_visc", "res_src", "res_bc", "rms"):
        return group_rel_err(a, ref, [[0] + sp, [1, 2], [3]])
    if what.endswith("grad"):
        g = lambda rows: [2 * r + d for r in rows for d in (0, 1)]
        return group_rel_err(a, ref, [g([0]), g([1, 2]), g([3]), g(sp)])
    if what.endswith("V"):
        return group_rel_err(a, ref, [[0], [1, 2], [3], [4], [5], [6], list(range(7, 7 + Ns))])
    return vec_rel_err(a, ref)


def pointwise_rel_err(a, ref, floor=1e-12):
    """The metric SURVEY.md 8(d) fixes:  max_p |a - ref| / max(|ref|, floor * max_p|ref[:, v]|)  per variable (column) v —
    every point is judged against its own reference value, with a per-variable floor so that numerically-zero entries do not
    divide by nothing.  Columns whose reference is identically zero are judged absolutely (must be below `floor` themselves)."""
    a = np.asarray(a, dtype=np.float64).reshape(ref.shape[0], -1)
    r = np.asarray(ref, dtype=np.float64).reshape(ref.shape[0], -1)
    colmax = np.abs(r).max(axis=0)
    den = np.maximum(np.abs(r), floor * colmax[None, :])
    err = np.abs(a - r)
    out = np.where(den > 0, err / np.where(den > 0, den, 1.0), err / floor)
    return out.max(axis=0)


def vec_rel_err(a, ref):
    """Per-variable vector-relative error  max_p|a-ref| / max_p|ref|  (columns = variables)."""
    a = np.asarray(a, dtype=np.float64).reshape(ref.shape[0], -1)
    r = np.asarray(ref, dtype=np.float64).reshape(ref.shape[0], -1)
    num = np.abs(a - r).max(axis=0)
    den = np.abs(r).max(axis=0)
    out = np.where(den > 0, num / np.where(den > 0, den, 1.0), num)
    return out


def mesh_dict(case):
    """Flat mesh dict (rx_mesh layout) of a golden case, for CReactiveNSSolverGPU."""
    mk = []
    for d in marker_list(case):
        mk.append(dict(kind=d["kind"], node=d["node"], normalNeighbor=d["normalNeighbor"], normal=d["normal"], T=d["T"],
                       Vmag=d["Vmag"], dir=d["dir"], Ys=d["Ys"], P=d["P"], Twall=d["Twall"], heatFlux=d["heatFlux"]))
    return dict(nPoint=case.nPoint, nPointDomain=case.nPointDomain, edgeNodes=case["geo_edge_nodes"],
                edgeNormal=case["geo_edge_normal"], coord=case["geo_coord"], volume=case["geo_volume"],
                nbPtr=case["geo_nb_ptr"], nbPoint=case["geo_nb_point"], markers=mk)


def load_package():
    import importlib
    import sys
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    return importlib.import_module("development-of-a-turbulent-numerical-solver-for-reactive-flows-in-su2_b200")


def gpu_solver(case, opts_override=None):
    pk = load_package()
    od = opts_dict(case)
    if opts_override:
        od.update(opts_override)
    return pk.CReactiveNSSolverGPU(mesh_dict(case), chem_arrays(case), od)
