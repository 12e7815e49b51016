"""Synthetic inputs for the benchmarks: the shipped 9-species / 2-reaction mechanism tables and a seeded,
thermodynamically consistent, non-uniform initial state on a channel mesh.  Input preparation only (numpy, host)."""
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
R_UNGAS = (6.02214129 * 1.0e23) * (1.3806488 * 1.0e-23) * 1.0e3


def load_mechanism(path=None):
    """Tables of Framework::ReactingModelLibrary for the shipped jet-reactor mechanism (rx_chem layout)."""
    z = np.load(path or os.path.join(HERE, "data", "mechanism_c4h6_9sp.npz"))
    ch = {k: z[k] for k in z.files if k != "meta_json"}
    for k in ("Ns", "Nr", "nKnots"):
        ch[k] = int(ch[k])
    ch["meta"] = json.loads(bytes(z["meta_json"]).decode())
    return ch


def default_opts(rans=False, cfl=0.1):
    """CConfig-derived constants of the shipped cfgs (my_combustion_first_chem_PaSR.cfg) with IGNITION off."""
    return dict(viscous=1, rans=int(rans), leastSquares=1, mInfty=0.01819802685932537, cfl=cfl, maxDt=1.0e6, Tmin=200.0, Tmax=6000.0,
                clipT=0, ignition=0, ignitionIter=0, ignitionT=1700.0, fuelIndex=0, oxidizerIndex=2, prandtlLam=0.72,
                prandtlTurb=0.9, lewisTurb=1.2, cMu=0.09, pasrLb=0.2, tkeInf=0.135)


def mixture_enthalpy(chem, T, Y):
    """h(T) = sum_s Y_s h_s(T)/M_s with the library's cubic splines (Common/src/Tools/spline.cpp:61-77), vectorised."""
    x = chem["x"]; n = chem["nKnots"]; Ns = chem["Ns"]
    h = x[1] - x[0]
    klo = np.clip(((T - x[0]) / h + 1).astype(np.int64), 1, n - 1)
    a = (x[klo] - T) / h; b = (T - x[klo - 1]) / h
    out = np.zeros_like(T)
    yt = chem["y"].reshape(5, Ns, n)[3]; y2 = chem["y2"].reshape(5, Ns, n)[3]
    for s in range(Ns):
        hs = a * yt[s][klo - 1] + b * yt[s][klo] + ((a ** 3 - a) * y2[s][klo - 1] + (b ** 3 - b) * y2[s][klo]) * (h * h) / 6.0
        out += Y[:, s] * hs / chem["mMasses"][s]
    return out


def synthetic_state(mesh, chem, seed=12345, rans=False, fuel=0, oxidizer=2, Tpeak=800.0, glob_xy=None):
    """Smooth shear-layer fields + seeded noise on (T, P, u, v, Y_s[, k, omega]); returns U [nP, 4+Ns], T guess, turb [nP, 6].
    `glob_xy` = (x0, x1, y0, y1) bounding box of the GLOBAL domain (for partitions)."""
    xy = mesh["coord"]; nP = xy.shape[0]; Ns = chem["Ns"]
    x0, x1, y0, y1 = glob_xy or (xy[:, 0].min(), xy[:, 0].max(), xy[:, 1].min(), xy[:, 1].max())
    xh = (xy[:, 0] - x0) / (x1 - x0); yh = (xy[:, 1] - y0) / (y1 - y0)
    # noise is a hash of the coordinates, so a partition sees the same values as the global mesh
    def noise(k):
        t = np.sin(xy[:, 0] * 12.9898e3 + xy[:, 1] * 78.233e3 + 37.719 * (k + seed % 97)) * 43758.5453
        return 2.0 * (t - np.floor(t)) - 1.0
    sh = 0.5 + 0.5 * np.tanh((yh - 0.5) * 6.0)
    T = 500.0 + Tpeak * np.exp(-((yh - 0.5) / 0.2) ** 2) * (0.6 + 0.4 * np.sin(2 * np.pi * xh)) + 8.0 * noise(0)
    P = 101325.0 * (1.0 + 0.05 * np.sin(3.0 * xh + yh)) * (1.0 + 1e-3 * noise(1))
    u = 30.0 * 4.0 * yh * (1.0 - yh) + 3.0 * np.sin(5.0 * xh) + 0.2 * noise(2)
    v = 2.0 * np.sin(2 * np.pi * xh) * np.cos(np.pi * yh) + 0.2 * noise(3)
    Y = np.stack([0.02 * (1.0 + np.sin(1.7 * s + 4.0 * xh + 2.0 * yh)) * (1.0 + 0.1 * noise(4 + s)) for s in range(Ns)], 1)
    Y[:, fuel] += 1.0 - sh
    Y[:, oxidizer] += sh
    Y[yh > 0.7, fuel] = 0.0
    Y[xh < 0.3, Ns - 1] = 0.0
    pure = (xh > 0.8) & (yh < 0.3)
    Y[pure] = 0.0; Y[pure, oxidizer] = 1.0
    Y /= Y.sum(1, keepdims=True)
    Rg = (Y * (R_UNGAS / chem["mMasses"])[None, :]).sum(1)
    rho = P / (Rg * T)
    turb = np.zeros((nP, 6)); k = np.zeros(nP)
    if rans:
        k = 0.05 + 5.0 * yh * (1.0 - yh) * (1.0 + 0.2 * noise(20))
        om = 2.0e3 * (1.0 + 0.8 * np.sin(3.0 * xh) * np.cos(2.0 * yh)) * (1.0 + 0.1 * noise(21))
        turb[:, 0] = rho * k / om; turb[:, 1] = k; turb[:, 2] = om; turb[:, 5] = 0.85
        turb[:, 3] = 5.0 * (1.0 - 2.0 * yh) / (y1 - y0) * 0.0; turb[:, 4] = 5.0 * (1.0 - 2.0 * yh) / (y1 - y0)
    hmix = mixture_enthalpy(chem, T, Y)
    U = np.empty((nP, 4 + Ns))
    U[:, 0] = rho; U[:, 1] = rho * u; U[:, 2] = rho * v; U[:, 3] = rho * (hmix + 0.5 * (u * u + v * v) + k) - P
    U[:, 4:] = rho[:, None] * Y
    return U, T.copy(), turb


def synthetic_points(chem, n, seed=12345, fuel=0, oxidizer=2):
    """Independent thermochemical states for the source-term microbenchmark: T ~ U[300, 2500] K,
    P ~ U[0.9, 1.4] 1e5 Pa, Y = normalised U[0,1]^Ns with 30 % of the points pure oxidiser and 10 % pure fuel (the 1e-30 /
    1e-15 / 1e-10 thresholds of the rate expressions), omega_turb ~ 10^U[2,7] 1/s.  Returns T, rho, Y[n][Ns], omega_turb, vol."""
    rng = np.random.default_rng(seed)
    Ns = chem["Ns"]
    T = rng.uniform(300.0, 2500.0, n)
    P = rng.uniform(0.9e5, 1.4e5, n)
    Y = rng.uniform(0.0, 1.0, (n, Ns))
    u = rng.uniform(0.0, 1.0, n)
    Y[u < 0.3] = 0.0; Y[u < 0.3, oxidizer] = 1.0
    pf = (u >= 0.3) & (u < 0.4)
    Y[pf] = 0.0; Y[pf, fuel] = 1.0
    Y /= Y.sum(1, keepdims=True)
    Rg = (Y * (R_UNGAS / chem["mMasses"])[None, :]).sum(1)
    rho = P / (Rg * T)
    omT = 10.0 ** rng.uniform(2.0, 7.0, n)
    vol = rng.uniform(0.5e-8, 2.0e-8, n)
    return T, rho, Y, omT, vol
