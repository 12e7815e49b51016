#!/usr/bin/env python
"""Distribution of the per-edge keys of the viscous kernel (BiCGSTAB iterations, +19 = closure QR ran) on the bench mesh, and
what a warp of 16 consecutive processing slots pays for it (max over the warp) in edge order and in the sorted order.
usage: python tools/kiters.py [--nx 1415 --ny 1415 --rans 0]"""
import argparse, importlib, json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pk = importlib.import_module("development-of-a-turbulent-numerical-solver-for-reactive-flows-in-su2_b200")
ap = argparse.ArgumentParser()
ap.add_argument("--nx", type=int, default=1415); ap.add_argument("--ny", type=int, default=1415); ap.add_argument("--rans", type=int, default=0)
a = ap.parse_args()
chem = pk.load_mechanism(); opts = pk.default_opts(rans=bool(a.rans), cfl=0.1)
mesh = pk.channel_mesh(a.nx, a.ny)
U0, T0, turb = pk.synthetic_state(mesh, chem, seed=12345, rans=bool(a.rans))
s = pk.CReactiveNSSolverGPU(mesh, chem, opts)
s.SetSolution(U0); s.SetTemperature(T0); s.SetTurbulence(turb)
for it in range(2):
    s.Iterate(it + 1, [0.66667, 0.66667, 1.0])
order, key = s.EdgeOrder()
it = (key % 19).astype(np.int64)
n = it.shape[0] // 16 * 16
out = {"edges": int(it.shape[0]), "hist": np.bincount(key).tolist(), "mean": float(it.mean()),
       "warp_max_edge_order": float(it[:n].reshape(-1, 16).max(1).mean()),
       "warp_max_current_order": float(it[order][:n].reshape(-1, 16).max(1).mean()),
       "warps_all_zero_edge_order": float((it[:n].reshape(-1, 16).max(1) == 0).mean()),
       "identity_order": bool((order == np.arange(order.shape[0])).all())}
print(json.dumps(out))
