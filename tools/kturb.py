#!/usr/bin/env python
"""Device time of the SST transport step (rx_turb_iterate) on the bench mesh.  usage: python tools/kturb.py [--nx 1415 --ny 1415 --reps 3]"""
import argparse, importlib, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pk = importlib.import_module("development-of-a-turbulent-numerical-solver-for-reactive-flows-in-su2_b200")
ap = argparse.ArgumentParser()
ap.add_argument("--nx", type=int, default=1415); ap.add_argument("--ny", type=int, default=1415); ap.add_argument("--reps", type=int, default=3)
a = ap.parse_args()
import torch
chem = pk.load_mechanism()
opts = pk.default_opts(rans=True, cfl=0.1)
mesh = pk.channel_mesh(a.nx, a.ny)
U0, T0, turb = pk.synthetic_state(mesh, chem, seed=12345, rans=True)
s = pk.CReactiveNSSolverGPU(mesh, chem, opts)
s.SetSolution(U0); s.SetTemperature(T0); s.SetTurbulence(turb)
s.ComputeWall_Distance()
s.TurbPostprocessing(np.ascontiguousarray(turb[:, 1:3]))
alphas = [0.66667, 0.66667, 1.0]
out = []
for it in range(a.reps):
    s.Iterate(it + 1, alphas, want_rms=False); torch.cuda.synchronize()
    t0 = time.perf_counter()
    rms, nlin = s.TurbIterate(0.135, 2.0e3)
    torch.cuda.synchronize()
    out.append((time.perf_counter() - t0) * 1e3)
print(json.dumps({"points": s.nPoint, "turb_step_ms": out, "fgmres_iterations": nlin, "rms": rms.tolist()}))
