#!/usr/bin/env python
"""Per-kernel CUDA-event times of the bench workload (rx_kernel_time), for one build of librx_cuda.so.
usage: [RX_CUDA_LIB=variants/librx_x.so] python tools/ktime.py [--nx 1415 --ny 1415 --rans 0 --reps 5]"""
import argparse, importlib, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pk = importlib.import_module("development-of-a-turbulent-numerical-solver-for-reactive-flows-in-su2_b200")
ap = argparse.ArgumentParser()
ap.add_argument("--nx", type=int, default=1415); ap.add_argument("--ny", type=int, default=1415)
ap.add_argument("--rans", type=int, default=0); ap.add_argument("--reps", type=int, default=5)
ap.add_argument("--iters", type=int, default=2)
a = ap.parse_args()
chem = pk.load_mechanism()
opts = pk.default_opts(rans=bool(a.rans), cfl=0.1)
mesh = pk.channel_mesh(a.nx, a.ny)
U0, T0, turb = pk.synthetic_state(mesh, chem, seed=12345, rans=bool(a.rans))
s = pk.CReactiveNSSolverGPU(mesh, chem, opts)
s.SetSolution(U0); s.SetTemperature(T0); s.SetTurbulence(turb)
for it in range(a.iters):
    rms = s.Iterate(it + 1, [0.66667, 0.66667, 1.0])
names = [("edge_flux_visc", 0), ("edge_flux_conv", 1), ("accumulate_source", 2), ("primitive", 3), ("transport", 7), ("gradient", 4), ("update", 6), ("source", 5)]
kt = {n: round(s.KernelTime(w, a.reps), 4) for n, w in names}
print(json.dumps({"lib": os.environ.get("RX_CUDA_LIB", "default"), "rans": a.rans, "edges": s.nEdge, "rms0": float(rms[0]), "kernel_ms": kt}))
