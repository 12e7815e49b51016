#!/usr/bin/env python
"""Device time of the stand-alone source kernel (k_source_points) over n points.  usage: [RX_CUDA_LIB=...] python tools/ksource.py [n]"""
import importlib, json, os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pk = importlib.import_module("development-of-a-turbulent-numerical-solver-for-reactive-flows-in-su2_b200")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 16_000_000
chem = pk.load_mechanism()
base = 2_000_000
T, rho, Y, omT, vol = pk.synthetic_points(chem, base, seed=12345)
dev = torch.device("cuda", 0)
reps = (n + base - 1) // base
t = lambda x: torch.as_tensor(np.ascontiguousarray(x), device=dev).repeat(reps)[:n].contiguous()
dT, dr, dO, dV = t(T), t(rho), t(omT), t(vol)
dY = torch.as_tensor(np.ascontiguousarray(Y.T), device=dev).repeat(1, reps)[:, :n].contiguous()
dR = torch.empty((chem["Ns"], n), dtype=torch.float64, device=dev)
s = pk.CReactiveNSSolverGPU(pk.channel_mesh(5, 4), chem, pk.default_opts(rans=True))
run = lambda: s.SourcePointsDev(n, dT.data_ptr(), dr.data_ptr(), dY.data_ptr(), dO.data_ptr(), dV.data_ptr(), dR.data_ptr())
for _ in range(3): run()
torch.cuda.synchronize()
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5): run()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
print(json.dumps({"lib": os.environ.get("RX_CUDA_LIB", "default"), "n": n, "ms": round(ms, 4), "Gpts_per_s": round(n / ms / 1e6, 3), "checksum": float(dR.sum())}))
