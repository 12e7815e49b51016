#!/usr/bin/env python
"""Time of the geometry preprocessing of geometry.py on a triangulated channel of --nx x --ny points: host build of the dual
mesh in the reference's numbering (numpy + rx_rcm_order) and, with a GPU, the wall distance (rx_wall_distance, host arrays in
and out).  usage: python tools/kgeom.py [--nx 1415 --ny 1415 --no-gpu]"""
import argparse, importlib, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pk = importlib.import_module("development-of-a-turbulent-numerical-solver-for-reactive-flows-in-su2_b200")
geo = importlib.import_module(pk.__name__ + ".geometry"); mm = importlib.import_module(pk.__name__ + ".mesh")
ap = argparse.ArgumentParser()
ap.add_argument("--nx", type=int, default=1415); ap.add_argument("--ny", type=int, default=1415); ap.add_argument("--no-gpu", action="store_true")
a = ap.parse_args()
tri = mm.channel_triangles(a.nx, a.ny)
su2 = dict(coord=mm.channel_points(a.nx, a.ny), etype=np.full(tri.shape[0], geo.TRIANGLE, dtype=np.int64),
           elems=np.concatenate([tri, np.full((tri.shape[0], 1), -1, dtype=np.int64)], axis=1), markers=list(mm.channel_boundaries(a.nx, a.ny)))
t0 = time.perf_counter(); mesh = geo.su2_dual_mesh(su2); t_build = time.perf_counter() - t0
out = {"points": int(mesh["nPoint"]), "edges": int(mesh["nEdge"]), "elements": int(tri.shape[0]), "dual_mesh_s": round(t_build, 3),
       "points_per_s": mesh["nPoint"] / t_build}
if not a.no_gpu:
    walls = ["lower_wall", "upper_wall"]
    geo.wall_distance(mesh, walls)
    t0 = time.perf_counter(); d = geo.wall_distance(mesh, walls); t_wd = time.perf_counter() - t0
    nw = sum(len(m["node"]) for m in mesh["markers"] if m["tag"] in walls)
    y = mesh["coord"][:, 1]
    out.update({"wall_nodes": int(nw), "wall_distance_s": round(t_wd, 4), "pairs_per_s": mesh["nPoint"] * nw / t_wd,
                "max_dist": float(d.max()), "half_height": float(0.5 * (y.max() - y.min()))})
print(json.dumps(out))
