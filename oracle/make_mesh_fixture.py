#!/usr/bin/env python
"""tests/golden/jet_mesh_su2.npz: the content of the reference's shipped jet-reactor mesh file (elements, points, marker
elements, in the file's numbering), read with the package's own reader.  Together with the `geo_*` arrays of
tests/golden/jet_rans_shipped.npz — the geometry the reference itself built from that file — it pins geometry.su2_dual_mesh.
Run in the build container (reads /root/reference); the GPU box only sees the fixture."""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
geo = importlib.import_module("development-of-a-turbulent-numerical-solver-for-reactive-flows-in-su2_b200.geometry")
SRC = "/root/reference/Test_Cases/TURBOLENT/TURBOLENT_COMBUSTION/mesh_stretched.su2"
su2 = geo.read_su2(SRC)
arrs = dict(coord=su2["coord"], etype=su2["etype"].astype(np.int8), elems=su2["elems"].astype(np.int32),
            marker_tags=np.array([t for t, _ in su2["markers"]]))
for k, (_, seg) in enumerate(su2["markers"]):
    arrs["marker%d" % k] = seg.astype(np.int32)
out = os.path.join(ROOT, "tests", "golden", "jet_mesh_su2.npz")
np.savez_compressed(out, **arrs)
print("wrote %s (%.2f MB)" % (out, os.path.getsize(out) / 1e6))
