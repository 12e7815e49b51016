"""Geometry preprocessing without the reference (package geometry.py + rx_wall_distance): from the content
of an SU2 mesh file to the flat dual mesh, in the reference's own numbering.  Oracle: the geometry the reference itself built
from the same files (`geo_*` arrays of the golden fixtures, exported by oracle/ref_harness.cpp) — every integer array must be
identical, every floating-point array bit-equal."""
import importlib
import json

import numpy as np
import pytest

from rxcase import GOLDEN, load_package


def _geo():
    pk = load_package()
    return pk, importlib.import_module(pk.__name__ + ".geometry")


def _jet_su2():
    z = np.load("%s/jet_mesh_su2.npz" % GOLDEN)
    tags = [str(t) for t in z["marker_tags"]]
    return dict(coord=z["coord"], etype=z["etype"].astype(np.int64), elems=z["elems"].astype(np.int64),
                markers=[(t, z["marker%d" % k].astype(np.int64)) for k, t in enumerate(tags)])


def _compare(mesh, z, tags):
    np.testing.assert_array_equal(mesh["globalIndex"], z["geo_global_index"])      # SetRCM_Ordering
    np.testing.assert_array_equal(mesh["coord"], z["geo_coord"])
    np.testing.assert_array_equal(mesh["nbPtr"], z["geo_nb_ptr"])                  # SetPoint_Connectivity
    np.testing.assert_array_equal(mesh["nbPoint"], z["geo_nb_point"])
    np.testing.assert_array_equal(mesh["edgeNodes"], z["geo_edge_nodes"])          # SetEdges
    np.testing.assert_array_equal(mesh["edgeNormal"], z["geo_edge_normal"])        # SetControlVolume, bit for bit
    np.testing.assert_array_equal(mesh["volume"], z["geo_volume"])
    for k, tag in enumerate(tags):
        m = [q for q in mesh["markers"] if q["tag"] == tag][0]
        np.testing.assert_array_equal(m["node"], z["geo_marker%d_node" % k])                      # SetVertex
        np.testing.assert_array_equal(m["normal"], z["geo_marker%d_normal" % k])                  # SetBoundControlVolume
        np.testing.assert_array_equal(m["normalNeighbor"], z["geo_marker%d_normal_neighbor" % k])  # FindNormal_Neighbor


def test_quadrilateral_jet_mesh_reproduces_reference_geometry():
    pk, geo = _geo()
    z = np.load("%s/jet_rans_shipped.npz" % GOLDEN)
    tags = [m["tag"] for m in json.loads(bytes(z["meta_json"]).decode())["markers"]]
    mesh = geo.su2_dual_mesh(_jet_su2())
    assert mesh["nEdge"] == 17810 and mesh["nPoint"] == 9000
    _compare(mesh, z, tags)


def test_triangulated_channel_reproduces_reference_geometry():
    """The 41 x 21 channel of the channel_* golden cases (oracle/make_golden.py: write_channel_su2), triangles."""
    pk, geo = _geo()
    mm = importlib.import_module(pk.__name__ + ".mesh")
    z = np.load("%s/channel_lam_rk.npz" % GOLDEN)
    tags = [m["tag"] for m in json.loads(bytes(z["meta_json"]).decode())["markers"]]
    nx, ny = 41, 21
    tri = mm.channel_triangles(nx, ny)
    su2 = dict(coord=mm.channel_points(nx, ny, stretch=1.5), etype=np.full(tri.shape[0], geo.TRIANGLE, dtype=np.int64),
               elems=np.concatenate([tri, np.full((tri.shape[0], 1), -1, dtype=np.int64)], axis=1),
               markers=[(t, s) for t, s in mm.channel_boundaries(nx, ny)])
    assert set(tags) == {t for t, _ in su2["markers"]}
    mesh = geo.su2_dual_mesh(su2)
    _compare(mesh, z, tags)


def test_su2_file_round_trip(tmp_path):
    pk, geo = _geo()
    su2 = _jet_su2()
    path = str(tmp_path / "jet.su2")
    geo.write_su2(path, su2)
    back = geo.read_su2(path)
    np.testing.assert_array_equal(back["coord"], su2["coord"])
    np.testing.assert_array_equal(back["etype"], su2["etype"])
    np.testing.assert_array_equal(back["elems"], su2["elems"])
    assert [t for t, _ in back["markers"]] == [t for t, _ in su2["markers"]]
    for (_, a), (_, b) in zip(back["markers"], su2["markers"]):
        np.testing.assert_array_equal(a, b)


def test_orientation_is_repaired():
    """Clockwise elements and inward boundary lines come out as the reference leaves them (Check_*_Orientation)."""
    pk, geo = _geo()
    su2 = _jet_su2()
    want = geo.su2_dual_mesh(su2)
    bad = dict(su2)
    el = su2["elems"].copy(); el[::3, 1], el[::3, 3] = su2["elems"][::3, 3], su2["elems"][::3, 1]     # every third quad clockwise
    bad["elems"] = el
    bad["markers"] = [(t, s[:, ::-1].copy() if k % 2 else s) for k, (t, s) in enumerate(su2["markers"])]
    got = geo.su2_dual_mesh(bad)
    # the numbering may differ (neighbour lists are formed before the orientation check, as in the reference); the geometry not
    c, el, nn = got["coord"], got["elems"], np.where(got["etype"] == geo.TRIANGLE, 3, 4)
    a = c[el[:, 1]] - c[el[:, 0]]; b2 = c[el[:, 2]] - c[el[:, 0]]
    assert (a[:, 0] * b2[:, 1] - a[:, 1] * b2[:, 0] > 0).all()                       # every element counter-clockwise

    def edge_map(m):
        g = m["globalIndex"]; out = {}
        for (i, j), n in zip(m["edgeNodes"], m["edgeNormal"]):
            gi, gj = int(g[i]), int(g[j])
            out[(min(gi, gj), max(gi, gj))] = n if gi < gj else -n
        return out
    A, B = edge_map(got), edge_map(want)
    assert A.keys() == B.keys()
    assert max(np.abs(A[k] - B[k]).max() for k in A) <= 1e-18
    vg = np.empty(got["nPoint"]); vg[got["globalIndex"]] = got["volume"]
    vw = np.empty(want["nPoint"]); vw[want["globalIndex"]] = want["volume"]
    np.testing.assert_allclose(vg, vw, rtol=1e-14)
    for ma, mb in zip(got["markers"], want["markers"]):                               # vertex normals as the reference orients them (into the domain)
        na = {int(got["globalIndex"][p]): n for p, n in zip(ma["node"], ma["normal"])}
        nb = {int(want["globalIndex"][p]): n for p, n in zip(mb["node"], mb["normal"])}
        assert na.keys() == nb.keys()
        assert max(np.abs(na[k] - nb[k]).max() for k in na) <= 1e-18


@pytest.mark.gpu
@pytest.mark.parametrize("name,walls", [("jet_rans_shipped", ["upper_wall", "lower_wall_pre", "lower_wall_post"]),
                                        ("channel_rans", ["lower_wall", "upper_wall"])])
def test_wall_distance_matches_reference(name, walls):
    """rx_wall_distance against geometry->node[]->GetWall_Distance() of the reference (column 2 of the `tpost` dump)."""
    pk, geo = _geo()
    from rxcase import Case, mesh_dict
    case = Case("%s/%s.npz" % (GOLDEN, name))
    ref = case[sorted(k for k in case.arr if k.endswith("_tpost"))[-1]][:, 2]
    mesh = mesh_dict(case)
    meta = case.meta["markers"]
    for m, mk in zip(mesh["markers"], meta):
        m["tag"] = mk["tag"]
    d = geo.wall_distance(mesh, walls)
    np.testing.assert_allclose(d, ref, rtol=1e-15, atol=0)


@pytest.mark.gpu
def test_solver_runs_on_the_mesh_built_from_the_file():
    """The dual mesh built here drives the CUDA path to the same residual norms as the geometry exported by the reference."""
    pk, geo = _geo()
    from rxcase import Case, chem_arrays, mesh_dict, opts_dict
    case = Case("%s/jet_rans_perturbed.npz" % GOLDEN)
    mine = geo.su2_dual_mesh(_jet_su2())
    ref = mesh_dict(case)
    for m, r in zip(mine["markers"], ref["markers"]):     # boundary data (kinds, states) from the case's cfg
        for k, v in r.items():
            if k not in ("node", "normal", "normalNeighbor"):
                m[k] = v
    it = sorted({int(k[2:6]) for k in case.arr if k.startswith("it")})[0]
    out = []
    for mesh in (mine, ref):
        s = pk.CReactiveNSSolverGPU(mesh, chem_arrays(case), opts_dict(case))
        s.SetSolution(case.stage(it, 0, "U0")); s.SetTemperature(case.stage(it, 0, "prev_V")[:, 0]); s.SetTurbulence(case.stage(it, 0, "turb"))
        out.append(np.array(s.Iterate(it, [1.0])))
    np.testing.assert_array_equal(out[0], out[1])


def test_mixed_triangle_quadrilateral_mesh_closes_every_control_volume():
    """A mesh of quadrilaterals with every other one split into two triangles (given in mixed orientation): the dual control
    volumes tile the domain and every one of them is closed — the outgoing edge normals around a point minus its (inward) boundary normals sum
    to zero — which is the discrete property the residual loop relies on (a constant flux leaves no residual)."""
    pk, geo = _geo()
    nx, ny = 9, 7
    xs = np.linspace(0.0, 2.0, nx) ** 1.3; ys = np.linspace(0.0, 1.0, ny) ** 0.8
    X, Y = np.meshgrid(xs, ys)
    coord = np.stack([X.ravel() + 0.01 * np.sin(7 * Y.ravel()), Y.ravel()], axis=1)
    etype, elems = [], []
    for j in range(ny - 1):
        for i in range(nx - 1):
            a = j * nx + i; b = a + 1; c = b + nx; d = a + nx
            if (i + j) % 2 == 0:
                etype.append(geo.QUADRILATERAL); elems.append([a, b, c, d] if (i % 3) else [a, d, c, b])     # some clockwise
            else:
                etype += [geo.TRIANGLE, geo.TRIANGLE]; elems += [[a, b, c, -1], [a, d, c, -1]]                # second one clockwise
    i = np.arange(nx - 1); j = np.arange(ny - 1)
    markers = [("bottom", np.stack([i + 1, i], 1)), ("right", np.stack([j * nx + nx - 1, (j + 1) * nx + nx - 1], 1)),
               ("top", np.stack([(ny - 1) * nx + i, (ny - 1) * nx + i + 1], 1)), ("left", np.stack([j * nx, (j + 1) * nx], 1))]
    su2 = dict(coord=coord, etype=np.array(etype, dtype=np.int64), elems=np.array(elems, dtype=np.int64), markers=markers)
    for rcm in (True, False):
        m = geo.su2_dual_mesh(su2, rcm=rcm)
        # area of the domain from the boundary polygon (shoelace), in the file's numbering
        ring = np.concatenate([np.arange(nx), np.arange(1, ny) * nx + nx - 1, (ny - 1) * nx + np.arange(nx - 2, -1, -1), np.arange(ny - 2, 0, -1) * nx])
        px, py = coord[ring, 0], coord[ring, 1]
        area = 0.5 * abs(np.dot(px, np.roll(py, -1)) - np.dot(py, np.roll(px, -1)))
        assert abs(m["volume"].sum() - area) <= 1e-14 * area
        assert (m["volume"] > 0).all()
        close = np.zeros((m["nPoint"], 2))
        np.add.at(close, m["edgeNodes"][:, 0], m["edgeNormal"]); np.add.at(close, m["edgeNodes"][:, 1], -m["edgeNormal"])
        for mk in m["markers"]:
            np.add.at(close, mk["node"], -mk["normal"])   # vertex normals point INTO the domain (the BC routines negate them, e.g. solver_direct_reactive.cpp BC_Inlet)
        assert np.abs(close).max() <= 1e-15
        assert m["nEdge"] == len({(min(a, b), max(a, b)) for a, b in m["edgeNodes"]})
        # neighbour lists and edges describe the same graph
        deg = np.diff(m["nbPtr"])
        assert deg.sum() == 2 * m["nEdge"]


def _mixed_su2(geo, nx=9, ny=7):
    xs = np.linspace(0.0, 0.2, nx) ** 1.3; ys = np.linspace(0.0, 0.05, ny) ** 0.8
    X, Y = np.meshgrid(xs, ys)
    coord = np.stack([X.ravel() + 1e-4 * np.sin(700 * Y.ravel()), Y.ravel()], axis=1)
    etype, elems = [], []
    for j in range(ny - 1):
        for i in range(nx - 1):
            a = j * nx + i; b = a + 1; c = b + nx; d = a + nx
            if (i + j) % 2 == 0:
                etype.append(geo.QUADRILATERAL); elems.append([a, b, c, d] if (i % 3) else [a, d, c, b])
            else:
                etype += [geo.TRIANGLE, geo.TRIANGLE]; elems += [[a, b, c, -1], [a, d, c, -1]]
    i = np.arange(nx - 1); j = np.arange(ny - 1)
    markers = [("bottom", np.stack([i + 1, i], 1)), ("right", np.stack([j * nx + nx - 1, (j + 1) * nx + nx - 1], 1)),
               ("top", np.stack([(ny - 1) * nx + i, (ny - 1) * nx + i + 1], 1)), ("left", np.stack([j * nx, (j + 1) * nx], 1))]
    return dict(coord=coord, etype=np.array(etype, dtype=np.int64), elems=np.array(elems, dtype=np.int64), markers=markers)


@pytest.mark.gpu
def test_uniform_state_leaves_no_convective_residual_on_a_mixed_mesh():
    """Free-stream preservation through the CUDA path on the mesh geometry.py builds from mixed, partly mis-oriented elements:
    with a uniform state the AUSM fluxes around every interior control volume cancel (closed volumes), to rounding."""
    pk, geo = _geo()
    mesh = geo.su2_dual_mesh(_mixed_su2(geo))
    chem = pk.load_mechanism()
    opts = pk.default_opts(rans=False, cfl=0.1)
    U, T, turb = pk.synthetic_state(mesh, chem, seed=3)
    U[:] = U[0]; T[:] = T[0]
    s = pk.CReactiveNSSolverGPU(mesh, chem, opts)
    s.SetSolution(U); s.SetTemperature(T); s.SetTurbulence(turb)
    s.Preprocessing(1, False, count_nonphysical=False)
    s.Upwind_Residual()
    R = s.GetLinSysRes()
    boundary = np.zeros(mesh["nPoint"], dtype=bool)
    for m in mesh["markers"]:
        boundary[m["node"]] = True
    # scale: the flux through one face, |F| * |n|
    nrm = np.sqrt((mesh["edgeNormal"] ** 2).sum(1)).max()
    rho, u, v, E = U[0, 0], U[0, 1] / U[0, 0], U[0, 2] / U[0, 0], U[0, 3]
    scale = np.array([rho * np.hypot(u, v), rho * (u * u + v * v) + 1e5, rho * (u * u + v * v) + 1e5, E * np.hypot(u, v) + 1e5 * np.hypot(u, v)] + [rho * np.hypot(u, v)] * chem["Ns"]) * nrm
    assert (~boundary).sum() > 10
    assert (np.abs(R[~boundary]) / scale).max() <= 1e-12
    assert np.abs(R[boundary]).max() > 0   # open boundary faces do carry flux until the BCs close them


def _same_mesh(a, b):
    for k in ("nPoint", "nPointDomain", "nEdge"):
        assert a[k] == b[k], k
    for k in ("globalIndex", "coord", "nbPtr", "nbPoint", "edgeNodes", "edgeNormal", "volume", "elems", "etype"):
        np.testing.assert_array_equal(a[k], b[k], err_msg=k)
    assert [m["tag"] for m in a["markers"]] == [m["tag"] for m in b["markers"]]
    for ma, mb in zip(a["markers"], b["markers"]):
        for k in ("node", "normal", "normalNeighbor"):
            np.testing.assert_array_equal(ma[k], mb[k], err_msg=ma["tag"] + " " + k)


def test_native_su2_preprocessing_equals_the_specification_and_the_reference(tmp_path):
    """rx_su2_mesh_read / rx_su2_mesh_build (csrc/host/rx_su2_mesh.cpp: parse, connectivity, RCM, orientation, edges, control volumes,
    vertices in one native pass) against su2_dual_mesh, array by array and bit for bit — on the shipped quadrilateral jet mesh (also against
    the geometry the reference built from it), a mixed mesh with mis-oriented elements, with and without the RCM numbering, from arrays and
    from a file."""
    pk, geo = _geo()
    jet = _jet_su2()
    z = np.load("%s/jet_rans_shipped.npz" % GOLDEN)
    tags = [m["tag"] for m in json.loads(bytes(z["meta_json"]).decode())["markers"]]
    nat = geo.su2_dual_mesh_native(jet)
    _compare(nat, z, tags)                                   # the reference's own geometry
    _same_mesh(nat, geo.su2_dual_mesh(jet))
    path = str(tmp_path / "jet.su2")
    geo.write_su2(path, jet)
    _same_mesh(geo.load_su2_mesh(path), nat)                 # through the file parser
    bad = dict(jet)
    el = jet["elems"].copy(); el[::3, 1], el[::3, 3] = jet["elems"][::3, 3], jet["elems"][::3, 1]
    bad["elems"] = el
    bad["markers"] = [(t, s[:, ::-1].copy() if k % 2 else s) for k, (t, s) in enumerate(jet["markers"])]
    _same_mesh(geo.su2_dual_mesh_native(bad), geo.su2_dual_mesh(bad))
    mixed = _mixed_su2(geo)
    for rcm in (True, False):
        _same_mesh(geo.su2_dual_mesh_native(mixed, rcm=rcm), geo.su2_dual_mesh(mixed, rcm=rcm))
    p2 = str(tmp_path / "mixed.su2")
    geo.write_su2(p2, mixed)
    _same_mesh(geo.load_su2_mesh(p2, rcm=False), geo.su2_dual_mesh(geo.read_su2(p2), rcm=False))


def test_native_su2_reader_refuses_malformed_files(tmp_path):
    pk, geo = _geo()
    with pytest.raises(ValueError, match="cannot open"):
        geo.load_su2_mesh(str(tmp_path / "absent.su2"))
    mixed = _mixed_su2(geo)
    good = str(tmp_path / "m.su2")
    geo.write_su2(good, mixed)
    txt = open(good).read()
    cases = {"3d.su2": (txt.replace("NDIME= 2", "NDIME= 3"), "2-D"),
             "cut.su2": (txt[:len(txt) // 3], "ends inside|no points"),
             "hex.su2": (txt.replace("\n9 ", "\n12 ", 1), "not a triangle or quadrilateral")}
    for name, (body, msg) in cases.items():
        p = str(tmp_path / name)
        open(p, "w").write(body)
        with pytest.raises(ValueError, match=msg):
            geo.load_su2_mesh(p)
    far = dict(mixed); far["markers"] = [("bottom", np.array([[0, mixed["coord"].shape[0] - 1]]))] + mixed["markers"][1:]
    with pytest.raises(ValueError, match="not a face"):
        geo.su2_dual_mesh_native(far)
    out = dict(mixed); e2 = mixed["elems"].copy(); e2[0, 0] = 10 ** 6; out["elems"] = e2
    with pytest.raises(ValueError, match="outside the mesh"):
        geo.su2_dual_mesh_native(out)
