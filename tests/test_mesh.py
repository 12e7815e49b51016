"""Host-side mesh builders (package mesh.py): the dual mesh against the reference's own geometry dump, and the strip
partitions against the global mesh (SU2's halo scheme: owned points first, one halo layer, edges with an owned end)."""
import numpy as np
import pytest

from rxcase import GOLDEN, Case, load_package, marker_list

import os


@pytest.fixture(scope="module")
def pk():
    return load_package()


def _edge_map(mesh, glob=None):
    """{(gi, gj): normal} with the normal pointing from gi to gj."""
    en = np.asarray(mesh["edgeNodes"]); nrm = np.asarray(mesh["edgeNormal"]).reshape(-1, 2)
    out = {}
    for (i, j), n in zip(en, nrm):
        gi, gj = (int(i), int(j)) if glob is None else (int(glob[i]), int(glob[j]))
        if gi < gj:
            out[(gi, gj)] = n.copy()
        else:
            out[(gj, gi)] = -n
    return out


def test_channel_mesh_matches_reference_geometry(pk):
    """41 x 21 stretched channel of tests/golden/channel_lam_rk.npz: same points, volumes, edges, normals, markers as the
    reference's CGeometry produced from the .su2 file (oracle/make_golden.py writes the same grid)."""
    case = Case(os.path.join(GOLDEN, "channel_lam_rk.npz"))
    st = case.meta.get("mesh", {}).get("stretch", 1.5)
    m = pk.channel_mesh(41, 21, stretch=st)
    assert m["nPoint"] == case.nPoint and len(m["edgeNodes"]) == case.nEdge
    # the reference renumbers the points (RCM); match them by coordinates
    key = lambda xy: {(round(float(x), 12), round(float(y), 12)): k for k, (x, y) in enumerate(xy)}
    km = key(m["coord"])
    r2m = np.array([km[(round(float(x), 12), round(float(y), 12))] for x, y in case["geo_coord"]])   # reference id -> mesh id
    assert len(set(r2m.tolist())) == case.nPoint
    np.testing.assert_allclose(m["volume"][r2m], case["geo_volume"], rtol=1e-12)
    ref = dict(edgeNodes=case["geo_edge_nodes"], edgeNormal=case["geo_edge_normal"])
    a, b = _edge_map(m), _edge_map(ref, r2m)
    assert a.keys() == b.keys()
    scale = np.abs(case["geo_edge_normal"]).max()
    assert max(np.abs(a[k] - b[k]).max() for k in a) < 1e-12 * scale
    for mk, rk in zip(m["markers"], marker_list(case)):
        rn = r2m[rk["node"]]
        o1, o2 = np.argsort(mk["node"]), np.argsort(rn)
        assert np.array_equal(np.asarray(mk["node"])[o1], rn[o2])
        assert np.array_equal(np.asarray(mk["normalNeighbor"])[o1], r2m[rk["normalNeighbor"]][o2])
        np.testing.assert_allclose(np.asarray(mk["normal"]).reshape(-1, 2)[o1], rk["normal"].reshape(-1, 2)[o2], atol=1e-12 * scale)


@pytest.mark.parametrize("nranks", [2, 3])
def test_strip_meshes_tile_the_global_mesh(pk, nranks):
    nxg, ny = 23, 9
    g = pk.channel_mesh(nxg, ny)
    gmap = _edge_map(g)
    seen_pts = np.zeros(g["nPoint"], dtype=int)
    counted_edges = 0
    for r in range(nranks):
        m, plan = pk.channel_strip_mesh(nxg, ny, r, nranks)
        gi, gj = plan["glob"]
        glob = gj * nxg + gi                       # global point id of the flat channel mesh
        nOwn = m["nPointDomain"]
        seen_pts[glob[:nOwn]] += 1
        np.testing.assert_allclose(m["coord"], g["coord"][glob], atol=1e-15)
        # owned control volumes are complete; halo volumes are not used by the path
        np.testing.assert_allclose(m["volume"][:nOwn], g["volume"][glob[:nOwn]], rtol=1e-12)
        lmap = _edge_map(m, glob)
        for k, n in lmap.items():
            assert k in gmap
            np.testing.assert_allclose(n, gmap[k], atol=1e-15 + 1e-12 * np.abs(gmap[k]).max())
        en = np.asarray(m["edgeNodes"])
        assert np.all((en[:, 0] < nOwn) | (en[:, 1] < nOwn))            # every kept edge has an owned end
        counted_edges += int((en[:, 0] < nOwn).sum())                     # bench.py's rule: counted by the owner of node 0
        # neighbours of owned points are complete (least-squares gradients need all of them)
        for p in range(0, nOwn, 7):
            nb_l = set(glob[m["nbPoint"][m["nbPtr"][p]:m["nbPtr"][p + 1]]].tolist())
            gp = glob[p]
            nb_g = set(g["nbPoint"][g["nbPtr"][gp]:g["nbPtr"][gp + 1]].tolist())
            assert nb_l == nb_g
        # halo send/recv lists pair up with the neighbour's
        for q, ids in plan["recv"].items():
            mq, pq = pk.channel_strip_mesh(nxg, ny, q, nranks)
            gq = pq["glob"][1] * nxg + pq["glob"][0]
            assert np.array_equal(glob[ids], gq[pq["send"][r]])
            assert np.all(np.asarray(ids) >= nOwn) and np.all(np.asarray(pq["send"][r]) < mq["nPointDomain"])
        # boundary vertices only on owned points, and all of them
        for mk, gk in zip(m["markers"], g["markers"]):
            assert np.all(np.asarray(mk["node"]) < nOwn)
    assert np.all(seen_pts == 1)
    assert counted_edges == len(g["edgeNodes"])
    nb = sum(len(pk.channel_strip_mesh(nxg, ny, r, nranks)[0]["markers"][k]["node"]) for r in range(nranks) for k in range(4))
    assert nb == sum(len(mk["node"]) for mk in g["markers"])


def test_strip_renumbering_short_side_slow_is_the_same_mesh_with_shorter_sweep_chains():
    """strip_short_side_slow: same points, edges, markers and exchange lists under another numbering; the dependency levels of a
    Gauss-Seidel sweep in index order (what rx_turb_iterate's LU-SGS follows) drop from about nx + 2 ny to ny + 2 nx."""
    import numpy as np
    from rxcase import load_package
    pk = load_package()
    m, plan = pk.channel_strip_mesh(160, 101, 1, 8)
    m2, p2 = pk.strip_short_side_slow(m, plan)
    n = m["nPointDomain"]
    assert m2["nPointDomain"] == n and m2["nPoint"] == m["nPoint"] and m2["nEdge"] == m["nEdge"]
    key = lambda mm, pp: {(int(a), int(b)) for a, b in zip(pp["glob"][0] * 1000 + pp["glob"][1], range(len(pp["glob"][0])))}
    ga, gb = plan["glob"][0] * 1000 + plan["glob"][1], p2["glob"][0] * 1000 + p2["glob"][1]
    assert sorted(ga[:n]) == sorted(gb[:n]) and list(ga[n:]) == list(gb[n:])            # owned set kept, halo layer untouched
    o = {int(g): i for i, g in enumerate(ga)}
    perm = np.array([o[int(g)] for g in gb])                                             # new -> old
    assert np.array_equal(m2["coord"], m["coord"][perm]) and np.array_equal(m2["volume"], m["volume"][perm])
    ea = {(int(ga[a]), int(ga[b])) for a, b in m["edgeNodes"]}; eb = {(int(gb[a]), int(gb[b])) for a, b in m2["edgeNodes"]}
    assert ea == eb
    for r in plan["send"]:
        assert list(ga[plan["send"][r]]) == list(gb[p2["send"][r]]) and list(ga[plan["recv"][r]]) == list(gb[p2["recv"][r]])

    def levels(mm):
        ptr, nb = mm["nbPtr"], mm["nbPoint"]
        lev = np.zeros(n, dtype=np.int64)
        for p in range(n):
            q = nb[ptr[p]:ptr[p + 1]]; q = q[q < p]
            lev[p] = lev[q].max() + 1 if len(q) else 0
        return int(lev.max()) + 1
    la, lb = levels(m), levels(m2)
    assert lb < 0.7 * la, (la, lb)


def test_native_dual_mesh_builder_equals_the_numpy_specification_bit_for_bit():
    """rx_dual_mesh_tri (csrc/host/rx_geometry.cpp: per-point buckets, one linear pass per array) against the numpy statements of
    mesh.dual_mesh, which restate CPhysicalGeometry::SetControlVolume (geometry_structure.cpp:10457): edges, normals, volumes, neighbour
    CSR and the boundary arrays derived from them, on stretched and jittered channels with shuffled element order and orientation."""
    import importlib
    import numpy as np
    from rxcase import load_package
    pk = load_package()
    mm = importlib.import_module(pk.__name__ + ".mesh")
    rng = np.random.default_rng(3)
    for nx, ny, st in [(41, 21, 1.5), (7, 5, 1.0), (64, 33, 1.0), (3, 3, 1.0)]:
        coord = mm.channel_points(nx, ny, stretch=st)
        inner = np.ones(nx * ny, dtype=bool).reshape(ny, nx); inner[0, :] = inner[-1, :] = False; inner[:, 0] = inner[:, -1] = False
        ys = np.unique(coord[:, 1])
        jit = (rng.random(coord.shape) - 0.5) * 0.2 * min(0.25 / nx, np.diff(ys).min())   # small against every spacing: no element inverts
        coord = coord + jit * inner.ravel()[:, None]                      # interior points jittered: no two contributions are equal by symmetry
        tri = mm.channel_triangles(nx, ny)
        tri = tri[rng.permutation(len(tri))]
        roll = rng.integers(0, 3, len(tri))
        tri = np.stack([tri[np.arange(len(tri)), (roll + k) % 3] for k in range(3)], axis=1)   # same orientation, another first node
        bnd = mm.channel_boundaries(nx, ny)
        a = mm.dual_mesh(coord, tri, bnd, native=True); b = mm.dual_mesh(coord, tri, bnd, native=False)
        for k in ["nEdge", "edgeNodes", "edgeNormal", "volume", "nbPtr", "nbPoint"]:
            assert np.array_equal(a[k], b[k]), (nx, ny, k)
        for ma, mb in zip(a["markers"], b["markers"]):
            for k in ["node", "normalNeighbor", "normal"]:
                assert np.array_equal(ma[k], mb[k]), (nx, ny, ma["tag"], k)
        assert abs(a["volume"].sum() - 0.25 * 0.025) <= 1e-12 * 0.25 * 0.025   # the dual volumes tile the channel
    L = pk.load_library()
    import ctypes as C
    assert L.rx_dual_mesh_tri(C.c_int64(0), None, C.c_int64(0), None, None) != 0   # argument errors are statuses, not crashes
