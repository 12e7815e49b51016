"""Flat-array builder of the vertex-centred median-dual mesh for synthetic triangulated 2-D channels.

Produces exactly the arrays the hot path reads from CGeometry (edge list + dual-face normals, control volumes,
neighbour lists, boundary vertex normals and normal neighbours) without SU2's object-per-point preprocessing, so the
4M / 32M / 256M-cell benchmark meshes of BASELINE.json can be built in seconds.  The formulas are SU2's:
  CPhysicalGeometry::SetControlVolume      Common/src/geometry_structure.cpp:10457   (CEdge::SetNodes_Coord / GetVolume,
                                            Common/src/dual_grid_structure.cpp:455-525)
  CPhysicalGeometry::SetBoundControlVolume geometry_structure.cpp:9595               (CVertex::SetNodes_Coord :618-636)
  CPhysicalGeometry::FindNormal_Neighbor   geometry_structure.cpp:12610
Numbering is our own (points row-major, edges sorted by (lower node, higher node)); tests/test_mesh.py checks the values
against the reference's geometry dump of the same .su2 mesh.
"""
import os

import numpy as np

KIND_INLET, KIND_OUTLET, KIND_ISOTHERMAL, KIND_HEATFLUX, KIND_EULER_WALL, KIND_INLET_TOTAL, KIND_INLET_MASSFLOW, KIND_SUPERSONIC_INLET, KIND_SUPERSONIC_OUTLET = 1, 2, 3, 4, 5, 6, 7, 8, 9


def channel_points(nx, ny, L=0.25, H=0.025, stretch=1.0):
    xs = np.linspace(0.0, L, nx)
    t = np.linspace(0.0, 1.0, ny)
    if stretch != 1.0:
        t = 0.5 * (1.0 + np.tanh(stretch * (2 * t - 1)) / np.tanh(stretch))
    ys = H * t
    X, Y = np.meshgrid(xs, ys)          # [ny, nx], point id = j*nx + i
    return np.stack([X.ravel(), Y.ravel()], axis=1)


def channel_triangles(nx, ny):
    """Every quad split in two with alternating diagonal (same as oracle/make_golden.py:write_channel_su2)."""
    i, j = np.meshgrid(np.arange(nx - 1), np.arange(ny - 1))
    i = i.ravel(); j = j.ravel()
    a = j * nx + i; b = a + 1; c = b + nx; d = a + nx
    even = ((i + j) % 2) == 0
    t0 = np.where(even[:, None], np.stack([a, b, c], 1), np.stack([a, b, d], 1))
    t1 = np.where(even[:, None], np.stack([a, c, d], 1), np.stack([b, c, d], 1))
    tri = np.empty((2 * len(a), 3), dtype=np.int64)
    tri[0::2] = t0; tri[1::2] = t1
    return tri


def dual_mesh(coord, tri, boundaries, nPointDomain=None, native=None):
    """coord [nP,2], tri [nT,3], boundaries = list of (tag, line elements [nL,2]) -> flat mesh dict (see rx_cuda.h: rx_mesh).
    Edges, normals, volumes and the neighbour CSR come from the library's host builder (rx_dual_mesh_tri, csrc/host/rx_geometry.cpp: one
    linear pass per array, 4-5 x faster on the 16 M-point meshes) when the library is built, else — or with native=False — from the numpy
    statements below, which are its specification: both give the same bits (tests/test_mesh.py)."""
    nP = coord.shape[0]
    x = coord[:, 0]; y = coord[:, 1]
    if native is None:
        native = os.environ.get("RX_MESH_NUMPY", "0") != "1"
    dm = None
    if native:
        try:
            from .rx_solver import dual_mesh_tri
            dm = dual_mesh_tri(coord, tri)
        except Exception:
            dm = None   # library not built yet (the CPU-only tools still work)
    if dm is not None:
        nE = dm["nEdge"]; ei = dm["edgeNodes"][:, 0].copy(); ej = dm["edgeNodes"][:, 1].copy()
        enx = dm["edgeNormal"][:, 0].copy(); eny = dm["edgeNormal"][:, 1].copy()
        vol = dm["volume"]; nbPtr = dm["nbPtr"]; nbPoint = dm["nbPoint"]
        return _dual_mesh_finish(coord, boundaries, nPointDomain, nP, x, y, nE, ei, ej, enx, eny, vol, nbPtr, nbPoint)
    # element CG as ((x0 + x1) + x2)/3.  The reference adds x_n/3 node by node (CPrimalGrid::SetCoord_CG); geometry.su2_dual_mesh
    # reproduces that bit for bit — this builder, used for the large synthetic meshes, is within one ulp of it.
    ecx = ((x[tri[:, 0]] + x[tri[:, 1]]) + x[tri[:, 2]]) / 3.0
    ecy = ((y[tri[:, 0]] + y[tri[:, 1]]) + y[tri[:, 2]]) / 3.0
    fa = np.concatenate([tri[:, 0], tri[:, 1], tri[:, 2]])
    fb = np.concatenate([tri[:, 1], tri[:, 2], tri[:, 0]])
    ex = np.concatenate([ecx, ecx, ecx]); ey = np.concatenate([ecy, ecy, ecy])
    key = np.minimum(fa, fb) * nP + np.maximum(fa, fb)
    ukey, inv = np.unique(key, return_inverse=True)
    del key                                                    # the large meshes are built per rank: keep the peak down
    nE = ukey.shape[0]
    ei = ukey // nP; ej = ukey % nP
    del ukey
    mx = 0.5 * (x[fa] + x[fb]); my = 0.5 * (y[fa] + y[fb])     # edge CG
    flip = fa > fb
    nxc = np.where(flip, my - ey, ey - my)
    enx = np.bincount(inv, weights=nxc, minlength=nE)
    del nxc
    nyc = np.where(flip, -(mx - ex), -(ex - mx))
    eny = np.bincount(inv, weights=nyc, minlength=nE)
    del nyc, inv, flip
    # control volumes: 0.5*|(E - P) x (M - P)| for both end points of every element face
    def area(p):
        xp = x[p]; yp = y[p]
        return 0.5 * np.abs((ex - xp) * (my - yp) - (ey - yp) * (mx - xp))
    vol = np.bincount(fa, weights=area(fa), minlength=nP) + np.bincount(fb, weights=area(fb), minlength=nP)
    del fa, fb, ex, ey, mx, my
    # neighbour CSR (sorted by neighbour index)
    src = np.concatenate([ei, ej]); dst = np.concatenate([ej, ei])
    order = np.lexsort((dst, src))
    nbPoint = dst[order]
    nbPtr = np.zeros(nP + 1, dtype=np.int64)
    np.cumsum(np.bincount(src, minlength=nP), out=nbPtr[1:])
    return _dual_mesh_finish(coord, boundaries, nPointDomain, nP, x, y, nE, ei, ej, enx, eny, vol, nbPtr, nbPoint)


def _dual_mesh_finish(coord, boundaries, nPointDomain, nP, x, y, nE, ei, ej, enx, eny, vol, nbPtr, nbPoint):
    """boundary markers (vertex lists, vertex normals, normal neighbours) and the mesh dict"""
    markers = []
    for tag, lines in boundaries:
        lines = np.asarray(lines, dtype=np.int64)
        seq = lines.ravel()                                   # vertex order = first appearance in the element loop
        _, first = np.unique(seq, return_index=True)
        node = seq[np.sort(first)]
        vid = np.full(nP, -1, dtype=np.int64); vid[node] = np.arange(len(node))
        cx = 0.5 * (x[lines[:, 0]] + x[lines[:, 1]]); cy = 0.5 * (y[lines[:, 0]] + y[lines[:, 1]])
        n0, n1 = lines[:, 0], lines[:, 1]
        vnx = np.zeros(len(node)); vny = np.zeros(len(node))
        # node 0: SetNodes_Coord(Elem_CG, Vertex) -> (Vy - Cy, -(Vx - Cx)); node 1: SetNodes_Coord(Vertex, Elem_CG) -> (Cy - Vy, -(Cx - Vx))
        np.add.at(vnx, vid[n0], y[n0] - cy); np.add.at(vny, vid[n0], -(x[n0] - cx))
        np.add.at(vnx, vid[n1], cy - y[n1]); np.add.at(vny, vid[n1], -(cx - x[n1]))
        # FindNormal_Neighbor: neighbour with the largest cosine to the vertex normal (last one on ties, `>=`)
        nn = np.empty(len(node), dtype=np.int64)
        for k, p in enumerate(node):
            nb = nbPoint[nbPtr[p]:nbPtr[p + 1]]
            dx = x[nb] - x[p]; dy = y[nb] - y[p]
            cos = (dx * vnx[k] + dy * vny[k]) / (np.sqrt(dx * dx + dy * dy) * np.sqrt(vnx[k] ** 2 + vny[k] ** 2))
            best = -1.0; arg = 0
            for q in range(len(nb)):
                if cos[q] >= best:
                    best = cos[q]; arg = q
            nn[k] = nb[arg]
        markers.append(dict(tag=tag, kind=0, node=node, normalNeighbor=nn, normal=np.stack([vnx, vny], 1)))
    return dict(nPoint=nP, nPointDomain=nP if nPointDomain is None else nPointDomain, nEdge=nE,
                edgeNodes=np.stack([ei, ej], 1), edgeNormal=np.stack([enx, eny], 1), coord=np.ascontiguousarray(coord),
                volume=vol, nbPtr=nbPtr, nbPoint=nbPoint, markers=markers)


def channel_boundaries(nx, ny):
    j = np.arange(ny - 1); i = np.arange(nx - 1)
    # line elements are oriented with the domain on their left, which is how CPhysicalGeometry::Check_BoundElem_Orientation leaves
    # them; the vertex normals SetBoundControlVolume forms from them point INTO the domain (the BC routines negate them)
    return [("inlet", np.stack([(j + 1) * nx, j * nx], 1)),
            ("outlet", np.stack([j * nx + nx - 1, (j + 1) * nx + nx - 1], 1)),
            ("lower_wall", np.stack([i, i + 1], 1)),
            ("upper_wall", np.stack([(ny - 1) * nx + i + 1, (ny - 1) * nx + i], 1))]


def channel_mesh(nx, ny, L=0.25, H=0.025, stretch=1.0, Ns=9, oxidizer=2):
    """The synthetic channel of SURVEY.md §8(d): inlet (TEMPERATURE_IMPOSE, pure oxidiser, 300 K, 20 m/s), outlet 101325 Pa,
    isothermal walls (300 K upper, 600 K lower) — the boundary set of tests/golden/channel_*.npz."""
    coord = channel_points(nx, ny, L, H, stretch)
    m = dual_mesh(coord, channel_triangles(nx, ny), channel_boundaries(nx, ny))
    ys = np.zeros(Ns); ys[oxidizer] = 1.0
    m["markers"][0].update(kind=KIND_INLET, T=300.0, Vmag=20.0, dir=[1.0, 0.0], Ys=ys)
    m["markers"][1].update(kind=KIND_OUTLET, P=101325.0)
    m["markers"][2].update(kind=KIND_ISOTHERMAL, Twall=600.0)
    m["markers"][3].update(kind=KIND_ISOTHERMAL, Twall=300.0)
    return m


def strip_owner(mesh, nParts):
    """Owner rank of every point for a strip partition in x (equal point counts)."""
    nP = mesh["nPoint"]
    order = np.argsort(mesh["coord"][:, 0], kind="stable")
    owner = np.empty(nP, dtype=np.int64)
    bounds = [(nP * r) // nParts for r in range(nParts + 1)]
    for r in range(nParts):
        owner[order[bounds[r]:bounds[r + 1]]] = r
    return owner


def graph_owner(mesh, nParts, refine_passes=8):
    """Owner rank of every point from the k-way graph partitioner of the library (rx_partition_kway: recursive inertial
    bisection + edge-cut refinement), the stand-in for the reference's ParMETIS_V3_PartKway call
    (Common/src/geometry_structure.cpp:11524).  Returns (owner, edge_cut)."""
    import ctypes as C
    from .rx_solver import load_library, RxError
    L = load_library()
    nP = int(mesh["nPoint"])
    ptr = np.ascontiguousarray(mesh["nbPtr"], dtype=np.int64); adj = np.ascontiguousarray(mesh["nbPoint"], dtype=np.int64)
    xy = np.ascontiguousarray(mesh["coord"], dtype=np.float64)
    owner = np.empty(nP, dtype=np.int64); cut = C.c_int64(0)
    I = C.POINTER(C.c_int64); D = C.POINTER(C.c_double)
    rc = L.rx_partition_kway(C.c_int64(nP), ptr.ctypes.data_as(I), adj.ctypes.data_as(I), xy.ctypes.data_as(D), int(nParts),
                             int(refine_passes), owner.ctypes.data_as(I), C.byref(cut))
    if rc != 0:
        raise RxError("rx_partition_kway failed (%d)" % rc)
    return owner, int(cut.value)


def rcm_order(mesh):
    """new -> old point permutation: reverse Cuthill-McKee over the owned points, halo points kept last (rx_rcm_order;
    role of CPhysicalGeometry::SetRCM_Ordering, geometry_structure.cpp:9200)."""
    import ctypes as C
    from .rx_solver import load_library, RxError
    L = load_library()
    ptr = np.ascontiguousarray(mesh["nbPtr"], dtype=np.int64); adj = np.ascontiguousarray(mesh["nbPoint"], dtype=np.int64)
    n2o = np.empty(int(mesh["nPoint"]), dtype=np.int64)
    I = C.POINTER(C.c_int64)
    rc = L.rx_rcm_order(C.c_int64(int(mesh["nPoint"])), C.c_int64(int(mesh["nPointDomain"])), ptr.ctypes.data_as(I), adj.ctypes.data_as(I),
                        n2o.ctypes.data_as(I))
    if rc != 0:
        raise RxError("rx_rcm_order failed (%d)" % rc)
    return n2o


def renumber_points(mesh, n2o, sort_edges=True):
    """The same flat mesh with points renumbered (new point q = old point n2o[q]).  Edges keep their orientation (the
    normal points from the first to the second node); with sort_edges they are re-sorted by (lower, higher) new index so
    that the edge kernels walk the point arrays front to back."""
    n2o = np.asarray(n2o, dtype=np.int64)
    nP = int(mesh["nPoint"])
    o2n = np.empty(nP, dtype=np.int64); o2n[n2o] = np.arange(nP)
    en = o2n[mesh["edgeNodes"]]
    normal = mesh["edgeNormal"]
    if sort_edges:
        eo = np.lexsort((en.max(axis=1), en.min(axis=1)))
        en = en[eo]; normal = normal[eo]
    cnt = (mesh["nbPtr"][1:] - mesh["nbPtr"][:-1])[n2o]
    nbPtr = np.zeros(nP + 1, dtype=np.int64); np.cumsum(cnt, out=nbPtr[1:])
    idx = np.repeat(mesh["nbPtr"][:-1][n2o] - nbPtr[:-1], cnt) + np.arange(nbPtr[-1])
    nb = o2n[mesh["nbPoint"][idx]]
    # neighbour lists stay sorted by neighbour index (the order the WLS sums run in is part of the numbering)
    seg = np.repeat(np.arange(nP), cnt)
    nb = nb[np.lexsort((nb, seg))]
    out = dict(mesh)
    out.update(edgeNodes=np.ascontiguousarray(en), edgeNormal=np.ascontiguousarray(normal), coord=mesh["coord"][n2o], volume=mesh["volume"][n2o],
               nbPtr=nbPtr, nbPoint=nb)
    mk = []
    for m in mesh["markers"]:
        d = dict(m); d.update(node=o2n[m["node"]], normalNeighbor=o2n[m["normalNeighbor"]]); mk.append(d)
    out["markers"] = mk
    return out, o2n


def partition_mesh(mesh, owner, rcm=False):
    """Splits a flat mesh by an owner array into ranks with one halo layer, SU2's scheme:
    owned points first, then halo points grouped by owner rank; every edge with at least one owned end is kept (halo-owned
    edges are duplicated on both sides); boundary vertices only on owned points.  Returns a list of dict(mesh=local flat
    mesh, glob=local->global point ids, send={rank: local owned ids}, recv={rank: local halo ids}) with matching send/recv
    orders.  rcm=True renumbers the owned points of every part by reverse Cuthill-McKee and re-sorts its edges."""
    owner = np.asarray(owner, dtype=np.int64)
    nParts = int(owner.max()) + 1
    nP = mesh["nPoint"]
    ei, ej = mesh["edgeNodes"][:, 0], mesh["edgeNodes"][:, 1]
    nbPtr, nbPoint = mesh["nbPtr"], mesh["nbPoint"]
    parts = []
    for r in range(nParts):
        own = np.flatnonzero(owner == r)
        keep = (owner[ei] == r) | (owner[ej] == r)
        e_i, e_j = ei[keep], ej[keep]
        touched = np.unique(np.concatenate([e_i, e_j]))
        halo = touched[owner[touched] != r]
        halo = halo[np.lexsort((halo, owner[halo]))]           # grouped by owner rank, ascending global id inside
        glob = np.concatenate([own, halo])
        g2l = np.full(nP, -1, dtype=np.int64); g2l[glob] = np.arange(len(glob))
        loc = dict(nPoint=len(glob), nPointDomain=len(own), nEdge=int(keep.sum()),
                   edgeNodes=np.stack([g2l[e_i], g2l[e_j]], 1), edgeNormal=mesh["edgeNormal"][keep],
                   coord=mesh["coord"][glob], volume=mesh["volume"][glob])
        # neighbour CSR restricted to local points (owned points keep all their neighbours by construction)
        cnt = nbPtr[glob + 1] - nbPtr[glob]
        off = np.zeros(len(glob) + 1, dtype=np.int64); np.cumsum(cnt, out=off[1:])
        idx = np.repeat(nbPtr[glob] - off[:-1], cnt) + np.arange(off[-1])
        nb = g2l[nbPoint[idx]]
        ok = nb >= 0
        seg = np.repeat(np.arange(len(glob)), cnt)[ok]
        loc["nbPtr"] = np.zeros(len(glob) + 1, dtype=np.int64)
        np.cumsum(np.bincount(seg, minlength=len(glob)), out=loc["nbPtr"][1:])
        loc["nbPoint"] = nb[ok]
        mk = []
        for m in mesh["markers"]:
            sel = owner[m["node"]] == r
            d = dict(m)
            d.update(node=g2l[m["node"][sel]], normalNeighbor=g2l[m["normalNeighbor"][sel]], normal=m["normal"][sel])
            mk.append(d)
        loc["markers"] = mk
        if rcm:
            loc, o2n = renumber_points(loc, rcm_order(loc))
            n2o = np.empty_like(o2n); n2o[o2n] = np.arange(len(o2n))
            glob = glob[n2o]
            g2l = np.full(nP, -1, dtype=np.int64); g2l[glob] = np.arange(len(glob))
        recv = {int(q): g2l[halo[owner[halo] == q]] for q in np.unique(owner[halo])}
        parts.append(dict(mesh=loc, glob=glob, recv=recv, halo_glob=halo))
    for r in range(nParts):
        send = {}
        g2l = np.full(nP, -1, dtype=np.int64); g2l[parts[r]["glob"]] = np.arange(len(parts[r]["glob"]))
        for q in range(nParts):
            if q != r and r in parts[q]["recv"]:
                hg = parts[q]["halo_glob"]
                want = hg[owner[hg] == r]                       # global ids rank q expects from r, in q's halo order
                send[q] = g2l[want]
        parts[r]["send"] = send
    for p in parts:
        p.pop("halo_glob")
    return parts


def partition_rank(mesh, owner, rank, rcm=False):
    """One rank's entry of partition_mesh(mesh, owner, rcm), built by the library's host code (rx_partition_mesh, csrc/host/rx_partition.cpp):
    every rank of a job builds its own part in one pass over the whole mesh instead of all of them.  Same arrays, same send / receive lists."""
    import ctypes as C
    from .rx_solver import load_library, RxError

    class _Part(C.Structure):
        _fields_ = [("nPoint", C.c_int64), ("nPointDomain", C.c_int64), ("nEdge", C.c_int64), ("glob", C.POINTER(C.c_int64)),
                    ("edgeNodes", C.POINTER(C.c_int64)), ("edgeNormal", C.POINTER(C.c_double)), ("coord", C.POINTER(C.c_double)),
                    ("volume", C.POINTER(C.c_double)), ("nbPtr", C.POINTER(C.c_int64)), ("nbPoint", C.POINTER(C.c_int64)), ("nMarker", C.c_int),
                    ("markerPtr", C.POINTER(C.c_int64)), ("markerNode", C.POINTER(C.c_int64)), ("markerNormalNeighbor", C.POINTER(C.c_int64)),
                    ("markerNormal", C.POINTER(C.c_double)), ("nNb", C.c_int), ("nbRank", C.POINTER(C.c_int)),
                    ("sendPtr", C.POINTER(C.c_int64)), ("sendPts", C.POINTER(C.c_int64)), ("recvPtr", C.POINTER(C.c_int64)), ("recvPts", C.POINTER(C.c_int64))]
    L = load_library()
    k64 = lambda a: np.ascontiguousarray(a, dtype=np.int64); kf = lambda a: np.ascontiguousarray(a, dtype=np.float64)
    I = lambda a: a.ctypes.data_as(C.POINTER(C.c_int64)); D = lambda a: a.ctypes.data_as(C.POINTER(C.c_double))
    en, nr, xy, vol = k64(mesh["edgeNodes"]), kf(mesh["edgeNormal"]), kf(mesh["coord"]), kf(mesh["volume"])
    ptr, nb, ow = k64(mesh["nbPtr"]), k64(mesh["nbPoint"]), k64(owner)
    mk = mesh["markers"]
    mptr = np.zeros(len(mk) + 1, dtype=np.int64)
    if mk:
        np.cumsum([len(m["node"]) for m in mk], out=mptr[1:])
    cat = lambda key, dt, shape: np.ascontiguousarray(np.concatenate([np.asarray(m[key], dtype=dt).reshape(shape) for m in mk]) if mk else np.zeros(shape, dt).reshape(shape)[:0])
    mnode, mnn, mnr = cat("node", np.int64, (-1,)), cat("normalNeighbor", np.int64, (-1,)), cat("normal", np.float64, (-1, 2))
    out = _Part()
    rc = L.rx_partition_mesh(C.c_int64(int(mesh["nPoint"])), C.c_int64(en.shape[0]), I(en), D(nr), D(xy), D(vol), I(ptr), I(nb), C.c_int(len(mk)), I(mptr),
                             I(mnode), I(mnn), D(mnr), I(ow), C.c_int(int(rank)), C.c_int(1 if rcm else 0), C.byref(out))
    if rc != 0:
        raise RxError("rx_partition_mesh failed (%d): rank %d owns no point, or the mesh arrays are inconsistent" % (rc, rank))
    try:
        arr = lambda p, n, shape=None: (np.ctypeslib.as_array(p, shape=(n,)).copy() if n else np.zeros(0, dtype=np.ctypeslib.as_array(p, shape=(1,)).dtype)).reshape(shape or (n,))
        nP, nE, nM, nN = out.nPoint, out.nEdge, out.nMarker, out.nNb
        lp = arr(out.markerPtr, nM + 1)
        nV = int(lp[-1])
        lnode, lnn, lnr = arr(out.markerNode, nV), arr(out.markerNormalNeighbor, nV), arr(out.markerNormal, 2 * nV, (nV, 2))
        markers = []
        for k, m in enumerate(mk):
            d = dict(m); a, b = int(lp[k]), int(lp[k + 1])
            d.update(node=lnode[a:b], normalNeighbor=lnn[a:b], normal=lnr[a:b]); markers.append(d)
        loc = dict(nPoint=nP, nPointDomain=out.nPointDomain, nEdge=nE, edgeNodes=arr(out.edgeNodes, 2 * nE, (nE, 2)), edgeNormal=arr(out.edgeNormal, 2 * nE, (nE, 2)),
                   coord=arr(out.coord, 2 * nP, (nP, 2)), volume=arr(out.volume, nP), nbPtr=arr(out.nbPtr, nP + 1), nbPoint=arr(out.nbPoint, int(out.nbPtr[nP])),
                   markers=markers)
        ranks = arr(out.nbRank, nN); sp, rp = arr(out.sendPtr, nN + 1), arr(out.recvPtr, nN + 1)
        spts, rpts = arr(out.sendPts, int(sp[-1])), arr(out.recvPts, int(rp[-1]))
        send = {int(q): spts[sp[k]:sp[k + 1]] for k, q in enumerate(ranks)}
        recv = {int(q): rpts[rp[k]:rp[k + 1]] for k, q in enumerate(ranks)}
        return dict(mesh=loc, glob=arr(out.glob, nP), recv=recv, send=send)
    finally:
        L.rx_part_free(C.byref(out))


def partition_strips(mesh, nParts):
    """Strip partition in x (deterministic, no graph partitioner): partition_mesh over strip_owner."""
    return partition_mesh(mesh, strip_owner(mesh, nParts))


def partition_graph(mesh, nParts, rcm=True, refine_passes=8):
    """k-way graph partition of any flat mesh (unstructured included) + RCM numbering inside every part."""
    owner, _ = graph_owner(mesh, nParts, refine_passes)
    return partition_mesh(mesh, owner, rcm=rcm)


def partition_graph_rank(mesh, nParts, rank, rcm=True, refine_passes=8):
    """partition_graph(mesh, nParts, rcm, refine_passes)[rank] without building the other ranks' parts (rx_partition_mesh): what each rank of a
    multi-GPU job calls — the k-way partition is deterministic, so every rank computes the same owner array."""
    owner, _ = graph_owner(mesh, nParts, refine_passes)
    return partition_rank(mesh, owner, rank, rcm=rcm)


def strip_short_side_slow(mesh, plan, min_ratio=3.0):
    """Renumber the owned points of a strip so that the SHORT side of the strip is the slow index (x-major when the strip is narrower
    than it is tall), halo points kept last, edges re-sorted.  The numbering is what the Gauss-Seidel sweeps of the SST step follow
    (CSysMatrix::ComputeLU_SGSPreconditioner walks the points in index order): their dependency levels number about
    (long side) + 2 (short side) this way instead of (short side) + 2 (long side) — 5 001 instead of 8 502 for a 500 x 4001 strip —
    the role SetRCM_Ordering (geometry_structure.cpp:9200) has in the reference.  Applied only to strips at least `min_ratio` times
    taller than wide: the x-major numbering costs the viscous edge kernel about 3 % (its gathers reach 4001 instead of 500 points
    away), which the shorter chains repay from that aspect ratio on (measured on the 32 M-cell case: 2 GPUs 267 -> 275 ms without
    the threshold, 4 GPUs 144 -> 138 ms, 8 GPUs 84 -> 75 ms).  Returns (mesh, plan) renumbered consistently."""
    n = int(mesh["nPointDomain"])
    gi, gj = plan["glob"]
    if min_ratio * len(np.unique(gi[:n])) > len(np.unique(gj[:n])):
        return mesh, plan
    own = np.lexsort((gj[:n], gi[:n]))
    n2o = np.concatenate([own, np.arange(n, int(mesh["nPoint"]))])
    out, o2n = renumber_points(mesh, n2o)
    p2 = dict(plan)
    p2["send"] = {r: o2n[v] for r, v in plan["send"].items()}
    p2["recv"] = {r: o2n[v] for r, v in plan["recv"].items()}
    p2["glob"] = (gi[n2o], gj[n2o])
    return out, p2


def channel_strip_mesh(nxg, ny, rank, nranks, L=0.25, H=0.025, stretch=1.0, Ns=9, oxidizer=2):
    """Local mesh of rank `rank` for a strip partition in x of the global nxg x ny channel, built WITHOUT the global mesh
    (weak scaling: every rank builds only its own strip).  SU2's scheme: owned points first, then one
    halo layer (left neighbour's column, then right neighbour's); every edge with at least one owned end is kept; boundary
    vertices only on owned points.  Returns (mesh, plan) with plan = dict(send={rank: local ids}, recv={rank: local ids},
    glob=(global i, j) of the local points, bbox=global bounding box); send/recv orders match (ascending j)."""
    c0 = (nxg * rank) // nranks; c1 = (nxg * (rank + 1)) // nranks
    hl = 1 if rank > 0 else 0; hr = 1 if rank < nranks - 1 else 0
    a0, a1 = c0 - hl, c1 + hr                      # local column range incl. halos
    nxl = a1 - a0
    xs = np.linspace(0.0, L, nxg)[a0:a1]
    t = np.linspace(0.0, 1.0, ny)
    if stretch != 1.0:
        t = 0.5 * (1.0 + np.tanh(stretch * (2 * t - 1)) / np.tanh(stretch))
    X, Y = np.meshgrid(xs, H * t)
    coord = np.stack([X.ravel(), Y.ravel()], axis=1)            # local id = j*nxl + il
    i, j = np.meshgrid(np.arange(nxl - 1), np.arange(ny - 1)); i = i.ravel(); j = j.ravel()
    a = j * nxl + i; b = a + 1; c = b + nxl; d = a + nxl
    even = (((i + a0) + j) % 2) == 0                            # diagonal parity of the GLOBAL grid
    tri = np.empty((2 * len(a), 3), dtype=np.int64)
    tri[0::2] = np.where(even[:, None], np.stack([a, b, c], 1), np.stack([a, b, d], 1))
    tri[1::2] = np.where(even[:, None], np.stack([a, c, d], 1), np.stack([b, c, d], 1))
    jj = np.arange(ny - 1); ii = np.arange(nxl - 1)
    bnd = []
    bnd.append(("inlet", np.stack([(jj + 1) * nxl, jj * nxl], 1) if rank == 0 else np.zeros((0, 2), dtype=np.int64)))
    bnd.append(("outlet", np.stack([jj * nxl + nxl - 1, (jj + 1) * nxl + nxl - 1], 1) if rank == nranks - 1 else np.zeros((0, 2), dtype=np.int64)))
    bnd.append(("lower_wall", np.stack([ii, ii + 1], 1)))
    bnd.append(("upper_wall", np.stack([(ny - 1) * nxl + ii + 1, (ny - 1) * nxl + ii], 1)))
    m = dual_mesh(coord, tri, bnd)
    il = np.tile(np.arange(nxl), ny); jl = np.repeat(np.arange(ny), nxl)
    owned = (il >= hl) & (il < nxl - hr)
    left = il < hl; right = il >= nxl - hr
    order = np.concatenate([np.flatnonzero(owned), np.flatnonzero(left), np.flatnonzero(right)])   # new -> old
    o2n = np.empty(nxl * ny, dtype=np.int64); o2n[order] = np.arange(nxl * ny)
    nOwn = int(owned.sum())
    ei, ej = m["edgeNodes"][:, 0], m["edgeNodes"][:, 1]
    keep = owned[ei] | owned[ej]
    # keep the edge orientation (normal points from first to second node); only the numbering changes
    mesh = dict(nPoint=nxl * ny, nPointDomain=nOwn, nEdge=int(keep.sum()),
                edgeNodes=np.stack([o2n[ei[keep]], o2n[ej[keep]]], 1), edgeNormal=m["edgeNormal"][keep],
                coord=m["coord"][order], volume=m["volume"][order])
    cnt = (m["nbPtr"][1:] - m["nbPtr"][:-1])[order]
    nbPtr = np.zeros(nxl * ny + 1, dtype=np.int64); np.cumsum(cnt, out=nbPtr[1:])
    starts = m["nbPtr"][:-1][order]
    idx = np.repeat(starts - nbPtr[:-1], cnt) + np.arange(nbPtr[-1])
    mesh["nbPtr"] = nbPtr; mesh["nbPoint"] = o2n[m["nbPoint"][idx]]
    mk = []
    for q in m["markers"]:
        sel = owned[q["node"]] if len(q["node"]) else np.zeros(0, dtype=bool)
        mk.append(dict(tag=q["tag"], kind=0, node=o2n[q["node"][sel]], normalNeighbor=o2n[q["normalNeighbor"][sel]], normal=q["normal"][sel]))
    ys = np.zeros(Ns); ys[oxidizer] = 1.0
    mk[0].update(kind=KIND_INLET, T=300.0, Vmag=20.0, dir=[1.0, 0.0], Ys=ys)
    mk[1].update(kind=KIND_OUTLET, P=101325.0)
    mk[2].update(kind=KIND_ISOTHERMAL, Twall=600.0)
    mk[3].update(kind=KIND_ISOTHERMAL, Twall=300.0)
    mesh["markers"] = mk
    send, recv = {}, {}
    if hl:
        send[rank - 1] = o2n[np.flatnonzero(il == hl)]          # my first owned column, ascending j
        recv[rank - 1] = o2n[np.flatnonzero(left)]
    if hr:
        send[rank + 1] = o2n[np.flatnonzero(il == nxl - hr - 1)]
        recv[rank + 1] = o2n[np.flatnonzero(right)]
    plan = dict(send=send, recv=recv, glob=(il[order] + a0, jl[order]), bbox=(0.0, L, 0.0, H))
    return mesh, plan
