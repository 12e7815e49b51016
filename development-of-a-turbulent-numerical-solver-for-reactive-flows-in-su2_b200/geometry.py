"""SU2 mesh file -> flat dual mesh in the REFERENCE's numbering (geometry preprocessing without the
reference's object-per-point structures).  2-D meshes of triangles and quadrilaterals.

Follows CDriver::Geometrical_Preprocessing (driver_structure.cpp) step by step, with flat arrays:
  read_su2                    CPhysicalGeometry::Read_SU2_Format (geometry_structure.cpp:4819): NDIME/NELEM/NPOIN/NMARK blocks
  point_connectivity          SetPoint_Connectivity (:9145): neighbours of a point in the order its elements (ascending) list them
                              (CTriangle/CQuadrilateral::Neighbor_Nodes, primal_grid_structure.cpp:194,260), first appearance kept
  rcm via rx_rcm_order        SetRCM_Ordering (:9200): lowest-degree start, neighbours appended by increasing degree, reversed
  orient_elements             Check_IntElem_Orientation (:8542), Check_BoundElem_Orientation (:8825) and ::Change_Orientation
  edge numbering              CGeometry::SetEdges: points ascending, neighbours in order, an edge is numbered at its lower point
  control volumes             SetCoord_CG (:9518), SetControlVolume (:10457), SetVertex (:9451), SetBoundControlVolume (:9595),
                              FindNormal_Neighbor — the formulae of mesh.dual_mesh, summed in the reference's element order
  wall_distance               ComputeWall_Distance (:8990): distance to the nearest NODE of the HEAT_FLUX / ISOTHERMAL markers
                              (on the GPU: rx_wall_distance, csrc/rx_cuda.cu)
The result feeds CReactiveNSSolverGPU directly; tests/test_geometry.py compares every array with the geometry the reference
itself built from the same file (golden fixtures `geo_*`)."""
import ctypes as C

import numpy as np

TRIANGLE, QUADRILATERAL, LINE = 5, 9, 3
_NNODE = {TRIANGLE: 3, QUADRILATERAL: 4}
_NEIGHBOR = {TRIANGLE: [[1, 2], [2, 0], [0, 1]], QUADRILATERAL: [[1, 3], [2, 0], [3, 1], [0, 2]]}
EPS = 1.0e-16


def read_su2(path):
    """-> dict(coord [nP,2], etype [nE], elems [nE,4] (-1 padded), markers [(tag, lines [n,2])]) in the file's numbering."""
    with open(path) as f:
        lines = [ln.strip() for ln in f]
    k = 0
    out = dict(markers=[])

    def val(s):
        return s.split("=")[1].strip()
    while k < len(lines):
        ln = lines[k]
        if ln.startswith("NDIME"):
            if int(val(ln)) != 2:
                raise ValueError("read_su2: 2-D meshes only")
        elif ln.startswith("NELEM"):
            n = int(val(ln))
            etype = np.empty(n, dtype=np.int64); elems = np.full((n, 4), -1, dtype=np.int64)
            for e in range(n):
                t = lines[k + 1 + e].split()
                etype[e] = int(t[0])
                nn = _NNODE.get(int(t[0]))
                if nn is None:
                    raise ValueError("read_su2: element type %s is not a triangle or quadrilateral" % t[0])
                elems[e, :nn] = [int(v) for v in t[1:1 + nn]]
            out["etype"], out["elems"] = etype, elems
            k += n
        elif ln.startswith("NPOIN"):
            n = int(val(ln).split()[0])
            out["coord"] = np.array([[float(v) for v in lines[k + 1 + p].split()[:2]] for p in range(n)])
            k += n
        elif ln.startswith("MARKER_TAG"):
            tag = val(ln)
            k += 1
            n = int(val(lines[k]))
            seg = np.array([[int(v) for v in lines[k + 1 + e].split()[1:3]] for e in range(n)], dtype=np.int64).reshape(n, 2)
            out["markers"].append((tag, seg))
            k += n
        k += 1
    return out


def point_connectivity(nP, etype, elems):
    """(nbPtr, nbPoint): CSR of the point neighbours in the reference's order (SetPoint_Connectivity)."""
    P, E, S, Q = [], [], [], []
    for t, nb in _NEIGHBOR.items():
        idx = np.nonzero(etype == t)[0]
        if idx.size == 0:
            continue
        for i, (a, b) in enumerate(nb):
            for s, j in enumerate((a, b)):
                P.append(elems[idx, i]); Q.append(elems[idx, j]); E.append(idx); S.append(np.full(idx.size, s))
    P = np.concatenate(P); Q = np.concatenate(Q); E = np.concatenate(E); S = np.concatenate(S)
    o = np.lexsort((S, E, P))
    P, Q = P[o], Q[o]
    _, first = np.unique(P * nP + Q, return_index=True)
    keep = np.sort(first)
    P, Q = P[keep], Q[keep]
    nbPtr = np.zeros(nP + 1, dtype=np.int64)
    np.cumsum(np.bincount(P, minlength=nP), out=nbPtr[1:])
    return nbPtr, np.ascontiguousarray(Q)


def rcm_new_to_old(nP, nbPtr, nbPoint):
    from .rx_solver import load_library
    L = load_library()
    n2o = np.empty(nP, dtype=np.int64)
    pi = lambda a: a.ctypes.data_as(C.POINTER(C.c_int64))
    rc = L.rx_rcm_order(C.c_int64(nP), C.c_int64(nP), pi(nbPtr), pi(nbPoint), pi(n2o))
    if rc != 0:
        raise RuntimeError("rx_rcm_order failed (%d)" % rc)
    return n2o


def _cross_half(c, p1, p2, p3):
    a = 0.5 * (c[p2] - c[p1]); b = 0.5 * (c[p3] - c[p1])
    return a[:, 0] * b[:, 1] - b[:, 0] * a[:, 1]


def orient_elements(coord, etype, elems, boundaries):
    """Counter-clockwise elements, boundary lines with the domain on their left — their vertex normals then point into the domain,
    SU2's convention (the BC routines negate them); returns (elems, boundaries)."""
    elems = elems.copy()
    tri = np.nonzero(etype == TRIANGLE)[0]
    if tri.size:
        flip = tri[_cross_half(coord, elems[tri, 0], elems[tri, 1], elems[tri, 2]) < 0.0]
        elems[flip, 0], elems[flip, 2] = elems[flip, 2].copy(), elems[flip, 0].copy()
    qd = np.nonzero(etype == QUADRILATERAL)[0]
    if qd.size:
        n = [elems[qd, i] for i in range(4)]
        t1 = _cross_half(coord, n[0], n[1], n[2]); t2 = _cross_half(coord, n[1], n[2], n[3])
        t3 = _cross_half(coord, n[2], n[3], n[0]); t4 = _cross_half(coord, n[3], n[0], n[2])
        flip = qd[(t1 < 0.0) & (t2 < 0.0) & (t3 < 0.0) & (t4 < 0.0)]
        elems[flip, 1], elems[flip, 3] = elems[flip, 3].copy(), elems[flip, 1].copy()
    # domain element of a boundary line: the element that holds both of its nodes (SetBoundVolume); the test point is the first
    # element node that is not on the line
    nP = coord.shape[0]
    nn = np.where(etype == TRIANGLE, 3, 4)
    fkey, felem = [], []
    for i in range(4):
        j = np.where(i + 1 < nn, i + 1, 0)
        ok = i < nn
        a = elems[ok, i]; b = elems[np.nonzero(ok)[0], j[ok]]
        fkey.append(np.minimum(a, b) * nP + np.maximum(a, b)); felem.append(np.nonzero(ok)[0])
    fkey = np.concatenate(fkey); felem = np.concatenate(felem)
    o = np.argsort(fkey, kind="stable"); fkey, felem = fkey[o], felem[o]
    out = []
    for tag, seg in boundaries:
        seg = np.asarray(seg, dtype=np.int64).copy()
        key = np.minimum(seg[:, 0], seg[:, 1]) * nP + np.maximum(seg[:, 0], seg[:, 1])
        pos = np.searchsorted(fkey, key)
        if np.any(fkey[np.minimum(pos, fkey.size - 1)] != key):
            raise ValueError("marker %s: a boundary element is not a face of any element" % tag)
        de = felem[pos]
        third = np.empty(seg.shape[0], dtype=np.int64)
        for m in range(seg.shape[0]):
            for v in elems[de[m], :nn[de[m]]]:
                if v != seg[m, 0] and v != seg[m, 1]:
                    third[m] = v
                    break
        flip = _cross_half(coord, seg[:, 0], seg[:, 1], third) < 0.0
        seg[flip] = seg[flip][:, ::-1]
        out.append((tag, seg))
    return elems, out


def su2_dual_mesh(su2, rcm=True):
    """read_su2() result -> flat mesh dict (the layout of mesh.dual_mesh / include/rx_cuda.h: rx_mesh) in the reference's point,
    edge, neighbour and vertex numbering; `globalIndex` = the file's index of every point (CPoint::GetGlobalIndex)."""
    coord, etype, elems = su2["coord"], su2["etype"], su2["elems"]
    boundaries = [(t, np.asarray(s, dtype=np.int64)) for t, s in su2["markers"]]
    nP = coord.shape[0]
    glob = np.arange(nP, dtype=np.int64)
    if rcm:
        nbPtr, nbPoint = point_connectivity(nP, etype, elems)
        n2o = rcm_new_to_old(nP, nbPtr, nbPoint)
        o2n = np.empty(nP, dtype=np.int64); o2n[n2o] = np.arange(nP)
        coord = np.ascontiguousarray(coord[n2o]); glob = n2o
        elems = np.where(elems >= 0, o2n[np.maximum(elems, 0)], -1)
        boundaries = [(t, o2n[s]) for t, s in boundaries]
    nbPtr, nbPoint = point_connectivity(nP, etype, elems)
    elems, boundaries = orient_elements(coord, etype, elems, boundaries)
    # SetEdges: an edge is numbered when its lower point meets it
    src = np.repeat(np.arange(nP, dtype=np.int64), np.diff(nbPtr))
    up = nbPoint > src
    ei, ej = src[up], nbPoint[up]
    nE = ei.shape[0]
    ekey = ei * nP + ej
    eo = np.argsort(ekey); ekey_s = ekey[eo]
    x, y = coord[:, 0], coord[:, 1]
    # element CG (CPrimalGrid::SetCoord_CG, primal_grid_structure.cpp: CG += coord/nNodes, node by node, from zero)
    nn = np.where(etype == TRIANGLE, 3, 4)
    e3 = np.maximum(elems[:, 3], 0)
    fn = nn.astype(np.float64)
    ecx = ((0.0 + x[elems[:, 0]] / fn) + x[elems[:, 1]] / fn) + x[elems[:, 2]] / fn
    ecy = ((0.0 + y[elems[:, 0]] / fn) + y[elems[:, 1]] / fn) + y[elems[:, 2]] / fn
    ecx = np.where(nn == 4, ecx + x[e3] / fn, ecx); ecy = np.where(nn == 4, ecy + y[e3] / fn, ecy)
    # faces in element-major order (the order SetControlVolume adds in)
    nEl = elems.shape[0]
    fa = elems.copy(); fb = np.empty_like(elems)
    for i in range(4):
        j = np.where(i + 1 < nn, i + 1, 0)
        fb[:, i] = elems[np.arange(nEl), j]
    valid = (np.arange(4)[None, :] < nn[:, None]).ravel()
    fa = fa.ravel()[valid]; fb = fb.ravel()[valid]
    ex = np.repeat(ecx, 4)[valid]; ey = np.repeat(ecy, 4)[valid]
    fe = eo[np.searchsorted(ekey_s, np.minimum(fa, fb) * nP + np.maximum(fa, fb))]
    mx = 0.5 * (x[fa] + x[fb]); my = 0.5 * (y[fa] + y[fb])     # edge CG
    flip = fa > fb
    enx = np.bincount(fe, weights=np.where(flip, my - ey, ey - my), minlength=nE)
    eny = np.bincount(fe, weights=np.where(flip, -(mx - ex), -(ex - mx)), minlength=nE)
    zero = np.sqrt(enx * enx + eny * eny) == 0.0
    enx[zero] = EPS * EPS; eny[zero] = EPS * EPS

    def area(p):
        # CEdge::GetVolume(point, edge CG, element CG) as SetControlVolume calls it: the arguments land on (Edge_CG, Elem_CG, Point),
        # so the vectors are taken from the element CG (dual_grid_structure.cpp:455-476)
        return 0.5 * np.abs((mx - ex) * (y[p] - ey) - (my - ey) * (x[p] - ex))
    pts = np.stack([fa, fb], 1).ravel(); ar = np.stack([area(fa), area(fb)], 1).ravel()   # i then j of every face
    vol = np.bincount(pts, weights=ar, minlength=nP)
    markers = []
    for tag, seg in boundaries:
        flat = seg.ravel()
        _, first = np.unique(flat, return_index=True)
        node = flat[np.sort(first)]                             # SetVertex: first appearance in the element loop
        vid = np.full(nP, -1, dtype=np.int64); vid[node] = np.arange(len(node))
        cx = 0.5 * (x[seg[:, 0]] + x[seg[:, 1]]); cy = 0.5 * (y[seg[:, 0]] + y[seg[:, 1]])
        n0, n1 = seg[:, 0], seg[:, 1]
        vnx = np.zeros(len(node)); vny = np.zeros(len(node))
        np.add.at(vnx, vid[n0], y[n0] - cy); np.add.at(vny, vid[n0], -(x[n0] - cx))
        np.add.at(vnx, vid[n1], cy - y[n1]); np.add.at(vny, vid[n1], -(cx - x[n1]))
        nnb = np.empty(len(node), dtype=np.int64)
        for k, p in enumerate(node):                            # FindNormal_Neighbor: largest cosine, last one on ties
            nb = nbPoint[nbPtr[p]:nbPtr[p + 1]]
            dx = x[nb] - x[p]; dy = y[nb] - y[p]
            cos = (dx * vnx[k] + dy * vny[k]) / (np.sqrt(dx * dx + dy * dy) * np.sqrt(vnx[k] ** 2 + vny[k] ** 2))
            best = -1.0; arg = 0
            for q in range(len(nb)):
                if cos[q] >= best:
                    best = cos[q]; arg = q
            nnb[k] = nb[arg]
        markers.append(dict(tag=tag, kind=0, node=node, normalNeighbor=nnb, normal=np.stack([vnx, vny], 1)))
    return dict(nPoint=nP, nPointDomain=nP, nEdge=nE, edgeNodes=np.stack([ei, ej], 1), edgeNormal=np.stack([enx, eny], 1),
                coord=np.ascontiguousarray(coord), volume=vol, nbPtr=nbPtr, nbPoint=nbPoint, markers=markers, globalIndex=glob,
                elems=elems, etype=etype)


class _Su2Mesh(C.Structure):
    _fields_ = [("nPoint", C.c_int64), ("nEdge", C.c_int64), ("nElem", C.c_int64), ("coord", C.POINTER(C.c_double)),
                ("globalIndex", C.POINTER(C.c_int64)), ("etype", C.POINTER(C.c_int64)), ("elems", C.POINTER(C.c_int64)),
                ("edgeNodes", C.POINTER(C.c_int64)), ("edgeNormal", C.POINTER(C.c_double)), ("volume", C.POINTER(C.c_double)),
                ("nbPtr", C.POINTER(C.c_int64)), ("nbPoint", C.POINTER(C.c_int64)), ("nMarker", C.c_int),
                ("markerTag", C.POINTER(C.c_char_p)), ("markerNVertex", C.POINTER(C.c_int64)),
                ("markerNode", C.POINTER(C.POINTER(C.c_int64))), ("markerNormalNeighbor", C.POINTER(C.POINTER(C.c_int64))),
                ("markerNormal", C.POINTER(C.POINTER(C.c_double)))]


def _su2_mesh_dict(L, out):
    """rx_su2_mesh -> the mesh dict of su2_dual_mesh (arrays copied, the struct released)."""
    try:
        nP, nE, nEl = out.nPoint, out.nEdge, out.nElem
        arr = lambda ptr, n, shape=None: np.ctypeslib.as_array(ptr, shape=(n,)).copy().reshape(shape or (n,))
        markers = []
        for m in range(out.nMarker):
            nv = out.markerNVertex[m]
            markers.append(dict(tag=out.markerTag[m].decode(), kind=0, node=arr(out.markerNode[m], nv) if nv else np.zeros(0, np.int64),
                                normalNeighbor=arr(out.markerNormalNeighbor[m], nv) if nv else np.zeros(0, np.int64),
                                normal=arr(out.markerNormal[m], 2 * nv, (nv, 2)) if nv else np.zeros((0, 2))))
        return dict(nPoint=nP, nPointDomain=nP, nEdge=nE, edgeNodes=arr(out.edgeNodes, 2 * nE, (nE, 2)), edgeNormal=arr(out.edgeNormal, 2 * nE, (nE, 2)),
                    coord=arr(out.coord, 2 * nP, (nP, 2)), volume=arr(out.volume, nP), nbPtr=arr(out.nbPtr, nP + 1), nbPoint=arr(out.nbPoint, 2 * nE),
                    markers=markers, globalIndex=arr(out.globalIndex, nP), elems=arr(out.elems, 4 * nEl, (nEl, 4)), etype=arr(out.etype, nEl))
    finally:
        L.rx_su2_mesh_free(C.byref(out))


def _su2_lib():
    from .rx_solver import load_library
    L = load_library()
    L.rx_su2_mesh_error.restype = C.c_char_p
    return L


def load_su2_mesh(path, rcm=True):
    """An ASCII .su2 file -> the flat mesh dict of su2_dual_mesh, by the library's host code (rx_su2_mesh_read, csrc/host/rx_su2_mesh.cpp):
    parse, point connectivity, RCM numbering, orientation, edges, control volumes, vertices and normal neighbours in one native pass —
    the same arrays, bit for bit, as su2_dual_mesh(read_su2(path)), which stays as the readable specification."""
    L = _su2_lib()
    out = _Su2Mesh()
    rc = L.rx_su2_mesh_read(str(path).encode(), C.c_int(1 if rcm else 0), C.byref(out))
    if rc != 0:
        raise ValueError(L.rx_su2_mesh_error().decode())
    return _su2_mesh_dict(L, out)


def su2_dual_mesh_native(su2, rcm=True):
    """su2_dual_mesh on a read_su2()-style dict through the library (rx_su2_mesh_build)."""
    L = _su2_lib()
    coord = np.ascontiguousarray(su2["coord"], dtype=np.float64); etype = np.ascontiguousarray(su2["etype"], dtype=np.int64)
    elems = np.ascontiguousarray(su2["elems"], dtype=np.int64)
    segs = [np.ascontiguousarray(s, dtype=np.int64).reshape(-1, 2) for _, s in su2["markers"]]
    ptr = np.zeros(len(segs) + 1, dtype=np.int64)
    if segs:
        np.cumsum([s.shape[0] for s in segs], out=ptr[1:])
    seg = np.ascontiguousarray(np.concatenate(segs) if segs else np.zeros((0, 2), np.int64))
    tags = (C.c_char_p * max(len(segs), 1))(*[t.encode() for t, _ in su2["markers"]])
    pi = lambda a: a.ctypes.data_as(C.POINTER(C.c_int64))
    out = _Su2Mesh()
    rc = L.rx_su2_mesh_build(C.c_int64(coord.shape[0]), coord.ctypes.data_as(C.POINTER(C.c_double)), C.c_int64(etype.shape[0]), pi(etype), pi(elems),
                             C.c_int(len(segs)), tags, pi(ptr), pi(seg), C.c_int(1 if rcm else 0), C.byref(out))
    if rc != 0:
        raise ValueError(L.rx_su2_mesh_error().decode())
    return _su2_mesh_dict(L, out)


def wall_distance(mesh, wall_tags, device=0):
    """ComputeWall_Distance on the GPU (rx_wall_distance): distance of every point to the nearest node of the named markers."""
    from .rx_solver import load_library, RxError
    L = load_library()
    nodes = [m["node"] for m in mesh["markers"] if m["tag"] in set(wall_tags)]
    xy = np.ascontiguousarray(mesh["coord"], dtype=np.float64)
    dist = np.zeros(xy.shape[0])
    if not nodes:
        return dist
    w = np.ascontiguousarray(xy[np.concatenate(nodes)])
    pd = lambda a: a.ctypes.data_as(C.POINTER(C.c_double))
    rc = L.rx_wall_distance(C.c_int64(xy.shape[0]), pd(xy), C.c_int64(w.shape[0]), pd(w), pd(dist), int(device))
    if rc != 0:
        raise RxError("rx_wall_distance failed (%d): %s" % (rc, L.rx_last_error(None).decode()))
    return dist


def write_su2(path, su2):
    """Inverse of read_su2 (the same blocks Read_SU2_Format parses); coordinates are written with 17 significant digits."""
    with open(path, "w") as f:
        f.write("NDIME= 2\nNELEM= %d\n" % su2["elems"].shape[0])
        for e, (t, row) in enumerate(zip(su2["etype"], su2["elems"])):
            f.write("%d %s %d\n" % (t, " ".join(str(int(v)) for v in row[:_NNODE[int(t)]]), e))
        f.write("NPOIN= %d\n" % su2["coord"].shape[0])
        for p, (x, y) in enumerate(su2["coord"]):
            f.write("%.17g %.17g %d\n" % (x, y, p))
        f.write("NMARK= %d\n" % len(su2["markers"]))
        for tag, seg in su2["markers"]:
            f.write("MARKER_TAG= %s\nMARKER_ELEMS= %d\n" % (tag, len(seg)))
            for a, b in seg:
                f.write("%d %d %d\n" % (LINE, a, b))
