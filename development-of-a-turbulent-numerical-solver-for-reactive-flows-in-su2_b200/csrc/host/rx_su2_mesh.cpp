// rx_su2_mesh.cpp — an ASCII .su2 mesh file (2-D, triangles and quadrilaterals) -> the flat dual mesh rx_mesh carries, in the REFERENCE's
// point, edge, neighbour and vertex numbering (geometry preprocessing without the reference's object-per-point
// structures; include/rx_cuda.h, "geometry").
//
// Follows CDriver::Geometrical_Preprocessing step by step (file:line in Common/src/geometry_structure.cpp unless noted):
//   parse                  CPhysicalGeometry::Read_SU2_Format :4819 — NDIME / NELEM / NPOIN / NMARK / MARKER_TAG / MARKER_ELEMS blocks
//   point connectivity     SetPoint_Connectivity :9145 — neighbours of a point in the order its elements (ascending) list them
//                          (CTriangle / CQuadrilateral::Neighbor_Nodes, primal_grid_structure.cpp:194,260), first appearance kept
//   numbering              SetRCM_Ordering :9200 (rx_rcm_order, rx_partition.cpp)
//   orientation            Check_IntElem_Orientation :8542, Check_BoundElem_Orientation :8825
//   edges                  CGeometry::SetEdges :8680 — points ascending, neighbours in order, an edge is numbered at its lower point
//   control volumes        SetCoord_CG :9518, SetControlVolume :10457, SetVertex :9451, SetBoundControlVolume :9595, FindNormal_Neighbor
// Every floating-point sum is taken in the reference's element order, which is also the order of the package's numpy specification
// (geometry.su2_dual_mesh): the arrays equal both bit for bit (tests/test_geometry.py).  Plain C ABI, no device code; the outputs are
// malloc'ed and handed back in one struct, released with rx_su2_mesh_free.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../../include/rx_cuda.h"

namespace {
const int TRIANGLE = 5, QUADRILATERAL = 9;
const double EPS = 1.0e-16;
const int NB3[3][2] = {{1, 2}, {2, 0}, {0, 1}};
const int NB4[4][2] = {{1, 3}, {2, 0}, {3, 1}, {0, 2}};

struct Raw {
  std::vector<double> coord;            // [nP][2]
  std::vector<int64_t> etype, elems;    // [nEl], [nEl][4] (-1 padded)
  std::vector<std::string> tags;
  std::vector<std::vector<int64_t>> seg;   // per marker [n][2]
};

inline int nnode(int64_t t) { return t == TRIANGLE ? 3 : 4; }

template <class T>
T* dup(const std::vector<T>& v) {
  T* p = (T*)malloc(sizeof(T) * std::max<size_t>(v.size(), 1));
  if (p && !v.empty()) memcpy(p, v.data(), sizeof(T) * v.size());
  return p;
}

// CSR of the point neighbours in the reference's order
void point_connectivity(int64_t nP, const std::vector<int64_t>& etype, const std::vector<int64_t>& elems, std::vector<int64_t>& nbPtr,
                        std::vector<int64_t>& nbPoint) {
  const int64_t nEl = (int64_t)etype.size();
  std::vector<int64_t> cnt((size_t)nP + 1, 0);
  for (int64_t e = 0; e < nEl; ++e)
    for (int i = 0; i < nnode(etype[e]); ++i) cnt[(size_t)elems[4 * e + i] + 1]++;
  for (int64_t p = 0; p < nP; ++p) cnt[(size_t)p + 1] += cnt[(size_t)p];
  std::vector<int64_t> inc((size_t)cnt[(size_t)nP]), fill(cnt.begin(), cnt.end() - 1);   // (element, local node) of every point, elements ascending
  for (int64_t e = 0; e < nEl; ++e)
    for (int i = 0; i < nnode(etype[e]); ++i) inc[(size_t)fill[(size_t)elems[4 * e + i]]++] = 4 * e + i;
  nbPtr.assign((size_t)nP + 1, 0);
  nbPoint.clear();
  for (int64_t p = 0; p < nP; ++p) {
    const size_t start = nbPoint.size();
    for (int64_t k = cnt[(size_t)p]; k < cnt[(size_t)p + 1]; ++k) {
      const int64_t e = inc[(size_t)k] / 4; const int i = (int)(inc[(size_t)k] % 4);
      for (int s = 0; s < 2; ++s) {
        const int j = etype[e] == TRIANGLE ? NB3[i][s] : NB4[i][s];
        const int64_t q = elems[4 * e + j];
        bool seen = false;
        for (size_t m = start; m < nbPoint.size(); ++m) if (nbPoint[m] == q) { seen = true; break; }
        if (!seen) nbPoint.push_back(q);
      }
    }
    nbPtr[(size_t)p + 1] = (int64_t)nbPoint.size();
  }
}

inline double cross_half(const std::vector<double>& c, int64_t p1, int64_t p2, int64_t p3) {
  const double ax = 0.5 * (c[2 * p2] - c[2 * p1]), ay = 0.5 * (c[2 * p2 + 1] - c[2 * p1 + 1]);
  const double bx = 0.5 * (c[2 * p3] - c[2 * p1]), by = 0.5 * (c[2 * p3 + 1] - c[2 * p1 + 1]);
  return ax * by - bx * ay;
}
}  // namespace

namespace {
thread_local std::string g_err;

int build(Raw& r, int rcm, rx_su2_mesh* out) {
  const int64_t nP = (int64_t)r.coord.size() / 2, nEl = (int64_t)r.etype.size();
  std::vector<int64_t> glob((size_t)nP);
  for (int64_t p = 0; p < nP; ++p) glob[(size_t)p] = p;
  std::vector<int64_t> nbPtr, nbPoint;
  if (rcm) {   // SetRCM_Ordering: new numbering of the points, everything that refers to points follows
    point_connectivity(nP, r.etype, r.elems, nbPtr, nbPoint);
    std::vector<int64_t> n2o((size_t)nP), o2n((size_t)nP);
    if (rx_rcm_order(nP, nP, nbPtr.data(), nbPoint.data(), n2o.data()) != RX_OK) { g_err = "rx_su2_mesh: rx_rcm_order failed"; return RX_ERR_ARG; }
    for (int64_t p = 0; p < nP; ++p) o2n[(size_t)n2o[(size_t)p]] = p;
    std::vector<double> c((size_t)2 * nP);
    for (int64_t p = 0; p < nP; ++p) { c[2 * p] = r.coord[2 * n2o[(size_t)p]]; c[2 * p + 1] = r.coord[2 * n2o[(size_t)p] + 1]; }
    r.coord.swap(c);
    glob = n2o;
    for (auto& v : r.elems) if (v >= 0) v = o2n[(size_t)v];
    for (auto& s : r.seg) for (auto& v : s) v = o2n[(size_t)v];
  }
  point_connectivity(nP, r.etype, r.elems, nbPtr, nbPoint);
  const std::vector<double>& c = r.coord;
  // ---- orientation: counter-clockwise elements, boundary lines with the domain on their left
  for (int64_t e = 0; e < nEl; ++e) {
    int64_t* n = &r.elems[4 * e];
    if (r.etype[e] == TRIANGLE) {
      if (cross_half(c, n[0], n[1], n[2]) < 0.0) std::swap(n[0], n[2]);
    } else {
      const double t1 = cross_half(c, n[0], n[1], n[2]), t2 = cross_half(c, n[1], n[2], n[3]);
      const double t3 = cross_half(c, n[2], n[3], n[0]), t4 = cross_half(c, n[3], n[0], n[2]);
      if (t1 < 0.0 && t2 < 0.0 && t3 < 0.0 && t4 < 0.0) std::swap(n[1], n[3]);
    }
  }
  // elements of every point (for the domain element of a boundary line)
  std::vector<int64_t> pePtr((size_t)nP + 1, 0), pe;
  for (int64_t e = 0; e < nEl; ++e) for (int i = 0; i < nnode(r.etype[e]); ++i) pePtr[(size_t)r.elems[4 * e + i] + 1]++;
  for (int64_t p = 0; p < nP; ++p) pePtr[(size_t)p + 1] += pePtr[(size_t)p];
  pe.resize((size_t)pePtr[(size_t)nP]);
  { std::vector<int64_t> f(pePtr.begin(), pePtr.end() - 1);
    for (int64_t e = 0; e < nEl; ++e) for (int i = 0; i < nnode(r.etype[e]); ++i) pe[(size_t)f[(size_t)r.elems[4 * e + i]]++] = e; }
  for (size_t m = 0; m < r.seg.size(); ++m) {
    std::vector<int64_t>& s = r.seg[m];
    for (size_t k = 0; k + 1 < s.size(); k += 2) {
      const int64_t a = s[k], b = s[k + 1];
      // the element that has (a, b) as a face; among several the one whose face has the lowest local index, then the lowest element index
      int64_t de = -1; int bestI = 99;
      for (int64_t q = pePtr[(size_t)a]; q < pePtr[(size_t)a + 1]; ++q) {
        const int64_t e = pe[(size_t)q]; const int nn = nnode(r.etype[e]);
        for (int i = 0; i < nn; ++i) {
          const int64_t u = r.elems[4 * e + i], v = r.elems[4 * e + (i + 1 < nn ? i + 1 : 0)];
          if (((u == a && v == b) || (u == b && v == a)) && (i < bestI || (i == bestI && e < de))) { bestI = i; de = e; }
        }
      }
      if (de < 0) { g_err = "rx_su2_mesh: marker " + r.tags[m] + ": a boundary element is not a face of any element"; return RX_ERR_ARG; }
      int64_t third = -1;
      for (int i = 0; i < nnode(r.etype[de]); ++i) { const int64_t v = r.elems[4 * de + i]; if (v != a && v != b) { third = v; break; } }
      if (cross_half(c, a, b, third) < 0.0) std::swap(s[k], s[k + 1]);
    }
  }
  // ---- SetEdges: an edge is numbered when its lower point meets it
  std::vector<int64_t> eStart((size_t)nP + 1, 0), ei, ej;
  for (int64_t p = 0; p < nP; ++p) {
    eStart[(size_t)p] = (int64_t)ei.size();
    for (int64_t k = nbPtr[(size_t)p]; k < nbPtr[(size_t)p + 1]; ++k) if (nbPoint[(size_t)k] > p) { ei.push_back(p); ej.push_back(nbPoint[(size_t)k]); }
  }
  const int64_t nE = (int64_t)ei.size();
  eStart[(size_t)nP] = nE;
  auto edge_of = [&](int64_t lo, int64_t hi) -> int64_t {
    for (int64_t e = eStart[(size_t)lo]; e < eStart[(size_t)lo + 1]; ++e) if (ej[(size_t)e] == hi) return e;
    return -1;
  };
  // ---- control volumes: faces in element-major order
  std::vector<double> enx((size_t)nE, 0.0), eny((size_t)nE, 0.0), vol((size_t)nP, 0.0);
  for (int64_t e = 0; e < nEl; ++e) {
    const int nn = nnode(r.etype[e]);
    const int64_t* n = &r.elems[4 * e];
    const double fn = (double)nn;
    double ecx = ((0.0 + c[2 * n[0]] / fn) + c[2 * n[1]] / fn) + c[2 * n[2]] / fn;       // CPrimalGrid::SetCoord_CG: CG += coord / nNodes, node by node
    double ecy = ((0.0 + c[2 * n[0] + 1] / fn) + c[2 * n[1] + 1] / fn) + c[2 * n[2] + 1] / fn;
    if (nn == 4) { ecx = ecx + c[2 * n[3]] / fn; ecy = ecy + c[2 * n[3] + 1] / fn; }
    for (int i = 0; i < nn; ++i) {
      const int64_t fa = n[i], fb = n[i + 1 < nn ? i + 1 : 0];
      const bool flip = fa > fb;
      const int64_t fe = edge_of(flip ? fb : fa, flip ? fa : fb);
      if (fe < 0) { g_err = "rx_su2_mesh: an element face is not an edge of the point graph"; return RX_ERR_ARG; }
      const double mx = 0.5 * (c[2 * fa] + c[2 * fb]), my = 0.5 * (c[2 * fa + 1] + c[2 * fb + 1]);   // edge CG
      enx[(size_t)fe] += flip ? my - ecy : ecy - my;
      eny[(size_t)fe] += flip ? -(mx - ecx) : -(ecx - mx);
      // CEdge::GetVolume(point, edge CG, element CG) as SetControlVolume calls it: the vectors are taken from the element CG (dual_grid_structure.cpp:455-476)
      vol[(size_t)fa] += 0.5 * std::fabs((mx - ecx) * (c[2 * fa + 1] - ecy) - (my - ecy) * (c[2 * fa] - ecx));
      vol[(size_t)fb] += 0.5 * std::fabs((mx - ecx) * (c[2 * fb + 1] - ecy) - (my - ecy) * (c[2 * fb] - ecx));
    }
  }
  for (int64_t e = 0; e < nE; ++e)
    if (std::sqrt(enx[(size_t)e] * enx[(size_t)e] + eny[(size_t)e] * eny[(size_t)e]) == 0.0) { enx[(size_t)e] = EPS * EPS; eny[(size_t)e] = EPS * EPS; }
  // ---- outputs
  memset(out, 0, sizeof *out);
  out->nPoint = nP; out->nEdge = nE; out->nElem = nEl;
  out->coord = dup(r.coord); out->globalIndex = dup(glob); out->etype = dup(r.etype); out->elems = dup(r.elems);
  std::vector<int64_t> en((size_t)2 * nE); std::vector<double> enr((size_t)2 * nE);
  for (int64_t e = 0; e < nE; ++e) { en[2 * e] = ei[(size_t)e]; en[2 * e + 1] = ej[(size_t)e]; enr[2 * e] = enx[(size_t)e]; enr[2 * e + 1] = eny[(size_t)e]; }
  out->edgeNodes = dup(en); out->edgeNormal = dup(enr); out->volume = dup(vol); out->nbPtr = dup(nbPtr); out->nbPoint = dup(nbPoint);
  const int nM = (int)r.seg.size();
  out->nMarker = nM;
  out->markerTag = (char**)calloc((size_t)std::max(nM, 1), sizeof(char*));
  out->markerNVertex = (int64_t*)calloc((size_t)std::max(nM, 1), sizeof(int64_t));
  out->markerNode = (int64_t**)calloc((size_t)std::max(nM, 1), sizeof(int64_t*));
  out->markerNormalNeighbor = (int64_t**)calloc((size_t)std::max(nM, 1), sizeof(int64_t*));
  out->markerNormal = (double**)calloc((size_t)std::max(nM, 1), sizeof(double*));
  if (!out->coord || !out->globalIndex || !out->etype || !out->elems || !out->edgeNodes || !out->edgeNormal || !out->volume || !out->nbPtr || !out->nbPoint ||
      !out->markerTag || !out->markerNVertex || !out->markerNode || !out->markerNormalNeighbor || !out->markerNormal) { g_err = "rx_su2_mesh: out of memory"; return RX_ERR_CUDA; }
  std::vector<int64_t> vid((size_t)nP, -1);
  for (int m = 0; m < nM; ++m) {
    const std::vector<int64_t>& s = r.seg[(size_t)m];
    std::vector<int64_t> node;   // SetVertex: first appearance in the boundary element loop
    for (int64_t v : s) if (vid[(size_t)v] < 0) { vid[(size_t)v] = (int64_t)node.size(); node.push_back(v); }
    const size_t nV = node.size();
    std::vector<double> vn(2 * nV, 0.0);
    // SetBoundControlVolume: half of every line, rotated, to its two vertices — all first nodes, then all second nodes (the order of the specification)
    for (size_t k = 0; k + 1 < s.size(); k += 2) {
      const int64_t n0 = s[k], n1 = s[k + 1];
      const double cx = 0.5 * (c[2 * n0] + c[2 * n1]), cy = 0.5 * (c[2 * n0 + 1] + c[2 * n1 + 1]);
      vn[2 * (size_t)vid[(size_t)n0]] += c[2 * n0 + 1] - cy; vn[2 * (size_t)vid[(size_t)n0] + 1] += -(c[2 * n0] - cx);
    }
    for (size_t k = 0; k + 1 < s.size(); k += 2) {
      const int64_t n0 = s[k], n1 = s[k + 1];
      const double cx = 0.5 * (c[2 * n0] + c[2 * n1]), cy = 0.5 * (c[2 * n0 + 1] + c[2 * n1 + 1]);
      vn[2 * (size_t)vid[(size_t)n1]] += cy - c[2 * n1 + 1]; vn[2 * (size_t)vid[(size_t)n1] + 1] += -(cx - c[2 * n1]);
    }
    std::vector<int64_t> nnb(nV);
    for (size_t k = 0; k < nV; ++k) {   // FindNormal_Neighbor: largest cosine, the last one on ties
      const int64_t p = node[k];
      double best = -1.0; int64_t arg = nbPoint[(size_t)nbPtr[(size_t)p]];
      for (int64_t q = nbPtr[(size_t)p]; q < nbPtr[(size_t)p + 1]; ++q) {
        const int64_t nb = nbPoint[(size_t)q];
        const double dx = c[2 * nb] - c[2 * p], dy = c[2 * nb + 1] - c[2 * p + 1];
        const double cs = (dx * vn[2 * k] + dy * vn[2 * k + 1]) / (std::sqrt(dx * dx + dy * dy) * std::sqrt(vn[2 * k] * vn[2 * k] + vn[2 * k + 1] * vn[2 * k + 1]));
        if (cs >= best) { best = cs; arg = nb; }
      }
      nnb[k] = arg;
    }
    for (int64_t v : node) vid[(size_t)v] = -1;
    out->markerTag[m] = strdup(r.tags[(size_t)m].c_str());
    out->markerNVertex[m] = (int64_t)nV;
    out->markerNode[m] = dup(node); out->markerNormalNeighbor[m] = dup(nnb); out->markerNormal[m] = dup(vn);
    if (!out->markerTag[m] || !out->markerNode[m] || !out->markerNormalNeighbor[m] || !out->markerNormal[m]) { g_err = "rx_su2_mesh: out of memory"; return RX_ERR_CUDA; }
  }
  return RX_OK;
}
}  // namespace

namespace {
// one pass over the file in memory; a keyword line is "KEY= value", the block of a counted keyword follows on the next lines
struct Lines {
  std::vector<char> buf; size_t pos = 0;
  bool next(char*& b, char*& e) {
    const size_t end = buf.size() - 1;   // buf[end] is the terminating zero appended after the file's bytes
    if (pos >= end) return false;
    b = &buf[pos];
    char* q = (char*)memchr(b, '\n', end - pos);
    e = q ? q : &buf[0] + end;
    pos = (size_t)(e - &buf[0]) + 1;
    *e = 0;
    while (b < e && (*b == ' ' || *b == '\t' || *b == '\r')) ++b;
    while (e > b && (e[-1] == ' ' || e[-1] == '\t' || e[-1] == '\r')) *--e = 0;
    return true;
  }
};
inline bool starts(const char* s, const char* key) { return strncmp(s, key, strlen(key)) == 0; }
inline const char* value_of(const char* s) { const char* q = strchr(s, '='); return q ? q + 1 : s + strlen(s); }

int parse(const char* path, Raw& r) {
  FILE* f = fopen(path, "rb");
  if (!f) { g_err = std::string("rx_su2_mesh_read: cannot open ") + path; return RX_ERR_ARG; }
  Lines L;
  fseek(f, 0, SEEK_END); const long sz = ftell(f); fseek(f, 0, SEEK_SET);
  L.buf.resize((size_t)std::max<long>(sz, 0) + 1);
  const size_t got = fread(L.buf.data(), 1, (size_t)std::max<long>(sz, 0), f);
  fclose(f);
  L.buf.resize(got + 1); L.buf[got] = 0;
  char *b, *e;
  auto fail = [&](const std::string& what) { g_err = "rx_su2_mesh_read: " + what; return RX_ERR_ARG; };
  while (L.next(b, e)) {
    if (starts(b, "NDIME")) {
      if (strtol(value_of(b), nullptr, 10) != 2) return fail("2-D meshes only");
    } else if (starts(b, "NELEM")) {
      const int64_t n = strtoll(value_of(b), nullptr, 10);
      if (n < 0) return fail("bad NELEM");
      r.etype.resize((size_t)n); r.elems.assign((size_t)4 * n, -1);
      for (int64_t k = 0; k < n; ++k) {
        if (!L.next(b, e)) return fail("the file ends inside the element block");
        char* q = b;
        const long t = strtol(q, &q, 10);
        if (t != TRIANGLE && t != QUADRILATERAL) return fail("element type " + std::to_string(t) + " is not a triangle or quadrilateral");
        r.etype[(size_t)k] = t;
        for (int i = 0; i < nnode(t); ++i) { char* q2; r.elems[4 * (size_t)k + i] = strtoll(q, &q2, 10); if (q2 == q) return fail("short element line"); q = q2; }
      }
    } else if (starts(b, "NPOIN")) {
      const int64_t n = strtoll(value_of(b), nullptr, 10);
      if (n <= 0) return fail("bad NPOIN");
      r.coord.resize((size_t)2 * n);
      for (int64_t p = 0; p < n; ++p) {
        if (!L.next(b, e)) return fail("the file ends inside the point block");
        char* q = b; char* q2;
        r.coord[2 * (size_t)p] = strtod(q, &q2); if (q2 == q) return fail("short point line"); q = q2;
        r.coord[2 * (size_t)p + 1] = strtod(q, &q2); if (q2 == q) return fail("short point line");
      }
    } else if (starts(b, "MARKER_TAG")) {
      const char* v = value_of(b);
      while (*v == ' ' || *v == '\t') ++v;
      r.tags.push_back(v);
      if (!L.next(b, e) || !starts(b, "MARKER_ELEMS")) return fail("MARKER_ELEMS does not follow MARKER_TAG= " + r.tags.back());
      const int64_t n = strtoll(value_of(b), nullptr, 10);
      if (n < 0) return fail("bad MARKER_ELEMS");
      std::vector<int64_t> s((size_t)2 * n);
      for (int64_t k = 0; k < n; ++k) {
        if (!L.next(b, e)) return fail("the file ends inside a marker block");
        char* q = b; char* q2;
        strtol(q, &q, 10);   // element type (a line)
        s[2 * (size_t)k] = strtoll(q, &q2, 10); if (q2 == q) return fail("short boundary element line"); q = q2;
        s[2 * (size_t)k + 1] = strtoll(q, &q2, 10); if (q2 == q) return fail("short boundary element line");
      }
      r.seg.push_back(std::move(s));
    }
  }
  if (r.coord.empty() || r.etype.empty()) return fail("no points or no elements in the file");
  return RX_OK;
}

int check(const Raw& r) {
  const int64_t nP = (int64_t)r.coord.size() / 2;
  for (size_t e = 0; e < r.etype.size(); ++e)
    for (int i = 0; i < nnode(r.etype[e]); ++i)
      if (r.elems[4 * e + i] < 0 || r.elems[4 * e + i] >= nP) { g_err = "rx_su2_mesh: an element refers to a point outside the mesh"; return RX_ERR_ARG; }
  for (const auto& s : r.seg) for (int64_t v : s) if (v < 0 || v >= nP) { g_err = "rx_su2_mesh: a boundary element refers to a point outside the mesh"; return RX_ERR_ARG; }
  return RX_OK;
}
}  // namespace

extern "C" const char* rx_su2_mesh_error(void) { return g_err.c_str(); }

extern "C" void rx_su2_mesh_free(rx_su2_mesh* m) {
  if (!m) return;
  free(m->coord); free(m->globalIndex); free(m->etype); free(m->elems); free(m->edgeNodes); free(m->edgeNormal); free(m->volume); free(m->nbPtr); free(m->nbPoint);
  for (int k = 0; k < m->nMarker; ++k) {
    if (m->markerTag) free(m->markerTag[k]);
    if (m->markerNode) free(m->markerNode[k]);
    if (m->markerNormalNeighbor) free(m->markerNormalNeighbor[k]);
    if (m->markerNormal) free(m->markerNormal[k]);
  }
  free(m->markerTag); free(m->markerNVertex); free(m->markerNode); free(m->markerNormalNeighbor); free(m->markerNormal);
  memset(m, 0, sizeof *m);
}

extern "C" int rx_su2_mesh_read(const char* path, int rcm, rx_su2_mesh* out) {
  if (!path || !out) { g_err = "rx_su2_mesh_read: NULL argument"; return RX_ERR_ARG; }
  memset(out, 0, sizeof *out);
  Raw r;
  int rc = parse(path, r);
  if (!rc) rc = check(r);
  if (!rc) rc = build(r, rcm, out);
  if (rc) rx_su2_mesh_free(out);
  return rc;
}

extern "C" int rx_su2_mesh_build(int64_t nPoint, const double* coord, int64_t nElem, const int64_t* etype, const int64_t* elems, int nMarker,
                                 const char* const* tags, const int64_t* segPtr, const int64_t* seg, int rcm, rx_su2_mesh* out) {
  if (!out) { g_err = "rx_su2_mesh_build: NULL argument"; return RX_ERR_ARG; }
  memset(out, 0, sizeof *out);
  if (nPoint <= 0 || nElem <= 0 || !coord || !etype || !elems || nMarker < 0 || (nMarker > 0 && (!tags || !segPtr || !seg))) { g_err = "rx_su2_mesh_build: bad argument"; return RX_ERR_ARG; }
  Raw r;
  r.coord.assign(coord, coord + 2 * nPoint);
  r.etype.assign(etype, etype + nElem);
  r.elems.assign(elems, elems + 4 * nElem);
  for (int64_t e = 0; e < nElem; ++e) if (etype[e] != TRIANGLE && etype[e] != QUADRILATERAL) { g_err = "rx_su2_mesh_build: element type is not a triangle or quadrilateral"; return RX_ERR_ARG; }
  for (int m = 0; m < nMarker; ++m) { r.tags.push_back(tags[m] ? tags[m] : ""); r.seg.emplace_back(seg + 2 * segPtr[m], seg + 2 * segPtr[m + 1]); }
  int rc = check(r);
  if (!rc) rc = build(r, rcm, out);
  if (rc) rx_su2_mesh_free(out);
  return rc;
}
