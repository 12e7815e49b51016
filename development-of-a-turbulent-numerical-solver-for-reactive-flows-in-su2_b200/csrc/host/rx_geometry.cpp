// rx_geometry.cpp — host-side geometry preprocessing of triangulated 2-D meshes into the flat arrays rx_mesh carries
// (include/rx_cuda.h, "geometry").
//
//   rx_dual_mesh_tri   the median-dual mesh of CPhysicalGeometry::SetEdges / SetControlVolume (Common/src/geometry_structure.cpp:
//                      edges :8680-8760, control volumes :10457-10560): one edge per pair of connected points, oriented from the lower
//                      to the higher index, its normal the sum over the adjacent elements of the rotated vectors (edge midpoint ->
//                      element centre), the dual volume of a point the sum of the triangles (point, edge midpoint, element centre),
//                      and the point-to-point CSR sorted by neighbour index.
//
// The large synthetic meshes (16 M points per GPU) are built per rank; this is the builder they go through.  It is one linear pass per
// array — per-point bucket lists instead of a global sort of the 3 nTri element faces — and every floating-point sum is taken in the
// order of the element faces (all faces 0-1, then all 1-2, then all 2-0, as mesh.py's reference builder lists them), so the arrays
// equal that builder's bit for bit (tests/test_mesh.py).  Plain C ABI, no device code; outputs are malloc'ed and handed back in one
// struct, released with rx_dual_mesh_free.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../../include/rx_cuda.h"

extern "C" void rx_dual_mesh_free(rx_dual_mesh* m) {
  if (!m) return;
  free(m->edgeNodes); free(m->edgeNormal); free(m->volume); free(m->nbPtr); free(m->nbPoint);
  memset(m, 0, sizeof *m);
}

extern "C" int rx_dual_mesh_tri(int64_t nPoint, const double* coord, int64_t nTri, const int64_t* tri, rx_dual_mesh* out) {
  if (!out) return RX_ERR_ARG;
  memset(out, 0, sizeof *out);
  if (nPoint <= 0 || nTri < 0 || !coord || (nTri > 0 && !tri)) return RX_ERR_ARG;
  for (int64_t k = 0; k < 3 * nTri; ++k)
    if (tri[k] < 0 || tri[k] >= nPoint) return RX_ERR_ARG;
  const int64_t nF = 3 * nTri;
  // face f of block b = f / nTri, element t = f % nTri: (tri[t][b], tri[t][(b + 1) % 3])
  auto faceA = [&](int64_t f) { const int64_t b = f / nTri, t = f - b * nTri; return tri[3 * t + b]; };
  auto faceB = [&](int64_t f) { const int64_t b = f / nTri, t = f - b * nTri; return tri[3 * t + (b + 1) % 3]; };
  // 1. per lower point: the higher end points of its faces, then sorted and made unique -> edges in (lower, higher) order
  std::vector<int64_t> cnt((size_t)nPoint + 1, 0);
  for (int64_t f = 0; f < nF; ++f) { const int64_t a = faceA(f), b = faceB(f); cnt[(size_t)std::min(a, b) + 1]++; }
  for (int64_t p = 0; p < nPoint; ++p) cnt[(size_t)p + 1] += cnt[(size_t)p];
  std::vector<int64_t> hi((size_t)nF), fill(cnt.begin(), cnt.end() - 1);
  for (int64_t f = 0; f < nF; ++f) { const int64_t a = faceA(f), b = faceB(f); hi[(size_t)fill[(size_t)std::min(a, b)]++] = std::max(a, b); }
  std::vector<int64_t> eStart((size_t)nPoint + 1, 0);
  int64_t nE = 0;
  for (int64_t p = 0; p < nPoint; ++p) {
    int64_t* b0 = hi.data() + cnt[(size_t)p]; int64_t* b1 = hi.data() + cnt[(size_t)p + 1];
    std::sort(b0, b1);
    int64_t* u = std::unique(b0, b1);
    eStart[(size_t)p] = nE;
    nE += u - b0;
    // keep the unique values at the front of the bucket; the tail is unused from here on
    cnt[(size_t)p] = cnt[(size_t)p];   // bucket start stays
    fill[(size_t)p] = u - b0;          // number of edges whose lower point is p
  }
  eStart[(size_t)nPoint] = nE;
  out->nEdge = nE;
  out->edgeNodes = (int64_t*)malloc(sizeof(int64_t) * 2 * (size_t)std::max<int64_t>(nE, 1));
  out->edgeNormal = (double*)calloc(2 * (size_t)std::max<int64_t>(nE, 1), sizeof(double));
  out->volume = (double*)malloc(sizeof(double) * (size_t)nPoint);
  out->nbPtr = (int64_t*)malloc(sizeof(int64_t) * ((size_t)nPoint + 1));
  out->nbPoint = (int64_t*)malloc(sizeof(int64_t) * 2 * (size_t)std::max<int64_t>(nE, 1));
  if (!out->edgeNodes || !out->edgeNormal || !out->volume || !out->nbPtr || !out->nbPoint) { rx_dual_mesh_free(out); return RX_ERR_CUDA; }
  for (int64_t p = 0; p < nPoint; ++p)
    for (int64_t k = 0; k < fill[(size_t)p]; ++k) {
      const int64_t e = eStart[(size_t)p] + k;
      out->edgeNodes[2 * e] = p; out->edgeNodes[2 * e + 1] = hi[(size_t)(cnt[(size_t)p] + k)];
    }
  // 2. normals and volumes, in face order (numpy.bincount adds its weights in array order, starting from 0.0)
  std::vector<double> volA((size_t)nPoint, 0.0), volB((size_t)nPoint, 0.0);
  for (int64_t f = 0; f < nF; ++f) {
    const int64_t blk = f / nTri, t = f - blk * nTri;
    const int64_t a = tri[3 * t + blk], b = tri[3 * t + (blk + 1) % 3];
    const int64_t t0 = tri[3 * t], t1 = tri[3 * t + 1], t2 = tri[3 * t + 2];
    const double ex = ((coord[2 * t0] + coord[2 * t1]) + coord[2 * t2]) / 3.0;
    const double ey = ((coord[2 * t0 + 1] + coord[2 * t1 + 1]) + coord[2 * t2 + 1]) / 3.0;
    const double mx = 0.5 * (coord[2 * a] + coord[2 * b]), my = 0.5 * (coord[2 * a + 1] + coord[2 * b + 1]);
    const bool flip = a > b;
    const int64_t lo = flip ? b : a, hv = flip ? a : b;
    const int64_t* b0 = hi.data() + cnt[(size_t)lo];
    const int64_t e = eStart[(size_t)lo] + (std::lower_bound(b0, b0 + fill[(size_t)lo], hv) - b0);
    out->edgeNormal[2 * e] += flip ? my - ey : ey - my;
    out->edgeNormal[2 * e + 1] += flip ? -(mx - ex) : -(ex - mx);
    {
      const double xp = coord[2 * a], yp = coord[2 * a + 1];
      volA[(size_t)a] += 0.5 * std::fabs((ex - xp) * (my - yp) - (ey - yp) * (mx - xp));
    }
    {
      const double xp = coord[2 * b], yp = coord[2 * b + 1];
      volB[(size_t)b] += 0.5 * std::fabs((ex - xp) * (my - yp) - (ey - yp) * (mx - xp));
    }
  }
  for (int64_t p = 0; p < nPoint; ++p) out->volume[p] = volA[(size_t)p] + volB[(size_t)p];
  // 3. neighbour CSR: the lower neighbours of a point arrive in ascending order when the edges are walked in order, its higher
  //    neighbours are its own bucket — together the list is sorted by neighbour index
  std::vector<int64_t> deg((size_t)nPoint + 1, 0);
  for (int64_t e = 0; e < nE; ++e) { deg[(size_t)out->edgeNodes[2 * e] + 1]++; deg[(size_t)out->edgeNodes[2 * e + 1] + 1]++; }
  out->nbPtr[0] = 0;
  for (int64_t p = 0; p < nPoint; ++p) out->nbPtr[p + 1] = out->nbPtr[p] + deg[(size_t)p + 1];
  std::vector<int64_t> pos((size_t)nPoint);
  for (int64_t p = 0; p < nPoint; ++p) pos[(size_t)p] = out->nbPtr[p];
  for (int64_t e = 0; e < nE; ++e) { const int64_t j = out->edgeNodes[2 * e + 1]; out->nbPoint[pos[(size_t)j]++] = out->edgeNodes[2 * e]; }   // lower neighbours first
  for (int64_t e = 0; e < nE; ++e) { const int64_t i = out->edgeNodes[2 * e]; out->nbPoint[pos[(size_t)i]++] = out->edgeNodes[2 * e + 1]; }
  return RX_OK;
}
