// rx_partition.cpp — host-side graph tools of the multi-GPU path (include/rx_cuda.h, "partitioning"):
//   rx_partition_kway  k-way partition of the point adjacency graph.  It stands where the reference calls
//                      ParMETIS_V3_PartKway on the same graph (Common/src/geometry_structure.cpp:11524, reached from
//                      SetColorGrid_Parallel, driver_structure.cpp:126); METIS itself is not linked.  Method: recursive
//                      inertial bisection (principal axis of the point cloud, split at the size-proportional quantile)
//                      followed by boundary refinement of the edge cut under a balance constraint — first per bisection,
//                      then k-way over the final parts.  Deterministic (ties are broken by point index).
//   rx_rcm_order       reverse Cuthill–McKee numbering of the owned points of one partition, halo points kept last — the
//                      role of CPhysicalGeometry::SetRCM_Ordering (geometry_structure.cpp:9200-9290; lowest-degree start,
//                      neighbours appended by increasing degree, result reversed).  Used to keep the edge kernels' point
//                      gathers local on unstructured partitions.
// Plain C ABI, no device code; the arrays are the caller's.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <numeric>
#include <utility>
#include <vector>
#include <dlfcn.h>
#include <cstdlib>
#include <cstring>
#include <string>

#include "../../../include/rx_cuda.h"

namespace {

struct Graph {
  int64_t n;
  const int64_t* ptr;
  const int64_t* adj;
};

// moves boundary points between side 0 and side 1 of subset `sub` while the cut shrinks (or balance improves at equal cut)
void refine_bisection(const Graph& g, const std::vector<int64_t>& ids, std::vector<int32_t>& label, int32_t sub,
                      std::vector<int8_t>& side, int64_t target0, int64_t slack, int passes) {
  int64_t n0 = 0;
  for (int64_t v : ids) n0 += side[v] == 0;
  const int64_t n = (int64_t)ids.size();
  for (int pass = 0; pass < passes; ++pass) {
    int64_t moved = 0;
    for (int64_t v : ids) {
      int64_t same = 0, other = 0;
      for (int64_t q = g.ptr[v]; q < g.ptr[v + 1]; ++q) {
        const int64_t w = g.adj[q];
        if (w >= g.n || label[w] != sub) continue;
        (side[w] == side[v] ? same : other)++;
      }
      if (other == 0) continue;
      const int64_t gain = other - same;
      const int64_t n0_after = n0 + (side[v] == 0 ? -1 : 1);
      const int64_t dev_now = std::llabs(n0 - target0), dev_after = std::llabs(n0_after - target0);
      const bool balanced = dev_after <= slack && n0_after > 0 && n0_after < n;
      if ((gain > 0 && (balanced || dev_after < dev_now)) || (gain == 0 && dev_after < dev_now && dev_now > slack)) {
        side[v] ^= 1; n0 = n0_after; ++moved;
      }
    }
    if (!moved) break;
  }
}

void bisect(const Graph& g, const double* xy, std::vector<int64_t>& ids, int32_t k0, int32_t k1, std::vector<int32_t>& label,
            std::vector<int8_t>& side, int32_t& nextSub, int passes, double tol, int64_t* owner) {
  if (k1 - k0 == 1) {
    for (int64_t v : ids) owner[v] = k0;
    return;
  }
  const int64_t n = (int64_t)ids.size();
  const int32_t kl = (k1 - k0) / 2;
  const int64_t target0 = (n * kl) / (k1 - k0);
  // principal axis of the subset
  double cx = 0, cy = 0;
  for (int64_t v : ids) { cx += xy[2 * v]; cy += xy[2 * v + 1]; }
  cx /= (double)std::max<int64_t>(n, 1); cy /= (double)std::max<int64_t>(n, 1);
  double sxx = 0, sxy = 0, syy = 0;
  for (int64_t v : ids) {
    const double dx = xy[2 * v] - cx, dy = xy[2 * v + 1] - cy;
    sxx += dx * dx; sxy += dx * dy; syy += dy * dy;
  }
  const double th = 0.5 * std::atan2(2.0 * sxy, sxx - syy);   // direction of the largest eigenvalue
  const double ax = std::cos(th), ay = std::sin(th);
  std::vector<std::pair<double, int64_t>> key(n);
  for (int64_t q = 0; q < n; ++q) key[q] = {ax * (xy[2 * ids[q]] - cx) + ay * (xy[2 * ids[q] + 1] - cy), ids[q]};
  if (target0 > 0 && target0 < n) std::nth_element(key.begin(), key.begin() + target0, key.end());
  const int32_t sub = nextSub++;
  for (int64_t q = 0; q < n; ++q) { label[key[q].second] = sub; side[key[q].second] = q < target0 ? 0 : 1; }
  const int64_t slack = std::max<int64_t>(1, (int64_t)(tol * (double)n / (double)(k1 - k0)));
  refine_bisection(g, ids, label, sub, side, target0, slack, passes);
  std::vector<int64_t> a, b;
  a.reserve(target0 + slack); b.reserve(n - target0 + slack);
  for (int64_t v : ids) (side[v] == 0 ? a : b).push_back(v);   // ids stay in ascending order
  std::vector<int64_t>().swap(ids);
  bisect(g, xy, a, k0, k0 + kl, label, side, nextSub, passes, tol, owner);
  bisect(g, xy, b, k0 + kl, k1, label, side, nextSub, passes, tol, owner);
}

// k-way boundary refinement: a point moves to the neighbouring part that holds most of its neighbours if that lowers the cut
void refine_kway(const Graph& g, int nParts, int64_t* owner, int passes, double tol) {
  std::vector<int64_t> size(nParts, 0);
  for (int64_t v = 0; v < g.n; ++v) size[owner[v]]++;
  const int64_t avg = (g.n + nParts - 1) / nParts;
  const int64_t hi = avg + std::max<int64_t>(1, (int64_t)(tol * (double)avg));
  const int64_t lo = std::max<int64_t>(1, avg - std::max<int64_t>(1, (int64_t)(tol * (double)avg)) - 1);
  std::vector<int64_t> cnt(nParts, 0);
  std::vector<int32_t> touched;
  for (int pass = 0; pass < passes; ++pass) {
    int64_t moved = 0;
    for (int64_t v = 0; v < g.n; ++v) {
      const int64_t me = owner[v];
      touched.clear();
      for (int64_t q = g.ptr[v]; q < g.ptr[v + 1]; ++q) {
        const int64_t w = g.adj[q];
        if (w >= g.n) continue;
        const int64_t p = owner[w];
        if (cnt[p]++ == 0) touched.push_back((int32_t)p);
      }
      int64_t best = me, bestCnt = cnt[me];
      std::sort(touched.begin(), touched.end());
      for (int32_t p : touched)
        if (p != me && cnt[p] > bestCnt && size[p] + 1 <= hi && size[me] - 1 >= lo) { best = p; bestCnt = cnt[p]; }
      for (int32_t p : touched) cnt[p] = 0;
      if (best != me) { owner[v] = best; size[me]--; size[best]++; ++moved; }
    }
    if (!moved) break;
  }
}

}  // namespace

// ---- optional METIS backend: METIS_PartGraphKway of the copy vendored in the reference tree (csrc/host/build_metis.sh ->
// librx_metis.so next to this library, or RX_METIS_LIB), bound with dlopen.  The reference partitions with ParMETIS_V3_PartKway
// (geometry_structure.cpp:11524); the serial multilevel k-way routine is the same algorithm family.  32-bit idx_t, 40 option words.
namespace {
typedef int (*metis_kway_fn)(int32_t*, int32_t*, int32_t*, int32_t*, int32_t*, int32_t*, int32_t*, int32_t*, float*, float*, int32_t*, int32_t*, int32_t*);
typedef int (*metis_opts_fn)(int32_t*);
struct MetisApi { void* h = nullptr; metis_kway_fn kway = nullptr; metis_opts_fn defaults = nullptr; bool tried = false; };
MetisApi& metis_api() {
  static MetisApi m;
  if (m.tried) return m;
  m.tried = true;
  std::string path;
  if (const char* e = std::getenv("RX_METIS_LIB")) path = e;
  else {
    Dl_info info;
    if (dladdr((void*)&metis_api, &info) && info.dli_fname) {
      path = info.dli_fname;
      const size_t k = path.find_last_of('/');
      path = (k == std::string::npos ? std::string(".") : path.substr(0, k)) + "/librx_metis.so";
    }
  }
  if (path.empty()) return m;
  m.h = dlopen(path.c_str(), RTLD_NOW | RTLD_LOCAL);
  if (!m.h) return m;
  m.kway = (metis_kway_fn)dlsym(m.h, "METIS_PartGraphKway");
  m.defaults = (metis_opts_fn)dlsym(m.h, "METIS_SetDefaultOptions");
  if (!m.kway || !m.defaults) { dlclose(m.h); m.h = nullptr; }
  return m;
}
bool want_metis() {
  const char* e = std::getenv("RX_PARTITIONER");
  if (e && !std::strcmp(e, "own")) return false;
  return metis_api().h != nullptr;
}
int64_t cut_of(int64_t nPoint, const int64_t* nbPtr, const int64_t* nbPoint, const int64_t* owner) {
  int64_t cut = 0;
  for (int64_t v = 0; v < nPoint; ++v)
    for (int64_t q = nbPtr[v]; q < nbPtr[v + 1]; ++q)
      if (nbPoint[q] > v && nbPoint[q] < nPoint && owner[nbPoint[q]] != owner[v]) ++cut;
  return cut;
}
}  // namespace

extern "C" int rx_partition_backend(void) { return want_metis() ? 1 : 0; }   // 1: METIS (librx_metis.so found), 0: the library's own partitioner

extern "C" int rx_partition_kway(int64_t nPoint, const int64_t* nbPtr, const int64_t* nbPoint, const double* coord, int nParts,
                                 int refinePasses, int64_t* owner, int64_t* edgeCut) {
  if (nPoint < 0 || nParts < 1 || !nbPtr || !nbPoint || !coord || !owner || (int64_t)nParts > std::max<int64_t>(nPoint, 1)) return RX_ERR_ARG;
  if (want_metis() && nParts > 1 && nPoint < (1LL << 31) && nbPtr[nPoint] < (1LL << 31)) {
    MetisApi& m = metis_api();
    std::vector<int32_t> xadj(nPoint + 1), adj(nbPtr[nPoint]), part(nPoint);
    for (int64_t v = 0; v <= nPoint; ++v) xadj[v] = (int32_t)nbPtr[v];
    for (int64_t q = 0; q < nbPtr[nPoint]; ++q) adj[q] = (int32_t)nbPoint[q];
    int32_t nv = (int32_t)nPoint, ncon = 1, np = nParts, cut = 0, opts[40];
    m.defaults(opts);
    if (m.kway(&nv, &ncon, xadj.data(), adj.data(), nullptr, nullptr, nullptr, &np, nullptr, nullptr, opts, &cut, part.data()) == 1) {   // METIS_OK
      for (int64_t v = 0; v < nPoint; ++v) owner[v] = part[v];
      if (edgeCut) *edgeCut = cut_of(nPoint, nbPtr, nbPoint, owner);
      return RX_OK;
    }
  }
  const Graph g{nPoint, nbPtr, nbPoint};
  const double tol = 0.02;
  {
    std::vector<int64_t> ids(nPoint);
    std::iota(ids.begin(), ids.end(), (int64_t)0);
    std::vector<int32_t> label(nPoint, -1);
    std::vector<int8_t> side(nPoint, 0);
    int32_t nextSub = 0;
    bisect(g, coord, ids, 0, nParts, label, side, nextSub, std::max(refinePasses, 0), tol, owner);
  }
  if (nParts > 2 && refinePasses > 0) refine_kway(g, nParts, owner, refinePasses, tol);
  if (edgeCut) *edgeCut = cut_of(nPoint, nbPtr, nbPoint, owner);
  return RX_OK;
}

extern "C" int rx_rcm_order(int64_t nPoint, int64_t nPointDomain, const int64_t* nbPtr, const int64_t* nbPoint, int64_t* newToOld) {
  if (nPoint < 0 || nPointDomain < 0 || nPointDomain > nPoint || !nbPtr || !nbPoint || !newToOld) return RX_ERR_ARG;
  const int64_t n = nPointDomain;
  auto degree = [&](int64_t v) { return nbPtr[v + 1] - nbPtr[v]; };
  std::vector<char> seen(n, 0);
  std::vector<int64_t> order;
  order.reserve(n);
  std::vector<int64_t> byDegree(n);
  std::iota(byDegree.begin(), byDegree.end(), (int64_t)0);
  std::stable_sort(byDegree.begin(), byDegree.end(), [&](int64_t a, int64_t b) { return degree(a) < degree(b); });
  std::vector<int64_t> nb;
  size_t head = 0;
  for (int64_t s : byDegree) {   // one breadth-first sweep per connected component, lowest-degree start
    if (seen[s]) continue;
    seen[s] = 1; order.push_back(s);
    while (head < order.size()) {
      const int64_t v = order[head++];
      nb.clear();
      for (int64_t q = nbPtr[v]; q < nbPtr[v + 1]; ++q) {
        const int64_t w = nbPoint[q];
        if (w < n && !seen[w]) { seen[w] = 1; nb.push_back(w); }
      }
      std::stable_sort(nb.begin(), nb.end(), [&](int64_t a, int64_t b) { return degree(a) < degree(b); });
      order.insert(order.end(), nb.begin(), nb.end());
    }
  }
  for (int64_t q = 0; q < n; ++q) newToOld[q] = order[n - 1 - q];
  for (int64_t q = n; q < nPoint; ++q) newToOld[q] = q;
  return RX_OK;
}

// ---- one rank's share of a partitioned mesh ---------------------------------------------------------------------------------------------
// rx_partition_mesh: from the whole flat mesh and an owner array, the local mesh of ONE rank in SU2's scheme (the role of the
// partitioned CPhysicalGeometry constructor and SetSendReceive, geometry_structure.cpp:540-1480, :2262-2540): owned points first (ascending global
// index), then one halo layer grouped by owner rank (ascending global index inside a group); every edge with at least one owned end, in the
// whole mesh's edge order; the neighbour lists restricted to local points; boundary vertices on owned points only; optionally the owned points
// renumbered by reverse Cuthill-McKee with the edges re-sorted by (lower, higher) local index and the neighbour lists by neighbour index.
// The send list towards a neighbour rank q holds the owned points that q's halo contains, in q's halo order (ascending global index), so that
// send and receive lists match without any negotiation.  Every rank builds its own part: the cost is one pass over the whole mesh per rank.
extern "C" void rx_part_free(rx_part* p) {
  if (!p) return;
  free(p->glob); free(p->edgeNodes); free(p->edgeNormal); free(p->coord); free(p->volume); free(p->nbPtr); free(p->nbPoint);
  free(p->markerPtr); free(p->markerNode); free(p->markerNormalNeighbor); free(p->markerNormal);
  free(p->nbRank); free(p->sendPtr); free(p->sendPts); free(p->recvPtr); free(p->recvPts);
  memset(p, 0, sizeof *p);
}

namespace {
template <class T>
T* dup_vec(const std::vector<T>& v) {
  T* p = (T*)malloc(sizeof(T) * std::max<size_t>(v.size(), 1));
  if (p && !v.empty()) memcpy(p, v.data(), sizeof(T) * v.size());
  return p;
}
}  // namespace

extern "C" int rx_partition_mesh(int64_t nPoint, int64_t nEdge, const int64_t* edgeNodes, const double* edgeNormal, const double* coord,
                                 const double* volume, const int64_t* nbPtr, const int64_t* nbPoint, int nMarker, const int64_t* markerPtr,
                                 const int64_t* markerNode, const int64_t* markerNormalNeighbor, const double* markerNormal,
                                 const int64_t* owner, int rank, int rcm, rx_part* out) {
  if (!out) return RX_ERR_ARG;
  memset(out, 0, sizeof *out);
  if (nPoint <= 0 || nEdge < 0 || !edgeNodes || !edgeNormal || !coord || !volume || !nbPtr || !nbPoint || !owner || rank < 0 || nMarker < 0 ||
      (nMarker > 0 && (!markerPtr || !markerNode || !markerNormalNeighbor || !markerNormal))) return RX_ERR_ARG;
  // owned points, kept edges, halo
  std::vector<int64_t> glob;
  for (int64_t p = 0; p < nPoint; ++p) if (owner[p] == rank) glob.push_back(p);
  const int64_t nOwn = (int64_t)glob.size();
  if (nOwn == 0) return RX_ERR_ARG;
  std::vector<char> touched((size_t)nPoint, 0);
  std::vector<int64_t> keep;
  for (int64_t e = 0; e < nEdge; ++e) {
    const int64_t i = edgeNodes[2 * e], j = edgeNodes[2 * e + 1];
    if (i < 0 || i >= nPoint || j < 0 || j >= nPoint) return RX_ERR_ARG;
    if (owner[i] == rank || owner[j] == rank) { keep.push_back(e); touched[(size_t)i] = 1; touched[(size_t)j] = 1; }
  }
  std::vector<int64_t> halo;
  for (int64_t p = 0; p < nPoint; ++p) if (touched[(size_t)p] && owner[p] != rank) halo.push_back(p);
  std::stable_sort(halo.begin(), halo.end(), [&](int64_t a, int64_t b) { return owner[a] < owner[b]; });   // ascending global index inside a group
  glob.insert(glob.end(), halo.begin(), halo.end());
  const int64_t nLoc = (int64_t)glob.size(), nE = (int64_t)keep.size();
  std::vector<int64_t> g2l((size_t)nPoint, -1);
  for (int64_t q = 0; q < nLoc; ++q) g2l[(size_t)glob[(size_t)q]] = q;
  // local arrays in the construction numbering
  std::vector<int64_t> en((size_t)2 * nE), lptr((size_t)nLoc + 1, 0), lnb;
  std::vector<double> enr((size_t)2 * nE), xy((size_t)2 * nLoc), vol((size_t)nLoc);
  for (int64_t k = 0; k < nE; ++k) {
    const int64_t e = keep[(size_t)k];
    en[2 * k] = g2l[(size_t)edgeNodes[2 * e]]; en[2 * k + 1] = g2l[(size_t)edgeNodes[2 * e + 1]];
    enr[2 * k] = edgeNormal[2 * e]; enr[2 * k + 1] = edgeNormal[2 * e + 1];
  }
  for (int64_t q = 0; q < nLoc; ++q) {
    const int64_t p = glob[(size_t)q];
    xy[2 * q] = coord[2 * p]; xy[2 * q + 1] = coord[2 * p + 1]; vol[(size_t)q] = volume[p];
    for (int64_t t = nbPtr[p]; t < nbPtr[p + 1]; ++t) { const int64_t l = g2l[(size_t)nbPoint[t]]; if (l >= 0) lnb.push_back(l); }
    lptr[(size_t)q + 1] = (int64_t)lnb.size();
  }
  std::vector<int64_t> mptr((size_t)nMarker + 1, 0), mnode, mnn;
  std::vector<double> mnr;
  for (int m = 0; m < nMarker; ++m) {
    for (int64_t k = markerPtr[m]; k < markerPtr[m + 1]; ++k)
      if (owner[markerNode[k]] == rank) {
        mnode.push_back(g2l[(size_t)markerNode[k]]); mnn.push_back(g2l[(size_t)markerNormalNeighbor[k]]);
        mnr.push_back(markerNormal[2 * k]); mnr.push_back(markerNormal[2 * k + 1]);
      }
    mptr[(size_t)m + 1] = (int64_t)mnode.size();
  }
  if (rcm) {
    std::vector<int64_t> n2o((size_t)nLoc), o2n((size_t)nLoc);
    if (rx_rcm_order(nLoc, nOwn, lptr.data(), lnb.data(), n2o.data()) != RX_OK) return RX_ERR_ARG;
    for (int64_t q = 0; q < nLoc; ++q) o2n[(size_t)n2o[(size_t)q]] = q;
    // edges keep their orientation and are re-sorted by (lower, higher) new index
    std::vector<int64_t> eo((size_t)nE);
    std::iota(eo.begin(), eo.end(), (int64_t)0);
    for (auto& v : en) v = o2n[(size_t)v];
    std::sort(eo.begin(), eo.end(), [&](int64_t a, int64_t b) {
      const int64_t la = std::min(en[2 * a], en[2 * a + 1]), lb = std::min(en[2 * b], en[2 * b + 1]);
      if (la != lb) return la < lb;
      const int64_t ha = std::max(en[2 * a], en[2 * a + 1]), hb = std::max(en[2 * b], en[2 * b + 1]);
      if (ha != hb) return ha < hb;
      return a < b;
    });
    std::vector<int64_t> en2((size_t)2 * nE); std::vector<double> enr2((size_t)2 * nE);
    for (int64_t k = 0; k < nE; ++k) { const int64_t e = eo[(size_t)k]; en2[2 * k] = en[2 * e]; en2[2 * k + 1] = en[2 * e + 1]; enr2[2 * k] = enr[2 * e]; enr2[2 * k + 1] = enr[2 * e + 1]; }
    en.swap(en2); enr.swap(enr2);
    std::vector<double> xy2((size_t)2 * nLoc), vol2((size_t)nLoc);
    std::vector<int64_t> glob2((size_t)nLoc), ptr2((size_t)nLoc + 1, 0), nb2;
    nb2.reserve(lnb.size());
    for (int64_t q = 0; q < nLoc; ++q) {
      const int64_t o = n2o[(size_t)q];
      xy2[2 * q] = xy[2 * o]; xy2[2 * q + 1] = xy[2 * o + 1]; vol2[(size_t)q] = vol[(size_t)o]; glob2[(size_t)q] = glob[(size_t)o];
      const size_t s0 = nb2.size();
      for (int64_t t = lptr[(size_t)o]; t < lptr[(size_t)o + 1]; ++t) nb2.push_back(o2n[(size_t)lnb[(size_t)t]]);
      std::sort(nb2.begin() + (std::ptrdiff_t)s0, nb2.end());   // neighbour lists stay sorted by neighbour index
      ptr2[(size_t)q + 1] = (int64_t)nb2.size();
    }
    xy.swap(xy2); vol.swap(vol2); glob.swap(glob2); lptr.swap(ptr2); lnb.swap(nb2);
    for (auto& v : mnode) v = o2n[(size_t)v];
    for (auto& v : mnn) v = o2n[(size_t)v];
    for (int64_t q = 0; q < nLoc; ++q) g2l[(size_t)glob[(size_t)q]] = q;
  }
  // neighbour ranks, receive lists (the halo groups) and send lists (owned points some point of the neighbour rank is joined to by an edge)
  std::vector<int> nbRank;
  std::vector<int64_t> recvPtr(1, 0), recvPts, sendPtr(1, 0), sendPts;
  for (size_t h = 0; h < halo.size();) {
    const int64_t q = owner[halo[h]];
    nbRank.push_back((int)q);
    for (; h < halo.size() && owner[halo[h]] == q; ++h) recvPts.push_back(g2l[(size_t)halo[h]]);
    recvPtr.push_back((int64_t)recvPts.size());
  }
  {
    std::vector<std::vector<int64_t>> lists(nbRank.size());
    std::vector<int> slotOf;                         // rank -> slot
    int maxRank = rank;
    for (int q : nbRank) maxRank = std::max(maxRank, q);
    slotOf.assign((size_t)maxRank + 1, -1);
    for (size_t s = 0; s < nbRank.size(); ++s) slotOf[(size_t)nbRank[s]] = (int)s;
    // an owned point p belongs to the halo of rank q exactly if an edge joins it to a point owned by q; walking the owned points in ascending
    // global index gives every list in the neighbour's halo order
    for (int64_t p = 0; p < nPoint; ++p) {
      if (owner[p] != rank) continue;
      for (int64_t t = nbPtr[p]; t < nbPtr[p + 1]; ++t) {
        const int64_t q = owner[nbPoint[t]];
        if (q == rank) continue;
        if (q < 0 || q > maxRank || slotOf[(size_t)q] < 0) return RX_ERR_ARG;   // the neighbour lists and the edges describe different graphs
        std::vector<int64_t>& l = lists[(size_t)slotOf[(size_t)q]];
        if (l.empty() || l.back() != p) l.push_back(p);   // the neighbours of p are walked together: p can only be the list's last entry
      }
    }
    for (size_t s = 0; s < nbRank.size(); ++s) {
      for (int64_t p : lists[s]) sendPts.push_back(g2l[(size_t)p]);
      sendPtr.push_back((int64_t)sendPts.size());
    }
  }
  out->nPoint = nLoc; out->nPointDomain = nOwn; out->nEdge = nE; out->nMarker = nMarker; out->nNb = (int)nbRank.size();
  out->glob = dup_vec(glob); out->edgeNodes = dup_vec(en); out->edgeNormal = dup_vec(enr); out->coord = dup_vec(xy); out->volume = dup_vec(vol);
  out->nbPtr = dup_vec(lptr); out->nbPoint = dup_vec(lnb);
  out->markerPtr = dup_vec(mptr); out->markerNode = dup_vec(mnode); out->markerNormalNeighbor = dup_vec(mnn); out->markerNormal = dup_vec(mnr);
  out->nbRank = dup_vec(nbRank); out->sendPtr = dup_vec(sendPtr); out->sendPts = dup_vec(sendPts); out->recvPtr = dup_vec(recvPtr); out->recvPts = dup_vec(recvPts);
  if (!out->glob || !out->edgeNodes || !out->edgeNormal || !out->coord || !out->volume || !out->nbPtr || !out->nbPoint || !out->markerPtr || !out->markerNode ||
      !out->markerNormalNeighbor || !out->markerNormal || !out->nbRank || !out->sendPtr || !out->sendPts || !out->recvPtr || !out->recvPts) { rx_part_free(out); return RX_ERR_CUDA; }
  return RX_OK;
}
