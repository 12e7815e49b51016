// tests/host_kernels.cpp — the one-thread-per-item CUDA kernels of csrc/rx_kernels.cuh and csrc/rx_kernels2.cuh run ON THE HOST, one simulated
// thread after another (tests/cuda_host_shim.h), in the launch order of rx_cuda.cu: k_primitive, k_transport, the gradient kernels, the limiters,
// the time-step kernels, k_edge_flux / k_edge_conv2, k_edge_visc_generic, k_accumulate, k_source and the boundary kernels — a whole staged
// residual evaluation of the path, laid out (structure of arrays, incident-edge lists, marker blocks) as rx_create lays it out.  TEST
// INFRASTRUCTURE for tests/test_host_kernels.py: the kernels' source, not only their per-item arithmetic, is compared with the reference's dumps
// where no GPU exists.  Not run here: the two-lanes-per-edge viscous kernel (warp shuffles; the thread-per-edge kernel of the same arithmetic
// stands in, as it does on the device for five and three species), the block reductions of the norms, the linear solve of the SST step (its sweeps synchronise inside a block), anything that needs the CUDA runtime.
// The SST solver's residual assembly (k_turb_rows, k_turb_dual_time, k_turb_bc_io, k_turb_bc_wall) and Postprocessing (k_turb_blend) are run.
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <vector>

#include "cuda_host_shim.h"
#include "rx_cuda.h"
#include "rx_kernels2.cuh"
#include "rx_turb.cuh"
#include "rx_turb_step.cuh"
#include "rx_visc_pair.cuh"   // compiled with -DRX_VP_NO_ASYNC_D: the 1/D prefetch through registers instead of cp.async (no PTX on the host)
#include <thread>

namespace rx { double smat[1 << 16]; double vp_sm[1 << 16]; }   // the dynamic shared memory of the boundary / thread-per-edge kernels: one thread-private matrix per lane

namespace {
using namespace rx;

template <class F>
void launch(long long n, int bs, F kernel) {
  blockDim.x = (unsigned)bs; gridDim.x = (unsigned)((n + bs - 1) / bs);
  for (unsigned b = 0; b < gridDim.x; ++b) {
    blockIdx.x = b;
    for (int t = 0; t < bs; ++t) { threadIdx.x = (unsigned)t; kernel(); }
  }
}

// the two-lanes-per-edge kernel: even lanes on this thread, odd lanes on a second one, pair after pair in the same order (cuda_host_shim.h)
template <class F>
void launch_pairs(long long nLanes, int bs, F kernel) {
  const unsigned nb = (unsigned)((nLanes + bs - 1) / bs);
  auto run = [&](int l) {
    blockDim.x = (unsigned)bs; gridDim.x = nb;
    for (unsigned b = 0; b < nb; ++b) {
      blockIdx.x = b;
      for (int t = l; t < bs; t += 2) { threadIdx.x = (unsigned)t; kernel(); emu::pair_barrier(); }
    }
  };
  std::thread odd(run, 1);
  run(0);
  odd.join();
}

struct CtxBase {
  virtual ~CtxBase() {}
  virtual void set_state(const double* U, const double* T, const double* turb6) = 0;
  virtual void set_old() = 0;
  virtual void preprocess(long long extIter, int output) = 0;
  virtual void time_step() = 0;
  virtual void residual(int what) = 0;   // 1 conv, 2 visc, 3 source
  virtual void bc() = 0;
  virtual void get(int what, double* out) = 0;   // 0 V, 1 aux6, 2 grad, 3 lim, 4 dt3, 5 R, 6 U
  virtual void update(double alpha) = 0;
  // SST solver: LinSysRes after Upwind + Viscous + Source (+ dual time) residual and the boundary conditions (tres); the Postprocessing fields (tpost)
  virtual void turb_residual(const double* kw, const double* blend3, const double* dist, double kineInf, double omegaInf, const double* kwN, const double* kwN1,
                             double* res2) = 0;
  virtual void turb_post(const double* kw, const double* dist, double sigmaK, double* post12) = 0;
  virtual void set_time_levels(const double* Un, const double* Un1) = 0;
  virtual void dual_time() = 0;
  long long intIter = 0;
  int usePair = 0;   // interior viscous edges of a nine-species context through k_edge_visc_pair instead of k_edge_visc_generic
  int nVar = 0, nPrim = 0, nGrad = 0;
  long long nP = 0;
};

template <int NS>
struct Ctx : CtxBase {
  ChemC<NS> cc;
  Opts o;
  Dev d;
  std::vector<MarkerDev> markers;
  // storage
  std::vector<double> sx; std::vector<Knot> stab;
  std::vector<int> ei, ej, nbPtr, nbIdx, incPtr, incEdge, nonPhys;
  std::vector<double> enx, eny, x, y, vol, U, Uold, V, X, Cp, gam, mu, kap, rD, grad, vort, strain, turb, dt, lamInv, lamVisc, R, lim, fluxC, fluxV;
  std::vector<double> Un, Un1; double dtND = 0.0;   // dual time stepping: the stored time levels and the physical time step
  std::vector<unsigned char> eKey;
  std::vector<unsigned long long> counters;
  std::vector<std::vector<int>> mNode, mNn; std::vector<std::vector<double>> mNx, mNy, mGhost;

  Ctx(const rx_mesh* m, const rx_chem* ch, const rx_opts* op) {
    nVar = 4 + NS; nPrim = 7 + NS; nGrad = 4 + NS; nP = m->nPoint;
    // ---- chemistry: the host side of fill_chem_ns (rx_cuda.cu)
    memset(&cc, 0, sizeof cc);
    cc.Nr = ch->Nr; cc.nKnots = ch->nKnots;
    sx.assign(ch->x, ch->x + ch->nKnots);
    stab.resize((size_t)5 * ch->nKnots * NS);
    for (int pr = 0; pr < 5; ++pr)
      for (int s = 0; s < NS; ++s)
        for (int k = 0; k < ch->nKnots; ++k) stab[((size_t)pr * ch->nKnots + k) * NS + s] = Knot{ch->y[((size_t)pr * NS + s) * ch->nKnots + k], ch->y2[((size_t)pr * NS + s) * ch->nKnots + k]};
    cc.x = sx.data(); cc.tab = stab.data();
    for (int s = 0; s < NS; ++s) { cc.M[s] = ch->mMasses[s]; cc.Ri[s] = R_UNGAS / cc.M[s]; cc.Mtot = cc.Mtot + cc.M[s]; cc.rM[s] = 0.0; }
    for (int i = 0; i < NS; ++i)
      for (int j = 0; j < NS; ++j) { cc.wPow[i][j] = std::pow(cc.M[j] / cc.M[i], 0.25); cc.wSq[i][j] = std::sqrt(8.0 * (1.0 + cc.M[i] / cc.M[j])); }
    for (int i = 0; i < NS; ++i) {
      const double mi = cc.M[i], dvi = std::cbrt(ch->diffVol[i]);
      for (int j = i; j < NS; ++j) { cc.Mij[i][j] = cc.Mij[j][i] = std::sqrt((mi * cc.M[j]) / (mi + cc.M[j])); cc.dvs[i][j] = cc.dvs[j][i] = dvi + std::cbrt(ch->diffVol[j]); }
    }
    for (int r = 0; r < NRMAX; ++r) {
      const bool on = r < ch->Nr;
      cc.As[r] = on ? ch->As[r] : 0; cc.Betas[r] = on ? ch->Betas[r] : 0; cc.Ta[r] = on ? ch->Ta[r] : 0;
      cc.AsB[r] = on ? ch->AsB[r] : 0; cc.BetasB[r] = on ? ch->BetasB[r] : 0; cc.TaB[r] = on ? ch->TaB[r] : 0;
      cc.rev[r] = on ? (ch->reversible[r] != 0.0) : 0; cc.avb[r] = on ? (ch->availBack[r] != 0.0) : 0;
      for (int s = 0; s < NS; ++s) {
        cc.nuR[r][s] = on ? ch->nuR[r * NS + s] : 0; cc.nuP[r][s] = on ? ch->nuP[r * NS + s] : 0;
        cc.expR[r][s] = on ? ch->expR[r * NS + s] : 0; cc.expP[r][s] = on ? ch->expP[r * NS + s] : 0;
      }
    }
    // ---- options as rx_create copies them
    memset(&o, 0, sizeof o);
    o.viscous = op->viscous; o.rans = op->rans; o.leastSquares = op->leastSquares;
    o.mInfty = op->mInfty; o.cfl = op->cfl; o.maxDt = op->maxDt; o.Tmin = op->Tmin; o.Tmax = op->Tmax;
    o.clipT = op->clipT; o.ignition = op->ignition; o.ignitionIter = op->ignitionIter; o.ignitionT = op->ignitionT;
    o.fuelIndex = op->fuelIndex; o.oxidizerIndex = op->oxidizerIndex;
    o.prandtlLam = op->prandtlLam; o.prandtlTurb = op->prandtlTurb; o.lewisTurb = op->lewisTurb; o.cMu = op->cMu; o.pasrLb = op->pasrLb; o.tkeInf = op->tkeInf;
    o.secondOrder = op->secondOrder; o.limiter = op->secondOrder ? op->limiter : 0; o.limCoeff = op->limCoeff; o.refElemLength = op->refElemLength;
    o.viscLimiter = op->viscLimiter ? 1 : 0;
    o.timeStepping = op->timeStepping; o.unstCfl = op->unstCfl; o.deltaUnstTime = op->deltaUnstTime; o.dualTime = op->dualTime; o.deltaUnstTimeND = op->deltaUnstTimeND;
    dtND = o.deltaUnstTimeND;
    // ---- mesh as rx_create lays it out (no padding here: np = nPoint, ne = nEdge)
    const long long nE = m->nEdge;
    memset(&d, 0, sizeof d);
    d.nPoint = nP; d.nPointDomain = m->nPointDomain; d.nEdge = nE; d.np = nP; d.ne = nE;
    ei.resize(nE); ej.resize(nE); enx.resize(nE); eny.resize(nE); x.resize(nP); y.resize(nP); vol.assign(m->volume, m->volume + nP);
    for (long long e = 0; e < nE; ++e) { ei[e] = (int)m->edgeNodes[2 * e]; ej[e] = (int)m->edgeNodes[2 * e + 1]; enx[e] = m->edgeNormal[2 * e]; eny[e] = m->edgeNormal[2 * e + 1]; }
    for (long long p = 0; p < nP; ++p) { x[p] = m->coord[2 * p]; y[p] = m->coord[2 * p + 1]; }
    nbPtr.resize(nP + 1); nbIdx.resize(m->nbPtr[nP]);
    for (long long p = 0; p <= nP; ++p) nbPtr[p] = (int)m->nbPtr[p];
    for (size_t k = 0; k < nbIdx.size(); ++k) nbIdx[k] = (int)m->nbPoint[k];
    incPtr.assign(nP + 1, 0); incEdge.resize(2 * nE);
    for (long long e = 0; e < nE; ++e) { incPtr[ei[e] + 1]++; incPtr[ej[e] + 1]++; }
    for (long long p = 0; p < nP; ++p) incPtr[p + 1] += incPtr[p];
    { std::vector<int> fill(incPtr.begin(), incPtr.end() - 1);
      for (long long e = 0; e < nE; ++e) { incEdge[fill[ei[e]]++] = (int)e; incEdge[fill[ej[e]]++] = ~(int)e; } }
    d.ei = ei.data(); d.ej = ej.data(); d.enx = enx.data(); d.eny = eny.data(); d.x = x.data(); d.y = y.data(); d.vol = vol.data();
    d.nbPtr = nbPtr.data(); d.nbIdx = nbIdx.data(); d.incPtr = incPtr.data(); d.incEdge = incEdge.data();
    auto A = [&](std::vector<double>& v, size_t n) { v.assign(std::max<size_t>(n, 1), 0.0); return v.data(); };
    d.U = A(U, nP * nVar); d.Uold = A(Uold, nP * nVar); d.R = A(R, nP * nVar); d.V = A(V, nP * nPrim); d.X = A(X, nP * NS);
    d.Cp = A(Cp, nP); d.gam = A(gam, nP); d.mu = A(mu, nP); d.kap = A(kap, nP); d.lim = A(lim, nP * 4);
    d.rD = o.viscous ? A(rD, nP * (size_t)(NS * (NS + 1) / 2)) : nullptr;
    d.grad = A(grad, nP * nGrad * 2); d.vort = A(vort, nP); d.strain = A(strain, nP); d.turb = A(turb, nP * 6);
    d.dt = A(dt, nP); d.lamInv = A(lamInv, nP); d.lamVisc = A(lamVisc, nP);
    d.fluxC = A(fluxC, nE * nVar); d.fluxV = o.viscous ? A(fluxV, nE * nVar) : nullptr;
    eKey.assign(std::max<long long>(nE, 1), 0); d.eKey = o.viscous ? eKey.data() : nullptr; d.eOrder = nullptr;
    nonPhys.assign(nP, 0); counters.assign(8, 0); d.nonPhys = nonPhys.data(); d.counters = counters.data();
    for (size_t k = 0; k < lim.size(); ++k) lim[k] = 0.0;
    // ---- markers
    mNode.resize(m->nMarker); mNn.resize(m->nMarker); mNx.resize(m->nMarker); mNy.resize(m->nMarker); mGhost.resize(m->nMarker);
    for (int k = 0; k < m->nMarker; ++k) {
      const rx_marker& s = m->markers[k];
      MarkerDev md{};
      md.kind = s.kind; md.nVertex = s.nVertex;
      mNode[k].resize(s.nVertex); mNn[k].resize(s.nVertex); mNx[k].resize(s.nVertex); mNy[k].resize(s.nVertex);
      for (int v = 0; v < s.nVertex; ++v) { mNode[k][v] = (int)s.node[v]; mNn[k][v] = (int)s.normalNeighbor[v]; mNx[k][v] = s.normal[2 * v]; mNy[k][v] = s.normal[2 * v + 1]; }
      mGhost[k].assign((size_t)std::max(s.nVertex, 1) * (7 + NS), 0.0);
      md.node = mNode[k].data(); md.nn = mNn[k].data(); md.nx = mNx[k].data(); md.ny = mNy[k].data(); md.ghost = mGhost[k].data();
      md.T = s.T; md.Vmag = s.Vmag; md.dir0 = s.dir[0]; md.dir1 = s.dir[1]; md.P = s.P; md.Twall = s.Twall; md.heatFlux = s.heatFlux;
      const bool inl = s.kind == RX_BC_INLET_TEMPERATURE_IMPOSE || s.kind == RX_BC_INLET_TOTAL_CONDITIONS || s.kind == RX_BC_INLET_MASS_FLOW || s.kind == RX_BC_SUPERSONIC_INLET;
      for (int q = 0; q < 16; ++q) md.Ys[q] = (inl && s.Ys && q < NS) ? s.Ys[q] : 0.0;
      markers.push_back(md);
    }
  }
  static bool is_inlet(int k) { return k == RX_BC_INLET_TEMPERATURE_IMPOSE || k == RX_BC_INLET_TOTAL_CONDITIONS || k == RX_BC_INLET_MASS_FLOW; }
  static bool is_super(int k) { return k == RX_BC_SUPERSONIC_INLET || k == RX_BC_SUPERSONIC_OUTLET; }

  void set_state(const double* Ua, const double* T, const double* t6) override {
    for (long long p = 0; p < nP; ++p) {
      for (int v = 0; v < nVar; ++v) U[v * nP + p] = Ua[p * nVar + v];
      V[V_T * nP + p] = T[p];
      for (int v = 0; v < 6; ++v) turb[v * nP + p] = t6 ? t6[p * 6 + v] : 0.0;
    }
  }
  void set_old() override { Uold = U; }
  void preprocess(long long extIter, int output) override {   // preprocess_primitive + preprocess_derived
    counters[0] = 0;
    launch(d.nPoint, 128, [&] { k_primitive<NS>(cc, o, d, extIter, output); });
    if (o.viscous) launch(d.nPoint, 128, [&] { k_transport<NS>(cc, d); });
    if (o.viscous || o.secondOrder) {
      if (o.leastSquares) launch(d.nPoint, 128, [&] { k_gradient_ls<NS>(d); });
      else {
        launch(d.nPointDomain, 128, [&] { k_gradient_gg_edges<NS>(d); });
        for (const MarkerDev& m : markers) if (m.nVertex) launch(m.nVertex, 64, [&] { k_gradient_gg_boundary<NS>(d, m); });
        launch(d.nPoint, 128, [&] { k_gradient_gg_finish<NS>(d); });
      }
    }
    if (o.secondOrder && o.limiter && !output) {
      if (o.limiter == 2) launch(d.nPoint, 128, [&] { k_limiter_bj(d); });
      else launch(d.nPoint, 128, [&] { k_limiter(o, d); });
    }
  }
  double global_min() {   // k_dt_min_partial / k_dt_min_final (block reductions on the device; a minimum does not depend on the order)
    double g = 1.0e6;
    for (long long p = 0; p < d.nPointDomain; ++p) if (d.vol[p] > EPS) g = fmin(g, d.dt[p]);
    return g;
  }
  void time_step() override {   // rx_time_step
    launch(d.nPointDomain, 128, [&] { k_timestep_edges(o, d); });
    for (const MarkerDev& m : markers) if (m.nVertex) launch(m.nVertex, 64, [&] { k_timestep_boundary(o, d, m); });
    Opts of = o;
    if (of.dualTime) of.timeStepping = 1;
    launch(d.nPointDomain, 128, [&] { k_timestep_finish(of, d); });
    if (o.timeStepping) {
      const double glob = global_min();
      if (!o.viscous) o.cfl = o.unstCfl;
      const int useFixed = (!o.viscous && o.cfl < EPS) ? 1 : 0;
      launch(d.nPointDomain, 256, [&] { k_dt_fill(d, &glob, o.deltaUnstTime, useFixed); });
    }
    if (o.dualTime) {
      if (intIter == 0 && o.unstCfl != 0.0) { const double glob = global_min(); k_dual_unst_dt(&glob, o.unstCfl, o.cfl, &dtND); }
      launch(d.nPointDomain, 256, [&] { k_dual_dt_clamp(d, o.maxDt, &dtND); });
    }
  }
  void set_time_levels(const double* a, const double* b) override {
    Un.assign((size_t)nP * nVar, 0.0); Un1.assign((size_t)nP * nVar, 0.0);
    for (long long p = 0; p < nP; ++p) for (int v = 0; v < nVar; ++v) { Un[v * nP + p] = a[p * nVar + v]; Un1[v * nP + p] = b[p * nVar + v]; }
  }
  void dual_time() override { launch(d.nPointDomain, 128, [&] { k_dual_time_residual(d, nVar, o.dualTime, &dtND, Un.data(), Un1.data()); }); }
  void residual(int what) override {
    if (what == 1) {
      if (o.secondOrder) launch(d.nEdge, 64, [&] { k_edge_conv2<NS>(cc, o, d); });
      else launch(d.nEdge, 128, [&] { k_edge_flux<NS>(cc, o, d); });
      launch(d.nPoint, 128, [&] { k_accumulate<NS, 1>(cc, o, d); });
    } else if (what == 2 && o.viscous) {
      if (usePair) launch_pair_kernel();
      else if (o.rans) launch(d.nEdge, EDGE_BS, [&] { k_edge_visc_generic<NS, true>(cc, o, d); });
      else launch(d.nEdge, EDGE_BS, [&] { k_edge_visc_generic<NS, false>(cc, o, d); });
      launch(d.nPoint, 128, [&] { k_accumulate<NS, 2>(cc, o, d); });
    } else if (what == 3) {
      launch(d.nPointDomain, 128, [&] { k_source<NS>(cc, o, d); });
    }
  }
  void launch_pair_kernel() {
    if constexpr (NS == 9) {
      if (o.viscLimiter) {
        if (o.rans) launch_pairs(2 * d.nEdge, VP_BS, [&] { k_edge_visc_pair<9, true, true>(cc, o, d); });
        else launch_pairs(2 * d.nEdge, VP_BS, [&] { k_edge_visc_pair<9, false, true>(cc, o, d); });
      } else if (o.rans) launch_pairs(2 * d.nEdge, VP_BS, [&] { k_edge_visc_pair<9, true>(cc, o, d); });
      else launch_pairs(2 * d.nEdge, VP_BS, [&] { k_edge_visc_pair<9, false>(cc, o, d); });
    }
  }
  void bc_marker(const MarkerDev& m) {
    if (!m.nVertex) return;
    if (is_super(m.kind)) launch(m.nVertex, EDGE_BS, [&] { k_bc_supersonic<NS>(cc, o, d, m); });
    else if (is_inlet(m.kind)) launch(m.nVertex, EDGE_BS, [&] { k_bc_inlet<NS>(cc, o, d, m); });
    else if (m.kind == RX_BC_OUTLET) launch(m.nVertex, EDGE_BS, [&] { k_bc_outlet<NS>(cc, o, d, m); });
    else if (m.kind == RX_BC_ISOTHERMAL) launch(m.nVertex, 64, [&] { k_bc_isothermal<NS>(cc, o, d, m); });
    else if (m.kind == RX_BC_HEATFLUX) launch(m.nVertex, 64, [&] { k_bc_heatflux(d, m); });
    else if (m.kind == RX_BC_EULER_WALL) launch(m.nVertex, 64, [&] { k_bc_euler_wall(o, d, m); });
  }
  void bc() override {   // rx_bc: weak conditions in marker order, then the walls
    for (const MarkerDev& m : markers) if (is_inlet(m.kind) || is_super(m.kind) || m.kind == RX_BC_OUTLET || m.kind == RX_BC_EULER_WALL) bc_marker(m);
    for (const MarkerDev& m : markers) if (m.kind == RX_BC_ISOTHERMAL || m.kind == RX_BC_HEATFLUX) bc_marker(m);
  }
  void update(double alpha) override {   // rx_update_explicit: the clipped update of every owned point; the block reductions of the norms are not formed here
    const long long nb = (d.nPointDomain + 127) / 128;
    std::vector<double> partial((size_t)std::max<long long>(nb, 1) * nVar);
    launch(nb * 128, 128, [&] { k_update<NS, false>(d, alpha, partial.data(), nullptr, nullptr); });
  }
  std::vector<double> tp, tres, tdiag, toff, tx, told, tkN, tkN1;
  void turb_load(const double* kw, const double* blend3, const double* dist) {
    tp.resize((size_t)TP_N * nP, 0.0);
    for (long long p = 0; p < nP; ++p) {
      tp[TP_K * nP + p] = kw[2 * p]; tp[TP_W * nP + p] = kw[2 * p + 1]; tp[TP_DIST * nP + p] = dist[p];
      if (blend3) { tp[TP_F1 * nP + p] = blend3[3 * p]; tp[TP_F2 * nP + p] = blend3[3 * p + 1]; tp[TP_CD * nP + p] = blend3[3 * p + 2]; }
    }
  }
  void turb_gradient() {
    double* t = tp.data();
    if (o.leastSquares) launch(d.nPoint, 128, [&] { k_turb_gradient_ls(d, t); });
    else {
      launch(d.nPoint, 128, [&] { k_turb_gradient_gg_edges(d, t); });
      for (const MarkerDev& m : markers) if (m.nVertex) launch(m.nVertex, 64, [&] { k_turb_gradient_gg_boundary(d, m, t); });
      launch(d.nPoint, 128, [&] { k_turb_gradient_gg_finish(d, t); });
    }
  }
  void turb_residual(const double* kw, const double* blend3, const double* dist, double kineInf, double omegaInf, const double* kwN, const double* kwN1,
                     double* res2) override {   // rx_turb_iterate up to the assembled residual
    turb_load(kw, blend3, dist);
    TurbC tc;
    tc.alfa_1 = tc.beta_1 / tc.beta_star - tc.sigma_om1 * 0.41 * 0.41 / std::sqrt(tc.beta_star);
    tc.alfa_2 = tc.beta_2 / tc.beta_star - tc.sigma_om2 * 0.41 * 0.41 / std::sqrt(tc.beta_star);
    tc.kineInf = kineInf; tc.omegaInf = omegaInf; tc.cflRed = 1.0; tc.relax = 1.0;
    TurbW w;
    w.nInc = (long long)incEdge.size();
    tres.assign((size_t)2 * nP, 0.0); tdiag.assign((size_t)2 * nP, 0.0); toff.assign((size_t)2 * std::max<long long>(w.nInc, 1), 0.0);
    tx.assign((size_t)2 * nP, 0.0); told.assign(tp.begin(), tp.begin() + 2 * nP);
    w.res = tres.data(); w.diag = tdiag.data(); w.off = toff.data(); w.x = tx.data(); w.old = told.data();
    turb_gradient();
    double* t = tp.data();
    launch(d.nPointDomain, 128, [&] { k_turb_rows(tc, d, t, w); });
    if (o.dualTime) {
      tkN.assign((size_t)2 * nP, 0.0); tkN1.assign((size_t)2 * nP, 0.0);
      for (long long p = 0; p < nP; ++p) for (int v = 0; v < 2; ++v) { tkN[v * nP + p] = kwN[2 * p + v]; tkN1[v * nP + p] = kwN1[2 * p + v]; }
      launch(d.nPointDomain, 128, [&] { k_turb_dual_time(d, t, w, o.dualTime, &dtND, Un.data(), Un1.data(), tkN.data(), tkN1.data()); });
    }
    for (const MarkerDev& m : markers)
      if (m.nVertex && (is_inlet(m.kind) || m.kind == RX_BC_OUTLET)) launch(m.nVertex, 64, [&] { k_turb_bc_io(tc, d, m, t, w, is_inlet(m.kind) ? 1 : 0); });
    for (const MarkerDev& m : markers)
      if (m.nVertex && (m.kind == RX_BC_ISOTHERMAL || m.kind == RX_BC_HEATFLUX)) launch(m.nVertex, 64, [&] { k_turb_bc_wall(tc, d, m, t, w); });
    for (long long p = 0; p < nP; ++p) { res2[2 * p] = tres[p]; res2[2 * p + 1] = tres[nP + p]; }
  }
  void turb_post(const double* kw, const double* dist, double sigmaK, double* post12) override {   // rx_turb_postprocess
    turb_load(kw, nullptr, dist);
    turb_gradient();
    double* t = tp.data();
    launch(d.nPoint, 128, [&] { k_turb_blend(d, t, sigmaK); });
    for (long long p = 0; p < nP; ++p) for (int v = 0; v < TP_N; ++v) post12[p * TP_N + v] = tp[v * nP + p];
  }
  void get(int what, double* out) override {
    auto aos = [&](const std::vector<double>& soa, int nv) { for (long long p = 0; p < nP; ++p) for (int v = 0; v < nv; ++v) out[p * nv + v] = soa[v * nP + p]; };
    if (what == 0) aos(V, nPrim);
    else if (what == 1) for (long long p = 0; p < nP; ++p) { double* a = out + 6 * p; a[0] = Cp[p]; a[1] = gam[p]; a[2] = mu[p]; a[3] = kap[p]; a[4] = vort[p]; a[5] = strain[p]; }
    else if (what == 2) aos(grad, nGrad * 2);
    else if (what == 3) aos(lim, 4);
    else if (what == 4) for (long long p = 0; p < nP; ++p) { out[3 * p] = dt[p]; out[3 * p + 1] = lamInv[p]; out[3 * p + 2] = lamVisc[p]; }
    else if (what == 5) aos(R, nVar);
    else if (what == 6) aos(U, nVar);
  }
};
}  // namespace

extern "C" void* hk_create(const rx_mesh* m, const rx_chem* ch, const rx_opts* op) {
  switch (ch->Ns) {
    case 9: return new Ctx<9>(m, ch, op);
    case 5: return new Ctx<5>(m, ch, op);
    case 3: return new Ctx<3>(m, ch, op);
    default: return nullptr;
  }
}
extern "C" void hk_destroy(void* h) { delete (CtxBase*)h; }
extern "C" void hk_set_state(void* h, const double* U, const double* T, const double* turb6) { ((CtxBase*)h)->set_state(U, T, turb6); }
extern "C" void hk_set_old(void* h) { ((CtxBase*)h)->set_old(); }
extern "C" void hk_preprocess(void* h, int64_t extIter, int output) { ((CtxBase*)h)->preprocess(extIter, output); }
extern "C" void hk_time_step(void* h) { ((CtxBase*)h)->time_step(); }
extern "C" void hk_residual(void* h, int what) { ((CtxBase*)h)->residual(what); }
extern "C" void hk_bc(void* h) { ((CtxBase*)h)->bc(); }
extern "C" void hk_get(void* h, int what, double* out) { ((CtxBase*)h)->get(what, out); }
extern "C" void hk_set_time_levels(void* h, const double* Un, const double* Un1) { ((CtxBase*)h)->set_time_levels(Un, Un1); }
extern "C" void hk_set_inner_iteration(void* h, int64_t it) { ((CtxBase*)h)->intIter = it; }
extern "C" void hk_dual_time(void* h) { ((CtxBase*)h)->dual_time(); }
extern "C" void hk_update(void* h, double alpha) { ((CtxBase*)h)->update(alpha); }
extern "C" void hk_turb_residual(void* h, const double* kw, const double* blend3, const double* dist, double kineInf, double omegaInf, const double* kwN,
                                 const double* kwN1, double* res2) { ((CtxBase*)h)->turb_residual(kw, blend3, dist, kineInf, omegaInf, kwN, kwN1, res2); }
extern "C" void hk_turb_post(void* h, const double* kw, const double* dist, double sigmaK, double* post12) { ((CtxBase*)h)->turb_post(kw, dist, sigmaK, post12); }
extern "C" void hk_use_pair(void* h, int on) { ((CtxBase*)h)->usePair = on; }

