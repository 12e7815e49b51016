// tests/host_emu.cpp — the per-point / per-edge arithmetic of the CUDA kernels, compiled FOR THE HOST (csrc/rx_device.cuh and csrc/rx_flux.cuh
// define their functions with RX_HD, which is plain `inline` outside nvcc) behind a small C interface, so that the CPU test tier can check the
// kernels' source arithmetic against the reference's golden dumps without a GPU (tests/test_host_emulation.py).  TEST INFRASTRUCTURE: nothing of
// the product links or loads this; the product path runs these functions on the device only.
//   emu_cons2prim   cons2prim<NS> as k_primitive calls it (rx_kernels.cuh), one point at a time, with the Solution_Old retry
//   emu_transport   transport<NS> (Wilke rules)
//   emu_ausm        ausm<NS>, one interior edge at a time
//   emu_source      mass_production<NS> as k_source_points scales it
//   emu_qr          colpiv_qr_solve<9, 2, 1> as k_qr_eval_serial calls it (the SST closure's pivoted Householder QR, serial form)
//   emu_viscous     viscous_flux<NS, EDGE = true> (Stefan-Maxwell / BiCGSTAB, SST closure / pivoted QR; the thread-per-edge form the boundary
//                   kernels and the five- and three-species edge kernels run), one interior edge at a time, with 1/D_st formed as k_transport does
// On the host exp / pow / log are libm's — the bits the device's glibc-exact restatements (rx_libm.cuh) reproduce — and every division is the
// plain IEEE one the device's fast paths are built to equal.
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#include "rx_cuda.h"
#include "rx_device.cuh"
#include "rx_flux.cuh"

namespace {
using namespace rx;

template <int NS>
struct HostChem {
  ChemC<NS> cc;
  std::vector<double> x;
  std::vector<Knot> tab;
  // the host-side part of fill_chem_ns (rx_cuda.cu); the refined reciprocals rM / rwSq belong to the device's fast division paths and stay zero
  explicit HostChem(const rx_chem* ch) {
    memset(&cc, 0, sizeof cc);
    cc.Nr = ch->Nr; cc.nKnots = ch->nKnots;
    x.assign(ch->x, ch->x + ch->nKnots);
    tab.resize((size_t)5 * ch->nKnots * NS);
    for (int pr = 0; pr < 5; ++pr)
      for (int s = 0; s < NS; ++s)
        for (int k = 0; k < ch->nKnots; ++k) {
          const size_t src = ((size_t)pr * NS + s) * ch->nKnots + k;
          tab[((size_t)pr * ch->nKnots + k) * NS + s] = Knot{ch->y[src], ch->y2[src]};
        }
    cc.x = x.data(); cc.tab = tab.data();
    cc.Mtot = 0.0;
    for (int s = 0; s < NS; ++s) { cc.M[s] = ch->mMasses[s]; cc.Ri[s] = R_UNGAS / cc.M[s]; cc.Mtot = cc.Mtot + cc.M[s]; }
    for (int i = 0; i < NS; ++i)
      for (int j = 0; j < NS; ++j) {
        cc.wPow[i][j] = std::pow(cc.M[j] / cc.M[i], 0.25);
        cc.wSq[i][j] = std::sqrt(8.0 * (1.0 + cc.M[i] / cc.M[j]));
      }
    for (int i = 0; i < NS; ++i) {
      const double mi = cc.M[i], dvi = std::cbrt(ch->diffVol[i]);
      for (int j = i; j < NS; ++j) {
        cc.Mij[i][j] = cc.Mij[j][i] = std::sqrt((mi * cc.M[j]) / (mi + cc.M[j]));
        cc.dvs[i][j] = cc.dvs[j][i] = dvi + std::cbrt(ch->diffVol[j]);
      }
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
  }
};

Opts opts_of(const rx_opts* op) {
  Opts o;
  memset(&o, 0, sizeof o);
  o.viscous = op->viscous; o.rans = op->rans; o.leastSquares = op->leastSquares;
  o.mInfty = op->mInfty; o.cfl = op->cfl; o.maxDt = op->maxDt; o.Tmin = op->Tmin; o.Tmax = op->Tmax;
  o.clipT = op->clipT; o.ignition = op->ignition; o.ignitionIter = op->ignitionIter; o.ignitionT = op->ignitionT;
  o.fuelIndex = op->fuelIndex; o.oxidizerIndex = op->oxidizerIndex;
  o.prandtlLam = op->prandtlLam; o.prandtlTurb = op->prandtlTurb; o.lewisTurb = op->lewisTurb; o.cMu = op->cMu;
  o.pasrLb = op->pasrLb; o.tkeInf = op->tkeInf;
  return o;
}

template <int NS>
int cons2prim_ns(const rx_chem* ch, const rx_opts* op, int64_t extIter, int64_t n, const double* U, const double* Uold, const double* T0, const double* ke,
                 double* V, int32_t* nonPhys) {
  HostChem<NS> h(ch);
  const Opts o = opts_of(op);
  constexpr int NV = 4 + NS, NP = 7 + NS;
  int failures = 0;
  for (int64_t p = 0; p < n; ++p) {
    double u[NV], v[NP];
    for (int k = 0; k < NV; ++k) u[k] = U[p * NV + k];
    for (int k = 0; k < NP; ++k) v[k] = 0.0;
    v[V_T] = T0[p];
    bool np = false;
    for (int attempt = 0; attempt < 2; ++attempt) {   // k_primitive: second pass from Solution_Old where the first state is non-physical
      bool failed = false;
      const bool bad = cons2prim(h.cc, o, (long long)extIter, u, v, o.rans ? ke[p] : 0.0, &failed);
      if (failed) ++failures;
      if (attempt == 0) np = bad;
      if (!(bad && extIter > 0) || attempt == 1 || !Uold) break;
      for (int k = 0; k < NV; ++k) u[k] = Uold[p * NV + k];
    }
    if (o.ignition && extIter < o.ignitionIter)
      if (v[V_Y + o.fuelIndex] > 0.4 && v[V_Y + o.oxidizerIndex] > 0.2 && v[V_T] < o.ignitionT) v[V_T] = o.ignitionT;
    for (int k = 0; k < NP; ++k) V[p * NP + k] = v[k];
    if (nonPhys) nonPhys[p] = np ? 1 : 0;
  }
  return failures;
}

template <int NS>
void transport_ns(const rx_chem* ch, int64_t n, const double* T, const double* Y, double* mu, double* kappa) {
  HostChem<NS> h(ch);
  for (int64_t p = 0; p < n; ++p) transport(h.cc, T[p], Y + p * NS, mu[p], kappa[p]);
}

template <int NS>
void ausm_ns(double mInfty, int64_t nE, const double* Vi, const double* Vj, const double* normal, double* res) {
  constexpr int NV = 4 + NS, NP = 7 + NS;
  for (int64_t e = 0; e < nE; ++e) ausm<NS>(mInfty, Vi + e * NP, Vj + e * NP, normal[2 * e], normal[2 * e + 1], res + e * NV);
}

template <int NS>
void source_ns(const rx_chem* ch, int rans, double cMu, double pasrLb, int64_t n, const double* T, const double* rho, const double* Y, const double* omT,
               const double* vol, double* res) {
  HostChem<NS> h(ch);
  for (int64_t p = 0; p < n; ++p) {
    double om[NS];
    mass_production<NS>(h.cc, rans != 0, cMu, pasrLb, T[p], rho[p], Y + p * NS, omT ? omT[p] : 0.0, om);
    const double Volume = vol[p];
    for (int s = 0; s < NS; ++s) res[p * NS + s] = om[s] * (-Volume / (1.0 / 1.0));   // k_source_points
  }
}

template <int NS>
void viscous_ns(const rx_chem* ch, const rx_opts* op, int64_t nE, const int64_t* en, const double* normal, int64_t nP, const double* V, const double* grad,
                const double* mu, const double* kap, const double* xy, const double* turb6, double* res) {
  HostChem<NS> h(ch);
  const Opts o = opts_of(op);
  constexpr int NV = 4 + NS, NP = 7 + NS, NG = 4 + NS, NQ = NS * (NS + 1) / 2;
  // per point: mole fractions (k_primitive) and 1/D_st, upper triangle (k_transport's plain-division form; GetDij_SM reacting_model_library.cpp:751-766)
  std::vector<double> X((size_t)nP * NS), rD((size_t)NQ * nP);
  for (int64_t p = 0; p < nP; ++p) {
    molar_from_mass(h.cc, V + p * NP + V_Y, &X[(size_t)p * NS]);
    const double T175 = 1.0e-3 * pow175(V[p * NP + V_T]);
    const double Patm = V[p * NP + V_P] * 1.0 / 101325.0;
    int q = 0;
    for (int s = 0; s < NS; ++s)
      for (int t = s; t < NS; ++t, ++q) {
        const double den = Patm * h.cc.Mij[s][t] * h.cc.dvs[s][t] * h.cc.dvs[s][t];
        rD[(size_t)q * nP + p] = 1.0 / ((T175 / den) / (1.0 * 1.0 * 1.0e4));
      }
  }
  auto load = [&](int64_t p, ViscEnd<NS>& E) {   // load_visc_end (rx_kernels2.cuh)
    for (int v = 0; v < NP; ++v) E.V[v] = V[p * NP + v];
    for (int k = 0; k < NG * 2; ++k) E.g[k] = grad[p * NG * 2 + k];
    for (int s = 0; s < NS; ++s) E.X[s] = X[(size_t)p * NS + s];
    E.mu = mu[p]; E.kap = kap[p];
    E.rD = rD.data() + p; E.rDs = nP;
    E.x[0] = xy[2 * p]; E.x[1] = xy[2 * p + 1];
    if (o.rans) { E.muT = turb6[6 * p]; E.ke = turb6[6 * p + 1]; E.gk[0] = turb6[6 * p + 3]; E.gk[1] = turb6[6 * p + 4]; }
    else { E.muT = 1.0; E.ke = 0.0; E.gk[0] = E.gk[1] = 0.0; }
  };
  std::vector<double> scratch((size_t)8 * NS * NS);
  for (int64_t e = 0; e < nE; ++e) {
    const int64_t i = en[2 * e], j = en[2 * e + 1];
    ViscEnd<NS> I, J;
    load(i, I); load(j, J);
    const double sigma_k = o.rans ? turb6[6 * i + 5] : 1.0;   // sigma_k of the first node (:5345)
    if (o.rans) viscous_flux<NS, true, 1, true>(h.cc, o, I, J, normal[2 * e], normal[2 * e + 1], sigma_k, scratch.data(), res + e * NV);
    else viscous_flux<NS, true, 1, false>(h.cc, o, I, J, normal[2 * e], normal[2 * e + 1], sigma_k, scratch.data(), res + e * NV);
  }
}
}  // namespace

#define EMU_NS(Ns, CALL)                \
  switch (Ns) {                         \
    case 9: { constexpr int NS = 9; CALL; break; } \
    case 5: { constexpr int NS = 5; CALL; break; } \
    case 3: { constexpr int NS = 3; CALL; break; } \
    default: return -1;                 \
  }

extern "C" int emu_cons2prim(const rx_chem* ch, const rx_opts* op, int64_t extIter, int64_t n, const double* U, const double* Uold, const double* T0,
                             const double* ke, double* V, int32_t* nonPhys) {
  int f = 0;
  EMU_NS(ch->Ns, f = cons2prim_ns<NS>(ch, op, extIter, n, U, Uold, T0, ke, V, nonPhys));
  return f;
}
extern "C" int emu_transport(const rx_chem* ch, int64_t n, const double* T, const double* Y, double* mu, double* kappa) {
  EMU_NS(ch->Ns, transport_ns<NS>(ch, n, T, Y, mu, kappa));
  return 0;
}
extern "C" int emu_ausm(int Ns, double mInfty, int64_t nE, const double* Vi, const double* Vj, const double* normal, double* res) {
  EMU_NS(Ns, ausm_ns<NS>(mInfty, nE, Vi, Vj, normal, res));
  return 0;
}
extern "C" int emu_source(const rx_chem* ch, int rans, double cMu, double pasrLb, int64_t n, const double* T, const double* rho, const double* Y,
                          const double* omT, const double* vol, double* res) {
  EMU_NS(ch->Ns, source_ns<NS>(ch, rans, cMu, pasrLb, n, T, rho, Y, omT, vol, res));
  return 0;
}
extern "C" int emu_viscous(const rx_chem* ch, const rx_opts* op, int64_t nE, const int64_t* en, const double* normal, int64_t nP, const double* V,
                           const double* grad, const double* mu, const double* kap, const double* xy, const double* turb6, double* res) {
  EMU_NS(ch->Ns, viscous_ns<NS>(ch, op, nE, en, normal, nP, V, grad, mu, kap, xy, turb6, res));
  return 0;
}
extern "C" int emu_qr(int64_t n, const double* A, const double* b, double* xs) {   // k_qr_eval_serial (rx_cuda.cu)
  constexpr int NS = 9;
  for (int64_t e = 0; e < n; ++e) {
    double Al[NS * NS], bl[2][NS], xl[2][NS];
    for (int k = 0; k < NS * NS; ++k) Al[k] = A[e * NS * NS + k];
    for (int dd = 0; dd < 2; ++dd) for (int i = 0; i < NS; ++i) bl[dd][i] = b[(e * 2 + dd) * NS + i];
    rx::colpiv_qr_solve<NS, 2, 1>(Al, bl, xl);
    for (int dd = 0; dd < 2; ++dd) for (int i = 0; i < NS; ++i) xs[(e * 2 + dd) * NS + i] = xl[dd][i];
  }
  return 0;
}
