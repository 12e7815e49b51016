// rx_kernels.cuh — CUDA kernels of the reactive residual path (sm_100a, FP64, no tensor cores: nothing here is a dense
// contraction).  State is structure-of-arrays: field f of point p lives at base[f*np + p] (np = padded point count), so
// a warp's loads of one field are one or two contiguous 128-byte lines.  Scatter is owner-computes: every point sums the
// fluxes of its incident edges in ascending edge order — the summation order of the reference's edge loops
// (LinSysRes.AddBlock/SubtractBlock, solver_direct_reactive.cpp:2760-2761,5374-5375) — so accumulation is deterministic
// and needs no atomics.
#pragma once
#include "rx_flux.cuh"
#ifndef RX_PR_MINB
#define RX_PR_MINB 4
#endif
#ifndef RX_GR_MINB
#define RX_GR_MINB 4
#endif
#ifndef RX_TR_MINB
#define RX_TR_MINB 3
#endif

namespace rx {

struct Dev {   // device pointers + sizes, passed by value to kernels
  long long nPoint, nPointDomain, nEdge, np, ne;   // np / ne: padded strides
  // mesh
  const int *ei, *ej; const double *enx, *eny;
  const double *x, *y, *vol;
  const int *nbPtr, *nbIdx;        // neighbour CSR (WLS gradient), reference order
  const int *incPtr, *incEdge;     // incident-edge CSR in ascending edge order; entry = e (point is node 0) or ~e (node 1)
  // state
  double *U, *Uold, *V, *X, *Cp, *gam, *mu, *kap, *rD, *grad, *vort, *strain, *turb, *dt, *lamInv, *lamVisc, *R, *lim;
  double *fluxC, *fluxV;
  const int* eOrder;        // processing order of the viscous edge kernel (rx_order.cuh); nullptr = edge order
  unsigned char* eKey;      // per edge: iterations of its last Stefan–Maxwell solve (+19 if the SST closure ran its QR)
  int* nonPhys; unsigned long long* counters;   // counters[0] = non-physical points of the last preprocess
};

struct MarkerDev { int kind, nVertex; const int *node, *nn; const double *nx, *ny; double T, Vmag, dir0, dir1, P, Twall, heatFlux; double Ys[16];
                   double* ghost; };   // ghost[v * nVertex + k]: primitive state imposed at vertex k (CharacPrimVar of the reference)

__global__ void k_aos_to_soa(const double* __restrict__ aos, double* __restrict__ soa, long long n, int nv, long long stride) {
  long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= n * nv) return;
  long long p = t / nv; int v = (int)(t - p * nv);
  soa[v * stride + p] = aos[t];
}
__global__ void k_soa_to_aos(const double* __restrict__ soa, double* __restrict__ aos, long long n, int nv, long long stride) {
  long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= n * nv) return;
  long long p = t / nv; int v = (int)(t - p * nv);
  aos[t] = soa[v * stride + p];
}
__global__ void k_gather_field(const double* __restrict__ soa, double* __restrict__ aos, long long n, int nvOut, int col) {
  long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (p < n) aos[p * nvOut + col] = soa[p];
}
__global__ void k_scatter_field(const double* __restrict__ aos, double* __restrict__ soa, long long n, int nvIn, int col) {
  long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (p < n) soa[p] = aos[p * nvIn + col];
}

__global__ void k_fill_recip(double* v, int n) {   // v[i] <- the refined reciprocal frecip() forms for v[i]
  const int i = threadIdx.x;
  if (i < n) v[i] = frecip(v[i]).r;
}

// Wilke mixing rules as transport<> in rx_device.cuh (ComputeEta :634-656, ComputeLambda :671-696), with the 27 quotients and
// 9 square roots of one species row issued as independent batches (rx_fdiv.cuh); same values bit for bit.
template <int NS>
__device__ __forceinline__ void transport_dev(const ChemC<NS>& cc, double T, const double* Y, double& muMix, double& kMix) {
  SplPos sp = spl_pos(cc, T);
  double mu[NS], yomC[NS], yomR[NS], muR[NS], kaR[NS], ycR[NS], yrR[NS];
#pragma unroll
  for (int s = 0; s < NS; ++s) {
    mu[s] = spl_eval(cc, P_MU, s, sp);
    kaR[s] = spl_eval(cc, P_KAPPA, s, sp);
    yomC[s] = clipy<NS>(Y[s]);
    yomR[s] = Y[s];
  }
  div_by_M(cc, yomC);
  div_by_M(cc, yomR);
#pragma unroll
  for (int s = 0; s < NS; ++s) { muR[s] = mu[s]; ycR[s] = yomC[s]; yrR[s] = yomR[s]; }
  Recip rmu[NS];
#pragma unroll
  for (int s = 0; s < NS; ++s) rmu[s] = frecip(mu[s]);
  // operand ranges of the quotients below, tested once per point instead of once per quotient (rx_fdiv.cuh: fdivu)
  bool inOk = true;
#pragma unroll
  for (int s = 0; s < NS; ++s) inOk = inOk && mid_range(mu[s]) && num_range(yomC[s]) && num_range(1.065 * yomR[s]);
  muMix = 0.0; kMix = 0.0;
#pragma unroll 1
  for (int i = 0; i < NS; ++i) {
    const double mui = muR[0], kai = kaR[0], yci = ycR[0], yri = yrR[0];
    double phiE = 0.0, phiL = 0.0, tj[NS], qe[NS], ql[NS];
    bool bad = !inOk;
#pragma unroll
    for (int j = 0; j < NS; ++j) tj[j] = fdivu(mui, rmu[j]);
#pragma unroll
    for (int j = 0; j < NS; ++j) tj[j] = 1.0 + fsqrt(tj[j], bad) * cc.wPow[i][j];
#pragma unroll
    for (int j = 0; j < NS; ++j) {
      const Recip rw{cc.wSq[i][j], cc.rwSq[i][j]};
      qe[j] = fdivu(yomC[j], rw);
      ql[j] = fdivu(1.065 * yomR[j], rw);
    }
    if (bad) {   // out-of-range operands only
#pragma unroll
      for (int j = 0; j < NS; ++j) {
        tj[j] = 1.0 + sqrt(mui / mu[j]) * cc.wPow[i][j];
        qe[j] = yomC[j] / cc.wSq[i][j];
        ql[j] = 1.065 * yomR[j] / cc.wSq[i][j];
      }
    }
#pragma unroll
    for (int j = 0; j < NS; ++j) {
      phiE += qe[j] * tj[j] * tj[j];
      const double add = ql[j] * tj[j] * tj[j];
      phiL = (j != i) ? phiL + add : phiL;
    }
    phiL += yri;
    muMix += mui * yci / phiE;
    kMix += kai * yri / phiL;
#pragma unroll
    for (int m = 0; m + 1 < NS; ++m) { muR[m] = muR[m + 1]; kaR[m] = kaR[m + 1]; ycR[m] = ycR[m + 1]; yrR[m] = yrR[m + 1]; }
  }
}

// ---- laminar transport of a point: Wilke mixture viscosity / conductivity and the inverse binary diffusion coefficients
// (CReactiveNSVariable::SetPrimVar, variable_direct_reactive.cpp:1188-1228).  Its own kernel: together with the
// temperature solve the code no longer fits the instruction cache.  Reads P, Y of the point as k_primitive left them, and
// the temperature from mu[p]: the ignition overwrite of V[T] comes after SetPrimVar in the reference (:1013-1024).
template <int NS>
__global__ void __launch_bounds__(128, RX_TR_MINB) k_transport(const __grid_constant__ ChemC<NS> cc, const Dev d) {
  const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (p >= d.nPoint) return;
  double Y[NS];
#pragma unroll
  for (int s = 0; s < NS; ++s) Y[s] = d.V[(V_Y + s) * d.np + p];
  const double T = d.mu[p], P = d.V[V_P * d.np + p];   // k_primitive parks the temperature SetPrimVar saw in mu[p]
  double mu, ka;
  transport_dev(cc, T, Y, mu, ka);
  d.mu[p] = mu; d.kap[p] = ka;
  // 1/D_st of this point, upper triangle (GetDij_SM :751-766; /(V_ref*L_ref*1e4) variable_direct_reactive.cpp:1212;
  // cwiseInverse numerics_direct_reactive.cpp:1447).  Stored per point so that an edge needs one division per pair.
  const double T175 = 1.0e-3 * pow175(T);
  const double Patm = P * 1.0 / 101325.0;
  const Recip r1e4 = frecip(1.0 * 1.0 * 1.0e4);
  const bool dOk = mid_range(T175) && mid_range(Patm);   // then every quotient below has mid-range operands (pair constants are O(1..100))
  constexpr int NQ = NS * (NS + 1) / 2, QB = NQ % 5 == 0 ? 5 : (NQ % 3 == 0 ? 3 : 1);   // nine species: 45 pairs in batches of five
  static_assert(NQ % QB == 0, "pair batches");
  int s = 0, t = 0;
#pragma unroll 1
  for (int q0 = 0; q0 < NQ; q0 += QB) {
    double den[QB], D[QB], rDv[QB];
    const bool bad = !dOk;
#pragma unroll
    for (int b = 0; b < QB; ++b) {
      den[b] = Patm * cc.Mij[s][t] * cc.dvs[s][t] * cc.dvs[s][t];
      if (++t == NS) { ++s; t = s; }
    }
#pragma unroll
    for (int b = 0; b < QB; ++b) D[b] = fdivu(fdivu(T175, frecip(den[b])), r1e4);
#pragma unroll
    for (int b = 0; b < QB; ++b) rDv[b] = fdivu(1.0, frecip(D[b]));
    if (bad) {
#pragma unroll
      for (int b = 0; b < QB; ++b) rDv[b] = 1.0 / ((T175 / den[b]) / (1.0 * 1.0 * 1.0e4));
    }
#pragma unroll
    for (int b = 0; b < QB; ++b) d.rD[(size_t)(q0 + b) * d.np + p] = rDv[b];
  }
}

// ---- SetPrimitive_Variables (solver_direct_reactive.cpp:985-1038) + CReactiveNSVariable::SetPrimVar (variable_direct_reactive.cpp:1188-1228)
template <int NS>
__global__ void __launch_bounds__(128, RX_PR_MINB) k_primitive(const __grid_constant__ ChemC<NS> cc, const __grid_constant__ Opts o, const Dev d,
                                                   long long extIter, int output) {
  const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (p >= d.nPoint) return;
  constexpr int NV = 4 + NS, NP = 7 + NS;
  double U[NV], V[NP];
#pragma unroll
  for (int v = 0; v < NV; ++v) U[v] = d.U[v * d.np + p];
  V[V_T] = d.V[V_T * d.np + p];
  const double ke = o.rans ? d.turb[1 * d.np + p] : 0.0;
  bool np = false;
#pragma unroll 1
  for (int attempt = 0; attempt < 2; ++attempt) {   // one copy of the temperature solve in the code; second pass = :1001-1008
    bool failed = false;
    const bool bad = cons2prim(cc, o, extIter, U, V, ke, &failed);
    if (failed) atomicAdd(d.counters + 4, 1ULL);   // the reference throws here (variable_direct_reactive.cpp:683, :724)
    if (attempt == 0) np = bad;
    if (!(bad && extIter > 0) || attempt == 1) break;
#pragma unroll
    for (int v = 0; v < NV; ++v) U[v] = d.Uold[v * d.np + p];
  }
  const double* Y = V + V_Y;
  const double Rg = rgas(cc, Y);
  const double a = V[V_A], T = V[V_T];
  const double Cp = (a * a * Rg) / (a * a - Rg * T);   // ComputeCP_FromSoundSpeed, reacting_model_library.hpp:408-411
  d.Cp[p] = Cp;
  d.gam[p] = (Cp / (Cp - Rg) - 1.0) + 1.0;             // dPdU[rhoE] = Gamma - 1 (:843-849); readers add the 1 back
  double X[NS];
  molar_from_mass(cc, Y, X);
#pragma unroll
  for (int s = 0; s < NS; ++s) d.X[s * d.np + p] = X[s];
  if (o.viscous) d.mu[p] = T;                          // for k_transport, which follows on the stream
  if (o.ignition && extIter < o.ignitionIter) {        // ignition hack: overwrite T only (:1013-1024)
    if (Y[o.fuelIndex] > 0.4 && Y[o.oxidizerIndex] > 0.2 && V[V_T] < o.ignitionT) V[V_T] = o.ignitionT;
  }
#pragma unroll
  for (int v = 0; v < NV; ++v) d.U[v * d.np + p] = U[v];
#pragma unroll
  for (int v = 0; v < NP; ++v) d.V[v * d.np + p] = V[v];
  d.nonPhys[p] = np ? 1 : 0;
  if (np) atomicAdd(d.counters, 1ULL);
  if (!output) {
#pragma unroll
    for (int v = 0; v < NV; ++v) d.R[v * d.np + p] = 0.0;
  }
}

template <int NS>
__device__ __forceinline__ void load_gradvars(const Dev& d, long long p, double* pv) {
  pv[G_T] = d.V[V_T * d.np + p]; pv[G_VX] = d.V[V_VX * d.np + p]; pv[G_VX + 1] = d.V[(V_VX + 1) * d.np + p]; pv[G_P] = d.V[V_P * d.np + p];
#pragma unroll
  for (int s = 0; s < NS; ++s) pv[G_X + s] = d.X[s * d.np + p];
}
template <int NS>
__device__ __forceinline__ void store_vort_strain(const Dev& d, long long p, const double* g /* [nG][2] */) {
  // CReactiveNSVariable::SetVorticity / SetStrainMag, variable_direct_reactive.cpp:1038-1097
  d.vort[p] = g[2 * 2 + 0] - g[1 * 2 + 1];
  double div = 0.0;
  div += g[1 * 2 + 0]; div += g[2 * 2 + 1];
  double sm = 0.0;
  double t = g[1 * 2 + 0] - 1.0 / 3.0 * div; sm += t * t;
  t = g[2 * 2 + 1] - 1.0 / 3.0 * div; sm += t * t;
  t = 0.5 * (g[1 * 2 + 1] + g[2 * 2 + 0]); sm += 2.0 * (t * t);
  d.strain[p] = sqrt(2.0 * sm);
}

// Vorticity / strain magnitude of listed points from their stored gradients.  The reference forms both for ALL points after
// Set_MPI_Primitive_Gradient (solver_direct_reactive.cpp:4716-4736, :5049): the gradient kernels above see a halo point before
// its gradient has arrived, so the halo unpack of the gradients (rx_halo_unpack, what = 1) re-forms them for the received points.
template <int NS>
__global__ void k_vort_strain_points(const Dev d, long long n, const long long* __restrict__ pts) {
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= n) return;
  const long long p = pts[t];
  double g[(4 + NS) * 2];
#pragma unroll
  for (int k = 2; k < 6; ++k) g[k] = d.grad[k * d.np + p];   // rows u, v
  store_vort_strain<NS>(d, p, g);
}

// ---- SetPrimitive_Gradient_LS (solver_direct_reactive.cpp:4887-5050), one thread per owned point.
// The reference divides 3 + 2*nPrimVarGrad products by the same squared distance per neighbour.  FAST: one refined
// reciprocal per neighbour serves all 29 quotients (rx_fdiv.cuh: correctly rounded where its range test passes, which is
// ORed into one flag for the whole point); if the flag came up the point is redone with plain divisions.
template <int NS, bool FAST>
__device__ __forceinline__ bool ls_sums(const Dev& d, long long p, const double* pi, double* C, double& r11, double& r12, double& r22) {
  constexpr int NG = 4 + NS;
  double pj[NG];
  bool bad = false;
#pragma unroll
  for (int k = 0; k < NG * 2; ++k) C[k] = 0.0;
  r11 = 0.0; r12 = 0.0; r22 = 0.0;
  const double xi = d.x[p], yi = d.y[p];
  for (int k = d.nbPtr[p]; k < d.nbPtr[p + 1]; ++k) {
    const int q = d.nbIdx[k];
    load_gradvars<NS>(d, q, pj);
    const double d0 = d.x[q] - xi, d1 = d.y[q] - yi;
    const double weight = 0.0 + d0 * d0 + d1 * d1;
    if (weight > EPS) {
      if (FAST) {
        // operand ranges instead of one test per quotient: offsets zero or in [2^-100, 2^101), differences zero or in
        // [2^-500, 2^201)  =>  weight in [2^-200, 2^203), numerators zero or in [2^-600, 2^302): fdivu is exact rounding
        const Recip rw = frecip(weight);
        bool ok = (exp_in<-100, 100>(d0) || d0 == 0.0) && (exp_in<-100, 100>(d1) || d1 == 0.0);
        r11 += fdivu(d0 * d0, rw);
        r12 += fdivu(d0 * d1, rw);
        r22 += fdivu(d1 * d1, rw);
#pragma unroll
        for (int v = 0; v < NG; ++v) {
          const double df = pj[v] - pi[v];
          ok = ok && (exp_in<-500, 200>(df) || df == 0.0);
          C[v * 2 + 0] += fdivu(d0 * df, rw);
          C[v * 2 + 1] += fdivu(d1 * df, rw);
        }
        bad = bad || !ok;
      } else {
        r11 += d0 * d0 / weight;
        r12 += d0 * d1 / weight;
        r22 += d1 * d1 / weight;
#pragma unroll
        for (int v = 0; v < NG; ++v) {
          C[v * 2 + 0] += d0 * (pj[v] - pi[v]) / weight;
          C[v * 2 + 1] += d1 * (pj[v] - pi[v]) / weight;
        }
      }
    }
  }
  return bad;
}

template <int NS>
__global__ void __launch_bounds__(128, RX_GR_MINB) k_gradient_ls(const Dev d) {
  const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (p >= d.nPoint) return;
  constexpr int NG = 4 + NS;
  double g[NG * 2];
  if (p < d.nPointDomain) {
    double pi[NG], C[NG * 2];
    load_gradvars<NS>(d, p, pi);
    double r11, r12, r22;
    if (ls_sums<NS, true>(d, p, pi, C, r11, r12, r22)) ls_sums<NS, false>(d, p, pi, C, r11, r12, r22);
    if (r11 > EPS) r11 = sqrt(r11); else r11 = 0.0;
    if (fabs(r11) > EPS) r12 = r12 / r11; else r12 = 0.0;
    if (r22 - r12 * r12 > EPS) r22 = sqrt(r22 - r12 * r12); else r22 = 0.0;
    double detR2 = (r11 * r22) * (r11 * r22);
    bool singular = false;
    if (fabs(detR2) < EPS) { detR2 = 1.0; singular = true; }
    double S00 = 0.0, S01 = 0.0, S11 = 0.0;
    if (!singular) {
      S00 = (r12 * r12 + r22 * r22) / detR2;
      S01 = -r11 * r12 / detR2;
      S11 = r11 * r11 / detR2;
    }
#pragma unroll
    for (int v = 0; v < NG; ++v) {
      g[v * 2 + 0] = (0.0 + C[v * 2 + 0] * S00) + C[v * 2 + 1] * S01;
      g[v * 2 + 1] = (0.0 + C[v * 2 + 0] * S01) + C[v * 2 + 1] * S11;
    }
#pragma unroll
    for (int k = 0; k < NG * 2; ++k) d.grad[k * d.np + p] = g[k];
  } else {
#pragma unroll
    for (int k = 0; k < NG * 2; ++k) g[k] = d.grad[k * d.np + p];   // halo gradients arrive by exchange
  }
  store_vort_strain<NS>(d, p, g);
}

// ---- SetPrimitive_Gradient_GG (solver_direct_reactive.cpp:4784-4880): interior-edge part, owner-computes.
// The species rows use the mole fractions of the edge's FIRST node for both ends (:4811-4814) — reproduced.
template <int NS>
__global__ void __launch_bounds__(128) k_gradient_gg_edges(const Dev d) {
  const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (p >= d.nPointDomain) return;
  constexpr int NG = 4 + NS;
  double g[NG * 2], pi[NG], pj[NG];
#pragma unroll
  for (int k = 0; k < NG * 2; ++k) g[k] = 0.0;
  for (int k = d.incPtr[p]; k < d.incPtr[p + 1]; ++k) {
    const int code = d.incEdge[k];
    const bool first = code >= 0;
    const int e = first ? code : ~code;
    const int i = d.ei[e], j = d.ej[e];
    load_gradvars<NS>(d, i, pi);
    load_gradvars<NS>(d, j, pj);
#pragma unroll
    for (int s = 0; s < NS; ++s) pj[G_X + s] = pi[G_X + s];
    const double nx = d.enx[e], ny = d.eny[e];
#pragma unroll
    for (int v = 0; v < NG; ++v) {
      const double avg = 0.5 * (pi[v] + pj[v]);
      if (first) { g[v * 2 + 0] += avg * nx; g[v * 2 + 1] += avg * ny; }
      else { g[v * 2 + 0] -= avg * nx; g[v * 2 + 1] -= avg * ny; }
    }
  }
#pragma unroll
  for (int k = 0; k < NG * 2; ++k) d.grad[k * d.np + p] = g[k];
}
template <int NS>
__global__ void k_gradient_gg_boundary(const Dev d, const MarkerDev m) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= m.nVertex) return;
  const int p = m.node[k];
  if (p >= d.nPointDomain) return;
  constexpr int NG = 4 + NS;
  double pi[NG];
  load_gradvars<NS>(d, p, pi);
#pragma unroll
  for (int v = 0; v < NG; ++v) {
    d.grad[(v * 2 + 0) * d.np + p] -= pi[v] * m.nx[k];
    d.grad[(v * 2 + 1) * d.np + p] -= pi[v] * m.ny[k];
  }
}
template <int NS>
__global__ void __launch_bounds__(128) k_gradient_gg_finish(const Dev d) {
  const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (p >= d.nPoint) return;
  constexpr int NG = 4 + NS;
  double g[NG * 2];
  const double vol = d.vol[p];
#pragma unroll
  for (int k = 0; k < NG * 2; ++k) {
    g[k] = d.grad[k * d.np + p];
    if (p < d.nPointDomain) { g[k] = g[k] / vol; d.grad[k * d.np + p] = g[k]; }
  }
  store_vort_strain<NS>(d, p, g);
}

// ---- SetTime_Step (solver_direct_reactive.cpp:5057-5298)
__device__ __forceinline__ void edge_lambda(const Dev& d, const Opts& o, int i, int j, double nx, double ny, double& lamI, double& lamV) {
  const double Area = sqrt(nx * nx + ny * ny);
  const double pvi = 0.0 + d.V[V_VX * d.np + i] * nx + d.V[(V_VX + 1) * d.np + i] * ny;
  const double pvj = 0.0 + d.V[V_VX * d.np + j] * nx + d.V[(V_VX + 1) * d.np + j] * ny;
  const double Mean_ProjVel = 0.5 * (pvi + pvj);
  const double Mean_SoundSpeed = 0.5 * (d.V[V_A * d.np + i] + d.V[V_A * d.np + j]);
  const double Mean_Density = 0.5 * (d.V[V_RHO * d.np + i] + d.V[V_RHO * d.np + j]);
  lamI = (fabs(Mean_ProjVel) + Mean_SoundSpeed) * Area;
  lamV = 0.0;
  if (!o.viscous) return;
  const double Mean_ThermalCond = 0.5 * (d.kap[i] + d.kap[j]);
  const double Mean_LaminarVisc = 0.5 * (d.mu[i] + d.mu[j]);
  const double Mean_CV = 0.5 * (d.Cp[i] / d.gam[i] + d.Cp[j] / d.gam[j]);
  double L1, L2;
  if (o.rans) {
    const double Mean_EddyVisc = 0.5 * (d.turb[i] + d.turb[j]);
    const double Gamma = d.gam[i];   // gamma of the edge's first node only (:5141)
    L1 = 4.0 / 3.0 * (Mean_LaminarVisc + Mean_EddyVisc);
    L2 = (1.0 + (o.prandtlLam / o.prandtlTurb) * (Mean_EddyVisc / Mean_LaminarVisc)) * (Gamma * Mean_LaminarVisc / o.prandtlLam);
  } else {
    L1 = 4.0 / 3.0 * Mean_LaminarVisc;
    L2 = Mean_ThermalCond / Mean_CV;
  }
  lamV = (L1 + L2) * Area * Area / Mean_Density;
}
__global__ void __launch_bounds__(128) k_timestep_edges(const __grid_constant__ Opts o, const Dev d) {
  const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (p >= d.nPointDomain) return;
  double li = 0.0, lv = 0.0;
  for (int k = d.incPtr[p]; k < d.incPtr[p + 1]; ++k) {
    const int code = d.incEdge[k];
    const int e = code >= 0 ? code : ~code;
    double a, b;
    edge_lambda(d, o, d.ei[e], d.ej[e], d.enx[e], d.eny[e], a, b);
    li += a; lv += b;
  }
  d.lamInv[p] = li; d.lamVisc[p] = lv;
}
__global__ void k_timestep_boundary(const __grid_constant__ Opts o, const Dev d, const MarkerDev m) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= m.nVertex) return;
  const int i = m.node[k];
  if (i >= d.nPointDomain) return;
  const double nx = m.nx[k], ny = m.ny[k];
  const double Area = sqrt(nx * nx + ny * ny);
  const double pv = 0.0 + d.V[V_VX * d.np + i] * nx + d.V[(V_VX + 1) * d.np + i] * ny;
  d.lamInv[i] += (fabs(pv) + d.V[V_A * d.np + i]) * Area;
  if (!o.viscous) return;
  const double mu = d.mu[i];
  double L1, L2;
  if (o.rans) {
    const double ev = d.turb[i];
    const double Gamma = d.gam[i];
    L1 = (4.0 / 3.0) * (mu + ev);
    L2 = (1.0 + (o.prandtlLam / o.prandtlTurb) * (ev / mu)) * (Gamma * mu / o.prandtlLam);
  } else {
    L1 = 4.0 / 3.0 * mu;
    L2 = d.kap[i] / (d.Cp[i] / d.gam[i]);
  }
  d.lamVisc[i] += (L1 + L2) * Area * Area / d.V[V_RHO * d.np + i];
}
__global__ void __launch_bounds__(128) k_timestep_finish(const __grid_constant__ Opts o, const Dev d) {
  const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (p >= d.nPointDomain) return;
  const double Vol = d.vol[p];
  double dt = 0.0;
  if (Vol > EPS) {
    dt = o.cfl * Vol / d.lamInv[p];
    if (o.viscous) dt = fmin(dt, o.cfl * 0.25 * Vol * Vol / d.lamVisc[p]);
    if (dt > o.maxDt && !o.timeStepping) dt = o.maxDt;   // time stepping: the global minimum is taken over the unclamped steps (:5227) and then overwrites them all
  }
  d.dt[p] = dt;
}
// UNSTEADY_SIMULATION = TIME_STEPPING (:5262-5272): minimum over the owned points with positive volume (the reference's loop skips the
// others), in two stages; then every owned point receives it.  A minimum does not depend on the order.
__global__ void __launch_bounds__(256) k_dt_min_partial(const Dev d, double* __restrict__ partial) {
  __shared__ double sh[256];
  const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  double v = 1.0e6;   // Global_Delta_Time starts at 1e6 (:5064)
  if (p < d.nPointDomain && d.vol[p] > EPS) v = fmin(v, d.dt[p]);
  sh[threadIdx.x] = v;
  __syncthreads();
  for (int w = 128; w > 0; w >>= 1) { if ((int)threadIdx.x < w) sh[threadIdx.x] = fmin(sh[threadIdx.x], sh[threadIdx.x + w]); __syncthreads(); }
  if (threadIdx.x == 0) partial[blockIdx.x] = sh[0];
}
__global__ void __launch_bounds__(256) k_dt_min_final(const double* __restrict__ partial, int n, double* __restrict__ out) {
  __shared__ double sh[256];
  double v = 1.0e6;
  for (int b = threadIdx.x; b < n; b += 256) v = fmin(v, partial[b]);
  sh[threadIdx.x] = v;
  __syncthreads();
  for (int w = 128; w > 0; w >>= 1) { if ((int)threadIdx.x < w) sh[threadIdx.x] = fmin(sh[threadIdx.x], sh[threadIdx.x + w]); __syncthreads(); }
  if (threadIdx.x == 0) out[0] = sh[0];
}
__global__ void k_dt_fill(const Dev d, const double* __restrict__ global, double fixed, int useFixed) {
  const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (p < d.nPointDomain) d.dt[p] = useFixed ? fixed : global[0];
}

// dual time stepping (:5274-5297, Euler :2141-2165).  k_dual_unst_dt: on pseudo-time iteration 0 with UNST_CFL_NUMBER set, the physical time
// step is UNST_CFL * (global minimum of the unclamped local steps) / CFL.  k_dual_dt_clamp: the local step, clamped to Max_DeltaTime as
// SetTime_Step does for every steady run, is then at most two thirds of the physical step.
__global__ void k_dual_unst_dt(const double* __restrict__ global, double unstCfl, double cfl, double* __restrict__ dtND) { dtND[0] = unstCfl * global[0] / cfl; }
__global__ void k_dual_dt_clamp(const Dev d, double maxDt, const double* __restrict__ dtND) {
  const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (p >= d.nPointDomain) return;
  double dt = d.dt[p];
  if (dt > maxDt) dt = maxDt;
  d.dt[p] = fmin((2.0 / 3.0) * dtND[0], dt);
}
// SetResidual_DualTime (:2172-2215): backward difference of the conserved variables in physical time, first or second order, times the
// dual volume, added to the residual of every owned point
__global__ void k_dual_time_residual(const Dev d, int nVar, int order, const double* __restrict__ dtND, const double* __restrict__ Un, const double* __restrict__ Un1) {
  const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (p >= d.nPointDomain) return;
  const double Vol = d.vol[p], TimeStep = dtND[0];
  for (int v = 0; v < nVar; ++v) {
    const size_t k = (size_t)v * d.np + p;
    const double up1 = d.U[k], un = Un[k];
    double r;
    if (order == 1) r = (up1 - un) * Vol / TimeStep;
    else r = (3.0 * up1 - 4.0 * un + Un1[k]) * Vol / (2.0 * TimeStep);
    d.R[k] += r;
  }
}

}  // namespace rx
