// rx_kernels2.cuh — edge-flux, owner-computes accumulation, source, boundary and update kernels.  See rx_kernels.cuh.
#pragma once
#include "rx_kernels.cuh"

namespace rx {

template <int NS>
__device__ __forceinline__ void load_V(const Dev& d, long long p, double* V) {
#pragma unroll
  for (int v = 0; v < 7 + NS; ++v) V[v] = d.V[v * d.np + p];
}
template <int NS>
__device__ __forceinline__ void load_visc_end(const Dev& d, const Opts& o, long long p, ViscEnd<NS>& E, bool withGrad) {
  load_V<NS>(d, p, E.V);
  if (withGrad) {
#pragma unroll
    for (int k = 0; k < (4 + NS) * 2; ++k) E.g[k] = d.grad[k * d.np + p];
  }
#pragma unroll
  for (int s = 0; s < NS; ++s) E.X[s] = d.X[s * d.np + p];
  E.mu = d.mu[p]; E.kap = d.kap[p];
  E.rD = d.rD + p; E.rDs = d.np;
  E.x[0] = d.x[p]; E.x[1] = d.y[p];
  if (o.rans) { E.muT = d.turb[p]; E.ke = d.turb[d.np + p]; E.gk[0] = d.turb[3 * d.np + p]; E.gk[1] = d.turb[4 * d.np + p]; }
  else { E.muT = 1.0; E.ke = 0.0; E.gk[0] = E.gk[1] = 0.0; }
}

// ---- Upwind_Residual (:2535-2785, 1st order): one thread per edge, AUSM flux to the edge array
constexpr int EDGE_BS = 64;   // threads per block of the boundary flux kernels (thread-private Stefan-Maxwell matrix in shared memory)
#ifndef RX_EF_MINB
#define RX_EF_MINB 6
#endif
// The kernel is bound by HBM (it reads both end states and writes 13 flux words per edge), so it is written for many resident warps:
// the interface quantities come from twelve scalars, then the rows stream through two registers each — the same operations as ausm<>.
template <int NS>
__global__ void __launch_bounds__(128, RX_EF_MINB) k_edge_flux(const __grid_constant__ ChemC<NS> cc, const __grid_constant__ Opts o, const Dev d) {
  const long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (e >= d.nEdge) return;
  const int i = d.ei[e], j = d.ej[e];
  const long long np = d.np, ne = d.ne;
  const double ui = d.V[V_VX * np + i], uj = d.V[V_VX * np + j], vi = d.V[(V_VX + 1) * np + i], vj = d.V[(V_VX + 1) * np + j];
  const double Hi = d.V[V_H * np + i], Hj = d.V[V_H * np + j];
  double Yi[NS], Yj[NS];
#pragma unroll
  for (int s = 0; s < NS; ++s) { Yi[s] = d.V[(V_Y + s) * np + i]; Yj[s] = d.V[(V_Y + s) * np + j]; }   // issued before the divisions of the interface
  const AusmIface f = ausm_iface(o.mInfty, d.V[V_RHO * np + i], ui, vi, d.V[V_P * np + i], d.V[V_A * np + i],
                                 d.V[V_RHO * np + j], uj, vj, d.V[V_P * np + j], d.V[V_A * np + j], d.enx[e], d.eny[e]);
  d.fluxC[U_RHO * ne + e] = ausm_row(f, 1.0, 1.0);
  double mx = ausm_row(f, ui, uj), my = ausm_row(f, vi, vj);
  d.fluxC[U_E * ne + e] = ausm_row(f, Hi, Hj);
#pragma unroll
  for (int s = 0; s < NS; ++s) d.fluxC[(U_S + s) * ne + e] = ausm_row(f, Yi[s], Yj[s]);
  mx += f.pLF * f.un0 * f.Area;
  my += f.pLF * f.un1 * f.Area;
  d.fluxC[U_MX * ne + e] = mx;
  d.fluxC[(U_MX + 1) * ne + e] = my;
}

// ---- SetPrimitive_Limiter (:1328-1523), Venkatakrishnan, owner-computes: one thread per point over its neighbours.
// The reference builds min/max of the neighbour differences and the minimum of the edge limiters in two edge loops;
// min and max do not depend on the order, so the per-point form gives the same bits.  Points whose neighbour list is
// incomplete (halos) receive their values through the limiter halo exchange, as in Set_MPI_Primitive_Limiter (:1756).
__global__ void __launch_bounds__(128) k_limiter(const __grid_constant__ Opts o, const Dev d) {
  const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (p >= d.nPoint) return;
  const int fld[4] = {V_T, V_VX, V_VX + 1, V_P};
  double vp[4], smax[4], smin[4], lim[4], g[4][2];
#pragma unroll
  for (int v = 0; v < 4; ++v) {
    vp[v] = d.V[fld[v] * d.np + p]; smax[v] = -EPS; smin[v] = EPS; lim[v] = 2.0;
    g[v][0] = d.grad[(2 * v) * d.np + p]; g[v][1] = d.grad[(2 * v + 1) * d.np + p];
  }
  const int k0 = d.nbPtr[p], k1 = d.nbPtr[p + 1];
  for (int k = k0; k < k1; ++k) {
    const int q = d.nbIdx[k];
#pragma unroll
    for (int v = 0; v < 4; ++v) {
      const double du = d.V[fld[v] * d.np + q] - vp[v];
      smin[v] = fmin(smin[v], du); smax[v] = fmax(smax[v], du);
    }
  }
  const double eps1 = o.limCoeff * o.refElemLength, eps2 = eps1 * eps1 * eps1;
  const double xp = d.x[p], yp = d.y[p];
  for (int k = k0; k < k1; ++k) {
    const int q = d.nbIdx[k];
    const double hx = 0.5 * (d.x[q] - xp), hy = 0.5 * (d.y[q] - yp);
#pragma unroll
    for (int v = 0; v < 4; ++v) {
      double dm = 0.0;
      dm += hx * g[v][0]; dm += hy * g[v][1];
      const double dp = dm > 0.0 ? smax[v] : smin[v];
      const double lv = (dp * dp + 2.0 * dp * dm + eps2) / (dp * dp + dp * dm + 2.0 * dm * dm + eps2);
      if (lv < lim[v]) lim[v] = lv;
    }
  }
#pragma unroll
  for (int v = 0; v < 4; ++v) d.lim[v * d.np + p] = lim[v];
}

// ---- SetPrimitive_Limiter, BARTH_JESPERSEN branch (:1380-1439) as coded.  The node-0 side of an edge takes the minimum; the
// node-1 side compares the solver's bool `limiter` (1.0) with the running value and ASSIGNS (:1425), so the value of a point
// depends on the order in which its edges are met.  Owner-computes over the incident-edge list in ascending edge order —
// the order of the reference's edge loop as one point sees it — gives the same sequence of updates, hence the same bits.
__global__ void __launch_bounds__(128) k_limiter_bj(const Dev d) {
  const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (p >= d.nPoint) return;
  const int fld[4] = {V_T, V_VX, V_VX + 1, V_P};
  double vp[4], smax[4], smin[4], lim[4], g[4][2];
#pragma unroll
  for (int v = 0; v < 4; ++v) {
    vp[v] = d.V[fld[v] * d.np + p]; smax[v] = -EPS; smin[v] = EPS; lim[v] = 2.0;
    g[v][0] = d.grad[(2 * v) * d.np + p]; g[v][1] = d.grad[(2 * v + 1) * d.np + p];
  }
  const int k0 = d.incPtr[p], k1 = d.incPtr[p + 1];
  for (int k = k0; k < k1; ++k) {
    const int code = d.incEdge[k];
    const int q = code >= 0 ? d.ej[code] : d.ei[~code];
#pragma unroll
    for (int v = 0; v < 4; ++v) {
      const double other = d.V[fld[v] * d.np + q];
      const double du = code >= 0 ? other - vp[v] : -(vp[v] - other);   // node 0 sees du, node 1 sees -du with du = V_j - V_i
      smin[v] = fmin(smin[v], du); smax[v] = fmax(smax[v], du);
    }
  }
  const double xp = d.x[p], yp = d.y[p];
  for (int k = k0; k < k1; ++k) {
    const int code = d.incEdge[k];
    const bool first = code >= 0;
    const int q = first ? d.ej[code] : d.ei[~code];
    const double hx = 0.5 * (d.x[q] - xp), hy = 0.5 * (d.y[q] - yp);
#pragma unroll
    for (int v = 0; v < 4; ++v) {
      double dm = 0.0;
      dm += hx * g[v][0]; dm += hy * g[v][1];
      double lv;
      if (dm < EPS) lv = 2.0;
      else lv = (dm > EPS ? smax[v] : smin[v]) / dm;
      if (first) { if (lv < lim[v]) lim[v] = lv; }
      else if (1.0 < lim[v]) lim[v] = lv;
    }
  }
#pragma unroll
  for (int v = 0; v < 4; ++v) { const double y = lim[v]; d.lim[v * d.np + p] = (y * y + 2.0 * y) / (y * y + y + 2.0); }
}

// ---- 2nd-order Upwind_Residual (:2554-2681): MUSCL reconstruction of [T,u,v,P] at the edge mid point, rebuild of rho, h, a
template <int NS>
__device__ __forceinline__ void reconstruct_state(const ChemC<NS>& cc, const Opts& o, const Dev& d, long long p, double hx, double hy,
                                                  bool useOtherP, double otherP, double* Vr, double& Precon) {
  double V[7 + NS];
  load_V<NS>(d, p, V);
  const int fld[4] = {V_T, V_VX, V_VX + 1, V_P};
  double rec[4];
#pragma unroll
  for (int v = 0; v < 4; ++v) {
    double proj = 0.0;
    proj += hx * d.grad[(2 * v) * d.np + p]; proj += hy * d.grad[(2 * v + 1) * d.np + p];
    rec[v] = V[fld[v]];
    if (o.limiter) rec[v] += d.lim[v * d.np + p] * proj; else rec[v] += proj;
  }
  Precon = rec[3];
  bool nonPhys = !(rec[0] > EPS);
  if (!nonPhys) nonPhys = !((useOtherP ? otherP : rec[3]) > EPS);   // the j side tests the pressure of the i side (:2612)
  if (nonPhys) {
#pragma unroll
    for (int v = 0; v < 7 + NS; ++v) Vr[v] = V[v];
    return;
  }
  const double* Ys = V + V_Y;
#pragma unroll
  for (int s = 0; s < NS; ++s) Vr[V_Y + s] = Ys[s];
  Vr[V_T] = rec[0]; Vr[V_VX] = rec[1]; Vr[V_VX + 1] = rec[2]; Vr[V_P] = rec[3];
  double rho = rec[3] / (rec[0] * rgas(cc, Ys));
  rho *= 1.0;
  Vr[V_RHO] = rho;
  bool ok = true;
  Vr[V_H] = enthalpy(cc, rec[0], Ys, ok) / 1.0;
  double sq = 0.0;
  sq += rec[1] * rec[1]; sq += rec[2] * rec[2];
  Vr[V_H] += 0.5 * sq;
  const double g = frozen_gamma(cc, rec[0], Ys);
  Vr[V_A] = sqrt(g * rec[3] / rho);
}
template <int NS>
__global__ void __launch_bounds__(64) k_edge_conv2(const __grid_constant__ ChemC<NS> cc, const __grid_constant__ Opts o, const Dev d) {
  const long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (e >= d.nEdge) return;
  constexpr int NV = 4 + NS;
  const int i = d.ei[e], j = d.ej[e];
  const double hx = 0.5 * (d.x[j] - d.x[i]), hy = 0.5 * (d.y[j] - d.y[i]);
  double Vi[7 + NS], Vj[7 + NS], res[NV], Pi, Pj;
  reconstruct_state<NS>(cc, o, d, i, hx, hy, false, 0.0, Vi, Pi);
  reconstruct_state<NS>(cc, o, d, j, -hx, -hy, true, Pi, Vj, Pj);
  ausm<NS>(o.mInfty, Vi, Vj, d.enx[e], d.eny[e], res);
#pragma unroll
  for (int v = 0; v < NV; ++v) d.fluxC[v * d.ne + e] = res[v];
}

#ifndef RX_SRC_MINB
#define RX_SRC_MINB 5
#endif
// ---- Source_Residual (:2792-2874) + CSourceReactive::ComputeChemistry (numerics_direct_reactive.cpp:1728-1825), per owned point
template <int NS>
__device__ __forceinline__ void point_source(const ChemC<NS>& cc, const Opts& o, const Dev& d, long long p, double* src) {
  double Y[NS];
#pragma unroll
  for (int s = 0; s < NS; ++s) Y[s] = d.V[(V_Y + s) * d.np + p];
  const double omT = o.rans ? d.turb[2 * d.np + p] : 0.0;
  mass_production<NS>(cc, o.rans != 0, o.cMu, o.pasrLb, d.V[V_T * d.np + p], d.V[V_RHO * d.np + p], Y, omT, src);
  const double Volume = d.vol[p];
#pragma unroll
  for (int s = 0; s < NS; ++s) src[s] = src[s] * (-Volume / (1.0 / 1.0));
}
template <int NS>
__global__ void __launch_bounds__(128, RX_SRC_MINB) k_source(const __grid_constant__ ChemC<NS> cc, const __grid_constant__ Opts o, const Dev d) {
  const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (p >= d.nPointDomain) return;
  double src[NS];
  point_source<NS>(cc, o, d, p, src);
  bool nanS = false;
#pragma unroll
  for (int s = 0; s < NS; ++s) { nanS = nanS || src[s] != src[s]; d.R[(U_S + s) * d.np + p] += src[s]; }
  if (nanS) atomicAdd(d.counters + 3, 1ULL);
}
// stand-alone form over independent points, SoA inputs Y[s*n+p], res[s*n+p]
template <int NS>
__global__ void __launch_bounds__(128, RX_SRC_MINB) k_source_points(const __grid_constant__ ChemC<NS> cc, int rans, double cMu, double pasrLb, long long n,
                                                       const double* __restrict__ T, const double* __restrict__ rho, const double* __restrict__ Y,
                                                       const double* __restrict__ omT, const double* __restrict__ vol, double* __restrict__ res) {
  const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (p >= n) return;
  double y[NS], om[NS];
#pragma unroll
  for (int s = 0; s < NS; ++s) y[s] = Y[s * n + p];
  mass_production<NS>(cc, rans != 0, cMu, pasrLb, T[p], rho[p], y, omT ? omT[p] : 0.0, om);
  const double Volume = vol[p];
#pragma unroll
  for (int s = 0; s < NS; ++s) res[s * n + p] = om[s] * (-Volume / (1.0 / 1.0));
}

// ---- owner-computes accumulation of edge fluxes into LinSysRes.  MODE: 1 conv (+i, -j), 2 visc (-i, +j), 3 conv, visc, then source
#ifndef RX_ACC_MINB
#define RX_ACC_MINB 4
#endif
template <int NS, int MODE>
__global__ void __launch_bounds__(128, RX_ACC_MINB) k_accumulate(const __grid_constant__ ChemC<NS> cc, const __grid_constant__ Opts o, const Dev d) {
  const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (p >= d.nPoint) return;
  constexpr int NV = 4 + NS;
  double r[NV];
#pragma unroll
  for (int v = 0; v < NV; ++v) r[v] = d.R[v * d.np + p];
  const int k0 = d.incPtr[p], k1 = d.incPtr[p + 1];
#ifdef RX_ACC_PREFETCH
  // the flux words of the next incident edge are in flight while those of the current one are added (same order of additions)
  auto sweep = [&](const double* __restrict__ flux, bool plusFirst) {
    if (k0 >= k1) return;
    double f[NV], g[NV];
    int code = d.incEdge[k0];
    {
      const long long e = code >= 0 ? code : ~code;
#pragma unroll
      for (int v = 0; v < NV; ++v) f[v] = flux[v * d.ne + e];
    }
    for (int k = k0; k < k1; ++k) {
      const bool add = (code >= 0) == plusFirst;
      int codeN = code;
      if (k + 1 < k1) {
        codeN = d.incEdge[k + 1];
        const long long e = codeN >= 0 ? codeN : ~codeN;
#pragma unroll
        for (int v = 0; v < NV; ++v) g[v] = flux[v * d.ne + e];
      }
#pragma unroll
      for (int v = 0; v < NV; ++v) { r[v] = add ? r[v] + f[v] : r[v] - f[v]; f[v] = g[v]; }
      code = codeN;
    }
  };
  if (MODE & 1) sweep(d.fluxC, true);
  if (MODE & 2) sweep(d.fluxV, false);
  bool nanC = false, nanV = false;
#else
  bool nanC = false, nanV = false;   // the reference tests every edge residual before adding it (:2744, :5361) and throws
  if (MODE & 1) {
    for (int k = k0; k < k1; ++k) {
      const int code = d.incEdge[k];
      const bool first = code >= 0;
      const long long e = first ? code : ~code;
#pragma unroll
      for (int v = 0; v < NV; ++v) { const double f = d.fluxC[v * d.ne + e]; nanC = nanC || f != f; r[v] = first ? r[v] + f : r[v] - f; }
    }
  }
  if (MODE & 2) {
    for (int k = k0; k < k1; ++k) {
      const int code = d.incEdge[k];
      const bool first = code >= 0;
      const long long e = first ? code : ~code;
#pragma unroll
      for (int v = 0; v < NV; ++v) { const double f = d.fluxV[v * d.ne + e]; nanV = nanV || f != f; r[v] = first ? r[v] - f : r[v] + f; }
    }
  }
#endif
  if (MODE == 3 && p < d.nPointDomain) {
    double src[NS];
    point_source<NS>(cc, o, d, p, src);
    bool nanS = false;
#pragma unroll
    for (int s = 0; s < NS; ++s) { nanS = nanS || src[s] != src[s]; r[U_S + s] += src[s]; }
    if (nanS) atomicAdd(d.counters + 3, 1ULL);   // "NaN found in the source residual" (:2861)
  }
  if (nanC) atomicAdd(d.counters + 1, 1ULL);     // "NaN found in the upwind residual" (:2772)
  if (nanV) atomicAdd(d.counters + 2, 1ULL);     // "NaN found in the viscous residual" (:5384)
#pragma unroll
  for (int v = 0; v < NV; ++v) d.R[v * d.np + p] = r[v];
}

// ---- boundary conditions: one thread per vertex of one marker (markers are issued in marker order on the stream)
template <int NS>
__device__ __forceinline__ void boundary_fluxes(const ChemC<NS>& cc, const Opts& o, const Dev& d, int i, const double* Vg, double nx, double ny) {
  constexpr int NV = 4 + NS;
  ViscEnd<NS> I, J;
  load_visc_end<NS>(d, o, i, I, o.viscous != 0);
  double res[NV];
  ausm<NS>(o.mInfty, I.V, Vg, nx, ny, res);
  double r[NV];
#pragma unroll
  for (int v = 0; v < NV; ++v) r[v] = d.R[v * d.np + i] + res[v];
  if (o.viscous) {
    J = I;   // gradient, transport, D_ij and turbulence of iPoint on both sides (:3583-3665, :4001-4110)
#pragma unroll
    for (int v = 0; v < 7 + NS; ++v) J.V[v] = Vg[v];
    extern __shared__ double smat[];
    const double sigma_k = o.rans ? d.turb[5 * d.np + i] : 1.0;
    if (o.rans) viscous_flux<NS, false, EDGE_BS, true>(cc, o, I, J, nx, ny, sigma_k, smat + threadIdx.x, res);
    else viscous_flux<NS, false, EDGE_BS, false>(cc, o, I, J, nx, ny, sigma_k, smat + threadIdx.x, res);
#pragma unroll
    for (int v = 0; v < NV; ++v) r[v] = r[v] - res[v];
  }
#pragma unroll
  for (int v = 0; v < NV; ++v) d.R[v * d.np + i] = r[v];
}
// ---- Viscous_Residual (:5283-5397) for the species counts the pair kernel is not laid out for: one thread per edge, the serial
// form of the same arithmetic (viscous_flux<NS, EDGE = true>: CAvgGradReactive_Flow::ComputeResidual with the edge correction)
template <int NS, bool RANS>
__global__ void __launch_bounds__(EDGE_BS) k_edge_visc_generic(const __grid_constant__ ChemC<NS> cc, const __grid_constant__ Opts o, const Dev d) {
  const long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (e >= d.nEdge) return;
  constexpr int NV = 4 + NS;
  const int i = d.ei[e], j = d.ej[e];
  ViscEnd<NS> I, J;
  load_visc_end<NS>(d, o, i, I, true);
  load_visc_end<NS>(d, o, j, J, true);
  if (o.viscLimiter) {   // VISCOUS_LIMITER_FLOW (numerics_direct_reactive.cpp:1482-1505): gradients of T, u, v times the limiter of their variable
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const double li = d.lim[r * d.np + i], lj = d.lim[r * d.np + j];
      const int g = (r == 0 ? G_T : G_VX + r - 1) * 2;
      I.g[g] = I.g[g] * li; I.g[g + 1] = I.g[g + 1] * li; J.g[g] = J.g[g] * lj; J.g[g + 1] = J.g[g + 1] * lj;
    }
  }
  extern __shared__ double smat[];
  const double sigma_k = RANS ? d.turb[5 * d.np + i] : 1.0;   // sigma_k of the first node (:5345)
  double res[NV];
  viscous_flux<NS, true, EDGE_BS, RANS>(cc, o, I, J, d.enx[e], d.eny[e], sigma_k, smat + threadIdx.x, res);
#pragma unroll
  for (int v = 0; v < NV; ++v) d.fluxV[v * d.ne + e] = res[v];
  if (d.eKey) d.eKey[e] = 0;
}

// BC_Inlet (solver_direct_reactive.cpp:3226-3674): TEMPERATURE_IMPOSE :3467-3500 (marker kind 1), TOTAL_CONDITIONS :3283-3409 (6), MASS_FLOW :3411-3465 (7)
template <int NS>
__global__ void __launch_bounds__(64) k_bc_inlet(const __grid_constant__ ChemC<NS> cc, const __grid_constant__ Opts o, const Dev d, const MarkerDev m) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= m.nVertex) return;
  const int i = m.node[k];
  if (i >= d.nPointDomain) return;
  const double nx = -m.nx[k], ny = -m.ny[k];
  double Vg[7 + NS], Ys[NS];
#pragma unroll
  for (int s = 0; s < NS; ++s) Ys[s] = m.Ys[s];
  bool ok = true;
  if (m.kind == 6) {
    // TOTAL_CONDITIONS (:3283-3409): static temperature from the outgoing Riemann invariant and the total enthalpy — secant from
    // (T_tot, T_tot + 1), at most 15 steps to 1e-9, else bisection on [300, T_tot], 100 steps to 1e-6 — then the isentropic state
    const double Area = sqrt(0.0 + m.nx[k] * m.nx[k] + m.ny[k] * m.ny[k]);
    const double un0 = nx / Area, un1 = ny / Area;
    const double Ttot = m.T / 1.0, Ptot = m.Vmag / 1.0;
    double Vn = 0.0;
    Vn += d.V[V_VX * d.np + i] * un0; Vn += d.V[(V_VX + 1) * d.np + i] * un1;
    const double SoundSpeed = d.V[V_A * d.np + i];
    double Gamma = d.gam[i];
    const double Gamma_Tot = frozen_gamma(cc, Ttot, Ys);
    Gamma = 2.0 / (1.0 / Gamma + 1.0 / Gamma_Tot);
    const double Gamma_Minus_One = Gamma - 1.0;
    const double Riemann = Vn + 2.0 * SoundSpeed / Gamma_Minus_One;
    double Tot_Enthalpy = enthalpy(cc, Ttot, Ys, ok);
    double alpha = 0.0;
    alpha = alpha + un0 * m.dir0; alpha = alpha + un1 * m.dir1;
    const double Rg = rgas(cc, Ys) / 1.0;
    auto f = [&](double T) {
      const double hb = enthalpy(cc, T, Ys, ok);
      const double cb = sqrt(Gamma * Rg * T);
      const double Vb = (Riemann - 2.0 * cb / Gamma_Minus_One) / alpha;
      return hb + 0.5 * Vb * Vb;
    };
    double Told = Ttot + 1.0, Tcurr = Ttot, Tnew;
    bool NRconvg = false;
#pragma unroll 1
    for (int it = 0; it < 15; ++it) {
      const double tmp = f(Tcurr);
      const double F = tmp - Tot_Enthalpy;
      const double dF = tmp - f(Told);
      Tnew = Tcurr - F * (Tcurr - Told) / dF;
      if (fabs(Tnew - Tcurr) < 1.0e-9) { NRconvg = true; break; }
      Told = Tcurr; Tcurr = Tnew;
    }
    if (NRconvg) Vg[V_T] = Tcurr;
    else {
      double Ta = 300.0 / 1.0, Tb = Ttot;
      bool Bconvg = false;
#pragma unroll 1
      for (int it = 0; it < 100; ++it) {
        Tcurr = (Ta + Tb) / 2.0;
        const double F = f(Tcurr) - Tot_Enthalpy;
        if (fabs(F) < 1.0e-6) { Vg[V_T] = Tcurr; Bconvg = true; break; }
        if (F > 0.0) Ta = Tcurr; else Tb = Tcurr;
      }
      if (!Bconvg) { Vg[V_T] = CUDART_NAN; atomicAdd(d.counters + 4, 1ULL); }   // the reference throws here (:3377); counted with the failed temperature solves
    }
    if (o.rans) Tot_Enthalpy += o.tkeInf;
    Vg[V_H] = Tot_Enthalpy;
    const double rho_tot = Ptot / (Rg * Ttot);
    Vg[V_RHO] = rho_tot * rx_pow(Vg[V_T] / Ttot, 1.0 / Gamma_Minus_One);
    Vg[V_P] = Vg[V_RHO] * Rg * Vg[V_T];
    Vg[V_A] = sqrt(Vg[V_T] * Gamma * Rg);
    const double Vel_Mag = fabs((Riemann - 2.0 * Vg[V_A] / Gamma_Minus_One) / alpha);
    Vg[V_VX] = Vel_Mag * m.dir0; Vg[V_VX + 1] = Vel_Mag * m.dir1;
  } else if (m.kind == 7) {
    // MASS_FLOW (:3411-3465): density and velocity imposed, the speed of sound from the outgoing Riemann invariant
    const double Area = sqrt(0.0 + m.nx[k] * m.nx[k] + m.ny[k] * m.ny[k]);
    const double un0 = nx / Area, un1 = ny / Area;
    const double Density = m.T / 1.0, Vel_Mag = m.Vmag / 1.0;
    double SoundSpeed = d.V[V_A * d.np + i];
    const double Gamma = d.gam[i];
    const double Gamma_Minus_One = Gamma - 1.0;
    double Vn = 0.0;
    Vn = Vn + d.V[V_VX * d.np + i] * un0; Vn = Vn + d.V[(V_VX + 1) * d.np + i] * un1;
    const double Riemann = Vn + 2.0 * SoundSpeed / Gamma_Minus_One;
    double alpha = 0.0;
    alpha = alpha + un0 * m.dir0; alpha = alpha + un1 * m.dir1;
    SoundSpeed = Riemann - Vel_Mag * alpha;
    SoundSpeed = fmax(0.0, 0.5 * Gamma_Minus_One * SoundSpeed);
    const double Pressure = SoundSpeed * SoundSpeed * Density / Gamma;
    Vg[V_T] = Pressure / (Density * rgas(cc, Ys)) * 1.0;
    Vg[V_VX] = Vel_Mag * m.dir0; Vg[V_VX + 1] = Vel_Mag * m.dir1;
    Vg[V_P] = Pressure;
    Vg[V_RHO] = Density;
    double Aux_Ent = enthalpy(cc, Vg[V_T] * 1.0, Ys, ok) / 1.0;
    if (o.rans) Aux_Ent += o.tkeInf;
    Vg[V_H] = Aux_Ent;
    Vg[V_H] += 0.5 * Vel_Mag * Vel_Mag;
    Vg[V_A] = SoundSpeed;
  } else {
    Vg[V_T] = m.T;
    Vg[V_VX] = m.Vmag * m.dir0; Vg[V_VX + 1] = m.Vmag * m.dir1;
    Vg[V_P] = d.V[V_P * d.np + i];
    const double Rg = rgas(cc, Ys);
    Vg[V_RHO] = Vg[V_P] / (Vg[V_T] * Rg) * 1.0;
    Vg[V_H] = enthalpy(cc, m.T, Ys, ok) / 1.0 + (o.rans ? 1 : 0) * o.tkeInf;
    Vg[V_H] += 0.5 * m.Vmag * m.Vmag;
    const double g = frozen_gamma(cc, m.T, Ys);
    Vg[V_A] = sqrt(g * Rg * m.T);
  }
#pragma unroll
  for (int s = 0; s < NS; ++s) Vg[V_Y + s] = Ys[s];
#pragma unroll
  for (int v = 0; v < 7 + NS; ++v) m.ghost[(size_t)v * m.nVertex + k] = Vg[v];   // CharacPrimVar (:3274), read by the turbulence BCs
  boundary_fluxes<NS>(cc, o, d, i, Vg, nx, ny);
}
// BC_Supersonic_Inlet (:2998-3214; kind 8: every primitive imposed, no turbulent kinetic energy in the enthalpy) and
// BC_Supersonic_Outlet (:3681-3807; kind 9: the outer state is the interior one)
template <int NS>
__global__ void __launch_bounds__(64) k_bc_supersonic(const __grid_constant__ ChemC<NS> cc, const __grid_constant__ Opts o, const Dev d, const MarkerDev m) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= m.nVertex) return;
  const int i = m.node[k];
  if (i >= d.nPointDomain) return;
  const double nx = -m.nx[k], ny = -m.ny[k];
  double Vb[7 + NS];
  if (m.kind == 8) {
    double Ys[NS];
#pragma unroll
    for (int s = 0; s < NS; ++s) Ys[s] = m.Ys[s];
    bool ok = true;
    const double Temperature = m.T, Pressure = m.P;
    const double Rg = rgas(cc, Ys);
    const double Density = Pressure / (Temperature * Rg) * 1.0;
    const double Enthalpy = enthalpy(cc, Temperature, Ys, ok);
    const double SoundSpeed = sqrt(frozen_gamma(cc, Temperature, Ys) * rgas(cc, Ys) * Temperature);
    Vb[V_T] = Temperature / 1.0;
    Vb[V_VX] = m.dir0 / 1.0; Vb[V_VX + 1] = m.dir1 / 1.0;
    Vb[V_P] = Pressure / 1.0; Vb[V_RHO] = Density / 1.0;
    double Velocity2 = 0.0;
    Velocity2 = Velocity2 + Vb[V_VX] * Vb[V_VX]; Velocity2 = Velocity2 + Vb[V_VX + 1] * Vb[V_VX + 1];
    Vb[V_H] = Enthalpy / 1.0 + 0.5 * Velocity2;
    Vb[V_A] = SoundSpeed / 1.0;
#pragma unroll
    for (int s = 0; s < NS; ++s) Vb[V_Y + s] = Ys[s];
  } else {
    load_V<NS>(d, i, Vb);
  }
#pragma unroll
  for (int v = 0; v < 7 + NS; ++v) m.ghost[(size_t)v * m.nVertex + k] = Vb[v];   // CharacPrimVar
  boundary_fluxes<NS>(cc, o, d, i, Vb, nx, ny);
}
// BC_Outlet (solver_direct_reactive.cpp:3808-4123)
template <int NS>
__global__ void __launch_bounds__(64) k_bc_outlet(const __grid_constant__ ChemC<NS> cc, const __grid_constant__ Opts o, const Dev d, const MarkerDev m) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= m.nVertex) return;
  const int i = m.node[k];
  if (i >= d.nPointDomain) return;
  double Nx = m.nx[k], Ny = m.ny[k];
  const double Area = sqrt(Nx * Nx + Ny * Ny);
  Nx = -Nx; Ny = -Ny;
  const double un0 = Nx / Area, un1 = Ny / Area;
  double Vd[7 + NS], Vo[7 + NS];
  load_V<NS>(d, i, Vd);
  const double Density = Vd[V_RHO];
  double Vel0 = Vd[V_VX], Vel1 = Vd[V_VX + 1];
  double Velocity2 = 0.0;
  Velocity2 += Vel0 * Vel0; Velocity2 += Vel1 * Vel1;
  const double Pressure = Vd[V_P];
  const double Gamma = d.gam[i];
  double SoundSpeed = sqrt(Gamma * Pressure / Density);
  const double Mach_Exit = sqrt(Velocity2) / SoundSpeed;
  if (Mach_Exit >= 1.0) {
#pragma unroll
    for (int v = 0; v < 7 + NS; ++v) Vo[v] = Vd[v];
  } else {
    const double Entropy = Pressure * rx_pow(1.0 / Density, Gamma);
    const double Vn = 0.0 + Vel0 * un0 + Vel1 * un1;
    const double Gamma_Minus_One = Gamma - 1.0;
    const double Riemann = Vn + 2.0 * SoundSpeed / Gamma_Minus_One;
    const double P_Exit = m.P / 1.0;
    Vo[V_P] = P_Exit;
    Vo[V_RHO] = rx_pow(P_Exit / Entropy, 1.0 / Gamma);
    SoundSpeed = sqrt(Gamma * P_Exit / Vo[V_RHO]);
    const double Vn_Exit = Riemann - 2.0 * SoundSpeed / Gamma_Minus_One;
    Velocity2 = 0.0;
    Vel0 += (Vn_Exit - Vn) * un0; Velocity2 += Vel0 * Vel0; Vo[V_VX] = Vel0;
    Vel1 += (Vn_Exit - Vn) * un1; Velocity2 += Vel1 * Vel1; Vo[V_VX + 1] = Vel1;
    const double* Ys = Vd + V_Y;
    Vo[V_T] = P_Exit / (Vo[V_RHO] * rgas(cc, Ys)) * 1.0;
    bool ok = true;
    Vo[V_H] = enthalpy(cc, Vo[V_T], Ys, ok) / 1.0 + (o.rans ? 1 : 0) * o.tkeInf;
    Vo[V_H] += 0.5 * Velocity2;
    Vo[V_A] = SoundSpeed;
#pragma unroll
    for (int s = 0; s < NS; ++s) Vo[V_Y + s] = Ys[s];
  }
#pragma unroll
  for (int v = 0; v < 7 + NS; ++v) m.ghost[(size_t)v * m.nVertex + k] = Vo[v];   // CharacPrimVar (:3844)
  boundary_fluxes<NS>(cc, o, d, i, Vo, Nx, Ny);
}
// BC_Isothermal_Wall (solver_direct_reactive.cpp:5393-5711), explicit, static wall
template <int NS>
__global__ void __launch_bounds__(64) k_bc_isothermal(const __grid_constant__ ChemC<NS> cc, const __grid_constant__ Opts o, const Dev d, const MarkerDev m) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= m.nVertex) return;
  const int i = m.node[k];
  if (i >= d.nPointDomain) return;
  const double nx = m.nx[k], ny = m.ny[k];
  const double Area = sqrt(nx * nx + ny * ny);
  const int pn = m.nn[k];
  double dij = 0.0;
  dij += (d.x[pn] - d.x[i]) * (d.x[pn] - d.x[i]);
  dij += (d.y[pn] - d.y[i]) * (d.y[pn] - d.y[i]);
  dij = sqrt(dij);
  const double rho = d.V[V_RHO * d.np + i];
  d.Uold[U_MX * d.np + i] = 0.0 * rho;          // SetVelocity_Old (variable_direct_reactive.cpp:950-957)
  d.Uold[(U_MX + 1) * d.np + i] = 0.0 * rho;
  d.R[U_MX * d.np + i] = 0.0;                   // LinSysRes.SetBlock_Zero(iPoint, RHOVX+iDim)
  d.R[(U_MX + 1) * d.np + i] = 0.0;
  const double Twall = m.Twall;
  const double Tj = d.V[V_T * d.np + pn];
  const double ktr = d.kap[i];
  double turb_closure = 0.0;
  if (o.rans) {
    const double eddy_v = d.turb[i];
    const SplPos sp = spl_pos(cc, Twall);
#pragma unroll
    for (int s = 0; s < NS; ++s) {
      const double cp = spl_eval(cc, P_CP, s, sp) / cc.M[s];
      const double aux_ys = d.U[(U_S + s) * d.np + i];   // partial density, not mass fraction (:5514-5517)
      turb_closure += eddy_v / o.prandtlTurb * cp * aux_ys * (Twall - Tj) / dij;
    }
  }
  const double dTdn = +(Twall - Tj) / dij;
  d.R[U_E * d.np + i] -= ktr * dTdn * Area + turb_closure * Area;
}

// BC_HeatFlux_Wall (solver_direct_reactive.cpp:5717-5918), explicit, static wall: no-slip + imposed heat flux
__global__ void __launch_bounds__(64) k_bc_heatflux(const Dev d, const MarkerDev m) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= m.nVertex) return;
  const int i = m.node[k];
  if (i >= d.nPointDomain) return;
  const double nx = m.nx[k], ny = m.ny[k];
  const double Area = sqrt(0.0 + nx * nx + ny * ny);
  const double rho = d.V[V_RHO * d.np + i];
  d.Uold[U_MX * d.np + i] = 0.0 * rho;          // SetVelocity_Old(0)
  d.Uold[(U_MX + 1) * d.np + i] = 0.0 * rho;
  d.R[U_MX * d.np + i] = 0.0;                   // LinSysRes.SetBlock_Zero(iPoint, RHOVX+iDim)
  d.R[(U_MX + 1) * d.np + i] = 0.0;
  d.R[U_E * d.np + i] -= m.heatFlux * Area;     // Res_Visc[rhoE] = Wall_HeatFlux*Area (:5775), SubtractBlock
}
// BC_Euler_Wall (solver_direct_reactive.cpp:2881-2997), explicit, static wall: pressure (+ 2/3 rho k) on the momentum rows
__global__ void __launch_bounds__(64) k_bc_euler_wall(const __grid_constant__ Opts o, const Dev d, const MarkerDev m) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= m.nVertex) return;
  const int i = m.node[k];
  if (i >= d.nPointDomain) return;
  const double nx = m.nx[k], ny = m.ny[k];
  const double Area = sqrt(0.0 + nx * nx + ny * ny);
  const double un0 = -nx / Area, un1 = -ny / Area;
  const double P = d.V[V_P * d.np + i], rho = d.V[V_RHO * d.np + i];
  const double ke = o.rans ? d.turb[1 * d.np + i] : 0.0;
  d.R[U_MX * d.np + i] += P * un0 * Area + 2.0 / 3.0 * rho * ke * un0 * Area;
  d.R[(U_MX + 1) * d.np + i] += P * un1 * Area + 2.0 / 3.0 * rho * ke * un1 * Area;
}

// ---- ExplicitEuler/RK_Iteration (:2414-2493): update + per-block partial sums of Res^2 (deterministic two-stage reduction)
// MAXR: also the largest |Res| of every variable and the point that holds it (CSolver::AddRes_Max, solver_structure.inl: the
// FIRST point in point order that attains the maximum) as per-block partials for k_resmax_finish.
template <int NS, bool MAXR>
__global__ void __launch_bounds__(128) k_update(const Dev d, double alpha, double* __restrict__ partial, double* __restrict__ pmax, int* __restrict__ pidx) {
  const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  constexpr int NV = 4 + NS;
  __shared__ double sh[4][NV];
  __shared__ double shm[MAXR ? 4 : 1][NV];
  __shared__ int shi[MAXR ? 4 : 1][NV];
  double sq[NV], ab[NV];
#pragma unroll
  for (int v = 0; v < NV; ++v) { sq[v] = 0.0; ab[v] = 0.0; }
  if (p < d.nPointDomain) {
    const double Vol = d.vol[p];
    const double Delta = Vol > EPS ? d.dt[p] / Vol : 0.0;
#pragma unroll
    for (int v = 0; v < NV; ++v) {
      const double Res = d.R[v * d.np + p] + 0.0;
      const double lo = (v == U_MX || v == U_MX + 1 || v == U_E) ? -1.0 / EPS : 0.0, hi = 1.0 / EPS;
      // CVariable::AddClippedSolution (variable_structure.cpp:207-212): std::min(std::max(x, lo), hi) — a NaN x stays NaN
      const double x = d.Uold[v * d.np + p] + (-Res * Delta * alpha);
      const double xl = (x < lo) ? lo : x;
      d.U[v * d.np + p] = (hi < xl) ? hi : xl;
      sq[v] = Res * Res;
      ab[v] = fabs(Res);
    }
  }
  if (MAXR) {
#pragma unroll
    for (int v = 0; v < NV; ++v) {
      double m = ab[v];
      int ip = (int)(p < d.nPointDomain ? p : 0x7fffffff);
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) {   // a higher lane is a later point: it wins only if strictly larger
        const double mo = __shfl_down_sync(0xffffffffu, m, off);
        const int io = __shfl_down_sync(0xffffffffu, ip, off);
        if (mo > m) { m = mo; ip = io; }
      }
      if ((threadIdx.x & 31) == 0) { shm[threadIdx.x >> 5][v] = m; shi[threadIdx.x >> 5][v] = ip; }
    }
  }
#pragma unroll
  for (int v = 0; v < NV; ++v) {
    double s = sq[v];
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) s += __shfl_down_sync(0xffffffffu, s, off);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5][v] = s;
  }
  __syncthreads();
  if (threadIdx.x < NV) {
    const int v = threadIdx.x;
    partial[(size_t)blockIdx.x * NV + v] = ((sh[0][v] + sh[1][v]) + (sh[2][v] + sh[3][v]));
    if (MAXR) {
      double m = shm[0][v]; int ip = shi[0][v];
#pragma unroll
      for (int w = 1; w < 4; ++w) if (shm[w][v] > m) { m = shm[w][v]; ip = shi[w][v]; }
      pmax[(size_t)blockIdx.x * NV + v] = m; pidx[(size_t)blockIdx.x * NV + v] = ip;
    }
  }
}
// largest residual per variable over the block partials: larger value wins, equal values go to the lower point index
__global__ void __launch_bounds__(256) k_resmax_finish(const double* __restrict__ pmax, const int* __restrict__ pidx, int nBlocks, int nv,
                                                       double* __restrict__ resMax, int* __restrict__ resMaxPoint) {
  const int v = blockIdx.x;
  __shared__ double sm[256];
  __shared__ int si[256];
  double m = 0.0; int ip = 0;   // SetRes_Max(iVar, 0.0, 0) before the loop (:2427)
  for (int b = threadIdx.x; b < nBlocks; b += 256) {
    const double mo = pmax[(size_t)b * nv + v]; const int io = pidx[(size_t)b * nv + v];
    if (mo > m || (mo == m && mo > 0.0 && io < ip)) { m = mo; ip = io; }
  }
  sm[threadIdx.x] = m; si[threadIdx.x] = ip;
  __syncthreads();
  for (int w = 128; w > 0; w >>= 1) {
    if (threadIdx.x < w) {
      const double mo = sm[threadIdx.x + w]; const int io = si[threadIdx.x + w];
      if (mo > sm[threadIdx.x] || (mo == sm[threadIdx.x] && mo > 0.0 && io < si[threadIdx.x])) { sm[threadIdx.x] = mo; si[threadIdx.x] = io; }
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) { resMax[v] = sm[0]; resMaxPoint[v] = si[0]; }
}
__global__ void __launch_bounds__(256) k_rms_finish(const double* __restrict__ partial, int nBlocks, int nv, long long nPointGlobal, double* __restrict__ rms, double* __restrict__ sumsq) {
  // one block per variable; fixed strided partial sums + fixed tree => bit-reproducible run to run
  const int v = blockIdx.x;
  __shared__ double sh[256];
  double s = 0.0;
  for (int b = threadIdx.x; b < nBlocks; b += 256) s += partial[(size_t)b * nv + v];
  sh[threadIdx.x] = s;
  __syncthreads();
  for (int w = 128; w > 0; w >>= 1) {
    if (threadIdx.x < w) sh[threadIdx.x] += sh[threadIdx.x + w];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    sumsq[v] = sh[0];
    rms[v] = fmax(EPS * EPS, sqrt(sh[0] / (double)nPointGlobal));   // CSolver::SetResidual_RMS, solver_structure.cpp:172-187
  }
}

// norms of a partitioned job: this partition's sums of squares + its owned point count -> (all-reduce) -> global RMS
__global__ void k_rms_pack(const double* __restrict__ sumsq, int nv, double nOwned, double* __restrict__ red) {
  const int v = threadIdx.x;
  if (v < nv) red[v] = sumsq[v];
  if (v == nv) red[nv] = nOwned;
}
__global__ void k_rms_global(const double* __restrict__ red, int nv, double* __restrict__ rms) {
  const int v = threadIdx.x;
  if (v < nv) rms[v] = fmax(EPS * EPS, sqrt(red[v] / red[nv]));
}

// ---- FP64 roof: eight independent DFMA chains per thread, nothing else in the loop
__global__ void __launch_bounds__(256) k_fp64_peak(int iters, double a, double* sink) {
  double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
  const double b = 1.0e-9;
#pragma unroll 4
  for (int i = 0; i < iters; ++i) {
    x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
    x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
  }
  const double s = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
  if (s == 12345.678) *sink = s;   // keeps the chains alive
}

// ---- halo plumbing: gather/scatter listed points of U (what=0) or of [grad, X, V, transport] (what=1)
__global__ void k_pack(const double* __restrict__ soa, long long stride, int nv, long long n, const long long* __restrict__ pts, double* __restrict__ buf) {
  long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= n * nv) return;
  const int v = (int)(t / n); const long long k = t - (long long)v * n;
  buf[t] = soa[v * stride + pts[k]];
}
__global__ void k_unpack(double* __restrict__ soa, long long stride, int nv, long long n, const long long* __restrict__ pts, const double* __restrict__ buf) {
  long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= n * nv) return;
  const int v = (int)(t / n); const long long k = t - (long long)v * n;
  soa[v * stride + pts[k]] = buf[t];
}

// ---- wall distance (CPhysicalGeometry::ComputeWall_Distance, geometry_structure.cpp:8990-9060): distance of every point to the
// nearest NODE of the no-slip markers.  The reference searches an alternating digital tree; here every point scans the wall
// nodes, staged through shared memory in tiles (nWall is O(sqrt(nPoint)) in 2-D): the minimum of the same squared distances.
constexpr int WD_TILE = 1024;
__global__ void __launch_bounds__(256) k_wall_distance(long long nPoint, const double* __restrict__ xy, long long nWall,
                                                       const double* __restrict__ wall, double* __restrict__ dist) {
  __shared__ double wx[WD_TILE], wy[WD_TILE];
  const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const bool in = p < nPoint;
  const double x = in ? xy[2 * p] : 0.0, y = in ? xy[2 * p + 1] : 0.0;
  double best = CUDART_INF;
  for (long long w0 = 0; w0 < nWall; w0 += WD_TILE) {
    const int n = (int)min((long long)WD_TILE, nWall - w0);
    __syncthreads();
    for (int k = threadIdx.x; k < n; k += blockDim.x) { wx[k] = wall[2 * (w0 + k)]; wy[k] = wall[2 * (w0 + k) + 1]; }
    __syncthreads();
    for (int k = 0; k < n; ++k) {
      const double dx = x - wx[k], dy = y - wy[k];
      double d2 = 0.0;
      d2 += dx * dx;
      d2 += dy * dy;
      best = d2 < best ? d2 : best;
    }
  }
  if (in) dist[p] = nWall > 0 ? sqrt(best) : 0.0;
}

// the same for a context: coordinates as the SoA fields of Dev, wall nodes given by their point indices
__global__ void __launch_bounds__(256) k_wall_distance_ctx(const Dev d, long long nWall, const int* __restrict__ wallNode, double* __restrict__ dist) {
  __shared__ double wx[WD_TILE], wy[WD_TILE];
  const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const bool in = p < d.nPoint;
  const double x = in ? d.x[p] : 0.0, y = in ? d.y[p] : 0.0;
  double best = CUDART_INF;
  for (long long w0 = 0; w0 < nWall; w0 += WD_TILE) {
    const int n = (int)min((long long)WD_TILE, nWall - w0);
    __syncthreads();
    for (int k = threadIdx.x; k < n; k += blockDim.x) { const int q = wallNode[w0 + k]; wx[k] = d.x[q]; wy[k] = d.y[q]; }
    __syncthreads();
    for (int k = 0; k < n; ++k) {
      const double dx = x - wx[k], dy = y - wy[k];
      double d2 = 0.0;
      d2 += dx * dx;
      d2 += dy * dy;
      best = d2 < best ? d2 : best;
    }
  }
  if (in) dist[p] = nWall > 0 ? sqrt(best) : 0.0;
}

// ... and for a partition of a multi-GPU job: the wall nodes of ALL partitions, gathered as coordinates (rx_compute_wall_distance)
__global__ void k_wall_gather(const Dev d, long long n, const int* __restrict__ wallNode, double* __restrict__ out) {
  const long long k = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (k < n) { const int q = wallNode[k]; out[2 * k] = d.x[q]; out[2 * k + 1] = d.y[q]; }
}
__global__ void __launch_bounds__(256) k_wall_distance_list(const Dev d, long long nWall, const double* __restrict__ wall, double* __restrict__ dist) {
  __shared__ double wx[WD_TILE], wy[WD_TILE];
  const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const bool in = p < d.nPoint;
  const double x = in ? d.x[p] : 0.0, y = in ? d.y[p] : 0.0;
  double best = CUDART_INF;
  for (long long w0 = 0; w0 < nWall; w0 += WD_TILE) {
    const int n = (int)min((long long)WD_TILE, nWall - w0);
    __syncthreads();
    for (int k = threadIdx.x; k < n; k += blockDim.x) { wx[k] = wall[2 * (w0 + k)]; wy[k] = wall[2 * (w0 + k) + 1]; }
    __syncthreads();
    for (int k = 0; k < n; ++k) {
      const double dx = x - wx[k], dy = y - wy[k];
      double d2 = 0.0;
      d2 += dx * dx;
      d2 += dy * dy;
      best = d2 < best ? d2 : best;
    }
  }
  if (in) dist[p] = nWall > 0 ? sqrt(best) : 0.0;
}

}  // namespace rx
