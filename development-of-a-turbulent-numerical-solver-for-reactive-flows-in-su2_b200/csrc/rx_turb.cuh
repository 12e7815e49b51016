// rx_turb.cuh — SST Postprocessing on the device (first slice): what CTurbSSTSolver::Postprocessing
// (solver_direct_turbulent.cpp:2953-3015) leaves for the flow path — eddy viscosity, grad k, sigma_k — from the turbulence
// solution (k, omega), the wall distance and the flow state already resident (rho, laminar viscosity, strain magnitude).
// With it the host hands over 2 doubles per point and iteration instead of the 6 of rx_set_turbulence.
//   k_turb_gradient_gg_* CSolver::SetSolution_Gradient_GG, solver_structure.cpp:519-579 (GREEN_GAUSS configurations)
//   k_turb_gradient_ls   CSolver::SetSolution_Gradient_LS, solver_structure.cpp:580-722 (2-D branch, nVar = 2).  Its tests are
//                        not those of the flow solver's own LS routine: weight != 0, r11 >= 0, |detR2| <= EPS.
//   k_turb_blend         CTurbSSTVariable::SetBlendingFunc (variable_direct_turbulent.cpp: CDkw, F1, F2) and the eddy viscosity
//                        (:2996-2999); constants of CTurbSSTSolver::CTurbSSTSolver :2716-2725.
// Plain IEEE arithmetic in the reference's order (the library is compiled with -fmad=false); pow and tanh are CUDA's (<= 2 ulp
// from glibc's), which is what the 1e-12 tolerance of the parity test absorbs.
// Layout of tp (SoA, stride np), the columns of the reference dump `tpost`:
//   0 k, 1 omega, 2 wall distance, 3 F1, 4 F2, 5 CDkw, 6 muT, 7 dk/dx, 8 dk/dy, 9 domega/dx, 10 domega/dy, 11 sigma_k
#pragma once
#include "rx_kernels.cuh"

namespace rx {

enum { TP_K = 0, TP_W = 1, TP_DIST = 2, TP_F1 = 3, TP_F2 = 4, TP_CD = 5, TP_MUT = 6, TP_GK = 7, TP_GW = 9, TP_SIGK = 11, TP_N = 12 };

__global__ void __launch_bounds__(128) k_turb_gradient_ls(const Dev d, double* __restrict__ tp) {
  const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (p >= d.nPoint) return;
  const long long np = d.np;
  if (p >= d.nPointDomain) {   // halo points receive theirs through the exchange (Set_MPI_Solution_Gradient, what = 3)
#pragma unroll
    for (int q = 0; q < 4; ++q) tp[(TP_GK + q) * np + p] = 0.0;
    return;
  }
  const double xi = d.x[p], yi = d.y[p];
  const double ki = tp[TP_K * np + p], wi = tp[TP_W * np + p];
  double C[2][2] = {{0.0, 0.0}, {0.0, 0.0}};
  double r11 = 0.0, r12 = 0.0, r22 = 0.0;
  for (int n = d.nbPtr[p]; n < d.nbPtr[p + 1]; ++n) {
    const int q = d.nbIdx[n];
    const double dx = d.x[q] - xi, dy = d.y[q] - yi;
    double weight = 0.0;
    weight += dx * dx;
    weight += dy * dy;
    if (weight != 0.0) {
      const double dk = tp[TP_K * np + q] - ki, dw = tp[TP_W * np + q] - wi;
      // the seven quotients share the divisor: one refined reciprocal, nvcc's own fast-path sequence per quotient and its
      // validity test ORed into one flag (rx_fdiv.cuh); the plain divisions are repeated if the flag comes up — same bits
      const double num[7] = {dx * dx, dx * dy, dy * dy, dx * dk, dy * dk, dx * dw, dy * dw};
      const Recip rw = frecip(weight);
      bool bad = false;
      double qv[7];
#pragma unroll
      for (int m = 0; m < 7; ++m) qv[m] = fdiv(num[m], rw, bad);
      if (bad) {
#pragma unroll
        for (int m = 0; m < 7; ++m) qv[m] = num[m] / weight;
      }
      r11 += qv[0]; r12 += qv[1]; r22 += qv[2];
      C[0][0] += qv[3]; C[0][1] += qv[4];
      C[1][0] += qv[5]; C[1][1] += qv[6];
    }
  }
  if (r11 >= 0.0) r11 = sqrt(r11); else r11 = 0.0;
  if (r11 != 0.0) r12 = r12 / r11; else r12 = 0.0;
  if (r22 - r12 * r12 >= 0.0) r22 = sqrt(r22 - r12 * r12); else r22 = 0.0;
  double detR2 = (r11 * r22) * (r11 * r22);
  bool singular = false;
  if (fabs(detR2) <= EPS) { detR2 = 1.0; singular = true; }
  double S00 = 0.0, S01 = 0.0, S11 = 0.0;
  if (!singular) {
    S00 = (r12 * r12 + r22 * r22) / detR2;
    S01 = -r11 * r12 / detR2;
    S11 = r11 * r11 / detR2;
  }
#pragma unroll
  for (int v = 0; v < 2; ++v) {
    double g0 = 0.0, g1 = 0.0;
    g0 += S00 * C[v][0]; g0 += S01 * C[v][1];
    g1 += S01 * C[v][0]; g1 += S11 * C[v][1];
    tp[(TP_GK + 2 * v) * np + p] = g0;
    tp[(TP_GK + 2 * v + 1) * np + p] = g1;
  }
}

// ---- CSolver::SetSolution_Gradient_GG (solver_structure.cpp:519-579) of (k, omega), owner-computes: a thread per owned point
// adds the interior-edge terms 0.5 (s_i + s_j) n in ascending edge order (+ at the edge's first node, - at its second), the
// marker kernels subtract s n_vertex in marker order, the last kernel divides by (volume + EPS) — the reference's order of sums.
__global__ void __launch_bounds__(128) k_turb_gradient_gg_edges(const Dev d, double* __restrict__ tp) {
  const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (p >= d.nPoint) return;
  const long long np = d.np;
  double g[4] = {0.0, 0.0, 0.0, 0.0};
  if (p < d.nPointDomain) {
    for (int k = d.incPtr[p]; k < d.incPtr[p + 1]; ++k) {
      const int code = d.incEdge[k];
      const bool first = code >= 0;
      const long long e = first ? code : ~code;
      const int i = d.ei[e], j = d.ej[e];
      const double nx = d.enx[e], ny = d.eny[e];
#pragma unroll
      for (int v = 0; v < 2; ++v) {
        const double avg = 0.5 * (tp[(TP_K + v) * np + i] + tp[(TP_K + v) * np + j]);
        const double px = avg * nx, py = avg * ny;
        g[2 * v] = first ? g[2 * v] + px : g[2 * v] - px;
        g[2 * v + 1] = first ? g[2 * v + 1] + py : g[2 * v + 1] - py;
      }
    }
  }
#pragma unroll
  for (int q = 0; q < 4; ++q) tp[(TP_GK + q) * np + p] = g[q];
}
__global__ void k_turb_gradient_gg_boundary(const Dev d, const MarkerDev m, double* __restrict__ tp) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= m.nVertex) return;
  const int p = m.node[k];
  if (p >= d.nPointDomain) return;
  const long long np = d.np;
#pragma unroll
  for (int v = 0; v < 2; ++v) {
    const double s = tp[(TP_K + v) * np + p];
    tp[(TP_GK + 2 * v) * np + p] -= s * m.nx[k];
    tp[(TP_GK + 2 * v + 1) * np + p] -= s * m.ny[k];
  }
}
__global__ void __launch_bounds__(128) k_turb_gradient_gg_finish(const Dev d, double* __restrict__ tp) {
  const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (p >= d.nPointDomain) return;
  const double a = d.vol[p] + EPS;
#pragma unroll
  for (int q = 0; q < 4; ++q) tp[(TP_GK + q) * d.np + p] = tp[(TP_GK + q) * d.np + p] / a;
}

// blending functions and eddy viscosity of every point (halo points included, as in the reference's loop over nPoint); fills
// the six fields the flow path reads (d.turb: muT, k, omega, dk/dx, dk/dy, sigma_k)
__global__ void __launch_bounds__(128) k_turb_blend(const Dev d, double* __restrict__ tp, double sigmaK) {
  const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (p >= d.nPoint) return;
  const long long np = d.np;
  const double sigma_om2 = 0.856, beta_star = 0.09, a1 = 0.31;
  const double rho = d.V[V_RHO * np + p], mu = d.mu[p], strMag = d.strain[p];
  const double kine = tp[TP_K * np + p], omega = tp[TP_W * np + p], dist = tp[TP_DIST * np + p];
  const double gk0 = tp[TP_GK * np + p], gk1 = tp[(TP_GK + 1) * np + p];
  const double gw0 = tp[TP_GW * np + p], gw1 = tp[(TP_GW + 1) * np + p];
  double CDkw = 0.0;
  CDkw += gk0 * gw0;
  CDkw += gk1 * gw1;
  CDkw *= 2.0 * rho * sigma_om2 / omega;
  CDkw = fmax(CDkw, 1.0e-20);
  const double arg2A = sqrt(kine) / (beta_star * omega * dist + EPS * EPS);
  const double arg2B = 500.0 * mu / (rho * dist * dist * omega + EPS * EPS);
  double arg2 = fmax(arg2A, arg2B);
  const double arg1 = fmin(arg2, 4.0 * rho * sigma_om2 * kine / (CDkw * dist * dist + EPS * EPS));
  const double F1 = tanh(rx_pow(arg1, 4.0));
  arg2 = fmax(2.0 * arg2A, arg2B);
  const double F2 = tanh(arg2 * arg2);   // g++ folds pow(x, 2.0) to x*x at -O2 (pow(x, 4.0) above stays a libm call)
  const double zeta = fmin(1.0 / omega, a1 / (strMag * F2));
  const double muT = fmin(fmax(rho * kine * zeta, 0.0), 1.0);
  tp[TP_F1 * np + p] = F1; tp[TP_F2 * np + p] = F2; tp[TP_CD * np + p] = CDkw; tp[TP_MUT * np + p] = muT; tp[TP_SIGK * np + p] = sigmaK;
  d.turb[0 * np + p] = muT; d.turb[1 * np + p] = kine; d.turb[2 * np + p] = omega;
  d.turb[3 * np + p] = gk0; d.turb[4 * np + p] = gk1; d.turb[5 * np + p] = sigmaK;
}

}  // namespace rx
