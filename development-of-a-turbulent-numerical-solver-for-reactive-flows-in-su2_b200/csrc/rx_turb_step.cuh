// rx_turb_step.cuh — the SST transport step on the device (the remainder): one
// CSingleGridIntegration::SingleGrid_Iteration of CTurbSSTSolver (integration_time.cpp:777-834) on the state the flow path left:
//   Preprocessing            solver_direct_turbulent.cpp:2923-2952   gradients of (k, omega) (rx_turb.cuh kernels), zero residual / Jacobian
//   Upwind_Residual          :429-520 + CUpwSca_TurbSST    numerics_direct_turbulent.cpp:865-922   (first order)
//   Viscous_Residual         :521-595 + CAvgGradCorrected_TurbSST :1080-1162
//   Source_Residual          :3018-3086 + CSourcePieceWise_TurbSST :1183-1245
//   BC_Inlet / BC_Outlet     :3270-3450 (CUpwSca_TurbSST + CAvgGrad_TurbSST :960-1040 with the flow path's boundary state)
//   BC_HeatFlux/Isothermal   :3087-3200 (omega_wall = 60 mu / (rho beta_1 d^2) imposed strongly, Jacobian row replaced by the identity)
//   ImplicitEuler_Iteration  :615-728: diagonal += Vol / (CFL_red dt_flow), rhs = -Res, FGMRES (Common/src/linear_solvers_structure.cpp:
//                            309-445, modified Gram-Schmidt :87-175) preconditioned by LU-SGS (matrix_structure.cpp ComputeLU_SGSPreconditioner),
//                            clipped conservative update (CVariable::AddConservativeSolution, variable_structure.cpp:214-220)
//   Postprocessing           rx_turb.cuh (first slice, round 1)
//
// Structure used: every Jacobian block of this model is DIAGONAL (the k and omega rows only couple through the flow), so the 2x2
// block-CSR matrix of the reference is stored as two scalar matrices on the incident-edge CSR the flow path already has:
// diag[v][p] and off[v][slot] (slot = position in incEdge: the entry of row p for the other end of that edge).  Rows are
// assembled owner-computes, per point, in the order the reference's edge / point / vertex loops add to them.
// The LU-SGS sweeps are Gauss-Seidel in point order — sequential by definition.  They are LEVEL-SCHEDULED: level(p) = 1 + max level of
// the lower-numbered neighbours (computed at create), all points of a level are independent and are relaxed together, levels in
// order, so the sweep produces exactly the sequential result; one thread block per variable walks the levels with a block barrier
// in between (the sweep is latency-bound: depth ~ nx + 2 ny for a row-major triangulated channel).
// Single partition only (the reference's sweep is block-Jacobi across partitions with an exchange in between; not built).
#pragma once
#include "rx_turb.cuh"

namespace rx {

struct TurbC {   // constants of CTurbSSTSolver::CTurbSSTSolver :2716-2735 and the cfg values of the implicit step
  double sigma_k1 = 0.85, sigma_k2 = 1.0, sigma_om1 = 0.5, sigma_om2 = 0.856, beta_1 = 0.075, beta_2 = 0.0828, beta_star = 0.09, a1 = 0.31;
  double alfa_1, alfa_2;
  double kineInf, omegaInf, cflRed, relax;
  double lo[2] = {1.0e-10, 1.0e-4}, hi[2] = {1.0e10, 1.0e15};
};

struct TurbW {   // work arrays of the step (SoA, stride np unless noted)
  double *res, *diag, *off;   // LinSysRes [2][np]; Jacobian diagonal [2][np]; off-diagonal [2][nInc] (nInc = 2 nEdge incident slots)
  double *x, *old;            // LinSysSol [2][np]; Solution_Old [2][np]
  long long nInc;
};

// CUpwSca_TurbSST::ComputeResidual for one face: a0, a1 (the Jacobians) and the two residual rows
__device__ __forceinline__ void turb_conv(double ui, double vi, double uj, double vj, double nx, double ny, double rhoi, double rhoj,
                                          const double* ti, const double* tj, double& a0, double& a1, double* r) {
  double q = 0.0;
  q += 0.5 * (ui + uj) * nx;
  q += 0.5 * (vi + vj) * ny;
  a0 = 0.5 * (q + fabs(q));
  a1 = 0.5 * (q - fabs(q));
  r[0] = a0 * rhoi * ti[0] + a1 * rhoj * tj[0];
  r[1] = a0 * rhoi * ti[1] + a1 * rhoj * tj[1];
}
// CAvgGradCorrected_TurbSST (CORR) / CAvgGrad_TurbSST (boundary faces): residual rows and the Jacobian entries ji[v], jj[v]
template <bool CORR>
__device__ __forceinline__ void turb_visc(const TurbC& tc, double xi, double yi, double xj, double yj, double nx, double ny, double rhoi, double rhoj,
                                          double mui, double muj, double mti, double mtj, double F1i, double F1j, const double* ti, const double* tj,
                                          const double* gi /* [2][2] */, const double* gj, double* r, double* ji, double* jj) {
  const double sk_i = F1i * tc.sigma_k1 + (1.0 - F1i) * tc.sigma_k2, sk_j = F1j * tc.sigma_k1 + (1.0 - F1j) * tc.sigma_k2;
  const double so_i = F1i * tc.sigma_om1 + (1.0 - F1i) * tc.sigma_om2, so_j = F1j * tc.sigma_om1 + (1.0 - F1j) * tc.sigma_om2;
  const double dik = mui + sk_i * mti, djk = muj + sk_j * mtj, dio = mui + so_i * mti, djo = muj + so_j * mtj;
  const double diff[2] = {0.5 * (dik + djk), 0.5 * (dio + djo)};
  const double ev[2] = {xj - xi, yj - yi}, nrm[2] = {nx, ny};
  double dist2 = 0.0, proj = 0.0;
#pragma unroll
  for (int d = 0; d < 2; ++d) { dist2 += ev[d] * ev[d]; proj += ev[d] * nrm[d]; }
  if (dist2 == 0.0) proj = 0.0; else proj = proj / dist2;
#pragma unroll
  for (int v = 0; v < 2; ++v) {
    double pn = 0.0, pe = 0.0;
#pragma unroll
    for (int d = 0; d < 2; ++d) {
      const double mg = 0.5 * (gi[v * 2 + d] + gj[v * 2 + d]);
      pn += mg * nrm[d];
      pe += mg * ev[d];
    }
    double pc = pn;
    if (CORR) pc -= pe * proj - (tj[v] - ti[v]) * proj;
    r[v] = diff[v] * pc;
    ji[v] = -diff[v] * proj / rhoi;
    jj[v] = diff[v] * proj / rhoj;
  }
}

// Upwind + Viscous + Source residual and Jacobian rows of every owned point, in the reference's accumulation order
__global__ void __launch_bounds__(128) k_turb_rows(const __grid_constant__ TurbC tc, const Dev d, const double* __restrict__ tp, const TurbW w) {
  const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (p >= d.nPointDomain) return;
  const long long np = d.np;
  auto U = [&](long long q) { return d.V[V_VX * np + q]; };
  auto Vv = [&](long long q) { return d.V[(V_VX + 1) * np + q]; };
  auto RHO = [&](long long q) { return d.V[V_RHO * np + q]; };
  double res[2] = {0.0, 0.0}, dg[2] = {0.0, 0.0};
  const double tpv[2] = {tp[TP_K * np + p], tp[TP_W * np + p]};
  const int k0 = d.incPtr[p], k1 = d.incPtr[p + 1];
  for (int k = k0; k < k1; ++k) {   // Upwind_Residual: edges in ascending order, AddBlock at node 0 / SubtractBlock at node 1
    const int code = d.incEdge[k];
    const bool first = code >= 0;
    const int e = first ? code : ~code;
    const int i = d.ei[e], j = d.ej[e];
    const double ti[2] = {tp[TP_K * np + i], tp[TP_W * np + i]}, tj[2] = {tp[TP_K * np + j], tp[TP_W * np + j]};
    double a0, a1, r[2];
    turb_conv(U(i), Vv(i), U(j), Vv(j), d.enx[e], d.eny[e], RHO(i), RHO(j), ti, tj, a0, a1, r);
#pragma unroll
    for (int v = 0; v < 2; ++v) {
      if (first) { res[v] += r[v]; dg[v] += a0; w.off[v * w.nInc + k] = 0.0 + a1; }
      else { res[v] -= r[v]; dg[v] -= a1; w.off[v * w.nInc + k] = 0.0 - a0; }
    }
  }
  for (int k = k0; k < k1; ++k) {   // Viscous_Residual: SubtractBlock at node 0 / AddBlock at node 1
    const int code = d.incEdge[k];
    const bool first = code >= 0;
    const int e = first ? code : ~code;
    const int i = d.ei[e], j = d.ej[e];
    const double ti[2] = {tp[TP_K * np + i], tp[TP_W * np + i]}, tj[2] = {tp[TP_K * np + j], tp[TP_W * np + j]};
    double gi[4], gj[4], r[2], ji[2], jj[2];
#pragma unroll
    for (int q = 0; q < 4; ++q) { gi[q] = tp[(TP_GK + q) * np + i]; gj[q] = tp[(TP_GK + q) * np + j]; }
    turb_visc<true>(tc, d.x[i], d.y[i], d.x[j], d.y[j], d.enx[e], d.eny[e], RHO(i), RHO(j), d.mu[i], d.mu[j], d.turb[i], d.turb[j],
                    tp[TP_F1 * np + i], tp[TP_F1 * np + j], ti, tj, gi, gj, r, ji, jj);
#pragma unroll
    for (int v = 0; v < 2; ++v) {
      if (first) { res[v] -= r[v]; dg[v] -= ji[v]; w.off[v * w.nInc + k] -= jj[v]; }
      else { res[v] += r[v]; dg[v] += jj[v]; w.off[v * w.nInc + k] += ji[v]; }
    }
  }
  {   // Source_Residual (CSourcePieceWise_TurbSST): SubtractBlock of the residual and of the diagonal Jacobian
    const double F1 = tp[TP_F1 * np + p], F2 = tp[TP_F2 * np + p], CDkw = tp[TP_CD * np + p], dist = tp[TP_DIST * np + p];
    const double rho = RHO(p), muT = d.turb[p], S = d.strain[p], Volume = d.vol[p];
    const double alfa_b = F1 * tc.alfa_1 + (1.0 - F1) * tc.alfa_2, beta_b = F1 * tc.beta_1 + (1.0 - F1) * tc.beta_2;
    double sr[2] = {0.0, 0.0}, sj[2] = {0.0, 0.0};
    if (dist > 1e-10) {
      double diverg = 0.0;
      diverg += d.grad[(G_VX * 2 + 0) * np + p];
      diverg += d.grad[((G_VX + 1) * 2 + 1) * np + p];
      double pk = muT * S * S - 2.0 / 3.0 * rho * tpv[0] * diverg;
      pk = fmin(pk, 20.0 * tc.beta_star * rho * tpv[1] * tpv[0]);
      pk = fmax(pk, 0.0);
      const double zeta = fmax(tpv[1], S * F2 / tc.a1);
      double pw = S * S - 2.0 / 3.0 * zeta * diverg;
      pw = fmax(pw, 0.0);
      sr[0] += pk * Volume;
      sr[1] += alfa_b * rho * pw * Volume;
      sr[0] -= tc.beta_star * rho * tpv[1] * tpv[0] * Volume;
      sr[1] -= beta_b * rho * tpv[1] * tpv[1] * Volume;
      sr[1] += (1.0 - F1) * CDkw * Volume;
      sj[0] = -tc.beta_star * tpv[1] * Volume;
      sj[1] = -2.0 * beta_b * tpv[1] * Volume;
    }
#pragma unroll
    for (int v = 0; v < 2; ++v) { res[v] -= sr[v]; dg[v] -= sj[v]; }
  }
#pragma unroll
  for (int v = 0; v < 2; ++v) { w.res[v * np + p] = res[v]; w.diag[v * np + p] = dg[v]; }
}

// BC_Inlet (inlet = 1: free-stream k, omega outside) / BC_Outlet (Neumann: the interior values outside) of one marker
__global__ void k_turb_bc_io(const __grid_constant__ TurbC tc, const Dev d, const MarkerDev m, const double* __restrict__ tp, const TurbW w, int inlet) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= m.nVertex) return;
  const int p = m.node[k];
  if (p >= d.nPointDomain) return;
  const long long np = d.np;
  const int pn = m.nn[k];
  const double nx = -m.nx[k], ny = -m.ny[k];
  const double ti[2] = {tp[TP_K * np + p], tp[TP_W * np + p]};
  const double tj[2] = {inlet ? tc.kineInf : ti[0], inlet ? tc.omegaInf : ti[1]};
  const double ug = m.ghost[(size_t)V_VX * m.nVertex + k], vg = m.ghost[(size_t)(V_VX + 1) * m.nVertex + k], rhog = m.ghost[(size_t)V_RHO * m.nVertex + k];
  const double rho = d.V[V_RHO * np + p];
  double a0, a1, r[2];
  turb_conv(d.V[V_VX * np + p], d.V[(V_VX + 1) * np + p], ug, vg, nx, ny, rho, rhog, ti, tj, a0, a1, r);
  double g[4], rv[2], ji[2], jj[2];
#pragma unroll
  for (int q = 0; q < 4; ++q) g[q] = tp[(TP_GK + q) * np + p];
  const double F1 = tp[TP_F1 * np + p];
  turb_visc<false>(tc, d.x[p], d.y[p], d.x[pn], d.y[pn], nx, ny, rho, rhog, d.mu[p], d.mu[p], d.turb[p], d.turb[p], F1, F1, ti, tj, g, g, rv, ji, jj);
#pragma unroll
  for (int v = 0; v < 2; ++v) {
    double rr = w.res[v * np + p], dd = w.diag[v * np + p];
    rr += r[v]; dd += a0;
    rr -= rv[v]; dd -= ji[v];
    w.res[v * np + p] = rr; w.diag[v * np + p] = dd;
  }
}
// BC_HeatFlux_Wall / BC_Isothermal_Wall: k = 0, omega = 60 mu / (rho beta_1 d^2) of the normal neighbour, imposed on Solution and
// Solution_Old; zero residual; the Jacobian rows become identity rows (DeleteValsRowi)
__global__ void k_turb_bc_wall(const __grid_constant__ TurbC tc, const Dev d, const MarkerDev m, double* __restrict__ tp, const TurbW w) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= m.nVertex) return;
  const int p = m.node[k];
  if (p >= d.nPointDomain) return;
  const long long np = d.np;
  const int q = m.nn[k];
  double dist = 0.0;
  dist += (d.x[p] - d.x[q]) * (d.x[p] - d.x[q]);
  dist += (d.y[p] - d.y[q]) * (d.y[p] - d.y[q]);
  dist = sqrt(dist);
  const double sol[2] = {0.0, 60.0 * d.mu[q] / (d.V[V_RHO * np + q] * tc.beta_1 * dist * dist)};
#pragma unroll
  for (int v = 0; v < 2; ++v) {
    tp[(TP_K + v) * np + p] = sol[v]; w.old[v * np + p] = sol[v];
    w.res[v * np + p] = 0.0; w.diag[v * np + p] = 1.0;
    for (int s = d.incPtr[p]; s < d.incPtr[p + 1]; ++s) w.off[v * w.nInc + s] = 0.0;
  }
}

// ImplicitEuler_Iteration, first loop: diagonal += Vol / (CFL_red dt_flow), rhs = -Res, x = 0, partial sums of rhs^2 (blocks of 128)
__global__ void __launch_bounds__(128) k_turb_implicit_rhs(const __grid_constant__ TurbC tc, const Dev d, const TurbW w, double* __restrict__ partial) {
  const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  __shared__ double sh[4][2];
  double sq[2] = {0.0, 0.0};
  if (p < d.nPointDomain) {
    const double Delta = d.vol[p] / (tc.cflRed * d.dt[p]);
#pragma unroll
    for (int v = 0; v < 2; ++v) {
      w.diag[v * d.np + p] += Delta;
      const double r = -w.res[v * d.np + p];
      w.res[v * d.np + p] = r; w.x[v * d.np + p] = 0.0;
      sq[v] = r * r;
    }
  }
#pragma unroll
  for (int v = 0; v < 2; ++v) {
    double s = sq[v];
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) s += __shfl_down_sync(0xffffffffu, s, off);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5][v] = s;
  }
  __syncthreads();
  if (threadIdx.x < 2) partial[(size_t)blockIdx.x * 2 + threadIdx.x] = (sh[0][threadIdx.x] + sh[1][threadIdx.x]) + (sh[2][threadIdx.x] + sh[3][threadIdx.x]);
}

// y = A x over the owned points (CSysMatrixVectorProduct): diagonal entry, then the incident slots in edge order
__global__ void __launch_bounds__(128) k_turb_matvec(const Dev d, const TurbW w, const double* __restrict__ x, double* __restrict__ y) {
  const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (p >= d.nPointDomain) return;
  const int k0 = d.incPtr[p], k1 = d.incPtr[p + 1];
#pragma unroll
  for (int v = 0; v < 2; ++v) {
    double s = w.diag[v * d.np + p] * x[v * d.np + p];
    for (int k = k0; k < k1; ++k) {
      const int code = d.incEdge[k];
      const int q = code >= 0 ? d.ej[code] : d.ei[~code];
      s += w.off[v * w.nInc + k] * x[v * d.np + q];
    }
    y[v * d.np + p] = s;
  }
}

// LU-SGS (CSysMatrix::ComputeLU_SGSPreconditioner): (D + L) x* = b by ascending levels, then (D + U) x = D x* by descending levels.
// One block per variable walks the level lists; points of a level only depend on earlier levels.
__global__ void __launch_bounds__(1024) k_turb_lusgs(const Dev d, const TurbW w, const int* __restrict__ lvlPtrF, const int* __restrict__ lvlPtsF, int nLvlF,
                                                    const int* __restrict__ lvlPtrB, const int* __restrict__ lvlPtsB, int nLvlB,
                                                    const double* __restrict__ b, double* __restrict__ out) {
  const int v = blockIdx.x;
  const long long np = d.np;
  const double* dg = w.diag + v * np;
  const double* of = w.off + v * w.nInc;
  double* x = out + v * np;
  const double* bb = b + v * np;
  for (int L = 0; L < nLvlF; ++L) {
    for (int t = lvlPtrF[L] + threadIdx.x; t < lvlPtrF[L + 1]; t += blockDim.x) {
      const int p = lvlPtsF[t];
      double s = 0.0;
      for (int k = d.incPtr[p]; k < d.incPtr[p + 1]; ++k) {
        const int code = d.incEdge[k];
        const int q = code >= 0 ? d.ej[code] : d.ei[~code];
        if (q < p) s += of[k] * x[q];
      }
      x[p] = (bb[p] - s) / dg[p];
    }
    __syncthreads();
  }
  for (int L = 0; L < nLvlB; ++L) {
    for (int t = lvlPtrB[L] + threadIdx.x; t < lvlPtrB[L + 1]; t += blockDim.x) {
      const int p = lvlPtsB[t];
      double s = 0.0;
      for (int k = d.incPtr[p]; k < d.incPtr[p + 1]; ++k) {
        const int code = d.incEdge[k];
        const int q = code >= 0 ? d.ej[code] : d.ei[~code];
        if (q > p) s += of[k] * x[q];
      }
      x[p] = (dg[p] * x[p] - s) / dg[p];
    }
    __syncthreads();
  }
}

// vector kernels over [2][nOwn] (stride np): dot product partials (blocks of 256, fixed tree), axpy, scale
__global__ void __launch_bounds__(256) k_turb_dot(long long n, long long np, const double* __restrict__ a, const double* __restrict__ b, double* __restrict__ partial) {
  __shared__ double sh[256];
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  double s = 0.0;
  if (t < 2 * n) { const long long v = t / n, p = t - v * n; s = a[v * np + p] * b[v * np + p]; }
  sh[threadIdx.x] = s;
  __syncthreads();
  for (int wd = 128; wd > 0; wd >>= 1) { if ((int)threadIdx.x < wd) sh[threadIdx.x] += sh[threadIdx.x + wd]; __syncthreads(); }
  if (threadIdx.x == 0) partial[blockIdx.x] = sh[0];
}
__global__ void __launch_bounds__(256) k_turb_sum(const double* __restrict__ partial, int nPart, int stride, int col, double* __restrict__ out) {
  __shared__ double sh[256];
  double s = 0.0;
  for (int b = threadIdx.x; b < nPart; b += 256) s += partial[(size_t)b * stride + col];
  sh[threadIdx.x] = s;
  __syncthreads();
  for (int wd = 128; wd > 0; wd >>= 1) { if ((int)threadIdx.x < wd) sh[threadIdx.x] += sh[threadIdx.x + wd]; __syncthreads(); }
  if (threadIdx.x == 0) out[col] = sh[0];
}
__global__ void k_turb_axpy(long long n, long long np, double a, const double* __restrict__ x, double* __restrict__ y) {   // y += a x
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= 2 * n) return;
  const long long v = t / n, p = t - v * n;
  y[v * np + p] += a * x[v * np + p];
}
__global__ void k_turb_scale_div(long long n, long long np, double den, double sign, const double* __restrict__ x, double* __restrict__ y) {   // y = (sign x) / den
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= 2 * n) return;
  const long long v = t / n, p = t - v * n;
  y[v * np + p] = (sign * x[v * np + p]) / den;
}

// update of (k, omega): AddConservativeSolution with the flow's old and new density, clipped (ImplicitEuler_Iteration :690-705)
__global__ void __launch_bounds__(128) k_turb_update(const __grid_constant__ TurbC tc, const Dev d, double* __restrict__ tp, const TurbW w) {
  const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (p >= d.nPointDomain) return;
  const double rhoOld = d.Uold[U_RHO * d.np + p], rho = d.V[V_RHO * d.np + p];
#pragma unroll
  for (int v = 0; v < 2; ++v) {
    const double val = (w.old[v * d.np + p] * rhoOld + tc.relax * w.x[v * d.np + p]) / rho;
    const double lo = (val < tc.lo[v]) ? tc.lo[v] : val;   // std::min(std::max(val, lower), upper)
    tp[(TP_K + v) * d.np + p] = (tc.hi[v] < lo) ? tc.hi[v] : lo;
  }
}

}  // namespace rx
