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

// CTurbSolver::SetResidual_DualTime (solver_direct_turbulent.cpp:730-815; static mesh, SST: conservative variables, the flow solver's density
// of the three time levels :778-780): AddBlock of the residual and of the diagonal Jacobian, after the source term, before the boundary conditions
__global__ void __launch_bounds__(128) k_turb_dual_time(const Dev d, const double* __restrict__ tp, const TurbW w, int order, const double* __restrict__ dtND,
                                                        const double* __restrict__ Un, const double* __restrict__ Un1,
                                                        const double* __restrict__ kwN, const double* __restrict__ kwN1) {
  const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (p >= d.nPointDomain) return;
  const long long np = d.np;
  const double Vol = d.vol[p], TimeStep = dtND[0];
  const double rP1 = d.U[p], rN = Un[p], rM1 = Un1[p];   // row 0 of the conserved variables: the mixture density
#pragma unroll
  for (int v = 0; v < 2; ++v) {
    double r, jd;
    if (order == 1) { r = (rP1 * tp[(TP_K + v) * np + p] - rN * kwN[v * np + p]) * Vol / TimeStep; jd = Vol / TimeStep; }
    else {
      r = (3.0 * rP1 * tp[(TP_K + v) * np + p] - 4.0 * rN * kwN[v * np + p] + 1.0 * rM1 * kwN1[v * np + p]) * Vol / (2.0 * TimeStep);
      jd = (Vol * 3.0) / (2.0 * TimeStep);
    }
    w.res[v * np + p] += r; w.diag[v * np + p] += jd;
  }
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
// One block per variable walks the level lists; points of a level only depend on earlier levels.  A sweep is a chain of ~(nx + 2 ny)
// dependent steps on ONE SM, so what counts per step is (1) the number of dependent memory round trips and (2) the bytes that SM's
// load/store path moves.  Hence the sweep runs in LEVEL-ORDER space: right-hand side, diagonal, coefficients, the neighbour lists and
// the unknown itself are held in the order in which the sweep visits the points (SweepList; permuted in and out by full-grid
// kernels), so every access of a level is a contiguous stream except the gathers of the neighbours' x, which land in the adjacent
// levels' contiguous ranges; measured on the bench mesh: 27.7 ms per application with point-order arrays (the SM's load path moves
// ~320 KB of 32-byte sectors per level for 8-byte words), see profiles/r2_summary.md for the level-order figure.
// The loop over the levels is software-pipelined, one stage per dependent load: while level L is relaxed, the level bounds of L+3,
// the list offsets of L+2 and the neighbour positions, coefficients, right-hand side and diagonal of L+1 are in flight.
struct SweepList {
  const int* lvlPtr;   // [nLvl + 1] offsets of the levels in level order
  const int* perm;     // [n] level-order position -> point
  const int* start;    // [n + 1] offsets of a position's (lower or upper) neighbours in colL / slot
  const int* colL;     // neighbour's POSITION in this list's level order
  const int* slot;     // index of the coefficient in TurbW::off (for the per-step gather into coefL)
  int nLvl, nEnt;      // levels; neighbour entries
  int nX;              // entries of x per variable: n, plus (backward list of a partition) the halo points, whose x* the upper
                       // triangle reads at positions n .. nPoint-1 (CSysMatrix::UpperProduct over columns received from the neighbours)
};
struct SweepData {     // per variable, in level order of one SweepList
  double *coefL, *dgL, *bL, *xL;   // [2][nEnt], [2][n], [2][n], [2][nX]
};
constexpr int SW_PRE = 3;   // neighbours carried in registers per point (a triangulated mesh has ~3 lower and ~3 upper ones; more: a plain loop)
struct SwS0 { int n0, n1; };
struct SwS1 { int t, a, cnt; };                                      // cnt < 0: no point for this thread at that level
struct SwS2 { int t, a, cnt; int col[SW_PRE]; double coef[SW_PRE]; double b, dg; };   // b: right-hand side (forward) or x* (backward)

template <bool BACK>
__device__ __forceinline__ void sweep_run(const SweepList& sl, const double* __restrict__ coefL, const double* __restrict__ dgL, const double* __restrict__ bL, double* xL) {
  const int nL = sl.nLvl;
  auto st0 = [&](int L) { SwS0 r; r.n0 = 0; r.n1 = 0; if (L < nL) { r.n0 = sl.lvlPtr[L]; r.n1 = sl.lvlPtr[L + 1]; } return r; };
  auto st1 = [&](const SwS0& q) { SwS1 r; r.t = q.n0 + (int)threadIdx.x; r.a = 0; r.cnt = -1;
                                  if (r.t < q.n1) { r.a = sl.start[r.t]; r.cnt = sl.start[r.t + 1] - r.a; } return r; };
  auto st2 = [&](const SwS1& q) { SwS2 r; r.t = q.t; r.a = q.a; r.cnt = q.cnt; r.b = 0.0; r.dg = 1.0;
#pragma unroll
                                  for (int j = 0; j < SW_PRE; ++j) { r.col[j] = 0; r.coef[j] = 0.0; }
                                  if (q.cnt >= 0) {
#pragma unroll
                                    for (int j = 0; j < SW_PRE; ++j) if (j < q.cnt) { r.col[j] = sl.colL[q.a + j]; r.coef[j] = coefL[q.a + j]; }
                                    r.dg = dgL[q.t];
                                    r.b = BACK ? xL[q.t] : bL[q.t];
                                  } return r; };
  SwS0 q0 = st0(2);
  SwS1 q1 = st1(st0(1));
  SwS2 q2 = st2(st1(st0(0)));
  for (int L = 0; L < nL; ++L) {
    const SwS0 n0 = st0(L + 3);
    const SwS1 n1 = st1(q0);
    const SwS2 n2 = st2(q1);
    if (q2.cnt >= 0) {
      double s = 0.0;
#pragma unroll
      for (int j = 0; j < SW_PRE; ++j)
        if (j < q2.cnt) s += q2.coef[j] * xL[q2.col[j]];
      for (int j = SW_PRE; j < q2.cnt; ++j) s += coefL[q2.a + j] * xL[sl.colL[q2.a + j]];
      xL[q2.t] = BACK ? (q2.dg * q2.b - s) / q2.dg : (q2.b - s) / q2.dg;
    }
    {   // levels wider than the block: the rest, unpipelined
      const int m0 = sl.lvlPtr[L], m1 = sl.lvlPtr[L + 1];
      for (int t = m0 + (int)blockDim.x + (int)threadIdx.x; t < m1; t += blockDim.x) {
        const int a = sl.start[t], e = sl.start[t + 1];
        double s = 0.0;
        for (int k = a; k < e; ++k) s += coefL[k] * xL[sl.colL[k]];
        xL[t] = BACK ? (dgL[t] * xL[t] - s) / dgL[t] : (bL[t] - s) / dgL[t];
      }
    }
    __syncthreads();
    q0 = n0; q1 = n1; q2 = n2;
  }
}
template <bool BACK>
__global__ void __launch_bounds__(1024) k_turb_sweep(const SweepList sl, const SweepData sd, long long n) {
  const int v = blockIdx.x;
  sweep_run<BACK>(sl, sd.coefL + (size_t)v * sl.nEnt, sd.dgL + v * n, sd.bL + v * n, sd.xL + (size_t)v * sl.nX);
}
// per step: coefficients and diagonal of the assembled rows into level order
__global__ void k_turb_sweep_coef(const SweepList sl, const SweepData sd, const TurbW w, long long n, long long np) {
  const long long k = blockIdx.x * (long long)blockDim.x + threadIdx.x;
#pragma unroll
  for (int v = 0; v < 2; ++v) {
    if (k < sl.nEnt) sd.coefL[(size_t)v * sl.nEnt + k] = w.off[v * w.nInc + sl.slot[k]];
    if (k < n) sd.dgL[v * n + k] = w.diag[v * np + sl.perm[k]];
  }
}
// per application: point order -> level order (dst[v][t] = src[v][perm[t]]), level order A -> level order B, level order -> point order
__global__ void k_turb_perm_in(const int* __restrict__ perm, long long n, long long np, const double* __restrict__ src, double* __restrict__ dst) {
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= n) return;
  const int p = perm[t];
  dst[t] = src[p]; dst[n + t] = src[np + p];
}
__global__ void k_turb_perm_ab(const int* __restrict__ permB, const int* __restrict__ posA, long long n, long long nXB, const double* __restrict__ xA, double* __restrict__ xB) {
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= n) return;
  const int s = posA[permB[t]];
  xB[t] = xA[s]; xB[nXB + t] = xA[n + s];
}
__global__ void k_turb_perm_out(const int* __restrict__ perm, long long n, long long nXsrc, long long np, const double* __restrict__ src, double* __restrict__ dst) {
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= n) return;
  const int p = perm[t];
  dst[p] = src[t]; dst[np + p] = src[nXsrc + t];
}
// partition: the neighbours' x* (received into the halo entries of a point-order vector) -> the tail of the backward list's x
__global__ void k_turb_halo_in(long long n, long long nPoint, long long np, const double* __restrict__ src, double* __restrict__ xB) {
  const long long h = n + blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (h >= nPoint) return;
  xB[h] = src[h]; xB[nPoint + h] = src[np + h];
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
