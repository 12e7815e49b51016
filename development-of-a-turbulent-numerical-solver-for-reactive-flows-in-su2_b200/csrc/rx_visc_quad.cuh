// rx_visc_quad.cuh — laminar viscous edge flux with FOUR LANES PER EDGE (eight edges per warp).
//
// Same arithmetic, bit for bit, as k_edge_visc_pair<NS,false> (rx_visc_pair.cuh) and viscous_flux<NS, EDGE=true> (rx_flux.cuh;
// CAvgGradReactive_Flow::ComputeResidual, numerics_direct_reactive.cpp:1425-1573; Solve_SM :451-470; GetGamma
// reacting_model_library.cpp:771-798).  The pair kernel is bound by the number of warps that fit at 168 registers; here Eigen's
// packet reduction of a 9-sum, ((a0+a4)+(a2+a6)) + ((a1+a5)+(a3+a7)), then + a8, is split once more:
//     lane 0: rows 0,4     lane 1: rows 2,6     lane 2: rows 1,5     lane 3: rows 3,7        (slot k < 2: row 4k + 2*(q&1) + (q>>1))
// so lane q forms a_r + a_{r+4}, the sums meet through shfl_xor 1 (the two halves of one packet lane) and shfl_xor 2 (even + odd
// packet lane) — IEEE addition commutes, so every lane holds the reference's value.  Row 8, the scalar tail, is slot 2 of ALL
// four lanes: each forms it from the same operands in the same (sequential) order, so it never has to be exchanged.
#pragma once
#include "rx_visc_pair.cuh"

namespace rx {

#ifndef RX_VQ_MINB
#define RX_VQ_MINB 4
#endif
#ifndef RX_VQ_MREG
#define RX_VQ_MREG 1
#endif
constexpr int VQ_BS = 128;            // threads per block (four per edge)
constexpr int VQ_MINB = RX_VQ_MINB;   // resident blocks per SM the register budget is set for
constexpr int VQ_MREG = RX_VQ_MREG;   // matrix slots held in registers during the solve
constexpr int VQ_SMW = 3 * 9 + 2 * 9; // shared-memory words per thread: three matrix rows + the staged 1/D words (rhs, r0, x later)

struct QuadLane {
  int q;      // 0..3
  int base;   // warp lane of the quad's lane 0
  __device__ __forceinline__ double x1(double v) const { return __shfl_xor_sync(VP_FULL, v, 1); }
  __device__ __forceinline__ double x2(double v) const { return __shfl_xor_sync(VP_FULL, v, 2); }
  __device__ __forceinline__ double from(double v, int k) const { return __shfl_sync(VP_FULL, v, base + k); }
};

// a[k] <- a[k] / M[row[k]] for the three slots of a lane
template <int NS>
__device__ __forceinline__ void div_by_M_rows3(const ChemC<NS>& cc, const int* row, double* a) {
  bool ok = true;
#pragma unroll
  for (int k = 0; k < 3; ++k) ok = ok && num_range(a[k]);
  if (ok) {
#pragma unroll
    for (int k = 0; k < 3; ++k) a[k] = fdivu(a[k], Recip{cc.M[row[k]], cc.rM[row[k]]});
  } else {
#pragma unroll
    for (int k = 0; k < 3; ++k) a[k] = a[k] / cc.M[row[k]];
  }
}

// sum of nine values spread over the quad in Eigen redux order
__device__ __forceinline__ double quad_redux(const QuadLane& L, const double* a) {
  const double s = a[0] + a[1];
  const double h = s + L.x1(s);
  const double tot = h + L.x2(h);
  return tot + a[2];
}
__device__ __forceinline__ double quad_dot(const QuadLane& L, const double* a, const double* b) {
  double t[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) t[k] = a[k] * b[k];
  return quad_redux(L, t);
}
// all nine entries of a vector spread over the quad: row r < 8 lives in slot r>>2 of lane ((r&3)>>1) + 2*(r&1)
__device__ __forceinline__ void quad_gather(const QuadLane& L, const double* a, double* f) {
#pragma unroll
  for (int r = 0; r < 8; ++r) f[r] = L.from(a[r >> 2], ((r & 3) >> 1) + 2 * (r & 1));
  f[8] = a[2];
}
// rows of the lane of A*f, Eigen GEMV order per row (rows < 8 the four-column tree, row 8 sequential)
template <class MatF>
__device__ __forceinline__ void quad_gemv(MatF A, const double* f, double* res) {
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    double r = 0.0;
    r = r + ((A(k, 0) * f[0] + A(k, 1) * f[1]) + (A(k, 2) * f[2] + A(k, 3) * f[3]));
    r = r + ((A(k, 4) * f[4] + A(k, 5) * f[5]) + (A(k, 6) * f[6] + A(k, 7) * f[7]));
    res[k] = A(k, 8) * f[8] + r;
  }
  {
    double r = 0.0;
#pragma unroll
    for (int c = 0; c < 9; ++c) r = A(2, c) * f[c] + r;
    res[2] = r;
  }
}

template <int NS>
__global__ void __launch_bounds__(VQ_BS, VQ_MINB) k_edge_visc_quad(const __grid_constant__ ChemC<NS> cc, const __grid_constant__ Opts o, const Dev d) {
  static_assert(NS == 9, "the row split {0,4 | 2,6 | 1,5 | 3,7} + 8 is laid out for nine species");
  extern __shared__ double vq_sm[];   // word w of thread t at vq_sm[w * VQ_BS + t]
  double* const sm = vq_sm + threadIdx.x;
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  QuadLane L;
  L.q = (int)(threadIdx.x & 3);
  L.base = (int)(threadIdx.x & 28);
  const int q4 = L.q;
  long long e = t >> 2;
  const bool keep = e < d.nEdge;
  if (!keep) e = d.nEdge - 1;   // the lane stays in the warp's shuffles and votes; nothing it computes is stored
  const long long np = d.np;
  const int pi = d.ei[e], pj = d.ej[e];
  const double nx = d.enx[e], ny = d.eny[e];
  const int r0 = 2 * (q4 & 1) + (q4 >> 1);
  const int row[3] = {r0, r0 + 4, 8};

  // ---- mean state, all lanes
  const double rho = 0.5 * (d.V[V_RHO * np + pi] + d.V[V_RHO * np + pj]);
  const double T = 0.5 * (d.V[V_T * np + pi] + d.V[V_T * np + pj]);
  const double um = 0.5 * (d.V[V_VX * np + pi] + d.V[V_VX * np + pj]);
  const double vm = 0.5 * (d.V[(V_VX + 1) * np + pi] + d.V[(V_VX + 1) * np + pj]);
  const double mu = 2.0 / (1.0 / d.mu[pi] + 1.0 / d.mu[pj]);
  const double ktr = 2.0 / (1.0 / d.kap[pi] + 1.0 / d.kap[pj]);
  double Ys[NS], Xs[NS];
#pragma unroll
  for (int s = 0; s < NS; ++s) Ys[s] = 0.5 * (d.V[(V_Y + s) * np + pi] + d.V[(V_Y + s) * np + pj]);
  molar_from_mass(cc, Ys, Xs);

  // ---- mean gradients with the edge correction (:1480-1500), three of the twelve rows per lane:
  // slot 0: T | u | v | X_8 on lanes 0..3,  slot 1: X_{r0},  slot 2: X_{r0+4}
  double G[3][2];
  {
    const double e0 = d.x[pj] - d.x[pi], e1 = d.y[pj] - d.y[pi];
    const double dist2 = 0.0 + e0 * e0 + e1 * e1;
    double gi0[3], gj0[3], gi1[3], gj1[3], vi[3], vj[3];   // all loads first: the divisions below end basic blocks
#pragma unroll
    for (int m = 0; m < 3; ++m) {
      int gr;                 // gradient row
      const double* val;      // field the row differentiates
      if (m == 0) {
        gr = q4 == 0 ? G_T : q4 == 1 ? G_VX : q4 == 2 ? G_VX + 1 : G_X + 8;
        val = q4 == 3 ? d.X + (size_t)8 * np : d.V + (size_t)(q4 == 0 ? V_T : q4 == 1 ? V_VX : V_VX + 1) * np;
      } else { gr = G_X + r0 + 4 * (m - 1); val = d.X + (size_t)(r0 + 4 * (m - 1)) * np; }
      const double* g = d.grad + (size_t)(gr * 2) * np;
      gi0[m] = g[pi]; gj0[m] = g[pj]; gi1[m] = g[np + pi]; gj1[m] = g[np + pj];
      vi[m] = val[pi]; vj[m] = val[pj];
    }
    const Recip rd2 = frecip(dist2);   // one refined reciprocal serves the six quotients
    bool bad = false;
    double c0[3], c1[3];
#pragma unroll
    for (int m = 0; m < 3; ++m) {
      const double g0 = 0.5 * (gi0[m] + gj0[m]);
      const double g1 = 0.5 * (gi1[m] + gj1[m]);
      const double proj = (0.0 + g0 * e0) + g1 * e1;
      const double diff = vj[m] - vi[m];
      c0[m] = (proj - diff) * e0; c1[m] = (proj - diff) * e1;
      G[m][0] = g0 - fdiv(c0[m], rd2, bad);
      G[m][1] = g1 - fdiv(c1[m], rd2, bad);
    }
    if (bad) {
#pragma unroll
      for (int m = 0; m < 3; ++m) {
        G[m][0] = 0.5 * (gi0[m] + gj0[m]) - c0[m] / dist2;
        G[m][1] = 0.5 * (gi1[m] + gj1[m]) - c1[m] / dist2;
      }
    }
  }
  // rows T, u, v, X_8 on all lanes
  double GT[2], GU[2], GV[2], GX8[2];
#pragma unroll
  for (int dd = 0; dd < 2; ++dd) {
    GT[dd] = L.from(G[0][dd], 0); GU[dd] = L.from(G[0][dd], 1); GV[dd] = L.from(G[0][dd], 2); GX8[dd] = L.from(G[0][dd], 3);
  }
  // viscous stresses and heat flux (SetLaminarTensorFlux :1099-1190)
  double FT[3][2];
  {
    const double Gv[2][2] = {{GU[0], GU[1]}, {GV[0], GV[1]}};   // Gv[c][dim]
    const double vel[2] = {um, vm};
    double tau[2][2];
    double div_vel = 0.0;
    div_vel += Gv[0][0]; div_vel += Gv[1][1];
#pragma unroll
    for (int a = 0; a < 2; ++a) {
      tau[a][0] = 0.0; tau[a][1] = 0.0;
#pragma unroll
      for (int b = 0; b < 2; ++b) tau[a][b] += mu * (Gv[b][a] + Gv[a][b]);
      tau[a][a] -= TWO3 * (mu * div_vel);
    }
#pragma unroll
    for (int a = 0; a < 2; ++a) {
      FT[2][a] = 0.0;
#pragma unroll
      for (int b = 0; b < 2; ++b) { FT[b][a] = tau[a][b]; FT[2][a] += tau[a][b] * vel[b]; }
      FT[2][a] += ktr * GT[a];
    }
  }
  // momentum rows are complete; of the energy row only the two stress/heat products have to outlive the solve
  const double fE0 = FT[2][0] * nx, fE1 = FT[2][1] * ny;
  if (keep && q4 == 0) {
    double pMx = 0.0, pMy = 0.0;
    pMx += FT[0][0] * nx; pMx += FT[0][1] * ny;
    pMy += FT[1][0] * nx; pMy += FT[1][1] * ny;
    d.fluxV[U_MX * d.ne + e] = pMx;
    d.fluxV[(U_MX + 1) * d.ne + e] = pMy;
  }
  // right-hand side of the lane's rows: -(grad X . n)
  double rhs[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    double gxn = 0.0;
    gxn += (k < 2 ? G[1 + k][0] : GX8[0]) * nx;
    gxn += (k < 2 ? G[1 + k][1] : GX8[1]) * ny;
    rhs[k] = -gxn;
  }

  // ---- Stefan–Maxwell matrix, the lane's rows (GetGamma :771-798, Gamma_tilde :463-465); see rx_visc_pair.cuh for the
  // formulae.  Slot k, column c at word k*NS + c of the thread's shared-memory words; the loop over the slots stays rolled.
  auto smA = [&](int k, int c) -> double& { return sm[(k * NS + c) * VQ_BS]; };
  double invd[3];
  {
    double sigma = 0.0, massTot = 0.0, yom9[NS];
#pragma unroll
    for (int s = 0; s < NS; ++s) sigma = sigma + Ys[s];
#pragma unroll
    for (int s = 0; s < NS; ++s) yom9[s] = Ys[s];
    div_by_M(cc, yom9);
#pragma unroll
    for (int s = 0; s < NS; ++s) massTot += yom9[s];
    massTot = 1.0 / massTot;
    const double smm = -sigma * massTot;
    const double* rdi = d.rD + pi;
    const double* rdj = d.rD + pj;
    double maxD = 0.0;
    double Xrow[3], Yrow[3];
    {
      const double xa = (q4 & 1) ? Xs[2] : Xs[0], xb = (q4 & 1) ? Xs[3] : Xs[1];
      const double xc = (q4 & 1) ? Xs[6] : Xs[4], xd = (q4 & 1) ? Xs[7] : Xs[5];
      const double ya = (q4 & 1) ? Ys[2] : Ys[0], yb = (q4 & 1) ? Ys[3] : Ys[1];
      const double yc = (q4 & 1) ? Ys[6] : Ys[4], yd = (q4 & 1) ? Ys[7] : Ys[5];
      Xrow[0] = (q4 >> 1) ? xb : xa; Xrow[1] = (q4 >> 1) ? xd : xc; Xrow[2] = Xs[8];
      Yrow[0] = (q4 >> 1) ? yb : ya; Yrow[1] = (q4 >> 1) ? yd : yc; Yrow[2] = Ys[8];
    }
    bool inOk = mid_range(rho);   // rho and the mole fractions enter every quotient of the assembly
#pragma unroll
    for (int s = 0; s < NS; ++s) inOk = inOk && num_range(Xs[s]);
    // 1/D_rc of the two end points for the slot being assembled, fetched one slot ahead by asynchronous copies straight into
    // 18 thread-private shared-memory words
    auto stageD = [&](int w) -> double* { return sm + (VQ_SMW - 2 * NS + w) * VQ_BS; };
    auto fetchD = [&](int k) {
      const int r = k < 2 ? r0 + 4 * k : 8;
      const int base_r = r * NS - r * (r - 1) / 2 - r;
#pragma unroll
      for (int c = 0; c < NS; ++c) {
        const int up = c * NS - c * (c - 1) / 2 - c + r;
        const size_t qq = (size_t)(c >= r ? base_r + c : up) * np;
        const unsigned sa = (unsigned)__cvta_generic_to_shared(stageD(c)), sb = (unsigned)__cvta_generic_to_shared(stageD(NS + c));
        asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(sa), "l"(rdi + qq) : "memory");
        asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(sb), "l"(rdj + qq) : "memory");
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
    };
    fetchD(0);
#pragma unroll 1
    for (int k = 0; k < 3; ++k) {
      const int r = k < 2 ? r0 + 4 * k : 8;
      const double Xr = Xrow[0];
      double sD[NS];
      asm volatile("cp.async.wait_group 0;" ::: "memory");
#pragma unroll
      for (int c = 0; c < NS; ++c) sD[c] = *stageD(c) + *stageD(NS + c);
      if (k < 2) fetchD(k + 1);
      double tk = 0.0, dm[NS], gq[NS], tq[NS];
      const double sx = smm * Xr;
      bool ok = inOk && num_range(sx);   // operand ranges of the 27 quotients (rx_fdiv.cuh: fdivu)
#pragma unroll
      for (int c = 0; c < NS; ++c) ok = ok && mid_range(sD[c]);
#pragma unroll
      for (int c = 0; c < NS; ++c) dm[c] = fdivu(2.0, frecip(sD[c]));   // D_rr takes part in the maximum only
#pragma unroll
      for (int c = 0; c < NS; ++c) {
        gq[c] = fdivu(sx, frecip(rho * cc.M[c] * dm[c]));
        tq[c] = fdivu(Xs[c], frecip(dm[c]));
      }
      if (!ok) {
#pragma unroll
        for (int c = 0; c < NS; ++c) {   // out-of-range values only
          dm[c] = 2.0 / sD[c];
          gq[c] = sx / (rho * cc.M[c] * dm[c]);
          tq[c] = Xs[c] / dm[c];
        }
      }
#pragma unroll
      for (int c = 0; c < NS; ++c) {
        maxD = dm[c] > maxD ? dm[c] : maxD;   // finite positive values: the maximum without fmax's NaN handling
        smA(k, c) = gq[c];
        tk = (c != r) ? tk + tq[c] : tk;
      }
      smA(k, r) = tk;   // the diagonal word carries tmp_r until the second pass
      Xrow[0] = Xrow[1]; Xrow[1] = Xrow[2];
    }
    maxD = fmax(maxD, L.x1(maxD));
    maxD = fmax(maxD, L.x2(maxD));
    const double alpha = 1.0 / (rho * maxD);
#pragma unroll 1
    for (int k = 0; k < 3; ++k) {
      const int r = k < 2 ? r0 + 4 * k : 8;
      const double ay = alpha * Yrow[0];
      const double dg = sigma * massTot * smA(k, r) / (rho * cc.M[r]);
      smA(k, r) = dg;
#pragma unroll
      for (int c = 0; c < NS; ++c) smA(k, c) = smA(k, c) + ay;
      Yrow[0] = Yrow[1]; Yrow[1] = Yrow[2];
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) { const double dg = smA(k, row[k]); invd[k] = dg != 0.0 ? 1.0 / dg : 1.0; }
  }
  // rows of the first VQ_MREG slots move to registers; rhs, r0 and x of the solve take the (now free) staging words
  double A[VQ_MREG > 0 ? VQ_MREG : 1][NS];
#pragma unroll
  for (int k = 0; k < VQ_MREG; ++k)
#pragma unroll
    for (int c = 0; c < NS; ++c) A[k][c] = smA(k, c);
  auto getA = [&](int k, int c) -> double { return k < VQ_MREG ? A[k][c] : sm[(k * NS + c) * VQ_BS]; };
  auto s_rhs = [&](int k) -> double& { return sm[(3 * NS + k) * VQ_BS]; };
  auto s_r0 = [&](int k) -> double& { return sm[(3 * NS + 3 + k) * VQ_BS]; };
  auto s_x = [&](int k) -> double& { return sm[(3 * NS + 6 + k) * VQ_BS]; };

  // ---- Eigen::BiCGSTAB with diagonal preconditioner, tol 1e-11, zero guess, at most 2N iterations (bicgstab<> in rx_flux.cuh
  // is the serial form and carries the citations; the loop below is the pair kernel's with three slots per lane).
  double r[3], v[3], p[3], tv[3], w1[3], f[NS];
#pragma unroll
  for (int k = 0; k < 3; ++k) { s_rhs(k) = rhs[k]; s_x(k) = 0.0; r[k] = rhs[k] - 0.0; s_r0(k) = r[k]; v[k] = 0.0; p[k] = 0.0; }
  double r0_sqnorm = quad_dot(L, r, r);
  const double rhs_sqnorm = r0_sqnorm;
  double w_b = 1;
  const double tol2 = 1.0e-11 * 1.0e-11 * rhs_sqnorm;
  const double eps2 = 2.220446049250313e-16 * 2.220446049250313e-16;
  int it = 0, restarts = 0;
  bool live = (rhs_sqnorm != 0) && (r0_sqnorm > tol2);   // first loop test: r == r0
  double rho_n = quad_dot(L, r, r);   // r0 == r
  double rhoOldS = 1.0, abS = 1.0, wbS = 1.0, q2 = 1.0;
  Recip Rold = frecip(1.0);
  bool badq2 = false;
#pragma unroll 1
  while (__any_sync(VP_FULL, live)) {
    const bool restart = live && (fabs(rho_n) < eps2 * r0_sqnorm);
    if (__any_sync(VP_FULL, restart)) {   // r = rhs - A x, r0 = r (rare)
#pragma unroll
      for (int k = 0; k < 3; ++k) w1[k] = s_x(k);
      quad_gather(L, w1, f);
      quad_gemv(getA, f, tv);
      double rn[3];
#pragma unroll
      for (int k = 0; k < 3; ++k) rn[k] = s_rhs(k) - tv[k];
      const double nn = quad_dot(L, rn, rn);
      if (restart) {
#pragma unroll
        for (int k = 0; k < 3; ++k) { r[k] = rn[k]; s_r0(k) = rn[k]; }
        rho_n = r0_sqnorm = nn;
        if (restarts++ == 0) it = 0;
      }
    }
    // a converged quad keeps running with the warp on benign scalars (beta = alpha = omega = 0); nothing of it is stored
    double beta;
    {
      bool bad = badq2;
      const double bn1 = live ? rho_n : 0.0;
      const double q1 = fdiv(bn1, Rold, bad);
      beta = bad ? (bn1 / rhoOldS) * (abS / wbS) : q1 * q2;
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) { p[k] = r[k] + beta * (p[k] - w_b * v[k]); w1[k] = invd[k] * p[k]; }
    quad_gather(L, w1, f);
    quad_gemv(getA, f, v);
#pragma unroll
    for (int k = 0; k < 3; ++k) w1[k] = s_r0(k);
    const double r0v = quad_dot(L, w1, v);
    const double alpha_n = (live ? rho_n : 0.0) / (live ? r0v : 1.0);
    double s[3], z[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) { s[k] = r[k] - alpha_n * v[k]; z[k] = invd[k] * s[k]; }
    quad_gather(L, z, f);
    quad_gemv(getA, f, tv);
    const double tt = quad_dot(L, tv, tv);
    const double ts = quad_dot(L, tv, s);
    const double w_n = (live && tt > 0.0) ? ts / tt : 0.0;
#pragma unroll
    for (int k = 0; k < 3; ++k) r[k] = s[k] - w_n * tv[k];
    const double rrn = quad_dot(L, r, r);
#pragma unroll
    for (int k = 0; k < 3; ++k) w1[k] = s_r0(k);
    const double rho_next = quad_dot(L, w1, r);
    if (live) {
#pragma unroll
      for (int k = 0; k < 3; ++k) s_x(k) = s_x(k) + (alpha_n * (invd[k] * p[k]) + w_n * z[k]);
      ++it;
      live = (rrn > tol2) && (it < 2 * NS);
    }
    rhoOldS = live ? rho_n : 1.0; abS = live ? alpha_n : 0.0; wbS = live ? w_n : 1.0;
    Rold = frecip(rhoOldS);
    badq2 = false;
    q2 = fdiv(abS, wbS, badq2);
    w_b = w_n;
    rho_n = rho_next;
  }
  double x[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) x[k] = s_x(k);

  // ---- assemble the projected flux (:1530-1573).  Proj[rho] = -sum Jd (Eigen sum), Proj[E] = sum_s -h_s Jd_s in species
  // order, then the stress/heat part; species rows go out from the lane that owns them, row 8 from lane 3.
  const double sumJ = quad_redux(L, x);
  const SplPos sp = spl_pos(cc, T);
  double hsr[3], hj[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) hsr[k] = spl_eval(cc, P_H, row[k], sp);
  div_by_M_rows3(cc, row, hsr);
#pragma unroll
  for (int k = 0; k < 3; ++k) hj[k] = -hsr[k] * x[k];
  quad_gather(L, hj, f);
  double pE = 0.0;
#pragma unroll
  for (int s2 = 0; s2 < NS; ++s2) pE += f[s2];
  pE += fE0;
  pE += fE1;
  if (!keep) return;
  const long long ne = d.ne;
  if (q4 == 0) {
    d.fluxV[U_RHO * ne + e] = -sumJ;
    d.fluxV[U_E * ne + e] = pE;
  }
#pragma unroll
  for (int k = 0; k < 2; ++k) d.fluxV[(U_S + row[k]) * ne + e] = -x[k];
  if (q4 == 3) d.fluxV[(U_S + 8) * ne + e] = -x[2];
}

}  // namespace rx
