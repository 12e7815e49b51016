// rx_flux.cuh — per-edge viscous flux (Stefan–Maxwell diffusion + SST closure) and per-point Arrhenius/PaSR source.
// See rx_device.cuh for conventions.  Reference citations are relative to /root/reference.
#pragma once
#include "rx_device.cuh"

namespace rx {

// ---- Eigen 3.3.7 reductions re-enacted (externals/Eigen/src/Core/Redux.h, SSE2 packets of 2 doubles) ----------
// sum/product of n coefficients: two packet accumulators over groups of four, combined, horizontal op, scalar tail.
template <int N, bool MUL>
RX_HD double eig_redux(const double* a) {
  constexpr int A2 = (N / 4) * 4, A1 = (N / 2) * 2;
  auto op = [](double x, double y) { return MUL ? x * y : x + y; };
  double res;
  if (A1 > 0) {
    double p00 = a[0], p01 = a[1];
    if (A1 > 2) {
      double p10 = a[2], p11 = a[3];
#pragma unroll
      for (int i = 4; i < A2; i += 4) {
        p00 = op(p00, a[i]); p01 = op(p01, a[i + 1]);
        p10 = op(p10, a[i + 2]); p11 = op(p11, a[i + 3]);
      }
      p00 = op(p00, p10); p01 = op(p01, p11);
      if (A1 > A2) { p00 = op(p00, a[A2]); p01 = op(p01, a[A2 + 1]); }
    }
    res = op(p00, p01);
#pragma unroll
    for (int i = A1; i < N; ++i) res = op(res, a[i]);
  } else {
    res = a[0];
  }
  return res;
}
template <int N>
RX_HD double eig_dot(const double* a, const double* b) {
  double t[N];
#pragma unroll
  for (int i = 0; i < N; ++i) t[i] = a[i] * b[i];
  return eig_redux<N, false>(t);
}
// res = A v for a column-major N x N matrix, in the order of Eigen's column-major GEMV kernel
// (externals/Eigen/src/Core/products/GeneralMatrixVector.h:88-327; SSE2, no FMA): columns four at a time with
// res + ((a0 v0 + a1 v1) + (a2 v2 + a3 v3)) for rows < (N & ~1), sequential products for the odd last row and for the
// N % 4 leftover columns.  A(i,j) is supplied by the functor so the matrix may live in registers, shared or local memory.
template <int N, class MatF>
RX_HD void eig_gemv(MatF A, const double* v, double* res) {
  constexpr int AS = N & ~1, CB = (N / 4) * 4;
#pragma unroll
  for (int j = 0; j < N; ++j) res[j] = 0.0;
#pragma unroll
  for (int i = 0; i < CB; i += 4) {
#pragma unroll
    for (int j = 0; j < AS; ++j)
      res[j] = res[j] + ((A(j, i) * v[i] + A(j, i + 1) * v[i + 1]) + (A(j, i + 2) * v[i + 2] + A(j, i + 3) * v[i + 3]));
#pragma unroll
    for (int j = AS; j < N; ++j) {
      res[j] = A(j, i) * v[i] + res[j];
      res[j] = A(j, i + 1) * v[i + 1] + res[j];
      res[j] = A(j, i + 2) * v[i + 2] + res[j];
      res[j] = A(j, i + 3) * v[i + 3] + res[j];
    }
  }
#pragma unroll
  for (int k = CB; k < N; ++k) {
#pragma unroll
    for (int j = 0; j < N; ++j) res[j] = A(j, k) * v[k] + res[j];
  }
}

// Eigen::BiCGSTAB<MatrixXd>, DiagonalPreconditioner, tol 1e-11, zero initial guess, maxIters = 2N
// (externals/Eigen/src/IterativeLinearSolvers/BiCGSTAB.h:28-108; call site numerics_direct_reactive.cpp:451-470).
// rhs, r0 and x live in caller-provided strided scratch (element k at ptr[k*S]; shared memory on the GPU) — they are touched
// a few times per iteration only — while r, p, v, t stay in registers.
template <int N, int S, class MatF>
RX_HD int bicgstab(MatF A, const double* rhs, double* r0, double* x) {
  // y = K^-1 p and z = K^-1 s are re-formed where they are used (same bits each time), and s shares the storage of r
  // (r is dead once s exists; the new r overwrites s).
  const double tol = 1.0e-11;
  double invd[N], r[N], v[N], p[N], t[N], w1[N];
#pragma unroll
  for (int i = 0; i < N; ++i) { const double d = A(i, i); invd[i] = d != 0.0 ? 1.0 / d : 1.0; x[i * S] = 0.0; }
#pragma unroll
  for (int i = 0; i < N; ++i) { r[i] = rhs[i * S] - 0.0; r0[i * S] = r[i]; }   // r = rhs - A*0
  double r0_sqnorm = eig_dot<N>(r, r);
  const double rhs_sqnorm = r0_sqnorm;   // same products, same order
  if (rhs_sqnorm == 0) return 0;
  double rho = 1, alpha = 1, w = 1;
#pragma unroll
  for (int i = 0; i < N; ++i) { v[i] = 0.0; p[i] = 0.0; }
  const double tol2 = tol * tol * rhs_sqnorm;
  const double eps2 = 2.220446049250313e-16 * 2.220446049250313e-16;
  int it = 0, restarts = 0;
#pragma unroll 1
  while (eig_dot<N>(r, r) > tol2 && it < 2 * N) {
    const double rho_old = rho;
#pragma unroll
    for (int k = 0; k < N; ++k) w1[k] = r0[k * S];
    rho = eig_dot<N>(w1, r);
    if (fabs(rho) < eps2 * r0_sqnorm) {
#pragma unroll
      for (int k = 0; k < N; ++k) w1[k] = x[k * S];
      eig_gemv<N>(A, w1, t);
#pragma unroll
      for (int k = 0; k < N; ++k) { r[k] = rhs[k * S] - t[k]; r0[k * S] = r[k]; }
      rho = r0_sqnorm = eig_dot<N>(r, r);
      if (restarts++ == 0) it = 0;
    }
    const double beta = (rho / rho_old) * (alpha / w);
#pragma unroll
    for (int k = 0; k < N; ++k) { p[k] = r[k] + beta * (p[k] - w * v[k]); w1[k] = invd[k] * p[k]; }
    eig_gemv<N>(A, w1, v);
#pragma unroll
    for (int k = 0; k < N; ++k) w1[k] = r0[k * S];
    alpha = rho / eig_dot<N>(w1, v);
#pragma unroll
    for (int k = 0; k < N; ++k) { r[k] = r[k] - alpha * v[k]; w1[k] = invd[k] * r[k]; }   // r now holds s, w1 holds z
    eig_gemv<N>(A, w1, t);
    const double tt = eig_dot<N>(t, t);
    if (tt > 0.0) w = eig_dot<N>(t, r) / tt; else w = 0.0;
#pragma unroll
    for (int k = 0; k < N; ++k) { x[k * S] = x[k * S] + (alpha * (invd[k] * p[k]) + w * w1[k]); r[k] = r[k] - w * t[k]; }
    ++it;
  }
  return it;
}

// ---- the summation orders of Eigen::ColPivHouseholderQR<MatrixXd>::solve on a 9 x 9 matrix (Eigen 3.3.7, SSE2, no FMA) ----------
// glibc's malloc hands Eigen 16-byte aligned storage, so entry (i, j) of the 9 x 9 factor is packet-aligned exactly when i + j is even;
// the orders below follow from that one fact and are checked bit for bit against the real headers (oracle/eigen_qr_check.cpp).
//  * squaredNorm / dot of a coefficient-wise expression: eig_redux<> above with packets from entry 0 (Core/Redux.h:208-256;
//    first_default_aligned() of an expression without direct access is 0);
//  * one entry of essential^T * bottom in applyHouseholderOnTheLeft (Householder/Householder.h:122): the row-major matrix-vector
//    kernel (Core/products/GeneralMatrixVector.h:363-612) with the essential vector as right-hand side; that vector starts one
//    entry below the diagonal (entry 10k + 1 of the storage: odd), so alignedStart = 1: one scalar product, packets {1,2},{3,4},...
//    accumulated from zero, reduced, then the scalar tail.  (The kernel's leading 0 + x additions are dropped: they matter to the
//    sign of a zero only.)
template <int M>
RX_HD double eig_gemv_row(const double* t) {
  constexpr int AS = 1 + ((M - 1) & ~1);
  double r = t[0];
  if (AS > 1) {
    double p0 = t[1], p1 = t[2];
#pragma unroll
    for (int j = 3; j < AS; j += 2) { p0 += t[j]; p1 += t[j + 1]; }
    r += p0 + p1;
  }
#pragma unroll
  for (int j = AS; j < M; ++j) r += t[j];
  return r;
}
// run-time lengths (serial factorisation below); `as` = entries before the first aligned one
template <class F>
RX_HD double eig_redux_n(int n, F term) {
  if (n == 0) return 0.0;
  const int n2 = n & ~1, n4 = n & ~3;
  if (n2 == 0) return term(0);
  double a0 = term(0), a1 = term(1);
  if (n2 > 2) {
    double b0 = term(2), b1 = term(3);
    for (int i = 4; i < n4; i += 4) { a0 += term(i); a1 += term(i + 1); b0 += term(i + 2); b1 += term(i + 3); }
    a0 += b0; a1 += b1;
    if (n2 > n4) { a0 += term(n4); a1 += term(n4 + 1); }
  }
  double r = a0 + a1;
  for (int i = n2; i < n; ++i) r += term(i);
  return r;
}
template <class F>
RX_HD double eig_gemv_row_n(int m, int as, F term) {
  if (as > m) as = m;
  const int aSize = as + ((m - as) & ~1);
  double t = 0.0;
  for (int j = 0; j < as; ++j) t += term(j);
  if (aSize > as) {
    double p0 = 0.0, p1 = 0.0;
    for (int j = as; j < aSize; j += 2) { p0 += term(j); p1 += term(j + 1); }
    t += p0 + p1;
  }
  for (int j = aSize; j < m; ++j) t += term(j);
  return t;
}

// Eigen::ColPivHouseholderQR<MatrixXd>(A).solve(b) for `nrhs` right-hand sides (externals/Eigen/src/QR/ColPivHouseholderQR.h:480-609,
// Householder/Householder.h:65-131, Core/products/TriangularSolverVector.h: back substitution in panels of 8 columns, the rows above a
// panel updated by the column-major matrix-vector kernel — for the single row above the panel of a full-rank 9 x 9 system, its
// columns in ascending order).  Same pivot rule, rank threshold, norm down-dating and, sum by sum, Eigen's own order: the result
// equals the reference's bit for bit.
// A is N x N column-major in caller-provided scratch (destroyed); b/x are [nrhs][N].
// Strided view of one column of a thread-private matrix: element i lives at base[i*S] (S = 1 for local memory,
// S = block size for the shared-memory layout in which consecutive threads own consecutive 8-byte words).
template <int S>
struct ColView {
  double* p;
  RX_HD double& operator[](int i) const { return p[i * S]; }
  RX_HD ColView operator+(int k) const { return ColView{p + k * S}; }
};
template <int N, int NRHS, int S>
RX_HD void colpiv_qr_solve(double* A, const double (*b)[N], double (*xout)[N]) {
  static_assert(N <= 9, "one row at most above the 8-column panel of the back substitution");
  double hC[N], nU[N], nD[N];
  int trans[N];
  auto col = [&](int j) { return ColView<S>{A + j * N * S}; };
  for (int k = 0; k < N; ++k) {
    const ColView<S> a = col(k);
    nD[k] = nU[k] = sqrt(eig_redux_n(N, [&](int i) { return a[i] * a[i]; }));
  }
  const double eps = 2.220446049250313e-16;
  double mx = nU[0];
  for (int k = 1; k < N; ++k) mx = fmax(mx, nU[k]);
  const double threshold_helper = (mx * eps) * (mx * eps) / (double)N;
  const double downdate = sqrt(eps);
  int nonzero = N;
  for (int k = 0; k < N; ++k) {
    int big = k;
    for (int j = k + 1; j < N; ++j) if (nU[j] > nU[big]) big = j;
    const double bigsq = nU[big] * nU[big];
    if (nonzero == N && bigsq < threshold_helper * (double)(N - k)) nonzero = k;
    trans[k] = big;
    if (k != big) {
      for (int i = 0; i < N; ++i) { const double tmp = col(k)[i]; col(k)[i] = col(big)[i]; col(big)[i] = tmp; }
      double tmp = nU[k]; nU[k] = nU[big]; nU[big] = tmp;
      tmp = nD[k]; nD[k] = nD[big]; nD[big] = tmp;
    }
    const ColView<S> v = col(k) + k;
    const int m = N - k;
    const double tailSq = eig_redux_n(m - 1, [&](int i) { return v[1 + i] * v[1 + i]; });
    const double c0 = v[0];
    double beta, tau;
    if (tailSq <= 2.2250738585072014e-308) {
      tau = 0.0; beta = c0;
      for (int i = 1; i < m; ++i) v[i] = 0.0;
    } else {
      beta = sqrt(c0 * c0 + tailSq);
      if (c0 >= 0.0) beta = -beta;
      for (int i = 1; i < m; ++i) v[i] = v[i] / (c0 - beta);
      tau = (beta - c0) / beta;
    }
    hC[k] = tau;
    v[0] = beta;
    if (m == 1) {
      for (int j = k + 1; j < N; ++j) col(j)[k] *= (1.0 - tau);
    } else if (tau != 0.0) {
      for (int j = k + 1; j < N; ++j) {
        const ColView<S> w = col(j) + k;
        double t = eig_gemv_row_n(m - 1, (k * (N + 1) + 1) & 1, [&](int i) { return w[1 + i] * v[1 + i]; });
        t += w[0];
        w[0] -= tau * t;
        for (int i = 1; i < m; ++i) w[i] -= tau * v[i] * t;
      }
    }
    for (int j = k + 1; j < N; ++j) {
      if (nU[j] != 0.0) {
        double temp = fabs(col(j)[k]) / nU[j];
        temp = (1.0 + temp) * (1.0 - temp);
        temp = temp < 0.0 ? 0.0 : temp;
        const double q = nU[j] / nD[j];
        const double temp2 = temp * (q * q);
        if (temp2 <= downdate) {
          const ColView<S> a = col(j) + (k + 1);
          nD[j] = sqrt(eig_redux_n(N - k - 1, [&](int i) { return a[i] * a[i]; })); nU[j] = nD[j];
        } else nU[j] *= sqrt(temp);
      }
    }
  }
  int perm[N];
  for (int i = 0; i < N; ++i) perm[i] = i;
  for (int k = 0; k < N; ++k) { const int t = perm[k]; perm[k] = perm[trans[k]]; perm[trans[k]] = t; }
  for (int q = 0; q < NRHS; ++q) {
    double c[N];
    if (nonzero == 0) { for (int i = 0; i < N; ++i) xout[q][i] = 0.0; continue; }
    for (int i = 0; i < N; ++i) c[i] = b[q][i];
    for (int k = 0; k < nonzero; ++k) {
      const int m = N - k;
      const ColView<S> v = col(k) + k;
      const double tau = hC[k];
      double* w = c + k;
      if (m == 1) w[0] *= (1.0 - tau);
      else if (tau != 0.0) {
        double t = eig_redux_n(m - 1, [&](int i) { return v[1 + i] * w[1 + i]; });   // 1 x m by m x 1: an inner product
        t += w[0];
        w[0] -= tau * t;
        for (int i = 1; i < m; ++i) w[i] -= tau * v[i] * t;
      }
    }
    for (int pi = nonzero; pi > 0; pi -= 8) {
      const int pw = pi < 8 ? pi : 8, start = pi - pw;
      for (int kk = 0; kk < pw; ++kk) {
        const int i = pi - kk - 1;
        c[i] = c[i] / col(i)[i];
        for (int r = start; r < i; ++r) c[r] -= c[i] * col(i)[r];
      }
      for (int r = 0; r < start; ++r)
        for (int j = start; j < pi; ++j) c[r] -= col(j)[r] * c[j];
    }
    for (int i = 0; i < nonzero; ++i) xout[q][perm[i]] = c[i];
    for (int i = nonzero; i < N; ++i) xout[q][perm[i]] = 0.0;
  }
}

// ---- viscous flux ------------------------------------------------------------------------------------------
// End state of an edge (or of a boundary vertex and its ghost): primitives, gradient, transport and SST inputs.
template <int NS>
struct ViscEnd {
  double V[7 + NS];
  double g[(4 + NS) * 2];   // rows [T,u,v,P,X_s] x 2 — filled only where needed
  double X[NS];             // mole fractions of the end state (GetMolarFromMass of V[Y])
  double mu, kap;
  const double* rD;         // 1/D_st of the point that owns the binary diffusion coefficients: upper triangle s <= t,
  long long rDs;            // entry q = tri_idx(s,t) at rD[q * rDs] (read where it is used; 45 values per end are too many to hold)
  double muT, ke, gk[2], x[2];
};

// CAvgGradReactive_Flow::ComputeResidual (numerics_direct_reactive.cpp:1425-1573) / _Boundary::ComputeResidual (:478-553),
// SetLaminarTensorFlux (:1099-1190), Solve_SM (:451-470) with GetGamma (reacting_model_library.cpp:771-798),
// SST_Reactive_ResidualClosure (:656-852), Get_Molar2MassGrad_Operator (:861-880).
// `scratch` holds NS*NS + 3*NS doubles with element k at scratch[k*S] (the S–M matrix, then the molar->mass operator): S = 1 for a
// thread-local array, S = block size for the shared-memory layout.
RX_HD constexpr int tri_idx(int NS, int s, int t) { return s * NS - s * (s - 1) / 2 + (t - s); }   // s <= t
template <int NS, bool EDGE, int S, bool RANS>
RX_HD void viscous_flux(const ChemC<NS>& cc, const Opts& o, const ViscEnd<NS>& I, const ViscEnd<NS>& J, double nx, double ny,
                        double sigma_k, double* scratch, double* res) {
  constexpr int NA = 3 + NS;
  const double mu = 2.0 / (1.0 / I.mu + 1.0 / J.mu);
  const double ktr = 2.0 / (1.0 / I.kap + 1.0 / J.kap);
  double Vm[7 + NS], G[NA][2];
#pragma unroll
  for (int v = 0; v < 7 + NS; ++v) Vm[v] = 0.5 * (I.V[v] + J.V[v]);
#pragma unroll
  for (int d = 0; d < 2; ++d) {
    G[0][d] = 0.5 * (I.g[G_T * 2 + d] + J.g[G_T * 2 + d]);
    G[1][d] = 0.5 * (I.g[(G_VX)*2 + d] + J.g[(G_VX)*2 + d]);
    G[2][d] = 0.5 * (I.g[(G_VX + 1) * 2 + d] + J.g[(G_VX + 1) * 2 + d]);
#pragma unroll
    for (int s = 0; s < NS; ++s) G[3 + s][d] = 0.5 * (I.g[(G_X + s) * 2 + d] + J.g[(G_X + s) * 2 + d]);
  }
  if (EDGE) {
    const double e0 = J.x[0] - I.x[0], e1 = J.x[1] - I.x[1];
    const double dist2 = 0.0 + e0 * e0 + e1 * e1;
#pragma unroll
    for (int r = 0; r < NA; ++r) {
      const double proj = (0.0 + G[r][0] * e0) + G[r][1] * e1;
      double diff;
      if (r == 0) diff = J.V[V_T] - I.V[V_T];
      else if (r < 3) diff = J.V[V_VX + r - 1] - I.V[V_VX + r - 1];
      else diff = J.X[r - 3] - I.X[r - 3];
      G[r][0] -= ((proj - diff) * e0) / dist2;
      G[r][1] -= ((proj - diff) * e1) / dist2;
    }
  }
  const double nrm[2] = {nx, ny};
  double Proj[4 + NS], FT[3][2], tau[2][2], hs[NS], Ys[NS], Xs[NS], gxn[NS], Jd[NS];   // FT rows: mom-x, mom-y, energy
  const double rho = Vm[V_RHO], T = Vm[V_T];
  const SplPos sp = spl_pos(cc, T);
#pragma unroll
  for (int s = 0; s < NS; ++s) Ys[s] = Vm[V_Y + s];
  molar_from_mass(cc, Ys, Xs);
  double div_vel = 0.0;
  div_vel += G[1][0]; div_vel += G[2][1];
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    tau[i][0] = 0.0; tau[i][1] = 0.0;
#pragma unroll
    for (int j = 0; j < 2; ++j) tau[i][j] += mu * (G[1 + j][i] + G[1 + i][j]);
    tau[i][i] -= TWO3 * (mu * div_vel);
  }
#pragma unroll
  for (int v = 0; v < 4 + NS; ++v) Proj[v] = 0.0;
#pragma unroll
  for (int s = 0; s < NS; ++s) gxn[s] = 0.0;
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    FT[2][i] = 0.0;
#pragma unroll
    for (int j = 0; j < 2; ++j) { FT[j][i] = tau[i][j]; FT[2][i] += tau[i][j] * Vm[V_VX + j]; }
    FT[2][i] += ktr * G[0][i];
#pragma unroll
    for (int s = 0; s < NS; ++s) gxn[s] += G[3 + s][i] * nrm[i];
  }
  // ---- Stefan–Maxwell system: mean D_ij (harmonic, element-wise :1447), Gamma (:771-798), Gamma_tilde (:463-465).
  // D and Gamma's D-dependent terms are symmetric bit for bit, so each unordered pair is evaluated once; the diagonal sums
  // tmp_i = sum_{k != i} X_k/D_ik still accumulate in ascending k (pairs (k,i), k < i, are met before pairs (i,j), j > i).
  {
    double* Gam = scratch;            // column-major like Eigen::MatrixXd: Gam(r,c) at scratch[(c*NS + r)*S]
    double* Xsm = scratch + NS * NS * S;   // copy of Xs for the rolled pair loop (dynamic index)
    double sigma = 0.0, massTot = 0.0;
#pragma unroll
    for (int s = 0; s < NS; ++s) sigma = sigma + Ys[s];
#pragma unroll
    for (int s = 0; s < NS; ++s) massTot += Ys[s] / cc.M[s];
    massTot = 1.0 / massTot;
#pragma unroll
    for (int s = 0; s < NS; ++s) { Xsm[s * S] = Xs[s]; Gam[(s * NS + s) * S] = 0.0; }   // diagonal slot holds tmp_s meanwhile
    double maxD = 0.0;
    const double sm = -sigma * massTot;
    const double* rdi = I.rD;
    const double* rdj = J.rD;
    // the pair loops stay rolled: 36 pairs x 5 IEEE divisions unrolled would be most of the kernel's code
#pragma unroll 1
    for (int i = 0; i < NS; ++i) {
      const double Xi = Xsm[i * S], Mi = cc.M[i];
      double ti = Gam[(i * NS + i) * S];
      maxD = fmax(maxD, 2.0 / (*rdi + *rdj));   // D_ii takes part in the maximum only
      rdi += I.rDs; rdj += J.rDs;
#pragma unroll 1
      for (int j = i + 1; j < NS; ++j) {
        const double dm = 2.0 / (*rdi + *rdj);   // 2/(1/D_i + 1/D_j)
        rdi += I.rDs; rdj += J.rDs;
        maxD = fmax(maxD, dm);
        const double Xj = Xsm[j * S];
        Gam[(j * NS + i) * S] = sm * Xi / (rho * cc.M[j] * dm);
        Gam[(i * NS + j) * S] = sm * Xj / (rho * Mi * dm);
        ti += Xj / dm;
        Gam[(j * NS + j) * S] += Xi / dm;
      }
      Gam[(i * NS + i) * S] = ti;
    }
    const double alpha = 1.0 / (rho * maxD);
#pragma unroll 1
    for (int i = 0; i < NS; ++i) Gam[(i * NS + i) * S] = sigma * massTot * Gam[(i * NS + i) * S] / (rho * cc.M[i]);
#pragma unroll 1
    for (int j = 0; j < NS; ++j) {
#pragma unroll
      for (int i = 0; i < NS; ++i) Gam[(j * NS + i) * S] = Gam[(j * NS + i) * S] + alpha * Ys[i];
    }
    double* rhs = Xsm;                        // Xs is no longer needed in scratch
    double* r0s = scratch + (NS * NS + NS) * S;
    double* xs = scratch + (NS * NS + 2 * NS) * S;
#pragma unroll
    for (int s = 0; s < NS; ++s) rhs[s * S] = -gxn[s];
    bicgstab<NS, S>([&](int r, int c) { return Gam[(c * NS + r) * S]; }, rhs, r0s, xs);
#pragma unroll
    for (int s = 0; s < NS; ++s) Jd[s] = xs[s * S];
  }
  Proj[U_RHO] = -eig_redux<NS, false>(Jd);
#pragma unroll
  for (int s = 0; s < NS; ++s) hs[s] = spl_eval(cc, P_H, s, sp) / cc.M[s];   // needed from here on only
#pragma unroll
  for (int s = 0; s < NS; ++s) { Proj[U_E] += -hs[s] * Jd[s]; Proj[U_S + s] = -Jd[s]; }
  // ---- SST closure
  if (RANS) {
    const double muT = 2.0 / (1.0 / I.muT + 1.0 / J.muT);
    const double ke = 0.5 * (I.ke + J.ke);
    const double gk[2] = {0.5 * (I.gk[0] + J.gk[0]), 0.5 * (I.gk[1] + J.gk[1])};
    const double PrT = o.prandtlTurb, LeT = o.lewisTurb;
    double Cps[NS], tt[2][2], rhsq[2][NS], MG[2][NS];
#pragma unroll
    for (int s = 0; s < NS; ++s) Cps[s] = spl_eval(cc, P_CP, s, sp) / cc.M[s];
    double dv = 0.0;
    dv += G[1][0]; dv += G[2][1];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      tt[i][0] = 0.0; tt[i][1] = 0.0;
#pragma unroll
      for (int j = 0; j < 2; ++j) tt[i][j] += muT * (G[1 + j][i] + G[1 + i][j]);
      tt[i][i] -= TWO3 * (muT * dv + ke * rho);
    }
    double sig = 0.0;
#pragma unroll
    for (int s = 0; s < NS; ++s) sig = sig + Xs[s];
    double* Mt = scratch;
    for (int i = 0; i < NS; ++i)
      for (int j = 0; j < NS; ++j)
        Mt[(j * NS + i) * S] = (i == j) ? cc.Mtot / cc.M[i] * (Ys[i] - Xs[i] + sig) * 1.0 + cc.Mtot * (Ys[i] / cc.M[i] - Xs[i] / cc.M[j]) * 0.0
                                  : cc.Mtot / cc.M[i] * (Ys[i] - Xs[i] + sig) * 0.0 + cc.Mtot * (Ys[i] / cc.M[i] - Xs[i] / cc.M[j]) * 1.0;
#pragma unroll
    for (int d = 0; d < 2; ++d)
#pragma unroll
      for (int s = 0; s < NS; ++s) rhsq[d][s] = G[3 + s][d];
    colpiv_qr_solve<NS, 2, S>(Mt, rhsq, MG);
#pragma unroll
    for (int d = 0; d < 2; ++d)
#pragma unroll
      for (int s = 0; s < NS; ++s) if (fabs(G[3 + s][d]) < 1e-8) MG[d][s] = 0.0;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
#pragma unroll
      for (int j = 0; j < 2; ++j) { FT[j][i] += tt[i][j]; FT[2][i] += tt[i][j] * Vm[V_VX + j]; }
#pragma unroll
      for (int s = 0; s < NS; ++s) Proj[U_S + s] += muT / (PrT * LeT) * MG[i][s] * nrm[i];
#pragma unroll
      for (int s = 0; s < NS; ++s) FT[2][i] += muT / (PrT * LeT) * hs[s] * Ys[s] * MG[i][s];
#pragma unroll
      for (int s = 0; s < NS; ++s) FT[2][i] += muT / PrT * Cps[s] * Ys[s] * G[0][i];
      FT[2][i] += (mu + muT / sigma_k) * gk[i];
    }
  }
#pragma unroll
  for (int d = 0; d < 2; ++d) {
    Proj[U_MX] += FT[0][d] * nrm[d];
    Proj[U_MX + 1] += FT[1][d] * nrm[d];
    Proj[U_E] += FT[2][d] * nrm[d];
  }
#pragma unroll
  for (int v = 0; v < 4 + NS; ++v) res[v] = Proj[v];
}

// ---- Arrhenius + PaSR source (reacting_model_library.cpp:99-227, 803-920; numerics_direct_reactive.cpp:1728-1825) ----
// std::pow(c, e) (Eigen's ArrayXd::pow, reacting_model_library.cpp:893-895, 910-912; std::pow(temp, beta) :840): exponents 0 and 1
// are exact in glibc (1 and c), everything else goes through the glibc-exact pow of rx_libm.cuh
RX_HD double pow_exp(double c, double e) {
  if (e == 0.0) return 1.0;
  if (e == 1.0) return c;
  return rx_pow(c, e);
}
template <int NS>
RX_HD void mass_production(const ChemC<NS>& cc, bool pasr, double cMu, double pasrLb, double T, double rho, const double* Yin,
                           double omegaTurb, double* omega) {
  const int Nr = cc.Nr;
  double Ys[NS], Cs[NS], tmp[NS];
#pragma unroll
  for (int s = 0; s < NS; ++s) omega[s] = 0.0;
  if (Nr == 0) return;
#pragma unroll
  for (int s = 0; s < NS; ++s) { Ys[s] = clipy<NS>(Yin[s]); Cs[s] = 1.0e3 * rho * Ys[s]; }   // SetConcentration :701-705
  div_by_M(cc, Cs);
  // the per-reaction production terms are folded into omega as they are formed, in reaction order — the order of the sums in
  // GetMassProductionTerm (:143-154 with the PaSR constants, :196-202 without) — so no per-reaction array is kept
  for (int r = 0; r < Nr; ++r) {
    // ComputeRateConstants :835-867 (K_eq from Gibbs only decides kb when no backward data are available)
    const double kf = cc.As[r] * pow_exp(T, cc.Betas[r]) * rx_exp(-cc.Ta[r] / T);
    double kb;
    if (!cc.avb[r]) {
      if (!cc.rev[r]) kb = 0.0;
      else {
        const SplPos sp = spl_pos(cc, T);
        double dG = 0.0, dnu = 0.0;
        for (int s = 0; s < NS; ++s) {
          const double dc = cc.nuP[r][s] - cc.nuR[r][s];
          if (dc != 0.0) { dG += dc * (spl_eval(cc, P_H, s, sp) - T * spl_eval(cc, P_S, s, sp)); dnu += dc; }
        }
        const double lnKp = -dG / (R_UNGAS * T);
        const double lnKc = lnKp - dnu * rx_log(R_UNGAS_ATM * T);
        kb = (rx_exp(lnKp) > 1.0e10) ? 0.0 : kf / rx_exp(lnKc);
      }
    } else kb = cc.AsB[r] * pow_exp(T, cc.BetasB[r]) * rx_exp(-cc.TaB[r] / T);
    // SetReactionRates :872-920
    double fr = 0.0, br = 0.0;
    bool zero = false;
#pragma unroll
    for (int s = 0; s < NS; ++s) zero = zero || (cc.expR[r][s] < 0.0 && Ys[s] < 1.0e-15);
    if (!zero) {
#pragma unroll
      for (int s = 0; s < NS; ++s) tmp[s] = pow_exp(Cs[s], cc.expR[r][s]);
      fr = eig_redux<NS, true>(tmp);
      fr *= kf;
    }
    zero = false;
#pragma unroll
    for (int s = 0; s < NS; ++s) zero = zero || (cc.expP[r][s] < 0.0 && Ys[s] < 1.0e-15);
    if (!zero) {
#pragma unroll
      for (int s = 0; s < NS; ++s) tmp[s] = pow_exp(Cs[s], cc.expP[r][s]);
      br = eig_redux<NS, true>(tmp);
      br *= kb;
    }
    double kpr = 1.0;
    if (pasr) {   // Set_DfrDrhos :122-136, GetTimeCombustion_r :208-227, AssemblePaSRConstant :161-190
      double hi = 0.0;
      bool first = true;
#pragma unroll
      for (int s = 0; s < NS; ++s)
        if (cc.nuP[r][s] != 0.0 || cc.nuR[r][s] != 0.0) {
          const double d = (Ys[s] > 1.0e-10) ? (fr * cc.expR[r][s] - br * cc.expP[r][s]) / (rho * Ys[s]) : 0.0;
          const double v = fabs(d * cc.M[s]);
          if (first || v > hi) { hi = v; first = false; }
        }
      const double tauMix = 1 / (cMu * omegaTurb);
      const double tauC = 1 / hi;
      if (isinf(tauC)) kpr = 1.0;
      else if ((tauC / (tauC + tauMix)) < pasrLb) kpr = pasrLb;
      else kpr = tauC / (tauC + tauMix);
    }
#pragma unroll
    for (int s = 0; s < NS; ++s) {
      const double w = 1.0e-3 * cc.M[s] * (cc.nuP[r][s] - cc.nuR[r][s]) * (fr - br);   // SetSourceTerm :99-114
      if (pasr) omega[s] = omega[s] + kpr * w;          // w = 0; w += kp[r] * wir[r][s]
      else omega[s] = (r == 0) ? w : omega[s] + w;      // w = wir[0][s]; w = w + wir[r][s]
    }
  }
}

}  // namespace rx
