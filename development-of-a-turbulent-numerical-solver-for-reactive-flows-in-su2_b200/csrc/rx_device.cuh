// rx_device.cuh — per-point / per-edge FP64 device arithmetic of the reactive residual path (sm_100a).
//
// One thread owns one point or one edge; all per-species work (NS is a template constant) is unrolled into registers.
// Compiled with -fmad=false: where the reference's result is tolerance-limited (secant temperature solve, BiCGSTAB
// Stefan–Maxwell solve) the operation order below is the reference's, so iteration counts agree with the CPU code.
// The same header compiles for the host (RX_HD plain inline) ONLY in tests/host_emu.cpp (tests/test_host_emulation.py): the arithmetic checked without a GPU.
//
// Reference citations are relative to /root/reference.
#pragma once
#include <math.h>
#include <stdint.h>
#ifdef __CUDACC__
#include "rx_fdiv.cuh"
#include "rx_libm.cuh"
#endif

#if defined(__CUDACC__)
#define RX_HD __host__ __device__ __forceinline__
#define RX_D __device__ __forceinline__
#else
#define RX_HD inline
#define RX_D inline
#endif

namespace rx {

constexpr int NRMAX = 4;            // reactions held in the by-value kernel parameter block
constexpr double EPS = 1.0e-16;     // Common/include/option_structure.hpp:134
constexpr double TWO3 = 2.0 / 3.0;  // option_structure.hpp:138
constexpr double R_UNGAS = (6.02214129 * 1.0e23) * (1.3806488 * 1.0e-23) * 1.0e3;  // physical_chemical_library.hpp:571-575
constexpr double R_UNGAS_ATM = 1.0e-3 * 0.082057338;                               // :579

enum { P_MU = 0, P_KAPPA = 1, P_CP = 2, P_H = 3, P_S = 4 };
// U = [rho, rho u, rho v, rho E, rho_s]; V = [T, u, v, P, rho, h, a, Y_s]; grad rows = [T, u, v, P, X_s]
enum { U_RHO = 0, U_MX = 1, U_E = 3, U_S = 4 };
enum { V_T = 0, V_VX = 1, V_P = 3, V_RHO = 4, V_H = 5, V_A = 6, V_Y = 7 };
enum { G_T = 0, G_VX = 1, G_P = 3, G_X = 4 };

// std::exp / std::pow / std::log of the reference: on the device the glibc-exact restatements of rx_libm.cuh (same bits as the
// host libm the CPU code calls), on the host (oracle-side emulation only) libm itself
RX_HD double rx_exp(double x) {
#ifdef __CUDA_ARCH__
  return gl_exp(x);
#else
  return exp(x);
#endif
}
RX_HD double rx_pow(double x, double y) {
#ifdef __CUDA_ARCH__
  return gl_pow(x, y);
#else
  return pow(x, y);
#endif
}
RX_HD double rx_log(double x) {
#ifdef __CUDA_ARCH__
  return gl_log(x);
#else
  return log(x);
#endif
}

struct Knot { double y, y2; };

// Mechanism + mixture constants, passed BY VALUE as a __grid_constant__ kernel parameter (lives in the constant bank).
// Host-side derived constants (pow/sqrt/cbrt of molar masses and diffusion volumes) are evaluated with the host libm,
// exactly the calls the reference makes per point (reacting_model_library.cpp:634-696,751-766).
template <int NS>
struct ChemC {
  int Nr, nKnots;
  const double* x;     // [nKnots] temperature grid (device)
  const Knot* tab;     // [5][nKnots][NS] interleaved {y, y2} (device; ~0.8 MB, L2 resident)
  double M[NS], Ri[NS];
  double rM[NS];         // refined reciprocal of M as frecip() forms it (filled on the device at create, rx_fdiv.cuh)
  double wPow[NS][NS];   // pow(M[j]/M[i], 0.25)
  double wSq[NS][NS];    // sqrt(8*(1 + M[i]/M[j]))
  double rwSq[NS][NS];   // refined reciprocal of wSq as frecip() forms it (filled on the device at create, rx_fdiv.cuh)
  double Mij[NS][NS];    // sqrt(M_i M_j/(M_i+M_j))
  double dvs[NS][NS];    // cbrt(dv_i) + cbrt(dv_j)
  double Mtot;           // sum of molar masses (numerics_direct_reactive.cpp:665)
  double As[NRMAX], Betas[NRMAX], Ta[NRMAX], AsB[NRMAX], BetasB[NRMAX], TaB[NRMAX];
  int rev[NRMAX], avb[NRMAX];
  double nuR[NRMAX][NS], nuP[NRMAX][NS], expR[NRMAX][NS], expP[NRMAX][NS];
};

struct Opts {
  int viscous, rans, leastSquares;
  double mInfty, cfl, maxDt, Tmin, Tmax;
  int clipT, ignition; long long ignitionIter; double ignitionT; int fuelIndex, oxidizerIndex;
  double prandtlLam, prandtlTurb, lewisTurb, cMu, pasrLb, tkeInf;
  int secondOrder, limiter;
  double limCoeff, refElemLength;
  int timeStepping; double unstCfl, deltaUnstTime;
  int dualTime; double deltaUnstTimeND;
  int viscLimiter;
};

// ---- spline (MathTools::GetSpline, Common/src/Tools/spline.cpp:61-77) ---------------------------
struct SplPos { int klo; double a, b, h; bool ok; };

template <int NS>
RX_HD SplPos spl_pos(const ChemC<NS>& cc, double value) {
  SplPos s;
  const double x0 = cc.x[0], xl = cc.x[cc.nKnots - 1];
  s.ok = !(value < x0 || value > xl);   // the reference throws std::out_of_range here (:63-64)
  if (!s.ok) value = fmin(fmax(value, x0), xl);
  s.h = cc.x[1] - x0;
  long long k = (long long)((value - x0) / s.h + 1);
  if (k > cc.nKnots - 1) k = cc.nKnots - 1;   // value == x.back(): the reference reads past the table; out of contract
  if (k < 1) k = 1;
  s.klo = (int)k;
  s.a = (cc.x[k] - value) / s.h;
  s.b = (value - cc.x[k - 1]) / s.h;
  return s;
}
template <int NS>
RX_HD double spl_eval(const ChemC<NS>& cc, int prop, int sp, const SplPos& s) {
  const Knot lo = cc.tab[((size_t)prop * cc.nKnots + (s.klo - 1)) * NS + sp];
  const Knot hi = cc.tab[((size_t)prop * cc.nKnots + s.klo) * NS + sp];
  return s.a * lo.y + s.b * hi.y + ((s.a * s.a * s.a - s.a) * lo.y2 + (s.b * s.b * s.b - s.b) * hi.y2) * (s.h * s.h) / 6.0;
}

// ---- mixture thermodynamics (reacting_model_library.cpp) -----------------------------------------
template <int NS> RX_HD double clipy(double y) { return y < 0.0 ? 1.0e-30 : y; }   // SetMassFractions :66-80

template <int NS>
RX_HD double rgas(const ChemC<NS>& cc, const double* Y) {   // ComputeRgas :27-40
  double r = 0.0;
#pragma unroll
  for (int s = 0; s < NS; ++s) r = r + clipy<NS>(Y[s]) * cc.Ri[s];
  return r;
}
// a[s] <- a[s] / M[s] for every species.  The reference divides per-kmol table values and mass fractions by the molar mass
// all over the path; on the device the quotients use the molar masses' reciprocals formed once at create (rx_fdiv.cuh:
// fdivu is the correctly rounded quotient for numerators that are zero or in [2^-600, 2^201); anything else divides plainly).
template <int NS>
RX_HD void div_by_M(const ChemC<NS>& cc, double* a) {
#ifdef __CUDA_ARCH__
  bool ok = true;
#pragma unroll
  for (int s = 0; s < NS; ++s) ok = ok && num_range(a[s]);
  if (ok) {
#pragma unroll
    for (int s = 0; s < NS; ++s) a[s] = fdivu(a[s], Recip{cc.M[s], cc.rM[s]});
    return;
  }
#endif
#pragma unroll
  for (int s = 0; s < NS; ++s) a[s] = a[s] / cc.M[s];
}
template <int NS>
RX_HD double enthalpy(const ChemC<NS>& cc, double T, const double* Y, bool& ok) {   // ComputeEnthalpy :519-523
  SplPos sp = spl_pos(cc, T);
  ok = ok && sp.ok;
  double h = 0.0, hs[NS];
#pragma unroll
  for (int s = 0; s < NS; ++s) hs[s] = spl_eval(cc, P_H, s, sp);
  div_by_M(cc, hs);
#pragma unroll
  for (int s = 0; s < NS; ++s) h = h + clipy<NS>(Y[s]) * hs[s];
  return h;
}
template <int NS>
RX_HD double frozen_gamma(const ChemC<NS>& cc, double T, const double* Y) {   // ComputeFrozenGamma :399-404, ComputeCP :612-619
  SplPos sp = spl_pos(cc, T);
  double cp = 0.0, cps[NS];
#pragma unroll
  for (int s = 0; s < NS; ++s) cps[s] = spl_eval(cc, P_CP, s, sp);
  div_by_M(cc, cps);
#pragma unroll
  for (int s = 0; s < NS; ++s) cp = cp + clipy<NS>(Y[s]) * cps[s];
  double R = rgas(cc, Y);
  double cv = cp - R;
  return cp / cv;
}
template <int NS>
RX_HD void molar_from_mass(const ChemC<NS>& cc, const double* Y, double* X) {   // SetMolarFromMass :84-93
  double sy = 0.0, sx = 0.0;
#pragma unroll
  for (int s = 0; s < NS; ++s) X[s] = clipy<NS>(Y[s]);
  div_by_M(cc, X);
#pragma unroll
  for (int s = 0; s < NS; ++s) sy = sy + clipy<NS>(Y[s]);
#pragma unroll
  for (int s = 0; s < NS; ++s) sx = sx + X[s];
  const double massTot = sy / sx;
#pragma unroll
  for (int s = 0; s < NS; ++s) X[s] = massTot * X[s];
}

// Wilke viscosity and conductivity mixing rules in one pass (ComputeEta :634-656, ComputeLambda :671-696).
// The reference evaluates the same pair factor t_ij twice (once per rule); it is computed once here — same bits.
// The outer species loop stays rolled (81 pair factors with a division, a square root and two more divisions each would
// otherwise be ~5k instructions); the i-indexed operands are rotated through registers so that no array is indexed dynamically.
template <int NS>
RX_HD void transport(const ChemC<NS>& cc, double T, const double* Y, double& muMix, double& kMix) {
  SplPos sp = spl_pos(cc, T);
  double mu[NS], yomC[NS], yomR[NS], muR[NS], kaR[NS], ycR[NS], yrR[NS];
#pragma unroll
  for (int s = 0; s < NS; ++s) {
    mu[s] = spl_eval(cc, P_MU, s, sp);
    kaR[s] = spl_eval(cc, P_KAPPA, s, sp);
    yomC[s] = clipy<NS>(Y[s]) / cc.M[s];   // ComputeEta uses the clipped Ys
    yomR[s] = Y[s] / cc.M[s];              // ComputeLambda uses the raw ys (:679)
    muR[s] = mu[s]; ycR[s] = yomC[s]; yrR[s] = yomR[s];
  }
  muMix = 0.0; kMix = 0.0;
#pragma unroll 1
  for (int i = 0; i < NS; ++i) {
    const double mui = muR[0], kai = kaR[0], yci = ycR[0], yri = yrR[0];
    double phiE = 0.0, phiL = 0.0;
#pragma unroll
    for (int j = 0; j < NS; ++j) {
      const double t = 1.0 + sqrt(mui / mu[j]) * cc.wPow[i][j];
      phiE += yomC[j] / cc.wSq[i][j] * t * t;
      const double add = 1.065 * yomR[j] / cc.wSq[i][j] * t * t;
      phiL = (j != i) ? phiL + add : phiL;
    }
    phiL += yri;
    muMix += mui * yci / phiE;
    kMix += kai * yri / phiL;
#pragma unroll
    for (int m = 0; m + 1 < NS; ++m) { muR[m] = muR[m + 1]; kaR[m] = kaR[m + 1]; ycR[m] = ycR[m + 1]; yrR[m] = yrR[m + 1]; }
  }
}

// std::pow(temp, 1.75) of GetDij_SM (reacting_model_library.cpp:751-766) with glibc's own bits (rx_libm.cuh).  A one-ulp change of
// the diffusion coefficients is amplified by the BiCGSTAB iteration of the Stefan-Maxwell solve to ~1e-11 of the diffusion flux;
// CUDA's pow (2 ulp) and even the correctly rounded value (glibc's differs from it for 0.08 % of temperatures) are not enough.
RX_HD double pow175(double a) { return rx_pow(a, 1.75); }

// ---- cons -> prim (CReactiveEulerVariable::Cons2PrimVar, SU2_CFD/src/variable_direct_reactive.cpp:550-778) ----
// U is modified in place where the reference modifies Solution (:571-583). V[V_T] holds the secant start value on entry.
template <int NS>
RX_HD bool cons2prim(const ChemC<NS>& cc, const Opts& o, long long extIter, double* U, double* V, double ke, bool* solveFailed = nullptr) {
  bool nonPhys = false;
#pragma unroll
  for (int s = 0; s < NS; ++s)
    if (U[U_S + s] < 0.0) { U[U_S + s] = 1.0e-30; nonPhys = true; }
  if (U[U_RHO] < EPS) { V[V_RHO] = U[U_RHO] = EPS; nonPhys = true; }
  else V[V_RHO] = U[U_RHO];
  double* Y = V + V_Y;
  double sumY = 0.0;
#pragma unroll
  for (int s = 0; s < NS; ++s) { Y[s] = U[U_S + s] / U[U_RHO]; }
#pragma unroll
  for (int s = 0; s < NS; ++s) sumY = sumY + Y[s];
  nonPhys = nonPhys || (fabs(sumY - 1.0) > 0.1);
  const double rho = U[U_RHO];
  const double rhoE = U[U_E] - rho * ke;
  double sqvel = 0.0;
#pragma unroll
  for (int d = 0; d < 2; ++d) { V[V_VX + d] = U[U_MX + d] / rho; sqvel += V[V_VX + d] * V[V_VX + d]; }
  const double Tmin = o.Tmin, Tmax = o.Tmax;
  const double NRtol = 1.0e-6, Btol = 1.0e-4;
  bool NRconvg = false;
  const double Rg = rgas(cc, Y);
  const double C1 = (-rhoE + 0.5 * rho * sqvel) / (rho * Rg);
  const double C2 = 1.0 / Rg;
  const double old_temp = V[V_T];
  double T = V[V_T], Told = T + 1.0, Tnew, f, df, hs, hs_old;
  for (int it = 0; it < 7; ++it) {
    bool ok = true;
    hs_old = enthalpy(cc, Told, Y, ok);
    hs = enthalpy(cc, T, Y, ok);
    if (!ok) {   // catch(std::out_of_range) :656-688 — bisection with up to 10000 halvings
      double Ta = Tmin, Tb = Tmax;
      for (int b = 0; b < 10000; ++b) {
        T = (Ta + Tb) / 2.0;
        bool dummy = true;
        hs = enthalpy(cc, T, Y, dummy);
        f = T - C1 - C2 * hs;
        if (fabs(f) < Btol) { NRconvg = true; break; }
        if (f > 0) Ta = T; else Tb = T;
      }
      if (!NRconvg && solveFailed) *solveFailed = true;   // "Convergence not achieved for bisection method after catching out of range" (:683)
      break;
    }
    f = T - C1 - C2 * hs;
    df = T - Told + C2 * (hs_old - hs);
    Tnew = T - f * (T - Told) / df;
    if (fabs(Tnew - T) < NRtol) { NRconvg = true; break; }
    Told = T; T = Tnew;
  }
  if (NRconvg) V[V_T] = T;
  else {
    double Ta = Tmin, Tb = Tmax;
    bool Bconvg = false;
    for (int b = 0; b < 32; ++b) {
      T = (Ta + Tb) / 2.0;
      bool dummy = true;
      hs = enthalpy(cc, T, Y, dummy);
      f = T - C1 - C2 * hs;
      if (fabs(f) < Btol) { V[V_T] = T; Bconvg = true; break; }
      if (f > 0) Ta = T; else Tb = T;
    }
    if (!Bconvg && solveFailed) *solveFailed = true;   // "Convergence not achieved for bisection method" (:724)
  }
  if (extIter > 0 && o.clipT) V[V_T] = fmin(fmax(V[V_T], 0.95 * old_temp), 1.05 * old_temp);
  if (V[V_T] < Tmin) { V[V_T] = Tmin; nonPhys = true; }
  else if (V[V_T] > Tmax) { V[V_T] = Tmax; nonPhys = true; }
  T = V[V_T];
  V[V_P] = rho * Rg * T;
  if (V[V_P] < EPS) { V[V_P] = EPS; nonPhys = true; }
  const double g = frozen_gamma(cc, T, Y);
  V[V_A] = sqrt(g * V[V_P] / rho);
  if (V[V_A] < EPS) { V[V_A] = EPS; nonPhys = true; }
  V[V_H] = (U[U_E] + V[V_P]) / rho;
  return nonPhys;
}

// ---- AUSM+-up (CUpwReactiveAUSM::ComputeResidual, SU2_CFD/src/numerics_direct_reactive.cpp:53-189) --------
// The interface quantities every row of the flux shares: mass flux M12, its modulus, the pressure flux, the unit normal and the area.
struct AusmIface { double M12, aM, pLF, un0, un1, Area; };
RX_HD AusmIface ausm_iface(double mInfty, double rho_i, double u_i, double v_i, double P_i, double a_i,
                           double rho_j, double u_j, double v_j, double P_j, double a_j, double nx, double ny) {
  AusmIface f;
  f.Area = sqrt(0.0 + nx * nx + ny * ny);
  f.un0 = nx / f.Area; f.un1 = ny / f.Area;
  double pv_i = 0.0, pv_j = 0.0;
  pv_i += u_i * f.un0; pv_j += u_j * f.un0;
  pv_i += v_i * f.un1; pv_j += v_j * f.un1;
  const double am = 0.5 * (a_i + a_j);
  const double mL = pv_i / am, mR = pv_j / am;
  const double mF2 = 0.5 * (mL * mL + mR * mR);
  const double mRef2 = fmin(1.0, fmax(mF2, mInfty * mInfty));
  const double mRef = sqrt(mRef2);
  const double fa = mRef * (2.0 - mRef);
  const double alpha = 3.0 / 16.0 * (5.0 * fa * fa - 4.0);
  const double beta = 0.125;
  double mLP, mRM, pLP, pRM;
  if (fabs(mL) < 1.0) {
    mLP = 0.25 * (mL + 1.0) * (mL + 1.0) + beta * (mL * mL - 1.0) * (mL * mL - 1.0);
    pLP = 0.25 * (mL + 1.0) * (mL + 1.0) * (2.0 - mL) + alpha * mL * (mL * mL - 1.0) * (mL * mL - 1.0);
  } else {
    mLP = 0.5 * (mL + fabs(mL));
    pLP = 0.5 * (1.0 + fabs(mL) / mL);
  }
  if (fabs(mR) < 1.0) {
    mRM = -0.25 * (mR - 1.0) * (mR - 1.0) - beta * (mR * mR - 1.0) * (mR * mR - 1.0);
    pRM = 0.25 * (mR - 1.0) * (mR - 1.0) * (2.0 + mR) - alpha * mR * (mR * mR - 1.0) * (mR * mR - 1.0);
  } else {
    mRM = 0.5 * (mR - fabs(mR));
    pRM = 0.5 * (1.0 - fabs(mR) / mR);
  }
  const double kP = 0.25, sigma = 1.0;
  double m12 = mLP + mRM;
  m12 -= kP / fa * fmax(1.0 - sigma * mF2, 0.0) * (P_j - P_i) / (0.5 * (rho_i + rho_j) * am * am);
  const double mLF = 0.5 * (m12 + fabs(m12));
  const double mRF = 0.5 * (m12 - fabs(m12));
  f.M12 = am * (mLF * rho_i + mRF * rho_j);
  f.aM = fabs(f.M12);
  const double Ku = 0.75;
  double pLF = pLP * P_i + pRM * P_j;
  pLF -= Ku * pLP * pRM * (rho_i + rho_j) * fa * am * (pv_j - pv_i);
  f.pLF = pLF;
  return f;
}
// one transported row: 0.5 (M12 (phi_i + phi_j) + |M12| (phi_i - phi_j)) Area
RX_HD double ausm_row(const AusmIface& f, double phi_i, double phi_j) { return 0.5 * (f.M12 * (phi_i + phi_j) + f.aM * (phi_i - phi_j)) * f.Area; }
template <int NS>
RX_HD void ausm(double mInfty, const double* Vi, const double* Vj, double nx, double ny, double* res) {
  const AusmIface f = ausm_iface(mInfty, Vi[V_RHO], Vi[V_VX], Vi[V_VX + 1], Vi[V_P], Vi[V_A], Vj[V_RHO], Vj[V_VX], Vj[V_VX + 1], Vj[V_P], Vj[V_A], nx, ny);
  res[U_RHO] = ausm_row(f, 1.0, 1.0);
  res[U_MX] = ausm_row(f, Vi[V_VX], Vj[V_VX]);
  res[U_MX + 1] = ausm_row(f, Vi[V_VX + 1], Vj[V_VX + 1]);
  res[U_E] = ausm_row(f, Vi[V_H], Vj[V_H]);
#pragma unroll
  for (int s = 0; s < NS; ++s) res[U_S + s] = ausm_row(f, Vi[V_Y + s], Vj[V_Y + s]);
  res[U_MX] += f.pLF * f.un0 * f.Area;
  res[U_MX + 1] += f.pLF * f.un1 * f.Area;
}

}  // namespace rx
