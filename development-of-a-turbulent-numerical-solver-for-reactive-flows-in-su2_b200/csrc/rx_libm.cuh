// rx_libm.cuh — exp, pow and log that return the SAME BITS as the CPU reference's libm calls.
//
// The reference evaluates the Arrhenius rates with std::exp / std::pow (reacting_model_library.cpp:835-920), the binary
// diffusion coefficients with std::pow(T, 1.75) (:751-766) and the subsonic outlet state with std::pow
// (solver_direct_reactive.cpp:3901-3935).  On the x86-64 hosts of this image those calls land in glibc 2.39's __exp_fma /
// __pow_fma / __log_fma (the ifunc variants chosen whenever the CPU has FMA + AVX2): table-driven algorithms accurate to
// ~0.51 ulp, NOT correctly rounded — so neither CUDA's exp/pow (1-2 ulp) nor a correctly rounded evaluation returns the same
// value at every point, and a one-ulp difference of a reaction rate or of D_ij is what separates two otherwise bit-identical
// 500-iteration runs (profiles/r2_longrun_drift.json).  The functions below restate the instruction sequences of those three
// glibc routines operation by operation (every fused multiply-add glibc's binary performs is an explicit fma() here, every
// separate multiply / add stays separate: the library is compiled with -fmad=false), over the same tables
// (rx_libm_tables.cuh, read out of the image's libm.so.6 by tools/extract_libm_tables.py).
//
// Scope: the main path of each routine (finite positive arguments, results in the normal range).  Arguments that glibc sends
// to its special-case code (zero / negative / subnormal / infinite bases, |y log x| >= 512 or < 2^-54, ...) fall back to
// CUDA's exp / pow / log; none of them occurs on the path for physical states.
// tests/test_gpu_libm.py checks bit equality with the host's libm over 10^6 arguments per function.
#pragma once
#include "rx_libm_tables.cuh"

namespace rx {

__device__ __forceinline__ double gl_u2d(unsigned long long u) { return __longlong_as_double((long long)u); }
__device__ __forceinline__ unsigned long long gl_d2u(double d) { return (unsigned long long)__double_as_longlong(d); }

// exp_inline of e_pow.c / the body of e_exp.c: exp(x + xtail) for 2^-54 <= |x| < 512; `ok` = false outside (caller falls back)
// `wide`: also 512 <= |x| < 1024 through specialcase() of e_exp.c (scale applied in two steps), where the result stays normal
__device__ __forceinline__ double gl_exp_core(double x, double xtail, bool& ok, bool wide = false) {
  const unsigned abstop = (unsigned)(gl_d2u(x) >> 52) & 0x7ffu;
  ok = (abstop - 0x3c9u) <= 0x3eu;
  const bool special = wide && !ok && abstop >= 0x408u && abstop <= 0x408u;   // 512 <= |x| < 1024
  const double InvLn2N = gl_u2d(GL_EXP_HEAD[0]), Shift = gl_u2d(GL_EXP_HEAD[1]), NegLn2hiN = gl_u2d(GL_EXP_HEAD[2]), NegLn2loN = gl_u2d(GL_EXP_HEAD[3]);
  const double C2 = gl_u2d(GL_EXP_HEAD[4]), C3 = gl_u2d(GL_EXP_HEAD[5]), C4 = gl_u2d(GL_EXP_HEAD[6]), C5 = gl_u2d(GL_EXP_HEAD[7]);
  const double z = fma(x, InvLn2N, Shift);
  const unsigned long long ki = gl_d2u(z);
  const double kd = __dsub_rn(z, Shift);
  double r = fma(kd, NegLn2hiN, x);
  r = fma(kd, NegLn2loN, r);
  r = __dadd_rn(r, xtail);                       // exp(): xtail = 0 (r + 0 = r exactly; r is never -0 here unless x is, excluded)
  const unsigned idx = 2u * (unsigned)(ki & 127ull);
  const unsigned long long top = ki << 45;
  const double tail = gl_u2d(GL_EXP_TAB[idx]);
  const unsigned long long sbits = GL_EXP_TAB[idx + 1] + top;
  const double p23 = fma(C3, r, C2);
  const double t3 = __dadd_rn(r, tail);
  const double r2 = __dmul_rn(r, r);
  const double p45 = fma(r, C5, C4);
  const double acc = fma(p23, r2, t3);
  const double r4 = __dmul_rn(r2, r2);
  const double tmp = fma(r4, p45, acc);
  if (special) {
    if ((ki & 0x80000000ull) == 0) {   // k > 0: the exponent of scale might have overflowed by <= 460
      const double scale = gl_u2d(sbits - (1009ull << 52));
      const double y = __dmul_rn(fma(scale, tmp, scale), gl_u2d(0x7f00000000000000ull));   // 0x1p1009
      ok = isfinite(y);
      return y;
    }
    const double scale = gl_u2d(sbits + (1022ull << 52));   // k < 0
    const double y = __dadd_rn(scale, __dmul_rn(scale, tmp));
    ok = y >= 1.0;   // below: glibc rounds once into the subnormal range (left to the fallback)
    return __dmul_rn(y, gl_u2d(0x0010000000000000ull));     // 0x1p-1022
  }
  const double scale = gl_u2d(sbits);
  return fma(scale, tmp, scale);
}

__device__ __forceinline__ double gl_exp(double x) {
  bool ok;
  const double y = gl_exp_core(x, 0.0, ok, true);
  return ok ? y : exp(x);
}

// pow(x, y) for positive normal x and 2^-65 <= |y| < 2^63 (log_inline + exp_inline of e_pow.c)
__device__ __forceinline__ double gl_pow(double x, double y) {
  const unsigned long long ix = gl_d2u(x), iy = gl_d2u(y);
  const unsigned topx = (unsigned)(ix >> 52), topy = (unsigned)(iy >> 52);
  if ((topx - 1u) > 0x7fdu || ((topy & 0x7ffu) - 0x3beu) > 0x7fu) return pow(x, y);
  const double Ln2hi = gl_u2d(GL_POWLOG_HEAD[0]), Ln2lo = gl_u2d(GL_POWLOG_HEAD[1]);
  const double A0 = gl_u2d(GL_POWLOG_HEAD[2]), A1 = gl_u2d(GL_POWLOG_HEAD[3]), A2 = gl_u2d(GL_POWLOG_HEAD[4]), A3 = gl_u2d(GL_POWLOG_HEAD[5]),
               A4 = gl_u2d(GL_POWLOG_HEAD[6]), A5 = gl_u2d(GL_POWLOG_HEAD[7]), A6 = gl_u2d(GL_POWLOG_HEAD[8]);
  const unsigned long long tmp = ix - 0x3fe6955500000000ull;
  const unsigned i = (unsigned)(tmp >> 45) & 127u;
  const int k = (int)((long long)tmp >> 52);
  const unsigned long long iz = ix - (tmp & 0xfff0000000000000ull);
  const double z = gl_u2d(iz), kd = (double)k;
  const double invc = gl_u2d(GL_POWLOG_TAB[4 * i]), logc = gl_u2d(GL_POWLOG_TAB[4 * i + 2]), logctail = gl_u2d(GL_POWLOG_TAB[4 * i + 3]);
  const double t1 = fma(kd, Ln2hi, logc);
  const double lo1 = fma(kd, Ln2lo, logctail);
  const double r = fma(z, invc, -1.0);
  const double ar = __dmul_rn(r, A0);
  const double p12 = fma(r, A2, A1);
  const double p34 = fma(r, A4, A3);
  const double t2 = __dadd_rn(r, t1);
  const double lo2 = __dadd_rn(__dsub_rn(t1, t2), r);
  const double ar2 = __dmul_rn(r, ar);
  const double ar3 = __dmul_rn(r, ar2);
  const double lo3 = fma(ar, r, -ar2);
  const double hi = __dadd_rn(t2, ar2);
  const double p56 = fma(r, A6, A5);
  const double lo4 = __dadd_rn(__dsub_rn(t2, hi), ar2);
  const double q = fma(p56, ar2, p34);
  const double p = fma(ar2, q, p12);
  double lo = __dadd_rn(lo1, lo2);
  lo = __dadd_rn(lo, lo3);
  lo = __dadd_rn(lo, lo4);
  lo = fma(ar3, p, lo);
  const double lhi = __dadd_rn(hi, lo);
  const double llo = __dadd_rn(__dsub_rn(hi, lhi), lo);
  const double ehi = __dmul_rn(y, lhi);
  const double e1 = fma(lhi, y, -ehi);
  const double elo = fma(y, llo, e1);
  bool ok;
  const double res = gl_exp_core(ehi, elo, ok, true);
  return ok ? res : pow(x, y);
}

// log(x) for positive normal x (e_log.c), both the table path and the polynomial path around 1
__device__ __forceinline__ double gl_log(double x) {
  const unsigned long long ix = gl_d2u(x);
  const unsigned top = (unsigned)(ix >> 48);
  if ((ix - 0x3fee000000000000ull) <= 0x308ffffffffffull) {   // 1 - 2^-4 <= x < 1 + 0x1.09p-4
    if (ix == 0x3ff0000000000000ull) return 0.0;
    double Bc[11];
#pragma unroll
    for (int q = 0; q < 11; ++q) Bc[q] = gl_u2d(GL_LOG_HEAD[7 + q]);
    const double r = __dsub_rn(x, 1.0);
    const double a = fma(r, Bc[2], Bc[1]);
    const double b = fma(r, Bc[5], Bc[4]);
    const double r2 = __dmul_rn(r, r);
    const double c = fma(r, Bc[8], Bc[7]);
    const double a2 = fma(r2, Bc[3], a);
    const double b2 = fma(r2, Bc[6], b);
    const double r3 = __dmul_rn(r, r2);
    double y = fma(r2, Bc[9], c);
    y = fma(r3, Bc[10], y);
    y = fma(y, r3, b2);
    y = fma(y, r3, a2);
    const double two27 = 134217728.0;
    const double rw = fma(r, two27, r);
    const double rhi = fma(-two27, r, rw);
    const double rhi2 = __dmul_rn(rhi, rhi);
    const double rlo = __dsub_rn(r, rhi);
    const double hi = fma(rhi2, Bc[0], r);
    const double d = __dsub_rn(r, hi);
    const double s = __dadd_rn(r, rhi);
    double lo = fma(rhi2, Bc[0], d);
    const double m = __dmul_rn(Bc[0], rlo);
    lo = fma(m, s, lo);
    y = fma(y, r3, lo);
    return __dadd_rn(hi, y);
  }
  if ((top - 0x10u) > 0x7fdfu) return log(x);   // zero, negative, subnormal, inf, nan
  const double Ln2hi = gl_u2d(GL_LOG_HEAD[0]), Ln2lo = gl_u2d(GL_LOG_HEAD[1]);
  const double A0 = gl_u2d(GL_LOG_HEAD[2]), A1 = gl_u2d(GL_LOG_HEAD[3]), A2 = gl_u2d(GL_LOG_HEAD[4]), A3 = gl_u2d(GL_LOG_HEAD[5]), A4 = gl_u2d(GL_LOG_HEAD[6]);
  const unsigned long long tmp = ix - 0x3fe6000000000000ull;
  const unsigned i = (unsigned)(tmp >> 45) & 127u;
  const int k = (int)((long long)tmp >> 52);
  const unsigned long long iz = ix - (tmp & 0xfff0000000000000ull);
  const double z = gl_u2d(iz), kd = (double)k;
  const double invc = gl_u2d(GL_LOG_TAB[2 * i]), logc = gl_u2d(GL_LOG_TAB[2 * i + 1]);
  const double w = fma(kd, Ln2hi, logc);
  const double r = fma(z, invc, -1.0);
  const double p12 = fma(r, A2, A1);
  const double hi = __dadd_rn(r, w);
  const double r2 = __dmul_rn(r, r);
  double lo = __dadd_rn(__dsub_rn(w, hi), r);
  lo = fma(kd, Ln2lo, lo);
  const double r3 = __dmul_rn(r, r2);
  const double p34 = fma(r, A4, A3);
  lo = fma(r2, A0, lo);
  const double q = fma(p34, r2, p12);
  const double y = fma(r3, q, lo);
  return __dadd_rn(y, hi);
}

}  // namespace rx
