// rx_visc_pair.cuh — laminar viscous edge flux with TWO LANES PER EDGE (sixteen edges per warp).
//
// Same arithmetic, bit for bit, as viscous_flux<NS, EDGE=true> in rx_flux.cuh (CAvgGradReactive_Flow::ComputeResidual,
// numerics_direct_reactive.cpp:1425-1573; Solve_SM :451-470; GetGamma reacting_model_library.cpp:771-798), arranged for the
// SM: the thread-per-edge form needs ~180 live doubles per thread for the 9x9 Stefan–Maxwell solve, which pins the kernel at
// 8 warps/SM with spills.  Here the two lanes of a pair split the nine species rows as
//     lane 0: rows 0,2,4,6        lane 1: rows 1,3,5,7 and 8          (row of slot k: 2k + lane, slot 4 = row 8)
// which is exactly how Eigen's packet reduction (Redux.h) splits a 9-sum: ((a0+a4)+(a2+a6)) + ((a1+a5)+(a3+a7)), then + a8.
// Each lane forms its half with the same expression, the halves meet through one shuffle (IEEE addition commutes), row 8 is
// the scalar tail on both.  The GEMV keeps Eigen's per-row order: rows < 8 use the four-column tree, row 8 — always slot 4 —
// the sequential form, so the choice is static.  Slot 4 of lane 0 mirrors row 8 and is never read by anything that is kept.
// Matrix rows live in registers (slots < MREG) or in thread-private shared memory words (slot >= MREG).
#pragma once
#include "rx_kernels.cuh"

namespace rx {

#ifndef RX_VP_BS
#define RX_VP_BS 128
#endif
#ifndef RX_VP_MINB
#define RX_VP_MINB 3
#endif
#ifndef RX_VP_MREG
#define RX_VP_MREG 2
#endif
constexpr int VP_BS = RX_VP_BS;       // threads per block (two per edge)
constexpr int VP_MINB = RX_VP_MINB;   // resident blocks per SM the register budget is set for
constexpr int VP_MREG = RX_VP_MREG;   // matrix slots held in registers; the others are read from shared memory
#ifndef RX_VP_NO_ASYNC_D
#define RX_VP_ASYNC_D 1   // 1/D prefetch through cp.async into shared memory (default); -DRX_VP_NO_ASYNC_D keeps it in registers
#endif
#ifdef RX_VP_ASYNC_D
constexpr int VP_SMW = 5 * 9 + 2 * 9;   // shared-memory words per thread: the lane's five matrix rows + the staged 1/D words
#else
constexpr int VP_SMW = 5 * 9;   // shared-memory words per thread: the lane's five matrix rows
#endif
constexpr int VP_SMW_RANS = 13 * 5;   // SST closure: the pair shares a 9 x 13 array (operator, two right-hand sides, norms)
constexpr unsigned VP_FULL = 0xffffffffu;

struct PairLane {
  int l;          // 0 / 1
  int odd;        // lane id of the odd lane of this pair within the warp
  __device__ __forceinline__ double other(double v) const { return __shfl_xor_sync(VP_FULL, v, 1); }
  __device__ __forceinline__ double fromOdd(double v) const { return __shfl_sync(VP_FULL, v, odd); }
  __device__ __forceinline__ double sel(double even, double oddv) const { return l ? oddv : even; }
};

// a[k] <- a[k] / M[row[k]] for the five slots of a lane (see div_by_M in rx_device.cuh)
template <int NS>
__device__ __forceinline__ void div_by_M_rows(const ChemC<NS>& cc, const int* row, double* a) {
  bool ok = true;
#pragma unroll
  for (int k = 0; k < 5; ++k) ok = ok && num_range(a[k]);
  if (ok) {
#pragma unroll
    for (int k = 0; k < 5; ++k) a[k] = fdivu(a[k], Recip{cc.M[row[k]], cc.rM[row[k]]});
  } else {
#pragma unroll
    for (int k = 0; k < 5; ++k) a[k] = a[k] / cc.M[row[k]];
  }
}

// sum of nine values spread over the pair (a[k] = value of the lane's slot k) in Eigen redux order
__device__ __forceinline__ double pair_redux(const PairLane& L, const double* a) {
  const double s = (a[0] + a[2]) + (a[1] + a[3]);
  const double tot = s + L.other(s);
  return tot + L.fromOdd(a[4]);
}
__device__ __forceinline__ double pair_dot(const PairLane& L, const double* a, const double* b) {
  double t[5];
#pragma unroll
  for (int k = 0; k < 5; ++k) t[k] = a[k] * b[k];
  return pair_redux(L, t);
}
// all nine entries of a vector spread over the pair
__device__ __forceinline__ void pair_gather(const PairLane& L, const double* a, double* f) {
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const double o = L.other(a[k]);
    f[2 * k] = L.l ? o : a[k];
    f[2 * k + 1] = L.l ? a[k] : o;
  }
  f[8] = L.fromOdd(a[4]);
}
// rows of the lane of A*f, Eigen GEMV order per row (see eig_gemv)
template <class MatF>
__device__ __forceinline__ void pair_gemv(MatF A, const double* f, double* res) {
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    double r = 0.0;
    r = r + ((A(k, 0) * f[0] + A(k, 1) * f[1]) + (A(k, 2) * f[2] + A(k, 3) * f[3]));
    r = r + ((A(k, 4) * f[4] + A(k, 5) * f[5]) + (A(k, 6) * f[6] + A(k, 7) * f[7]));
    res[k] = A(k, 8) * f[8] + r;
  }
  {
    double r = 0.0;
#pragma unroll
    for (int c = 0; c < 9; ++c) r = A(4, c) * f[c] + r;
    res[4] = r;
  }
}


// ---- Eigen::ColPivHouseholderQR(Mt).solve(b) for two right-hand sides, worked by the two lanes of a pair on ONE copy of
// the matrix in shared memory (colpiv_qr_solve<> in rx_flux.cuh is the serial form and carries the citations; this is the
// same sequence of operations in the same order — pivot rule, rank threshold, norm down-dating, reflectors, and every sum in
// Eigen's own order — so the result equals the serial routine's, and the reference's, bit for bit).  The work is split by
// COLUMNS: after the reflector of step k is formed, lane l applies it to columns k+1+l, k+3+l, ... and down-dates their norms;
// the two right-hand sides are columns 9 and 10, one per lane at every step (the lane whose column parity it matches), and each
// lane back-substitutes one right-hand side.  No sum crosses lanes.  The step loop stays rolled (nine instantiated steps were
// measured 9 % slower: instruction fetch); entries at or above the diagonal take part as zeros, which no sum order notices.
// Storage (pair-shared: the words of both lanes): entry (i, j) at smPair[(i & 1) + (j*5 + (i >> 1)) * VP_BS], j = 0..10;
// updated / direct column norms at column slots 11 and 12 (i = column index).
template <int NS>
struct PairMat {
  double* base;   // smPair
  __device__ __forceinline__ double& operator()(int i, int j) const { return base[(i & 1) + (j * 5 + (i >> 1)) * VP_BS]; }
  __device__ __forceinline__ double* col(int j) const { return base + (j * 5) * VP_BS; }
  __device__ __forceinline__ static constexpr int off(int i) { return (i & 1) + (i >> 1) * VP_BS; }   // entry i of a column
};
constexpr int VP_QR_COLS = 13;   // 9 matrix columns, 2 right-hand sides, 2 norm vectors

// Eigen's packet-ordered sum (eig_redux<8 - k> of rx_flux.cuh) over the entries k+1 .. 8 of a zero-padded T[0..8] (T[i] = 0 for
// i <= k).  Written out for the eight lengths, the trees are
//   k = 0: ((T1+T5)+(T3+T7)) + ((T2+T6)+(T4+T8))        k = 1: (((T2+T4)+T6) + ((T3+T5)+T7)) + T8      k = 2: ((T3+T5)+T7) + ((T4+T6)+T8)
//   k = 3: ((T4+T6)+(T5+T7)) + T8     k = 4: (T5+T7)+(T6+T8)     k = 5: (T6+T7)+T8     k = 6: T7+T8     k = 7: T8
// and with the padding zeros in place every odd k is (B + A) + T8 and every even k >= 2 is A + C, where A = (T3+T5)+T7,
// B = (T2+T4)+T6, C = (T4+T6)+T8 (adding a zero changes nothing).  k is uniform across the warp.
__device__ __forceinline__ double eig_redux_tail(int k, const double* T) {
  if (k == 0) return ((T[1] + T[5]) + (T[3] + T[7])) + ((T[2] + T[6]) + (T[4] + T[8]));
  const double A = (T[3] + T[5]) + T[7], e46 = T[4] + T[6];
  return (k & 1) ? (((T[2] + T[4]) + T[6]) + A) + T[8] : A + (e46 + T[8]);
}

template <int NS>
__device__ __forceinline__ void pair_colpiv_qr_solve(const PairLane& L, double* smPair, double (*mg)[5]) {
  static_assert(NS == 9, "storage laid out for nine species");
  using PM = PairMat<NS>;
  const PM M{smPair};
  const int l = L.l;
  double* const nU = M.col(NS + 2);
  double* const nD = M.col(NS + 3);
  __syncwarp();   // the operator rows and right-hand sides were written by both lanes
  for (int j = l; j < NS; j += 2) {
    const double* cj = M.col(j);
    double sq[NS];
#pragma unroll
    for (int i = 0; i < NS; ++i) sq[i] = cj[PM::off(i)] * cj[PM::off(i)];
    const double n = sqrt(eig_redux<NS, false>(sq));
    nU[PM::off(j)] = n; nD[PM::off(j)] = n;
  }
  __syncwarp();
  const double eps = 2.220446049250313e-16;
  double mx = nU[PM::off(0)];
#pragma unroll
  for (int j = 1; j < NS; ++j) mx = fmax(mx, nU[PM::off(j)]);
  const double threshold_helper = (mx * eps) * (mx * eps) / (double)NS;
  const double downdate = sqrt(eps);
  int nonzero = NS;
  unsigned long long perm = 0x876543210ULL;   // nibble i = original column now at position i
#pragma unroll 1
  for (int k = 0; k < NS; ++k) {
    // pivot: first column of largest updated norm (both lanes, same result)
    int big = k;
    double nb = nU[PM::off(k)];
    for (int j = k + 1; j < NS; ++j) { const double nj = nU[PM::off(j)]; if (nj > nb) { nb = nj; big = j; } }
    if (nonzero == NS && nb * nb < threshold_helper * (double)(NS - k)) nonzero = k;
    __syncwarp();   // both lanes have read the norms (the scan above, and before the loop) before lane 0 swaps them below
    if (big != k) {
      double* ck = M.col(k); double* cb = M.col(big);
#pragma unroll
      for (int h = 0; h < 5; ++h) {   // lane l swaps the rows of its parity
        const int i = 2 * h + l;
        if (i < NS) { const double t = ck[PM::off(i)]; ck[PM::off(i)] = cb[PM::off(i)]; cb[PM::off(i)] = t; }
      }
      if (l == 0) {
        nU[PM::off(big)] = nU[PM::off(k)]; nU[PM::off(k)] = nb;
        const double t = nD[PM::off(k)]; nD[PM::off(k)] = nD[PM::off(big)]; nD[PM::off(big)] = t;
      }
      const unsigned long long pk = (perm >> (4 * k)) & 15ULL, pb = (perm >> (4 * big)) & 15ULL;
      perm = (perm & ~((15ULL << (4 * k)) | (15ULL << (4 * big)))) | (pb << (4 * k)) | (pk << (4 * big));
    }
    __syncwarp();
    // reflector of column k: both lanes form the scalars, each normalises the entries of its parity
    double* const ck = M.col(k);
    double vv[NS];   // entries below the diagonal, zero elsewhere
    vv[0] = 0.0;
#pragma unroll
    for (int i = 1; i < NS; ++i) vv[i] = i > k ? ck[PM::off(i)] : 0.0;
    const double c0 = M(k, k);
    __syncwarp();   // both lanes hold the raw column before either overwrites its half
    double T[NS];
    T[0] = 0.0;
#pragma unroll
    for (int i = 1; i < NS; ++i) T[i] = vv[i] * vv[i];
    const double tailSq = eig_redux_tail(k, T);
    double beta, tau;
    if (tailSq <= 2.2250738585072014e-308) {
      tau = 0.0; beta = c0;
#pragma unroll
      for (int i = 1; i < NS; ++i) vv[i] = 0.0;
    } else {
      beta = sqrt(c0 * c0 + tailSq);
      if (c0 >= 0.0) beta = -beta;
      const double den = c0 - beta;
      bool bad = !mid_range(den);   // operand ranges once per reflector instead of a validity test per quotient (rx_fdiv.cuh)
#pragma unroll
      for (int i = 1; i < NS; ++i) bad = bad || !num_range(vv[i]);
      const Recip rden = frecip(den);
      double vq[NS];
#pragma unroll
      for (int i = 1; i < NS; ++i) vq[i] = fdivu(vv[i], rden);
      if (bad) {
#pragma unroll
        for (int i = 1; i < NS; ++i) vq[i] = vv[i] / den;
      }
#pragma unroll
      for (int i = 1; i < NS; ++i) vv[i] = vq[i];
      tau = (beta - c0) / beta;
    }
#pragma unroll
    for (int i = 1; i < NS; ++i) if (i > k && (i & 1) == l) ck[PM::off(i)] = vv[i];
    if (l == (k & 1)) M(k, k) = beta;
    // columns of this lane: t = v.w + w_k,  w_k -= tau t,  w_i -= (tau v_i) t;  then the norm down-date
    if (tau != 0.0) {   // never at k = 8
      double tvv[NS];
#pragma unroll
      for (int i = 1; i < NS; ++i) tvv[i] = tau * vv[i];
      if (nonzero == NS) {   // solve() applies the first `nonzero` reflectors to the right-hand sides: an inner product each
        double* const cj = M.col(((k + 1 + l) & 1) ? NS : NS + 1);   // one right-hand side per lane at every step
        double w[NS];
#pragma unroll
        for (int i = 1; i < NS; ++i) { w[i] = cj[PM::off(i)]; T[i] = vv[i] * w[i]; }
        const double wk = cj[PM::off(k)];
        const double t = eig_redux_tail(k, T) + wk;
        cj[PM::off(k)] = wk - tau * t;
#pragma unroll
        for (int i = 1; i < NS; ++i) if (i > k) cj[PM::off(i)] = w[i] - tvv[i] * t;
      }
      // matrix columns: the first entry of the essential vector is the row kernel's scalar head (eig_gemv_row, rx_flux.cuh); it
      // is taken out of vv[] so that the two packet lanes are plain sums over the even and the odd entries
      double vv0 = 0.0;
#pragma unroll
      for (int i = 1; i < NS; ++i) { const bool head = i == k + 1; vv0 = head ? vv[i] : vv0; vv[i] = head ? 0.0 : vv[i]; }
      const int offHead = PM::off(k + 1);
      const bool kOdd = k & 1;
      for (int j = k + 1 + l; j < NS; j += 2) {
        double* const cj = M.col(j);
        double w[NS];
#pragma unroll
        for (int i = 1; i < NS; ++i) w[i] = cj[PM::off(i)];
        const double t0 = cj[offHead] * vv0;
        const double se6 = (w[2] * vv[2] + w[4] * vv[4]) + w[6] * vv[6], so = (w[3] * vv[3] + w[5] * vv[5]) + w[7] * vv[7];
        const double t8 = w[8] * vv[8];
        double t = kOdd ? t0 + (so + (se6 + t8)) : (t0 + (se6 + so)) + t8;
        const double wk = cj[PM::off(k)];
        t += wk;
        cj[PM::off(k)] = wk - tau * t;
#pragma unroll
        for (int i = 1; i < NS; ++i) if (i > k) cj[PM::off(i)] = w[i] - tvv[i] * t;
      }
    }
    for (int j = k + 1 + l; j < NS; j += 2) {
      const double nu = nU[PM::off(j)];
      if (nu != 0.0) {
        const double* cj = M.col(j);
        const double ndj = nD[PM::off(j)];
        const double akj = fabs(cj[PM::off(k)]);
        bool bad = false;
        double temp = fdiv(akj, nu, bad);
        double qq = fdiv(nu, ndj, bad);
        if (bad) { temp = akj / nu; qq = nu / ndj; }
        temp = (1.0 + temp) * (1.0 - temp);
        temp = temp < 0.0 ? 0.0 : temp;
        const double temp2 = temp * (qq * qq);
        if (temp2 <= downdate) {
#pragma unroll
          for (int i = 1; i < NS; ++i) { const double a = i > k ? cj[PM::off(i)] : 0.0; T[i] = a * a; }
          const double n = sqrt(eig_redux_tail(k, T));
          nD[PM::off(j)] = n; nU[PM::off(j)] = n;
        } else {
          bool bad2 = false;
          double rt = fsqrt(temp, bad2);
          if (bad2) rt = sqrt(temp);
          nU[PM::off(j)] = nu * rt;
        }
      }
    }
    __syncwarp();
  }
  // back substitution with R (upper triangle) as Eigen's triangular solver does it: within the panel of the last 8 columns a
  // solved unknown is subtracted from the rows above it in the panel; row 0 of a full-rank system (above the panel) then takes
  // the panel's columns in ascending order.  Lane l owns right-hand side l (column 9 + l); entries >= nonzero are zero.
  {
    double* const cr = M.col(NS + l);
    const int start = nonzero == NS ? 1 : 0;
    for (int i = nonzero - 1; i >= start; --i) {
      const double* ci = M.col(i);
      const double x = cr[PM::off(i)] / ci[PM::off(i)];
      cr[PM::off(i)] = x;
#pragma unroll
      for (int r = 0; r < NS - 1; ++r) if (r >= start && r < i) cr[PM::off(r)] = cr[PM::off(r)] - x * ci[PM::off(r)];
    }
    if (start) {
      double c0 = cr[PM::off(0)];
#pragma unroll
      for (int j = 1; j < NS; ++j) c0 = c0 - M.col(j)[PM::off(0)] * cr[PM::off(j)];
      cr[PM::off(0)] = c0 / M.col(0)[PM::off(0)];
    }
    // undo the column permutation into the (dead) norm slots: X(dd = l, perm[i]) = c[i]
    double* const xo = M.col(NS + 2 + l);
    __syncwarp();
    for (int i = 0; i < NS; ++i) {
      const int tr = (int)((perm >> (4 * i)) & 15ULL);
      xo[PM::off(tr)] = i < nonzero ? cr[PM::off(i)] : 0.0;
    }
    __syncwarp();
  }
#pragma unroll
  for (int dd = 0; dd < 2; ++dd) {
    const double* xo = M.col(NS + 2 + dd);
#pragma unroll
    for (int q2 = 0; q2 < 4; ++q2) mg[dd][q2] = xo[PM::off(2 * q2 + l)];
    mg[dd][4] = xo[PM::off(8)];
  }
}

// upper-triangle index of entry (r, c) of the symmetric 1/D array (tri_idx of rx_flux.cuh with the smaller index first)
struct TriQ { int q[81]; };
constexpr TriQ make_triq() {
  TriQ t{};
  for (int r = 0; r < 9; ++r)
    for (int c = 0; c < 9; ++c) { const int a = r < c ? r : c, b = r < c ? c : r; t.q[r * 9 + c] = a * 9 - a * (a - 1) / 2 + (b - a); }
  return t;
}
__constant__ TriQ c_triq = make_triq();

template <int NS, bool RANS, bool VLIM = false>   // VLIM: VISCOUS_LIMITER_FLOW (its own instantiation: the default kernel is untouched)
#ifdef RX_VP_MAXNREG   // explicit register cap instead of a resident-block target (tuning hook)
__global__ void __maxnreg__(RX_VP_MAXNREG) k_edge_visc_pair(
#else
__global__ void __launch_bounds__(VP_BS, VP_MINB) k_edge_visc_pair(
#endif
    const __grid_constant__ ChemC<NS> cc, const __grid_constant__ Opts o, const Dev d) {
  static_assert(NS == 9, "the row split {0,2,4,6 | 1,3,5,7,8} is laid out for nine species");
  extern __shared__ double vp_sm[];   // word w of thread t at vp_sm[w * VP_BS + t]
  double* const sm = vp_sm + threadIdx.x;
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  PairLane L;
  L.l = (int)(threadIdx.x & 1);
  L.odd = (int)(threadIdx.x & 31) | 1;
  long long e = t >> 1;
  const bool keep = e < d.nEdge;
  if (!keep) e = d.nEdge - 1;   // the lane stays in the warp's shuffles and votes; nothing it computes is stored
  if (d.eOrder) e = d.eOrder[e];   // processing slot -> edge (rx_order.cuh); every store below goes by edge index
  const int l = L.l;
  const long long np = d.np;
  const int pi = d.ei[e], pj = d.ej[e];
  const double nx = d.enx[e], ny = d.eny[e];
  int row[5];
#pragma unroll
  for (int k = 0; k < 4; ++k) row[k] = 2 * k + l;
  row[4] = 8;
  const double* rdi = d.rD + pi;
  const double* rdj = d.rD + pj;
#ifdef RX_VP_ASYNC_D
  // 1/D_rc of the two end points for the Stefan–Maxwell slot being assembled, fetched one slot ahead by asynchronous copies straight
  // into 18 thread-private shared-memory words (cp.async: no registers are held across the division batches of the current slot).
  // Slot 0 is requested HERE, before anything else of the edge is loaded, so that its latency lies under the mean state and the
  // gradients; the upper-triangle index of (row, column) comes from a constant table (one LDC + one IMAD.WIDE per address).
  const unsigned npB = (unsigned)(np * 8);   // np < 2^29 points
  auto stageD = [&](int w) -> double* { return sm + (VP_SMW - 2 * NS + w) * VP_BS; };
  auto fetchD = [&](int k) {
    const int r = k < 4 ? 2 * k + l : 8;
    const int* tq = c_triq.q + r * NS;
#pragma unroll
    for (int c = 0; c < NS; ++c) {
      const unsigned long long off = (unsigned long long)(unsigned)tq[c] * npB;
      const unsigned sa = (unsigned)__cvta_generic_to_shared(stageD(c)), sb = (unsigned)__cvta_generic_to_shared(stageD(NS + c));
      asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(sa), "l"((const char*)rdi + off) : "memory");
      asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(sb), "l"((const char*)rdj + off) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  fetchD(0);
#endif

  // ---- mean state, both lanes
  const double rho = 0.5 * (d.V[V_RHO * np + pi] + d.V[V_RHO * np + pj]);
  const double T = 0.5 * (d.V[V_T * np + pi] + d.V[V_T * np + pj]);
  const double um = 0.5 * (d.V[V_VX * np + pi] + d.V[V_VX * np + pj]);
  const double vm = 0.5 * (d.V[(V_VX + 1) * np + pi] + d.V[(V_VX + 1) * np + pj]);
  double mu, ktr;   // harmonic means; the six quotients as two batches of independent fast-path divisions (rx_fdiv.cuh), same bits
  {
    const double a0 = d.mu[pi], a1 = d.mu[pj], b0 = d.kap[pi], b1 = d.kap[pj];
    bool bad = false;
    const double i0 = fdiv(1.0, a0, bad), i1 = fdiv(1.0, a1, bad), j0 = fdiv(1.0, b0, bad), j1 = fdiv(1.0, b1, bad);
    mu = fdiv(2.0, i0 + i1, bad); ktr = fdiv(2.0, j0 + j1, bad);
    if (bad) { mu = 2.0 / (1.0 / a0 + 1.0 / a1); ktr = 2.0 / (1.0 / b0 + 1.0 / b1); }
  }
  double Ys[NS], Xs[NS];
#pragma unroll
  for (int s = 0; s < NS; ++s) Ys[s] = 0.5 * (d.V[(V_Y + s) * np + pi] + d.V[(V_Y + s) * np + pj]);
  molar_from_mass(cc, Ys, Xs);

  // ---- mean gradients with the edge correction (:1480-1500), six rows per lane:
  // slot 0: T | v,  slot 1: u | X_8,  slot 2+k: X_{2k+lane}
  double G[6][2];
  {
    const double e0 = d.x[pj] - d.x[pi], e1 = d.y[pj] - d.y[pi];
    const double dist2 = 0.0 + e0 * e0 + e1 * e1;
    double gi0[6], gj0[6], gi1[6], gj1[6], vi[6], vj[6];   // all loads first: the divisions below end basic blocks
#pragma unroll
    for (int q = 0; q < 6; ++q) {
      int gr;                 // gradient row
      const double* val;      // field the row differentiates
      if (q == 0) { gr = l ? G_VX + 1 : G_T; val = d.V + (size_t)(l ? V_VX + 1 : V_T) * np; }
      else if (q == 1) { gr = l ? G_X + 8 : G_VX; val = l ? d.X + (size_t)8 * np : d.V + (size_t)V_VX * np; }
      else { gr = G_X + 2 * (q - 2) + l; val = d.X + (size_t)(2 * (q - 2) + l) * np; }
      const double* g = d.grad + (size_t)(gr * 2) * np;
      gi0[q] = g[pi]; gj0[q] = g[pj]; gi1[q] = g[np + pi]; gj1[q] = g[np + pj];
      vi[q] = val[pi]; vj[q] = val[pj];
    }
    if constexpr (VLIM) {
      // numerics_direct_reactive.cpp:1482-1505: the gradients of T, u, v enter the mean times the limiter of their variable (Limiter_Primitive rows
      // T, u, v); mole-fraction rows are not limited.  Slot 0 holds T | v, slot 1 holds u | X_8.
      const int r0 = l ? 2 : 0;
      const double li0 = d.lim[r0 * np + pi], lj0 = d.lim[r0 * np + pj];
      gi0[0] = gi0[0] * li0; gi1[0] = gi1[0] * li0; gj0[0] = gj0[0] * lj0; gj1[0] = gj1[0] * lj0;
      if (!l) {
        const double li1 = d.lim[1 * np + pi], lj1 = d.lim[1 * np + pj];
        gi0[1] = gi0[1] * li1; gi1[1] = gi1[1] * li1; gj0[1] = gj0[1] * lj1; gj1[1] = gj1[1] * lj1;
      }
    }
    const Recip rd2 = frecip(dist2);   // one refined reciprocal serves the twelve quotients
    bool bad = false;
    double c0[6], c1[6];
#pragma unroll
    for (int q = 0; q < 6; ++q) {
      const double g0 = 0.5 * (gi0[q] + gj0[q]);
      const double g1 = 0.5 * (gi1[q] + gj1[q]);
      const double proj = (0.0 + g0 * e0) + g1 * e1;
      const double diff = vj[q] - vi[q];
      c0[q] = (proj - diff) * e0; c1[q] = (proj - diff) * e1;
      G[q][0] = g0 - fdiv(c0[q], rd2, bad);
      G[q][1] = g1 - fdiv(c1[q], rd2, bad);
    }
    if (bad) {
#pragma unroll
      for (int q = 0; q < 6; ++q) {
        G[q][0] = 0.5 * (gi0[q] + gj0[q]) - c0[q] / dist2;
        G[q][1] = 0.5 * (gi1[q] + gj1[q]) - c1[q] / dist2;
      }
    }
  }
  // rows T, u, v on both lanes
  double GT[2], GU[2], GV[2];
#pragma unroll
  for (int dd = 0; dd < 2; ++dd) {
    const double a = L.other(G[0][dd]);   // lane 0 receives v, lane 1 receives T
    const double b = L.other(G[1][dd]);   // lane 1 receives u
    GT[dd] = l ? a : G[0][dd];
    GV[dd] = l ? G[0][dd] : a;
    GU[dd] = l ? b : G[1][dd];
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
  // SST closure, part 1 (SST_Reactive_ResidualClosure :656-852): Boussinesq stresses join the tensor before the projection
  double muT = 0.0;
  if (RANS) {
    muT = 2.0 / (1.0 / d.turb[pi] + 1.0 / d.turb[pj]);
    const double ke = 0.5 * (d.turb[np + pi] + d.turb[np + pj]);
    const double Gv[2][2] = {{GU[0], GU[1]}, {GV[0], GV[1]}};
    const double vel[2] = {um, vm};
    double dv = 0.0;
    dv += Gv[0][0]; dv += Gv[1][1];
#pragma unroll
    for (int a = 0; a < 2; ++a) {
      double tt[2] = {0.0, 0.0};
#pragma unroll
      for (int b = 0; b < 2; ++b) tt[b] += muT * (Gv[b][a] + Gv[a][b]);
      tt[a] -= TWO3 * (muT * dv + ke * rho);
#pragma unroll
      for (int b = 0; b < 2; ++b) { FT[b][a] += tt[b]; FT[2][a] += tt[b] * vel[b]; }
    }
  }
  // momentum rows are complete; of the energy row only the two stress/heat products have to outlive the solve
  // (RANS: the raw tensor entries, which still receive the species and k terms of the closure)
  const double fE0 = RANS ? FT[2][0] : FT[2][0] * nx, fE1 = RANS ? FT[2][1] : FT[2][1] * ny;
  if (keep && l == 0) {
    double pMx = 0.0, pMy = 0.0;
    pMx += FT[0][0] * nx; pMx += FT[0][1] * ny;
    pMy += FT[1][0] * nx; pMy += FT[1][1] * ny;
    d.fluxV[U_MX * d.ne + e] = pMx;
    d.fluxV[(U_MX + 1) * d.ne + e] = pMy;
  }
  // right-hand side of the lane's rows: -(grad X . n)
  double rhs[5];
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    const int q = k < 4 ? k + 2 : 1;
    double gxn = 0.0;
    gxn += G[q][0] * nx;
    gxn += G[q][1] * ny;
    rhs[k] = -gxn;
  }
  if (l == 0) rhs[4] = 0.0;   // slot 4 of the even lane is a mirror; keep it finite

  // ---- Stefan–Maxwell matrix, the lane's rows (GetGamma :771-798, Gamma_tilde :463-465).  For row r and column c != r:
  //   D = 2/(1/D_rc(i) + 1/D_rc(j)),  Gamma(r,c) = -sigma*massTot*X_r/(rho*M_c*D),  tmp_r += X_c/D  (ascending c),
  //   Gamma(r,r) = sigma*massTot*tmp_r/(rho*M_r);  every entry of row r then gets alpha*Y_r, alpha = 1/(rho*max D).
  // Columns 2k and 2k+1 of slot k < 4 hold the diagonal on one lane and an ordinary entry on the other.
  // The matrix is assembled in thread-private shared-memory words (slot k, column c at word k*NS + c) by loops that stay
  // rolled over the slots — 135 IEEE divisions unrolled would be most of the kernel's code — with the per-slot operands
  // rotated through registers so that no register array is indexed dynamically.
  auto smA = [&](int k, int c) -> double& { return sm[(k * NS + c) * VP_BS]; };
  double invd[5] = {1.0, 1.0, 1.0, 1.0, 1.0};
  // A zero right-hand side (uniform composition along the edge) makes BiCGSTAB return x = 0 without touching the matrix; a warp
  // whose sixteen edges are all of that kind skips the assembly and the solve (x = 0 below, as the loop would have left it).
  const bool anyRhs = __any_sync(VP_FULL, pair_dot(L, rhs, rhs) != 0);
  if (anyRhs) {
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
    double maxD = 0.0;
    double Xrow[5], Yrow[5];
#pragma unroll
    for (int k = 0; k < 4; ++k) { Xrow[k] = L.sel(Xs[2 * k], Xs[2 * k + 1]); Yrow[k] = L.sel(Ys[2 * k], Ys[2 * k + 1]); }
    Xrow[4] = Xs[8]; Yrow[4] = Ys[8];
    bool inOk = mid_range(rho);   // rho and the mole fractions enter every quotient of the assembly
#pragma unroll
    for (int s = 0; s < NS; ++s) inOk = inOk && num_range(Xs[s]);
#ifndef RX_VP_ASYNC_D
    double ra[NS], rb[NS];   // 1/D_rc of the two end points for the slot being assembled, fetched one slot ahead
    auto fetchD = [&](int k) {
      const int r = k < 4 ? 2 * k + l : 8;
      // upper-triangle index of (min, max): row r from its diagonal on is contiguous (base_r + c); above it, column r of
      // row c (a compile-time offset per c, plus r)
      const int base_r = r * NS - r * (r - 1) / 2 - r;
#pragma unroll
      for (int c = 0; c < NS; ++c) {
        const int up = c * NS - c * (c - 1) / 2 - c + r;
        const size_t q = (size_t)(c >= r ? base_r + c : up) * np;
        ra[c] = rdi[q]; rb[c] = rdj[q];
      }
    };
    fetchD(0);
#endif
#pragma unroll 1
    for (int k = 0; k < 5; ++k) {
      const int r = k < 4 ? 2 * k + l : 8;
      const double Xr = Xrow[0];
      double sD[NS];
#ifdef RX_VP_ASYNC_D
      asm volatile("cp.async.wait_group 0;" ::: "memory");
#pragma unroll
      for (int c = 0; c < NS; ++c) sD[c] = *stageD(c) + *stageD(NS + c);
#else
#pragma unroll
      for (int c = 0; c < NS; ++c) sD[c] = ra[c] + rb[c];
#endif
      if (k < 4) fetchD(k + 1);
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
#pragma unroll
      for (int m = 0; m < 4; ++m) Xrow[m] = Xrow[m + 1];
    }
    maxD = fmax(maxD, L.other(maxD));
    const double alpha = 1.0 / (rho * maxD);
#pragma unroll 1
    for (int k = 0; k < 5; ++k) {
      const int r = k < 4 ? 2 * k + l : 8;
      const double ay = alpha * Yrow[0];
      const double dg = sigma * massTot * smA(k, r) / (rho * cc.M[r]);
      smA(k, r) = dg;
#pragma unroll
      for (int c = 0; c < NS; ++c) smA(k, c) = smA(k, c) + ay;
#pragma unroll
      for (int m = 0; m < 4; ++m) Yrow[m] = Yrow[m + 1];
    }
#pragma unroll
    for (int k = 0; k < 5; ++k) { const double dg = smA(k, row[k]); invd[k] = dg != 0.0 ? 1.0 / dg : 1.0; }
  }
#ifdef RX_VP_ASYNC_D
  asm volatile("cp.async.wait_group 0;" ::: "memory");   // a warp that skipped the assembly still has slot 0's copies in flight
#endif
  // rows of the first VP_MREG slots move to registers; their shared-memory words then hold rhs, r0 and x of the solve
  double A[VP_MREG > 0 ? VP_MREG : 1][NS];
#pragma unroll
  for (int k = 0; k < VP_MREG; ++k)
#pragma unroll
    for (int c = 0; c < NS; ++c) A[k][c] = smA(k, c);
  auto getA = [&](int k, int c) -> double { return k < VP_MREG ? A[k][c] : sm[(k * NS + c) * VP_BS]; };
#ifdef RX_VP_ASYNC_D
  constexpr int SW0 = VP_SMW - 2 * NS;   // rhs, r0 and x of the solve take the 1/D staging words, free once the matrix stands
#else
  constexpr int SW0 = 0;
  static_assert(VP_MREG * NS >= 15, "rhs, r0 and x reuse the words of the register-held rows");
#endif
  auto s_rhs = [&](int k) -> double& { return sm[(SW0 + k) * VP_BS]; };
  auto s_r0 = [&](int k) -> double& { return sm[(SW0 + 5 + k) * VP_BS]; };
  auto s_x = [&](int k) -> double& { return sm[(SW0 + 10 + k) * VP_BS]; };

  // ---- Eigen::BiCGSTAB with diagonal preconditioner, tol 1e-11, zero guess, at most 2N iterations (see bicgstab<> in
  // rx_flux.cuh for the serial form and its citations).  All scalars are identical on the two lanes of a pair.  Pairs of
  // one warp run the loop together until the last one has converged; a pair that is done keeps its state (`live`).
  double r[5], v[5], p[5], tv[5], w1[5], f[NS];
#pragma unroll
  for (int k = 0; k < 5; ++k) { s_rhs(k) = rhs[k]; s_x(k) = 0.0; r[k] = rhs[k] - 0.0; s_r0(k) = r[k]; v[k] = 0.0; p[k] = 0.0; }
  double r0_sqnorm = pair_dot(L, r, r);
  const double rhs_sqnorm = r0_sqnorm;
  double w_b = 1;
  const double tol2 = 1.0e-11 * 1.0e-11 * rhs_sqnorm;
  const double eps2 = 2.220446049250313e-16 * 2.220446049250313e-16;
  int it = 0, restarts = 0;
  bool live = (rhs_sqnorm != 0) && (r0_sqnorm > tol2);   // first loop test: r == r0
  // Software-pipelined scalars: rho = r0.r of the coming iteration is formed next to |r|^2 at the end of the previous one, and
  // the factors of beta = (rho/rho_old)(alpha/omega) that do not need the new rho — 1/rho_old and alpha/omega — as soon as
  // their operands exist.  Same operations on the same values; two reductions and a division leave the critical path.
  double rho_n = pair_dot(L, r, r);   // r0 == r
  double rhoOldS = 1.0, abS = 1.0, wbS = 1.0, q2 = 1.0;   // sanitised copies (a converged pair: 1, 0, 1) for the fallback
  Recip Rold = frecip(1.0);
  bool badq2 = false;
#pragma unroll 1
  while (__any_sync(VP_FULL, live)) {
    const bool restart = live && (fabs(rho_n) < eps2 * r0_sqnorm);
    if (__any_sync(VP_FULL, restart)) {   // r = rhs - A x, r0 = r (rare)
#pragma unroll
      for (int k = 0; k < 5; ++k) w1[k] = s_x(k);
      pair_gather(L, w1, f);
      pair_gemv(getA, f, tv);
      double rn[5];
#pragma unroll
      for (int k = 0; k < 5; ++k) rn[k] = s_rhs(k) - tv[k];
      const double nn = pair_dot(L, rn, rn);
      if (restart) {
#pragma unroll
        for (int k = 0; k < 5; ++k) { r[k] = rn[k]; s_r0(k) = rn[k]; }
        rho_n = r0_sqnorm = nn;
        if (restarts++ == 0) it = 0;
      }
    }
    // A pair that has converged keeps running with the warp; nothing of it is stored any more.  Instead of guarding its
    // vectors (twenty selects and as many moves per iteration) its SCALARS are made benign — beta = alpha = omega = 0 —
    // which freezes r, p = r, v = A M^-1 r at finite values.  Live pairs see exactly the arithmetic of before.
    double beta;
    {
      bool bad = badq2;
      const double bn1 = live ? rho_n : 0.0;
      const double q1 = fdiv(bn1, Rold, bad);
      beta = bad ? (bn1 / rhoOldS) * (abS / wbS) : q1 * q2;
    }
#pragma unroll
    for (int k = 0; k < 5; ++k) { p[k] = r[k] + beta * (p[k] - w_b * v[k]); w1[k] = invd[k] * p[k]; }
    pair_gather(L, w1, f);
    pair_gemv(getA, f, v);
#pragma unroll
    for (int k = 0; k < 5; ++k) w1[k] = s_r0(k);
    const double r0v = pair_dot(L, w1, v);
    const double alpha_n = (live ? rho_n : 0.0) / (live ? r0v : 1.0);
    double s[5], z[5];
#pragma unroll
    for (int k = 0; k < 5; ++k) { s[k] = r[k] - alpha_n * v[k]; z[k] = invd[k] * s[k]; }
    pair_gather(L, z, f);
    pair_gemv(getA, f, tv);
    const double tt = pair_dot(L, tv, tv);
    const double ts = pair_dot(L, tv, s);
    const double w_n = (live && tt > 0.0) ? ts / tt : 0.0;
#pragma unroll
    for (int k = 0; k < 5; ++k) r[k] = s[k] - w_n * tv[k];
    const double rrn = pair_dot(L, r, r);
#pragma unroll
    for (int k = 0; k < 5; ++k) w1[k] = s_r0(k);
    const double rho_next = pair_dot(L, w1, r);
    if (live) {
#pragma unroll
      for (int k = 0; k < 5; ++k) s_x(k) = s_x(k) + (alpha_n * (invd[k] * p[k]) + w_n * z[k]);
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
  double x[5];
#pragma unroll
  for (int k = 0; k < 5; ++k) x[k] = s_x(k);

  // ---- assemble the projected flux (:1530-1573).  Proj[rho] = -sum Jd (Eigen sum), Proj[E] = sum_s -h_s Jd_s in species
  // order, then the stress/heat part; species rows go out from the lane that owns them.
  const double sumJ = pair_redux(L, x);
  const SplPos sp = spl_pos(cc, T);
  double hsr[5], hj[5];
#pragma unroll
  for (int k = 0; k < 5; ++k) hsr[k] = spl_eval(cc, P_H, row[k], sp);
  div_by_M_rows(cc, row, hsr);
#pragma unroll
  for (int k = 0; k < 5; ++k) hj[k] = -hsr[k] * x[k];
  pair_gather(L, hj, f);
  double pE = 0.0;
#pragma unroll
  for (int s2 = 0; s2 < NS; ++s2) pE += f[s2];
  double pS[5];
#pragma unroll
  for (int k = 0; k < 5; ++k) pS[k] = -x[k];
  bool ranQR = false;
  if (RANS) {
    // ---- SST closure, part 2: mass-fraction gradients from the mole-fraction ones through the molar->mass operator
    // (Get_Molar2MassGrad_Operator :861-880, solved with Eigen's ColPivHouseholderQR :743), then the turbulent species,
    // enthalpy and k fluxes in the reference's order of additions (:800-830; viscous_flux<> in rx_flux.cuh).
    double sig = 0.0;
#pragma unroll
    for (int s2 = 0; s2 < NS; ++s2) sig = sig + Xs[s2];
    double Yr[5], Xr[5];
#pragma unroll
    for (int k = 0; k < 4; ++k) { Yr[k] = L.sel(Ys[2 * k], Ys[2 * k + 1]); Xr[k] = L.sel(Xs[2 * k], Xs[2 * k + 1]); }
    Yr[4] = Ys[8]; Xr[4] = Xs[8];
    double* const smPair = sm - l;
    const PairMat<NS> Mt{smPair};
    __syncwarp();   // the pair's words are about to be shared: both lanes are done with the solve's private words
    // the divisors are the molar masses: their refined reciprocals were formed once at create (ChemC::rM, the value frecip()
    // returns); the numerators' ranges are tested once for the whole operator (rx_fdiv.cuh: fdivu)
    bool numOk = num_range(cc.Mtot);
#pragma unroll
    for (int s2 = 0; s2 < NS; ++s2) numOk = numOk && num_range(Ys[s2]) && num_range(Xs[s2]);
#pragma unroll 1
    for (int k = 0; k < 5; ++k) {
      const int r = k < 4 ? 2 * k + l : 8;
      const double Mr = cc.M[r];
      const Recip rMr{Mr, cc.rM[r]};
      double xom[NS];
      const bool bad = !numOk;
      double mom = fdivu(cc.Mtot, rMr), yom = fdivu(Yr[0], rMr);
#pragma unroll
      for (int c = 0; c < NS; ++c) xom[c] = fdivu(Xr[0], Recip{cc.M[c], cc.rM[c]});
      if (bad) {
        mom = cc.Mtot / Mr; yom = Yr[0] / Mr;
#pragma unroll
        for (int c = 0; c < NS; ++c) xom[c] = Xr[0] / cc.M[c];
      }
      const double a = mom * (Yr[0] - Xr[0] + sig);
      if (k < 4 || l == 1) {   // row 8 is written by the odd lane
        // m_tilde(i,j) = a*(i==j) + b*(i!=j) (:872-873): a*1.0 + b*0.0 on the diagonal, a*0.0 + b*1.0 off it
#pragma unroll
        for (int c = 0; c < NS; ++c) {
          const double b = cc.Mtot * (yom - xom[c]);
          Mt(r, c) = (c == r) ? a + b * 0.0 : a * 0.0 + b;
        }
      }
#pragma unroll
      for (int m = 0; m < 4; ++m) { Yr[m] = Yr[m + 1]; Xr[m] = Xr[m + 1]; }
    }
    double MG[2][5];
    bool need = false;   // entries below the threshold are zeroed after the solve; if all are, the solve is moot
#pragma unroll
    for (int dd = 0; dd < 2; ++dd) {
#pragma unroll
      for (int k = 0; k < 4; ++k) { Mt(2 * k + l, NS + dd) = G[2 + k][dd]; MG[dd][k] = 0.0; need = need || !(fabs(G[2 + k][dd]) < 1e-8); }
      MG[dd][4] = 0.0;
      if (l) { Mt(8, NS + dd) = G[1][dd]; need = need || !(fabs(G[1][dd]) < 1e-8); }   // X_8 lives on the odd lane
    }
    ranQR = need;
    if (__any_sync(VP_FULL, need)) pair_colpiv_qr_solve<NS>(L, smPair, MG);
#pragma unroll
    for (int dd = 0; dd < 2; ++dd) {
#pragma unroll
      for (int k = 0; k < 4; ++k) if (fabs(G[2 + k][dd]) < 1e-8) MG[dd][k] = 0.0;
      if (l == 0 || fabs(G[1][dd]) < 1e-8) MG[dd][4] = 0.0;
    }
    const double c1 = muT / (o.prandtlTurb * o.lewisTurb), c2 = muT / o.prandtlTurb;
    const double sigma_k = d.turb[5 * np + pi];   // sigma_k of the first node (:5345)
    const double nrm[2] = {nx, ny};
    double FE[2] = {fE0, fE1};
#pragma unroll
    for (int k = 0; k < 4; ++k) Yr[k] = L.sel(Ys[2 * k], Ys[2 * k + 1]);
    Yr[4] = Ys[8];
    double cps[5];
#pragma unroll
    for (int k = 0; k < 5; ++k) cps[k] = spl_eval(cc, P_CP, row[k], sp);
    div_by_M_rows(cc, row, cps);
#pragma unroll
    for (int a = 0; a < 2; ++a) {
      double t1[5], t2[5];
#pragma unroll
      for (int k = 0; k < 5; ++k) {
        pS[k] += c1 * MG[a][k] * nrm[a];
        t1[k] = c1 * hsr[k] * Yr[k] * MG[a][k];
        t2[k] = c2 * cps[k] * Yr[k] * GT[a];
      }
      pair_gather(L, t1, f);
#pragma unroll
      for (int s2 = 0; s2 < NS; ++s2) FE[a] += f[s2];
      pair_gather(L, t2, f);
#pragma unroll
      for (int s2 = 0; s2 < NS; ++s2) FE[a] += f[s2];
      const double gk = 0.5 * (d.turb[(3 + a) * np + pi] + d.turb[(3 + a) * np + pj]);
      FE[a] += (mu + muT / sigma_k) * gk;
    }
    pE += FE[0] * nx;
    pE += FE[1] * ny;
  } else {
    pE += fE0;
    pE += fE1;
  }
  if (!keep) return;
  const long long ne = d.ne;
  if (l == 0) {
    d.fluxV[U_RHO * ne + e] = -sumJ;
    d.fluxV[U_E * ne + e] = pE;
    d.eKey[e] = (unsigned char)(it + (ranQR ? 19 : 0));
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) d.fluxV[(U_S + row[k]) * ne + e] = pS[k];
  if (l == 1) d.fluxV[(U_S + 8) * ne + e] = pS[4];
}

}  // namespace rx
