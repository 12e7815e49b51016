// TEST INFRASTRUCTURE — pins the oracle's restatement of Eigen::ColPivHouseholderQR<MatrixXd>::solve (rx_oracle.cpp: colPivQRSolve)
// and of the BiCGSTAB sums against the REAL Eigen 3.3.7 headers the reference vendors (externals/Eigen), compiled with the
// reference's own flags (oracle/build_ref.sh).  Built into oracle/_ref/eigen_qr_check by build_ref.sh when /root/reference is present;
// run by tests/test_oracle_golden.py::test_qr_restatement_equals_eigen_bit_for_bit.  Exit status 0 = every solution bit-identical.
//
// Matrices: (1) the reference's molar->mass gradient operator (numerics_direct_reactive.cpp:860-880) for random compositions of the
// nine species, which is what the SST closure factorises (:743); (2) dense random matrices with entries over several decades;
// (3) rank-deficient and zero matrices (pivot threshold, tau = 0 reflectors); (4) orders 1..9.
#include <Eigen/Dense>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>
#include "rx_oracle.h"

static long check(const Eigen::MatrixXd& M, const Eigen::MatrixXd& G, int row0, long& worst_ulps) {
  const int n = (int)M.rows();
  long bad = 0;
  for (int d = 0; d < G.cols(); ++d) {
    Eigen::VectorXd x = M.colPivHouseholderQr().solve(G.col(d).segment(row0, n));   // the call of numerics_direct_reactive.cpp:743
    std::vector<double> b(n), y(n);
    for (int i = 0; i < n; ++i) b[i] = G(row0 + i, d);
    rxo_colpiv_qr_solve(n, M.data(), b.data(), y.data());
    for (int i = 0; i < n; ++i) {
      long long a, c;
      std::memcpy(&a, &x[i], 8); std::memcpy(&c, &y[i], 8);
      if (a != c) { ++bad; long u = (long)std::llabs(a - c); if (u > worst_ulps) worst_ulps = u; }
    }
  }
  return bad;
}

int main(int argc, char** argv) {
  const long N = argc > 1 ? atol(argv[1]) : 20000;
  std::mt19937_64 rng(12345);
  std::uniform_real_distribution<double> U(0.0, 1.0);
  const double Mm[9] = {2.016, 31.998, 1.008, 15.999, 17.007, 33.006, 34.014, 18.015, 28.014};
  long bad = 0, worst = 0, cases = 0;
  for (long it = 0; it < N; ++it) {   // (1)
    const int n = 9;
    double Y[9], X[9], sy = 0, sx = 0;
    for (int s = 0; s < n; ++s) { Y[s] = std::pow(10.0, -12.0 * U(rng) * (U(rng) < 0.5)) * U(rng); sy += Y[s]; }
    for (int s = 0; s < n; ++s) { Y[s] /= sy; X[s] = Y[s] / Mm[s]; sx += X[s]; }
    for (int s = 0; s < n; ++s) X[s] /= sx;
    double sigma = 0, mt = 0;
    for (int s = 0; s < n; ++s) { sigma += X[s]; mt += Mm[s]; }
    Eigen::MatrixXd M(n, n);
    for (int i = 0; i < n; ++i) for (int j = 0; j < n; ++j)
      M(i, j) = i == j ? mt / Mm[i] * (Y[i] - X[i] + sigma) : mt * (Y[i] / Mm[i] - X[i] / Mm[j]);
    Eigen::MatrixXd G(13, 2);
    for (int i = 0; i < 13; ++i) for (int d = 0; d < 2; ++d) G(i, d) = (U(rng) - 0.5) * std::pow(10.0, 6.0 * U(rng) - 3.0);
    bad += check(M, G, 3, worst); ++cases;
  }
  for (long it = 0; it < N; ++it) {   // (2) and (4)
    const int n = 1 + (int)(it % 9);
    Eigen::MatrixXd M(n, n), G(n + 2, 2);
    for (int i = 0; i < n; ++i) for (int j = 0; j < n; ++j) M(i, j) = (U(rng) - 0.5) * std::pow(10.0, 8.0 * U(rng) - 4.0);
    for (int i = 0; i < n + 2; ++i) for (int d = 0; d < 2; ++d) G(i, d) = U(rng) - 0.5;
    bad += check(M, G, (int)(it % 3), worst); ++cases;
  }
  for (long it = 0; it < N / 4; ++it) {   // (3)
    const int n = 2 + (int)(it % 8), r = (int)(it % n);
    Eigen::MatrixXd B(n, std::max(r, 1)), C(std::max(r, 1), n), G(n, 2);
    for (int i = 0; i < n; ++i) for (int j = 0; j < B.cols(); ++j) { B(i, j) = U(rng) - 0.5; C(j, i) = U(rng) - 0.5; }
    Eigen::MatrixXd M = r == 0 ? Eigen::MatrixXd::Zero(n, n) : Eigen::MatrixXd(B * C);
    if (it % 5 == 0) M.col(n - 1).setZero();
    if (it % 7 == 0) M.row(0).setZero();
    for (int i = 0; i < n; ++i) for (int d = 0; d < 2; ++d) G(i, d) = U(rng) - 0.5;
    bad += check(M, G, 0, worst); ++cases;
  }
  std::printf("EIGEN_QR_CHECK cases=%ld mismatched_entries=%ld worst_ulps=%ld eigen=%d.%d.%d\n", cases, bad, worst,
              EIGEN_WORLD_VERSION, EIGEN_MAJOR_VERSION, EIGEN_MINOR_VERSION);
  return bad ? 1 : 0;
}
