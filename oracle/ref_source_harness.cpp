// TEST INFRASTRUCTURE — not on the product path.
//
// ref_source_harness: the direct-call oracle of SURVEY.md 7.1 step 0(i).  It drives the reference's own
// Framework::ReactingModelLibrary (linked from oracle/_ref/libsu2ref.a) over independent points exactly as
// CSourceReactive::ComputeChemistry does (SU2_CFD/src/numerics_direct_reactive.cpp:1728-1825):
//     SetSourceTerm(T, rho, Ys)                                   (reacting_model_library.cpp:99-114)
//     RANS:    Set_DfrDrhos(T, rho); AssemblePaSRConstant(omega_turb, C_mu, PaSR_lb); GetMassProductionTerm(i)   (:122-190, :143-154)
//     laminar: GetMassProductionTerm()                            (:196-202)
//     res[s] = omega[s] * (-Volume / (Density_Ref / Time_Ref)),   reference values 1 (DIMENSIONAL)
// and pins rxo_source_points (oracle/rx_oracle.cpp) and k_source_points on the states the source microbenchmark is timed on.
//
// usage (cwd = a case directory holding Mixture/ Chemistry/ Thermo/ Transp/): ref_source_harness <config_lib_file> <in.bin> <out.bin>
//   in.bin : int64 n, int64 rans, double cMu, double pasrLb, then T[n], rho[n], Y[n][Ns], omegaTurb[n], vol[n] (doubles)
//   out.bin: res[n][Ns] doubles
#include "../../Common/include/Framework/reacting_model_library.hpp"
#include <cstdint>
#include <cstdio>
#include <vector>

int main(int argc, char** argv) {
  if (argc < 4) { fprintf(stderr, "usage: ref_source_harness <config_lib_file> <in.bin> <out.bin>\n"); return 1; }
  using namespace Framework;
  ReactingModelLibrary lib(argv[1], "");
  lib.Setup();
  const unsigned short Ns = lib.nSpecies;
  FILE* f = fopen(argv[2], "rb");
  if (!f) { perror(argv[2]); return 2; }
  int64_t n = 0, rans = 0; double cMu = 0, lb = 0;
  if (fread(&n, 8, 1, f) != 1 || fread(&rans, 8, 1, f) != 1 || fread(&cMu, 8, 1, f) != 1 || fread(&lb, 8, 1, f) != 1) return 3;
  std::vector<double> T(n), rho(n), Y((size_t)n * Ns), om(n), vol(n), res((size_t)n * Ns);
  if (fread(T.data(), 8, n, f) != (size_t)n || fread(rho.data(), 8, n, f) != (size_t)n || fread(Y.data(), 8, (size_t)n * Ns, f) != (size_t)n * Ns ||
      fread(om.data(), 8, n, f) != (size_t)n || fread(vol.data(), 8, n, f) != (size_t)n) return 3;
  fclose(f);
  std::vector<double> Ys(Ns), omega(Ns);
  for (int64_t p = 0; p < n; ++p) {
    for (unsigned short s = 0; s < Ns; ++s) Ys[s] = Y[(size_t)p * Ns + s];
    lib.SetSourceTerm(T[p], rho[p], Ys);
    if (rans) {
      lib.Set_DfrDrhos(T[p], rho[p]);
      lib.AssemblePaSRConstant(om[p], cMu, lb);
      for (unsigned short s = 0; s < Ns; ++s) omega[s] = lib.GetMassProductionTerm(s);
    } else {
      Eigen::VectorXd::Map(&omega[0], Ns) = lib.GetMassProductionTerm();
    }
    for (unsigned short s = 0; s < Ns; ++s) {
      double r = omega[s];
      r *= -vol[p] / (1.0 / 1.0);
      res[(size_t)p * Ns + s] = r;
    }
  }
  f = fopen(argv[3], "wb");
  if (!f) { perror(argv[3]); return 2; }
  fwrite(res.data(), 8, res.size(), f);
  fclose(f);
  printf("ref_source_harness: %lld points, Ns %d, %s\n", (long long)n, (int)Ns, rans ? "PaSR" : "laminar");
  return 0;
}
