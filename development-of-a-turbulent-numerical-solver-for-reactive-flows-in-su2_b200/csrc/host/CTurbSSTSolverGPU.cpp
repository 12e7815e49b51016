// CTurbSSTSolverGPU — see the header.  Compiled with the reference's include paths (oracle/build_ref.sh; -fno-access-control for
// the blending-function members of CTurbSSTVariable, for which upstream would add three setters).
#include "CTurbSSTSolverGPU.hpp"
#include "CReactiveNSSolverGPU.hpp"

#include "variable_structure.hpp"

#include <cstdlib>
#include <stdexcept>
#include <string>
#include <vector>

namespace {
void check_rc(rx_ctx* ctx, int rc, const char* where) {
  if (rc != RX_OK) throw std::runtime_error(std::string(where) + ": " + rx_last_error(ctx) + " (librx_cuda, status " + std::to_string(rc) + ")");
}
CReactiveNSSolverGPU* flow_of(CSolver** solver_container) {
  CReactiveNSSolverGPU* flow = dynamic_cast<CReactiveNSSolverGPU*>(solver_container[FLOW_SOL]);
  if (!flow) throw std::runtime_error("CTurbSSTSolverGPU: the flow solver is not a CReactiveNSSolverGPU (the turbulence step runs on its device state)");
  return flow;
}
}  // namespace

void CTurbSSTSolverGPU::SetWallDistance(rx_ctx* ctx, CGeometry* geometry) {
  if (wallDistSet) return;
  std::vector<double> dist(nPoint);
  for (unsigned long p = 0; p < nPoint; ++p) dist[p] = geometry->node[p]->GetWall_Distance();
  check_rc(ctx, rx_set_wall_distance(ctx, dist.data()), "rx_set_wall_distance");
  wallDistSet = true;
}

// mirrors of (k, omega, F1, F2, CDkw, mu_t, gradients) for output / restart / monitors, and the flow solver's host copy of mu_t
void CTurbSSTSolverGPU::MirrorFromDevice(rx_ctx* ctx, CReactiveNSSolverGPU* flow) {
  std::vector<double> post((size_t)nPoint * 12);
  check_rc(ctx, rx_get_turbulence(ctx, post.data()), "rx_get_turbulence");
  flow->SetTurbulenceOnDevice(post.data());   // the fields the flow path reads are current on the device: no upload at its next Preprocessing
  for (unsigned long p = 0; p < nPoint; ++p) {
    const double* q = &post[p * 12];
    CTurbSSTVariable* n = static_cast<CTurbSSTVariable*>(node[p]);
    n->SetSolution(0, q[0]); n->SetSolution(1, q[1]);
    n->F1 = q[3]; n->F2 = q[4]; n->CDkw = q[5];
    n->SetmuT(q[6]);
    n->SetGradient(0, 0, q[7]); n->SetGradient(0, 1, q[8]); n->SetGradient(1, 0, q[9]); n->SetGradient(1, 1, q[10]);
  }
}

void CTurbSSTSolverGPU::Postprocessing(CGeometry* geometry, CSolver** solver_container, CConfig*, unsigned short) {
  if (stepped) return;   // every step ends with it on the device
  CReactiveNSSolverGPU* flow = flow_of(solver_container);
  rx_ctx* ctx = flow->Context();
  SetWallDistance(ctx, geometry);
  std::vector<double> kw((size_t)nPoint * 2);
  for (unsigned long p = 0; p < nPoint; ++p) { kw[2 * p] = node[p]->GetSolution(0); kw[2 * p + 1] = node[p]->GetSolution(1); }
  check_rc(ctx, rx_turb_postprocess(ctx, kw.data(), node[0]->Get_Sigmak(), 0), "rx_turb_postprocess");
  MirrorFromDevice(ctx, flow);
  hostDirty = false;
}

void CTurbSSTSolverGPU::ImplicitEuler_Iteration(CGeometry* geometry, CSolver** solver_container, CConfig* config) {
  CReactiveNSSolverGPU* flow = flow_of(solver_container);
  rx_ctx* ctx = flow->Context();
  auto check = [&](int rc, const char* where) { check_rc(ctx, rc, where); };
  if (const char* e = std::getenv("SU2_RX_MIRROR")) mirror = std::atoi(e);
  SetWallDistance(ctx, geometry);
  if (hostDirty) {   // after a restart / a change of the host state: the host state is the truth
    std::vector<double> kw((size_t)nPoint * 2), bl((size_t)nPoint * 3);
    for (unsigned long p = 0; p < nPoint; ++p) {
      kw[2 * p] = node[p]->GetSolution(0); kw[2 * p + 1] = node[p]->GetSolution(1);
      bl[3 * p] = node[p]->GetF1blending(); bl[3 * p + 1] = node[p]->GetF2blending(); bl[3 * p + 2] = node[p]->GetCrossDiff();
    }
    check(rx_turb_set_state(ctx, kw.data(), bl.data()), "rx_turb_set_state");
    hostDirty = false;
  }
  if (dualTime) {
    // node[] holds (k, omega) after every step (MirrorFromDevice), so CIntegration::SetDualTime_Solver shifts the right values there; the two
    // levels — 4 doubles per point, against the 12 the mirror brings back — go up with every step, which needs no hook in the integration class
    std::vector<double> a((size_t)nPoint * 2), b((size_t)nPoint * 2);
    for (unsigned long p = 0; p < nPoint; ++p)
      for (unsigned short v = 0; v < 2; ++v) { a[2 * p + v] = node[p]->GetSolution_time_n()[v]; b[2 * p + v] = node[p]->GetSolution_time_n1()[v]; }
    check(rx_turb_set_time_levels(ctx, a.data(), b.data()), "rx_turb_set_time_levels");
  }
  rx_turb_opts o{};
  o.kineInf = kine_Inf; o.omegaInf = omega_Inf;
  o.cflRed = config->GetCFLRedCoeff_Turb(); o.relax = config->GetRelaxation_Factor_Turb();
  o.linIter = (int)config->GetLinear_Solver_Iter(); o.linTol = config->GetLinear_Solver_Error();
  double rms[2];
  int nLin = 0;
  check(rx_turb_iterate(ctx, &o, rms, &nLin), "rx_turb_iterate");
  stepped = true;
  for (unsigned short v = 0; v < nVar; ++v) SetRes_RMS(v, rms[v]);
  MirrorFromDevice(ctx, flow);
  if (mirror >= 2) {
    std::vector<double> r((size_t)nPoint * 2), x((size_t)nPoint * 2);
    check(rx_turb_get_step(ctx, r.data(), x.data()), "rx_turb_get_step");
    for (unsigned long k = 0; k < nPoint * 2; ++k) { LinSysRes[k] = r[k]; LinSysSol[k] = x[k]; }
  }
}
