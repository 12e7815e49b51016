// TEST INFRASTRUCTURE — not on the product path.
//
// ref_harness: drives the *reference's own* classes (linked from oracle/_ref/libsu2ref.a, built by
// oracle/build_ref.sh from /root/reference) through one solver iteration at a time, in exactly the
// call order of CMultiGridIntegration::MultiGrid_Cycle / MultiGrid_Iteration
// (SU2_CFD/src/integration_time.cpp:42-181) and CIntegration::Space_Integration
// (SU2_CFD/src/integration_structure.cpp:53-193), and dumps every intermediate array of the hot path
// as .npy files.  These dumps are the golden vectors that pin oracle/rx_oracle.cpp and the CUDA path.
//
// Built with -fno-access-control so that protected containers of CDriver and the private tables of
// Framework::ReactingModelLibrary can be read without editing any reference source.
//
// usage: ref_harness <cfg> <nIter> <outdir> [dump_from_iter] [perturb_seed]
#include "../../SU2_CFD/include/driver_structure.hpp"
#include "../../SU2_CFD/include/solver_reactive.hpp"
#include "../../Common/include/Framework/reacting_model_library.hpp"
#ifdef RX_WITH_GPU
// ref_harness_gpu: the same driver loop and dumps, with the flow solver object replaced by the CUDA-backed subclass
// (package csrc/host/CReactiveNSSolverGPU) when SU2_RX_BACKEND=cuda — the drop-in demonstration of INTEGRATION.md.
#include "CReactiveNSSolverGPU.hpp"
#include "CTurbSSTSolverGPU.hpp"
#include <cstring>
#endif
#include <cstdio>
#include <cstring>
#include <cstdint>
#include <string>
#include <vector>
#include <sstream>
#include <sys/stat.h>
#include <time.h>

static std::string g_out;

static void npy_write(const std::string& name, const void* data, const char* descr, size_t elsize,
                      const std::vector<size_t>& shape) {
  std::string path = g_out + "/" + name + ".npy";
  FILE* f = fopen(path.c_str(), "wb");
  if (!f) { perror(path.c_str()); exit(2); }
  std::ostringstream hd;
  hd << "{'descr': '" << descr << "', 'fortran_order': False, 'shape': (";
  size_t n = 1;
  for (size_t i = 0; i < shape.size(); ++i) { hd << shape[i] << (shape.size() == 1 || i + 1 < shape.size() ? "," : ""); n *= shape[i]; }
  hd << "), }";
  std::string h = hd.str();
  size_t total = 10 + h.size() + 1;
  size_t pad = (64 - total % 64) % 64;
  h.append(pad, ' ');
  h.push_back('\n');
  unsigned char pre[10] = {0x93, 'N', 'U', 'M', 'P', 'Y', 1, 0, (unsigned char)(h.size() & 0xff), (unsigned char)(h.size() >> 8)};
  fwrite(pre, 1, 10, f);
  fwrite(h.data(), 1, h.size(), f);
  fwrite(data, elsize, n, f);
  fclose(f);
}
static void dump_f64(const std::string& name, const std::vector<double>& v, std::vector<size_t> shape) {
  npy_write(name, v.data(), "<f8", 8, shape);
}
static void dump_i64(const std::string& name, const std::vector<int64_t>& v, std::vector<size_t> shape) {
  npy_write(name, v.data(), "<i8", 8, shape);
}

struct Ctx {
  CFluidDriver* drv;
  CGeometry* geo;
  CConfig* cfg;
  CSolver** sol;
  CReactiveEulerSolver* flow;
  CNumerics** num;
  CIntegration* integ;
  unsigned long nPoint, nPointDomain;
  unsigned short nVar, nPrimVar, nPrimVarGrad, nDim, nSpecies;
  bool rans, viscous;
};

static std::string tag(unsigned long it, int rk, const char* what) {
  char b[128];
  snprintf(b, sizeof b, "it%04lu_rk%d_%s", it, rk, what);
  return b;
}

static void dump_solution(Ctx& c, const std::string& name) {
  std::vector<double> U(c.nPoint * c.nVar);
  for (unsigned long p = 0; p < c.nPoint; ++p)
    for (unsigned short v = 0; v < c.nVar; ++v) U[p * c.nVar + v] = c.flow->node[p]->GetSolution(v);
  dump_f64(name, U, {c.nPoint, c.nVar});
}
static void dump_residual(Ctx& c, const std::string& name) {
  std::vector<double> R(c.nPoint * c.nVar);
  for (unsigned long p = 0; p < c.nPoint; ++p) {
    su2double* b = c.flow->LinSysRes.GetBlock(p);
    for (unsigned short v = 0; v < c.nVar; ++v) R[p * c.nVar + v] = b[v];
  }
  dump_f64(name, R, {c.nPoint, c.nVar});
}
static void dump_primitive(Ctx& c, const std::string& pre) {
  std::vector<double> V(c.nPoint * c.nPrimVar), aux(c.nPoint * 6), G;
  for (unsigned long p = 0; p < c.nPoint; ++p) {
    CVariable* n = c.flow->node[p];
    for (unsigned short v = 0; v < c.nPrimVar; ++v) V[p * c.nPrimVar + v] = n->GetPrimitive(v);
    aux[p * 6 + 0] = n->GetSpecificHeatCp();
    aux[p * 6 + 1] = n->GetdPdU()[c.nDim + 1] + 1.0;  // Gamma = dPdU[rhoE] + 1
    aux[p * 6 + 2] = c.viscous ? n->GetLaminarViscosity() : 0.0;
    aux[p * 6 + 3] = c.viscous ? n->GetThermalConductivity() : 0.0;
    aux[p * 6 + 4] = c.viscous ? n->GetVorticity()[2] : 0.0;
    aux[p * 6 + 5] = c.viscous ? n->GetStrainMag() : 0.0;
  }
  dump_f64(pre + "V", V, {c.nPoint, c.nPrimVar});
  if (c.flow->limiter) {
    std::vector<double> L(c.nPoint * 4);
    for (unsigned long p = 0; p < c.nPoint; ++p)
      for (unsigned short v = 0; v < 4; ++v) L[p * 4 + v] = c.flow->node[p]->GetLimiter_Primitive()[v];
    dump_f64(pre + "lim", L, {c.nPoint, 4});
  }
  dump_f64(pre + "aux", aux, {c.nPoint, 6});
  if (c.viscous) {
    G.resize(c.nPoint * c.nPrimVarGrad * c.nDim);
    for (unsigned long p = 0; p < c.nPoint; ++p) {
      su2double** g = c.flow->node[p]->GetGradient_Primitive();
      for (unsigned short v = 0; v < c.nPrimVarGrad; ++v)
        for (unsigned short d = 0; d < c.nDim; ++d) G[(p * c.nPrimVarGrad + v) * c.nDim + d] = g[v][d];
    }
    dump_f64(pre + "grad", G, {c.nPoint, c.nPrimVarGrad, c.nDim});
  }
}
static void dump_turb(Ctx& c, const std::string& name) {
  // columns: muT, k, omega, dk/dx, dk/dy, sigma_k
  std::vector<double> T(c.nPoint * 6, 0.0);
  if (c.rans) {
    CSolver* ts = c.sol[TURB_SOL];
    for (unsigned long p = 0; p < c.nPoint; ++p) {
      T[p * 6 + 0] = ts->node[p]->GetmuT();
      T[p * 6 + 1] = ts->node[p]->GetSolution(0);
      T[p * 6 + 2] = ts->node[p]->GetSolution(1);
      T[p * 6 + 3] = ts->node[p]->GetGradient()[0][0];
      T[p * 6 + 4] = ts->node[p]->GetGradient()[0][1];
      T[p * 6 + 5] = ts->node[p]->Get_Sigmak();
    }
  }
  dump_f64(name, T, {c.nPoint, 6});
}
static void dump_turb_post(Ctx& c, const std::string& name) {
  // state of CTurbSSTSolver right after its iteration, i.e. after CTurbSSTSolver::Postprocessing (solver_direct_turbulent.cpp:2953):
  // columns k, omega, wall distance, F1, F2, CDkw, muT, dk/dx, dk/dy, domega/dx, domega/dy, sigma_k
  std::vector<double> T(c.nPoint * 12, 0.0);
  if (c.rans) {
    CSolver* ts = c.sol[TURB_SOL];
    for (unsigned long p = 0; p < c.nPoint; ++p) {
      CVariable* n = ts->node[p];
      double* t = &T[p * 12];
      t[0] = n->GetSolution(0); t[1] = n->GetSolution(1); t[2] = c.geo->node[p]->GetWall_Distance();
      t[3] = n->GetF1blending(); t[4] = n->GetF2blending(); t[5] = n->GetCrossDiff(); t[6] = n->GetmuT();
      t[7] = n->GetGradient()[0][0]; t[8] = n->GetGradient()[0][1]; t[9] = n->GetGradient()[1][0]; t[10] = n->GetGradient()[1][1];
      t[11] = n->Get_Sigmak();
    }
  }
  dump_f64(name, T, {c.nPoint, 12});
}
static void dump_dt(Ctx& c, const std::string& name) {
  std::vector<double> D(c.nPoint * 3, 0.0);
  for (unsigned long p = 0; p < c.nPointDomain; ++p) {
    D[p * 3 + 0] = c.flow->node[p]->GetDelta_Time();
    D[p * 3 + 1] = c.flow->node[p]->GetMax_Lambda_Inv();
    D[p * 3 + 2] = c.viscous ? c.flow->node[p]->GetMax_Lambda_Visc() : 0.0;
  }
  dump_f64(name, D, {c.nPoint, 3});
}

// ---- geometry: everything the path reads from CGeometry, as flat arrays -------------------------
static void dump_geometry(Ctx& c, std::ostringstream& js) {
  CGeometry* g = c.geo;
  unsigned long nEdge = g->GetnEdge();
  std::vector<int64_t> en(nEdge * 2), gi(c.nPoint), dom(c.nPoint);
  std::vector<double> enrm(nEdge * 2), xy(c.nPoint * 2), vol(c.nPoint);
  for (unsigned long e = 0; e < nEdge; ++e) {
    en[2 * e] = g->edge[e]->GetNode(0); en[2 * e + 1] = g->edge[e]->GetNode(1);
    enrm[2 * e] = g->edge[e]->GetNormal()[0]; enrm[2 * e + 1] = g->edge[e]->GetNormal()[1];
  }
  std::vector<int64_t> nb_ptr(c.nPoint + 1, 0), nb_pt, nb_edge;
  for (unsigned long p = 0; p < c.nPoint; ++p) {
    xy[2 * p] = g->node[p]->GetCoord(0); xy[2 * p + 1] = g->node[p]->GetCoord(1);
    vol[p] = g->node[p]->GetVolume();
    gi[p] = g->node[p]->GetGlobalIndex();
    dom[p] = g->node[p]->GetDomain() ? 1 : 0;
    for (unsigned short k = 0; k < g->node[p]->GetnPoint(); ++k) {
      nb_pt.push_back(g->node[p]->GetPoint(k));
      nb_edge.push_back(g->node[p]->GetEdge(k));
    }
    nb_ptr[p + 1] = nb_pt.size();
  }
  dump_i64("geo_edge_nodes", en, {nEdge, 2});
  dump_f64("geo_edge_normal", enrm, {nEdge, 2});
  dump_f64("geo_coord", xy, {c.nPoint, 2});
  dump_f64("geo_volume", vol, {c.nPoint});
  dump_i64("geo_global_index", gi, {c.nPoint});
  dump_i64("geo_domain", dom, {c.nPoint});
  dump_i64("geo_nb_ptr", nb_ptr, {c.nPoint + 1});
  dump_i64("geo_nb_point", nb_pt, {nb_pt.size()});
  dump_i64("geo_nb_edge", nb_edge, {nb_edge.size()});
  js << "\"nPoint\": " << c.nPoint << ", \"nPointDomain\": " << c.nPointDomain << ", \"nEdge\": " << nEdge << ",\n";
  js << "\"markers\": [";
  for (unsigned short m = 0; m < g->GetnMarker(); ++m) {
    unsigned long nv = g->nVertex[m];
    std::vector<int64_t> vn(nv), vnn(nv);
    std::vector<double> vnrm(nv * 2);
    for (unsigned long v = 0; v < nv; ++v) {
      vn[v] = g->vertex[m][v]->GetNode();
      vnn[v] = g->vertex[m][v]->GetNormal_Neighbor();
      vnrm[2 * v] = g->vertex[m][v]->GetNormal()[0]; vnrm[2 * v + 1] = g->vertex[m][v]->GetNormal()[1];
    }
    std::string ms = "geo_marker" + std::to_string(m);
    dump_i64(ms + "_node", vn, {nv});
    dump_i64(ms + "_normal_neighbor", vnn, {nv});
    dump_f64(ms + "_normal", vnrm, {nv, 2});
    std::string t = c.cfg->GetMarker_All_TagBound(m);
    unsigned short kind = c.cfg->GetMarker_All_KindBC(m);
    js << (m ? ",\n  " : "\n  ") << "{\"tag\": \"" << t << "\", \"kind\": " << kind << ", \"nVertex\": " << nv;
    js.precision(17);
    if (kind == INLET_FLOW) {
      js << ", \"bc\": \"inlet\", \"T\": " << c.cfg->GetInlet_Ttotal(t) << ", \"Vmag\": " << c.cfg->GetInlet_Ptotal(t)
         << ", \"dir\": [" << c.cfg->GetInlet_FlowDir(t)[0] << ", " << c.cfg->GetInlet_FlowDir(t)[1] << "], \"Ys\": [";
      for (unsigned short s = 0; s < c.nSpecies; ++s) js << (s ? ", " : "") << c.cfg->GetInlet_MassFrac(t)[s];
      js << "]";
    } else if (kind == SUPERSONIC_INLET) {   // BC_Supersonic_Inlet (:2998): temperature, pressure, velocity vector, mass fractions
      js << ", \"bc\": \"supersonic_inlet\", \"T\": " << c.cfg->GetInlet_Temperature(t) << ", \"P\": " << c.cfg->GetInlet_Pressure(t)
         << ", \"dir\": [" << c.cfg->GetInlet_Velocity(t)[0] << ", " << c.cfg->GetInlet_Velocity(t)[1] << "], \"Ys\": [";
      for (unsigned short s = 0; s < c.nSpecies; ++s) js << (s ? ", " : "") << c.cfg->GetInlet_MassFrac(t)[s];
      js << "]";
    } else if (kind == SUPERSONIC_OUTLET) {
      js << ", \"bc\": \"supersonic_outlet\"";
    } else if (kind == OUTLET_FLOW) {
      js << ", \"bc\": \"outlet\", \"P\": " << c.cfg->GetOutlet_Pressure(t);
    } else if (kind == ISOTHERMAL) {
      js << ", \"bc\": \"isothermal\", \"Twall\": " << c.cfg->GetIsothermal_Temperature(t);
    } else if (kind == HEAT_FLUX) {
      js << ", \"bc\": \"heatflux\", \"q\": " << c.cfg->GetWall_HeatFlux(t);
    } else if (kind == EULER_WALL) {
      js << ", \"bc\": \"euler_wall\"";
    } else {
      js << ", \"bc\": \"other\"";
    }
    js << "}";
  }
  js << "],\n";
}

// ---- chemistry library internals (Common/src/Framework/reacting_model_library.cpp) ---------------
static void dump_chem(Ctx& c, std::ostringstream& js) {
  using namespace Framework;
  ReactingModelLibrary lib(c.cfg->GetConfigLibFile(), c.cfg->GetLibraryPath());
  lib.Setup();
  unsigned short Ns = lib.nSpecies, Nr = lib.nReactions;
  js.precision(17);
  js << "\"chem\": {\"Ns\": " << Ns << ", \"Nr\": " << Nr << ", \"species\": [";
  std::vector<std::string> names(Ns);
  for (auto& kv : lib.Species_Names) names[kv.second] = kv.first;
  for (unsigned short s = 0; s < Ns; ++s) js << (s ? ", " : "") << "\"" << names[s] << "\"";
  auto vec = [&](const char* k, const std::vector<double>& v) {
    js << "], \"" << k << "\": [";
    for (size_t i = 0; i < v.size(); ++i) js << (i ? ", " : "") << v[i];
  };
  vec("mMasses", lib.mMasses); vec("Ri", lib.Ri); vec("Diff_Volumes", lib.Diff_Volumes);
  vec("Formation_Enthalpies", lib.Formation_Enthalpies);
  if (Nr) {
    vec("As", lib.As); vec("Betas", lib.Betas); vec("Temps_Activation", lib.Temps_Activation);
    std::vector<double> rev(Nr), avb(Nr, 0), Ab(Nr, 0), Bb(Nr, 0), Tb(Nr, 0);
    for (unsigned short r = 0; r < Nr; ++r) {
      rev[r] = lib.Reversible_Reactions[r];
      if (lib.Available_Backward_Rate.count(r)) { avb[r] = 1; Ab[r] = lib.As_back[r]; Bb[r] = lib.Betas_back[r]; Tb[r] = lib.Temps_Activation_back[r]; }
    }
    vec("Reversible", rev); vec("Available_Backward", avb); vec("As_back", Ab); vec("Betas_back", Bb); vec("Temps_Activation_back", Tb);
    // matrices, all written as [reaction][species]
    auto mat_sr = [&](const char* k, const Eigen::MatrixXd& M, bool species_rows) {
      js << "], \"" << k << "\": [";
      for (unsigned short r = 0; r < Nr; ++r)
        for (unsigned short s = 0; s < Ns; ++s) js << ((r || s) ? ", " : "") << (species_rows ? M(s, r) : M(r, s));
    };
    mat_sr("Stoich_Reactants", lib.Stoich_Coeffs_Reactants, true);
    mat_sr("Stoich_Products", lib.Stoich_Coeffs_Products, true);
    mat_sr("Stoich_Reactants_Exp", lib.Stoich_Coeffs_Reactants_Exp, false);
    mat_sr("Stoich_Products_Exp", lib.Stoich_Coeffs_Products_Exp, false);
  }
  js << "]},\n";
  // spline tables: [prop][species][knot], prop order mu, kappa, cp, h, s; x grid shared
  typedef decltype(lib.Mu_Spline) SplTab;
  const SplTab* tabs[5] = {&lib.Mu_Spline, &lib.Kappa_Spline, &lib.Cp_Spline, &lib.Enth_Spline, &lib.Entr_Spline};
  size_t n = std::get<0>(lib.Mu_Spline[0]).size();
  std::vector<double> X(std::get<0>(lib.Mu_Spline[0])), Y(5 * Ns * n), Y2(5 * Ns * n);
  for (int t = 0; t < 5; ++t)
    for (unsigned short s = 0; s < Ns; ++s) {
      const auto& tp = (*tabs[t])[s];
      if (std::get<0>(tp).size() != n) { fprintf(stderr, "spline grids differ\n"); exit(3); }
      for (size_t k = 0; k < n; ++k) {
        Y[(t * Ns + s) * n + k] = std::get<1>(tp)[k];
        Y2[(t * Ns + s) * n + k] = std::get<2>(tp).data()[k];  // y2[n-1] lives past size() (spline.cpp:54); read via data()
      }
    }
  dump_f64("chem_x", X, {n});
  dump_f64("chem_y", Y, {5, (size_t)Ns, n});
  dump_f64("chem_y2", Y2, {5, (size_t)Ns, n});
}

static void dump_config(Ctx& c, std::ostringstream& js) {
  CConfig* f = c.cfg;
  js.precision(17);
  js << "\"cfg\": {\"kind_solver\": " << f->GetKind_Solver() << ", \"rans\": " << (c.rans ? 1 : 0) << ", \"viscous\": " << (c.viscous ? 1 : 0)
     << ", \"turb_model\": " << f->GetKind_Turb_Model()
     << ", \"mach\": " << f->GetMach() << ", \"cfl\": " << f->GetCFL(MESH_0) << ", \"max_dt\": " << f->GetMax_DeltaTime()
     << ", \"Tmin\": " << f->GetTemperatureMin() << ", \"Tmax\": " << f->GetTemperatureMax()
     << ", \"clip_T\": " << (f->GetClipping_Temp() ? 1 : 0)
     << ", \"ignition\": " << (f->GetIgnition() ? 1 : 0) << ", \"ignition_iter\": " << f->GetIgnitionIter()
     << ", \"ignition_T\": " << f->GetIgnitionTemperature() << ", \"fuel_index\": " << f->GetFuelIndex()
     << ", \"oxidizer_index\": " << f->GetOxidizerIndex()
     << ", \"prandtl_lam\": " << f->GetPrandtl_Lam() << ", \"prandtl_turb\": " << f->GetPrandtl_Turb()
     << ", \"lewis_turb\": " << f->GetLewis_Turb() << ", \"c_mu\": " << f->Get_Cmu() << ", \"pasr_lb\": " << f->Get_PaSR_LB()
     << ", \"tke_inf\": " << (c.rans ? c.flow->GetTke_Inf() : 0.0)
     << ", \"kine_inf\": " << (c.rans ? static_cast<CTurbSSTSolver*>(c.sol[TURB_SOL])->kine_Inf : 0.0) << ", \"omega_inf\": " << (c.rans ? static_cast<CTurbSSTSolver*>(c.sol[TURB_SOL])->omega_Inf : 0.0)
     << ", \"cfl_red_turb\": " << f->GetCFLRedCoeff_Turb() << ", \"relax_turb\": " << f->GetRelaxation_Factor_Turb()
     << ", \"lin_iter\": " << f->GetLinear_Solver_Iter() << ", \"lin_tol\": " << f->GetLinear_Solver_Error()
     << ", \"spatial_order_turb\": " << f->GetSpatialOrder_Turb()
     << ", \"grad_method\": " << f->GetKind_Gradient_Method() << ", \"time_scheme\": " << f->GetKind_TimeIntScheme_Flow()
     << ", \"kind_inlet\": " << f->GetKind_Inlet()
     << ", \"time_stepping\": " << (f->GetUnsteady_Simulation() == TIME_STEPPING ? 1 : 0) << ", \"unst_cfl\": " << f->GetUnst_CFL()
     << ", \"unst_dt\": " << f->GetDelta_UnstTime()
     << ", \"dual_time\": " << (f->GetUnsteady_Simulation() == DT_STEPPING_1ST ? 1 : f->GetUnsteady_Simulation() == DT_STEPPING_2ND ? 2 : 0)
     << ", \"visc_limiter\": " << (f->GetViscous_Limiter_Flow() ? 1 : 0) << ", \"unst_dt_nd\": " << f->GetDelta_UnstTimeND() << ", \"dual_inner\": " << (getenv("RX_DUAL_INNER") ? atol(getenv("RX_DUAL_INNER")) : 2)
     << ", \"second_order\": " << (c.flow->second_order ? 1 : 0) << ", \"limiter\": " << (c.flow->limiter ? 1 : 0)
     << ", \"limiter_kind\": " << f->GetKind_SlopeLimit_Flow() << ", \"limiter_coeff\": " << f->GetLimiterCoeff()
     << ", \"ref_elem_length\": " << f->GetRefElemLength() << ", \"limiter_iter\": " << f->GetLimiterIter()
     << ", \"nRK\": " << f->GetnRKStep() << ", \"rk_alpha\": [";
  for (unsigned short k = 0; k < f->GetnRKStep(); ++k) js << (k ? ", " : "") << f->Get_Alpha_RKStep(k);
  js << "], \"lower\": [";
  for (unsigned short v = 0; v < c.nVar; ++v) js << (v ? ", " : "") << c.flow->Lower_Limit[v];
  js << "], \"upper\": [";
  for (unsigned short v = 0; v < c.nVar; ++v) js << (v ? ", " : "") << c.flow->Upper_Limit[v];
  js << "]}\n";
}

// ---- boundary loop, same order as CIntegration::Space_Integration (integration_structure.cpp:89-190)
static void apply_bcs(Ctx& c) {
  CSolver* s = c.flow;
  for (unsigned short m = 0; m < c.cfg->GetnMarker_All(); ++m) {
    switch (c.cfg->GetMarker_All_KindBC(m)) {
      case EULER_WALL: s->BC_Euler_Wall(c.geo, c.sol, c.num[CONV_BOUND_TERM], c.cfg, m); break;
      case INLET_FLOW: s->BC_Inlet(c.geo, c.sol, c.num[CONV_BOUND_TERM], c.num[VISC_BOUND_TERM], c.cfg, m); break;
      case OUTLET_FLOW: s->BC_Outlet(c.geo, c.sol, c.num[CONV_BOUND_TERM], c.num[VISC_BOUND_TERM], c.cfg, m); break;
      case SUPERSONIC_INLET: s->BC_Supersonic_Inlet(c.geo, c.sol, c.num[CONV_BOUND_TERM], c.num[VISC_BOUND_TERM], c.cfg, m); break;
      case SUPERSONIC_OUTLET: s->BC_Supersonic_Outlet(c.geo, c.sol, c.num[CONV_BOUND_TERM], c.num[VISC_BOUND_TERM], c.cfg, m); break;
      case SYMMETRY_PLANE: s->BC_Sym_Plane(c.geo, c.sol, c.num[CONV_BOUND_TERM], c.num[VISC_BOUND_TERM], c.cfg, m); break;
      default: break;
    }
  }
  for (unsigned short m = 0; m < c.cfg->GetnMarker_All(); ++m) {
    switch (c.cfg->GetMarker_All_KindBC(m)) {
      case ISOTHERMAL: s->BC_Isothermal_Wall(c.geo, c.sol, c.num[CONV_BOUND_TERM], c.num[VISC_BOUND_TERM], c.cfg, m); break;
      case HEAT_FLUX: s->BC_HeatFlux_Wall(c.geo, c.sol, c.num[CONV_BOUND_TERM], c.num[VISC_BOUND_TERM], c.cfg, m); break;
      default: break;
    }
  }
}


// ---- synthetic non-uniform initial state (so that every term of the path is exercised) ----------
// Smooth fields + seeded noise on (T, P, u, v, Y_s, k, omega); conserved variables are rebuilt with the
// reference's own library (ComputeEnthalpy / ComputeRgas), so the state is thermodynamically consistent.
#include <random>
static double g_noise = 1.0;  // env RX_PERTURB_NOISE: scale of the point-wise random part (0: smooth fields, for boundary conditions that need a consistent state)
static double g_tamp = 800.0;   // env RX_PERTURB_TAMP: peak of the temperature bump (2000 => burning-state temperatures up to 2500 K)
static void perturb_state(Ctx& c, unsigned long seed) {
  using namespace Framework;
  ReactingModelLibrary lib(c.cfg->GetConfigLibFile(), c.cfg->GetLibraryPath());
  lib.Setup();
  unsigned short Ns = c.nSpecies;
  double x0 = 1e300, x1 = -1e300, y0 = 1e300, y1 = -1e300;
  for (unsigned long p = 0; p < c.nPoint; ++p) {
    double x = c.geo->node[p]->GetCoord(0), y = c.geo->node[p]->GetCoord(1);
    x0 = std::min(x0, x); x1 = std::max(x1, x); y0 = std::min(y0, y); y1 = std::max(y1, y);
  }
  std::vector<double> U(c.nVar);
  for (unsigned long p = 0; p < c.nPoint; ++p) {
    unsigned long gi = c.geo->node[p]->GetGlobalIndex();
    std::mt19937_64 gen(seed * 1000003ULL + gi);
    std::uniform_real_distribution<double> un(-1.0, 1.0);
    double xh = (c.geo->node[p]->GetCoord(0) - x0) / (x1 - x0), yh = (c.geo->node[p]->GetCoord(1) - y0) / (y1 - y0);
    double sh = 0.5 + 0.5 * std::tanh((yh - 0.5) * 6.0);
    double T = 500.0 + g_tamp * std::exp(-std::pow((yh - 0.5) / 0.2, 2)) * (0.6 + 0.4 * std::sin(6.283185307179586 * xh)) + 8.0 * g_noise * un(gen);
    double P = 101325.0 * (1.0 + 0.05 * std::sin(3.0 * xh + yh)) * (1.0 + 1e-3 * g_noise * un(gen));
    double u = 30.0 * 4.0 * yh * (1.0 - yh) + 3.0 * std::sin(5.0 * xh) + 0.2 * g_noise * un(gen);
    double v = 2.0 * std::sin(6.283185307179586 * xh) * std::cos(3.141592653589793 * yh) + 0.2 * g_noise * un(gen);
    std::vector<double> Y(Ns, 0.0);
    for (unsigned short s = 0; s < Ns; ++s) Y[s] = 0.02 * (1.0 + std::sin(1.7 * s + 4.0 * xh + 2.0 * yh)) * (1.0 + 0.1 * g_noise * un(gen));
    Y[c.cfg->GetFuelIndex()] += 1.0 - sh;
    Y[c.cfg->GetOxidizerIndex()] += sh;
    if (yh > 0.7) Y[c.cfg->GetFuelIndex()] = 0.0;          // exact zeros: exercise the 1e-15 / 1e-10 thresholds
    if (xh < 0.3) Y[Ns - 1] = 0.0;
    if (xh > 0.8 && yh < 0.3) { for (unsigned short s = 0; s < Ns; ++s) Y[s] = 0.0; Y[c.cfg->GetOxidizerIndex()] = 1.0; }
    double sum = 0.0; for (double w : Y) sum += w;
    for (double& w : Y) w /= sum;
    double k = 0.0;
    if (c.rans) {
      k = 0.05 + 5.0 * yh * (1.0 - yh) * (1.0 + 0.2 * g_noise * un(gen));
      double om = 2.0e3 * (1.0 + 0.8 * std::sin(3.0 * xh) * std::cos(2.0 * yh)) * (1.0 + 0.1 * g_noise * un(gen));
      CVariable* t = c.sol[TURB_SOL]->node[p];
      double rho_ = P / (lib.ComputeRgas(Y) * T);
      t->SetSolution(0, k); t->SetSolution(1, om); t->SetSolution_Old(0, k); t->SetSolution_Old(1, om);
      t->SetmuT(rho_ * k / om);
    }
    double R = lib.ComputeRgas(Y), rho = P / (R * T), h = lib.ComputeEnthalpy(T, Y);
    U[0] = rho; U[1] = rho * u; U[2] = rho * v; U[3] = rho * (h + 0.5 * (u * u + v * v) + k) - P;
    for (unsigned short s = 0; s < Ns; ++s) U[4 + s] = rho * Y[s];
    for (unsigned short q = 0; q < c.nVar; ++q) { c.flow->node[p]->SetSolution(q, U[q]); c.flow->node[p]->SetSolution_Old(q, U[q]); }
  }
#ifdef RX_WITH_GPU
  if (CReactiveNSSolverGPU* g = dynamic_cast<CReactiveNSSolverGPU*>(c.flow)) g->MarkHostSolutionDirty();
  if (c.rans) if (CTurbSSTSolverGPU* t = dynamic_cast<CTurbSSTSolverGPU*>(c.sol[TURB_SOL])) t->MarkHostSolutionDirty();
#endif
}

// ---- edge cases of the cons->prim path that a smooth state never reaches ("What's weak" #2).  Applied to the
// host state right before iteration `it` >= 1 (env RX_POKE_ITER / RX_POKE_MODE), on disjoint point sets chosen by global index:
//   'a'  secant start value 201 K: depending on composition and temperature the secant converges late, leaves the spline table
//        (std::out_of_range -> 10000-step bisection, variable_direct_reactive.cpp:656-688) or fails to converge in 7 steps
//        (32-step bisection, :698-724)
//   'b'  secant start value Tmax - 0.4: T_old = T + 1 is outside the table at the first evaluation (exception at iteration 0)
//   'c'  species densities scaled by 0.8: |sum Y - 1| > 0.1 -> non-physical -> Solution_Old is restored (:296-301)
//   'd'  start value 1.2 x the stored temperature: with CLIPPING_TEMPRATURE= YES the solved T is clipped to 0.95 x that (:728-739)
static void poke_state(Ctx& c, const std::string& mode) {
  const bool A = mode.find('a') != std::string::npos, B = mode.find('b') != std::string::npos,
             C = mode.find('c') != std::string::npos, D = mode.find('d') != std::string::npos;
  unsigned long n[4] = {0, 0, 0, 0};
  for (unsigned long p = 0; p < c.nPoint; ++p) {
    const unsigned long gi = c.geo->node[p]->GetGlobalIndex();
    CVariable* nd = c.flow->node[p];
    if (A && gi % 7 == 3) { nd->GetPrimitive()[0] = 201.0; ++n[0]; }
    else if (B && gi % 7 == 5) { nd->GetPrimitive()[0] = c.cfg->GetTemperatureMax() - 0.4; ++n[1]; }
    else if (C && gi % 11 == 2) { for (unsigned short s = 0; s < c.nSpecies; ++s) nd->SetSolution(4 + s, 0.8 * nd->GetSolution(4 + s)); ++n[2]; }
    else if (D && gi % 13 == 4) { nd->GetPrimitive()[0] = 1.2 * nd->GetPrimitive()[0]; ++n[3]; }
  }
  printf("HARNESS poke '%s': %lu / %lu / %lu / %lu points (a/b/c/d)\n", mode.c_str(), n[0], n[1], n[2], n[3]);
#ifdef RX_WITH_GPU
  if (CReactiveNSSolverGPU* g = dynamic_cast<CReactiveNSSolverGPU*>(c.flow)) g->MarkHostSolutionDirty();
#endif
}
// dual time stepping: the two stored time levels U^n, U^(n-1) (CVariable::Solution_time_n / _time_n1)
static void dump_time_levels(Ctx& c, const std::string& base) {
  std::vector<double> A(c.nPoint * c.nVar), B(c.nPoint * c.nVar);
  for (unsigned long p = 0; p < c.nPoint; ++p)
    for (unsigned short v = 0; v < c.nVar; ++v) { A[p * c.nVar + v] = c.flow->node[p]->GetSolution_time_n()[v]; B[p * c.nVar + v] = c.flow->node[p]->GetSolution_time_n1()[v]; }
  dump_f64(base + "Un", A, {c.nPoint, c.nVar});
  dump_f64(base + "Un1", B, {c.nPoint, c.nVar});
}
static void dump_solution_old(Ctx& c, const std::string& name) {
  std::vector<double> U(c.nPoint * c.nVar);
  for (unsigned long p = 0; p < c.nPoint; ++p)
    for (unsigned short v = 0; v < c.nVar; ++v) U[p * c.nVar + v] = c.flow->node[p]->GetSolution_Old()[v];
  dump_f64(name, U, {c.nPoint, c.nVar});
}
// long runs (env RX_CKPT="n1,n2,..."): after n completed iterations dump the conserved state, the temperature and (RANS) k, omega, muT
static void dump_checkpoint(Ctx& c, unsigned long nDone) {
  char b[64]; snprintf(b, sizeof b, "ckpt%06lu_", nDone);
  dump_solution(c, std::string(b) + "U");
  std::vector<double> T(c.nPoint);
  for (unsigned long p = 0; p < c.nPoint; ++p) T[p] = c.flow->node[p]->GetPrimitive(0);
  dump_f64(std::string(b) + "T", T, {c.nPoint});
  if (c.rans) {
    std::vector<double> K(c.nPoint * 3);
    for (unsigned long p = 0; p < c.nPoint; ++p) {
      CVariable* t = c.sol[TURB_SOL]->node[p];
      K[3 * p] = t->GetSolution(0); K[3 * p + 1] = t->GetSolution(1); K[3 * p + 2] = t->GetmuT();
    }
    dump_f64(std::string(b) + "kw", K, {c.nPoint, 3});
  }
}

int main(int argc, char** argv) {
  if (argc < 4) { fprintf(stderr, "usage: ref_harness <cfg> <nIter> <outdir> [dump_from_iter] [perturb_seed]\n"); return 1; }
  char cfgname[MAX_STRING_SIZE];
  strcpy(cfgname, argv[1]);
  unsigned long nIter = atol(argv[2]);
  g_out = argv[3];
  unsigned long dump_from = argc > 4 ? atol(argv[4]) : 0;
  mkdir(g_out.c_str(), 0755);

  SU2_Comm comm(0);
  CConfig* cfg0 = new CConfig(cfgname, SU2_CFD);
  unsigned short nZone = CConfig::GetnZone(cfg0->GetMesh_FileName(), cfg0->GetMesh_FileFormat(), cfg0);
  unsigned short nDim = CConfig::GetnDim(cfg0->GetMesh_FileName(), cfg0->GetMesh_FileFormat());
  delete cfg0;
  CFluidDriver* drv = new CFluidDriver(cfgname, nZone, nDim, comm);

  Ctx c;
  c.drv = drv;
  c.geo = drv->geometry_container[ZONE_0][MESH_0];
  c.cfg = drv->config_container[ZONE_0];
  c.sol = drv->solver_container[ZONE_0][MESH_0];
#ifdef RX_WITH_GPU
  if (getenv("SU2_RX_BACKEND") && !strcmp(getenv("SU2_RX_BACKEND"), "cuda")) {
    // the factory switch of driver_structure.cpp:795-801, applied after the driver built its containers
    c.sol[FLOW_SOL] = new CReactiveNSSolverGPU(c.geo, c.cfg, MESH_0);
    printf("HARNESS flow solver: CReactiveNSSolverGPU (librx_cuda)\n");
    // the factory switch of driver_structure.cpp:817: the SST solver's iteration on the device as well (SU2_RX_TURB=device)
    if (getenv("SU2_RX_TURB") && !strcmp(getenv("SU2_RX_TURB"), "device") && c.cfg->GetKind_Solver() == REACTIVE_RANS) {
      c.sol[TURB_SOL] = new CTurbSSTSolverGPU(c.geo, c.cfg, MESH_0);
      // what the driver does right after constructing the SST solver (driver_structure.cpp:818-819)
      c.sol[FLOW_SOL]->Preprocessing(c.geo, c.sol, c.cfg, MESH_0, NO_RK_ITER, RUNTIME_REACTIVE_SYS, false);
      c.sol[TURB_SOL]->Postprocessing(c.geo, c.sol, c.cfg, MESH_0);
      printf("HARNESS turbulence solver: CTurbSSTSolverGPU (rx_turb_iterate)\n");
    }
  }
#endif
  c.flow = dynamic_cast<CReactiveEulerSolver*>(c.sol[FLOW_SOL]);
  if (!c.flow) { fprintf(stderr, "not a reactive solver\n"); return 4; }
  c.num = drv->numerics_container[ZONE_0][MESH_0][FLOW_SOL];
  c.integ = drv->integration_container[ZONE_0][FLOW_SOL];
  c.nPoint = c.geo->GetnPoint(); c.nPointDomain = c.geo->GetnPointDomain();
  c.nDim = nDim; c.nSpecies = c.flow->nSpecies; c.nVar = c.flow->nVar; c.nPrimVar = c.flow->nPrimVar;
  c.nPrimVarGrad = c.flow->nPrimVarGrad;
  unsigned short kind = c.cfg->GetKind_Solver();
  c.rans = (kind == REACTIVE_RANS);
  c.viscous = (kind == REACTIVE_RANS || kind == REACTIVE_NAVIER_STOKES);

  std::ostringstream js;
  js << "{\n\"nDim\": " << nDim << ", \"nVar\": " << c.nVar << ", \"nPrimVar\": " << c.nPrimVar << ", \"nPrimVarGrad\": " << c.nPrimVarGrad
     << ", \"nIter\": " << nIter << ",\n";
  dump_geometry(c, js);
  dump_chem(c, js);
  dump_config(c, js);
  js << "}\n";
  { FILE* f = fopen((g_out + "/meta.json").c_str(), "w"); fputs(js.str().c_str(), f); fclose(f); }

  if (getenv("RX_PERTURB_TAMP")) g_tamp = atof(getenv("RX_PERTURB_TAMP"));
  if (getenv("RX_PERTURB_NOISE")) g_noise = atof(getenv("RX_PERTURB_NOISE"));
  if (argc > 5) perturb_state(c, atol(argv[5]));
  const long pokeIter = getenv("RX_POKE_ITER") ? atol(getenv("RX_POKE_ITER")) : -1;
  const std::string pokeMode = getenv("RX_POKE_MODE") ? getenv("RX_POKE_MODE") : "";
  std::vector<unsigned long> ckpt;
  if (getenv("RX_CKPT")) { std::stringstream ss(getenv("RX_CKPT")); std::string tok; while (std::getline(ss, tok, ',')) ckpt.push_back(atol(tok.c_str())); }
  const bool quiet = getenv("RX_QUIET") != nullptr;

  for (unsigned long n : ckpt) if (n == 0) dump_checkpoint(c, 0);   // the state the run starts from

  // dual time stepping: every harness iteration is one pseudo-time iteration; after every RX_DUAL_INNER of them (default 2) the physical
  // time step is closed as CFluidIteration::Update does (iteration_structure.cpp:623: CIntegration::SetDualTime_Solver shifts the time levels)
  const bool dualTime = c.cfg->GetUnsteady_Simulation() == DT_STEPPING_1ST || c.cfg->GetUnsteady_Simulation() == DT_STEPPING_2ND;
  const unsigned long dualInner = getenv("RX_DUAL_INNER") ? atol(getenv("RX_DUAL_INNER")) : 2;
  std::vector<double> hist, histAll;
  for (unsigned long it = 0; it < nIter; ++it) {
    bool dmp = it >= dump_from;
    struct timespec ts0, ts1; clock_gettime(CLOCK_MONOTONIC, &ts0);
    drv->PreprocessExtIter(it);
    c.cfg->SetGlobalParam(kind, RUNTIME_REACTIVE_SYS, it);
    if ((long)it == pokeIter) poke_state(c, pokeMode);
    if (dmp) {
      dump_solution(c, tag(it, 0, "U0"));
      if (pokeIter >= 0) dump_solution_old(c, tag(it, 0, "Uold"));
      if (dualTime) dump_time_levels(c, tag(it, 0, ""));
      dump_primitive(c, tag(it, 0, "prev_"));   // V[T] here is the secant start value (variable_direct_reactive.cpp:622)
      dump_turb(c, tag(it, 0, "turb"));
    }
    unsigned short nRK = (c.cfg->GetKind_TimeIntScheme() == RUNGE_KUTTA_EXPLICIT) ? c.cfg->GetnRKStep() : 1;
    for (unsigned short rk = 0; rk < nRK; ++rk) {
      c.flow->Preprocessing(c.geo, c.sol, c.cfg, MESH_0, rk, RUNTIME_REACTIVE_SYS, false);
      if (dmp) dump_primitive(c, tag(it, rk, "pre_"));
      if (rk == 0) {
        c.flow->Set_OldSolution(c.geo);
        c.flow->SetTime_Step(c.geo, c.sol, c.cfg, MESH_0, dualTime ? it % dualInner : it);   // dual time: the pseudo-time iteration index of the physical step (CFluidIteration::Iterate: IntIter)
        if (dmp) dump_dt(c, tag(it, rk, "dt"));
      }
      c.flow->Upwind_Residual(c.geo, c.sol, c.num[CONV_TERM], c.cfg, MESH_0);
      if (dmp) dump_residual(c, tag(it, rk, "res_conv"));
      c.flow->Viscous_Residual(c.geo, c.sol, c.num[VISC_TERM], c.cfg, MESH_0, rk);
      if (dmp) dump_residual(c, tag(it, rk, "res_visc"));
      c.flow->Source_Residual(c.geo, c.sol, c.num[SOURCE_FIRST_TERM], c.num[SOURCE_SECOND_TERM], c.cfg, MESH_0);
      if (dmp) dump_residual(c, tag(it, rk, "res_src"));
      if (dualTime) {   // CIntegration::Space_Integration :86-87: the dual time source term comes between the source term and the boundary conditions
        c.flow->SetResidual_DualTime(c.geo, c.sol, c.cfg, rk, MESH_0, RUNTIME_REACTIVE_SYS);
        if (dmp) dump_residual(c, tag(it, rk, "res_dual"));
      }
      apply_bcs(c);
      if (dmp) dump_residual(c, tag(it, rk, "res_bc"));
      c.integ->Time_Integration(c.geo, c.sol, c.cfg, rk, RUNTIME_REACTIVE_SYS, it);
      if (dmp) {
        dump_solution(c, tag(it, rk, "U1"));
        std::vector<double> rms(c.nVar);
        for (unsigned short v = 0; v < c.nVar; ++v) rms[v] = c.flow->GetRes_RMS(v);
        dump_f64(tag(it, rk, "rms"), rms, {c.nVar});
      }
      c.flow->Postprocessing(c.geo, c.sol, c.cfg, MESH_0);
    }
    c.flow->Preprocessing(c.geo, c.sol, c.cfg, MESH_0, NO_RK_ITER, RUNTIME_REACTIVE_SYS, true);
    if (dmp) dump_primitive(c, tag(it, 0, "out_"));
    clock_gettime(CLOCK_MONOTONIC, &ts1);   // flow iteration only (the hot path); the SST step below is outside it
    if (!quiet) printf("HARNESS_TIME it %lu flow_seconds %.6f\n", it, (ts1.tv_sec - ts0.tv_sec) + 1e-9 * (ts1.tv_nsec - ts0.tv_nsec));
    if (c.rans) {
      // CSingleGridIntegration::SingleGrid_Iteration (integration_time.cpp:777-834) for the SST solver, stage by stage so that the
      // residual before the solve, the solution of the linear system and the updated (k, omega) can be dumped (the convergence
      // monitor at its end only sets flags)
      c.cfg->SetGlobalParam(REACTIVE_RANS, RUNTIME_TURB_SYS, it);
      CSolver* ts = c.sol[TURB_SOL];
      CIntegration* ti = drv->integration_container[ZONE_0][TURB_SOL];
      if (dmp) {   // what the step reads from the previous Postprocessing besides (k, omega): F1, F2, cross diffusion
        std::vector<double> B(c.nPoint * 3);
        for (unsigned long p = 0; p < c.nPoint; ++p) { B[3 * p] = ts->node[p]->GetF1blending(); B[3 * p + 1] = ts->node[p]->GetF2blending(); B[3 * p + 2] = ts->node[p]->GetCrossDiff(); }
        dump_f64(tag(it, 0, "tblend"), B, {c.nPoint, 3});
        if (dualTime) {   // the turbulence solver's own time levels (read by CTurbSolver::SetResidual_DualTime, solver_direct_turbulent.cpp:762-764)
          std::vector<double> A1(c.nPoint * 2), A2(c.nPoint * 2);
          for (unsigned long p = 0; p < c.nPoint; ++p)
            for (int v = 0; v < 2; ++v) { A1[2 * p + v] = ts->node[p]->GetSolution_time_n()[v]; A2[2 * p + v] = ts->node[p]->GetSolution_time_n1()[v]; }
          dump_f64(tag(it, 0, "tUn"), A1, {c.nPoint, 2});
          dump_f64(tag(it, 0, "tUn1"), A2, {c.nPoint, 2});
        }
      }
      ts->Preprocessing(c.geo, c.sol, c.cfg, MESH_0, NO_RK_ITER, RUNTIME_TURB_SYS, false);
      ts->Set_OldSolution(c.geo);
      ts->SetTime_Step(c.geo, c.sol, c.cfg, MESH_0, dualTime ? it % dualInner : it);
      ti->Space_Integration(c.geo, c.sol, drv->numerics_container[ZONE_0][MESH_0][TURB_SOL], c.cfg, MESH_0, NO_RK_ITER, RUNTIME_TURB_SYS);
      if (dmp) {
        std::vector<double> R(c.nPoint * 2);
        for (unsigned long p = 0; p < c.nPoint; ++p) { R[2 * p] = ts->LinSysRes[2 * p]; R[2 * p + 1] = ts->LinSysRes[2 * p + 1]; }
        dump_f64(tag(it, 0, "tres"), R, {c.nPoint, 2});   // LinSysRes after Upwind, Viscous, Source and the boundary conditions
      }
      ti->Time_Integration(c.geo, c.sol, c.cfg, NO_RK_ITER, RUNTIME_TURB_SYS, it);
      if (dmp) {
        std::vector<double> X(c.nPoint * 2), S(c.nPoint * 2), Q(2);
        for (unsigned long p = 0; p < c.nPoint; ++p) {
          X[2 * p] = ts->LinSysSol[2 * p]; X[2 * p + 1] = ts->LinSysSol[2 * p + 1];
          S[2 * p] = ts->node[p]->GetSolution(0); S[2 * p + 1] = ts->node[p]->GetSolution(1);
        }
        dump_f64(tag(it, 0, "tdx"), X, {c.nPoint, 2});    // solution of the implicit system (FGMRES + LU-SGS)
        dump_f64(tag(it, 0, "tsol"), S, {c.nPoint, 2});   // (k, omega) after the clipped conservative update
        Q[0] = ts->GetRes_RMS(0); Q[1] = ts->GetRes_RMS(1);
        dump_f64(tag(it, 0, "trms"), Q, {2});
      }
      ts->Postprocessing(c.geo, c.sol, c.cfg, MESH_0);
      if (dmp) dump_turb_post(c, tag(it, 0, "tpost"));   // flow state at this point: the out_ dump above
    }
    if (dualTime && (it + 1) % dualInner == 0) {
      // CFluidIteration::Update (iteration_structure.cpp:623-640): the flow solver's levels, then the turbulence solver's
      c.integ->SetDualTime_Solver(c.geo, c.flow, c.cfg, MESH_0);
      if (c.rans) drv->integration_container[ZONE_0][TURB_SOL]->SetDualTime_Solver(c.geo, c.sol[TURB_SOL], c.cfg, MESH_0);
#ifdef RX_WITH_GPU
      // the optional device-side shift; without it the next SetTime_Step uploads the levels the line above shifted in node[]
      if (getenv("RX_DUAL_SHIFT_DEVICE") && atoi(getenv("RX_DUAL_SHIFT_DEVICE")))
        if (CReactiveNSSolverGPU* g = dynamic_cast<CReactiveNSSolverGPU*>(c.flow)) g->ShiftTimeLevels();
#endif
    }
    double r0 = log10(c.flow->GetRes_RMS(0)), r3 = log10(c.flow->GetRes_RMS(nDim + 1));
    if (!quiet || it % 100 == 0) printf("HARNESS it %lu  Res[Rho] %.6f  Res[RhoE] %.6f\n", it, r0, r3);
    hist.push_back(r0); hist.push_back(r3);
    for (unsigned short v = 0; v < c.nVar; ++v) histAll.push_back(c.flow->GetRes_RMS(v));
    for (unsigned long n : ckpt) if (n == it + 1) dump_checkpoint(c, n);
  }
  if (getenv("RX_RESTART_OUT")) {   // the reference's own writer: COutput::SetResult_Files -> MergeCoordinates, MergeSolution, SetRestart (output_structure.cpp:7397-7470, :3858)
    COutput* out = drv->output;
    out->MergeCoordinates(c.cfg, c.geo);
    out->MergeSolution(c.cfg, c.geo, c.sol, ZONE_0);
    out->SetRestart(c.cfg, c.geo, c.sol, ZONE_0);
    printf("HARNESS wrote %s\n", c.cfg->GetRestart_FlowFileName().c_str());
    if (getenv("RX_TECPLOT_OUT")) {   // the volume solution file as well: MergeConnectivity + SetTecplotASCII (output_tecplot.cpp:46-420), "flow.dat"
      out->MergeConnectivity(c.cfg, c.geo, ZONE_0);
      out->SetTecplotASCII(c.cfg, c.geo, c.sol, ZONE_0, 1, false);
      printf("HARNESS wrote %s.dat\n", c.cfg->GetFlow_FileName().c_str());
    }
#ifdef RX_WITH_GPU
    if (CReactiveNSSolverGPU* g = dynamic_cast<CReactiveNSSolverGPU*>(c.flow)) {   // the same file straight from the device state
      g->WriteRestart(c.geo, c.sol, c.cfg, "restart_flow_gpu.dat");
      printf("HARNESS wrote restart_flow_gpu.dat (rx_restart_write)\n");
      if (getenv("RX_TECPLOT_OUT")) { g->WriteTecplot(c.geo, c.sol, c.cfg, "flow_gpu.dat"); printf("HARNESS wrote flow_gpu.dat (rx_tecplot_write)\n"); }
    }
#endif
  }
  {   // peak resident set of this process (bench.py reports it next to the CPU timing: the working set is far beyond the caches)
    FILE* f = fopen("/proc/self/status", "r");
    char line[256];
    while (f && fgets(line, sizeof line, f)) if (!strncmp(line, "VmHWM:", 6)) printf("HARNESS_RSS_KB %ld\n", atol(line + 6));
    if (f) fclose(f);
  }
  dump_f64("history", hist, {(size_t)nIter, 2});
  dump_f64("history_rms", histAll, {(size_t)nIter, (size_t)c.nVar});   // CSolver::GetRes_RMS of every variable after every iteration
  return 0;
}
