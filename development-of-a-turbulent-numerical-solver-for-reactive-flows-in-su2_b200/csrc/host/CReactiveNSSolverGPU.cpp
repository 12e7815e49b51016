// CReactiveNSSolverGPU — see the header.  Compiled with the reference's include paths (oracle/build_ref.sh does so, with
// -fno-access-control for the private tables of Framework::ReactingModelLibrary; upstream would add one `friend` line).
#include "CReactiveNSSolverGPU.hpp"

#include "reacting_model_library.hpp"

#include <sstream>
#include <cstdlib>
#include <stdexcept>
#include <string>

namespace {
int env_int(const char* name, int dflt) {
  const char* s = std::getenv(name);
  return s ? std::atoi(s) : dflt;
}
}  // namespace

void CReactiveNSSolverGPU::Check(int rc, const char* where) {
  if (rc == RX_OK) return;
  // the reference reports trouble on this path by throwing (solver_direct_reactive.cpp:2772,2861,5384)
  throw std::runtime_error(std::string(where) + ": " + rx_last_error(ctx) + " (librx_cuda, status " + std::to_string(rc) + ")");
}

CReactiveNSSolverGPU::CReactiveNSSolverGPU(CGeometry* geometry, CConfig* config, unsigned short iMesh)
    : CReactiveNSSolver(geometry, config, iMesh) {
  using namespace Framework;
  mirror = env_int("SU2_RX_MIRROR", 1);
  turbPost = env_int("SU2_RX_TURB_POST", 0) != 0;
  const unsigned short kind = config->GetKind_Solver();
  ransMode = (kind == REACTIVE_RANS);
  viscousMode = (kind == REACTIVE_RANS || kind == REACTIVE_NAVIER_STOKES);
  if (implicit) throw std::runtime_error("CReactiveNSSolverGPU: EULER_IMPLICIT is not on the CUDA path (explicit Euler / RK only)");
  if (limiter && config->GetKind_SlopeLimit_Flow() != VENKATAKRISHNAN && config->GetKind_SlopeLimit_Flow() != BARTH_JESPERSEN)
    throw std::runtime_error("CReactiveNSSolverGPU: only the VENKATAKRISHNAN and BARTH_JESPERSEN limiters are on the CUDA path");
  if (grid_movement) throw std::runtime_error("CReactiveNSSolverGPU: moving grids are not on the CUDA path");
  if (nDim != 2) throw std::runtime_error("CReactiveNSSolverGPU: two space dimensions only");
  // one grid level (MGLEVEL= 0 in every shipped cfg): the coarse levels' restriction / prolongation (integration_time.cpp:183-330) work on node[] only
  if (config->GetnMGLevels() > 0 || iMesh != MESH_0) throw std::runtime_error("CReactiveNSSolverGPU: multigrid (MGLEVEL > 0) is not on the CUDA path");
  // the reference computes gradients for GREEN_GAUSS and WEIGHTED_LEAST_SQUARES only (Preprocessing :4713-4716; any other kind leaves them unset)
  if (config->GetKind_Gradient_Method() != GREEN_GAUSS && config->GetKind_Gradient_Method() != WEIGHTED_LEAST_SQUARES)
    throw std::runtime_error("CReactiveNSSolverGPU: NUM_METHOD_GRAD must be GREEN_GAUSS or WEIGHTED_LEAST_SQUARES");
  // every *_Ref is 1 on the CUDA path, as in all shipped cfgs (REF_DIMENSIONALIZATION= DIMENSIONAL, SI units: solver_direct_reactive.cpp:846-887)
  if (config->GetRef_NonDim() != DIMENSIONAL) throw std::runtime_error("CReactiveNSSolverGPU: REF_DIMENSIONALIZATION other than DIMENSIONAL is not on the CUDA path");
  // VISCOUS_LIMITER_FLOW: CAvgGradReactive_Flow multiplies the gradients of T, u, v by the limiters (numerics_direct_reactive.cpp:1482-1505), which
  // Viscous_Residual hands over only under SPATIAL_ORDER_FLOW= 2ND_ORDER_LIMITER (:5326); without it the reference dereferences an unset array
  if (config->GetViscous_Limiter_Flow() && config->GetSpatialOrder_Flow() != SECOND_ORDER_LIMITER)
    throw std::runtime_error("CReactiveNSSolverGPU: VISCOUS_LIMITER_FLOW needs SPATIAL_ORDER_FLOW= 2ND_ORDER_LIMITER (the reference reads an unset limiter array otherwise)");
  if (config->GetSystemMeasurements() == US) throw std::runtime_error("CReactiveNSSolverGPU: SYSTEM_MEASUREMENTS= US is not on the CUDA path");
  // A partitioned geometry needs the device halo exchange behind Set_MPI_Solution / _Primitive_Gradient / _Primitive_Limiter
  // (solver_reactive.hpp:162,176,183), which this class does not override: the base class would exchange host node[] only and the
  // device halos would go stale without any error.  Refuse instead (the Python host, parallel.py, drives partitioned runs).
#ifndef HAVE_MPI
  if (nPoint != nPointDomain) throw std::runtime_error("CReactiveNSSolverGPU: partitioned geometry (halo points present) in a build without MPI");
  for (unsigned short m = 0; m < config->GetnMarker_All(); ++m)
    if (config->GetMarker_All_KindBC(m) == SEND_RECEIVE)
      throw std::runtime_error("CReactiveNSSolverGPU: SEND_RECEIVE markers present in a build without MPI");
#endif

  // ---- geometry -> rx_mesh (flat arrays; dual_grid_structure.hpp:140,747,884)
  const unsigned long nEdge = geometry->GetnEdge();
  std::vector<int64_t> en(2 * nEdge), nbPtr(nPoint + 1, 0), nbPt;
  std::vector<double> enrm(2 * nEdge), xy(2 * nPoint), vol(nPoint);
  for (unsigned long e = 0; e < nEdge; ++e) {
    en[2 * e] = geometry->edge[e]->GetNode(0); en[2 * e + 1] = geometry->edge[e]->GetNode(1);
    enrm[2 * e] = geometry->edge[e]->GetNormal()[0]; enrm[2 * e + 1] = geometry->edge[e]->GetNormal()[1];
  }
  for (unsigned long p = 0; p < nPoint; ++p) {
    xy[2 * p] = geometry->node[p]->GetCoord(0); xy[2 * p + 1] = geometry->node[p]->GetCoord(1);
    vol[p] = geometry->node[p]->GetVolume();
    for (unsigned short k = 0; k < geometry->node[p]->GetnPoint(); ++k) nbPt.push_back(geometry->node[p]->GetPoint(k));
    nbPtr[p + 1] = (int64_t)nbPt.size();
  }
  const unsigned short nMk = config->GetnMarker_All();
  std::vector<rx_marker> mk(nMk);
  std::vector<std::vector<int64_t>> vn(nMk), vnn(nMk);
  std::vector<std::vector<double>> vnrm(nMk), vys(nMk);
  for (unsigned short m = 0; m < nMk; ++m) {
    const unsigned long nv = geometry->nVertex[m];
    vn[m].resize(nv); vnn[m].resize(nv); vnrm[m].resize(2 * nv); vys[m].assign(nSpecies, 0.0);
    for (unsigned long v = 0; v < nv; ++v) {
      vn[m][v] = geometry->vertex[m][v]->GetNode();
      vnn[m][v] = geometry->vertex[m][v]->GetNormal_Neighbor();
      vnrm[m][2 * v] = geometry->vertex[m][v]->GetNormal()[0]; vnrm[m][2 * v + 1] = geometry->vertex[m][v]->GetNormal()[1];
    }
    rx_marker& r = mk[m];
    r = rx_marker{};
    r.nVertex = (int)nv; r.node = vn[m].data(); r.normalNeighbor = vnn[m].data(); r.normal = vnrm[m].data(); r.Ys = vys[m].data();
    const std::string tag = config->GetMarker_All_TagBound(m);
    switch (config->GetMarker_All_KindBC(m)) {
      case INLET_FLOW:
        switch (config->GetKind_Inlet()) {   // INLET_TYPE: the three kinds BC_Inlet implements (:3280-3500)
          case TEMPERATURE_IMPOSE: r.kind = RX_BC_INLET_TEMPERATURE_IMPOSE; break;
          case TOTAL_CONDITIONS: r.kind = RX_BC_INLET_TOTAL_CONDITIONS; break;
          case MASS_FLOW: r.kind = RX_BC_INLET_MASS_FLOW; break;
          default: throw std::runtime_error("CReactiveNSSolverGPU: INLET_TYPE not implemented by the reference's BC_Inlet either");
        }
        r.T = config->GetInlet_Ttotal(tag); r.Vmag = config->GetInlet_Ptotal(tag);
        r.dir[0] = config->GetInlet_FlowDir(tag)[0]; r.dir[1] = config->GetInlet_FlowDir(tag)[1];
        for (unsigned short s = 0; s < nSpecies; ++s) vys[m][s] = config->GetInlet_MassFrac(tag)[s];
        break;
      case SUPERSONIC_INLET:
        if (ransMode) throw std::runtime_error("CReactiveNSSolverGPU: BC_Supersonic_Inlet leaves the turbulence inputs of the viscous numerics unset in the reference; RANS is not supported with it");
        r.kind = RX_BC_SUPERSONIC_INLET;
        r.T = config->GetInlet_Temperature(tag); r.P = config->GetInlet_Pressure(tag);
        r.dir[0] = config->GetInlet_Velocity(tag)[0]; r.dir[1] = config->GetInlet_Velocity(tag)[1];
        for (unsigned short s = 0; s < nSpecies; ++s) vys[m][s] = config->GetInlet_MassFrac(tag)[s];
        break;
      case SUPERSONIC_OUTLET:
        if (ransMode) throw std::runtime_error("CReactiveNSSolverGPU: BC_Supersonic_Outlet leaves the turbulence inputs of the viscous numerics unset in the reference; RANS is not supported with it");
        r.kind = RX_BC_SUPERSONIC_OUTLET; break;
      case OUTLET_FLOW: r.kind = RX_BC_OUTLET; r.P = config->GetOutlet_Pressure(tag); break;
      case ISOTHERMAL: r.kind = RX_BC_ISOTHERMAL; r.Twall = config->GetIsothermal_Temperature(tag); break;
      case HEAT_FLUX: r.kind = RX_BC_HEATFLUX; r.heatFlux = config->GetWall_HeatFlux(tag); break;
      case EULER_WALL: r.kind = RX_BC_EULER_WALL; break;
      case SEND_RECEIVE: case INTERNAL_BOUNDARY: r.kind = RX_BC_NONE; break;
      default: throw std::runtime_error("CReactiveNSSolverGPU: boundary condition of marker " + tag + " is not on the CUDA path");
    }
  }
  rx_mesh mesh{};
  mesh.nPoint = (int64_t)nPoint; mesh.nPointDomain = (int64_t)nPointDomain; mesh.nEdge = (int64_t)nEdge;
  mesh.edgeNodes = en.data(); mesh.edgeNormal = enrm.data(); mesh.coord = xy.data(); mesh.volume = vol.data();
  mesh.nbPtr = nbPtr.data(); mesh.nbPoint = nbPt.data(); mesh.nMarker = nMk; mesh.markers = mk.data();

  // ---- chemistry tables of the library object the reference solver already set up (reacting_model_library.hpp:655-766)
  ReactingModelLibrary& lib = *std::dynamic_pointer_cast<ReactingModelLibrary>(library);
  const int Ns = lib.nSpecies, Nr = lib.nReactions;
  typedef decltype(lib.Mu_Spline) SplTab;
  const SplTab* tabs[5] = {&lib.Mu_Spline, &lib.Kappa_Spline, &lib.Cp_Spline, &lib.Enth_Spline, &lib.Entr_Spline};
  const size_t nK = std::get<0>(lib.Mu_Spline[0]).size();
  std::vector<double> X(std::get<0>(lib.Mu_Spline[0])), Y(5 * Ns * nK), Y2(5 * Ns * nK);
  for (int t = 0; t < 5; ++t)
    for (int s = 0; s < Ns; ++s) {
      const auto& tp = (*tabs[t])[s];
      for (size_t k = 0; k < nK; ++k) {
        Y[(t * Ns + s) * nK + k] = std::get<1>(tp)[k];
        Y2[(t * Ns + s) * nK + k] = std::get<2>(tp).data()[k];   // the last second derivative lives past size() (spline.cpp:54)
      }
    }
  std::vector<double> rev(Nr ? Nr : 1, 0.0), avb(rev), Ab(rev), Bb(rev), Tb(rev);
  std::vector<double> nuR(Nr * Ns + 1), nuP(nuR), eR(nuR), eP(nuR);
  for (int r = 0; r < Nr; ++r) {
    rev[r] = lib.Reversible_Reactions[r];
    if (lib.Available_Backward_Rate.count(r)) { avb[r] = 1; Ab[r] = lib.As_back[r]; Bb[r] = lib.Betas_back[r]; Tb[r] = lib.Temps_Activation_back[r]; }
    for (int s = 0; s < Ns; ++s) {
      nuR[r * Ns + s] = lib.Stoich_Coeffs_Reactants(s, r); nuP[r * Ns + s] = lib.Stoich_Coeffs_Products(s, r);
      eR[r * Ns + s] = lib.Stoich_Coeffs_Reactants_Exp(r, s); eP[r * Ns + s] = lib.Stoich_Coeffs_Products_Exp(r, s);
    }
  }
  rx_chem chem{};
  chem.Ns = Ns; chem.Nr = Nr; chem.nKnots = (int)nK;
  chem.mMasses = lib.mMasses.data(); chem.diffVol = lib.Diff_Volumes.data(); chem.x = X.data(); chem.y = Y.data(); chem.y2 = Y2.data();
  chem.As = lib.As.data(); chem.Betas = lib.Betas.data(); chem.Ta = lib.Temps_Activation.data();
  chem.reversible = rev.data(); chem.availBack = avb.data(); chem.AsB = Ab.data(); chem.BetasB = Bb.data(); chem.TaB = Tb.data();
  chem.nuR = nuR.data(); chem.nuP = nuP.data(); chem.expR = eR.data(); chem.expP = eP.data();

  // ---- constants the solver / numerics constructors read from CConfig (numerics_direct_reactive.cpp:17-19,395-399,1417,1690-1694)
  rx_opts o{};
  o.viscous = viscousMode; o.rans = ransMode; o.leastSquares = least_squares ? 1 : 0;
  o.mInfty = config->GetMach(); o.cfl = config->GetCFL(iMesh); o.maxDt = config->GetMax_DeltaTime();
  o.Tmin = config->GetTemperatureMin(); o.Tmax = config->GetTemperatureMax(); o.clipT = config->GetClipping_Temp() ? 1 : 0;
  o.ignition = config->GetIgnition() ? 1 : 0; o.ignitionIter = (int64_t)config->GetIgnitionIter(); o.ignitionT = config->GetIgnitionTemperature();
  o.fuelIndex = config->GetFuelIndex(); o.oxidizerIndex = config->GetOxidizerIndex();
  o.prandtlLam = config->GetPrandtl_Lam(); o.prandtlTurb = config->GetPrandtl_Turb(); o.lewisTurb = config->GetLewis_Turb();
  o.cMu = config->Get_Cmu(); o.pasrLb = config->Get_PaSR_LB(); o.tkeInf = ransMode ? GetTke_Inf() : 0.0;
  o.secondOrder = second_order ? 1 : 0; o.limiter = limiter ? (config->GetKind_SlopeLimit_Flow() == BARTH_JESPERSEN ? 2 : 1) : 0;
  o.limCoeff = config->GetLimiterCoeff(); o.refElemLength = config->GetRefElemLength();
  o.timeStepping = config->GetUnsteady_Simulation() == TIME_STEPPING ? 1 : 0; o.unstCfl = config->GetUnst_CFL(); o.deltaUnstTime = config->GetDelta_UnstTime();
  // dual time stepping with explicit pseudo-time iterations (SetResidual_DualTime :2172-2215, the pseudo-time step clamp :5274-5297)
  dualTime = config->GetUnsteady_Simulation() == DT_STEPPING_1ST ? 1 : config->GetUnsteady_Simulation() == DT_STEPPING_2ND ? 2 : 0;
  // (REACTIVE_RANS: the reference's host SST solver keeps its own time levels and reads the flow density of the three levels from node[],
  // where CIntegration::SetDualTime_Solver keeps them current; the device SST solver refuses dual time stepping)
  o.dualTime = dualTime; o.deltaUnstTimeND = config->GetDelta_UnstTimeND();
  o.viscLimiter = (viscousMode && config->GetViscous_Limiter_Flow()) ? 1 : 0;
  o.device = env_int("SU2_RX_DEVICE", 0);

  const int rc = rx_create(&ctx, &mesh, &chem, &o);
  if (rc != RX_OK) throw std::runtime_error(std::string("CReactiveNSSolverGPU: ") + rx_last_error(nullptr));
  bufU.resize((size_t)nPoint * nVar); bufR.resize((size_t)nPoint * nVar);
  bufV.resize((size_t)nPoint * nPrimVar); bufAux.resize((size_t)nPoint * 6);
  bufGrad.resize((size_t)nPoint * nPrimVarGrad * nDim); bufT.resize((size_t)nPoint * 6);
#ifdef RX_COMM_BLOCK
  InitComm(geometry, config);
#endif
}

#ifdef RX_COMM_BLOCK
// One rank per GPU.  The neighbour lists are the vertex lists of the SEND_RECEIVE marker pairs, in the order Set_MPI_Solution walks them
// (solver_direct_reactive.cpp:1541-1560): marker S (SendRecv > 0) goes to rank SendRecv - 1, marker S + 1 receives from |SendRecv| - 1.
// Compiled in MPI builds; this container has no MPI, so oracle/build_ref.sh compiles the block once more against the five single-rank
// stand-ins of RX_MPI_STUB (CReactiveNSSolverGPU.hpp) as a syntax and type check, and tests/test_gpu_nccl.py drives the same C entry points.
void CReactiveNSSolverGPU::InitComm(CGeometry* geometry, CConfig* config) {
  int rank = 0, size = 1;
  MPI_Comm_rank(MPI_COMM_WORLD, &rank); MPI_Comm_size(MPI_COMM_WORLD, &size);
  if (size == 1) return;
  std::vector<int> nb;
  std::vector<int64_t> sp(1, 0), rp(1, 0), spts, rpts;
  for (unsigned short m = 0; m < geometry->GetnMarker(); ++m) {
    if (config->GetMarker_All_KindBC(m) != SEND_RECEIVE || config->GetMarker_All_SendRecv(m) <= 0) continue;
    const unsigned short ms = m, mr = m + 1;
    nb.push_back(config->GetMarker_All_SendRecv(ms) - 1);
    for (unsigned long v = 0; v < geometry->nVertex[ms]; ++v) spts.push_back((int64_t)geometry->vertex[ms][v]->GetNode());
    for (unsigned long v = 0; v < geometry->nVertex[mr]; ++v) rpts.push_back((int64_t)geometry->vertex[mr][v]->GetNode());
    sp.push_back((int64_t)spts.size()); rp.push_back((int64_t)rpts.size());
  }
  char id[128];
  if (rank == MASTER_NODE) Check(rx_comm_unique_id(id), "rx_comm_unique_id");
  MPI_Bcast(id, 128, MPI_CHAR, MASTER_NODE, MPI_COMM_WORLD);
  Check(rx_comm_init(ctx, id, rank, size, (int)nb.size(), nb.data(), sp.data(), spts.data(), rp.data(), rpts.data()), "rx_comm_init");
  commReady = true;
}
#endif

void CReactiveNSSolverGPU::Set_MPI_Solution(CGeometry*, CConfig*) {
  if (commReady) Check(rx_halo_exchange(ctx, 0), "rx_halo_exchange(U)");
}
void CReactiveNSSolverGPU::Set_MPI_Primitive_Gradient(CGeometry*, CConfig*) {
  if (commReady) Check(rx_halo_exchange(ctx, 1), "rx_halo_exchange(gradients)");
}
void CReactiveNSSolverGPU::Set_MPI_Primitive_Limiter(CGeometry*, CConfig*) {
  if (commReady) Check(rx_halo_exchange(ctx, 2), "rx_halo_exchange(limiters)");
}

CReactiveNSSolverGPU::~CReactiveNSSolverGPU() { rx_destroy(ctx); }

// node[] -> device: conserved variables and the previous temperature (start value of the secant solve)
void CReactiveNSSolverGPU::UploadState() {
  std::vector<double> T(nPoint);
  for (unsigned long p = 0; p < nPoint; ++p) {
    for (unsigned short v = 0; v < nVar; ++v) bufU[p * nVar + v] = node[p]->GetSolution(v);
    T[p] = node[p]->GetPrimitive(T_INDEX_PRIM);
  }
  Check(rx_set_solution(ctx, bufU.data()), "rx_set_solution");
  Check(rx_set_temperature(ctx, T.data()), "rx_set_temperature");
  hostDirty = false;
}

// what the viscous closure, the time step and the PaSR source read from CTurbSSTSolver (solver_direct_reactive.cpp:5333-5353, :2829)
void CReactiveNSSolverGPU::UploadTurbulence(CSolver** solver_container) {
  CSolver* ts = solver_container[TURB_SOL];
  for (unsigned long p = 0; p < nPoint; ++p) {
    double* t = &bufT[p * 6];
    t[0] = ts->node[p]->GetmuT(); t[1] = ts->node[p]->GetSolution(0); t[2] = ts->node[p]->GetSolution(1);
    t[3] = ts->node[p]->GetGradient()[0][0]; t[4] = ts->node[p]->GetGradient()[0][1]; t[5] = ts->node[p]->Get_Sigmak();
  }
  Check(rx_set_turbulence(ctx, bufT.data()), "rx_set_turbulence");
}

// The same fields formed on the device (rx_turb_postprocess = CTurbSSTSolver::Postprocessing, solver_direct_turbulent.cpp:2953-3015)
// from the turbulence solution alone.  Called once per iteration, before the first Preprocessing of the flow step: the device
// still holds the primitive state of the previous Output pass, which is the state the host-side Postprocessing saw.
void CReactiveNSSolverGPU::DeviceTurbPostprocessing(CGeometry* geometry, CSolver** solver_container) {
  CSolver* ts = solver_container[TURB_SOL];
  if (!wallDistSet) {
    std::vector<double> dist(nPoint);
    for (unsigned long p = 0; p < nPoint; ++p) dist[p] = geometry->node[p]->GetWall_Distance();
    Check(rx_set_wall_distance(ctx, dist.data()), "rx_set_wall_distance");
    wallDistSet = true;
  }
  std::vector<double> kw((size_t)nPoint * 2), post((size_t)nPoint * 12);
  for (unsigned long p = 0; p < nPoint; ++p) { kw[2 * p] = ts->node[p]->GetSolution(0); kw[2 * p + 1] = ts->node[p]->GetSolution(1); }
  Check(rx_turb_postprocess(ctx, kw.data(), ts->node[0]->Get_Sigmak(), 0), "rx_turb_postprocess");
  // host copy of the six fields, as UploadTurbulence leaves it: MirrorPrimitives hands the eddy viscosity to the flow variables
  Check(rx_get_turbulence(ctx, post.data()), "rx_get_turbulence");
  for (unsigned long p = 0; p < nPoint; ++p) {
    double* t = &bufT[p * 6]; const double* q = &post[p * 12];
    t[0] = q[6]; t[1] = q[0]; t[2] = q[1]; t[3] = q[7]; t[4] = q[8]; t[5] = q[11];
  }
}

void CReactiveNSSolverGPU::MirrorPrimitives() {
  Check(rx_get_primitive(ctx, bufV.data(), bufAux.data(), viscousMode ? bufGrad.data() : nullptr), "rx_get_primitive");
  for (unsigned long p = 0; p < nPoint; ++p) {
    CReactiveNSVariable* n = static_cast<CReactiveNSVariable*>(node[p]);
    su2double* V = n->GetPrimitive();
    for (unsigned short v = 0; v < nPrimVar; ++v) V[v] = bufV[p * nPrimVar + v];
    const double* a = &bufAux[p * 6];
    n->Cp = a[0];
    n->GetdPdU()[nDim + 1] = a[1] - 1.0;
    if (ransMode) n->SetEddyViscosity(bufT[p * 6]);   // SetPrimVar(eddy_visc, ...) stores it in the flow variable (:1193)
    if (viscousMode) {
      n->SetLaminarViscosity(a[2]); n->SetThermalConductivity(a[3]);
      n->GetVorticity()[0] = 0.0; n->GetVorticity()[1] = 0.0; n->GetVorticity()[2] = a[4];
      n->StrainMag = a[5];
      su2double** g = n->GetGradient_Primitive();
      for (unsigned short v = 0; v < nPrimVarGrad; ++v)
        for (unsigned short d = 0; d < nDim; ++d) g[v][d] = bufGrad[(p * nPrimVarGrad + v) * nDim + d];
    }
  }
}

void CReactiveNSSolverGPU::MirrorLimiter() {
  std::vector<double> L((size_t)nPoint * 4);
  Check(rx_get_limiter(ctx, L.data()), "rx_get_limiter");
  for (unsigned long p = 0; p < nPoint; ++p)
    for (unsigned short v = 0; v < 4; ++v) node[p]->SetLimiter_Primitive(v, L[p * 4 + v]);
}

void CReactiveNSSolverGPU::MirrorResidual() {
  Check(rx_get_residual(ctx, bufR.data()), "rx_get_residual");
  for (unsigned long p = 0; p < nPoint; ++p) LinSysRes.SetBlock(p, &bufR[p * nVar]);
}

void CReactiveNSSolverGPU::MirrorTimeStep() {
  std::vector<double> dt((size_t)nPoint * 3);
  Check(rx_get_dt(ctx, dt.data()), "rx_get_dt");
  for (unsigned long p = 0; p < nPointDomain; ++p) {
    node[p]->SetDelta_Time(dt[p * 3]);
    node[p]->SetMax_Lambda_Inv(dt[p * 3 + 1]);
    if (viscousMode) node[p]->SetMax_Lambda_Visc(dt[p * 3 + 2]);
  }
}

void CReactiveNSSolverGPU::Preprocessing(CGeometry* geometry, CSolver** solver_container, CConfig* config, unsigned short iMesh,
                                         unsigned short iRKStep, unsigned short RunTime_EqSystem, bool Output) {
  extIter = config->GetExtIter();
  const bool wasDirty = hostDirty;   // a state set on the host (restart, initial condition) has no device primitives yet
  if (hostDirty) UploadState();
  if (ransMode && turbOnDevice) {
    // CTurbSSTSolverGPU keeps (mu_t, k, omega, grad k, sigma_k) current on the device: nothing to hand over
  } else if (ransMode) {
    const bool onDevice = turbPost && turbPostReady && !wasDirty;
    if (!onDevice) UploadTurbulence(solver_container);
    else if (iRKStep == 0 && !Output) DeviceTurbPostprocessing(geometry, solver_container);
    // else: later RK stages and the Output pass keep the fields of stage 0 (the host copies are the same all iteration long)
  }
  int64_t bad = 0;
  const bool count = (config->GetConsole_Output_Verb() == VERB_HIGH);
  Check(rx_preprocess(ctx, (int64_t)extIter, Output ? 1 : 0, count ? &bad : nullptr), "rx_preprocess");
  if (count && iMesh == MESH_0) config->SetNonphysical_Points((unsigned long)bad);
  if (commReady) {   // the reference exchanges inside its Preprocessing (solver_direct_reactive.cpp:4716-4760)
    if (viscousMode || second_order) Set_MPI_Primitive_Gradient(geometry, config);
    if (limiter && !Output) Set_MPI_Primitive_Limiter(geometry, config);
  }
  if (Output) turbPostReady = true;   // the device now holds the primitive state the next SST Postprocessing refers to
  if (mirror >= 1) MirrorPrimitives();
  if (mirror >= 2 && limiter && !Output) MirrorLimiter();
  if (mirror >= 2 && !Output) MirrorResidual();
}

void CReactiveNSSolverGPU::SetTime_Step(CGeometry* geometry, CSolver** solver_container, CConfig* config, unsigned short iMesh,
                                        unsigned long Iteration) {
  // MultiGrid_Cycle calls the (non-virtual) CSolver::Set_OldSolution immediately before SetTime_Step at RK stage 0
  // (integration_time.cpp:158-163); the device copy of U_old is taken here, at the same point of the sequence.
  Check(rx_set_old_solution(ctx), "rx_set_old_solution");
  Check(rx_set_cfl(ctx, config->GetCFL(iMesh)), "rx_set_cfl");   // read at every call, as the reference does: CFL_ADAPT changes it between iterations
  if (dualTime) {
    // `Iteration` is the pseudo-time iteration index of the running physical time step.  On its iteration 0 the stored time levels are
    // taken from node[] (CIntegration::SetDualTime_Solver, a restart or the initial condition wrote them there; Solution itself is
    // current on the host at every mirror level) unless ShiftTimeLevels() already moved them on the device.
    Check(rx_set_inner_iteration(ctx, (int64_t)Iteration), "rx_set_inner_iteration");
    if (Iteration == 0) {   // a new physical time step begins
      if (!levelsOnDevice) UploadTimeLevels();
      levelsOnDevice = false; levelsUploaded = true;
    } else if (!levelsUploaded) { UploadTimeLevels(); levelsUploaded = true; }
  }
  Check(rx_time_step(ctx), "rx_time_step");
  if (dualTime && Iteration == 0 && config->GetUnst_CFL() != 0.0 && iMesh == MESH_0) {   // :5276-5289
    double dtND = 0.0;
    Check(rx_get_delta_unst_time(ctx, &dtND), "rx_get_delta_unst_time");
    config->SetDelta_UnstTimeND(dtND);
  }
  // the Euler solver's SetTime_Step replaces the CFL by UNST_CFL_NUMBER in a time-accurate run (:2129); the library did the same inside
  if (!viscousMode && config->GetUnsteady_Simulation() == TIME_STEPPING) config->SetCFL(iMesh, config->GetUnst_CFL());
  if (mirror >= 2 || (mirror >= 1 && ransMode)) MirrorTimeStep();   // the SST implicit step uses the flow time step
}

void CReactiveNSSolverGPU::UploadTimeLevels() {
  std::vector<double> A((size_t)nPoint * nVar), B((size_t)nPoint * nVar);
  for (unsigned long p = 0; p < nPoint; ++p)
    for (unsigned short v = 0; v < nVar; ++v) { A[p * nVar + v] = node[p]->GetSolution_time_n()[v]; B[p * nVar + v] = node[p]->GetSolution_time_n1()[v]; }
  Check(rx_set_time_levels(ctx, A.data(), B.data()), "rx_set_time_levels");
}
void CReactiveNSSolverGPU::ShiftTimeLevels() {
  if (!dualTime) return;
  if (!levelsUploaded) UploadTimeLevels();   // nothing on the device yet: the shift starts from what node[] holds
  Check(rx_shift_time_levels(ctx), "rx_shift_time_levels");
  levelsOnDevice = true; levelsUploaded = true;
}
void CReactiveNSSolverGPU::SetResidual_DualTime(CGeometry*, CSolver**, CConfig*, unsigned short, unsigned short, unsigned short) {
  if (!levelsUploaded) { UploadTimeLevels(); levelsUploaded = true; }
  Check(rx_residual_dual_time(ctx), "rx_residual_dual_time");
  if (mirror >= 2) MirrorResidual();
}
void CReactiveNSSolverGPU::Upwind_Residual(CGeometry*, CSolver**, CNumerics*, CConfig*, unsigned short) {
  Check(rx_residual_conv(ctx), "rx_residual_conv");
  if (mirror >= 2) { Check(rx_get_status(ctx, nullptr), "Upwind_Residual"); MirrorResidual(); }   // the reference throws in this stage on a NaN residual
}
void CReactiveNSSolverGPU::Viscous_Residual(CGeometry*, CSolver**, CNumerics*, CConfig*, unsigned short, unsigned short) {
  Check(rx_residual_visc(ctx), "rx_residual_visc");
  if (mirror >= 2) { Check(rx_get_status(ctx, nullptr), "Viscous_Residual"); MirrorResidual(); }   // the reference throws in this stage on a NaN residual
}
void CReactiveNSSolverGPU::Source_Residual(CGeometry*, CSolver**, CNumerics*, CNumerics*, CConfig*, unsigned short) {
  Check(rx_residual_source(ctx), "rx_residual_source");
  if (mirror >= 2) { Check(rx_get_status(ctx, nullptr), "Source_Residual"); MirrorResidual(); }   // the reference throws in this stage on a NaN residual
}
// the state imposed at the boundary is what CTurbSSTSolver::BC_Inlet / BC_Outlet read as GetCharacPrimVar
void CReactiveNSSolverGPU::MirrorBoundaryState(CGeometry* geometry, unsigned short val_marker) {
  const unsigned long nv = geometry->nVertex[val_marker];
  std::vector<double> V((size_t)nv * nPrimVar);
  Check(rx_get_bc_state(ctx, val_marker, V.data()), "rx_get_bc_state");
  for (unsigned long k = 0; k < nv; ++k) {
    if (!geometry->node[geometry->vertex[val_marker][k]->GetNode()]->GetDomain()) continue;
    for (unsigned short v = 0; v < nPrimVar; ++v) CharacPrimVar[val_marker][k][v] = V[k * nPrimVar + v];
  }
}
void CReactiveNSSolverGPU::BC_Inlet(CGeometry* geometry, CSolver**, CNumerics*, CNumerics*, CConfig*, unsigned short val_marker) {
  Check(rx_bc_marker(ctx, val_marker), "rx_bc_marker");
  if (mirror >= 1 && ransMode) MirrorBoundaryState(geometry, val_marker);
  if (mirror >= 2) MirrorResidual();
}
void CReactiveNSSolverGPU::BC_Outlet(CGeometry* geometry, CSolver**, CNumerics*, CNumerics*, CConfig*, unsigned short val_marker) {
  Check(rx_bc_marker(ctx, val_marker), "rx_bc_marker");
  if (mirror >= 1 && ransMode) MirrorBoundaryState(geometry, val_marker);
  if (mirror >= 2) MirrorResidual();
}
void CReactiveNSSolverGPU::BC_Supersonic_Inlet(CGeometry*, CSolver**, CNumerics*, CNumerics*, CConfig*, unsigned short val_marker) {
  Check(rx_bc_marker(ctx, val_marker), "rx_bc_marker");
  if (mirror >= 2) MirrorResidual();
}
void CReactiveNSSolverGPU::BC_Supersonic_Outlet(CGeometry*, CSolver**, CNumerics*, CNumerics*, CConfig*, unsigned short val_marker) {
  Check(rx_bc_marker(ctx, val_marker), "rx_bc_marker");
  if (mirror >= 2) MirrorResidual();
}
void CReactiveNSSolverGPU::BC_HeatFlux_Wall(CGeometry*, CSolver**, CNumerics*, CNumerics*, CConfig*, unsigned short val_marker) {
  Check(rx_bc_marker(ctx, val_marker), "rx_bc_marker");
  if (mirror >= 2) MirrorResidual();
}
void CReactiveNSSolverGPU::BC_Euler_Wall(CGeometry*, CSolver**, CNumerics*, CConfig*, unsigned short val_marker) {
  Check(rx_bc_marker(ctx, val_marker), "rx_bc_marker");
  if (mirror >= 2) MirrorResidual();
}
void CReactiveNSSolverGPU::BC_Isothermal_Wall(CGeometry*, CSolver**, CNumerics*, CNumerics*, CConfig*, unsigned short val_marker) {
  Check(rx_bc_marker(ctx, val_marker), "rx_bc_marker");
  if (mirror >= 2) MirrorResidual();
}

// U = clip(U_old - alpha * Res * dt / Vol), residual norms; mirrors what output, restart and the host-side
// CSolver::Set_OldSolution read (solver_direct_reactive.cpp:2414-2493, solver_structure.cpp:172-260)
void CReactiveNSSolverGPU::Update(CGeometry* geometry, double alpha, bool lastStage) {
  std::vector<double> rms(nVar);
  Check(rx_update_explicit(ctx, alpha, rms.data()), "rx_update_explicit");
  if (commReady) {   // Set_MPI_Solution and the global norms, as ExplicitEuler/RK_Iteration do (:2442-2448)
    Set_MPI_Solution(geometry, nullptr);
    Check(rx_comm_allreduce_rms(ctx, rms.data()), "rx_comm_allreduce_rms");
  }
  std::vector<double> rmax(nVar);
  std::vector<int64_t> pmax(nVar);
  Check(rx_get_residual_max(ctx, rmax.data(), pmax.data()), "rx_get_residual_max");
  for (unsigned short v = 0; v < nVar; ++v) {
    if (rms[v] != rms[v]) throw std::runtime_error("SU2 has diverged (NaN detected in the residual norms)");
    SetRes_RMS(v, rms[v]);
    SetRes_Max(v, 0.0, 0);   // then what the reference's point loop leaves behind (AddRes_Max :2436)
    if (rmax[v] > 0.0)
      AddRes_Max(v, rmax[v], geometry->node[pmax[v]]->GetGlobalIndex(), geometry->node[pmax[v]]->GetCoord());
  }
  // node[]->Solution is what output, restart and monitors read — after the iteration; the intermediate RK stages are read by nobody on
  // the host (mirror level 2 keeps them for the stage-by-stage parity dumps)
  if (lastStage || mirror >= 2) {
    Check(rx_get_solution(ctx, bufU.data()), "rx_get_solution");
    for (unsigned long p = 0; p < nPoint; ++p) {
      su2double* sol = node[p]->GetSolution();
      for (unsigned short v = 0; v < nVar; ++v) sol[v] = bufU[p * nVar + v];
    }
  }
}
void CReactiveNSSolverGPU::ExplicitEuler_Iteration(CGeometry* geometry, CSolver**, CConfig*) { Update(geometry, 1.0, true); }
void CReactiveNSSolverGPU::ExplicitRK_Iteration(CGeometry* geometry, CSolver**, CConfig* config, unsigned short iRKStep) {
  Update(geometry, config->Get_Alpha_RKStep(iRKStep), iRKStep + 1 == config->GetnRKStep());
}

void CReactiveNSSolverGPU::WriteRestart(CGeometry* geometry, CSolver** solver_container, CConfig* config, const std::string& filename) {
  std::vector<int64_t> gi(nPoint);
  for (unsigned long p = 0; p < nPoint; ++p) gi[p] = (int64_t)geometry->node[p]->GetGlobalIndex();
  std::vector<double> kw;
  if (ransMode) {
    kw.resize((size_t)nPoint * 2);
    for (unsigned long p = 0; p < nPoint; ++p) { kw[2 * p] = solver_container[TURB_SOL]->node[p]->GetSolution(0); kw[2 * p + 1] = solver_container[TURB_SOL]->node[p]->GetSolution(1); }
  }
  rx_restart_info info{};
  info.aoa = config->GetAoA() - config->GetAoA_Offset(); info.sideslip = config->GetAoS() - config->GetAoS_Offset();
  info.bcThrust = config->GetInitial_BCThrust(); info.dCD_dCL = config->GetdCD_dCL();
  info.extIter = (int64_t)(config->GetExtIter() + config->GetExtIter_OffSet() + 1);
  Check(rx_restart_write(ctx, filename.c_str(), gi.data(), ransMode ? kw.data() : nullptr, &info, 0), "rx_restart_write");
}

// COutput::SetTecplotASCII for the volume solution (output_tecplot.cpp:46-420) straight from the device state: triangles of the
// (single-partition) geometry in element order, global point indices as MergeConnectivity writes them (:881)
void CReactiveNSSolverGPU::WriteTecplot(CGeometry* geometry, CSolver** solver_container, CConfig* config, const std::string& filename) {
  std::vector<int64_t> gi(nPoint), tria, quad;
  for (unsigned long p = 0; p < nPoint; ++p) gi[p] = (int64_t)geometry->node[p]->GetGlobalIndex();
  for (unsigned long e = 0; e < geometry->GetnElem(); ++e) {
    const unsigned short vtk = geometry->elem[e]->GetVTK_Type();
    if (vtk != TRIANGLE && vtk != QUADRILATERAL) throw std::runtime_error("CReactiveNSSolverGPU::WriteTecplot: 2-D meshes of triangles and quadrilaterals only");
    std::vector<int64_t>& dst = vtk == TRIANGLE ? tria : quad;
    for (unsigned short k = 0; k < (vtk == TRIANGLE ? 3 : 4); ++k) dst.push_back((int64_t)geometry->node[geometry->elem[e]->GetNode(k)]->GetGlobalIndex());
  }
  std::vector<double> kw;
  if (ransMode) {
    kw.resize((size_t)nPoint * 2);
    for (unsigned long p = 0; p < nPoint; ++p) { kw[2 * p] = solver_container[TURB_SOL]->node[p]->GetSolution(0); kw[2 * p + 1] = solver_container[TURB_SOL]->node[p]->GetSolution(1); }
  }
  std::ostringstream zone;   // unsteady runs label the zone (output_tecplot.cpp:277-279), with the stream's default formatting
  if (config->GetUnsteady_Simulation() && config->GetWrt_Unsteady())
    zone << "STRANDID=" << SU2_TYPE::Int(config->GetExtIter() + 1) << ", SOLUTIONTIME=" << config->GetDelta_UnstTime() * config->GetExtIter() << ", ";
  const std::string zp = zone.str();
  Check(rx_tecplot_write(ctx, filename.c_str(), gi.data(), ransMode ? kw.data() : nullptr, (int64_t)(tria.size() / 3), tria.data(), (int64_t)(quad.size() / 4), quad.data(),
                         zp.empty() ? nullptr : zp.c_str()), "rx_tecplot_write");
}

void CReactiveNSSolverGPU::SetTurbulenceOnDevice(const double* post12) {
  turbOnDevice = true;
  for (unsigned long p = 0; p < nPoint; ++p) {
    double* t = &bufT[p * 6]; const double* q = &post12[p * 12];
    t[0] = q[6]; t[1] = q[0]; t[2] = q[1]; t[3] = q[7]; t[4] = q[8]; t[5] = q[11];
  }
}
