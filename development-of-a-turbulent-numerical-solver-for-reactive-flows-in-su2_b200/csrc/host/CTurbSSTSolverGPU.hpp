// CTurbSSTSolverGPU — the turbulence half of the drop-in: a CTurbSSTSolver whose whole iteration (Preprocessing, Upwind / Viscous /
// Source residual, boundary conditions, implicit Euler step with FGMRES + LU-SGS, Postprocessing) runs on the device state of the
// flow solver (a CReactiveNSSolverGPU in solver_container[FLOW_SOL]) through rx_turb_iterate (include/rx_cuda.h,
// csrc/rx_turb_step.cuh).  CSingleGridIntegration::SingleGrid_Iteration (integration_time.cpp:777-834) drives it unchanged: the residual
// and boundary virtuals are no-ops, ImplicitEuler_Iteration does the step, the (k, omega, mu_t, F1, F2, CDkw, gradients) mirrors in
// node[] are refreshed for output and restart.  With it a RANS run hands nothing per point across PCIe between the two solvers.
// Factory switch: the `new CTurbSSTSolver(...)` of driver_structure.cpp:817.
#pragma once
#include <stdexcept>
#include "solver_structure.hpp"
#include "rx_cuda.h"

class CReactiveNSSolverGPU;

class CTurbSSTSolverGPU : public CTurbSSTSolver {
 public:
  CTurbSSTSolverGPU(CGeometry* geometry, CConfig* config, unsigned short iMesh) : CTurbSSTSolver(geometry, config, iMesh) {
    dualTime = config->GetUnsteady_Simulation() == DT_STEPPING_1ST || config->GetUnsteady_Simulation() == DT_STEPPING_2ND;
    // the device step is the first-order scheme of every shipped cfg (SPATIAL_ORDER_TURB defaults to 1ST_ORDER, config_structure.cpp:1189); the
    // reference's second-order branch (solver_direct_turbulent.cpp:463-520) indexes the four flow limiters with thirteen rows, and the compiled
    // reference ends in a segmentation fault on the channel case with SPATIAL_ORDER_TURB= 2ND_ORDER
    // rx_turb_iterate solves with FGMRES preconditioned by LU-SGS (CSysSolve::Solve, linear_solvers_structure.cpp:626-665), the pair of the shipped cfgs
    if (config->GetKind_Linear_Solver() != FGMRES || config->GetKind_Linear_Solver_Prec() != LU_SGS)
      throw std::runtime_error("CTurbSSTSolverGPU: LINEAR_SOLVER / LINEAR_SOLVER_PREC other than FGMRES / LU_SGS are not on the CUDA path (SU2_RX_TURB=host keeps the reference's own turbulence solver)");
    if (config->GetSpatialOrder_Turb() != FIRST_ORDER)
      throw std::runtime_error("CTurbSSTSolverGPU: SPATIAL_ORDER_TURB other than 1ST_ORDER is not on the CUDA path (SU2_RX_TURB=host keeps the reference's own turbulence solver)");
  }

  void Preprocessing(CGeometry*, CSolver**, CConfig*, unsigned short, unsigned short, unsigned short, bool) override {}
  void Upwind_Residual(CGeometry*, CSolver**, CNumerics*, CConfig*, unsigned short) override {}
  void Viscous_Residual(CGeometry*, CSolver**, CNumerics*, CConfig*, unsigned short, unsigned short) override {}
  void Source_Residual(CGeometry*, CSolver**, CNumerics*, CNumerics*, CConfig*, unsigned short) override {}
  void SetResidual_DualTime(CGeometry*, CSolver**, CConfig*, unsigned short, unsigned short, unsigned short) override {}   // inside rx_turb_iterate
  void BC_HeatFlux_Wall(CGeometry*, CSolver**, CNumerics*, CNumerics*, CConfig*, unsigned short) override {}
  void BC_Isothermal_Wall(CGeometry*, CSolver**, CNumerics*, CNumerics*, CConfig*, unsigned short) override {}
  void BC_Inlet(CGeometry*, CSolver**, CNumerics*, CNumerics*, CConfig*, unsigned short) override {}
  void BC_Outlet(CGeometry*, CSolver**, CNumerics*, CNumerics*, CConfig*, unsigned short) override {}
  void ImplicitEuler_Iteration(CGeometry* geometry, CSolver** solver_container, CConfig* config) override;
  // done on the device at the end of every step; the one call the driver makes right after constructing the solver
  // (driver_structure.cpp:817-819: flow Preprocessing, then this) forms the initial blending functions and eddy viscosity on the device
  void Postprocessing(CGeometry* geometry, CSolver** solver_container, CConfig* config, unsigned short iMesh) override;

  // node[]->Solution was changed on the host (restart, initial condition): hand (k, omega, F1, F2, CDkw) to the device before the next step
  void MarkHostSolutionDirty() { hostDirty = true; }

 private:
  void MirrorFromDevice(rx_ctx* ctx, CReactiveNSSolverGPU* flow);
  void SetWallDistance(rx_ctx* ctx, CGeometry* geometry);
  bool hostDirty = true, wallDistSet = false, stepped = false;
  bool dualTime = false;   // DUAL_TIME_STEPPING: the stored levels (k, omega)^n, (k, omega)^(n-1) go to the device before every step
  int mirror = 1;   // env SU2_RX_MIRROR as for the flow solver: 2 also mirrors LinSysRes / LinSysSol (parity dumps)
};
