// CReactiveNSSolverGPU — host side of the drop-in: a CSolver the reference's driver, integration and output code use
// unchanged, whose residual path runs in librx_cuda.so (include/rx_cuda.h).
//
// It derives from the reference's CReactiveNSSolver (SU2_CFD/include/solver_reactive.hpp:398-560), so construction
// (free-stream state, node[] allocation, limits, restart load) is the reference's own; the overrides below replace the
// virtuals that MultiGrid_Cycle / Space_Integration / Time_Integration call (integration_time.cpp:143-181,
// integration_structure.cpp:66-190, :307-330), with the signatures of solver_reactive.hpp.  The CNumerics* arguments are
// accepted and ignored: the numerics' constants were read from CConfig at construction, as the numerics constructors do.
//
// This translation unit is compiled WITH the reference's headers and linked into its SU2_CFD; nothing of
// the reference is copied here.
#pragma once
#include "solver_reactive.hpp"
#include "rx_cuda.h"

#include <string>
#include <vector>

// The communicator set-up (InitComm) needs five MPI symbols.  In an MPI build they are the real ones; RX_MPI_STUB supplies single-rank stand-ins
// so that the block can be compiled (syntax and types) where no MPI exists (oracle/build_ref.sh does, into a discarded object).
#if defined(HAVE_MPI)
#define RX_COMM_BLOCK 1
#elif defined(RX_MPI_STUB)
#define RX_COMM_BLOCK 1
typedef int MPI_Comm;
typedef int MPI_Datatype;
static const MPI_Comm MPI_COMM_WORLD = 0;
static const MPI_Datatype MPI_CHAR = 1;
static inline int MPI_Comm_rank(MPI_Comm, int* rank) { *rank = 0; return 0; }
static inline int MPI_Comm_size(MPI_Comm, int* size) { *size = 1; return 0; }
static inline int MPI_Bcast(void*, int, MPI_Datatype, int, MPI_Comm) { return 0; }
#endif

class CReactiveNSSolverGPU : public CReactiveNSSolver {
 public:
  // mirror levels (env SU2_RX_MIRROR, default 1):
  //   0  conserved variables and residual norms after each update (what output and restart read)
  //   1  + primitives, gradients, transport, vorticity/strain after each Preprocessing (what CTurbSSTSolver reads)
  //   2  + LinSysRes after each residual / boundary stage and the time steps (stage-by-stage parity dumps)
  CReactiveNSSolverGPU(CGeometry* geometry, CConfig* config, unsigned short iMesh);
  ~CReactiveNSSolverGPU() override;

  void Preprocessing(CGeometry* geometry, CSolver** solver_container, CConfig* config, unsigned short iMesh,
                     unsigned short iRKStep, unsigned short RunTime_EqSystem, bool Output) override;
  void SetTime_Step(CGeometry* geometry, CSolver** solver_container, CConfig* config, unsigned short iMesh,
                    unsigned long Iteration) override;
  void Upwind_Residual(CGeometry* geometry, CSolver** solver_container, CNumerics* numerics, CConfig* config,
                       unsigned short iMesh) override;
  void Viscous_Residual(CGeometry* geometry, CSolver** solver_container, CNumerics* numerics, CConfig* config,
                        unsigned short iMesh, unsigned short iRKStep) override;
  void Source_Residual(CGeometry* geometry, CSolver** solver_container, CNumerics* numerics, CNumerics* second_numerics,
                       CConfig* config, unsigned short iMesh) override;
  void SetResidual_DualTime(CGeometry* geometry, CSolver** solver_container, CConfig* config, unsigned short iRKStep,
                            unsigned short iMesh, unsigned short RunTime_EqSystem) override;
  void BC_Inlet(CGeometry* geometry, CSolver** solver_container, CNumerics* conv_numerics, CNumerics* visc_numerics,
                CConfig* config, unsigned short val_marker) override;
  void BC_Outlet(CGeometry* geometry, CSolver** solver_container, CNumerics* conv_numerics, CNumerics* visc_numerics,
                 CConfig* config, unsigned short val_marker) override;
  void BC_Supersonic_Inlet(CGeometry* geometry, CSolver** solver_container, CNumerics* conv_numerics, CNumerics* visc_numerics,
                           CConfig* config, unsigned short val_marker) override;
  void BC_Supersonic_Outlet(CGeometry* geometry, CSolver** solver_container, CNumerics* conv_numerics, CNumerics* visc_numerics,
                            CConfig* config, unsigned short val_marker) override;
  void BC_Isothermal_Wall(CGeometry* geometry, CSolver** solver_container, CNumerics* conv_numerics,
                          CNumerics* visc_numerics, CConfig* config, unsigned short val_marker) override;
  void BC_HeatFlux_Wall(CGeometry* geometry, CSolver** solver_container, CNumerics* conv_numerics,
                        CNumerics* visc_numerics, CConfig* config, unsigned short val_marker) override;
  void BC_Euler_Wall(CGeometry* geometry, CSolver** solver_container, CNumerics* numerics, CConfig* config,
                     unsigned short val_marker) override;
  void ExplicitEuler_Iteration(CGeometry* geometry, CSolver** solver_container, CConfig* config) override;
  void ExplicitRK_Iteration(CGeometry* geometry, CSolver** solver_container, CConfig* config,
                            unsigned short iRKStep) override;

  // Halo exchanges (solver_reactive.hpp:162,176,183).  With a communicator (MPI builds: InitComm at construction) they run inside
  // the library over NCCL between the device states (rx_halo_exchange); in a serial build there is no SEND_RECEIVE marker and they
  // are no-ops, as the base versions are.
  void Set_MPI_Solution(CGeometry* geometry, CConfig* config) override;
  void Set_MPI_Primitive_Gradient(CGeometry* geometry, CConfig* config) override;
  void Set_MPI_Primitive_Limiter(CGeometry* geometry, CConfig* config) override;

  // COutput::SetRestart for this solver, straight from the device state (rx_restart_write): the same bytes the reference's
  // MergeSolution + SetRestart produce (output_structure.cpp:1991-2130, 3858-4060) without mirroring the state into node[].
  // One call in COutput::SetResult_Files replaces the pair for a CReactiveNSSolverGPU flow solver.
  void WriteRestart(CGeometry* geometry, CSolver** solver_container, CConfig* config, const std::string& filename);
  void WriteTecplot(CGeometry* geometry, CSolver** solver_container, CConfig* config, const std::string& filename);   // rx_tecplot_write

  // Dual time stepping.  CIntegration::SetDualTime_Solver (integration_structure.cpp:568-584) is not virtual: it shifts the time levels in
  // node[], and the next physical time step's first SetTime_Step uploads them.  Calling ShiftTimeLevels() next to it (one line,
  // INTEGRATION.md 4f) shifts them on the device instead and saves that upload.
  void ShiftTimeLevels();
  // node[]->Solution was changed on the host (restart, initial condition): upload before the next stage
  void MarkHostSolutionDirty() { hostDirty = true; turbOnDevice = false; levelsUploaded = levelsOnDevice = false; }   // the turbulence fields are taken from the host again, too
  rx_ctx* Context() { return ctx; }
  // called by CTurbSSTSolverGPU after its step: the six turbulence fields the flow path reads are current on the device (no upload at
  // the next Preprocessing); post12 = rx_get_turbulence rows, for the host mirror of the eddy viscosity
  void SetTurbulenceOnDevice(const double* post12);

 private:
  void Check(int rc, const char* where);
  void UploadState();
  void UploadTurbulence(CSolver** solver_container);
  void DeviceTurbPostprocessing(CGeometry* geometry, CSolver** solver_container);
  void MirrorPrimitives();
  void MirrorResidual();
  void UploadTimeLevels();
  void MirrorLimiter();
  void MirrorTimeStep();
  void MirrorBoundaryState(CGeometry* geometry, unsigned short val_marker);
  void Update(CGeometry* geometry, double alpha, bool lastStage);

#ifdef RX_COMM_BLOCK
  void InitComm(CGeometry* geometry, CConfig* config);   // neighbour lists from the SEND_RECEIVE marker pairs, NCCL id by MPI_Bcast
#endif
  rx_ctx* ctx = nullptr;
  bool commReady = false;
  bool hostDirty = true, ransMode = false, viscousMode = true;
  int dualTime = 0;                                        // 0, or the order of DUAL_TIME_STEPPING
  bool levelsUploaded = false, levelsOnDevice = false;     // the device holds time levels at all / holds the ones of the coming physical step
  int mirror = 1;
  // SU2_RX_TURB_POST=1: the fields the flow path reads from the SST solver (muT, grad k, sigma_k) are formed on the device from
  // (k, omega) by rx_turb_postprocess instead of being copied from CTurbSSTSolver::Postprocessing's results
  bool turbPost = false, turbPostReady = false, wallDistSet = false, turbOnDevice = false;
  unsigned long extIter = 0;
  std::vector<double> bufU, bufV, bufAux, bufGrad, bufR, bufT;
};
