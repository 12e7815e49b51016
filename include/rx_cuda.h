/* rx_cuda.h — C ABI of librx_cuda.so: the B200-native (sm_100a, FP64 CUDA) implementation of the reactive-flow
 * residual hot path of the SU2 fork LorenzoVallisa/Development-of-a-turbulent-numerical-solver-for-reactive-flows-in-SU2.
 *
 * This is the drop-in boundary: a host-side CSolver subclass (csrc/host/CReactiveNSSolverGPU.*, or the
 * Python mirror rx_solver.py) overrides the CSolver virtuals declared in SU2_CFD/include/solver_reactive.hpp:127-363,441-554
 * and forwards each to one entry point below.  Plain pointers and sizes only; no C++ or torch types.
 *
 * Conventions
 *  - every `const double*` / `double*` argument is a HOST pointer unless the name ends in `_dev`; the library copies at
 *    the call and never retains host pointers.  Arrays are array-of-structures in SU2's point numbering, exactly the
 *    layout of CVariable::Solution / CSysVector (Common/src/vector_structure.cpp:352-395):
 *        U[p][nVar]   = [rho, rho u, rho v, rho E, rho_1..rho_Ns]        (variable_direct_reactive.cpp:4-17)
 *        V[p][nPrim]  = [T, u, v, P, rho, h, a, Y_1..Y_Ns]
 *        grad[p][nGrad][2], rows [T, u, v, P, X_1..X_Ns]
 *  - every call returns 0 or a negative rx_status; rx_last_error() gives the text.  No exception crosses the ABI.
 *  - one host thread per rx_ctx; contexts are independent.  All work of a context is issued on one CUDA stream
 *    (rx_set_stream; default = the legacy default stream) and the stage calls do NOT synchronise the host.
 *  - two space dimensions (all configurations of BASELINE.json are 2-D); Ns must be one of the compiled instantiations
 *    (rx_supported_species).  There is no CPU fallback: without a CUDA device rx_create fails with RX_ERR_NO_DEVICE.
 */
#ifndef RX_CUDA_H
#define RX_CUDA_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct rx_ctx rx_ctx;

enum rx_status { RX_OK = 0, RX_ERR_ARG = -1, RX_ERR_NO_DEVICE = -2, RX_ERR_CUDA = -3, RX_ERR_UNSUPPORTED = -4, RX_ERR_NAN = -5 };
/* the three kinds of BC_Inlet's switch (INLET_TYPE, solver_direct_reactive.cpp:3280-3500): TEMPERATURE_IMPOSE reads T = temperature, Vmag = velocity;
 * TOTAL_CONDITIONS T = total temperature, Vmag = total pressure; MASS_FLOW T = density, Vmag = velocity (the second and third value of MARKER_INLET) */
enum rx_bc_kind { RX_BC_NONE = 0, RX_BC_INLET_TEMPERATURE_IMPOSE = 1, RX_BC_OUTLET = 2, RX_BC_ISOTHERMAL = 3, RX_BC_HEATFLUX = 4, RX_BC_EULER_WALL = 5,
                  RX_BC_INLET_TOTAL_CONDITIONS = 6, RX_BC_INLET_MASS_FLOW = 7,
                  RX_BC_SUPERSONIC_INLET = 8,    /* BC_Supersonic_Inlet :2998-3214: T, P, velocity vector in dir[], Ys; laminar / inviscid contexts */
                  RX_BC_SUPERSONIC_OUTLET = 9 }; /* BC_Supersonic_Outlet :3681-3807 */

/* Tables held by Framework::ReactingModelLibrary after Setup() (Common/src/Framework/reacting_model_library.cpp:1520-1586,
 * members reacting_model_library.hpp:655-766).  Matrices are [reaction][species]. */
typedef struct {
  int Ns, Nr, nKnots;
  const double* mMasses;   /* [Ns] g/mol                     (mMasses)            */
  const double* diffVol;   /* [Ns] Fuller diffusion volumes  (Diff_Volumes)       */
  const double* x;         /* [nKnots] equispaced temperature grid of the splines */
  const double* y;         /* [5][Ns][nKnots]: Mu, Kappa, Cp, Enth, Entr _Spline values   */
  const double* y2;        /* [5][Ns][nKnots]: their second derivatives (Tools/spline.cpp:10-58) */
  const double *As, *Betas, *Ta;                 /* [Nr] forward Arrhenius (SI)                 */
  const double *reversible, *availBack;          /* [Nr] 0/1: Reversible_Reactions, Available_Backward_Rate */
  const double *AsB, *BetasB, *TaB;              /* [Nr] backward Arrhenius where available     */
  const double *nuR, *nuP, *expR, *expP;         /* [Nr][Ns] Stoich_Coeffs_{Reactants,Products}[_Exp] */
} rx_chem;

/* One boundary marker: geometry->vertex[m][v]->{GetNode, GetNormal, GetNormal_Neighbor} and the marker's cfg values
 * (MARKER_INLET + INLET_MASS_FRAC / MARKER_OUTLET / MARKER_ISOTHERMAL / MARKER_HEATFLUX / MARKER_EULER). */
typedef struct {
  int kind;                 /* rx_bc_kind */
  int nVertex;
  const int64_t* node;            /* [nVertex] */
  const int64_t* normalNeighbor;  /* [nVertex] */
  const double* normal;           /* [nVertex][2] */
  double T, Vmag, dir[2];   /* inlet: GetInlet_Ttotal, GetInlet_Ptotal (velocity magnitude), GetInlet_FlowDir */
  const double* Ys;         /* inlet: [Ns] GetInlet_MassFrac */
  double P;                 /* outlet: GetOutlet_Pressure */
  double Twall;             /* isothermal: GetIsothermal_Temperature */
  double heatFlux;          /* heat-flux wall: GetWall_HeatFlux */
} rx_marker;

/* Flat export of CGeometry (Common/include/dual_grid_structure.hpp:140,747,884). Points [0,nPointDomain) are owned,
 * [nPointDomain,nPoint) are halos (geometry->node[p]->GetDomain()). */
typedef struct {
  int64_t nPoint, nPointDomain, nEdge;
  const int64_t* edgeNodes;   /* [nEdge][2]  edge[e]->GetNode(0/1) */
  const double* edgeNormal;   /* [nEdge][2]  edge[e]->GetNormal()  */
  const double* coord;        /* [nPoint][2] */
  const double* volume;       /* [nPoint]    */
  const int64_t* nbPtr;       /* [nPoint+1]  neighbour CSR, node[p]->GetPoint(k) order (WLS gradient) */
  const int64_t* nbPoint;
  int nMarker;
  const rx_marker* markers;
} rx_mesh;

/* Constants the reference reads from CConfig in the solver / numerics constructors
 * (numerics_direct_reactive.cpp:17-19,395-399,1417,1690-1694; solver_direct_reactive.cpp:4131-4383). */
typedef struct {
  int viscous, rans, leastSquares;
  double mInfty, cfl, maxDt, Tmin, Tmax;
  int clipT, ignition; int64_t ignitionIter; double ignitionT; int fuelIndex, oxidizerIndex;
  double prandtlLam, prandtlTurb, lewisTurb, cMu, pasrLb, tkeInf;
  int secondOrder, limiter;   /* SPATIAL_ORDER_FLOW = 2ND_ORDER[_LIMITER] (MUSCL on T,u,v,P :2554-2681); limiter: 1 = VENKATAKRISHNAN (:1443-1523), 2 = BARTH_JESPERSEN as coded (:1380-1439) */
  double limCoeff, refElemLength;   /* LIMITER_COEFF, REF_ELEM_LENGTH */
  int device;               /* CUDA device ordinal */
  int timeStepping;         /* UNSTEADY_SIMULATION = TIME_STEPPING: rx_time_step gives every point the global minimum time step (over all partitions
                               when a communicator exists), SetTime_Step :5262-5272; inviscid contexts follow the Euler solver's variant :2121-2139 */
  double unstCfl, deltaUnstTime;   /* UNST_CFL_NUMBER, UNST_TIMESTEP (the Euler variant replaces the CFL by the former and, when it is zero, uses the latter) */
  int dualTime;             /* UNSTEADY_SIMULATION = DUAL_TIME_STEPPING-1ST_ORDER (1) / -2ND_ORDER (2) with explicit pseudo-time iterations: rx_time_step
                               clamps the pseudo-time step to 2/3 of the physical one (:5274-5297), rx_residual_dual_time adds the physical-time source term */
  double deltaUnstTimeND;   /* the physical time step, CConfig::GetDelta_UnstTimeND; when unstCfl != 0 rx_time_step recomputes it on pseudo-time iteration 0 */
  int viscLimiter;          /* VISCOUS_LIMITER_FLOW: the interior viscous flux forms its mean gradients of T, u, v from gradients times the limiter of their
                               variable (numerics_direct_reactive.cpp:1482-1505); needs limiter != 0 (the reference reads an unset limiter array otherwise) */
} rx_opts;

int rx_supported_species(int Ns);            /* 1 if kernels for this species count are compiled in */
int rx_create(rx_ctx** out, const rx_mesh*, const rx_chem*, const rx_opts*);   /* replaces the CReactiveNSSolver ctor, solver_direct_reactive.cpp:4131 */
void rx_destroy(rx_ctx*);
const char* rx_last_error(rx_ctx*);          /* ctx may be NULL: error of the last failed rx_create */
int rx_set_stream(rx_ctx*, void* cudaStream);
int rx_synchronize(rx_ctx*);

/* mirrors of node[]->Solution / Primitive / LinSysRes (host <-> device, AoS, SU2 numbering) */
int rx_set_solution(rx_ctx*, const double* U);
int rx_set_temperature(rx_ctx*, const double* T);         /* V[T] = start value of the secant solve (variable_direct_reactive.cpp:622) */
int rx_set_turbulence(rx_ctx*, const double* turb6);      /* [nPoint][6]: muT, k, omega, dk/dx, dk/dy, sigma_k from CTurbSSTSolver */
/* SST Postprocessing on the device (first slice; csrc/rx_turb.cuh): replaces the host-side
 * CTurbSSTSolver::Postprocessing (solver_direct_turbulent.cpp:2953-3015) + rx_set_turbulence.  From the turbulence solution
 * (k, omega) and the wall distance (CGeometry::ComputeWall_Distance, geometry_structure.cpp:8990) it forms grad k, grad omega
 * (CSolver::SetSolution_Gradient_LS / _GG, solver_structure.cpp:580-722 / 519-579), CDkw, F1, F2 (CTurbSSTVariable::SetBlendingFunc) and
 * muT = min(max(rho k min(1/omega, a1/(S F2)), 0), 1) with rho, laminar viscosity and strain magnitude of the last rx_preprocess,
 * and leaves muT, k, omega, grad k, sigma_k where rx_set_turbulence would have put them.  RANS contexts; the gradient method is
 * the context's (WLS, or Green-Gauss: SetSolution_Gradient_GG :519-579).
 * stage: 0 = all of it (single partition); 1 = upload + gradients only, 2 = blending + muT only — between the two a
 * partitioned run exchanges the gradients of its halo points (rx_halo_pack/unpack with what = 3, Set_MPI_Solution_Gradient
 * solver_direct_turbulent.cpp:227). */
int rx_wall_distance(int64_t nPoint, const double* xy /* [nPoint][2] */, int64_t nWall, const double* wallXy /* [nWall][2]: nodes of the HEAT_FLUX / ISOTHERMAL markers */,
                     double* dist /* [nPoint] */, int device);   /* CPhysicalGeometry::ComputeWall_Distance geometry_structure.cpp:8990 (nearest wall NODE), host arrays, no context */
int rx_set_wall_distance(rx_ctx*, const double* dist /* [nPoint] */);
int rx_compute_wall_distance(rx_ctx*);   /* the same from the context's own mesh, on the device: nearest node of its ISOTHERMAL / HEAT_FLUX markers; a partition with a communicator (rx_comm_init) gathers the wall nodes of all partitions first, as the reference's MPI build does (collective: every rank calls it) */
int rx_turb_postprocess(rx_ctx*, const double* kOmega /* [nPoint][2]; NULL: the (k, omega) uploaded by the previous call */, double sigmaK /* CTurbSSTVariable::sigma_k = constants[0] */, int stage);
/* The SST transport step on the device (remainder; csrc/rx_turb_step.cuh): one
 * CSingleGridIntegration::SingleGrid_Iteration of CTurbSSTSolver (integration_time.cpp:777-834) — Preprocessing
 * (solver_direct_turbulent.cpp:2923), Upwind / Viscous / Source_Residual (:429, :521, :3018 with CUpwSca_TurbSST,
 * CAvgGradCorrected_TurbSST, CSourcePieceWise_TurbSST), BC_Inlet / BC_Outlet / BC_HeatFlux_Wall / BC_Isothermal_Wall (:3087-3450),
 * ImplicitEuler_Iteration (:615: FGMRES + LU-SGS of Common/src/linear_solvers_structure.cpp:309 and matrix_structure.cpp, clipped
 * conservative update) and Postprocessing — on the flow state the last rx_iterate / rx_preprocess left (primitives, gradients,
 * time steps, the boundary states of the inlet / outlet markers).  First-order upwind (SPATIAL_ORDER_TURB default).  On a partition
 * (a context with halo points) the step needs the library's communicator (rx_comm_init) and performs the reference's MPI steps over
 * NCCL on the context's stream: Set_MPI_Solution_Gradient after Preprocessing and Postprocessing (solver_direct_turbulent.cpp:227), the
 * neighbours' x* between the two LU-SGS sweeps and the preconditioned vector after them (CSysMatrix::SendReceive_Solution,
 * matrix_structure.cpp:1689, 1707), Set_MPI_Solution of (k, omega) after the update (:707), one all-reduce per inner product and for
 * the residual norms (dotProd, vector_structure.cpp:397-419); every rank must call it.
 * rx_turb_set_state: the turbulence solution (k, omega) [nPoint][2] and, optionally, F1, F2, CDkw [nPoint][3] of the previous
 * Postprocessing (a restart or a test hands them over; in a run they stay resident).  rx_turb_get_step: LinSysRes before the solve
 * and the solution of the linear system of the last step (parity tests).  resRms2 / linIters may be NULL. */
typedef struct { double kineInf, omegaInf;   /* free-stream k, omega (CTurbSSTSolver ctor :2751-2752) for the inlet */
                 double cflRed, relax;       /* CFL_REDUCTION_TURB, RELAXATION_FACTOR_TURB */
                 int linIter; double linTol; /* LINEAR_SOLVER_ITER, LINEAR_SOLVER_ERROR (FGMRES, LU_SGS) */ } rx_turb_opts;
int rx_turb_set_state(rx_ctx*, const double* kOmega, const double* blend3);
int rx_turb_iterate(rx_ctx*, const rx_turb_opts*, double* resRms2, int* linIters);
int rx_turb_get_step(rx_ctx*, double* res2 /* [nPoint][2] */, double* dx2 /* [nPoint][2] */);
/* dual time stepping of the SST solver: in a dual-time RANS context rx_turb_iterate adds CTurbSolver::SetResidual_DualTime
 * (solver_direct_turbulent.cpp:730-815: rho k, rho omega with the flow density of the three time levels) to the residual and the diagonal
 * Jacobian.  Its stored levels (k, omega)^n, (k, omega)^(n-1), [nPoint][2] each, live on the device like the flow solver's. */
int rx_turb_set_time_levels(rx_ctx*, const double* kwN, const double* kwN1);
int rx_turb_get_time_levels(rx_ctx*, double* kwN, double* kwN1);   /* either may be NULL */
int rx_turb_shift_time_levels(rx_ctx*);                            /* CIntegration::SetDualTime_Solver on the SST solver (iteration_structure.cpp:631) */
int rx_get_turbulence(rx_ctx*, double* post12 /* [nPoint][12]: k, omega, dist, F1, F2, CDkw, muT, dk/dx, dk/dy, domega/dx, domega/dy, sigma_k */);
int rx_get_solution(rx_ctx*, double* U);
int rx_get_primitive(rx_ctx*, double* V, double* aux6, double* grad);   /* aux: Cp, gamma, mu, kappa, vorticity_z, strain mag; any may be NULL */
int rx_get_dt(rx_ctx*, double* dt3);                      /* [nPoint][3]: Delta_Time, Max_Lambda_Inv, Max_Lambda_Visc */
int rx_get_residual(rx_ctx*, double* R);                  /* LinSysRes */
int rx_get_limiter(rx_ctx*, double* lim4);                /* [nPoint][4]: Limiter_Primitive of T, u, v, P */

/* stages; each replaces the CSolver virtual named on the right (solver_direct_reactive.cpp) */
int rx_preprocess(rx_ctx*, int64_t extIter, int output, int64_t* nonPhysical /* may be NULL: no host sync */);  /* Preprocessing :4693 */
int rx_set_old_solution(rx_ctx*);                         /* CSolver::Set_OldSolution, solver_structure.inl:898 */
int rx_time_step(rx_ctx*);                                /* SetTime_Step :5057 */
/* dual time stepping.  The stored time levels U^n, U^(n-1) (CVariable::Solution_time_n / _time_n1) live on the device; until they are set or
 * shifted for the first time they equal the current solution (what the reference's SetInitialCondition leaves). */
int rx_set_time_levels(rx_ctx*, const double* Un, const double* Un1);   /* [nPoint][nVar] each */
int rx_get_time_levels(rx_ctx*, double* Un, double* Un1);               /* either may be NULL */
int rx_shift_time_levels(rx_ctx*);                        /* CIntegration::SetDualTime_Solver, integration_structure.cpp:568-584: U^(n-1) <- U^n, U^n <- U */
int rx_set_inner_iteration(rx_ctx*, int64_t intIter);     /* the `Iteration` argument of SetTime_Step (the pseudo-time iteration index, :5057) */
int rx_get_delta_unst_time(rx_ctx*, double* dtND);        /* the physical time step in use (config->SetDelta_UnstTimeND, :5288); synchronises the stream */
int rx_residual_dual_time(rx_ctx*);                       /* SetResidual_DualTime :2172-2215 (static mesh, explicit) */
int rx_set_cfl(rx_ctx*, double cfl);                      /* config->GetCFL(iMesh) as SetTime_Step reads it at every call: CFL_ADAPT (COutput::SetCFL_Number,
                                                              output_structure.cpp:5990-6030, driver_structure.cpp:2804) changes it between iterations */
int rx_residual_conv(rx_ctx*);                            /* Upwind_Residual :2535 */
int rx_residual_visc(rx_ctx*);                            /* Viscous_Residual :5305 */
int rx_residual_source(rx_ctx*);                          /* Source_Residual :2792 */
int rx_bc(rx_ctx*);                                       /* BC_Inlet :3226, BC_Outlet :3808, BC_Euler_Wall :2881, then BC_Isothermal_Wall :5393, BC_HeatFlux_Wall :5717 (integration_structure.cpp:89-190) */
int rx_bc_marker(rx_ctx*, int marker);                    /* one BC_* call of the reference: the boundary condition of marker `marker` (order is the caller's) */
int rx_get_bc_state(rx_ctx*, int marker, double* V /* [nVertex][nPrim] */);   /* state imposed by the last inlet/outlet BC of the marker = CharacPrimVar (:3274, :3844), read by the turbulence BCs */
int rx_update_explicit(rx_ctx*, double rkAlpha, double* resRms /* [nVar] or NULL: no host sync */);  /* ExplicitEuler/RK_Iteration :2414,:2456 */

/* Events at which the reference throws (solver_direct_reactive.cpp:2772 "NaN found in the upwind residual", :5384 viscous, :2861 source;
 * variable_direct_reactive.cpp:683, :724 "Convergence not achieved for bisection method"), counted on the device:
 * status5 = { non-physical points of the last rx_preprocess, NaN conv flux, NaN visc flux, NaN source, failed temperature solves }.
 * Entries 1..4 stay set until read (this call, or rx_update_explicit / rx_iterate with resRms, which return RX_ERR_NAN with the
 * reference's message in rx_last_error).  Synchronises the host.  status5 may be NULL. */
int rx_get_status(rx_ctx*, int64_t* status5);

int rx_get_residual_max(rx_ctx*, double* resMax /* [nVar] */, int64_t* point /* [nVar], local point index */);   /* CSolver::AddRes_Max of the last update that returned norms: largest |LinSysRes| per variable and the first point attaining it (:2436) */
int rx_residual_all(rx_ctx*);                             /* Upwind + Viscous + Source residual in one pass (same sums, same order) */

/* one full MultiGrid_Cycle (integration_time.cpp:143-181) with everything resident: for each RK stage
 * preprocess -> (stage 0: set_old, time_step) -> fused conv+visc+source residual -> bc -> update; then the Output
 * preprocess pass (integration_time.cpp:110-112).  resRms may be NULL. */
int rx_iterate(rx_ctx*, int64_t extIter, int nRK, const double* rkAlpha, double* resRms);

/* diagnostic of the viscous kernel's processing order (no reference counterpart; csrc/rx_order.cuh): order[nEdge] = edge
 * handled by processing slot k (identity until the first rx_set_old_solution, which re-forms it once per iteration; on by default for the laminar path, off for RANS; env RX_VISC_ORDER=0/1 forces), key[nEdge] =
 * BiCGSTAB iterations of the edge's last Stefan-Maxwell solve (Solve_SM :451-470), + 19 if the SST closure ran its QR.
 * Either pointer may be NULL.  Results never depend on the order: fluxes are stored and summed by edge index. */
int rx_get_edge_order(rx_ctx*, int32_t* order, uint8_t* key);

/* stand-alone Arrhenius(+PaSR) source over independent points (CSourceReactive::ComputeChemistry,
 * numerics_direct_reactive.cpp:1728-1825).  Host-pointer and device-pointer forms. */
int rx_source_points(const rx_chem*, int rans, double cMu, double pasrLb, int64_t n, const double* T, const double* rho,
                     const double* Y, const double* omegaTurb, const double* vol, double* res, int device);
int rx_source_points_dev(rx_ctx*, int64_t n, const double* T_dev, const double* rho_dev, const double* Y_dev /* [Ns][n] SoA */,
                         const double* omegaTurb_dev, const double* vol_dev, double* res_dev /* [Ns][n] SoA */);

/* halo exchange plumbing for one-process-per-GPU runs (replaces Set_MPI_Solution :1530, Set_MPI_Primitive_Gradient :1877):
 * pack `what` (0 = conserved U, 1 = primitive gradients, 2 = limiters of T,u,v,P: Set_MPI_Primitive_Limiter :1756, 3 = grad k and grad omega of rx_turb_postprocess: Set_MPI_Solution_Gradient solver_direct_turbulent.cpp:227, 4 = the turbulence solution k, omega: CTurbSolver::Set_MPI_Solution :111) of the listed points into buf_dev[nVarOfWhat][n],
 * or unpack into them; the transport in between (NCCL send/recv) belongs to the caller. */
int rx_halo_pack(rx_ctx*, int what, int64_t n, const int64_t* points_dev, double* buf_dev);
int rx_halo_unpack(rx_ctx*, int what, int64_t n, const int64_t* points_dev, const double* buf_dev);
int rx_halo_width(rx_ctx*, int what);

/* The same exchange INSIDE the library, over NCCL (csrc/rx_comm.cuh; NCCL is bound at run time with dlopen, env RX_NCCL_LIB overrides
 * the name): one process per GPU, one rx_ctx per process.
 *   rx_comm_unique_id : ncclGetUniqueId on one rank; the 128 bytes travel to the others by the host's own means (MPI_Bcast in SU2).
 *   rx_comm_init      : ncclCommInitRank + the neighbour lists of this partition: for neighbour k (rank neighborRank[k]) the local points
 *                       sendPoints[sendPtr[k] .. sendPtr[k+1]) go out and recvPoints[recvPtr[k] .. recvPtr[k+1]) (halo points) come in, in the
 *                       order both sides agreed on — the vertex order of the SEND_RECEIVE marker pair in SU2 (geometry->vertex[MarkerS/R]).
 *   rx_halo_exchange  : Set_MPI_Solution (what 0, solver_direct_reactive.cpp:1530), Set_MPI_Primitive_Gradient (1, :1877),
 *                       Set_MPI_Primitive_Limiter (2, :1756), Set_MPI_Solution_Gradient of (k, omega) (3), Set_MPI_Solution of (k, omega) (4): pack, grouped ncclSend/ncclRecv,
 *                       unpack — enqueued on the context's stream, no host synchronisation.  rx_iterate calls it at the reference's points
 *                       once a communicator exists, and returns the norms of the whole job.
 *   rx_comm_allreduce_rms : residual norms of the whole job from every partition's sums of squares of the last rx_update_explicit
 *                       (CSolver::SetResidual_RMS under MPI, solver_structure.cpp:188-260): one ncclAllReduce of nVar + 1 doubles. */
int rx_comm_unique_id(void* id128);
int rx_comm_init(rx_ctx*, const void* id128, int rank, int nRanks, int nNeighbors, const int* neighborRank,
                 const int64_t* sendPtr, const int64_t* sendPoints, const int64_t* recvPtr, const int64_t* recvPoints);
int rx_halo_exchange(rx_ctx*, int what);
int rx_comm_allreduce_rms(rx_ctx*, double* resRms /* [nVar] host, or NULL */);

/* measurement hook: average device time in ms (CUDA events on the context's stream) of `reps` launches of one kernel on the
 * current state; which = 0 viscous edge flux (convective if inviscid), 1 convective edge flux, 2 accumulation+source,
 * 3 cons->prim, 4 gradient, 5 source, 6 explicit update, 7 laminar transport */
int rx_kernel_time(rx_ctx*, int which, int reps, double* msAvg);

/* measurement hook: FP64 throughput of the device the context lives on, from a register-resident DFMA loop (8 independent
 * chains per thread, every SM filled), in TFLOP/s with one fused multiply-add counted as two operations.  The compute
 * roof of the viscous and source kernels (MEASURED_PEAKS.json carries no FP64 entry). */
int rx_fp64_peak(rx_ctx*, double* tflops);

/* diagnostic hook: the device's glibc-exact exp (which = 0), pow (1; y required) and log (2) of csrc/rx_libm.cuh over host arrays —
 * the functions the kernels call where the reference calls std::exp / std::pow / std::log (reacting_model_library.cpp:751-766,
 * 835-920; solver_direct_reactive.cpp:3901-3935).  tests/test_gpu_libm.py compares them bit for bit with the host's libm. */
int rx_libm_eval(int which, int64_t n, const double* x, const double* y, double* out, int device);

/* diagnostic hook: x = Eigen::ColPivHouseholderQR<MatrixXd>(A).solve(b) for n independent 9 x 9 systems (A[n][81] column-major, two
 * right-hand sides b[n][2][9]) — the call of the SST closure, numerics_direct_reactive.cpp:743 — by both device forms: the serial routine
 * (boundary vertices) into xSerial and the two-lanes-per-system routine of the viscous edge kernel into xPair.  Both follow Eigen 3.3.7's
 * own summation orders (externals/Eigen/src/QR/ColPivHouseholderQR.h:480-609); tests/test_gpu_qr.py compares them bit for bit with the
 * oracle's restatement, which oracle/eigen_qr_check.cpp pins against the real Eigen headers. */
int rx_qr_eval(int64_t n, const double* A, const double* b, double* xSerial, double* xPair, int device);

/* partitioning (host-side graph tools, no device work; csrc/host/rx_partition.cpp).
 * rx_partition_kway: k-way partition of the point adjacency graph (CSR nbPtr/nbPoint, the same arrays rx_mesh carries) into
 * nParts parts of equal size (2 % tolerance) with a small edge cut: recursive inertial bisection on coord[nPoint][2] +
 * boundary refinement (refinePasses sweeps; 0 = geometric split only).  owner[nPoint] receives the part of every point,
 * *edgeCut (may be NULL) the number of cut edges.  When librx_metis.so is present the partition is METIS's (coord and refinePasses unused).
 * Stands where the reference calls ParMETIS_V3_PartKway
 * (Common/src/geometry_structure.cpp:11524 via SetColorGrid_Parallel, driver_structure.cpp:126).
 * rx_rcm_order: reverse Cuthill-McKee numbering of the first nPointDomain (owned) points, halo points stay last;
 * newToOld[nPoint].  Role of CPhysicalGeometry::SetRCM_Ordering (geometry_structure.cpp:9200, called driver_structure.cpp:560). */
int rx_partition_kway(int64_t nPoint, const int64_t* nbPtr, const int64_t* nbPoint, const double* coord, int nParts,
                      int refinePasses, int64_t* owner, int64_t* edgeCut);
/* One rank's share of a partitioned mesh in SU2's scheme (owned points, one halo layer grouped by owner rank, every edge with an owned end,
 * boundary vertices on owned points, matching send / receive lists per neighbour rank; rcm != 0: owned points renumbered by rx_rcm_order, edges
 * re-sorted) — the role of the partitioned CPhysicalGeometry constructor and SetSendReceive (geometry_structure.cpp:540-1480, :2262-2540).
 * Markers of the whole mesh are passed concatenated: vertices of marker m = [markerPtr[m], markerPtr[m+1]).  All local indices refer to the
 * part's numbering; glob[] maps them back.  Arrays are malloc'ed; release with rx_part_free. */
typedef struct {
  int64_t nPoint, nPointDomain, nEdge;
  int64_t* glob;          /* [nPoint] local -> global point index */
  int64_t* edgeNodes; double* edgeNormal; double* coord; double* volume; int64_t* nbPtr; int64_t* nbPoint;   /* as rx_mesh */
  int nMarker;
  int64_t* markerPtr;     /* [nMarker+1] */
  int64_t* markerNode; int64_t* markerNormalNeighbor; double* markerNormal;
  int nNb;                /* neighbour ranks, ascending */
  int* nbRank;            /* [nNb] */
  int64_t* sendPtr; int64_t* sendPts;   /* [nNb+1], owned local points to send to neighbour k, in that neighbour's halo order */
  int64_t* recvPtr; int64_t* recvPts;   /* [nNb+1], local halo points received from neighbour k */
} rx_part;
int rx_partition_mesh(int64_t nPoint, int64_t nEdge, const int64_t* edgeNodes, const double* edgeNormal, const double* coord, const double* volume,
                      const int64_t* nbPtr, const int64_t* nbPoint, int nMarker, const int64_t* markerPtr, const int64_t* markerNode,
                      const int64_t* markerNormalNeighbor, const double* markerNormal, const int64_t* owner, int rank, int rcm, rx_part* out);
void rx_part_free(rx_part*);
int rx_partition_backend(void);   /* 1: rx_partition_kway calls METIS_PartGraphKway of librx_metis.so (METIS 5 as vendored in the reference tree, built by
                                    csrc/host/build_metis.sh; found next to the library or through RX_METIS_LIB); 0: the library's own partitioner
                                    (also forced by RX_PARTITIONER=own) */
int rx_rcm_order(int64_t nPoint, int64_t nPointDomain, const int64_t* nbPtr, const int64_t* nbPoint, int64_t* newToOld);

/* geometry (host-side, no device work; csrc/host/rx_geometry.cpp).
 * rx_dual_mesh_tri: the median-dual mesh of a triangulated 2-D point set — what CGeometry::SetEdges and CPhysicalGeometry::SetControlVolume
 * (Common/src/geometry_structure.cpp:10457) leave in edge[] / node[]: edges (lower, higher point) in that order, edge normals, dual volumes
 * and the point-to-point CSR sorted by neighbour index — as flat arrays in the layout of rx_mesh.  The arrays are malloc'ed by the
 * library; release them with rx_dual_mesh_free. */
typedef struct {
  int64_t nEdge;
  int64_t* edgeNodes;   /* [nEdge][2] */
  double* edgeNormal;   /* [nEdge][2] */
  double* volume;       /* [nPoint]   */
  int64_t* nbPtr;       /* [nPoint+1] */
  int64_t* nbPoint;     /* [2 nEdge]  */
} rx_dual_mesh;
int rx_dual_mesh_tri(int64_t nPoint, const double* coord /* [nPoint][2] */, int64_t nTri, const int64_t* tri /* [nTri][3] */, rx_dual_mesh* out);
void rx_dual_mesh_free(rx_dual_mesh*);

/* An ASCII .su2 mesh file (2-D, triangles and quadrilaterals) -> the flat dual mesh in the REFERENCE's point, edge, neighbour and vertex
 * numbering (csrc/host/rx_su2_mesh.cpp): CPhysicalGeometry::Read_SU2_Format (geometry_structure.cpp:4819), SetPoint_Connectivity :9145,
 * SetRCM_Ordering :9200 (rcm != 0), Check_IntElem_/Check_BoundElem_Orientation :8542/:8825, SetEdges :8680, SetControlVolume :10457, SetVertex :9451,
 * SetBoundControlVolume :9595, FindNormal_Neighbor.  rx_su2_mesh_build takes the parsed arrays instead of a file (elems [nElem][4], -1 padded;
 * boundary lines of marker m = seg[segPtr[m] .. segPtr[m+1]) as node pairs).  Arrays are malloc'ed; release with rx_su2_mesh_free.
 * rx_su2_mesh_error: message of the calling thread's last failed call. */
typedef struct {
  int64_t nPoint, nEdge, nElem;
  double* coord;          /* [nPoint][2], in the new numbering */
  int64_t* globalIndex;   /* [nPoint]: the file's index of every point (CPoint::GetGlobalIndex) */
  int64_t* etype;         /* [nElem]: 5 triangle, 9 quadrilateral */
  int64_t* elems;         /* [nElem][4], oriented, -1 padded */
  int64_t* edgeNodes;     /* [nEdge][2] */
  double* edgeNormal;     /* [nEdge][2] */
  double* volume;         /* [nPoint] */
  int64_t* nbPtr;         /* [nPoint+1] */
  int64_t* nbPoint;       /* [2 nEdge] */
  int nMarker;
  char** markerTag;                 /* [nMarker] */
  int64_t* markerNVertex;           /* [nMarker] */
  int64_t** markerNode;             /* [nMarker][nVertex] */
  int64_t** markerNormalNeighbor;   /* [nMarker][nVertex] */
  double** markerNormal;            /* [nMarker][nVertex][2], pointing into the domain (SU2's convention) */
} rx_su2_mesh;
int rx_su2_mesh_read(const char* path, int rcm, rx_su2_mesh* out);
int rx_su2_mesh_build(int64_t nPoint, const double* coord, int64_t nElem, const int64_t* etype, const int64_t* elems, int nMarker,
                      const char* const* tags, const int64_t* segPtr, const int64_t* seg, int rcm, rx_su2_mesh* out);
void rx_su2_mesh_free(rx_su2_mesh*);
const char* rx_su2_mesh_error(void);

/* SU2 native ASCII restart files straight from / into the device-resident state (csrc/rx_restart.cuh).
 * rx_restart_write replaces COutput::SetRestart for the reactive solvers (output_structure.cpp:3858-4060; columns of MergeSolution
 * :1991-2130): "PointID" x y Conservative_1..nVar [k omega] Pressure Temperature Mach [Laminar_Viscosity [mu_t]], every value as
 * C++ `scientific` with precision 15, then the trailer (AOA, SIDESLIP_ANGLE, INITIAL_BCTHRUST, DCD_DCL_VALUE, EXT_ITER) — byte for byte the
 * file the reference writes for the same state (tests/test_gpu_restart.py).  Primitive columns are those of the last rx_preprocess.
 * globalIndex[nPoint] (NULL = identity) gives the PointID of every local point; kOmega[nPoint][2] is the turbulence solution (RANS;
 * it lives in CTurbSSTSolver) ; piece != 0 writes the owned points' lines only (no header / trailer) for per-partition writers,
 * merged by PointID afterwards.  The fields are staged through pinned memory in chunks: no host mirror of the state is needed.
 * rx_restart_read replaces CReactiveNSSolver::Load_Restart (solver_direct_reactive.cpp:4390-4520) and the turbulence reader
 * (solver_direct_turbulent.cpp:1046-1064): conserved variables of every local point (halo points included) are set from the file,
 * kOmega (may be NULL) receives (k, omega); fails with the reference's message if the file does not cover the owned points. */
typedef struct { double aoa, sideslip, bcThrust, dCD_dCL; int64_t extIter; } rx_restart_info;
int rx_restart_write(rx_ctx*, const char* path, const int64_t* globalIndex, const double* kOmega, const rx_restart_info* info /* NULL: zeros, INITIAL_BCTHRUST 4000 */, int piece);
/* rx_tecplot_write replaces COutput::SetTecplotASCII for the volume solution of the reactive solvers (SU2_CFD/src/output_tecplot.cpp:46-420,
 * fed by MergeConnectivity / MergeSolution): title, the VARIABLES line as the reference prints it, the ZONE line, one line per point
 * (x, y and the restart file's columns in C++ `scientific` with the stream's default precision) and the 1-based connectivity of the
 * nTria triangles tria[nTria][3] (point indices in the numbering of globalIndex, or local if NULL).  Single partition (a partition
 * writes a restart piece; the merged restart goes through SU2_SOL).  Byte-identical to the reference's file for the laminar case. */
int rx_tecplot_write(rx_ctx*, const char* path, const int64_t* globalIndex, const double* kOmega, int64_t nTria, const int64_t* tria,
                     int64_t nQuad, const int64_t* quad /* [nQuad][4], written after the triangles */,
                     const char* zonePrefix /* NULL, or what an unsteady run puts between "ZONE " and "NODES=": "STRANDID=3, SOLUTIONTIME=2e-07, " (:277-279) */);
int rx_restart_read(rx_ctx*, const char* path, const int64_t* globalIndex, int64_t nGlobal /* points in the file; <= 0: nPointDomain */, double* kOmega, rx_restart_info* info /* may be NULL */);

/* measurement hook (no reference counterpart): save (restore = 0) or restore (restore = 1) the evolving state of the context — U,
 * U_old, V (with the secant start temperature), the turbulence fields, the viscous kernel's edge order — device to device, on the
 * context's stream, so that the same iteration can be timed from the same state any number of times. */
int rx_state_snapshot(rx_ctx*, int restore);

/* number of kernels this context has launched so far (bench.py's gpu_launches) */
int64_t rx_launch_count(rx_ctx*);

/* Allocation-free residual norms.  NO COUNTERPART IN THE REFERENCE: added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The same
 * reduction rx_update_explicit performs for CSolver::SetResidual_RMS — rms[v] = max(EPS^2, sqrt(sum_{p < nPointDomain}
 * R[v][p]^2 / nPointGlobal)), same partial sums in the same order, so bit-equal to that entry point for the same LinSysRes —
 * but context-free and with its scratch (per-block partial sums) in a file-scope static __device__ array of fixed size
 * (4 MiB per GPU that loads the library, held until exit) instead of rx_ctx memory.  It calls no cudaMalloc / cudaFree / allocating
 * routine.  Inputs and outputs are the caller's DEVICE pointers: R_dev[nVar][ld] structure-of-arrays (variable v of point p at
 * v * ld + p, ld >= nPointDomain), rms_dev[nVar]; work is enqueued on cudaStream of `device` and the host is not synchronised.
 * Limits: nPointDomain <= RX_STATIC_RMS_MAX_POINTS and nVar <= RX_STATIC_RMS_MAX_VARS, else RX_ERR_UNSUPPORTED with the
 * limits in rx_last_error(NULL).  All calls on a GPU share the one array, which keeps the last call's contents (each call
 * writes every slot it reads, nothing needs clearing): calls that overlap on the same GPU race — serialise them. */
#define RX_STATIC_RMS_MAX_POINTS 4194304   /* 4 Mi owned points */
#define RX_STATIC_RMS_MAX_VARS 16
int rx_residual_rms_static(const double* R_dev, int64_t nPointDomain, int64_t ld, int nVar, int64_t nPointGlobal,
                           double* rms_dev, int device, void* cudaStream);

#ifdef __cplusplus
}
#endif
#endif
