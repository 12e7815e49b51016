/* TEST INFRASTRUCTURE — CPU restatement ("oracle") of the reactive-flow residual hot path.
 *
 * Nothing on the product path may include, link or call this.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs use it, as the checker.
 *
 * Parity status: PINNED.  Every stage below is checked in tests/test_oracle_golden.py against dumps
 * produced by the reference's own classes (oracle/ref_harness.cpp linked to oracle/_ref/libsu2ref.a,
 * built from /root/reference by oracle/build_ref.sh) and committed under tests/golden/.
 *
 * Layout mirrors the reference (array-of-structures, SU2 point/edge numbering):
 *   U[p][nVar]      = [rho, rho u, rho v, rho E, rho_1..rho_Ns]
 *   V[p][nPrim]     = [T, u, v, P, rho, h, a, Y_1..Y_Ns]
 *   grad[p][nGrad][2], rows = [T, u, v, P, X_1..X_Ns]   (mole fractions)
 * Two space dimensions only (all BASELINE configs are 2-D).
 */
#ifndef RX_ORACLE_H
#define RX_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct rxo_ctx rxo_ctx;

/* chemistry tables exactly as Framework::ReactingModelLibrary holds them after Setup()
 * (Common/src/Framework/reacting_model_library.cpp:1520-1586). Matrices are [reaction][species]. */
typedef struct {
  int Ns, Nr, nKnots;
  const double* mMasses;      /* [Ns] g/mol */
  const double* diffVol;      /* [Ns] */
  const double* x;            /* [nKnots] temperature grid (equispaced) */
  const double* y;            /* [5][Ns][nKnots]  mu, kappa, cp, h, s */
  const double* y2;           /* [5][Ns][nKnots] */
  const double* As; const double* Betas; const double* Ta;          /* [Nr] */
  const double* reversible; const double* availBack;                /* [Nr] 0/1 */
  const double* AsB; const double* BetasB; const double* TaB;       /* [Nr] */
  const double* nuR; const double* nuP; const double* expR; const double* expP; /* [Nr][Ns] */
} rxo_chem;

typedef struct {
  int kind;            /* 1 inlet (TEMPERATURE_IMPOSE), 2 outlet, 3 isothermal wall, 4 heat-flux wall, 5 Euler wall, 0 ignored */
  int nVertex;
  const int64_t* node; const int64_t* normalNeighbor; const double* normal; /* [nVertex], [nVertex], [nVertex][2] */
  double T, Vmag, dir[2]; const double* Ys;   /* inlet */
  double P;                                   /* outlet */
  double Twall;                               /* isothermal */
  double heatFlux;                            /* heat-flux wall */
} rxo_marker;

typedef struct {
  int64_t nPoint, nPointDomain, nEdge;
  const int64_t* edgeNodes;   /* [nEdge][2] */
  const double* edgeNormal;   /* [nEdge][2] */
  const double* coord;        /* [nPoint][2] */
  const double* volume;       /* [nPoint] */
  const int64_t* nbPtr; const int64_t* nbPoint;   /* neighbour CSR in CPoint::GetPoint order */
  int nMarker; const rxo_marker* markers;
} rxo_mesh;

typedef struct {
  int viscous, rans, leastSquares;     /* REACTIVE_NAVIER_STOKES / REACTIVE_RANS / WEIGHTED_LEAST_SQUARES */
  double mInfty, cfl, maxDt, Tmin, Tmax;
  int clipT, ignition; int64_t ignitionIter; double ignitionT; int fuelIndex, oxidizerIndex;
  double prandtlLam, prandtlTurb, lewisTurb, cMu, pasrLb, tkeInf;
  int secondOrder, limiter;            /* SPATIAL_ORDER_FLOW = 2ND_ORDER[_LIMITER]; limiter: 1 = VENKATAKRISHNAN */
  double limCoeff, refElemLength;      /* LIMITER_COEFF, REF_ELEM_LENGTH */
  int timeStepping;                    /* UNSTEADY_SIMULATION = TIME_STEPPING: every point advances with the global minimum time step */
  double unstCfl, deltaUnstTime;       /* UNST_CFL_NUMBER, UNST_TIMESTEP (read by the Euler solver's variant only) */
  int dualTime;                        /* UNSTEADY_SIMULATION = DUAL_TIME_STEPPING-1ST_ORDER (1) / -2ND_ORDER (2), explicit pseudo-time iterations */
  double deltaUnstTimeND;              /* the physical time step (CConfig::GetDelta_UnstTimeND); recomputed from unstCfl when that is not zero */
  int viscLimiter;                     /* VISCOUS_LIMITER_FLOW: the mean gradients of T, u, v of the interior viscous flux are formed from limited gradients */
} rxo_opts;

rxo_ctx* rxo_create(const rxo_mesh*, const rxo_chem*, const rxo_opts*);
void rxo_destroy(rxo_ctx*);

/* state in/out (AoS, SU2 numbering) */
void rxo_set_solution(rxo_ctx*, const double* U);
void rxo_set_temperature(rxo_ctx*, const double* T);      /* secant start value V[T] (variable_direct_reactive.cpp:622) */
void rxo_set_turbulence(rxo_ctx*, const double* turb6);   /* [nPoint][6] = muT,k,omega,dk/dx,dk/dy,sigma_k */
void rxo_get_solution(rxo_ctx*, double* U);
void rxo_get_primitive(rxo_ctx*, double* V, double* aux6, double* grad);  /* aux = Cp,gamma,mu,kappa,vort,strain */
void rxo_get_dt(rxo_ctx*, double* dt3);                   /* dt, lambda_inv, lambda_visc */
void rxo_get_residual(rxo_ctx*, double* R);
void rxo_get_limiter(rxo_ctx*, double* lim4);            /* [nPoint][4]: limiter of T, u, v, P */

/* stages, same order as CMultiGridIntegration::MultiGrid_Cycle (integration_time.cpp:143-181) */
int64_t rxo_preprocess(rxo_ctx*, int64_t extIter, int output);
void rxo_set_old_solution(rxo_ctx*);
void rxo_time_step(rxo_ctx*);
/* dual time stepping: the stored time levels U^n, U^(n-1), the pseudo-time iteration index SetTime_Step receives (the physical time step
   follows UNST_CFL_NUMBER on iteration 0), the source term of SetResidual_DualTime and CIntegration::SetDualTime_Solver's level shift */
void rxo_set_time_levels(rxo_ctx*, const double* Un, const double* Un1);
void rxo_get_time_levels(rxo_ctx*, double* Un, double* Un1);
void rxo_set_inner_iteration(rxo_ctx*, int64_t intIter);
double rxo_get_delta_unst_time(rxo_ctx*);
void rxo_dual_time_residual(rxo_ctx*);
void rxo_shift_time_levels(rxo_ctx*);
/* the SST solver's time levels (k, omega)^n, (k, omega)^(n-1), [nPoint][2] each: read by rxo_turb_iterate in a dual-time context */
void rxo_turb_set_time_levels(rxo_ctx*, const double* kwN, const double* kwN1);
void rxo_upwind_residual(rxo_ctx*);
void rxo_viscous_residual(rxo_ctx*);
void rxo_source_residual(rxo_ctx*);
void rxo_boundary_conditions(rxo_ctx*);
void rxo_explicit_update(rxo_ctx*, double rkAlpha, double* resRms);

/* stand-alone leaf entry points (microbenchmarks / unit parity) */
void rxo_source_points(const rxo_chem*, int rans, double cMu, double pasrLb, int64_t n, const double* T, const double* rho,
                       const double* Y /*[n][Ns]*/, const double* omegaTurb, const double* vol, double* res /*[n][Ns]*/);
int rxo_bicgstab(int n, const double* A /*col-major*/, const double* rhs, double* x, int* iters);
/* Eigen::ColPivHouseholderQR<MatrixXd>(A).solve(b) in Eigen 3.3.7's own summation order (n <= 9) */
void rxo_colpiv_qr_solve(int n, const double* A /*col-major*/, const double* b, double* x);
double rxo_spline(const rxo_chem*, int prop, int species, double T);

/* SST transport step restated (CTurbSSTSolver's SingleGrid_Iteration on the context's flow state; see rx_oracle.cpp): kw[nPoint][2],
 * blend3[nPoint][3] = F1, F2, CDkw of the previous Postprocessing, dist[nPoint]; outputs (any may be NULL): res2 = LinSysRes before the solve,
 * dx2 = LinSysSol, rms2[2], linIters, post12[nPoint][12] = the state after Postprocessing. */
int rxo_turb_iterate(rxo_ctx*, const double* kw, const double* blend3, const double* dist, double sigmaK, double kineInf, double omegaInf,
                     double cflRed, double relax, int linIter, double linTol, double* res2, double* dx2, double* rms2, int* linIters, double* post12);

#ifdef __cplusplus
}
#endif
#endif
