// rx_cuda.cu — host side of librx_cuda.so: context, mesh flattening, and the C ABI of include/rx_cuda.h.
// Every entry point enqueues kernels on the context's stream; there is NO CPU fallback anywhere in this library.
#include "../../include/rx_cuda.h"
#include "rx_kernels2.cuh"
#include "rx_visc_pair.cuh"
#include "rx_order.cuh"
#include "rx_turb.cuh"
#include "rx_turb_step.cuh"

#include <cuda_runtime.h>
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <string>
#include <vector>

#include "rx_comm.cuh"

using namespace rx;

static constexpr size_t SM_PAIR = (size_t)VP_SMW * VP_BS * sizeof(double);
static constexpr size_t SM_PAIR_RANS = (size_t)VP_SMW_RANS * VP_BS * sizeof(double);
static constexpr size_t SM_VISC = (size_t)(9 * 9 + 3 * 9) * EDGE_BS * sizeof(double);   // thread-private 9x9 matrix + three 9-vectors in shared memory

static thread_local std::string g_create_error;
static inline bool is_supersonic(int kind) { return kind == RX_BC_SUPERSONIC_INLET || kind == RX_BC_SUPERSONIC_OUTLET; }
static inline bool is_inlet(int kind) { return kind == RX_BC_INLET_TEMPERATURE_IMPOSE || kind == RX_BC_INLET_TOTAL_CONDITIONS || kind == RX_BC_INLET_MASS_FLOW; }

struct rx_ctx {
  int device = 0, Ns = 0, nVar = 0, nPrim = 0, nGrad = 0, nMarker = 0;
  cudaStream_t stream = 0;
  std::string err;
  Dev d{};
  Opts o{};
  ChemC<9> cc9{}; ChemC<5> cc5{}; ChemC<3> cc3{};   // the one matching Ns is filled (RX_NS below)
  std::vector<MarkerDev> markers;
  std::vector<void*> allocs;
  // SST transport step (rx_turb_step.cuh): work arrays, Krylov vectors, level lists of the LU-SGS sweeps
  TurbW tw{};
  bool turbStepReady = false;
  SweepList swF{}, swB{};
  SweepData sdF{}, sdB{};
  int* posF = nullptr;              // point -> position in the forward sweep's level order
  std::vector<double*> kw, kz;      // FGMRES w[0..m], z[0..m]
  double *tPartial = nullptr, *tScal = nullptr;
  int tBlocks = 0, tDotBlocks = 0;
  std::vector<int> hIncPtr, hNbOfSlot;   // host copy of the incident CSR (neighbour of every slot) for the level lists
  bool snapPending = false;         // outputPending at the time of the snapshot
  bool outputPending = false;       // rx_iterate ran only the cons->prim half of the Output Preprocessing pass; transport / gradients on demand (ensure_output)
  Comm* comm = nullptr;             // in-library halo exchange over NCCL (rx_comm_init); nullptr: single partition or caller-side transport
  std::vector<void*> snap;          // rx_state_snapshot: device copies of the evolving state, same order as snapshot_fields()
  bool allocFailed = false;         // any cudaMalloc / cudaMemset / cudaMemcpy of dalloc / dupload failed (checked at the end of rx_create)
  double *stage = nullptr, *partial = nullptr, *rms = nullptr, *sumsq = nullptr, *pmax = nullptr, *resMax = nullptr;
  int *pidx = nullptr, *resMaxPoint = nullptr;
  bool haveResMax = false;
  size_t stageBytes = 0;
  int updBlocks = 0;
  long long launches = 0;
  cudaStream_t bcStream = nullptr;   // second stream for boundary kernels of markers that share no vertex
  cudaEvent_t evFork = nullptr, evJoin = nullptr;
  bool bcConcurrent = false;
  bool orderEdges = true;           // viscous kernel processes the edges grouped by their last iteration count (rx_order.cuh); default: laminar on, RANS off; env RX_VISC_ORDER=0/1 forces
  int* eOrderBuf = nullptr; int nOrdTiles = 0;
  double* eddyFlow = nullptr;       // RANS: eddy viscosity as the flow variables hold it (copied at every rx_preprocess, CReactiveNSVariable::SetPrimVar :1193) — what COutput writes
  double* tpost = nullptr;          // SST Postprocessing fields (rx_turb.cuh), allocated for RANS contexts
  bool haveWallDist = false; double sigmaK = 0.85;
  int sweepThreads = 1024;           // block size of the LU-SGS sweep kernels (test hook RX_SWEEP_THREADS: a small block sends the levels through the wide-level loop)
  double* dtPartial = nullptr; int dtBlocks = 0;   // TIME_STEPPING: block minima of the local time steps + the global one
  double* Un = nullptr; double* Un1 = nullptr;     // dual time stepping: U^n, U^(n-1), laid out like U
  double* dtND = nullptr;                          // dual time stepping: the physical time step, on the device (rx_time_step may recompute it)
  bool timeLevelsSet = false; long long intIter = 0;
  double* tkwN = nullptr; double* tkwN1 = nullptr; bool turbLevelsSet = false;   // ... and the SST solver's (k, omega)^n, (k, omega)^(n-1) (RANS contexts)
};

extern "C" {
static int halo_exchange_fork(rx_ctx* c, int what);
static int ensure_time_levels(rx_ctx* c, const char* who);
static int comm_join(rx_ctx* c);
static int ensure_output(rx_ctx* c);
static int halo_exchange_field(rx_ctx* c, double* field, int nv, bool vortStrain, cudaStream_t s);
static int comm_allreduce_sum(rx_ctx* c, double* dev, int n);
}

#define RX_CK(call)                                                                                   \
  do {                                                                                                \
    cudaError_t e_ = (call);                                                                          \
    if (e_ != cudaSuccess) {                                                                          \
      c->err = std::string(#call) + ": " + cudaGetErrorString(e_);                                    \
      return RX_ERR_CUDA;                                                                             \
    }                                                                                                 \
  } while (0)

// Kernels are instantiated for the species counts of the mixtures the reference ships: 9 (jet reactor, TURBOLENT_COMBUSTION), 5 and 3
// (TURBOLENT_FLAT_PLATE: test.txt, test_air.txt).  A launch site names its kernel once, with NS and cc bound by the dispatch.
template <int N> struct ChemOf;
template <> struct ChemOf<9> { static ChemC<9>& get(rx_ctx* c) { return c->cc9; } };
template <> struct ChemOf<5> { static ChemC<5>& get(rx_ctx* c) { return c->cc5; } };
template <> struct ChemOf<3> { static ChemC<3>& get(rx_ctx* c) { return c->cc3; } };
#define RX_NS_CASE(N, ...) case N: { constexpr int NS = N; const ChemC<N>& cc = ChemOf<N>::get(c); (void)cc; (void)NS; __VA_ARGS__; } break;
#define RX_NS(...) do { switch (c->Ns) { RX_NS_CASE(9, __VA_ARGS__) RX_NS_CASE(5, __VA_ARGS__) RX_NS_CASE(3, __VA_ARGS__) default: break; } } while (0)

namespace {
template <class T>
T* dalloc(rx_ctx* c, size_t n) {
  void* p = nullptr;
  if (cudaMalloc(&p, std::max<size_t>(n, 1) * sizeof(T)) != cudaSuccess) { c->allocFailed = true; return nullptr; }
  c->allocs.push_back(p);
  if (cudaMemset(p, 0, std::max<size_t>(n, 1) * sizeof(T)) != cudaSuccess) { c->allocFailed = true; return nullptr; }
  return (T*)p;
}
template <class T>
T* dupload(rx_ctx* c, const std::vector<T>& h) {
  T* p = dalloc<T>(c, h.size());
  if (p && !h.empty() && cudaMemcpy(p, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice) != cudaSuccess) { c->allocFailed = true; return nullptr; }
  if (!p) c->allocFailed = true;
  return p;
}
inline unsigned nblk(long long n, int bs) { return (unsigned)((n + bs - 1) / bs); }

// host-side constants of the mechanism; libm calls are the ones the reference makes (reacting_model_library.cpp:634-696,751-766)
template <int NS>
int fill_chem_ns(rx_ctx* c, const rx_chem* ch) {
  if (ch->Nr > NRMAX) return RX_ERR_UNSUPPORTED;
  ChemC<NS>& cc = ChemOf<NS>::get(c);
  cc.Nr = ch->Nr; cc.nKnots = ch->nKnots;
  std::vector<double> x(ch->x, ch->x + ch->nKnots);
  cc.x = dupload(c, x);
  std::vector<Knot> tab((size_t)5 * ch->nKnots * NS);
  for (int pr = 0; pr < 5; ++pr)
    for (int s = 0; s < NS; ++s)
      for (int k = 0; k < ch->nKnots; ++k) {
        const size_t src = ((size_t)pr * NS + s) * ch->nKnots + k;
        tab[((size_t)pr * ch->nKnots + k) * NS + s] = Knot{ch->y[src], ch->y2[src]};
      }
  cc.tab = dupload(c, tab);
  if (!cc.x || !cc.tab) return RX_ERR_CUDA;
  cc.Mtot = 0.0;
  for (int s = 0; s < NS; ++s) { cc.M[s] = ch->mMasses[s]; cc.Ri[s] = R_UNGAS / cc.M[s]; cc.Mtot = cc.Mtot + cc.M[s]; }
  for (int i = 0; i < NS; ++i)
    for (int j = 0; j < NS; ++j) {
      cc.wPow[i][j] = std::pow(cc.M[j] / cc.M[i], 0.25);
      cc.wSq[i][j] = std::sqrt(8.0 * (1.0 + cc.M[i] / cc.M[j]));
    }
  {   // reciprocals of the Wilke denominators exactly as the kernels would refine them
    double* dv = dupload(c, std::vector<double>(&cc.wSq[0][0], &cc.wSq[0][0] + NS * NS));
    if (!dv) return RX_ERR_CUDA;
    k_fill_recip<<<1, NS * NS>>>(dv, NS * NS);
    if (cudaMemcpy(&cc.rwSq[0][0], dv, sizeof(double) * NS * NS, cudaMemcpyDeviceToHost) != cudaSuccess) return RX_ERR_CUDA;
    double* dm = dupload(c, std::vector<double>(&cc.M[0], &cc.M[0] + NS));
    if (!dm) return RX_ERR_CUDA;
    k_fill_recip<<<1, NS>>>(dm, NS);
    if (cudaMemcpy(&cc.rM[0], dm, sizeof(double) * NS, cudaMemcpyDeviceToHost) != cudaSuccess) return RX_ERR_CUDA;
  }
  for (int i = 0; i < NS; ++i) {
    const double mi = cc.M[i], dvi = std::cbrt(ch->diffVol[i]);
    for (int j = i; j < NS; ++j) {
      cc.Mij[i][j] = cc.Mij[j][i] = std::sqrt((mi * cc.M[j]) / (mi + cc.M[j]));
      cc.dvs[i][j] = cc.dvs[j][i] = dvi + std::cbrt(ch->diffVol[j]);
    }
  }
  for (int r = 0; r < NRMAX; ++r) {
    const bool on = r < ch->Nr;
    cc.As[r] = on ? ch->As[r] : 0; cc.Betas[r] = on ? ch->Betas[r] : 0; cc.Ta[r] = on ? ch->Ta[r] : 0;
    cc.AsB[r] = on ? ch->AsB[r] : 0; cc.BetasB[r] = on ? ch->BetasB[r] : 0; cc.TaB[r] = on ? ch->TaB[r] : 0;
    cc.rev[r] = on ? (ch->reversible[r] != 0.0) : 0; cc.avb[r] = on ? (ch->availBack[r] != 0.0) : 0;
    for (int s = 0; s < NS; ++s) {
      cc.nuR[r][s] = on ? ch->nuR[r * NS + s] : 0; cc.nuP[r][s] = on ? ch->nuP[r * NS + s] : 0;
      cc.expR[r][s] = on ? ch->expR[r * NS + s] : 0; cc.expP[r][s] = on ? ch->expP[r * NS + s] : 0;
    }
  }
  return RX_OK;
}

int fill_chem(rx_ctx* c, const rx_chem* ch) {
  switch (ch->Ns) {
    case 9: return fill_chem_ns<9>(c, ch);
    case 5: return fill_chem_ns<5>(c, ch);
    case 3: return fill_chem_ns<3>(c, ch);
    default: return RX_ERR_UNSUPPORTED;
  }
}

int ensure_stage(rx_ctx* c, size_t bytes) {
  if (bytes <= c->stageBytes) return RX_OK;
  if (c->stage) cudaFree(c->stage);
  c->stage = nullptr; c->stageBytes = 0;
  RX_CK(cudaMalloc((void**)&c->stage, bytes));
  c->stageBytes = bytes;
  return RX_OK;
}
int upload_aos(rx_ctx* c, const double* h, double* soa, int nv) {
  const long long n = c->d.nPoint;
  int rc = ensure_stage(c, (size_t)n * nv * sizeof(double));
  if (rc) return rc;
  RX_CK(cudaMemcpyAsync(c->stage, h, (size_t)n * nv * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  k_aos_to_soa<<<nblk(n * nv, 256), 256, 0, c->stream>>>(c->stage, soa, n, nv, c->d.np);
  c->launches++;
  RX_CK(cudaGetLastError());
  RX_CK(cudaStreamSynchronize(c->stream));   // the caller may reuse its host buffer on return
  return RX_OK;
}
int download_aos(rx_ctx* c, const double* soa, double* h, int nv) {
  const long long n = c->d.nPoint;
  int rc = ensure_stage(c, (size_t)n * nv * sizeof(double));
  if (rc) return rc;
  k_soa_to_aos<<<nblk(n * nv, 256), 256, 0, c->stream>>>(soa, c->stage, n, nv, c->d.np);
  c->launches++;
  RX_CK(cudaGetLastError());
  RX_CK(cudaMemcpyAsync(h, c->stage, (size_t)n * nv * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  RX_CK(cudaStreamSynchronize(c->stream));
  return RX_OK;
}
}  // namespace

// viscous edge flux: nine species (the mechanism every BASELINE configuration uses) go through the two-lanes-per-edge kernel, whose row
// split and storage are laid out for nine rows; the other instantiated species counts take the thread-per-edge form of the same
// arithmetic (viscous_flux<NS, EDGE = true>, rx_flux.cuh), with the thread-private Stefan–Maxwell scratch in shared memory
template <int NS>
static void launch_edge_visc(rx_ctx* c, const ChemC<NS>& cc, cudaStream_t s) {
  const Dev& d = c->d;
  if constexpr (NS == 9) {
    if (c->o.viscLimiter) {   // VISCOUS_LIMITER_FLOW: the same kernel with the limited mean gradients compiled in
      if (c->o.rans) k_edge_visc_pair<9, true, true><<<nblk(2 * d.nEdge, VP_BS), VP_BS, SM_PAIR_RANS, s>>>(cc, c->o, d);
      else k_edge_visc_pair<9, false, true><<<nblk(2 * d.nEdge, VP_BS), VP_BS, SM_PAIR, s>>>(cc, c->o, d);
    }
    else if (c->o.rans) k_edge_visc_pair<9, true><<<nblk(2 * d.nEdge, VP_BS), VP_BS, SM_PAIR_RANS, s>>>(cc, c->o, d);
    else k_edge_visc_pair<9, false><<<nblk(2 * d.nEdge, VP_BS), VP_BS, SM_PAIR, s>>>(cc, c->o, d);
  } else {
    if (c->o.rans) k_edge_visc_generic<NS, true><<<nblk(d.nEdge, EDGE_BS), EDGE_BS, SM_VISC, s>>>(cc, c->o, d);
    else k_edge_visc_generic<NS, false><<<nblk(d.nEdge, EDGE_BS), EDGE_BS, SM_VISC, s>>>(cc, c->o, d);
  }
}
template <int NS>
static cudaError_t set_kernel_attributes() {
  const int smv = (int)SM_VISC;
  const cudaFuncAttribute at = cudaFuncAttributeMaxDynamicSharedMemorySize;
  cudaError_t e = cudaSuccess;
  if constexpr (NS == 9) {
    e = cudaFuncSetAttribute(k_edge_visc_pair<9, true>, at, (int)SM_PAIR_RANS);
    if (e == cudaSuccess && SM_PAIR > 48 * 1024) e = cudaFuncSetAttribute(k_edge_visc_pair<9, false>, at, (int)SM_PAIR);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_edge_visc_pair<9, true, true>, at, (int)SM_PAIR_RANS);
    if (e == cudaSuccess && SM_PAIR > 48 * 1024) e = cudaFuncSetAttribute(k_edge_visc_pair<9, false, true>, at, (int)SM_PAIR);
  } else {
    e = cudaFuncSetAttribute(k_edge_visc_generic<NS, true>, at, smv);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_edge_visc_generic<NS, false>, at, smv);
  }
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_bc_inlet<NS>, at, smv);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_bc_outlet<NS>, at, smv);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_bc_supersonic<NS>, at, smv);
  return e;
}

// launch of the edge-flux kernels: the AUSM flux from its own bandwidth-bound kernel (or the MUSCL kernel), the viscous flux
// from the two-lanes-per-edge kernel (rx_visc_pair.cuh; the RANS instantiation carries the SST closure).
template <bool CONV, bool VISC>
static void launch_edge_flux(rx_ctx* c, cudaStream_t s) {
  const Dev& d = c->d;
  if (CONV) {
    if (c->o.secondOrder) RX_NS(k_edge_conv2<NS><<<nblk(d.nEdge, 64), 64, 0, s>>>(cc, c->o, d));
    else RX_NS(k_edge_flux<NS><<<nblk(d.nEdge, 128), 128, 0, s>>>(cc, c->o, d));
    c->launches++;
  }
  if (VISC) {
    RX_NS(launch_edge_visc<NS>(c, cc, s));
    c->launches++;
  }
}

// conv + visc accumulation and the source term in one kernel
static void launch_accumulate_visc(rx_ctx* c, cudaStream_t s) {
  const Dev& d = c->d;
  RX_NS(k_accumulate<NS, 3><<<nblk(d.nPoint, 128), 128, 0, s>>>(cc, c->o, d));
  c->launches++;
}

extern "C" {

int rx_supported_species(int Ns) { return Ns == 9 || Ns == 5 || Ns == 3; }

const char* rx_last_error(rx_ctx* c) { return c ? c->err.c_str() : g_create_error.c_str(); }

int rx_create(rx_ctx** out, const rx_mesh* m, const rx_chem* ch, const rx_opts* op) {
  if (!out || !m || !ch || !op) { g_create_error = "rx_create: null argument"; return RX_ERR_ARG; }
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    g_create_error = "rx_create: no CUDA device (this library has no CPU path)";
    return RX_ERR_NO_DEVICE;
  }
  if (!rx_supported_species(ch->Ns)) { g_create_error = "rx_create: kernels are instantiated for Ns = 9, 5 and 3 (the mixtures the reference ships)"; return RX_ERR_UNSUPPORTED; }
  if (ch->Nr > NRMAX) { g_create_error = "rx_create: more reactions than NRMAX"; return RX_ERR_UNSUPPORTED; }
  if (m->nPoint >= (1LL << 31) || m->nEdge >= (1LL << 31)) { g_create_error = "rx_create: mesh exceeds int32 indices"; return RX_ERR_UNSUPPORTED; }
  rx_ctx* c = new rx_ctx();
  c->device = op->device;
  {   // measured on the bench mesh: laminar 6.08 -> 5.96 ms per launch, RANS 14.51 -> 14.75 ms (its closure gathers more per point)
    const char* ev = getenv("RX_VISC_ORDER");
    c->orderEdges = ev && (ev[0] == '0' || ev[0] == '1') ? ev[0] == '1' : !op->rans;
  }
  if (cudaSetDevice(c->device) != cudaSuccess) { g_create_error = "rx_create: cudaSetDevice failed"; delete c; return RX_ERR_CUDA; }
  const int Ns = ch->Ns;
  c->Ns = Ns; c->nVar = Ns + 4; c->nPrim = Ns + 7; c->nGrad = Ns + 4;
  {   // the thread-private Stefan–Maxwell scratch exceeds the default 48 KB of dynamic shared memory per block
    cudaError_t e = cudaSuccess;
    RX_NS(e = set_kernel_attributes<NS>());
    if (e != cudaSuccess) { g_create_error = std::string("rx_create: cudaFuncSetAttribute: ") + cudaGetErrorString(e); delete c; return RX_ERR_CUDA; }
  }
  if (Ns != 9) c->orderEdges = false;   // the processing order by iteration count belongs to the pair kernel
  Opts& o = c->o;
  o.viscous = op->viscous; o.rans = op->rans; o.leastSquares = op->leastSquares;
  o.mInfty = op->mInfty; o.cfl = op->cfl; o.maxDt = op->maxDt; o.Tmin = op->Tmin; o.Tmax = op->Tmax;
  o.clipT = op->clipT; o.ignition = op->ignition; o.ignitionIter = op->ignitionIter; o.ignitionT = op->ignitionT;
  o.fuelIndex = op->fuelIndex; o.oxidizerIndex = op->oxidizerIndex;
  o.prandtlLam = op->prandtlLam; o.prandtlTurb = op->prandtlTurb; o.lewisTurb = op->lewisTurb; o.cMu = op->cMu;
  o.pasrLb = op->pasrLb; o.tkeInf = op->tkeInf;
  o.timeStepping = op->timeStepping; o.unstCfl = op->unstCfl; o.deltaUnstTime = op->deltaUnstTime;
  o.dualTime = op->dualTime; o.deltaUnstTimeND = op->deltaUnstTimeND;
  o.viscLimiter = op->viscLimiter ? 1 : 0;
  if (o.viscLimiter && !(op->viscous && op->secondOrder && op->limiter)) { g_create_error = "rx_create: viscLimiter (VISCOUS_LIMITER_FLOW) needs a viscous context with secondOrder and limiter set: the limiters are those of SPATIAL_ORDER_FLOW= 2ND_ORDER_LIMITER"; delete c; return RX_ERR_ARG; }
  if (o.dualTime < 0 || o.dualTime > 2 || (o.dualTime && o.timeStepping)) { g_create_error = "rx_create: dualTime must be 0, 1 or 2 and excludes timeStepping"; delete c; return RX_ERR_ARG; }
  if (o.dualTime && !(o.deltaUnstTimeND > 0.0) && o.unstCfl == 0.0) { g_create_error = "rx_create: dual time stepping needs a physical time step (deltaUnstTimeND) or an unsteady CFL number"; delete c; return RX_ERR_ARG; }
  o.secondOrder = op->secondOrder; o.limiter = op->secondOrder ? op->limiter : 0; o.limCoeff = op->limCoeff; o.refElemLength = op->refElemLength;
  int rc = fill_chem(c, ch);
  if (rc) { g_create_error = "rx_create: chemistry upload failed"; rx_destroy(c); return rc; }

  Dev& d = c->d;
  const long long nP = m->nPoint, nE = m->nEdge;
  d.nPoint = nP; d.nPointDomain = m->nPointDomain; d.nEdge = nE;
  d.np = (nP + 15) / 16 * 16; d.ne = (nE + 15) / 16 * 16;
  std::vector<int> ei(nE), ej(nE);
  std::vector<double> enx(nE), eny(nE), x(nP), y(nP), vol(m->volume, m->volume + nP);
  for (long long e = 0; e < nE; ++e) {
    ei[e] = (int)m->edgeNodes[2 * e]; ej[e] = (int)m->edgeNodes[2 * e + 1];
    enx[e] = m->edgeNormal[2 * e]; eny[e] = m->edgeNormal[2 * e + 1];
  }
  for (long long p = 0; p < nP; ++p) { x[p] = m->coord[2 * p]; y[p] = m->coord[2 * p + 1]; }
  std::vector<int> nbPtr(nP + 1), nbIdx(m->nbPtr[nP]);
  for (long long p = 0; p <= nP; ++p) nbPtr[p] = (int)m->nbPtr[p];
  for (size_t k = 0; k < nbIdx.size(); ++k) nbIdx[k] = (int)m->nbPoint[k];
  // incident-edge CSR in ascending edge order (counting sort keeps the order)
  std::vector<int> incPtr(nP + 1, 0), incEdge(2 * nE);
  for (long long e = 0; e < nE; ++e) { incPtr[ei[e] + 1]++; incPtr[ej[e] + 1]++; }
  for (long long p = 0; p < nP; ++p) incPtr[p + 1] += incPtr[p];
  {
    std::vector<int> fill(incPtr.begin(), incPtr.end() - 1);
    for (long long e = 0; e < nE; ++e) { incEdge[fill[ei[e]]++] = (int)e; incEdge[fill[ej[e]]++] = ~(int)e; }
  }
  if (op->rans) {   // neighbour of every incident slot, for the level lists of the turbulence step's LU-SGS sweeps
    c->hIncPtr = incPtr;
    c->hNbOfSlot.resize(incEdge.size());
    for (size_t k = 0; k < incEdge.size(); ++k) { const int code = incEdge[k]; c->hNbOfSlot[k] = code >= 0 ? ej[code] : ei[~code]; }
  }
  d.ei = dupload(c, ei); d.ej = dupload(c, ej); d.enx = dupload(c, enx); d.eny = dupload(c, eny);
  d.x = dupload(c, x); d.y = dupload(c, y); d.vol = dupload(c, vol);
  d.nbPtr = dupload(c, nbPtr); d.nbIdx = dupload(c, nbIdx); d.incPtr = dupload(c, incPtr); d.incEdge = dupload(c, incEdge);
  const size_t np = d.np, ne = d.ne;
  d.U = dalloc<double>(c, np * c->nVar); d.Uold = dalloc<double>(c, np * c->nVar); d.R = dalloc<double>(c, np * c->nVar);
  d.V = dalloc<double>(c, np * c->nPrim); d.X = dalloc<double>(c, np * Ns);
  d.Cp = dalloc<double>(c, np); d.gam = dalloc<double>(c, np); d.mu = dalloc<double>(c, np); d.kap = dalloc<double>(c, np);
  d.lim = dalloc<double>(c, np * 4);
  d.rD = o.viscous ? dalloc<double>(c, np * (size_t)(Ns * (Ns + 1) / 2)) : nullptr;
  if (o.dualTime) { c->Un = dalloc<double>(c, np * c->nVar); c->Un1 = dalloc<double>(c, np * c->nVar); c->dtND = dalloc<double>(c, 1);
    if (c->dtND && cudaMemcpy(c->dtND, &o.deltaUnstTimeND, sizeof(double), cudaMemcpyHostToDevice) != cudaSuccess) c->allocFailed = true;
    if (o.rans) { c->tkwN = dalloc<double>(c, np * 2); c->tkwN1 = dalloc<double>(c, np * 2); } }
  d.grad = dalloc<double>(c, np * c->nGrad * 2); d.vort = dalloc<double>(c, np); d.strain = dalloc<double>(c, np);
  d.turb = dalloc<double>(c, np * 6); c->tpost = o.rans ? dalloc<double>(c, np * TP_N) : nullptr; c->eddyFlow = o.rans ? dalloc<double>(c, np) : nullptr; d.dt = dalloc<double>(c, np); d.lamInv = dalloc<double>(c, np); d.lamVisc = dalloc<double>(c, np);
  d.fluxC = dalloc<double>(c, ne * c->nVar);
  d.fluxV = o.viscous ? dalloc<double>(c, ne * c->nVar) : nullptr;
  d.eOrder = nullptr;
  d.eKey = o.viscous ? dalloc<unsigned char>(c, ne) : nullptr;
  if (o.viscous && c->orderEdges) {
    c->nOrdTiles = (int)((d.nEdge + ORD_TILE - 1) / ORD_TILE);
    c->eOrderBuf = dalloc<int>(c, ne);
  }
  d.nonPhys = dalloc<int>(c, np); d.counters = dalloc<unsigned long long>(c, 8);   // [0] non-physical points of the last preprocess; sticky until read: [1..3] NaN in the conv / visc / source residual, [4] temperature solves that did not converge
  c->updBlocks = (int)nblk(d.nPointDomain, 128);
  c->partial = dalloc<double>(c, (size_t)c->updBlocks * c->nVar);
  c->rms = dalloc<double>(c, c->nVar); c->sumsq = dalloc<double>(c, c->nVar);
  c->pmax = dalloc<double>(c, (size_t)c->updBlocks * c->nVar); c->pidx = dalloc<int>(c, (size_t)c->updBlocks * c->nVar);
  c->resMax = dalloc<double>(c, c->nVar); c->resMaxPoint = dalloc<int>(c, c->nVar);
  if (c->allocFailed || !d.counters || !c->sumsq || !d.fluxC || !d.grad || (o.viscous && !d.fluxV)) {
    g_create_error = "rx_create: device allocation failed"; rx_destroy(c); return RX_ERR_CUDA;
  }
  c->nMarker = m->nMarker;
  for (int k = 0; k < m->nMarker; ++k) {
    const rx_marker& s = m->markers[k];
    if (s.kind < RX_BC_NONE || s.kind > RX_BC_SUPERSONIC_OUTLET) { g_create_error = "rx_create: unknown marker kind"; rx_destroy(c); return RX_ERR_ARG; }
    if (is_supersonic(s.kind) && o.rans) {   // the reference's two supersonic conditions never hand the turbulence inputs to the viscous numerics (:3150-3190)
      g_create_error = "rx_create: supersonic inlet / outlet markers are supported for laminar and inviscid contexts only"; rx_destroy(c); return RX_ERR_UNSUPPORTED;
    }
    MarkerDev md{};
    md.kind = s.kind; md.nVertex = s.nVertex;
    std::vector<int> node(s.nVertex), nn(s.nVertex);
    std::vector<double> nx(s.nVertex), ny(s.nVertex);
    for (int v = 0; v < s.nVertex; ++v) {
      node[v] = (int)s.node[v]; nn[v] = (int)s.normalNeighbor[v]; nx[v] = s.normal[2 * v]; ny[v] = s.normal[2 * v + 1];
    }
    md.node = dupload(c, node); md.nn = dupload(c, nn); md.nx = dupload(c, nx); md.ny = dupload(c, ny);
    md.ghost = dalloc<double>(c, (size_t)std::max(s.nVertex, 1) * (7 + Ns));
    md.T = s.T; md.Vmag = s.Vmag; md.dir0 = s.dir[0]; md.dir1 = s.dir[1]; md.P = s.P; md.Twall = s.Twall; md.heatFlux = s.heatFlux;
    for (int q = 0; q < 16; ++q) md.Ys[q] = ((is_inlet(s.kind) || s.kind == RX_BC_SUPERSONIC_INLET) && s.Ys && q < Ns) ? s.Ys[q] : 0.0;
    c->markers.push_back(md);
  }
  {   // may the first-group boundary kernels overlap?  Only if no point carries two of those markers.
    std::vector<char> seen(nP, 0);
    int nFirst = 0; bool disjoint = true;
    for (int k = 0; k < m->nMarker; ++k) {
      const rx_marker& s = m->markers[k];
      if (!(is_inlet(s.kind) || is_supersonic(s.kind) || s.kind == RX_BC_OUTLET || s.kind == RX_BC_EULER_WALL) || !s.nVertex) continue;
      ++nFirst;
      for (int v = 0; v < s.nVertex; ++v) { if (seen[s.node[v]]) disjoint = false; seen[s.node[v]] = 1; }
    }
    const char* ev = getenv("RX_BC_SERIAL");
    if (nFirst >= 2 && disjoint && !(ev && ev[0] == '1')) {
      if (cudaStreamCreateWithFlags(&c->bcStream, cudaStreamNonBlocking) == cudaSuccess &&
          cudaEventCreateWithFlags(&c->evFork, cudaEventDisableTiming) == cudaSuccess &&
          cudaEventCreateWithFlags(&c->evJoin, cudaEventDisableTiming) == cudaSuccess) c->bcConcurrent = true;
    }
  }
  if (cudaDeviceSynchronize() != cudaSuccess) { g_create_error = "rx_create: upload failed"; rx_destroy(c); return RX_ERR_CUDA; }
  *out = c;
  return RX_OK;
}

void rx_destroy(rx_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  for (void* p : c->allocs) cudaFree(p);
  if (c->stage) cudaFree(c->stage);
  if (c->evFork) cudaEventDestroy(c->evFork);
  if (c->evJoin) cudaEventDestroy(c->evJoin);
  if (c->bcStream) cudaStreamDestroy(c->bcStream);
  if (c->comm) {
    std::string why;
    NcclApi* a = nccl_api(why);
    if (a && c->comm->comm) a->CommDestroy(c->comm->comm);
    if (c->comm->cs) cudaStreamDestroy(c->comm->cs);
    if (c->comm->evReady) cudaEventDestroy(c->comm->evReady);
    if (c->comm->evDone) cudaEventDestroy(c->comm->evDone);
    delete c->comm;
  }
  delete c;
}

// The evolving state of a context: what one iteration reads from the previous one.  Saved / restored device-to-device so that a
// measurement can run the SAME iteration from the SAME state any number of times (bench.py: the timed value must not depend on
// how many steps are timed).  No reference counterpart (the reference has no such need).
static std::vector<std::pair<void*, size_t>> snapshot_fields(rx_ctx* c) {
  const Dev& d = c->d;
  const size_t np = d.np, ne = d.ne, w = sizeof(double);
  std::vector<std::pair<void*, size_t>> f = {{d.U, np * c->nVar * w}, {d.Uold, np * c->nVar * w}, {d.V, np * c->nPrim * w}, {d.turb, np * 6 * w}};
  if (d.mu) f.push_back({d.mu, np * w});   // while the Output pass is pending mu holds the temperature k_primitive parked for k_transport
  if (d.eKey) f.push_back({d.eKey, ne});
  if (c->eOrderBuf) f.push_back({c->eOrderBuf, ne * sizeof(int)});
  if (c->tpost) f.push_back({c->tpost, np * TP_N * w});
  if (c->Un) { f.push_back({c->Un, np * c->nVar * w}); f.push_back({c->Un1, np * c->nVar * w}); f.push_back({c->dtND, w}); }
  if (c->tkwN) { f.push_back({c->tkwN, np * 2 * w}); f.push_back({c->tkwN1, np * 2 * w}); }
  return f;
}
int rx_state_snapshot(rx_ctx* c, int restore) {
  const auto f = snapshot_fields(c);
  if (!restore && c->snap.empty()) {
    for (const auto& q : f) {
      void* p = nullptr;
      RX_CK(cudaMalloc(&p, q.second));
      c->allocs.push_back(p); c->snap.push_back(p);
    }
  }
  if (c->snap.size() != f.size()) { c->err = "rx_state_snapshot: restore before any save"; return RX_ERR_ARG; }
  for (size_t k = 0; k < f.size(); ++k)
    RX_CK(cudaMemcpyAsync(restore ? f[k].first : c->snap[k], restore ? c->snap[k] : f[k].first, f[k].second, cudaMemcpyDeviceToDevice, c->stream));
  if (restore) c->outputPending = c->snapPending; else c->snapPending = c->outputPending;
  return RX_OK;
}

int rx_set_stream(rx_ctx* c, void* s) { c->stream = (cudaStream_t)s; return RX_OK; }
int rx_synchronize(rx_ctx* c) { RX_CK(cudaStreamSynchronize(c->stream)); return RX_OK; }
int64_t rx_launch_count(rx_ctx* c) { return c->launches; }

int rx_fp64_peak(rx_ctx* c, double* tflops) {
  int nsm = 0;
  RX_CK(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, c->device));
  double* sink = nullptr;
  RX_CK(cudaMalloc(&sink, sizeof(double)));
  const int iters = 1 << 15, blocks = nsm * 8, threads = 256;
  cudaEvent_t e0, e1;
  RX_CK(cudaEventCreate(&e0)); RX_CK(cudaEventCreate(&e1));
  float best = 1e30f;
  for (int rep = 0; rep < 4; ++rep) {   // rep 0 warms up
    RX_CK(cudaEventRecord(e0, c->stream));
    k_fp64_peak<<<blocks, threads, 0, c->stream>>>(iters, 1.0000001, sink);
    RX_CK(cudaEventRecord(e1, c->stream));
    RX_CK(cudaEventSynchronize(e1));
    float ms = 0.f;
    RX_CK(cudaEventElapsedTime(&ms, e0, e1));
    if (rep > 0 && ms < best) best = ms;
    c->launches++;
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(sink);
  if (tflops) *tflops = 2.0 * 8.0 * (double)iters * (double)blocks * threads / (best * 1e-3) / 1e12;
  return RX_OK;
}

int rx_set_solution(rx_ctx* c, const double* U) { return upload_aos(c, U, c->d.U, c->nVar); }
int rx_set_turbulence(rx_ctx* c, const double* t) { return upload_aos(c, t, c->d.turb, 6); }
// SST Postprocessing on the device (rx_turb.cuh)
int rx_set_wall_distance(rx_ctx* c, const double* dist) {
  if (!c->tpost) { c->err = "rx_set_wall_distance: not a RANS context"; return RX_ERR_ARG; }
  const int rc = upload_aos(c, dist, c->tpost + (size_t)TP_DIST * c->d.np, 1);
  if (rc == RX_OK) c->haveWallDist = true;
  return rc;
}
// wall distance of the context's own mesh, on the device: nodes of the ISOTHERMAL / HEAT_FLUX markers in marker order
// (CPhysicalGeometry::ComputeWall_Distance :8990-9060).  Single-partition meshes: a partition sees only its own wall nodes.
int rx_compute_wall_distance(rx_ctx* c) {
  if (!c->tpost) { c->err = "rx_compute_wall_distance: not a RANS context"; return RX_ERR_ARG; }
  if (c->d.nPoint != c->d.nPointDomain && !c->comm) { c->err = "rx_compute_wall_distance: a partition sees only its own wall nodes: give it the communicator (rx_comm_init), or use rx_set_wall_distance"; return RX_ERR_UNSUPPORTED; }
  const Dev& d = c->d;
  long long nWall = 0;
  for (const MarkerDev& m : c->markers) if (m.kind == 3 || m.kind == 4) nWall += m.nVertex;
  int* wn = nullptr;
  RX_CK(cudaMalloc(&wn, (size_t)std::max<long long>(nWall, 1) * sizeof(int)));
  long long off = 0;
  for (const MarkerDev& m : c->markers)
    if ((m.kind == 3 || m.kind == 4) && m.nVertex) {
      cudaMemcpyAsync(wn + off, m.node, (size_t)m.nVertex * sizeof(int), cudaMemcpyDeviceToDevice, c->stream);
      off += m.nVertex;
    }
  if (c->comm) {
    // the wall nodes of every partition (the reference gathers them over the ranks, geometry_structure.cpp:9010-9040).  Two sum-reductions
    // do the gather: the counts, each rank in its own slot, then the coordinates, each rank in its own segment of a zeroed list — a sum
    // with zeros is exact, and a minimum does not depend on the order of the list
    const int nr = c->comm->nRanks, me = c->comm->rank;
    double* cnt = nullptr;
    int rc = RX_OK;
    if (cudaMalloc(&cnt, (size_t)nr * sizeof(double)) != cudaSuccess) { cudaFree(wn); c->err = "rx_compute_wall_distance: device allocation failed"; return RX_ERR_CUDA; }
    std::vector<double> h((size_t)nr, 0.0);
    h[(size_t)me] = (double)nWall;
    cudaMemcpyAsync(cnt, h.data(), (size_t)nr * sizeof(double), cudaMemcpyHostToDevice, c->stream);
    rc = comm_allreduce_sum(c, cnt, nr);
    const char* cudaFail = "rx_compute_wall_distance: CUDA error while gathering the wall nodes over the partitions";
    if (!rc && cudaMemcpyAsync(h.data(), cnt, (size_t)nr * sizeof(double), cudaMemcpyDeviceToHost, c->stream) != cudaSuccess) { rc = RX_ERR_CUDA; c->err = cudaFail; }
    if (!rc && cudaStreamSynchronize(c->stream) != cudaSuccess) { rc = RX_ERR_CUDA; c->err = cudaFail; }
    cudaFree(cnt);
    long long total = 0, mine = 0;
    for (int r = 0; r < nr; ++r) { if (r == me) mine = total; total += (long long)h[(size_t)r]; }
    double* wall = nullptr;
    if (!rc && 2 * total > 0x7fffffffLL) { rc = RX_ERR_UNSUPPORTED; c->err = "rx_compute_wall_distance: more than 2^30 wall nodes"; }
    if (!rc && cudaMalloc(&wall, (size_t)std::max<long long>(2 * total, 1) * sizeof(double)) != cudaSuccess) { rc = RX_ERR_CUDA; c->err = cudaFail; }
    if (!rc) {
      cudaMemsetAsync(wall, 0, (size_t)std::max<long long>(2 * total, 1) * sizeof(double), c->stream);
      if (nWall) { k_wall_gather<<<nblk(nWall, 256), 256, 0, c->stream>>>(d, nWall, wn, wall + 2 * mine); c->launches++; }
      if (total) rc = comm_allreduce_sum(c, wall, (int)(2 * total));
    }
    if (!rc) {
      k_wall_distance_list<<<nblk(d.nPoint, 256), 256, 0, c->stream>>>(d, total, wall, c->tpost + (size_t)TP_DIST * d.np);
      c->launches++;
      if (cudaStreamSynchronize(c->stream) != cudaSuccess || cudaGetLastError() != cudaSuccess) { rc = RX_ERR_CUDA; c->err = cudaFail; }
    }
    cudaFree(wall); cudaFree(wn);
    if (rc) return rc;
    c->haveWallDist = true;
    return RX_OK;
  }
  k_wall_distance_ctx<<<nblk(d.nPoint, 256), 256, 0, c->stream>>>(d, nWall, wn, c->tpost + (size_t)TP_DIST * d.np);
  c->launches++;
  const cudaError_t e = cudaStreamSynchronize(c->stream);
  cudaFree(wn);
  RX_CK(e);
  RX_CK(cudaGetLastError());
  c->haveWallDist = true;
  return RX_OK;
}
int rx_turb_postprocess(rx_ctx* c, const double* kOmega, double sigmaK, int stage) {
  if (!c->tpost) { c->err = "rx_turb_postprocess: not a RANS context"; return RX_ERR_ARG; }
  if (!c->haveWallDist) { c->err = "rx_turb_postprocess: rx_set_wall_distance has not been called"; return RX_ERR_ARG; }
  if (stage < 0 || stage > 2) { c->err = "rx_turb_postprocess: bad stage"; return RX_ERR_ARG; }
  { const int rc = ensure_output(c); if (rc) return rc; }
  const Dev& d = c->d;
  if (stage != 2) {
    if (kOmega) { const int rc = upload_aos(c, kOmega, c->tpost, 2); if (rc) return rc; }   // NULL: the (k, omega) of the previous call
    if (c->o.leastSquares) { k_turb_gradient_ls<<<nblk(d.nPoint, 128), 128, 0, c->stream>>>(d, c->tpost); c->launches++; }
    else {
      k_turb_gradient_gg_edges<<<nblk(d.nPoint, 128), 128, 0, c->stream>>>(d, c->tpost); c->launches++;
      for (const MarkerDev& m : c->markers)
        if (m.nVertex) { k_turb_gradient_gg_boundary<<<nblk(m.nVertex, 64), 64, 0, c->stream>>>(d, m, c->tpost); c->launches++; }
      k_turb_gradient_gg_finish<<<nblk(d.nPoint, 128), 128, 0, c->stream>>>(d, c->tpost); c->launches++;
    }
  }
  c->sigmaK = sigmaK;
  if (stage == 0 && c->comm) {   // Set_MPI_Solution_Gradient (solver_direct_turbulent.cpp:227) between the two halves
    const int rc = halo_exchange_field(c, c->tpost + (size_t)TP_GK * d.np, 4, false, c->stream);
    if (rc) return rc;
  }
  if (stage != 1) { k_turb_blend<<<nblk(d.nPoint, 128), 128, 0, c->stream>>>(d, c->tpost, sigmaK); c->launches++; }
  RX_CK(cudaGetLastError());
  return RX_OK;
}
// ---- SST transport step (rx_turb_step.cuh)
static int turb_step_setup(rx_ctx* c, int m) {
  if (c->turbStepReady && (int)c->kw.size() >= m + 1) return RX_OK;
  const Dev& d = c->d;
  const size_t np = d.np;
  if (!c->turbStepReady) {
    c->tw.nInc = (long long)c->hNbOfSlot.size();
    c->tw.res = dalloc<double>(c, 2 * np); c->tw.diag = dalloc<double>(c, 2 * np); c->tw.x = dalloc<double>(c, 2 * np); c->tw.old = dalloc<double>(c, 2 * np);
    c->tw.off = dalloc<double>(c, 2 * (size_t)std::max<long long>(c->tw.nInc, 1));
    c->tBlocks = (int)nblk(d.nPointDomain, 128); c->tDotBlocks = (int)nblk(2 * d.nPointDomain, 256);
    c->tPartial = dalloc<double>(c, (size_t)std::max(2 * c->tBlocks, c->tDotBlocks) + 2);
    c->tScal = dalloc<double>(c, 8);
    // level lists: forward sweep level(p) = 1 + max over lower-numbered neighbours; backward sweep likewise over higher-numbered ones
    const long long n = d.nPointDomain;
    std::vector<int> lev(n, 0);
    auto build = [&](bool fwd, SweepList& sl, SweepData& sd) {
      int nl = 0;
      for (long long t = 0; t < n; ++t) {
        const long long p = fwd ? t : n - 1 - t;
        int l = 0;
        for (int k = c->hIncPtr[p]; k < c->hIncPtr[p + 1]; ++k) {
          const int q = c->hNbOfSlot[k];
          if (q < n && (fwd ? q < p : q > p)) l = std::max(l, lev[q] + 1);
        }
        lev[p] = l; nl = std::max(nl, l + 1);
      }
      std::vector<int> ptr(nl + 1, 0), pts(n), pos(n), start(n + 1, 0), col, slot;
      for (long long p = 0; p < n; ++p) ptr[lev[p] + 1]++;
      for (int l = 0; l < nl; ++l) ptr[l + 1] += ptr[l];
      std::vector<int> fill(ptr.begin(), ptr.end() - 1);
      for (long long p = 0; p < n; ++p) { pos[p] = fill[lev[p]]; pts[fill[lev[p]]++] = (int)p; }
      for (long long t = 0; t < n; ++t) {   // the point's lower (upper) neighbours in slot order, as POSITIONS in this level order
        const int p = pts[t];
        for (int k = c->hIncPtr[p]; k < c->hIncPtr[p + 1]; ++k) {
          const int q = c->hNbOfSlot[k];
          if (fwd ? q < p : q > p) { col.push_back(q < n ? pos[q] : q); slot.push_back(k); }   // a halo column: its x* sits at position q of the backward list's x
        }
        start[t + 1] = (int)col.size();
      }
      sl.nEnt = (int)col.size();
      if (col.empty()) { col.push_back(0); slot.push_back(0); }
      sl.nLvl = nl; sl.lvlPtr = dupload(c, ptr); sl.perm = dupload(c, pts); sl.start = dupload(c, start); sl.colL = dupload(c, col); sl.slot = dupload(c, slot);
      sl.nX = (int)(fwd ? n : d.nPoint);
      sd.coefL = dalloc<double>(c, 2 * (size_t)std::max(sl.nEnt, 1)); sd.dgL = dalloc<double>(c, 2 * (size_t)n); sd.bL = dalloc<double>(c, 2 * (size_t)n); sd.xL = dalloc<double>(c, 2 * (size_t)sl.nX);
      if (fwd) c->posF = dupload(c, pos);
    };
    build(true, c->swF, c->sdF);
    build(false, c->swB, c->sdB);
    if (c->allocFailed) { c->err = "rx_turb_iterate: device allocation failed"; return RX_ERR_CUDA; }
    if (const char* ev = getenv("RX_SWEEP_THREADS")) { const int t = atoi(ev); if (t >= 32 && t <= 1024 && t % 32 == 0) c->sweepThreads = t; }
    c->turbStepReady = true;
  }
  while ((int)c->kw.size() < m + 1) {
    c->kw.push_back(dalloc<double>(c, 2 * np)); c->kz.push_back(dalloc<double>(c, 2 * np));
    if (!c->kw.back() || !c->kz.back()) { c->err = "rx_turb_iterate: device allocation failed"; return RX_ERR_CUDA; }
  }
  return RX_OK;
}
// dot product of two [2][nOwn] vectors: fixed block partials + fixed tree, result on the host (FGMRES needs it there)
static int turb_dot(rx_ctx* c, const double* a, const double* b, double* out) {
  const Dev& d = c->d;
  k_turb_dot<<<c->tDotBlocks, 256, 0, c->stream>>>(d.nPointDomain, d.np, a, b, c->tPartial);
  k_turb_sum<<<1, 256, 0, c->stream>>>(c->tPartial, c->tDotBlocks, 1, 0, c->tScal);
  c->launches += 2;
  if (c->comm) { const int rc = comm_allreduce_sum(c, c->tScal, 1); if (rc) return rc; }
  RX_CK(cudaMemcpyAsync(out, c->tScal, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  RX_CK(cudaStreamSynchronize(c->stream));
  return RX_OK;
}
int rx_turb_set_state(rx_ctx* c, const double* kOmega, const double* blend3) {
  if (!c->tpost) { c->err = "rx_turb_set_state: not a RANS context"; return RX_ERR_ARG; }
  int rc = RX_OK;
  if (kOmega) rc = upload_aos(c, kOmega, c->tpost, 2);
  if (!rc && blend3) rc = upload_aos(c, blend3, c->tpost + (size_t)TP_F1 * c->d.np, 3);   // F1, F2, CDkw are consecutive columns
  return rc;
}
// dual time stepping: the SST solver's stored time levels.  Until set or shifted for the first time they equal the current (k, omega).
static int ensure_turb_levels(rx_ctx* c, const char* who) {
  if (!c->tkwN) { c->err = std::string(who) + ": not a RANS context created for dual time stepping"; return RX_ERR_ARG; }
  if (!c->turbLevelsSet) {
    const size_t b = 2 * (size_t)c->d.np * sizeof(double);   // k, omega are the first two columns of the Postprocessing block
    RX_CK(cudaMemcpyAsync(c->tkwN, c->tpost, b, cudaMemcpyDeviceToDevice, c->stream));
    RX_CK(cudaMemcpyAsync(c->tkwN1, c->tpost, b, cudaMemcpyDeviceToDevice, c->stream));
    c->turbLevelsSet = true;
  }
  return RX_OK;
}
int rx_turb_set_time_levels(rx_ctx* c, const double* kwN, const double* kwN1) {
  if (!c->tkwN) { c->err = "rx_turb_set_time_levels: not a RANS context created for dual time stepping"; return RX_ERR_ARG; }
  if (!kwN || !kwN1) { c->err = "rx_turb_set_time_levels: NULL argument"; return RX_ERR_ARG; }
  int rc = upload_aos(c, kwN, c->tkwN, 2);
  if (!rc) rc = upload_aos(c, kwN1, c->tkwN1, 2);
  if (!rc) c->turbLevelsSet = true;
  return rc;
}
int rx_turb_get_time_levels(rx_ctx* c, double* kwN, double* kwN1) {
  int rc = ensure_turb_levels(c, "rx_turb_get_time_levels");
  if (!rc && kwN) rc = download_aos(c, c->tkwN, kwN, 2);
  if (!rc && kwN1) rc = download_aos(c, c->tkwN1, kwN1, 2);
  return rc;
}
int rx_turb_shift_time_levels(rx_ctx* c) {
  const int rc = ensure_turb_levels(c, "rx_turb_shift_time_levels");
  if (rc) return rc;
  const size_t b = 2 * (size_t)c->d.np * sizeof(double);
  RX_CK(cudaMemcpyAsync(c->tkwN1, c->tkwN, b, cudaMemcpyDeviceToDevice, c->stream));
  RX_CK(cudaMemcpyAsync(c->tkwN, c->tpost, b, cudaMemcpyDeviceToDevice, c->stream));
  return RX_OK;
}
int rx_turb_get_step(rx_ctx* c, double* res2, double* dx2) {
  if (!c->turbStepReady) { c->err = "rx_turb_get_step: no turbulence step has run"; return RX_ERR_ARG; }
  int rc = RX_OK;
  if (res2) {   // the work array holds the right-hand side -Res of the implicit system
    rc = download_aos(c, c->tw.res, res2, 2);
    for (long long k = 0; k < 2 * c->d.nPoint; ++k) res2[k] = -res2[k];
  }
  if (!rc && dx2) rc = download_aos(c, c->tw.x, dx2, 2);
  return rc;
}
int rx_turb_iterate(rx_ctx* c, const rx_turb_opts* to, double* resRms2, int* linIters) {
  if (!c->tpost) { c->err = "rx_turb_iterate: not a RANS context"; return RX_ERR_ARG; }
  if (!c->haveWallDist) { c->err = "rx_turb_iterate: the wall distance has not been set (rx_set_wall_distance / rx_compute_wall_distance)"; return RX_ERR_ARG; }
  if (c->d.nPoint != c->d.nPointDomain && !c->comm) { c->err = "rx_turb_iterate: a partition needs the in-library communicator (rx_comm_init): the step exchanges its Krylov and LU-SGS vectors"; return RX_ERR_UNSUPPORTED; }
  if (!to || to->linIter < 1 || to->linIter > 64) { c->err = "rx_turb_iterate: bad options"; return RX_ERR_ARG; }
  const int m = to->linIter;
  int rc = ensure_output(c);
  if (!rc) rc = turb_step_setup(c, m);
  if (rc) return rc;
  const Dev& d = c->d;
  cudaStream_t s = c->stream;
  double* tp = c->tpost;
  TurbC tc;
  tc.alfa_1 = tc.beta_1 / tc.beta_star - tc.sigma_om1 * 0.41 * 0.41 / std::sqrt(tc.beta_star);
  tc.alfa_2 = tc.beta_2 / tc.beta_star - tc.sigma_om2 * 0.41 * 0.41 / std::sqrt(tc.beta_star);
  tc.kineInf = to->kineInf; tc.omegaInf = to->omegaInf; tc.cflRed = to->cflRed; tc.relax = to->relax;
  const TurbW& w = c->tw;
  const long long n = d.nPointDomain;
  // Preprocessing: gradients of (k, omega); Set_OldSolution
  if (c->o.leastSquares) { k_turb_gradient_ls<<<nblk(d.nPoint, 128), 128, 0, s>>>(d, tp); c->launches++; }
  else {
    k_turb_gradient_gg_edges<<<nblk(d.nPoint, 128), 128, 0, s>>>(d, tp); c->launches++;
    for (const MarkerDev& mk : c->markers)
      if (mk.nVertex) { k_turb_gradient_gg_boundary<<<nblk(mk.nVertex, 64), 64, 0, s>>>(d, mk, tp); c->launches++; }
    k_turb_gradient_gg_finish<<<nblk(d.nPoint, 128), 128, 0, s>>>(d, tp); c->launches++;
  }
  const bool par = c->comm != nullptr;   // partition of a multi-GPU job: the exchanges of the reference's MPI build, over NCCL on the context's stream
  if (par && (rc = halo_exchange_field(c, tp + (size_t)TP_GK * d.np, 4, false, s))) return rc;
  RX_CK(cudaMemcpyAsync(w.old, tp, 2 * (size_t)d.np * sizeof(double), cudaMemcpyDeviceToDevice, s));
  // Space_Integration: rows, then the boundary conditions (weak ones first, walls last: integration_structure.cpp:89-190)
  k_turb_rows<<<nblk(n, 128), 128, 0, s>>>(tc, d, tp, w); c->launches++;
  if (c->o.dualTime) {   // SetResidual_DualTime comes between the source term and the boundary conditions (integration_structure.cpp:86-87)
    if ((rc = ensure_time_levels(c, "rx_turb_iterate")) || (rc = ensure_turb_levels(c, "rx_turb_iterate"))) return rc;
    k_turb_dual_time<<<nblk(n, 128), 128, 0, s>>>(d, tp, w, c->o.dualTime, c->dtND, c->Un, c->Un1, c->tkwN, c->tkwN1); c->launches++;
  }
  for (const MarkerDev& mk : c->markers)
    if (mk.nVertex && (is_inlet(mk.kind) || mk.kind == RX_BC_OUTLET)) {
      k_turb_bc_io<<<nblk(mk.nVertex, 64), 64, 0, s>>>(tc, d, mk, tp, w, is_inlet(mk.kind) ? 1 : 0); c->launches++;
    }
  for (const MarkerDev& mk : c->markers)
    if (mk.nVertex && (mk.kind == RX_BC_ISOTHERMAL || mk.kind == RX_BC_HEATFLUX)) { k_turb_bc_wall<<<nblk(mk.nVertex, 64), 64, 0, s>>>(tc, d, mk, tp, w); c->launches++; }
  // ImplicitEuler_Iteration
  k_turb_implicit_rhs<<<c->tBlocks, 128, 0, s>>>(tc, d, w, c->tPartial);
  k_turb_sum<<<1, 256, 0, s>>>(c->tPartial, c->tBlocks, 2, 0, c->tScal + 2);
  k_turb_sum<<<1, 256, 0, s>>>(c->tPartial, c->tBlocks, 2, 1, c->tScal + 2);
  c->launches += 3;
  double ss[3];
  if (par) {   // sums of squares and the owned-point count over the ranks (SetResidual_RMS with MPI, solver_structure.cpp:188-260)
    const double nOwn = (double)n;
    RX_CK(cudaMemcpyAsync(c->tScal + 4, &nOwn, sizeof(double), cudaMemcpyHostToDevice, s));
    if ((rc = comm_allreduce_sum(c, c->tScal + 2, 3))) return rc;
  }
  RX_CK(cudaMemcpyAsync(ss, c->tScal + 2, 3 * sizeof(double), cudaMemcpyDeviceToHost, s));
  RX_CK(cudaStreamSynchronize(s));
  if (resRms2) for (int v = 0; v < 2; ++v) resRms2[v] = std::max(EPS * EPS, std::sqrt(ss[v] / (par ? ss[2] : (double)d.nPoint)));   // SetResidual_RMS
  // FGMRES (linear_solvers_structure.cpp:309-445), x0 = 0: w0 = (A x0 - b) / (-beta) = b / beta
  // level-order copies of the assembled rows for the sweeps (once per step), then per application: permute in, forward sweep,
  // re-order to the backward sweep's level order, backward sweep, permute out
  k_turb_sweep_coef<<<nblk(std::max<long long>(c->swF.nEnt, n), 256), 256, 0, s>>>(c->swF, c->sdF, w, n, d.np);
  k_turb_sweep_coef<<<nblk(std::max<long long>(c->swB.nEnt, n), 256), 256, 0, s>>>(c->swB, c->sdB, w, n, d.np);
  c->launches += 2;
  auto precond = [&](const double* in, double* out) -> int {
    k_turb_perm_in<<<nblk(n, 256), 256, 0, s>>>(c->swF.perm, n, d.np, in, c->sdF.bL);
    k_turb_sweep<false><<<2, c->sweepThreads, 0, s>>>(c->swF, c->sdF, n);
    if (par && d.nPoint > n) {   // x* of the neighbours' points -> the halo tail of the backward list (SendReceive_Solution between the sweeps, :1689)
      k_turb_perm_out<<<nblk(n, 256), 256, 0, s>>>(c->swF.perm, n, n, d.np, c->sdF.xL, out);
      const int e = halo_exchange_field(c, out, 2, false, s);
      if (e) return e;
      k_turb_halo_in<<<nblk(d.nPoint - n, 256), 256, 0, s>>>(n, d.nPoint, d.np, out, c->sdB.xL);
      c->launches += 2;
    }
    k_turb_perm_ab<<<nblk(n, 256), 256, 0, s>>>(c->swB.perm, c->posF, n, c->swB.nX, c->sdF.xL, c->sdB.xL);
    k_turb_sweep<true><<<2, c->sweepThreads, 0, s>>>(c->swB, c->sdB, n);
    k_turb_perm_out<<<nblk(n, 256), 256, 0, s>>>(c->swB.perm, n, c->swB.nX, d.np, c->sdB.xL, out);
    c->launches += 5;
    return par ? halo_exchange_field(c, out, 2, false, s) : RX_OK;   // the matrix-vector product reads the neighbours' entries (:1707)
  };
  const unsigned gv = nblk(2 * n, 256);
  std::vector<double> g(m + 1, 0.0), sn(m + 1, 0.0), cs(m + 1, 0.0), y(m, 0.0);
  std::vector<std::vector<double>> H(m + 1, std::vector<double>(m, 0.0));
  double bb;
  if ((rc = turb_dot(c, w.res, w.res, &bb))) return rc;
  double norm0 = std::sqrt(bb), beta = norm0;
  int it = 0;
  if (!(beta < to->linTol * norm0 || beta < 2.220446049250313e-16)) {   // eps = machine epsilon (matrix_structure.hpp:49)
    k_turb_scale_div<<<gv, 256, 0, s>>>(n, d.np, -beta, -1.0, w.res, c->kw[0]); c->launches++;
    g[0] = beta; norm0 = beta;
    auto applyGivens = [](double sg, double cg, double& h1, double& h2) { const double t = cg * h1 + sg * h2; h2 = cg * h2 - sg * h1; h1 = t; };
    auto sign = [](double x, double yv) { return yv == 0.0 ? 0.0 : (yv < 0 ? -std::fabs(x) : std::fabs(x)); };
    for (it = 0; it < m; ++it) {
      if (beta < to->linTol * norm0) break;
      if ((rc = precond(c->kw[it], c->kz[it]))) return rc;
      k_turb_matvec<<<nblk(n, 128), 128, 0, s>>>(d, w, c->kz[it], c->kw[it + 1]); c->launches++;
      // ModGramSchmidt (:87-175)
      double nrm;
      if ((rc = turb_dot(c, c->kw[it + 1], c->kw[it + 1], &nrm))) return rc;
      double thr = nrm * 0.98;
      if (nrm <= 0.0 || nrm != nrm) { c->err = "SU2 has diverged (FGMRES: the Krylov vector of the turbulence system has no norm)"; return RX_ERR_NAN; }
      for (int k = 0; k < it + 1; ++k) {
        double prod;
        if ((rc = turb_dot(c, c->kw[it + 1], c->kw[k], &prod))) return rc;
        H[k][it] = prod;
        k_turb_axpy<<<gv, 256, 0, s>>>(n, d.np, -prod, c->kw[k], c->kw[it + 1]); c->launches++;
        if (prod * prod > thr) {
          if ((rc = turb_dot(c, c->kw[it + 1], c->kw[k], &prod))) return rc;
          H[k][it] += prod;
          k_turb_axpy<<<gv, 256, 0, s>>>(n, d.np, -prod, c->kw[k], c->kw[it + 1]); c->launches++;
        }
        nrm -= H[k][it] * H[k][it];
        if (nrm < 0.0) nrm = 0.0;
        thr = nrm * 0.98;
      }
      if ((rc = turb_dot(c, c->kw[it + 1], c->kw[it + 1], &nrm))) return rc;
      nrm = std::sqrt(nrm);
      H[it + 1][it] = nrm;
      k_turb_scale_div<<<gv, 256, 0, s>>>(n, d.np, nrm, 1.0, c->kw[it + 1], c->kw[it + 1]); c->launches++;
      for (int k = 0; k < it; ++k) applyGivens(sn[k], cs[k], H[k][it], H[k + 1][it]);
      {   // GenerateGivens (:44-71)
        double& dx = H[it][it]; double& dy = H[it + 1][it]; double& sg = sn[it]; double& cg = cs[it];
        if (dx == 0.0 && dy == 0.0) { cg = 1.0; sg = 0.0; }
        else if (std::fabs(dy) > std::fabs(dx)) { const double t = dx / dy; dx = std::sqrt(1.0 + t * t); sg = sign(1.0 / dx, dy); cg = t * sg; }
        else if (std::fabs(dy) <= std::fabs(dx)) { const double t = dy / dx; dy = std::sqrt(1.0 + t * t); cg = sign(1.0 / dy, dx); sg = t * cg; }
        else { dx = 0.0; dy = 0.0; cg = 1.0; sg = 0.0; }
        dx = std::fabs(dx * dy); dy = 0.0;
      }
      applyGivens(sn[it], cs[it], g[it], g[it + 1]);
      beta = std::fabs(g[it + 1]);
    }
    // SolveReduced + x += sum y_k z_k
    for (int k = 0; k < it; ++k) y[k] = g[k];
    for (int k = it - 1; k >= 0; --k) { y[k] /= H[k][k]; for (int j = k - 1; j >= 0; --j) y[j] -= H[j][k] * y[k]; }
    for (int k = 0; k < it; ++k) { k_turb_axpy<<<gv, 256, 0, s>>>(n, d.np, y[k], c->kz[k], w.x); c->launches++; }
  }
  if (linIters) *linIters = it;
  k_turb_update<<<nblk(n, 128), 128, 0, s>>>(tc, d, tp, w); c->launches++;
  RX_CK(cudaGetLastError());
  if (par && (rc = halo_exchange_field(c, tp + (size_t)TP_K * d.np, 2, false, s))) return rc;   // Set_MPI_Solution (solver_direct_turbulent.cpp:707)
  // Postprocessing: gradients of the new (k, omega), blending functions, eddy viscosity -> the fields the flow path reads
  return rx_turb_postprocess(c, nullptr, c->sigmaK, 0);
}

int rx_get_turbulence(rx_ctx* c, double* post12) {
  if (!c->tpost) { c->err = "rx_get_turbulence: not a RANS context"; return RX_ERR_ARG; }
  return download_aos(c, c->tpost, post12, TP_N);
}
int rx_set_temperature(rx_ctx* c, const double* T) { return upload_aos(c, T, c->d.V + V_T * c->d.np, 1); }
int rx_get_solution(rx_ctx* c, double* U) { return download_aos(c, c->d.U, U, c->nVar); }
int rx_get_residual(rx_ctx* c, double* R) { return download_aos(c, c->d.R, R, c->nVar); }
int rx_get_limiter(rx_ctx* c, double* lim4) { return download_aos(c, c->d.lim, lim4, 4); }
int rx_get_primitive(rx_ctx* c, double* V, double* aux, double* grad) {
  int rc = ensure_output(c);
  if (rc) return rc;
  if (V) rc = download_aos(c, c->d.V, V, c->nPrim);
  if (!rc && grad) rc = download_aos(c, c->d.grad, grad, c->nGrad * 2);
  if (!rc && aux) {
    const long long n = c->d.nPoint;
    rc = ensure_stage(c, (size_t)n * 6 * sizeof(double));
    if (rc) return rc;
    const double* f[6] = {c->d.Cp, c->d.gam, c->d.mu, c->d.kap, c->d.vort, c->d.strain};
    for (int k = 0; k < 6; ++k) { k_gather_field<<<nblk(n, 256), 256, 0, c->stream>>>(f[k], c->stage, n, 6, k); c->launches++; }
    RX_CK(cudaGetLastError());
    RX_CK(cudaMemcpyAsync(aux, c->stage, (size_t)n * 6 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    RX_CK(cudaStreamSynchronize(c->stream));
  }
  return rc;
}
int rx_get_dt(rx_ctx* c, double* dt3) {
  const long long n = c->d.nPoint;
  int rc = ensure_stage(c, (size_t)n * 3 * sizeof(double));
  if (rc) return rc;
  const double* f[3] = {c->d.dt, c->d.lamInv, c->d.lamVisc};
  for (int k = 0; k < 3; ++k) { k_gather_field<<<nblk(n, 256), 256, 0, c->stream>>>(f[k], c->stage, n, 3, k); c->launches++; }
  RX_CK(cudaGetLastError());
  RX_CK(cudaMemcpyAsync(dt3, c->stage, (size_t)n * 3 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  RX_CK(cudaStreamSynchronize(c->stream));
  return RX_OK;
}

// Preprocessing in two halves: cons->prim (the stateful temperature solve: it must run exactly as often as in the reference) and
// everything derived from it (transport, gradients, vorticity / strain, limiters), which is a pure function of the first half.
static int preprocess_primitive(rx_ctx* c, int64_t extIter, int output) {
  const Dev& d = c->d;
  cudaStream_t s = c->stream;
  c->outputPending = false;
  RX_CK(cudaMemsetAsync(d.counters, 0, sizeof(unsigned long long), s));
  RX_NS(k_primitive<NS><<<nblk(d.nPoint, 128), 128, 0, s>>>(cc, c->o, d, (long long)extIter, output));
  c->launches++;
  if (c->eddyFlow) RX_CK(cudaMemcpyAsync(c->eddyFlow, d.turb, (size_t)d.np * sizeof(double), cudaMemcpyDeviceToDevice, s));
  return RX_OK;
}
static int preprocess_derived(rx_ctx* c, int output) {
  const Dev& d = c->d;
  cudaStream_t s = c->stream;
  if (c->o.viscous) { RX_NS(k_transport<NS><<<nblk(d.nPoint, 128), 128, 0, s>>>(cc, d)); c->launches++; }
  if (c->o.viscous || c->o.secondOrder) {
    if (c->o.leastSquares) {
      RX_NS(k_gradient_ls<NS><<<nblk(d.nPoint, 128), 128, 0, s>>>(d));
      c->launches++;
    } else {
      RX_NS(k_gradient_gg_edges<NS><<<nblk(d.nPointDomain, 128), 128, 0, s>>>(d));
      c->launches++;
      for (const MarkerDev& m : c->markers)
        if (m.nVertex) { RX_NS(k_gradient_gg_boundary<NS><<<nblk(m.nVertex, 64), 64, 0, s>>>(d, m)); c->launches++; }
      RX_NS(k_gradient_gg_finish<NS><<<nblk(d.nPoint, 128), 128, 0, s>>>(d));
      c->launches++;
    }
  }
  if (c->o.secondOrder && c->o.limiter && !output) {   // :4739-4742
    if (c->o.limiter == 2) k_limiter_bj<<<nblk(d.nPoint, 128), 128, 0, s>>>(d);
    else k_limiter<<<nblk(d.nPoint, 128), 128, 0, s>>>(c->o, d);
    c->launches++;
  }
  RX_CK(cudaGetLastError());
  return RX_OK;
}
// The Output pass of rx_iterate is lazy: the next iteration's first Preprocessing recomputes transport and gradients from its own
// temperature solve, so the Output pass's copies are only formed when somebody reads them in between (rx_get_primitive, the
// turbulence step, the restart writer, a gradient halo pack) — same values either way, 1.2 ms per laminar step saved on the bench mesh.
static int ensure_output(rx_ctx* c) {
  if (!c->outputPending) return RX_OK;
  c->outputPending = false;
  int rc = preprocess_derived(c, 1);
  if (!rc && c->comm && (c->o.viscous || c->o.secondOrder)) rc = rx_halo_exchange(c, 1);
  return rc;
}
int rx_preprocess(rx_ctx* c, int64_t extIter, int output, int64_t* nonPhysical) {
  int rc = preprocess_primitive(c, extIter, output);
  if (!rc) rc = preprocess_derived(c, output);
  if (rc) return rc;
  if (nonPhysical) {
    unsigned long long h = 0;
    RX_CK(cudaMemcpyAsync(&h, c->d.counters, sizeof(h), cudaMemcpyDeviceToHost, c->stream));
    RX_CK(cudaStreamSynchronize(c->stream));
    *nonPhysical = (int64_t)h;
  }
  return RX_OK;
}

// new processing order of the viscous edge kernel from the keys its last launch left (rx_order.cuh).  Called once per
// iteration, from rx_set_old_solution — the one call every driver of the path (rx_iterate, the staged calls of the halo-
// exchanging loop) makes exactly once per iteration before the first viscous residual.
static int order_edges(rx_ctx* c) {
  Dev& d = c->d;
  if (!c->orderEdges || !c->eOrderBuf || d.nEdge == 0) return RX_OK;
  k_order_tiles<<<c->nOrdTiles, ORD_BS, 0, c->stream>>>(d.eKey, d.nEdge, c->eOrderBuf);
  c->launches += 1;
  RX_CK(cudaGetLastError());
  d.eOrder = c->eOrderBuf;
  return RX_OK;
}

int rx_set_old_solution(rx_ctx* c) {
  { const int rc = order_edges(c); if (rc) return rc; }
  RX_CK(cudaMemcpyAsync(c->d.Uold, c->d.U, (size_t)c->d.np * c->nVar * sizeof(double), cudaMemcpyDeviceToDevice, c->stream));
  return RX_OK;
}

int rx_set_cfl(rx_ctx* c, double cfl) {
  if (!(cfl >= 0.0)) { c->err = "rx_set_cfl: the CFL number must be non-negative"; return RX_ERR_ARG; }
  c->o.cfl = cfl;
  return RX_OK;
}
// minimum of the local time steps over the owned points of every partition (Global_Delta_Time and its MPI_MIN reduction, :5227, :5264-5269):
// two-stage minimum on the device, then one ncclAllReduce when a communicator exists; *glob points at the result on the device
static int global_dt_min(rx_ctx* c, double** glob) {
  const Dev& d = c->d;
  cudaStream_t s = c->stream;
  if (!c->dtPartial) {
    c->dtBlocks = (int)nblk(d.nPointDomain, 256);
    c->dtPartial = dalloc<double>(c, (size_t)c->dtBlocks + 1);
    if (!c->dtPartial) { c->err = "rx_time_step: device allocation failed"; return RX_ERR_CUDA; }
  }
  *glob = c->dtPartial + c->dtBlocks;
  k_dt_min_partial<<<c->dtBlocks, 256, 0, s>>>(d, c->dtPartial);
  k_dt_min_final<<<1, 256, 0, s>>>(c->dtPartial, c->dtBlocks, *glob);
  c->launches += 2;
  if (c->comm) {
    std::string why;
    NcclApi* a = nccl_api(why);
    const int e = a->AllReduce(*glob, *glob, 1, NCCL_FLOAT64, NCCL_MIN, c->comm->comm, s);
    if (e) { c->err = std::string("rx_time_step: NCCL all-reduce: ") + a->GetErrorString(e); return RX_ERR_CUDA; }
  }
  return RX_OK;
}
int rx_time_step(rx_ctx* c) {
  const Dev& d = c->d;
  cudaStream_t s = c->stream;
  k_timestep_edges<<<nblk(d.nPointDomain, 128), 128, 0, s>>>(c->o, d);
  c->launches++;
  for (const MarkerDev& m : c->markers)
    if (m.nVertex) { k_timestep_boundary<<<nblk(m.nVertex, 64), 64, 0, s>>>(c->o, d, m); c->launches++; }
  {
    Opts of = c->o;
    if (of.dualTime) of.timeStepping = 1;   // leave the steps unclamped: k_dual_dt_clamp applies Max_DeltaTime after the global minimum has been taken
    k_timestep_finish<<<nblk(d.nPointDomain, 128), 128, 0, s>>>(of, d);
  }
  c->launches++;
  if (c->o.timeStepping) {
    double* glob = nullptr;
    { const int rc = global_dt_min(c, &glob); if (rc) return rc; }
    // the Euler solver's variant (:2121-2139): from now on the CFL is UNST_CFL_NUMBER; if that is zero every point takes UNST_TIMESTEP
    if (!c->o.viscous) c->o.cfl = c->o.unstCfl;
    const int useFixed = (!c->o.viscous && c->o.cfl < EPS) ? 1 : 0;
    k_dt_fill<<<nblk(d.nPointDomain, 256), 256, 0, s>>>(d, glob, c->o.deltaUnstTime, useFixed);
    c->launches++;
  }
  if (c->o.dualTime) {
    if (c->intIter == 0 && c->o.unstCfl != 0.0) {   // the physical step follows UNST_CFL_NUMBER (:5276-5289): global minimum of the unclamped local steps
      double* glob = nullptr;
      { const int rc = global_dt_min(c, &glob); if (rc) return rc; }
      k_dual_unst_dt<<<1, 1, 0, s>>>(glob, c->o.unstCfl, c->o.cfl, c->dtND);
      c->launches++;
    }
    k_dual_dt_clamp<<<nblk(d.nPointDomain, 256), 256, 0, s>>>(d, c->o.maxDt, c->dtND);
    c->launches++;
  }
  RX_CK(cudaGetLastError());
  return RX_OK;
}

// ---- dual time stepping ---------------------------------------------------------------------------------------------------------
static int ensure_time_levels(rx_ctx* c, const char* who) {
  if (!c->o.dualTime) { c->err = std::string(who) + ": the context was not created for dual time stepping (rx_opts.dualTime)"; return RX_ERR_ARG; }
  if (!c->timeLevelsSet) {   // SetInitialCondition leaves U^n = U^(n-1) = U
    const size_t b = (size_t)c->d.np * c->nVar * sizeof(double);
    RX_CK(cudaMemcpyAsync(c->Un, c->d.U, b, cudaMemcpyDeviceToDevice, c->stream));
    RX_CK(cudaMemcpyAsync(c->Un1, c->d.U, b, cudaMemcpyDeviceToDevice, c->stream));
    c->timeLevelsSet = true;
  }
  return RX_OK;
}
int rx_set_time_levels(rx_ctx* c, const double* Un, const double* Un1) {
  if (!c->o.dualTime) { c->err = "rx_set_time_levels: the context was not created for dual time stepping (rx_opts.dualTime)"; return RX_ERR_ARG; }
  if (!Un || !Un1) { c->err = "rx_set_time_levels: NULL argument"; return RX_ERR_ARG; }
  int rc = upload_aos(c, Un, c->Un, c->nVar);
  if (!rc) rc = upload_aos(c, Un1, c->Un1, c->nVar);
  if (!rc) c->timeLevelsSet = true;
  return rc;
}
int rx_get_time_levels(rx_ctx* c, double* Un, double* Un1) {
  int rc = ensure_time_levels(c, "rx_get_time_levels");
  if (!rc && Un) rc = download_aos(c, c->Un, Un, c->nVar);
  if (!rc && Un1) rc = download_aos(c, c->Un1, Un1, c->nVar);
  return rc;
}
int rx_shift_time_levels(rx_ctx* c) {
  const int rc = ensure_time_levels(c, "rx_shift_time_levels");
  if (rc) return rc;
  const size_t b = (size_t)c->d.np * c->nVar * sizeof(double);
  RX_CK(cudaMemcpyAsync(c->Un1, c->Un, b, cudaMemcpyDeviceToDevice, c->stream));
  RX_CK(cudaMemcpyAsync(c->Un, c->d.U, b, cudaMemcpyDeviceToDevice, c->stream));
  return RX_OK;
}
int rx_set_inner_iteration(rx_ctx* c, int64_t intIter) {
  if (intIter < 0) { c->err = "rx_set_inner_iteration: negative iteration index"; return RX_ERR_ARG; }
  c->intIter = intIter;
  return RX_OK;
}
int rx_get_delta_unst_time(rx_ctx* c, double* dtND) {
  if (!c->o.dualTime || !dtND) { c->err = "rx_get_delta_unst_time: not a dual time stepping context, or NULL argument"; return RX_ERR_ARG; }
  RX_CK(cudaMemcpyAsync(dtND, c->dtND, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  RX_CK(cudaStreamSynchronize(c->stream));
  return RX_OK;
}
int rx_residual_dual_time(rx_ctx* c) {
  const int rc = ensure_time_levels(c, "rx_residual_dual_time");
  if (rc) return rc;
  const Dev& d = c->d;
  k_dual_time_residual<<<nblk(d.nPointDomain, 128), 128, 0, c->stream>>>(d, c->nVar, c->o.dualTime, c->dtND, c->Un, c->Un1);
  c->launches++;
  RX_CK(cudaGetLastError());
  return RX_OK;
}

int rx_residual_conv(rx_ctx* c) {
  const Dev& d = c->d;
  launch_edge_flux<true, false>(c, c->stream);
  RX_NS(k_accumulate<NS, 1><<<nblk(d.nPoint, 128), 128, 0, c->stream>>>(cc, c->o, d));
  c->launches++;
  RX_CK(cudaGetLastError());
  return RX_OK;
}
int rx_residual_visc(rx_ctx* c) {
  if (!c->o.viscous) return RX_OK;
  const Dev& d = c->d;
  launch_edge_flux<false, true>(c, c->stream);
  RX_NS(k_accumulate<NS, 2><<<nblk(d.nPoint, 128), 128, 0, c->stream>>>(cc, c->o, d));
  c->launches++;
  RX_CK(cudaGetLastError());
  return RX_OK;
}
int rx_residual_source(rx_ctx* c) {
  const Dev& d = c->d;
  RX_NS(k_source<NS><<<nblk(d.nPointDomain, 128), 128, 0, c->stream>>>(cc, c->o, d));
  c->launches++;
  RX_CK(cudaGetLastError());
  return RX_OK;
}
static void launch_bc(rx_ctx* c, const MarkerDev& m, cudaStream_t s) {
  const Dev& d = c->d;
  if (!m.nVertex) return;
  if (is_supersonic(m.kind)) { RX_NS(k_bc_supersonic<NS><<<nblk(m.nVertex, EDGE_BS), EDGE_BS, SM_VISC, s>>>(cc, c->o, d, m)); c->launches++; }
  else if (is_inlet(m.kind)) { RX_NS(k_bc_inlet<NS><<<nblk(m.nVertex, EDGE_BS), EDGE_BS, SM_VISC, s>>>(cc, c->o, d, m)); c->launches++; }
  else if (m.kind == RX_BC_OUTLET) { RX_NS(k_bc_outlet<NS><<<nblk(m.nVertex, EDGE_BS), EDGE_BS, SM_VISC, s>>>(cc, c->o, d, m)); c->launches++; }
  else if (m.kind == RX_BC_ISOTHERMAL) { RX_NS(k_bc_isothermal<NS><<<nblk(m.nVertex, 64), 64, 0, s>>>(cc, c->o, d, m)); c->launches++; }
  else if (m.kind == RX_BC_HEATFLUX) { k_bc_heatflux<<<nblk(m.nVertex, 64), 64, 0, s>>>(d, m); c->launches++; }
  else if (m.kind == RX_BC_EULER_WALL) { k_bc_euler_wall<<<nblk(m.nVertex, 64), 64, 0, s>>>(c->o, d, m); c->launches++; }
}
int rx_bc(rx_ctx* c) {
  // First group (inlet, outlet, Euler wall).  Each of these kernels is one partial wave of long serial threads (a Stefan-Maxwell
  // solve per vertex), so markers that share no vertex — established at create — alternate between two streams; every
  // residual row is still written by one kernel only, in the reference's order per point.
  const bool fork = c->bcConcurrent;
  if (fork) { RX_CK(cudaEventRecord(c->evFork, c->stream)); RX_CK(cudaStreamWaitEvent(c->bcStream, c->evFork, 0)); }
  int nth = 0;
  for (const MarkerDev& m : c->markers)
    if (is_inlet(m.kind) || is_supersonic(m.kind) || m.kind == RX_BC_OUTLET || m.kind == RX_BC_EULER_WALL) {
      if (!m.nVertex) continue;
      launch_bc(c, m, (fork && (nth & 1)) ? c->bcStream : c->stream);
      ++nth;
    }
  if (fork) { RX_CK(cudaEventRecord(c->evJoin, c->bcStream)); RX_CK(cudaStreamWaitEvent(c->stream, c->evJoin, 0)); }
  for (const MarkerDev& m : c->markers)
    if (m.kind == RX_BC_ISOTHERMAL || m.kind == RX_BC_HEATFLUX) launch_bc(c, m, c->stream);
  RX_CK(cudaGetLastError());
  return RX_OK;
}
int rx_bc_marker(rx_ctx* c, int marker) {
  if (marker < 0 || marker >= (int)c->markers.size()) { c->err = "rx_bc_marker: marker index out of range"; return RX_ERR_ARG; }
  launch_bc(c, c->markers[marker], c->stream);
  RX_CK(cudaGetLastError());
  return RX_OK;
}
int rx_get_bc_state(rx_ctx* c, int marker, double* V) {
  if (marker < 0 || marker >= (int)c->markers.size()) { c->err = "rx_get_bc_state: marker index out of range"; return RX_ERR_ARG; }
  const MarkerDev& m = c->markers[marker];
  if (!m.nVertex) return RX_OK;
  const int np_ = c->nPrim;
  std::vector<double> soa((size_t)m.nVertex * np_);
  RX_CK(cudaMemcpyAsync(soa.data(), m.ghost, soa.size() * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  RX_CK(cudaStreamSynchronize(c->stream));
  for (int k = 0; k < m.nVertex; ++k)
    for (int v = 0; v < np_; ++v) V[(size_t)k * np_ + v] = soa[(size_t)v * m.nVertex + k];
  return RX_OK;
}
// The reference throws where a residual turns NaN or the temperature solve fails (solver_direct_reactive.cpp:2772, 2861, 5384;
// variable_direct_reactive.cpp:683, 724).  The kernels count those events in d.counters[1..4]; they stay set until a host read.
static int read_status(rx_ctx* c, unsigned long long* h5) {
  RX_CK(cudaMemcpyAsync(h5, c->d.counters, 5 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
  RX_CK(cudaStreamSynchronize(c->stream));
  if (h5[1] || h5[2] || h5[3] || h5[4]) RX_CK(cudaMemsetAsync(c->d.counters + 1, 0, 4 * sizeof(unsigned long long), c->stream));
  return RX_OK;
}
static int status_error(rx_ctx* c, const unsigned long long* h) {
  if (h[4]) { c->err = "Convergence not achieved for bisection method (temperature solve, " + std::to_string(h[4]) + " points)"; return RX_ERR_NAN; }
  if (h[1]) { c->err = "NaN found in the upwind residual"; return RX_ERR_NAN; }
  if (h[2]) { c->err = "NaN found in the viscous residual"; return RX_ERR_NAN; }
  if (h[3]) { c->err = "NaN found in the source residual"; return RX_ERR_NAN; }
  return RX_OK;
}
int rx_get_status(rx_ctx* c, int64_t* status5) {
  unsigned long long h[5];
  const int rc = read_status(c, h);
  if (rc) return rc;
  if (status5) for (int k = 0; k < 5; ++k) status5[k] = (int64_t)h[k];
  return status_error(c, h);
}

int rx_update_explicit(rx_ctx* c, double alpha, double* resRms) {
  const Dev& d = c->d;
  cudaStream_t s = c->stream;
  if (resRms) {   // norms wanted on the host: the location of the largest residual comes with them (rx_get_residual_max)
    RX_NS(k_update<NS, true><<<c->updBlocks, 128, 0, s>>>(d, alpha, c->partial, c->pmax, c->pidx));
    k_resmax_finish<<<c->nVar, 256, 0, s>>>(c->pmax, c->pidx, c->updBlocks, c->nVar, c->resMax, c->resMaxPoint);
    c->launches++;
    c->haveResMax = true;
  } else {
    RX_NS(k_update<NS, false><<<c->updBlocks, 128, 0, s>>>(d, alpha, c->partial, nullptr, nullptr));
  }
  k_rms_finish<<<c->nVar, 256, 0, s>>>(c->partial, c->updBlocks, c->nVar, d.nPoint, c->rms, c->sumsq);
  c->launches += 2;
  RX_CK(cudaGetLastError());
  if (resRms) {   // the host synchronises here anyway: the NaN / failed-solve flags come with the norms
    unsigned long long h[5];
    RX_CK(cudaMemcpyAsync(resRms, c->rms, c->nVar * sizeof(double), cudaMemcpyDeviceToHost, s));
    const int rc = read_status(c, h);
    if (rc) return rc;
    return status_error(c, h);
  }
  return RX_OK;
}

int rx_get_residual_max(rx_ctx* c, double* resMax, int64_t* point) {
  if (!c->haveResMax) { c->err = "rx_get_residual_max: no update with residual norms has run yet"; return RX_ERR_ARG; }
  std::vector<double> m(c->nVar); std::vector<int> ip(c->nVar);
  RX_CK(cudaMemcpyAsync(m.data(), c->resMax, c->nVar * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  RX_CK(cudaMemcpyAsync(ip.data(), c->resMaxPoint, c->nVar * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  RX_CK(cudaStreamSynchronize(c->stream));
  for (int v = 0; v < c->nVar; ++v) { if (resMax) resMax[v] = m[v]; if (point) point[v] = ip[v]; }
  return RX_OK;
}

int rx_residual_all(rx_ctx* c) {
  const Dev& d = c->d;
  cudaStream_t s = c->stream;
  int rc;
  if (c->o.viscous) {
    if (c->o.secondOrder && (rc = comm_join(c))) return rc;   // the MUSCL flux reads halo gradients and limiters
    launch_edge_flux<true, false>(c, s);
    if ((rc = comm_join(c))) return rc;                       // a forked gradient exchange (rx_iterate) must have landed before the viscous flux
    launch_edge_flux<false, true>(c, s);
    launch_accumulate_visc(c, s);
  } else {
    if ((rc = comm_join(c))) return rc;
    launch_edge_flux<true, false>(c, s);
    RX_NS(k_accumulate<NS, 1><<<nblk(d.nPoint, 128), 128, 0, s>>>(cc, c->o, d));
    RX_NS(k_source<NS><<<nblk(d.nPointDomain, 128), 128, 0, s>>>(cc, c->o, d));
    c->launches += 2;
  }
  RX_CK(cudaGetLastError());
  return RX_OK;
}

// average device time (ms, CUDA events on the context's stream) of `reps` launches of one kernel of the path, on the
// context's current state.  which: 0 viscous edge flux (convective if inviscid), 1 convective edge flux, 2 owner-computes
// accumulation + source, 3 cons->prim, 4 gradient, 5 source only, 6 explicit update, 7 laminar transport.
int rx_kernel_time(rx_ctx* c, int which, int reps, double* msAvg) {
  { const int rc = ensure_output(c); if (rc) return rc; }
  const Dev& d = c->d;
  cudaStream_t s = c->stream;
  cudaEvent_t e0, e1;
  RX_CK(cudaEventCreate(&e0)); RX_CK(cudaEventCreate(&e1));
  for (int pass = 0; pass < 2; ++pass) {   // pass 0 = warm-up
    const int n = pass == 0 ? 1 : reps;
    RX_CK(cudaEventRecord(e0, s));
    for (int r = 0; r < n; ++r) {
      switch (which) {
        case 0: if (c->o.viscous) launch_edge_flux<false, true>(c, s);
                else launch_edge_flux<true, false>(c, s);
                c->launches--; break;   // the helper counts its own launch
        case 1: launch_edge_flux<true, false>(c, s); c->launches--; break;
        case 2: if (c->o.viscous) { launch_accumulate_visc(c, s); c->launches--; }
                else RX_NS(k_accumulate<NS, 1><<<nblk(d.nPoint, 128), 128, 0, s>>>(cc, c->o, d)); break;
        case 3: RX_NS(k_primitive<NS><<<nblk(d.nPoint, 128), 128, 0, s>>>(cc, c->o, d, 1LL, 1)); break;
        case 7: RX_NS(k_transport<NS><<<nblk(d.nPoint, 128), 128, 0, s>>>(cc, d)); break;
        case 4: if (c->o.leastSquares) RX_NS(k_gradient_ls<NS><<<nblk(d.nPoint, 128), 128, 0, s>>>(d));
                else RX_NS(k_gradient_gg_edges<NS><<<nblk(d.nPointDomain, 128), 128, 0, s>>>(d)); break;
        case 5: RX_NS(k_source<NS><<<nblk(d.nPointDomain, 128), 128, 0, s>>>(cc, c->o, d)); break;
        case 8: if (!c->tpost) { cudaEventDestroy(e0); cudaEventDestroy(e1); c->err = "rx_kernel_time: not a RANS context"; return RX_ERR_ARG; }
                k_turb_gradient_ls<<<nblk(d.nPoint, 128), 128, 0, s>>>(d, c->tpost); break;
        case 9: if (!c->tpost) { cudaEventDestroy(e0); cudaEventDestroy(e1); c->err = "rx_kernel_time: not a RANS context"; return RX_ERR_ARG; }
                k_turb_blend<<<nblk(d.nPoint, 128), 128, 0, s>>>(d, c->tpost, c->sigmaK); break;
        case 6: RX_NS(k_update<NS, false><<<c->updBlocks, 128, 0, s>>>(d, 0.0, c->partial, nullptr, nullptr)); break;
        default: cudaEventDestroy(e0); cudaEventDestroy(e1); c->err = "rx_kernel_time: unknown kernel id"; return RX_ERR_ARG;
      }
      c->launches++;
    }
    RX_CK(cudaEventRecord(e1, s));
    RX_CK(cudaEventSynchronize(e1));
    RX_CK(cudaGetLastError());
    float ms = 0.f;
    RX_CK(cudaEventElapsedTime(&ms, e0, e1));
    if (pass == 1 && msAvg) *msAvg = (double)ms / n;
  }
  if (which == 3 && c->o.viscous) { RX_NS(k_transport<NS><<<nblk(d.nPoint, 128), 128, 0, s>>>(cc, d)); c->launches++; }   // mu[p] held T
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  return RX_OK;
}

int rx_get_edge_order(rx_ctx* c, int32_t* order, uint8_t* key) {
  const Dev& d = c->d;
  if (!d.eKey) { c->err = "rx_get_edge_order: inviscid context"; return RX_ERR_ARG; }
  RX_CK(cudaStreamSynchronize(c->stream));
  if (key) RX_CK(cudaMemcpy(key, d.eKey, (size_t)d.nEdge, cudaMemcpyDeviceToHost));
  if (order) {
    if (d.eOrder) RX_CK(cudaMemcpy(order, d.eOrder, (size_t)d.nEdge * sizeof(int), cudaMemcpyDeviceToHost));
    else for (long long e = 0; e < d.nEdge; ++e) order[e] = (int32_t)e;
  }
  return RX_OK;
}

int rx_iterate(rx_ctx* c, int64_t extIter, int nRK, const double* rkAlpha, double* resRms) {
  // With a communicator (rx_comm_init) the exchanges sit where the reference has them: gradients (and limiters) after every
  // Preprocessing — Set_MPI_Primitive_Gradient / _Limiter, solver_direct_reactive.cpp:4716-4760 — and the conserved variables after
  // every update — Set_MPI_Solution, :2442 / :2487; the norms of the last stage are then those of the whole job.
  int rc;
  const bool par = c->comm != nullptr;
  for (int rk = 0; rk < nRK; ++rk) {
    if ((rc = rx_preprocess(c, extIter, 0, nullptr))) return rc;
    // forked onto the communicator's stream: U_old copy, time step and the first-order convective flux do not read halo gradients
    if (par && (c->o.viscous || c->o.secondOrder) && (rc = halo_exchange_fork(c, 1))) return rc;
    if (par && c->o.secondOrder && c->o.limiter && (rc = halo_exchange_fork(c, 2))) return rc;
    if (rk == 0) {
      if ((rc = rx_set_old_solution(c))) return rc;
      if ((rc = rx_time_step(c))) return rc;
    }
    if ((rc = rx_residual_all(c))) return rc;
    if (c->o.dualTime && (rc = rx_residual_dual_time(c))) return rc;   // CIntegration::Space_Integration :86-87: after the source term, before the boundary conditions
    if ((rc = rx_bc(c))) return rc;
    const bool last = rk == nRK - 1;
    if (par && last && resRms) {
      if ((rc = rx_update_explicit(c, rkAlpha ? rkAlpha[rk] : 1.0, nullptr))) return rc;
      if ((rc = rx_halo_exchange(c, 0))) return rc;
      if ((rc = rx_comm_allreduce_rms(c, resRms))) return rc;
    } else {
      if ((rc = rx_update_explicit(c, rkAlpha ? rkAlpha[rk] : 1.0, last ? resRms : nullptr))) return rc;
      if (par && (rc = rx_halo_exchange(c, 0))) return rc;
    }
  }
  if ((rc = preprocess_primitive(c, extIter, 1))) return rc;   // the Output pass (integration_time.cpp:110-112): temperature solve now, the rest on demand
  c->outputPending = true;
  return RX_OK;
}

int rx_source_points_dev(rx_ctx* c, int64_t n, const double* T, const double* rho, const double* Y, const double* omT, const double* vol, double* res) {
  RX_NS(k_source_points<NS><<<nblk(n, 128), 128, 0, c->stream>>>(cc, c->o.rans, c->o.cMu, c->o.pasrLb, (long long)n, T, rho, Y, omT, vol, res));
  c->launches++;
  RX_CK(cudaGetLastError());
  return RX_OK;
}


// host-pointer form: builds a throw-away context-free parameter block, stages the inputs (AoS Y[n][Ns] -> SoA on the host side
// of the copy), runs k_source_points and copies the result back (res[n][Ns]).
int rx_source_points(const rx_chem* ch, int rans, double cMu, double pasrLb, int64_t n, const double* T, const double* rho,
                     const double* Y, const double* omT, const double* vol, double* res, int device) {
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { g_create_error = "rx_source_points: no CUDA device (no CPU path)"; return RX_ERR_NO_DEVICE; }
  if (!rx_supported_species(ch->Ns) || ch->Nr > NRMAX) { g_create_error = "rx_source_points: unsupported mechanism size"; return RX_ERR_UNSUPPORTED; }
  rx_ctx tmp; rx_ctx* c = &tmp;
  c->device = device;
  const int Ns = ch->Ns;
  c->Ns = Ns;   // the species-count dispatch of the launch below reads it
  RX_CK(cudaSetDevice(device));
  int rc = fill_chem(c, ch);
  double *dT = dalloc<double>(c, n), *dr = dalloc<double>(c, n), *dY = dalloc<double>(c, n * Ns), *dO = dalloc<double>(c, n),
         *dV = dalloc<double>(c, n), *dR = dalloc<double>(c, n * Ns), *dA = dalloc<double>(c, n * Ns);
  if (rc == RX_OK && !(dT && dr && dY && dO && dV && dR && dA)) rc = RX_ERR_CUDA;
  if (rc == RX_OK) {
    cudaMemcpy(dT, T, n * sizeof(double), cudaMemcpyHostToDevice);
    cudaMemcpy(dr, rho, n * sizeof(double), cudaMemcpyHostToDevice);
    cudaMemcpy(dV, vol, n * sizeof(double), cudaMemcpyHostToDevice);
    if (omT) cudaMemcpy(dO, omT, n * sizeof(double), cudaMemcpyHostToDevice);
    cudaMemcpy(dA, Y, n * Ns * sizeof(double), cudaMemcpyHostToDevice);
    k_aos_to_soa<<<nblk(n * Ns, 256), 256>>>(dA, dY, n, Ns, n);
    RX_NS(k_source_points<NS><<<nblk(n, 128), 128>>>(cc, rans, cMu, pasrLb, (long long)n, dT, dr, dY, omT ? dO : nullptr, dV, dR));
    k_soa_to_aos<<<nblk(n * Ns, 256), 256>>>(dR, dA, n, Ns, n);
    if (cudaMemcpy(res, dA, n * Ns * sizeof(double), cudaMemcpyDeviceToHost) != cudaSuccess || cudaGetLastError() != cudaSuccess) {
      g_create_error = "rx_source_points: kernel or copy failed"; rc = RX_ERR_CUDA;
    }
  }
  for (void* p : c->allocs) cudaFree(p);
  return rc;
}

// CPhysicalGeometry::ComputeWall_Distance for host arrays (no context needed): dist[p] = distance to the nearest wall node
int rx_wall_distance(int64_t nPoint, const double* xy, int64_t nWall, const double* wallXy, double* dist, int device) {
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { g_create_error = "rx_wall_distance: no CUDA device (no CPU path)"; return RX_ERR_NO_DEVICE; }
  if (nPoint < 0 || nWall < 0 || !xy || !dist || (nWall > 0 && !wallXy)) { g_create_error = "rx_wall_distance: bad arguments"; return RX_ERR_ARG; }
  if (nPoint == 0) return RX_OK;
  if (cudaSetDevice(device) != cudaSuccess) { g_create_error = "rx_wall_distance: cudaSetDevice failed"; return RX_ERR_CUDA; }
  double *dXy = nullptr, *dW = nullptr, *dD = nullptr;
  int rc = RX_OK;
  if (cudaMalloc(&dXy, (size_t)nPoint * 2 * sizeof(double)) != cudaSuccess || cudaMalloc(&dD, (size_t)nPoint * sizeof(double)) != cudaSuccess ||
      cudaMalloc(&dW, (size_t)std::max<int64_t>(nWall, 1) * 2 * sizeof(double)) != cudaSuccess) rc = RX_ERR_CUDA;
  if (rc == RX_OK) {
    cudaMemcpy(dXy, xy, (size_t)nPoint * 2 * sizeof(double), cudaMemcpyHostToDevice);
    if (nWall > 0) cudaMemcpy(dW, wallXy, (size_t)nWall * 2 * sizeof(double), cudaMemcpyHostToDevice);
    k_wall_distance<<<nblk(nPoint, 256), 256>>>((long long)nPoint, dXy, (long long)nWall, dW, dD);
    if (cudaMemcpy(dist, dD, (size_t)nPoint * sizeof(double), cudaMemcpyDeviceToHost) != cudaSuccess || cudaGetLastError() != cudaSuccess) {
      g_create_error = "rx_wall_distance: kernel or copy failed"; rc = RX_ERR_CUDA;
    }
  } else g_create_error = "rx_wall_distance: out of device memory";
  cudaFree(dXy); cudaFree(dW); cudaFree(dD);
  return rc;
}

// ---- in-library communication (rx_comm.cuh)
int rx_comm_unique_id(void* id128) {
  std::string why;
  NcclApi* a = nccl_api(why);
  if (!a) { g_create_error = "rx_comm_unique_id: " + why; return RX_ERR_UNSUPPORTED; }
  nccl_uid id;
  const int e = a->GetUniqueId(&id);
  if (e) { g_create_error = std::string("ncclGetUniqueId: ") + a->GetErrorString(e); return RX_ERR_CUDA; }
  memcpy(id128, &id, sizeof id);
  return RX_OK;
}
int rx_comm_init(rx_ctx* c, const void* id128, int rank, int nRanks, int nNeighbors, const int* neighborRank,
                 const int64_t* sendPtr, const int64_t* sendPoints, const int64_t* recvPtr, const int64_t* recvPoints) {
  if (c->comm) { c->err = "rx_comm_init: the context already has a communicator"; return RX_ERR_ARG; }
  if (!id128 || rank < 0 || rank >= nRanks || nNeighbors < 0 || (nNeighbors && (!neighborRank || !sendPtr || !recvPtr))) { c->err = "rx_comm_init: bad arguments"; return RX_ERR_ARG; }
  std::string why;
  NcclApi* a = nccl_api(why);
  if (!a) { c->err = "rx_comm_init: " + why; return RX_ERR_UNSUPPORTED; }
  RX_CK(cudaSetDevice(c->device));
  Comm* m = new Comm();
  m->rank = rank; m->nRanks = nRanks;
  m->nbRank.assign(neighborRank, neighborRank + nNeighbors);
  m->sendPtr.assign(1, 0); m->recvPtr.assign(1, 0);
  for (int k = 0; k < nNeighbors; ++k) { m->sendPtr.push_back(sendPtr[k + 1]); m->recvPtr.push_back(recvPtr[k + 1]); }
  const long long nS = m->sendPtr.back(), nR = m->recvPtr.back();
  for (long long k = 0; k < nS; ++k) if (sendPoints[k] < 0 || sendPoints[k] >= c->d.nPoint) { delete m; c->err = "rx_comm_init: send point out of range"; return RX_ERR_ARG; }
  for (long long k = 0; k < nR; ++k) if (recvPoints[k] < 0 || recvPoints[k] >= c->d.nPoint) { delete m; c->err = "rx_comm_init: receive point out of range"; return RX_ERR_ARG; }
  m->sendPts = dupload(c, std::vector<long long>(sendPoints, sendPoints + nS));
  m->recvPts = dupload(c, std::vector<long long>(recvPoints, recvPoints + nR));
  m->sendBuf = dalloc<double>(c, (size_t)std::max<long long>(nS, 1) * HALO_MAXW);
  m->recvBuf = dalloc<double>(c, (size_t)std::max<long long>(nR, 1) * HALO_MAXW);
  m->redBuf = dalloc<double>(c, 32);
  if (c->allocFailed || !m->sendPts || !m->recvPts || !m->sendBuf || !m->recvBuf || !m->redBuf) { delete m; c->err = "rx_comm_init: device allocation failed"; return RX_ERR_CUDA; }
  nccl_uid id;
  memcpy(&id, id128, sizeof id);
  const int e = a->CommInitRank(&m->comm, nRanks, id, rank);
  if (e) { c->err = std::string("ncclCommInitRank: ") + a->GetErrorString(e); delete m; return RX_ERR_CUDA; }
  if (cudaStreamCreateWithFlags(&m->cs, cudaStreamNonBlocking) != cudaSuccess || cudaEventCreateWithFlags(&m->evReady, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&m->evDone, cudaEventDisableTiming) != cudaSuccess) { a->CommDestroy(m->comm); delete m; c->err = "rx_comm_init: stream / event creation failed"; return RX_ERR_CUDA; }
  c->comm = m;
  return RX_OK;
}
// pack -> grouped ncclSend / ncclRecv with every neighbour -> unpack, on the context's stream (Set_MPI_* of the reference)
static int halo_exchange_on(rx_ctx* c, int what, cudaStream_t s);
int rx_halo_exchange(rx_ctx* c, int what) {
  if (!c->comm) { c->err = "rx_halo_exchange: no communicator (rx_comm_init)"; return RX_ERR_ARG; }
  if (what < 0 || what > 4 || (what >= 3 && !c->tpost)) { c->err = "rx_halo_exchange: bad `what`"; return RX_ERR_ARG; }
  return halo_exchange_on(c, what, c->stream);
}
// the same exchange forked onto the communicator's stream (everything enqueued so far on the context's stream is waited for); the
// context's stream continues and joins with comm_join() before the first kernel that reads the exchanged halo values
static int halo_exchange_fork(rx_ctx* c, int what) {
  Comm* m = c->comm;
  if (!m->pending) {
    RX_CK(cudaEventRecord(m->evReady, c->stream));
    RX_CK(cudaStreamWaitEvent(m->cs, m->evReady, 0));
  }
  const int rc = halo_exchange_on(c, what, m->cs);
  m->pending = true;
  return rc;
}
static int comm_join(rx_ctx* c) {
  Comm* m = c->comm;
  if (!m || !m->pending) return RX_OK;
  RX_CK(cudaEventRecord(m->evDone, m->cs));
  RX_CK(cudaStreamWaitEvent(c->stream, m->evDone, 0));
  m->pending = false;
  return RX_OK;
}
static double* halo_field(rx_ctx* c, int what);
static int halo_exchange_on(rx_ctx* c, int what, cudaStream_t s) {
  return halo_exchange_field(c, halo_field(c, what), rx_halo_width(c, what), what == 1 && c->o.viscous, s);
}
// `nv` columns (stride np) of any point field: the turbulence step also exchanges its Krylov / LU-SGS vectors this way
// (CSysMatrix::SendReceive_Solution, matrix_structure.cpp:794)
static int halo_exchange_field(rx_ctx* c, double* field, int nv, bool vortStrain, cudaStream_t s) {
  Comm* m = c->comm;
  std::string why;
  NcclApi* a = nccl_api(why);
  const int nNb = (int)m->nbRank.size();
  for (int k = 0; k < nNb; ++k) {
    const long long n = m->sendPtr[k + 1] - m->sendPtr[k];
    if (n <= 0) continue;
    k_pack<<<nblk(n * nv, 256), 256, 0, s>>>(field, c->d.np, nv, n, m->sendPts + m->sendPtr[k], m->sendBuf + (size_t)m->sendPtr[k] * nv);
    c->launches++;
  }
  RX_CK(cudaGetLastError());
  int e = a->GroupStart();
  for (int k = 0; k < nNb && !e; ++k) {
    const long long ns = m->sendPtr[k + 1] - m->sendPtr[k], nr = m->recvPtr[k + 1] - m->recvPtr[k];
    if (ns > 0) e = a->Send(m->sendBuf + (size_t)m->sendPtr[k] * nv, (size_t)ns * nv, NCCL_FLOAT64, m->nbRank[k], m->comm, s);
    if (nr > 0 && !e) e = a->Recv(m->recvBuf + (size_t)m->recvPtr[k] * nv, (size_t)nr * nv, NCCL_FLOAT64, m->nbRank[k], m->comm, s);
    m->bytesMoved += (ns + nr) * nv * 8;
  }
  if (!e) e = a->GroupEnd();
  if (e) { c->err = std::string("rx_halo_exchange: NCCL: ") + a->GetErrorString(e); return RX_ERR_CUDA; }
  for (int k = 0; k < nNb; ++k) {
    const long long n = m->recvPtr[k + 1] - m->recvPtr[k];
    if (n <= 0) continue;
    k_unpack<<<nblk(n * nv, 256), 256, 0, s>>>(field, c->d.np, nv, n, m->recvPts + m->recvPtr[k], m->recvBuf + (size_t)m->recvPtr[k] * nv);
    c->launches++;
    if (vortStrain) { RX_NS(k_vort_strain_points<NS><<<nblk(n, 128), 128, 0, s>>>(c->d, n, m->recvPts + m->recvPtr[k])); c->launches++; }
  }
  RX_CK(cudaGetLastError());
  return RX_OK;
}
// in-place sum over the ranks of n <= 8 doubles on the device (dotProd / SetResidual_RMS with MPI: vector_structure.cpp:397-419)
static int comm_allreduce_sum(rx_ctx* c, double* dev, int n) {
  Comm* m = c->comm;
  std::string why;
  NcclApi* a = nccl_api(why);
  const int e = a->AllReduce(dev, dev, (size_t)n, NCCL_FLOAT64, NCCL_SUM, m->comm, c->stream);
  if (e) { c->err = std::string("NCCL all-reduce (sum): ") + a->GetErrorString(e); return RX_ERR_CUDA; }
  return RX_OK;
}
// norms of the whole job from the partitions' sums of squares (SetResidual_RMS with MPI, solver_structure.cpp:188-260): one all-reduce
int rx_comm_allreduce_rms(rx_ctx* c, double* resRms) {
  Comm* m = c->comm;
  if (!m) { c->err = "rx_comm_allreduce_rms: no communicator (rx_comm_init)"; return RX_ERR_ARG; }
  std::string why;
  NcclApi* a = nccl_api(why);
  cudaStream_t s = c->stream;
  k_rms_pack<<<1, 32, 0, s>>>(c->sumsq, c->nVar, (double)c->d.nPointDomain, m->redBuf);
  const int e = a->AllReduce(m->redBuf, m->redBuf, (size_t)c->nVar + 1, NCCL_FLOAT64, NCCL_SUM, m->comm, s);
  if (e) { c->err = std::string("rx_comm_allreduce_rms: NCCL: ") + a->GetErrorString(e); return RX_ERR_CUDA; }
  k_rms_global<<<1, 32, 0, s>>>(m->redBuf, c->nVar, c->rms);
  c->launches += 2;
  RX_CK(cudaGetLastError());
  if (resRms) {
    unsigned long long h[5];
    RX_CK(cudaMemcpyAsync(resRms, c->rms, c->nVar * sizeof(double), cudaMemcpyDeviceToHost, s));
    const int rc = read_status(c, h);
    if (rc) return rc;
    return status_error(c, h);
  }
  return RX_OK;
}

int rx_halo_width(rx_ctx* c, int what) { return what == 0 ? c->nVar : what == 1 ? c->nGrad * 2 : (what == 2 || what == 3) ? 4 : what == 4 ? 2 : 0; }   // 2: limiter, 3: grad k, grad omega, 4: k, omega
static int halo_check(rx_ctx* c, int what, const void* pts, const void* buf) {
  if (what < 0 || what > 4) { c->err = "rx_halo_pack/unpack: `what` must be 0 (U), 1 (gradients), 2 (limiters), 3 (grad k, grad omega) or 4 (k, omega)"; return RX_ERR_ARG; }
  if (what >= 3 && !c->tpost) { c->err = "rx_halo_pack/unpack: what = 3, 4 need a RANS context"; return RX_ERR_ARG; }
  if (!pts || !buf) { c->err = "rx_halo_pack/unpack: null device pointer"; return RX_ERR_ARG; }
  return RX_OK;
}
static double* halo_field(rx_ctx* c, int what) { return what == 0 ? c->d.U : what == 1 ? c->d.grad : what == 2 ? c->d.lim : what == 3 ? c->tpost + (size_t)TP_GK * c->d.np : c->tpost + (size_t)TP_K * c->d.np; }
int rx_halo_pack(rx_ctx* c, int what, int64_t n, const int64_t* pts, double* buf) {
  if (n <= 0) return RX_OK;
  if (what == 1) { const int rc = ensure_output(c); if (rc) return rc; }
  { const int rc = halo_check(c, what, pts, buf); if (rc) return rc; }
  const int nv = rx_halo_width(c, what);
  k_pack<<<nblk(n * nv, 256), 256, 0, c->stream>>>(halo_field(c, what), c->d.np, nv, n, (const long long*)pts, buf);
  c->launches++;
  RX_CK(cudaGetLastError());
  return RX_OK;
}
int rx_halo_unpack(rx_ctx* c, int what, int64_t n, const int64_t* pts, const double* buf) {
  if (n <= 0) return RX_OK;
  { const int rc = halo_check(c, what, pts, buf); if (rc) return rc; }
  const int nv = rx_halo_width(c, what);
  k_unpack<<<nblk(n * nv, 256), 256, 0, c->stream>>>(halo_field(c, what), c->d.np, nv, n, (const long long*)pts, buf);
  c->launches++;
  if (what == 1 && c->o.viscous) {   // SetVorticity / SetStrainMag of the received points (:4716-4736 run after Set_MPI_Primitive_Gradient)
    RX_NS(k_vort_strain_points<NS><<<nblk(n, 128), 128, 0, c->stream>>>(c->d, (long long)n, (const long long*)pts));
    c->launches++;
  }
  RX_CK(cudaGetLastError());
  return RX_OK;
}

}  // extern "C"

// diagnostic: the glibc-exact exp / pow / log of rx_libm.cuh over host arrays (tests/test_gpu_libm.py compares with the host's libm)
namespace rx {
__global__ void k_libm_eval(int which, long long n, const double* __restrict__ x, const double* __restrict__ y, double* __restrict__ out) {
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= n) return;
  out[t] = which == 0 ? gl_exp(x[t]) : which == 1 ? gl_pow(x[t], y[t]) : gl_log(x[t]);
}
}  // namespace rx
extern "C" int rx_libm_eval(int which, int64_t n, const double* x, const double* y, double* out, int device) {
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { g_create_error = "rx_libm_eval: no CUDA device (no CPU path)"; return RX_ERR_NO_DEVICE; }
  if (which < 0 || which > 2 || n < 0 || !x || !out || (which == 1 && !y)) { g_create_error = "rx_libm_eval: bad arguments"; return RX_ERR_ARG; }
  if (n == 0) return RX_OK;
  if (cudaSetDevice(device) != cudaSuccess) { g_create_error = "rx_libm_eval: cudaSetDevice failed"; return RX_ERR_CUDA; }
  double *dx = nullptr, *dy = nullptr, *dout = nullptr;
  int rc = RX_OK;
  const size_t nb = (size_t)n * sizeof(double);
  if (cudaMalloc(&dx, nb) != cudaSuccess || cudaMalloc(&dout, nb) != cudaSuccess || (which == 1 && cudaMalloc(&dy, nb) != cudaSuccess)) rc = RX_ERR_CUDA;
  if (rc == RX_OK && (cudaMemcpy(dx, x, nb, cudaMemcpyHostToDevice) != cudaSuccess || (which == 1 && cudaMemcpy(dy, y, nb, cudaMemcpyHostToDevice) != cudaSuccess))) rc = RX_ERR_CUDA;
  if (rc == RX_OK) {
    rx::k_libm_eval<<<(unsigned)((n + 255) / 256), 256>>>(which, (long long)n, dx, dy, dout);
    if (cudaMemcpy(out, dout, nb, cudaMemcpyDeviceToHost) != cudaSuccess || cudaGetLastError() != cudaSuccess) rc = RX_ERR_CUDA;
  }
  if (rc != RX_OK) g_create_error = "rx_libm_eval: allocation, copy or kernel failed";
  cudaFree(dx); cudaFree(dy); cudaFree(dout);
  return rc;
}

// diagnostic hook: the two device forms of Eigen::ColPivHouseholderQR<MatrixXd>::solve (serial colpiv_qr_solve<9, 2, 1>, used at boundary
// vertices, and the two-lanes-per-system pair_colpiv_qr_solve<9> of the viscous edge kernel) over host arrays of 9 x 9 systems
namespace rx {
__global__ void __launch_bounds__(VP_BS) k_qr_eval(long long n, const double* __restrict__ A, const double* __restrict__ b,
                                                   double* __restrict__ xp) {
  constexpr int NS = 9;
  extern __shared__ double vp_sm[];
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  PairLane L;
  L.l = (int)(threadIdx.x & 1);
  L.odd = (int)(threadIdx.x & 31) | 1;
  const int l = L.l;
  long long e = t >> 1;
  const bool keep = e < n;
  if (!keep) e = n - 1;   // whole warps take part in the pair routine's warp synchronisations
  double* const smPair = vp_sm + threadIdx.x - l;
  const PairMat<NS> Mt{smPair};
  const double* Ae = A + e * NS * NS;
  const double* be = b + e * 2 * NS;
  for (int i = l; i < NS; i += 2) {
    for (int j = 0; j < NS; ++j) Mt(i, j) = Ae[j * NS + i];
    for (int dd = 0; dd < 2; ++dd) Mt(i, NS + dd) = be[dd * NS + i];
  }
  double MG[2][5];
  pair_colpiv_qr_solve<NS>(L, smPair, MG);
  if (keep) {
    for (int dd = 0; dd < 2; ++dd) {
      for (int q = 0; q < 4; ++q) xp[(e * 2 + dd) * NS + 2 * q + l] = MG[dd][q];
      if (l) xp[(e * 2 + dd) * NS + 8] = MG[dd][4];
    }
  }
}
__global__ void k_qr_eval_serial(long long n, const double* __restrict__ A, const double* __restrict__ b, double* __restrict__ xs) {
  constexpr int NS = 9;
  const long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (e >= n) return;
  double Al[NS * NS], bl[2][NS], xl[2][NS];
  for (int k = 0; k < NS * NS; ++k) Al[k] = A[e * NS * NS + k];
  for (int dd = 0; dd < 2; ++dd) for (int i = 0; i < NS; ++i) bl[dd][i] = b[(e * 2 + dd) * NS + i];
  colpiv_qr_solve<NS, 2, 1>(Al, bl, xl);
  for (int dd = 0; dd < 2; ++dd) for (int i = 0; i < NS; ++i) xs[(e * 2 + dd) * NS + i] = xl[dd][i];
}
}  // namespace rx
extern "C" int rx_qr_eval(int64_t n, const double* A, const double* b, double* xSerial, double* xPair, int device) {
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { g_create_error = "rx_qr_eval: no CUDA device (no CPU path)"; return RX_ERR_NO_DEVICE; }
  if (n < 0 || !A || !b || !xSerial || !xPair) { g_create_error = "rx_qr_eval: bad arguments"; return RX_ERR_ARG; }
  if (n == 0) return RX_OK;
  if (cudaSetDevice(device) != cudaSuccess) { g_create_error = "rx_qr_eval: cudaSetDevice failed"; return RX_ERR_CUDA; }
  double *dA = nullptr, *db = nullptr, *dxs = nullptr, *dxp = nullptr;
  int rc = RX_OK;
  const size_t nA = (size_t)n * 81 * sizeof(double), nb = (size_t)n * 18 * sizeof(double);
  if (cudaMalloc(&dA, nA) != cudaSuccess || cudaMalloc(&db, nb) != cudaSuccess || cudaMalloc(&dxs, nb) != cudaSuccess || cudaMalloc(&dxp, nb) != cudaSuccess) rc = RX_ERR_CUDA;
  if (rc == RX_OK && (cudaMemcpy(dA, A, nA, cudaMemcpyHostToDevice) != cudaSuccess || cudaMemcpy(db, b, nb, cudaMemcpyHostToDevice) != cudaSuccess)) rc = RX_ERR_CUDA;
  if (rc == RX_OK) {
    const size_t smem = (size_t)rx::VP_SMW_RANS * rx::VP_BS * sizeof(double);
    if (cudaFuncSetAttribute(rx::k_qr_eval, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) rc = RX_ERR_CUDA;
    if (rc == RX_OK) {
      rx::k_qr_eval<<<(unsigned)((2 * n + rx::VP_BS - 1) / rx::VP_BS), rx::VP_BS, smem>>>((long long)n, dA, db, dxp);
      rx::k_qr_eval_serial<<<(unsigned)((n + 63) / 64), 64>>>((long long)n, dA, db, dxs);
      if (cudaMemcpy(xSerial, dxs, nb, cudaMemcpyDeviceToHost) != cudaSuccess || cudaMemcpy(xPair, dxp, nb, cudaMemcpyDeviceToHost) != cudaSuccess ||
          cudaGetLastError() != cudaSuccess) rc = RX_ERR_CUDA;
    }
  }
  if (rc != RX_OK) g_create_error = "rx_qr_eval: allocation, copy or kernel failed";
  cudaFree(dA); cudaFree(db); cudaFree(dxs); cudaFree(dxp);
  return rc;
}

// SU2 native restart files from / into the device-resident state (COutput::SetRestart, Load_Restart)
#include "rx_restart.cuh"

// allocation-free residual norms over a static device scratch array (its own entry point; no existing path uses it)
#include "rx_noalloc.cuh"
