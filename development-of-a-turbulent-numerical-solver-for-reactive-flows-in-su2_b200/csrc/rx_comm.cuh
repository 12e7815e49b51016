// rx_comm.cuh — halo exchange INSIDE the library, over NCCL (one process per GPU; NVLink 5 / NVSwitch on a B200 node).
//
// Replaces the transport half of CReactiveEulerSolver::Set_MPI_Solution (solver_direct_reactive.cpp:1530-1636),
// Set_MPI_Primitive_Gradient (:1877-1993), Set_MPI_Primitive_Limiter (:1756-1870) and, for the device SST Postprocessing,
// Set_MPI_Solution_Gradient (solver_direct_turbulent.cpp:227): per neighbour ONE message buf[var][vertex] — the reference's own
// packing — sent with ncclSend / ncclRecv grouped over the neighbours, packed and unpacked by the kernels of rx_kernels2.cuh, all on
// the context's stream so that a whole iteration (rx_iterate) runs without the host in the loop.  The path has no other
// collective; the residual norms of a partition are combined by one ncclAllReduce of nVar + 1 doubles (rx_comm_allreduce_rms).
//
// NCCL is bound at run time (dlopen of libnccl.so.2: the copy the process already has — torch's — or the system one), so the
// library loads and runs single-GPU without it.  Prototypes below restate the public NCCL C API (nccl.h) for the eight calls used.
#pragma once
#include <dlfcn.h>

namespace rx {

typedef struct ncclComm* nccl_comm_t;
struct nccl_uid { char internal[128]; };
enum { NCCL_FLOAT64 = 8, NCCL_SUM = 0, NCCL_MIN = 3 };   // ncclDataType_t / ncclRedOp_t values of nccl.h

struct NcclApi {
  void* h = nullptr;
  int (*GetUniqueId)(nccl_uid*) = nullptr;
  int (*CommInitRank)(nccl_comm_t*, int, nccl_uid, int) = nullptr;
  int (*CommDestroy)(nccl_comm_t) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  int (*Send)(const void*, size_t, int, int, nccl_comm_t, cudaStream_t) = nullptr;
  int (*Recv)(void*, size_t, int, int, nccl_comm_t, cudaStream_t) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, nccl_comm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};

inline NcclApi* nccl_api(std::string& why) {
  static NcclApi api;
  static bool tried = false;
  if (api.h) return &api;
  if (tried) { why = "libnccl.so.2 could not be loaded"; return nullptr; }
  tried = true;
  const char* env = getenv("RX_NCCL_LIB");
  const char* names[] = {env, "libnccl.so.2", "libnccl.so"};
  for (const char* n : names) {
    if (!n) continue;
    api.h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (api.h) break;
  }
  if (!api.h) { why = std::string("dlopen(libnccl.so.2): ") + (dlerror() ? dlerror() : "not found") + " (set RX_NCCL_LIB)"; return nullptr; }
  bool ok = true;
  auto sym = [&](const char* s) { void* p = dlsym(api.h, s); ok = ok && p; return p; };
  api.GetUniqueId = (int (*)(nccl_uid*))sym("ncclGetUniqueId");
  api.CommInitRank = (int (*)(nccl_comm_t*, int, nccl_uid, int))sym("ncclCommInitRank");
  api.CommDestroy = (int (*)(nccl_comm_t))sym("ncclCommDestroy");
  api.GroupStart = (int (*)())sym("ncclGroupStart");
  api.GroupEnd = (int (*)())sym("ncclGroupEnd");
  api.Send = (int (*)(const void*, size_t, int, int, nccl_comm_t, cudaStream_t))sym("ncclSend");
  api.Recv = (int (*)(void*, size_t, int, int, nccl_comm_t, cudaStream_t))sym("ncclRecv");
  api.AllReduce = (int (*)(const void*, void*, size_t, int, int, nccl_comm_t, cudaStream_t))sym("ncclAllReduce");
  api.GetErrorString = (const char* (*)(int))sym("ncclGetErrorString");
  if (!ok) { why = "libnccl.so.2 lacks an expected symbol"; dlclose(api.h); api.h = nullptr; return nullptr; }
  return &api;
}

// neighbour lists and buffers of one context
struct Comm {
  nccl_comm_t comm = nullptr;
  int rank = 0, nRanks = 1;
  std::vector<int> nbRank;
  std::vector<long long> sendPtr, recvPtr;   // [nNb + 1] offsets into the concatenated point lists
  long long* sendPts = nullptr;              // device, concatenated over neighbours
  long long* recvPts = nullptr;
  double* sendBuf = nullptr;                 // device, HALO_MAXW doubles per listed point
  double* recvBuf = nullptr;
  double* redBuf = nullptr;                  // device, nVar + 1 doubles for the norm reduction
  long long bytesMoved = 0;
  // overlap: the gradient / limiter exchange of rx_iterate runs on its own stream while the context's stream goes on with the
  // kernels that do not read halo gradients (U_old copy, time step, first-order convective flux); joined before the viscous kernel
  cudaStream_t cs = nullptr;
  cudaEvent_t evReady = nullptr, evDone = nullptr;
  bool pending = false;
};
constexpr int HALO_MAXW = 26;   // widest exchange: 13 gradient rows x 2

}  // namespace rx
