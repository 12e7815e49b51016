// rx_noalloc.cuh — rx_residual_rms_static: the residual-norm reduction of rx_update_explicit (k_update's block partials +
// k_rms_finish, rx_kernels2.cuh) as ONE context-free entry point that allocates nothing.  No counterpart in the reference;
// added for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
// serving loops, harnesses that forbid it).
//
// The block partials — the scratch rx_update_explicit keeps in rx_ctx::partial (cudaMalloc at rx_create) — live here in a
// file-scope __device__ array of fixed size: RX_STATIC_RMS_MAX_POINTS / 128 blocks x RX_STATIC_RMS_MAX_VARS doubles (+ one row) = 4 MiB,
// held by every GPU that loads the library until the process exits.  Host code never uses the array's name as an address
// (in host code it is only a stand-in): the device address is taken once per device with cudaGetSymbolAddress, that device
// current, and handed to the kernels.
//
// Every call shares the array and it keeps the last call's contents.  Nothing has to start at zero: a call writes each of
// the nBlocks x nVar partials it later reads (blocks past the last point write the sum of an empty block, 0.0) and reads no
// other slot, so stale contents of an earlier, larger call are never seen.  Calls that overlap on the GPU (two streams, two
// host threads, two processes' worth of graphs replayed concurrently on one device) race on the array: serialise them.
//
// Same partial sums in the same order as rx_update_explicit (128 points per block, shuffle tree per warp, (w0+w1)+(w2+w3),
// then 256 strided sums + tree per variable), so the norms are bit-equal to that entry point's for the same LinSysRes.
#pragma once
#include <cuda_runtime.h>
#include <mutex>

namespace rx {

constexpr int NA_BS = 128;
constexpr long long NA_MAX_POINTS = RX_STATIC_RMS_MAX_POINTS;
constexpr int NA_MAX_VARS = RX_STATIC_RMS_MAX_VARS;
constexpr long long NA_MAX_BLOCKS = NA_MAX_POINTS / NA_BS;
constexpr int NA_MAX_DEVICES = 64;

static __device__ __align__(256) double g_naPartial[NA_MAX_BLOCKS * NA_MAX_VARS + NA_MAX_VARS];   // block partials, then the nVar sums of squares k_rms_finish also leaves

// per-block sums of squares of R[v * ld + p], p < n: the reduction half of k_update
__global__ void __launch_bounds__(NA_BS) k_sumsq_partial_static(const double* __restrict__ R, long long n, long long ld, int nv, double* __restrict__ partial) {
  const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  __shared__ double sh[4][NA_MAX_VARS];
  for (int v = 0; v < nv; ++v) {
    double s = 0.0;
    if (p < n) { const double Res = R[v * ld + p] + 0.0; s = Res * Res; }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) s += __shfl_down_sync(0xffffffffu, s, off);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5][v] = s;
  }
  __syncthreads();
  if ((int)threadIdx.x < nv) {
    const int v = threadIdx.x;
    partial[(size_t)blockIdx.x * nv + v] = ((sh[0][v] + sh[1][v]) + (sh[2][v] + sh[3][v]));
  }
}

// the device address of g_naPartial on `device` (current on entry), taken once per device
static int na_scratch(int device, double** out, const char** why) {
  static std::mutex mu;
  static double* ptr[NA_MAX_DEVICES] = {};
  std::lock_guard<std::mutex> lock(mu);
  if (!ptr[device]) {
    double* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_naPartial);
    if (e != cudaSuccess || !p) { *why = cudaGetErrorString(e); return RX_ERR_CUDA; }
    ptr[device] = p;
  }
  *out = ptr[device];
  return RX_OK;
}

}  // namespace rx

extern "C" int rx_residual_rms_static(const double* R_dev, int64_t nPointDomain, int64_t ld, int nVar, int64_t nPointGlobal,
                                      double* rms_dev, int device, void* cudaStream) {
  using namespace rx;
  char msg[256];
  if (!R_dev || !rms_dev || nPointDomain < 0 || ld < nPointDomain || nVar < 1 || nPointGlobal < 1) {
    g_create_error = "rx_residual_rms_static: bad arguments"; return RX_ERR_ARG;
  }
  if (nPointDomain > NA_MAX_POINTS || nVar > NA_MAX_VARS) {
    snprintf(msg, sizeof msg, "rx_residual_rms_static: %lld points x %d variables exceed the static scratch (at most %lld points, %d variables); use rx_update_explicit",
             (long long)nPointDomain, nVar, NA_MAX_POINTS, NA_MAX_VARS);
    g_create_error = msg; return RX_ERR_UNSUPPORTED;
  }
  if (device < 0 || device >= NA_MAX_DEVICES) { g_create_error = "rx_residual_rms_static: device ordinal out of range"; return RX_ERR_ARG; }
  int prev = -1;
  if (cudaGetDevice(&prev) != cudaSuccess) { g_create_error = "rx_residual_rms_static: no CUDA device (this library has no CPU path)"; return RX_ERR_NO_DEVICE; }
  if (prev != device && cudaSetDevice(device) != cudaSuccess) { g_create_error = "rx_residual_rms_static: cudaSetDevice failed"; return RX_ERR_CUDA; }
  int rc = RX_OK;
  double* scratch = nullptr;
  const char* why = "";
  if ((rc = na_scratch(device, &scratch, &why)) != RX_OK) {
    snprintf(msg, sizeof msg, "rx_residual_rms_static: cudaGetSymbolAddress: %s", why);
    g_create_error = msg;
  } else {
    cudaStream_t s = (cudaStream_t)cudaStream;
    const int nBlocks = (int)((nPointDomain + NA_BS - 1) / NA_BS);
    if (nBlocks > 0) k_sumsq_partial_static<<<nBlocks, NA_BS, 0, s>>>(R_dev, nPointDomain, ld, nVar, scratch);
    k_rms_finish<<<nVar, 256, 0, s>>>(scratch, nBlocks, nVar, nPointGlobal, rms_dev, scratch + (size_t)NA_MAX_BLOCKS * NA_MAX_VARS);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
      snprintf(msg, sizeof msg, "rx_residual_rms_static: launch failed: %s", cudaGetErrorString(e));
      g_create_error = msg; rc = RX_ERR_CUDA;
    }
  }
  if (prev != device) cudaSetDevice(prev);
  return rc;
}
