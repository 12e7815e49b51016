// tests/cuda_host_shim.h — just enough of the CUDA execution model to compile the one-thread-per-item kernels of csrc/rx_kernels.cuh and
// csrc/rx_kernels2.cuh FOR THE HOST (tests/host_kernels.cpp): the qualifiers vanish, the thread / block indices are thread-local variables a
// launch loop sets, the fast division paths of rx_fdiv.cuh become the IEEE operations they are built to equal.  TEST INFRASTRUCTURE.
// Kernels that cooperate inside a block (block reductions through shared memory with barriers) are not run through this shim.  The one cooperation
// that IS emulated is the pair of lanes of the two-lanes-per-edge viscous kernel: the two lanes of a pair run as two host threads in lockstep, the
// shuffles and votes of the warp become exchanges between the two (a vote over the pair instead of the warp: a lane that has converged only waits for
// the others, so the result of an edge does not depend on which edges share its warp — the property tests/test_gpu_order.py pins on the device).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <atomic>
#include <thread>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __launch_bounds__(...)
#define __maxnreg__(...)
#define __grid_constant__
#define __restrict__
#define __shared__
#define __constant__
#define CUDART_INF (std::numeric_limits<double>::infinity())
#define CUDART_NAN (std::numeric_limits<double>::quiet_NaN())

struct emu_dim3 { unsigned x = 0, y = 0, z = 0; };
static thread_local emu_dim3 threadIdx, blockIdx, blockDim, gridDim;
inline void __syncthreads() {}
template <class T> inline T __shfl_down_sync(unsigned, T v, int) { return v; }   // only so that the block reductions compile; they are not run here
inline unsigned long long atomicAdd(unsigned long long* a, unsigned long long v) { const unsigned long long o = *a; *a = o + v; return o; }
inline int __double2hiint(double x) { int64_t b; memcpy(&b, &x, 8); return (int)(b >> 32); }
using std::fmin; using std::fmax; using std::fabs; using std::sqrt;
template <class T> inline T min(T a, T b) { return a < b ? a : b; }
template <class T> inline T max(T a, T b) { return a > b ? a : b; }

namespace emu {   // two host threads in lockstep = the two lanes of one pair
struct PairSync { std::atomic<int> arrived{0}; std::atomic<int> gen{0}; unsigned long long slot[32]; };
inline PairSync& ps() { static PairSync s; return s; }
inline void pair_barrier() {
  PairSync& s = ps();
  const int g = s.gen.load(std::memory_order_acquire);
  if (s.arrived.fetch_add(1, std::memory_order_acq_rel) == 1) { s.arrived.store(0, std::memory_order_relaxed); s.gen.store(g + 1, std::memory_order_release); }
  else { int spins = 0; while (s.gen.load(std::memory_order_acquire) == g) if (++spins > 2000) { std::this_thread::yield(); spins = 0; } }   // yields on a box with fewer cores than lanes
}
template <class T> inline T exchange(T v, int srcLane) {
  static_assert(sizeof(T) <= 8, "one word per lane");
  PairSync& s = ps();
  memcpy(&s.slot[threadIdx.x & 31], &v, sizeof(T));
  pair_barrier();
  T r; memcpy(&r, &s.slot[srcLane & 31], sizeof(T));
  pair_barrier();
  return r;
}
}  // namespace emu
template <class T> inline T __shfl_xor_sync(unsigned, T v, int m) { return emu::exchange(v, (int)(threadIdx.x & 31) ^ m); }
template <class T> inline T __shfl_sync(unsigned, T v, int src) { return emu::exchange(v, src); }
inline int __any_sync(unsigned, int p) { const int q = emu::exchange(p ? 1 : 0, (int)(threadIdx.x & 31) ^ 1); return (p ? 1 : 0) | q; }
inline void __syncwarp() { emu::pair_barrier(); }

namespace rx {
// rx_fdiv.cuh on the host: the quotients and roots the fast paths return where their range tests pass are the correctly rounded ones
struct Recip { double b, r; };
inline Recip frecip(double b) { return Recip{b, 0.0}; }
inline double fdiv(double a, const Recip& R, bool&) { return a / R.b; }
inline double fdiv(double a, double b, bool&) { return a / b; }
inline double fdivu(double a, const Recip& R) { return a / R.b; }
template <int LO, int HI> inline bool exp_in(double x) { const unsigned e = ((unsigned)__double2hiint(x) >> 20) & 0x7ffu; return e - (unsigned)(1023 + LO) <= (unsigned)(HI - LO); }
inline bool mid_range(double x) { const unsigned e = ((unsigned)__double2hiint(x) >> 20) & 0x7ffu; return e - (1023u - 200u) <= 400u; }
inline bool num_range(double x) { const unsigned e = ((unsigned)__double2hiint(x) >> 20) & 0x7ffu; return (e - (1023u - 600u) <= 800u) || x == 0.0; }
inline double fsqrt(double x, bool&) { return std::sqrt(x); }
}  // namespace rx
