// tests/cuda_host_shim.h — just enough of the CUDA execution model to compile the one-thread-per-item kernels of csrc/rx_kernels.cuh and
// csrc/rx_kernels2.cuh FOR THE HOST (tests/host_kernels.cpp): the qualifiers vanish, the thread / block indices are thread-local variables a
// launch loop sets, the fast division paths of rx_fdiv.cuh become the IEEE operations they are built to equal.  TEST INFRASTRUCTURE.
// Kernels that cooperate inside a block (shuffles, block reductions through shared memory with barriers) are not run through this shim.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __launch_bounds__(...)
#define __maxnreg__(...)
#define __grid_constant__
#define __restrict__
#define __shared__
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
