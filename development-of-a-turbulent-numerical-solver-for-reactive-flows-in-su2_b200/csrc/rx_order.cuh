// rx_order.cuh — processing order of the viscous edge kernel.
//
// The lanes of one warp of k_edge_visc_pair run the Stefan–Maxwell BiCGSTAB loop until the LAST of their sixteen edges has
// converged, and skip the matrix assembly (and, RANS, the pivoted QR) only if ALL sixteen may.  Iteration counts are 0, 5, 6 or
// 7 on the bench state and interleave in space, so a warp in natural edge order runs max(count) = 5.9 iterations for a mean of
// 5.25.  The kernel therefore records a small key per edge (iterations of its last solve, + 19 if the closure needed the QR)
// and the next iteration processes the edges in the order of a STABLE counting sort by that key: warps become uniform, edges
// of one key stay in ascending order (their point gathers stay local).  Fluxes are still stored by edge index and summed by
// k_accumulate in edge order, so the result does not depend on the processing order — same bits with and without.
//
// Three launches per iteration: tile histograms, one exclusive scan of the bin-major table, stable scatter.
#pragma once
#include "rx_kernels.cuh"

namespace rx {

constexpr int ORD_NB = 40;        // bins: keys 0..18 (iterations), +19 when the SST closure ran its QR
constexpr int ORD_BS = 256;       // threads per block
constexpr int ORD_ROUNDS = 8;     // a block orders ORD_BS * ORD_ROUNDS consecutive edges
constexpr int ORD_TILE = ORD_BS * ORD_ROUNDS;

// hist[bin * nTiles + tile] = number of edges of the tile with that key
__global__ void __launch_bounds__(ORD_BS) k_order_hist(const unsigned char* __restrict__ key, long long nEdge, int nTiles, int* __restrict__ hist) {
  __shared__ int h[ORD_NB];
  if (threadIdx.x < ORD_NB) h[threadIdx.x] = 0;
  __syncthreads();
  const long long e0 = (long long)blockIdx.x * ORD_TILE;
#pragma unroll
  for (int r = 0; r < ORD_ROUNDS; ++r) {
    const long long e = e0 + r * ORD_BS + threadIdx.x;
    if (e < nEdge) atomicAdd(&h[min((int)key[e], ORD_NB - 1)], 1);
  }
  __syncthreads();
  if (threadIdx.x < ORD_NB) hist[(size_t)threadIdx.x * nTiles + blockIdx.x] = h[threadIdx.x];
}

// in-place exclusive scan of v[0..n) by ONE block of 1024 threads (thread i owns a contiguous chunk)
__global__ void __launch_bounds__(1024) k_order_scan(int* __restrict__ v, long long n) {
  __shared__ int part[1024];
  const long long chunk = (n + 1023) / 1024;
  const long long a = min(n, threadIdx.x * chunk), b = min(n, a + chunk);
  int s = 0;
  for (long long i = a; i < b; ++i) s += v[i];
  part[threadIdx.x] = s;
  __syncthreads();
  for (int off = 1; off < 1024; off <<= 1) {   // Hillis–Steele inclusive scan of the 1024 partial sums
    const int add = threadIdx.x >= off ? part[threadIdx.x - off] : 0;
    __syncthreads();
    part[threadIdx.x] += add;
    __syncthreads();
  }
  int run = part[threadIdx.x] - s;
  for (long long i = a; i < b; ++i) { const int x = v[i]; v[i] = run; run += x; }
}

// order[start(bin, tile) + rank of the edge among the tile's edges of that bin] = edge; ranks follow the edge index
__global__ void __launch_bounds__(ORD_BS) k_order_scatter(const unsigned char* __restrict__ key, long long nEdge, int nTiles,
                                                         const int* __restrict__ start, int* __restrict__ order) {
  constexpr int NW = ORD_BS / 32;
  __shared__ int base[ORD_NB + 1];
  __shared__ int cnt[NW][ORD_NB + 1];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x < ORD_NB) base[threadIdx.x] = start[(size_t)threadIdx.x * nTiles + blockIdx.x];
  const long long e0 = (long long)blockIdx.x * ORD_TILE;
  for (int r = 0; r < ORD_ROUNDS; ++r) {
    for (int i = threadIdx.x; i < NW * (ORD_NB + 1); i += ORD_BS) (&cnt[0][0])[i] = 0;
    __syncthreads();
    const long long e = e0 + r * ORD_BS + threadIdx.x;
    const int k = e < nEdge ? min((int)key[e], ORD_NB - 1) : ORD_NB;   // bin ORD_NB: past the end, never written out
    const unsigned peers = __match_any_sync(0xffffffffu, k);
    const int rank = __popc(peers & ((1u << lane) - 1u));
    if (rank == 0) cnt[w][k] = __popc(peers);
    __syncthreads();
    if (k < ORD_NB) {
      int off = base[k] + rank;
      for (int q = 0; q < w; ++q) off += cnt[q][k];
      order[off] = (int)e;
    }
    __syncthreads();
    if (threadIdx.x < ORD_NB) {
      int tot = 0;
#pragma unroll
      for (int q = 0; q < NW; ++q) tot += cnt[q][threadIdx.x];
      base[threadIdx.x] += tot;
    }
    __syncthreads();
  }
}

}  // namespace rx
