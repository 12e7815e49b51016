// rx_order.cuh — processing order of the viscous edge kernel.
//
// The lanes of one warp of k_edge_visc_pair run the Stefan–Maxwell BiCGSTAB loop until the LAST of their sixteen edges has
// converged, and skip the matrix assembly (and, RANS, the pivoted QR) only if ALL sixteen may.  Iteration counts are 0, 5, 6 or
// 7 on the bench state and interleave in space, so a warp in natural edge order runs max(count) = 5.9 iterations for a mean of
// 5.25.  The kernel therefore records a small key per edge (iterations of its last solve, + 19 if the closure needed the QR)
// and the next iteration processes the edges of every TILE of ORD_TILE consecutive edges in the order of a STABLE counting sort
// by that key: warps become uniform (at most one mixed warp per key and tile), a tile's edges are still handled together by
// neighbouring blocks — sorting over the whole mesh was measured and gave the iterations back as cache misses, every line of
// edge and point data being fetched once per key group — and edges of one key stay in ascending order.  Fluxes are still
// stored by edge index and summed by k_accumulate in edge order, so the result does not depend on the processing order —
// same bits with and without.  One launch per iteration.
#pragma once
#include "rx_kernels.cuh"

namespace rx {

constexpr int ORD_NB = 40;        // bins: keys 0..18 (iterations), +19 when the SST closure ran its QR
constexpr int ORD_BS = 256;       // threads per block
#ifndef RX_ORD_ROUNDS
#define RX_ORD_ROUNDS 8
#endif
constexpr int ORD_ROUNDS = RX_ORD_ROUNDS;   // a block orders the ORD_BS * ORD_ROUNDS consecutive edges of its tile
constexpr int ORD_TILE = ORD_BS * ORD_ROUNDS;

// order[tile start + number of the tile's edges with a smaller key + rank among those of the same key] = edge
__global__ void __launch_bounds__(ORD_BS) k_order_tiles(const unsigned char* __restrict__ key, long long nEdge, int* __restrict__ order) {
  constexpr int NW = ORD_BS / 32;
  __shared__ int base[ORD_NB + 1];
  __shared__ int cnt[NW][ORD_NB + 1];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long e0 = (long long)blockIdx.x * ORD_TILE;
  if (threadIdx.x <= ORD_NB) base[threadIdx.x] = 0;
  __syncthreads();
  int kk[ORD_ROUNDS];
#pragma unroll
  for (int r = 0; r < ORD_ROUNDS; ++r) {
    const long long e = e0 + r * ORD_BS + threadIdx.x;
    kk[r] = e < nEdge ? min((int)key[e], ORD_NB - 1) : ORD_NB;   // bin ORD_NB: past the end, never written out
    atomicAdd(&base[kk[r]], 1);
  }
  __syncthreads();
  if (threadIdx.x == 0) {   // exclusive scan of the forty counts
    int run = (int)e0;
    for (int b = 0; b <= ORD_NB; ++b) { const int x = base[b]; base[b] = run; run += x; }
  }
#pragma unroll
  for (int r = 0; r < ORD_ROUNDS; ++r) {
    for (int i = threadIdx.x; i < NW * (ORD_NB + 1); i += ORD_BS) (&cnt[0][0])[i] = 0;
    __syncthreads();
    const int k = kk[r];
    const unsigned peers = __match_any_sync(0xffffffffu, k);
    const int rank = __popc(peers & ((1u << lane) - 1u));
    if (rank == 0) cnt[w][k] = __popc(peers);
    __syncthreads();
    if (k < ORD_NB) {
      int off = base[k] + rank;
      for (int q = 0; q < w; ++q) off += cnt[q][k];
      order[off] = (int)(e0 + r * ORD_BS + threadIdx.x);
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
