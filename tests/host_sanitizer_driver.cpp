// Driver of tests/test_host_sanitizers.py: the host-only parts of the library (csrc/host: dual-mesh builder, RCM, k-way partitioner, .su2 reader and
// geometry preprocessing) compiled with -fsanitize=address,undefined and run on jittered triangulations of several sizes, on the mesh files
// named on the command line and on malformed input.  Prints one line per case; any sanitizer report makes the process fail.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "rx_cuda.h"

static int builders() {
  int bad = 0;
  for (int trial = 0; trial < 6; ++trial) {
    const int nx = 7 + 13 * trial, ny = 5 + 9 * trial;
    const int64_t nP = (int64_t)nx * ny;
    std::vector<double> xy(2 * nP);
    srand(trial + 1);
    for (int j = 0; j < ny; ++j)
      for (int i = 0; i < nx; ++i) {
        xy[2 * (j * nx + i)] = i + 0.2 * (rand() / (double)RAND_MAX - 0.5);
        xy[2 * (j * nx + i) + 1] = j + 0.2 * (rand() / (double)RAND_MAX - 0.5);
      }
    std::vector<int64_t> tri;
    for (int j = 0; j + 1 < ny; ++j)
      for (int i = 0; i + 1 < nx; ++i) {
        const int64_t a = j * nx + i, b = a + 1, c = b + nx, d = a + nx;
        if ((i + j) & 1) { tri.insert(tri.end(), {a, b, c}); tri.insert(tri.end(), {a, c, d}); }
        else { tri.insert(tri.end(), {a, b, d}); tri.insert(tri.end(), {b, c, d}); }
      }
    rx_dual_mesh m;
    const int rc = rx_dual_mesh_tri(nP, xy.data(), (int64_t)tri.size() / 3, tri.data(), &m);
    std::vector<int64_t> n2o(nP), part(nP);
    int64_t cut = 0;
    const int rc2 = rc ? rc : rx_rcm_order(nP, nP, m.nbPtr, m.nbPoint, n2o.data());
    int rc3 = 0;
    for (int k : {1, 2, 3, 8}) if (!rc) rc3 |= rx_partition_kway(nP, m.nbPtr, m.nbPoint, xy.data(), k, 8, part.data(), &cut);
    // this rank's share of the mesh for every rank of the last (8-way) partition, with and without the RCM renumbering; no markers
    int rc4 = 0;
    if (!rc && !rc3)
      for (int r = 0; r < 8; ++r)
        for (int rcm = 0; rcm < 2; ++rcm) {
          bool owns = false;
          for (int64_t p = 0; p < nP; ++p) owns = owns || part[p] == r;
          rx_part q;
          const int e = rx_partition_mesh(nP, m.nEdge, m.edgeNodes, m.edgeNormal, xy.data(), m.volume, m.nbPtr, m.nbPoint, 0, nullptr, nullptr, nullptr, nullptr,
                                          part.data(), r, rcm, &q);
          if (owns ? e != 0 : e == 0) rc4 = 1;
          if (!e && (q.nPointDomain <= 0 || q.nPoint < q.nPointDomain || q.sendPtr[q.nNb] <= 0 || q.recvPtr[q.nNb] != q.nPoint - q.nPointDomain)) rc4 = 1;
          rx_part_free(&q);
        }
    printf("builders nP %lld nE %lld rc %d %d %d %d\n", (long long)nP, (long long)m.nEdge, rc, rc2, rc3, rc4);
    bad |= rc | rc2 | rc3 | rc4;
    rx_dual_mesh_free(&m);
  }
  rx_dual_mesh m;
  if (rx_dual_mesh_tri(0, nullptr, 0, nullptr, &m) == 0 || rx_dual_mesh_tri(3, nullptr, 1, nullptr, &m) == 0) bad = 1;   // bad arguments are refused
  return bad;
}

int main(int argc, char** argv) {
  int bad = builders();
  for (int k = 1; k < argc; ++k)
    for (int rcm = 0; rcm < 2; ++rcm) {
      rx_su2_mesh m;
      const int rc = rx_su2_mesh_read(argv[k], rcm, &m);
      printf("su2 %s rcm %d rc %d nP %lld nE %lld %s\n", argv[k], rcm, rc, (long long)m.nPoint, (long long)m.nEdge, rc ? rx_su2_mesh_error() : "");
      rx_su2_mesh_free(&m);
    }
  printf("DONE %d\n", bad);
  return bad;
}
