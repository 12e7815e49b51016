// rx_restart.cuh — SU2 native ASCII restart files straight from / into the device-resident state.
//
// Writer: byte-compatible with COutput::SetRestart for the reactive solvers (SU2_CFD/src/output_structure.cpp:3858-4060, data
// layout of COutput::MergeSolution :1991-2130): header line of quoted names, one line per owned point
//     PointID \t x \t y \t Conservative_1..nVar [\t k \t omega] \t Pressure \t Temperature \t Mach [\t Laminar_Viscosity [\t mu_t]] \t \n
// every value in C++ `scientific` with precision 15 (= printf "%.15e"), then the five trailer lines (AOA, SIDESLIP_ANGLE,
// INITIAL_BCTHRUST, DCD_DCL_VALUE, EXT_ITER).  The fields are gathered on the device into a pinned staging buffer in chunks
// (no full host mirror of the state) and formatted by a few host threads.
// Reader: CReactiveNSSolver::Load_Restart (solver_direct_reactive.cpp:4390-4520; Euler :566) and the turbulence reader
// (solver_direct_turbulent.cpp:1046-1064): PointID, two coordinates (skipped), nVar conserved values, then (k, omega).
// Both are included into rx_cuda.cu (they need rx_ctx).
#pragma once
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

namespace rx {

constexpr long long RST_CHUNK = 1 << 18;   // points per staging chunk

// one chunk of fields, SoA on the device -> [field][point] in the pinned buffer
__global__ void k_restart_gather(const Dev d, const double* __restrict__ eddyFlow, int rans, int viscous, long long p0, long long n, int nv, const long long* __restrict__ order,
                                 double* __restrict__ out) {
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= n) return;
  const long long p = order ? order[p0 + t] : p0 + t;   // order: owned points by ascending PointID
  int f = 0;
  out[(f++) * n + t] = d.x[p];
  out[(f++) * n + t] = d.y[p];
  for (int v = 0; v < nv; ++v) out[(f++) * n + t] = d.U[v * d.np + p];
  f += rans ? 2 : 0;   // k, omega are filled on the host (the caller's turbulence solution)
  out[(f++) * n + t] = d.V[V_P * d.np + p];
  out[(f++) * n + t] = d.V[V_T * d.np + p];
  {   // sqrt(GetVelocity2()) / GetSoundSpeed(), GetVelocity2 = inner_product from 0.0 (variable_reactive.hpp:633)
    double v2 = 0.0;
    v2 = v2 + d.V[V_VX * d.np + p] * d.V[V_VX * d.np + p];
    v2 = v2 + d.V[(V_VX + 1) * d.np + p] * d.V[(V_VX + 1) * d.np + p];
    out[(f++) * n + t] = sqrt(v2) / d.V[V_A * d.np + p];
  }
  if (viscous) out[(f++) * n + t] = d.mu[p];
  if (rans) out[(f++) * n + t] = eddyFlow[p];   // eddy viscosity as the flow variable holds it (set at the last Preprocessing), not the turbulence solver's newer one
}

// tecplot: no PointID column and the stream's default precision (C++ `scientific` = printf "%e"), COutput::SetTecplotASCII (output_tecplot.cpp:299-318)
inline void restart_format(const double* buf, long long n, int nf, const long long* ids, bool tecplot, std::string& out) {
  char tmp[64];
  out.clear();
  out.reserve((size_t)n * (nf * 23 + 12));
  for (long long t = 0; t < n; ++t) {
    int m;
    if (!tecplot) { m = snprintf(tmp, sizeof tmp, "%lld\t", ids[t]); out.append(tmp, m); }
    for (int f = 0; f < nf; ++f) {
      m = snprintf(tmp, sizeof tmp, tecplot ? "%e\t" : "%.15e\t", buf[(size_t)f * n + t]);
      out.append(tmp, m);
    }
    out.push_back('\n');
  }
}

}  // namespace rx

// the per-point lines of a restart (tecplot = false) or Tecplot volume file (true) of the owned points, by ascending PointID
static int restart_points(rx_ctx* c, FILE* fo, const int64_t* globalIndex, const double* kOmega, bool tecplot) {
  using namespace rx;
  const Dev& d = c->d;
  const bool rans = c->o.rans != 0, visc = c->o.viscous != 0;
  const long long nOwn = d.nPointDomain;
  const int nCons = c->nVar + (rans ? 2 : 0);
  const int nf = 2 + nCons + 3 + (visc ? 1 : 0) + (rans ? 1 : 0);
  // lines go out by ascending PointID (the reference's master writes global index 0, 1, 2, ...; SU2 renumbers the points of a
  // partition — SetRCM_Ordering — so the local order is not the global one)
  std::vector<long long> order;
  long long* dOrder = nullptr;
  if (globalIndex) {
    bool sorted = true;
    for (long long p = 1; p < nOwn && sorted; ++p) sorted = globalIndex[p] > globalIndex[p - 1];
    if (!sorted) {
      order.resize((size_t)nOwn);
      for (long long p = 0; p < nOwn; ++p) order[p] = p;
      std::sort(order.begin(), order.end(), [&](long long a, long long b) { return globalIndex[a] < globalIndex[b]; });
      if (cudaMalloc((void**)&dOrder, (size_t)nOwn * sizeof(long long)) != cudaSuccess ||
          cudaMemcpy(dOrder, order.data(), (size_t)nOwn * sizeof(long long), cudaMemcpyHostToDevice) != cudaSuccess) {
        if (dOrder) cudaFree(dOrder);
        c->err = "rx_restart_write: staging allocation failed"; return RX_ERR_CUDA;
      }
    }
  }
  const long long chunk = std::min<long long>(RST_CHUNK, std::max<long long>(nOwn, 1));
  double *pin = nullptr, *dev = nullptr;
  if (cudaMallocHost((void**)&pin, (size_t)chunk * nf * sizeof(double)) != cudaSuccess || cudaMalloc((void**)&dev, (size_t)chunk * nf * sizeof(double)) != cudaSuccess) {
    if (pin) cudaFreeHost(pin);
    if (dOrder) cudaFree(dOrder);
    c->err = "rx_restart_write: staging allocation failed"; return RX_ERR_CUDA;
  }
  const int nThreads = (int)std::max(1u, std::min(8u, std::thread::hardware_concurrency()));
  std::vector<std::string> parts(nThreads);
  std::vector<long long> ids((size_t)chunk);
  int rc = RX_OK;
  for (long long p0 = 0; p0 < nOwn && rc == RX_OK; p0 += chunk) {
    const long long n = std::min(chunk, nOwn - p0);
    k_restart_gather<<<nblk(n, 256), 256, 0, c->stream>>>(d, c->eddyFlow, rans ? 1 : 0, visc ? 1 : 0, p0, n, c->nVar, dOrder, dev);
    c->launches++;
    if (cudaMemcpyAsync(pin, dev, (size_t)n * nf * sizeof(double), cudaMemcpyDeviceToHost, c->stream) != cudaSuccess ||
        cudaStreamSynchronize(c->stream) != cudaSuccess) { c->err = "rx_restart_write: device copy failed"; rc = RX_ERR_CUDA; break; }
    if (rans)
      for (long long t = 0; t < n; ++t) {
        const long long p = order.empty() ? p0 + t : order[p0 + t];
        pin[(size_t)(2 + c->nVar) * n + t] = kOmega[2 * p]; pin[(size_t)(3 + c->nVar) * n + t] = kOmega[2 * p + 1];
      }
    for (long long t = 0; t < n; ++t) { const long long p = order.empty() ? p0 + t : order[p0 + t]; ids[t] = globalIndex ? globalIndex[p] : p; }
    // the gather wrote fields with stride n; format slices of points in parallel, write them in order
    std::vector<std::thread> th;
    const long long per = (n + nThreads - 1) / nThreads;
    for (int k = 0; k < nThreads; ++k) {
      const long long a = std::min(n, k * per), b = std::min(n, a + per);
      th.emplace_back([&, k, a, b]() {
        // a sub-range of a [field][n] buffer is not contiguous per field: copy the slice into a compact [field][m] block first
        const long long m = b - a;
        std::vector<double> loc((size_t)m * nf);
        for (int f = 0; f < nf; ++f) memcpy(&loc[(size_t)f * m], pin + (size_t)f * n + a, (size_t)m * sizeof(double));
        restart_format(loc.data(), m, nf, ids.data() + a, tecplot, parts[k]);
      });
    }
    for (auto& t : th) t.join();
    for (int k = 0; k < nThreads; ++k)
      if (!parts[k].empty() && fwrite(parts[k].data(), 1, parts[k].size(), fo) != parts[k].size()) { c->err = "rx_restart_write: short write"; rc = RX_ERR_ARG; }
  }
  cudaFreeHost(pin); cudaFree(dev);
  if (dOrder) cudaFree(dOrder);
  return rc;
}

extern "C" int rx_restart_write(rx_ctx* c, const char* path, const int64_t* globalIndex, const double* kOmega, const rx_restart_info* info, int piece) {
  using namespace rx;
  if (!c || !path) { if (c) c->err = "rx_restart_write: null argument"; return RX_ERR_ARG; }
  { const int rc = ensure_output(c); if (rc) return rc; }   // laminar viscosity of the Output pass
  const bool rans = c->o.rans != 0, visc = c->o.viscous != 0;
  if (rans && !kOmega) { c->err = "rx_restart_write: a RANS restart carries the turbulence solution (k, omega): kOmega must not be NULL"; return RX_ERR_ARG; }
  const int nCons = c->nVar + (rans ? 2 : 0);
  FILE* fo = fopen(path, "w");
  if (!fo) { c->err = std::string("rx_restart_write: cannot open ") + path; return RX_ERR_ARG; }
  if (!piece) {
    std::string h = "\"PointID\"\t\"x\"\t\"y\"";
    for (int v = 0; v < nCons; ++v) h += "\t\"Conservative_" + std::to_string(v + 1) + "\"";
    h += "\t\"Pressure\"\t\"Temperature\"\t\"Mach\"";
    if (visc) h += "\t\"Laminar_Viscosity\"";
    if (rans) h += "\t\"<greek>m</greek><sub>t</sub>\"";
    h += "\n";
    fwrite(h.data(), 1, h.size(), fo);
  }
  int rc = restart_points(c, fo, globalIndex, kOmega, false);
  if (rc == RX_OK && !piece) {
    rx_restart_info z{};
    z.bcThrust = 4000.0;   // CConfig default of INITIAL_BCTHRUST
    const rx_restart_info& q = info ? *info : z;
    fprintf(fo, "AOA= %.15e\nSIDESLIP_ANGLE= %.15e\nINITIAL_BCTHRUST= %.15e\nDCD_DCL_VALUE= %.15e\nEXT_ITER= %lld\n", q.aoa, q.sideslip, q.bcThrust, q.dCD_dCL,
            (long long)q.extIter);
  }
  if (fclose(fo) != 0 && rc == RX_OK) { c->err = "rx_restart_write: close failed"; rc = RX_ERR_ARG; }
  return rc;
}

// Tecplot ASCII volume solution file (COutput::SetTecplotASCII, SU2_CFD/src/output_tecplot.cpp:46-420, for the reactive solvers): title,
// the VARIABLES line exactly as the reference prints it (for REACTIVE_RANS it names more variables than it writes, :163-178 — kept,
// the file is what SU2 users' scripts read), the ZONE line (2-D volume zones are FEQUADRILATERAL, triangles repeat their third node),
// one line per point — x, y and the restart file's columns, default `scientific` precision — then the 1-based connectivity, all triangles
// before all quadrilaterals as MergeConnectivity collects them (output_structure.cpp:881).
extern "C" int rx_tecplot_write(rx_ctx* c, const char* path, const int64_t* globalIndex, const double* kOmega, int64_t nTria, const int64_t* tria,
                                int64_t nQuad, const int64_t* quad, const char* zonePrefix) {
  using namespace rx;
  if (!c || !path || (nTria > 0 && !tria) || (nQuad > 0 && !quad) || nTria < 0 || nQuad < 0) { if (c) c->err = "rx_tecplot_write: bad argument"; return RX_ERR_ARG; }
  if (c->d.nPoint != c->d.nPointDomain) { c->err = "rx_tecplot_write: one partition holds part of the zone only; merge restart pieces and post-process (SU2_SOL) instead"; return RX_ERR_UNSUPPORTED; }
  { const int rc = ensure_output(c); if (rc) return rc; }
  const bool rans = c->o.rans != 0, visc = c->o.viscous != 0;
  if (rans && !kOmega) { c->err = "rx_tecplot_write: a RANS file carries the turbulence solution (k, omega): kOmega must not be NULL"; return RX_ERR_ARG; }
  const int nCons = c->nVar + (rans ? 2 : 0);
  FILE* fo = fopen(path, "w");
  if (!fo) { c->err = std::string("rx_tecplot_write: cannot open ") + path; return RX_ERR_ARG; }
  std::string h = "TITLE = \"Visualization of the volumetric solution\"\nVARIABLES = \"x\",\"y\"";
  for (int v = 0; v < nCons; ++v) h += ",\"Conservative_" + std::to_string(v + 1) + "\"";
  h += ",\"Pressure\",\"Temperature\",\"Mach\"";
  if (visc) h += ", \"<greek>m</greek>\"";
  if (rans) h += ",\"Pressure\",\"Temperature\",\"C<sub>p</sub>\",\"Mach\", \"<greek>m</greek>\", \"C<sub>f</sub>_x\", \"C<sub>f</sub>_y\", \"h\", \"y<sup>+</sup>\", \"<greek>m</greek><sub>t</sub>\"";
  h += std::string("\nZONE ") + (zonePrefix ? zonePrefix : "") + "NODES= " + std::to_string((long long)c->d.nPointDomain) + ", ELEMENTS= " + std::to_string((long long)(nTria + nQuad)) + ", DATAPACKING=POINT, ZONETYPE=FEQUADRILATERAL\n";
  fwrite(h.data(), 1, h.size(), fo);
  int rc = restart_points(c, fo, globalIndex, kOmega, true);
  if (rc == RX_OK) {
    std::string body;
    body.reserve((size_t)(nTria + nQuad) * 32);
    char tmp[96];
    for (int64_t e = 0; e < nTria; ++e) {
      const int m = snprintf(tmp, sizeof tmp, "%lld\t%lld\t%lld\t%lld\n", (long long)tria[3 * e] + 1, (long long)tria[3 * e + 1] + 1, (long long)tria[3 * e + 2] + 1, (long long)tria[3 * e + 2] + 1);
      body.append(tmp, m);
    }
    for (int64_t e = 0; e < nQuad; ++e) {
      const int m = snprintf(tmp, sizeof tmp, "%lld\t%lld\t%lld\t%lld\n", (long long)quad[4 * e] + 1, (long long)quad[4 * e + 1] + 1, (long long)quad[4 * e + 2] + 1, (long long)quad[4 * e + 3] + 1);
      body.append(tmp, m);
    }
    if (fwrite(body.data(), 1, body.size(), fo) != body.size()) { c->err = "rx_tecplot_write: short write"; rc = RX_ERR_ARG; }
  }
  if (fclose(fo) != 0 && rc == RX_OK) { c->err = "rx_tecplot_write: close failed"; rc = RX_ERR_ARG; }
  return rc;
}

extern "C" int rx_restart_read(rx_ctx* c, const char* path, const int64_t* globalIndex, int64_t nGlobal, double* kOmega, rx_restart_info* info) {
  using namespace rx;
  if (!c || !path) { if (c) c->err = "rx_restart_read: null argument"; return RX_ERR_ARG; }
  const Dev& d = c->d;
  const bool rans = c->o.rans != 0;
  const long long nP = d.nPoint;
  if (nGlobal <= 0) nGlobal = d.nPointDomain;
  FILE* fi = fopen(path, "r");
  if (!fi) { c->err = std::string("There is no flow restart file!! ") + path; return RX_ERR_ARG; }
  std::string txt;
  {
    fseek(fi, 0, SEEK_END);
    const long sz = ftell(fi);
    fseek(fi, 0, SEEK_SET);
    txt.resize((size_t)std::max(sz, 0L));
    if (sz > 0 && fread(&txt[0], 1, (size_t)sz, fi) != (size_t)sz) { fclose(fi); c->err = "rx_restart_read: short read"; return RX_ERR_ARG; }
    fclose(fi);
  }
  // global index -> local points (Global2Local of the reference; halo points of a partition are filled too)
  std::vector<long long> g2l, nextSame;
  if (globalIndex) {
    g2l.assign((size_t)nGlobal, -1);
    nextSame.assign((size_t)nP, -1);   // further local copies of one global point, chained in ascending local index
    for (long long p = nP - 1; p >= 0; --p) {
      const long long g = globalIndex[p];
      if (g < 0 || g >= nGlobal) { c->err = "rx_restart_read: global index out of range"; return RX_ERR_ARG; }
      nextSame[p] = g2l[g]; g2l[g] = p;
    }
  }
  const int nv = c->nVar;
  std::vector<double> U((size_t)nP * nv, 0.0);
  std::vector<char> seen((size_t)nP, 0);
  const char* s = txt.c_str();
  const char* end = s + txt.size();
  const char* nl = (const char*)memchr(s, '\n', end - s);   // header line
  if (!nl) { c->err = "rx_restart_read: empty restart file"; return RX_ERR_ARG; }
  s = nl + 1;
  long long found = 0;
  for (long long line = 0; line < nGlobal; ++line) {
    if (s >= end) break;
    nl = (const char*)memchr(s, '\n', end - s);
    const char* le = nl ? nl : end;
    char* q = nullptr;
    const long long id = strtoll(s, &q, 10);
    if (q == s) break;   // not a data line (trailer reached early)
    const long long p = globalIndex ? (id >= 0 && id < nGlobal ? g2l[id] : -1) : (id < d.nPointDomain ? id : -1);
    if (p >= 0) {
      const char* r = q;
      strtod(r, &q); r = q; strtod(r, &q); r = q;   // coordinates
      for (int v = 0; v < nv; ++v) { U[(size_t)p * nv + v] = strtod(r, &q); if (q == r) { c->err = "rx_restart_read: truncated line"; return RX_ERR_ARG; } r = q; }
      if (rans && kOmega) { kOmega[2 * p] = strtod(r, &q); r = q; kOmega[2 * p + 1] = strtod(r, &q); r = q; }
      if (!seen[p]) { seen[p] = 1; if (p < d.nPointDomain) ++found; }
      if (globalIndex)
        for (long long pp = nextSame[p]; pp >= 0; pp = nextSame[pp]) {
          memcpy(&U[(size_t)pp * nv], &U[(size_t)p * nv], nv * sizeof(double));
          if (rans && kOmega) { kOmega[2 * pp] = kOmega[2 * p]; kOmega[2 * pp + 1] = kOmega[2 * p + 1]; }
          seen[pp] = 1;
        }
    }
    s = le + 1;
  }
  if (found < d.nPointDomain) { c->err = std::string("The solution file ") + path + " doesn't match with the mesh file!"; return RX_ERR_ARG; }
  if (info) {
    rx_restart_info z{};
    auto grab = [&](const char* key, double& v) { const char* k = strstr(s < end ? s : txt.c_str(), key); if (k) v = strtod(k + strlen(key), nullptr); };
    grab("AOA=", z.aoa); grab("SIDESLIP_ANGLE=", z.sideslip); grab("INITIAL_BCTHRUST=", z.bcThrust); grab("DCD_DCL_VALUE=", z.dCD_dCL);
    double it = 0.0; grab("EXT_ITER=", it); z.extIter = (int64_t)it;
    *info = z;
  }
  return upload_aos(c, U.data(), c->d.U, nv);
}
