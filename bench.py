#!/usr/bin/env python
"""bench.py — edge-flux + source updates/s of the reactive residual hot path on B200 (BASELINE.json metric).

Own arm (default):   python bench.py --gpus N --steps K --warmup W      (N>1 under torchrun, one rank per GPU)
  workload = BASELINE.json configs[1]: laminar multispecies reactive NS, shipped 9-species mechanism, synthetic
  triangulated 2-D channel of 1415x1415 points per GPU (4.0M cells, 2.0M points, 6.0M edges), explicit 3-stage RK,
  AUSM 1st order + Stefan-Maxwell viscous flux + Arrhenius source, WLS gradients.  One step = one solver iteration
  (MultiGrid_Cycle: 3 RK stages of preprocess/gradients/time step/residual/BCs/update + the Output preprocessing pass).
  Units of work per step = nRK * (interior edges + owned points)   [edge-flux updates + source-point updates].
Reference arm:       python bench.py --impl reference --steps K --warmup W
  the reference's own CPU classes (oracle/_ref/ref_harness, compiled from the reference sources) on the same workload
  shape at a bounded size, one serial process per host core (no MPI runtime on the box), aggregate throughput.
"""
import argparse
import json
import os
import re
import shutil
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
PKG = "development-of-a-turbulent-numerical-solver-for-reactive-flows-in-su2_b200"
RK_ALPHA = [0.66667, 0.66667, 1.0]
CONV_KERNEL_NAME = "k_edge_flux<9> (AUSM+-up edge flux)"
METRIC = "edge-flux+source updates/s"
REF_CASE = os.path.join(ROOT, "oracle", "_ref", "cases", "bench_ref")
HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_harness")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region (B200_PROFILING.md)."""
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu):
        self.gpu = gpu; self.p = None; self.lines = []

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=lambda: [self.lines.append(l) for l in self.p.stdout], daemon=True)
            self.t.start()
        except Exception:
            self.p = None

    def stop(self):
        if not self.p:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        self.t.join(timeout=2)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for l in self.lines:
            f = [x.strip() for x in l.split(",")]
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
                for n, v in zip(names, f[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


_ALL_CPUS = None   # affinity of the process at start (bind_to_gpu_numa narrows it for the GPU arm; the CPU reference gets all of it back)


def bind_to_gpu_numa(local):
    """Multi-GPU runs: keep this rank (and the pinned host buffers it allocates) on the CPUs NVML reports as local to its GPU, so that
    eight ranks do not pull 8 x 416 MB per step through one socket's memory.  Returns a note for the JSON line, or None."""
    global _ALL_CPUS
    try:
        import pynvml
        import torch
        _ALL_CPUS = os.sched_getaffinity(0)
        pynvml.nvmlInit()
        pr = torch.cuda.get_device_properties(local)
        bus = "%08x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        h = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
        ncpu = os.cpu_count() or 1
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        cpus = {i for i in range(ncpu) if (mask[i // 64] >> (i % 64)) & 1} & _ALL_CPUS
        if cpus and cpus != _ALL_CPUS:
            os.sched_setaffinity(0, cpus)
            return "rank bound to the %d CPUs NVML reports as local to its GPU" % len(cpus)
    except Exception:
        pass
    return None


def run_reference(steps, warmup, procs=None):
    """Times oracle/_ref/ref_harness (the reference's CReactiveNSSolver & co.) on the bounded bench case."""
    if not (os.path.exists(HARNESS) and os.path.isdir(REF_CASE)):
        return None
    if _ALL_CPUS:
        os.sched_setaffinity(0, _ALL_CPUS)   # the CPU arm uses the whole host, whatever the GPU arm bound itself to
    procs = procs or min(physical_cores(), REF_PROCS)
    nit = warmup + steps
    tmp = tempfile.mkdtemp(prefix="rxref_")
    ps = []
    for k in range(procs):
        d = os.path.join(tmp, "p%d" % k)
        shutil.copytree(REF_CASE, d)
        ps.append(subprocess.Popen([HARNESS, "case.cfg", str(nit), "dump", str(10 ** 9), "11"], cwd=d, stdout=subprocess.PIPE,
                                   stderr=subprocess.STDOUT, text=True))
    times, counts = [], None
    for p in ps:
        out = p.communicate()[0]
        t = [float(m.group(1)) for m in re.finditer(r"HARNESS_TIME it \d+ flow_seconds ([0-9.eE+-]+)", out)]
        if p.returncode != 0 or len(t) < nit:
            shutil.rmtree(tmp, ignore_errors=True)
            raise RuntimeError("reference harness failed:\n" + out[-2000:])
        times.append(t[warmup:])
        rss = [int(m.group(1)) for m in re.finditer(r"HARNESS_RSS_KB (\d+)", out)]
    meta = json.load(open(os.path.join(tmp, "p0", "dump", "meta.json")))
    shutil.rmtree(tmp, ignore_errors=True)
    nE, nP = meta["nEdge"], meta["nPointDomain"]
    units = len(RK_ALPHA) * (nE + nP)
    per_proc = [units * steps / sum(t) for t in times]
    ms = 1e3 * max(sum(t) for t in times) / steps
    return dict(value=float(sum(per_proc)), cores=procs, ms_per_step=ms, nEdge=nE, nPoint=nP, rss_mb=(rss[0] / 1024.0 if rss else None),
                sample="%d independent serial reference processes, one per physical core up to %d (no MPI runtime on the box), each the 3-stage-RK laminar "
                       "reactive NS iteration of the GPU arm's cfg on a %d-point / %d-edge triangulated channel (peak resident set %s MB per process: far "
                       "beyond the caches), %d timed iterations; aggregate updates/s" % (procs, REF_PROCS, nP, nE, ("%.0f" % (rss[0] / 1024.0)) if rss else "?", steps))


def bench_source(a, pk, chem, rank, world, local, dev, K, W):
    """configs[3]: the Arrhenius + PaSR source term alone over independent points (rx_source_points_dev; every rank its own
    points, no exchange).  value: device-timed passes over HBM-resident SoA inputs; e2e: the host-buffer entry point
    rx_source_points (allocation, H2D, kernel, D2H) on a bounded number of points."""
    import torch
    import torch.distributed as dist
    n = int(a.points)
    base = min(n, 4_000_000)                     # seeded states, tiled on the device up to n
    T, rho, Y, omT, vol = pk.synthetic_points(chem, base, seed=12345 + rank)
    reps = (n + base - 1) // base
    t = lambda x: torch.as_tensor(np.ascontiguousarray(x), device=dev)
    tile = lambda x: t(x).repeat(reps)[:n].contiguous()
    dT, dr, dO, dV = tile(T), tile(rho), tile(omT), tile(vol)
    dY = t(Y.T).repeat(1, reps)[:, :n].contiguous()
    dR = torch.empty((chem["Ns"], n), dtype=torch.float64, device=dev)
    opts = pk.default_opts(rans=True)
    solver = pk.CReactiveNSSolverGPU(pk.channel_mesh(5, 4), chem, opts, device=local)

    def step():
        solver.SourcePointsDev(n, dT.data_ptr(), dr.data_ptr(), dY.data_ptr(), dO.data_ptr(), dV.data_ptr(), dR.data_ptr())

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(W, 3)):
        step()
    l0 = solver.LaunchCount()
    sampler = ClockSampler(local); sampler.start()
    barrier()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K):
        step()
    e1.record()
    barrier()
    ms_dev = e0.elapsed_time(e1) / K
    clocks = sampler.stop()
    launches = solver.LaunchCount() - l0
    if not bool(torch.isfinite(dR).all()):
        raise SystemExit("bench.py: non-finite source term")
    ne = min(n, 8_000_000)
    hT, hr, hY, hO, hV = T, rho, Y, omT, vol
    if ne > base:
        r2 = (ne + base - 1) // base
        hT, hr, hO, hV = [np.tile(x, r2)[:ne] for x in (T, rho, omT, vol)]; hY = np.tile(Y, (r2, 1))[:ne]
    else:
        hT, hr, hO, hV, hY = T[:ne], rho[:ne], omT[:ne], vol[:ne], Y[:ne]
    pk.source_points(chem, hT, hr, hY, hV, omegaTurb=hO, rans=True, device=local)
    barrier()
    t0 = time.perf_counter()
    Ke = 3
    for _ in range(Ke):
        pk.source_points(chem, hT, hr, hY, hV, omegaTurb=hO, rans=True, device=local)
    barrier()
    ms_e2e = 1e3 * (time.perf_counter() - t0) / Ke
    fp64_peak = solver.Fp64Peak()
    tt = torch.tensor([ms_dev, ms_e2e], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    ms_dev, ms_e2e = float(tt[0]), float(tt[1])
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    flops_per_point = 3.5e3      # FP64-equivalent operations of Arrhenius + PaSR for 2 reactions, 9 species
    bytes_per_point = 8.0 * (4 + 2 * chem["Ns"])
    ach = flops_per_point * n / (ms_dev * 1e-3) / 1e12
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    line = {"metric": "source-point updates/s", "value": world * n / (ms_dev * 1e-3), "unit": "updates/s", "n_gpus": world, "steps": K, "warmup": max(W, 3),
            "ms_per_step": ms_dev, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "isolated Arrhenius + PaSR source term (9 species, 2 reactions) over independent points, SoA inputs resident in HBM",
                       "points_per_gpu": n, "cache": "inputs %.1f GB per GPU, larger than L2" % (n * bytes_per_point / 1e9)},
            "e2e": {"value": world * ne / (ms_e2e * 1e-3), "unit": "updates/s", "ms_per_step": ms_e2e, "points_per_call": ne,
                    "h2d_bytes_per_step": int(ne * 8 * (4 + chem["Ns"])), "d2h_bytes_per_step": int(ne * 8 * chem["Ns"])},
            "gpu_launches": int(launches), "clocks": clocks,
            "roofline": {"kernel": "k_source_points<9>", "bound": "fp64", "achieved": ach, "peak": fp64_peak, "unit": "TFLOP/s",
                         "frac": ach / fp64_peak if fp64_peak else None, "traffic": None, "flops_per_point": flops_per_point,
                         "peak_source": "DFMA loop measured in this run (rx_fp64_peak)",
                         "hbm_view": {"achieved": bytes_per_point * n / (ms_dev * 1e-3) / 1e9, "peak": hbm, "unit": "GB/s",
                                      "frac": bytes_per_point * n / (ms_dev * 1e-3) / 1e9 / hbm, "bytes_per_point": bytes_per_point}}}
    if world == 1 and not a.no_cpu_baseline:
        try:   # the CPU restatement (oracle, one core) on a bounded sample of the same states
            sys.path.insert(0, os.path.join(ROOT, "tests"))
            import ctypes as C
            import rxcase
            lib = C.CDLL(rxcase.build_oracle()) if os.path.exists(os.path.join(ROOT, "oracle", "rx_oracle.cpp")) else None
            ns = 200_000
            keep = []
            k = lambda x: (keep.append(np.ascontiguousarray(x, dtype=np.float64)), keep[-1])[1]
            ch = rxcase._Chem(chem["Ns"], chem["Nr"], chem["nKnots"], *[rxcase._p(k(np.asarray(chem[q]).reshape(-1))) for q in
                              ["mMasses", "diffVol", "x", "y", "y2", "As", "Betas", "Ta", "reversible", "availBack", "AsB", "BetasB", "TaB", "nuR", "nuP", "expR", "expP"]])
            res = np.empty((ns, chem["Ns"]))
            lib.rxo_source_points.restype = None
            t0 = time.perf_counter(); nrep = 0
            while time.perf_counter() - t0 < 10.0:
                lib.rxo_source_points(C.byref(ch), 1, C.c_double(0.09), C.c_double(0.2), C.c_int64(ns), rxcase._p(k(T[:ns])), rxcase._p(k(rho[:ns])),
                                      rxcase._p(k(Y[:ns])), rxcase._p(k(omT[:ns])), rxcase._p(k(vol[:ns])), rxcase._p(res))
                nrep += 1
            line["cpu_baseline"] = {"value": nrep * ns / (time.perf_counter() - t0), "unit": "updates/s", "cores": 1, "kind": "port",
                                    "sample": "%d passes of the CPU restatement (oracle/rx_oracle.cpp: rxo_source_points) over the first %d points, one core" % (nrep, ns)}
        except Exception as e:
            line["cpu_baseline"] = {"value": None, "unit": "updates/s", "cores": 0, "kind": "port", "sample": "failed: %s" % str(e)[:200]}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def bench_turbpost(a, pk, chem, rank, world, local, dev, K, W):
    """First slice of SURVEY.md 8(f) rank 1: CTurbSSTSolver::Postprocessing on the device (rx_turb_postprocess) over the points of
    the bench mesh (every rank its own 1415 x 1415 strip-sized channel; no exchange is timed: one partition per rank).
    value: device-timed passes of the two kernels on resident (k, omega); e2e: the host-buffer call — H2D of (k, omega),
    kernels, D2H of the twelve fields (rx_get_turbulence)."""
    import torch
    import torch.distributed as dist
    opts = pk.default_opts(rans=True, cfl=0.1)
    mesh = pk.channel_mesh(a.nx, a.ny)
    U0, T0, turb = pk.synthetic_state(mesh, chem, seed=12345 + rank, rans=True)
    solver = pk.CReactiveNSSolverGPU(mesh, chem, opts, device=local)
    solver.SetSolution(U0); solver.SetTemperature(T0); solver.SetTurbulence(turb)
    solver.Preprocessing(1, True, count_nonphysical=False)
    xy = np.asarray(mesh["coord"]); x, y = xy[:, 0], xy[:, 1]
    kw = np.ascontiguousarray(np.stack([1.0 + 0.5 * np.sin(40.0 * x) * np.cos(90.0 * y), 2.0e3 * (1.0 + 0.3 * np.cos(25.0 * x + 60.0 * y))], axis=1))
    wall = np.minimum(y - y.min(), y.max() - y) + 1e-6
    solver.SetWall_Distance(wall)
    n = solver.nPoint

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    solver.TurbPostprocessing(kw)
    for _ in range(max(W, 3)):
        solver.TurbPostprocessing(None)     # (k, omega) resident
    l0 = solver.LaunchCount()
    sampler = ClockSampler(local); sampler.start()
    barrier()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K):
        solver.TurbPostprocessing(None)
    e1.record()
    barrier()
    ms_dev = e0.elapsed_time(e1) / K
    launches = solver.LaunchCount() - l0
    solver.TurbPostprocessing(kw); out = solver.GetTurbulence()
    if not np.all(np.isfinite(out)):
        raise SystemExit("bench.py: non-finite turbulence fields")
    barrier()
    t0 = time.perf_counter()
    for _ in range(K):
        solver.TurbPostprocessing(kw); out = solver.GetTurbulence()
    barrier()
    ms_e2e = 1e3 * (time.perf_counter() - t0) / K
    clocks = sampler.stop()
    kt = {"k_turb_gradient_ls": solver.KernelTime(8, 5), "k_turb_blend": solver.KernelTime(9, 5)}
    tt = torch.tensor([ms_dev, ms_e2e], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    ms_dev, ms_e2e = float(tt[0]), float(tt[1])
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    # compulsory bytes per point: gradient kernel k, omega, x, y (32) + neighbour CSR (4 + 6 * 4) + 4 gradient words out (32);
    # blending kernel rho, mu, strain, k, omega, dist (48) + 4 gradient words (32) + 5 + 6 words out (88)
    bytes_per_point = 32.0 + 28.0 + 32.0 + 48.0 + 32.0 + 88.0
    ach = bytes_per_point * n / (ms_dev * 1e-3) / 1e9
    line = {"metric": "turbulence-postprocessing point updates/s", "value": world * n / (ms_dev * 1e-3), "unit": "updates/s", "n_gpus": world, "steps": K,
            "warmup": max(W, 3), "ms_per_step": ms_dev, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "SST Postprocessing (gradients of k and omega by weighted least squares, blending functions, eddy viscosity) over the points of the synthetic triangulated channel",
                       "points_per_gpu": int(n), "cache": "fields %.2f GB per GPU, larger than L2" % (n * 8 * 30 / 1e9)},
            "e2e": {"value": world * n / (ms_e2e * 1e-3), "unit": "updates/s", "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": int(n * 16), "d2h_bytes_per_step": int(n * 96)},
            "gpu_launches": int(launches), "clocks": clocks, "kernel_ms": kt,
            "roofline": {"kernel": "k_turb_gradient_ls + k_turb_blend", "bound": "hbm", "achieved": ach, "peak": hbm, "unit": "GB/s", "frac": ach / hbm,
                         "traffic": None, "bytes_per_point": bytes_per_point,
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs (measured)" if "hbm_gbs" in peaks else "fallback of B200_PROFILING.md"}}
    if world == 1 and not a.no_cpu_baseline:
        try:   # the CPU restatement (oracle/rx_oracle.cpp: rxo_turb_postprocess, one core) on a smaller mesh of the same kind
            sys.path.insert(0, os.path.join(ROOT, "tests"))
            import rxcase
            ms = pk.channel_mesh(501, 501)
            Us, Ts, tbs = pk.synthetic_state(ms, chem, seed=12345, rans=True)
            o = rxcase.Oracle.from_mesh(ms, chem, opts)
            o.set_solution(Us); o.set_temperature(Ts); o.set_turbulence(tbs); o.preprocess(1, output=True)
            xs = np.asarray(ms["coord"]); kws = np.ascontiguousarray(np.stack([1.0 + 0.5 * np.sin(40.0 * xs[:, 0]), 2.0e3 + 0.0 * xs[:, 0]], axis=1))
            ds = np.minimum(xs[:, 1] - xs[:, 1].min(), xs[:, 1].max() - xs[:, 1]) + 1e-6
            o.turb_postprocess(kws, ds)
            t0 = time.perf_counter(); reps = 0
            while time.perf_counter() - t0 < 10.0:
                o.turb_postprocess(kws, ds); reps += 1
            dt = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": reps * kws.shape[0] / dt, "unit": "updates/s", "cores": 1, "kind": "port",
                                    "sample": "%d passes of the CPU restatement (oracle/rx_oracle.cpp: rxo_turb_postprocess) over a 501 x 501 channel, one core" % reps}
        except Exception as ex:   # the oracle is test infrastructure: its absence must not hide the GPU numbers
            line["cpu_baseline"] = {"value": None, "unit": "updates/s", "cores": 0, "kind": "port", "sample": "unavailable: %s" % ex}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def physical_cores():
    """Physical core count of the host (unique (package, core) pairs of /proc/cpuinfo; logical count / 2 if that fails)."""
    try:
        pairs, pid, cid = set(), None, None
        for l in open("/proc/cpuinfo"):
            if l.startswith("physical id"):
                pid = l.split(":")[1].strip()
            elif l.startswith("core id"):
                cid = l.split(":")[1].strip()
            elif not l.strip():
                if pid is not None and cid is not None:
                    pairs.add((pid, cid))
                pid = cid = None
        if pairs:
            return len(pairs)
    except Exception:
        pass
    return max(1, (os.cpu_count() or 2) // 2)


# Extrapolation anchor: the compiled reference on the GPU arm's own 1415 x 1415 mesh (2.0 M points, 6.0 M edges), one
# serial process in the build container (same image, 8-core host), RUNGE-KUTTA 3 stages: 602.3 s and 508.8 s for two iterations =
# 47.2 k updates/s per core, against 50.8 k updates/s per core for the 151 x 151 sample mesh on the same host the same hour — the bounded
# sample overstates the CPU's per-update rate at full size by 7 %.  Measured once (it takes 25 minutes); quoted in every reference line.
ANCHOR_4M = {"mesh": "1415 x 1415 points, 6001016 edges (the GPU arm's mesh)", "seconds_per_iteration": [602.27, 508.81], "updates_per_s_one_core": 47189.0,
             "sample_mesh_updates_per_s_one_core_same_host": 50800.0, "ratio_full_size_to_sample": 0.93, "where": "build container, one serial process, measured once"}
REF_PROCS = 8   # fixed: one serial reference process per physical core, at most 8, so that every box times the same configuration


def measure_flow(pk, par, chem, mesh, plan, rans, rank, world, local, dev, K, W, full):
    """One flow workload (laminar or RANS) on this rank's mesh: W warm-up iterations from the synthetic state, then K timed
    iterations that ALL start from the same warmed state (rx_state_snapshot: device-to-device restore outside the CUDA events),
    so the timed work does not depend on K.  Returns a dict of local measurements (the caller reduces over ranks)."""
    import torch
    import torch.distributed as dist
    opts = pk.default_opts(rans=rans, cfl=0.1)
    U0, T0, turb = pk.synthetic_state(mesh, chem, seed=12345, rans=rans, glob_xy=plan.get("bbox"))
    solver = pk.CReactiveNSSolverGPU(mesh, chem, opts, device=local)
    ex = None
    if world > 1:
        par.library_comm_init(solver, plan, dev)   # halo exchange and global norms inside the library over NCCL (rx_comm_init): rx_iterate is the whole step
    nOwn, nE = solver.nPointDomain, solver.nEdge
    own = np.zeros(solver.nPoint, dtype=bool); own[:nOwn] = True
    en = mesh["edgeNodes"]
    nE_counted = int(own[en[:, 0]].sum())   # an edge shared with a neighbour partition is counted once (by the owner of its first node)
    units_local = len(RK_ALPHA) * (nE_counted + nOwn)
    Uh = torch.empty((solver.nPoint, solver.nVar), dtype=torch.float64).pin_memory()
    Uh.numpy()[:] = U0
    solver.SetSolutionPtr(Uh.data_ptr()); solver.SetTemperature(T0); solver.SetTurbulence(turb)

    # RANS: the SST transport step runs on the device after every flow iteration (rx_turb_iterate: residuals, boundary conditions,
    # FGMRES + LU-SGS, update, Postprocessing), as SU2's iteration does — on a partition too: the step exchanges its gradients, LU-SGS
    # and Krylov vectors and (k, omega) over the library's communicator and all-reduces its dot products (the reference's MPI points)
    turb_dev = bool(rans)
    lin_its = []
    if turb_dev:
        if world == 1:
            solver.ComputeWall_Distance()
            solver.TurbPostprocessing(np.ascontiguousarray(turb[:, 1:3]))
        else:   # a partition sees only its own wall nodes: the channel's walls are the lines y = y_min, y_max of the WHOLE mesh, each
                # point has a wall node straight below and above it, so the nearest-wall-node distance is the vertical one
            xy = np.asarray(mesh["coord"], dtype=np.float64)
            y0, y1 = plan["bbox"][2], plan["bbox"][3]
            solver.SetWall_Distance(np.ascontiguousarray(np.minimum(xy[:, 1] - y0, y1 - xy[:, 1])))
            par.distributed_turb_postprocess(solver, None, np.ascontiguousarray(turb[:, 1:3]))

    def step(it, want_rms=False):
        if world > 1:
            r = par.distributed_iterate(solver, ex, it, RK_ALPHA, last_rms=want_rms)
        else:
            r = solver.Iterate(it, RK_ALPHA, want_rms=want_rms)
        if turb_dev:
            lin_its.append(solver.TurbIterate(0.135, 2.0e3)[1])
        return r

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    it = 1
    for _ in range(max(W, 3)):
        step(it); it += 1
    solver.SaveState()                       # the state every timed step starts from
    l0 = solver.LaunchCount()
    sampler = ClockSampler(local); sampler.start()
    barrier()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    for e0, e1 in ev:
        solver.RestoreState()
        e0.record()
        step(it)
        e1.record()
    barrier()
    per_step = [e0.elapsed_time(e1) for e0, e1 in ev]
    ms_dev = float(sum(per_step) / K)
    clocks = sampler.stop()
    launches = solver.LaunchCount() - l0
    out = dict(solver=solver, units_local=units_local, ms_dev=ms_dev, ms_step_min=float(min(per_step)), ms_step_max=float(max(per_step)),
               launches=int(launches), clocks=clocks, nOwn=nOwn, nE=nE, nPoint=solver.nPoint, nVar=solver.nVar)
    solver.RestoreState()
    rms = step(it, want_rms=True)
    if not np.all(np.isfinite(rms)):
        raise SystemExit("bench.py: non-finite residual after the timed region")
    out["rms_rho"] = float(rms[0])
    out["turbulence"] = ("SST transport step on the device after every flow iteration (rx_turb_iterate; FGMRES iterations per step: %s), inside the timed region" % sorted(set(lin_its))) if turb_dev else \
                        None
    if solver.opts["viscous"]:
        _o, key = solver.EdgeOrder()         # BiCGSTAB iterations of every edge's Stefan-Maxwell solve in the timed state: the work is fixed
        hist = np.bincount(key.astype(np.int64) % 19, minlength=19)
        out["sm_iterations_histogram"] = {str(k): int(v) for k, v in enumerate(hist) if v}
    if not full:
        return out
    # End to end through the public call with HOST buffers, the same iteration from the same state: the secant start values and the
    # edge order are restored on the device, the conserved state comes from pinned host memory (rx_set_solution, H2D), the iteration
    # runs, the residual norms AND the updated conserved state go back to pinned host memory (rx_get_solution, D2H).
    solver.RestoreState()
    solver.GetSolutionPtr(Uh.data_ptr())     # host copy of the timed state
    Ur = torch.empty((solver.nPoint, solver.nVar), dtype=torch.float64).pin_memory()
    for _ in range(2):
        solver.RestoreState(); solver.SetSolutionPtr(Uh.data_ptr()); step(it, want_rms=True); solver.GetSolutionPtr(Ur.data_ptr())
    barrier()
    t0 = time.perf_counter()
    Ke = max(3, min(K, 10))
    for _ in range(Ke):
        solver.RestoreState()
        solver.SetSolutionPtr(Uh.data_ptr())
        step(it, want_rms=True)
        solver.GetSolutionPtr(Ur.data_ptr())
    barrier()
    out["ms_e2e"] = 1e3 * (time.perf_counter() - t0) / Ke
    out["fp64_peak"] = solver.Fp64Peak()
    out["kt"] = {n: solver.KernelTime(w, 5) for n, w in [("edge_flux_visc", 0), ("edge_flux_conv", 1), ("accumulate_source", 2), ("primitive", 3), ("transport", 7), ("gradient", 4), ("update", 6)]}
    return out


def reduce_flow(m, world, dev):
    import torch
    import torch.distributed as dist
    tt = torch.tensor([m["ms_dev"], m.get("ms_e2e", 0.0), float(m["units_local"])], dtype=torch.float64, device=dev)
    if world > 1:
        mx = tt.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = tt.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        return float(mx[0]), float(mx[1]), float(sm[2])
    return float(tt[0]), float(tt[1]), float(tt[2])


def measure_source(pk, chem, rank, world, local, dev, n, K):
    """configs[3] in short form for the default line: device-timed passes of k_source_points over n resident points per GPU."""
    import torch
    base = min(n, 4_000_000)
    T, rho, Y, omT, vol = pk.synthetic_points(chem, base, seed=12345 + rank)
    reps = (n + base - 1) // base
    t = lambda x: torch.as_tensor(np.ascontiguousarray(x), device=dev)
    tile = lambda x: t(x).repeat(reps)[:n].contiguous()
    dT, dr, dO, dV = tile(T), tile(rho), tile(omT), tile(vol)
    dY = t(Y.T).repeat(1, reps)[:, :n].contiguous()
    dR = torch.empty((chem["Ns"], n), dtype=torch.float64, device=dev)
    solver = pk.CReactiveNSSolverGPU(pk.channel_mesh(5, 4), chem, pk.default_opts(rans=True), device=local)
    f = lambda: solver.SourcePointsDev(n, dT.data_ptr(), dr.data_ptr(), dY.data_ptr(), dO.data_ptr(), dV.data_ptr(), dR.data_ptr())
    for _ in range(3):
        f()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K):
        f()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / K
    ok = bool(torch.isfinite(dR).all())
    solver.close()
    return ms, ok


def bench_dropin(a, K, W):
    """The real cost of the C++ drop-in: the reference's own driver / integration classes with the flow solver
    replaced by CReactiveNSSolverGPU (oracle/_ref/ref_harness_gpu) on a 251 001-point channel, at SU2_RX_MIRROR=0 (conserved variables and
    norms mirrored into node[] after every update) and 1 (+ primitives, gradients, transport after every Preprocessing), next to the
    compiled CPU reference on the same mesh.  Times are the harness's flow-iteration wall times (HARNESS_TIME)."""
    d = os.path.join(ROOT, "oracle", "_ref", "cases", "dropin_250k")
    hg = os.path.join(ROOT, "oracle", "_ref", "ref_harness_gpu")
    if not (os.path.exists(hg) and os.path.exists(os.path.join(d, "case.cfg"))):
        print(json.dumps({"metric": "dropin_e2e", "unavailable": "oracle/_ref/ref_harness_gpu or the dropin_250k case is not in this snapshot"}))
        return

    def run(binary, nit, env):
        t0 = time.perf_counter()
        out = subprocess.run([binary, "case.cfg", str(nit), "dump_dropin", str(10 ** 9), "11"], cwd=d, env=dict(os.environ, **env),
                             stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        shutil.rmtree(os.path.join(d, "dump_dropin"), ignore_errors=True)
        if out.returncode != 0:
            raise RuntimeError(out.stdout[-2000:])
        t = [float(m.group(1)) for m in re.finditer(r"HARNESS_TIME it \d+ flow_seconds ([0-9.eE+-]+)", out.stdout)]
        return t, time.perf_counter() - t0

    nit = W + K
    res = {}
    for mirror in ("0", "1"):
        t, wall = run(hg, nit, {"SU2_RX_BACKEND": "cuda", "SU2_RX_MIRROR": mirror})
        res["mirror%s" % mirror] = {"ms_per_iteration": 1e3 * float(np.mean(t[W:])), "iterations": K, "process_wall_s": wall}
    tc, wallc = run(HARNESS, 2, {})
    meta_pts, meta_edges = 251001, 751000
    units = len(RK_ALPHA) * (meta_edges + meta_pts)
    cpu_ms = 1e3 * float(tc[-1])
    line = {"metric": "dropin_e2e", "unit": "ms per flow iteration", "higher_is_better": False, "n_gpus": 1, "steps": K, "warmup": W, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "the reference's CFluidDriver / CMultiGridIntegration on CReactiveNSSolverGPU (ref_harness_gpu), laminar reactive NS, 3-stage RK, 501 x 501 channel",
                       "points": meta_pts, "edges": meta_edges},
            "dropin": res, "cpu_reference": {"ms_per_iteration": cpu_ms, "cores": 1, "iterations_timed": 1},
            "updates_per_s": {k: units / (v["ms_per_iteration"] * 1e-3) for k, v in res.items()},
            "speedup_vs_one_core": {k: cpu_ms / v["ms_per_iteration"] for k, v in res.items()}}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--nx", type=int, default=1415)
    ap.add_argument("--ny", type=int, default=1415)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra-workloads", action="store_true", help="skip the short RANS and source-term measurements embedded in the default line")
    ap.add_argument("--workload", default="laminar", choices=["laminar", "rans", "source", "turbpost", "dropin"],
                    help="laminar = BASELINE.json configs[1] (the bench line); rans = the RANS + PaSR variant of the same mesh (configs[2]/[4] per-GPU share); "
                         "source = isolated Arrhenius + PaSR source term over --points independent points per GPU (configs[3]); "
                         "turbpost = SST Postprocessing on the device (first slice of SURVEY.md 8(f) rank 1)")
    ap.add_argument("--points", type=int, default=64_000_000)
    ap.add_argument("--strong", action="store_true", help="strong scaling: --nx x --ny is the WHOLE mesh, split over the ranks (default: weak, --nx x --ny per rank)")
    ap.add_argument("--partition", default="strips", choices=["strips", "graph"],
                    help="strips: x-strips built per rank without the global mesh; graph: every rank builds the global mesh, partitions its point graph k-way "
                         "(rx_partition_kway + RCM numbering, the ParMETIS / SetRCM_Ordering stand-ins) and keeps its own part")
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    W = max(a.warmup, 0); K = max(a.steps, 1)
    rans = a.workload == "rans"
    WL = {False: "laminar reactive NS (9 species, 2 reactions), synthetic triangulated 2-D channel, explicit 3-stage RK, AUSM + Stefan-Maxwell viscous + Arrhenius source, WLS gradients",
          True: "reactive RANS (9 species, 2 reactions, SST), synthetic triangulated 2-D channel, explicit 3-stage RK, AUSM + Stefan-Maxwell viscous + SST closure + Arrhenius/PaSR source, WLS gradients"}
    workload = WL[rans]

    if a.workload == "dropin":
        if rank == 0:
            bench_dropin(a, K, W)
        return
    if a.impl == "reference":
        if rank != 0:
            return
        if rans or a.workload in ("source", "turbpost"):
            print(json.dumps({"impl": "reference", "unavailable": "the reference arm is prepared for the laminar bench workload (configs[1]) only"}))
            return
        r = run_reference(K, W, procs=min(physical_cores(), REF_PROCS))
        if r is None:
            print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/ref_harness or its case directory is missing (run oracle/build_ref.sh and oracle/make_golden.py in the build container)"}))
            return
        print(json.dumps({
            "impl": "reference", "metric": METRIC, "value": r["value"], "unit": "updates/s", "n_gpus": a.gpus, "steps": K, "warmup": W,
            "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload, "points": r["nPoint"], "edges": r["nEdge"], "rk_stages": len(RK_ALPHA),
                       "sample_note": "same cfg and mesh generator as the GPU arm at a bounded size (the CPU reference needs ~20 us per update: the 4.0M-cell mesh takes minutes per iteration); throughput is per update, see cpu_baseline.sample"},
            "cpu_baseline": {"value": r["value"], "unit": "updates/s", "cores": r["cores"], "kind": "reference", "sample": r["sample"],
                             "physical_cores_on_host": physical_cores(), "rss_mb_per_process": r.get("rss_mb"), "anchor_4m_cells": ANCHOR_4M},
            "e2e": {"value": r["value"], "unit": "updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return

    import importlib
    import torch
    pk = importlib.import_module(PKG)
    par = importlib.import_module(PKG + ".parallel")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the hot path has no CPU fallback")
    torch.cuda.set_device(local)
    numa_note = bind_to_gpu_numa(local) if world > 1 else None
    dev = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # stdout carries the one JSON line only
        dist.init_process_group("nccl", device_id=dev)
    chem = pk.load_mechanism()
    if a.workload == "source":
        return bench_source(a, pk, chem, rank, world, local, dev, K, W)
    if a.workload == "turbpost":
        return bench_turbpost(a, pk, chem, rank, world, local, dev, K, W)
    nxg = a.nx if a.strong else a.nx * world   # weak scaling: every rank owns an nx x ny strip; strong: nx x ny is the whole mesh
    t_mesh = time.perf_counter()
    if world > 1 and a.partition == "graph":
        gm = pk.channel_mesh(nxg, a.ny)
        part = pk.partition_graph_rank(gm, world, rank, rcm=True)   # this rank's part only (rx_partition_mesh)
        mesh, plan = part["mesh"], dict(send=part["send"], recv=part["recv"], bbox=(0.0, 0.25, 0.0, 0.025), glob=part["glob"])
        del gm, part
    elif world > 1:
        mesh, plan = pk.channel_strip_mesh(nxg, a.ny, rank, world)
        if a.strong:   # narrow strips: number the points with the short side slow, which shortens the level chains of the SST step's sweeps
            mesh, plan = pk.strip_short_side_slow(mesh, plan)
    else:
        mesh = pk.channel_mesh(nxg, a.ny); plan = dict(send={}, recv={}, bbox=None)
    t_mesh = time.perf_counter() - t_mesh
    m = measure_flow(pk, par, chem, mesh, plan, rans, rank, world, local, dev, K, W, full=True)
    ms_dev, ms_e2e, units = reduce_flow(m, world, dev)
    solver = m["solver"]; kt = m["kt"]; nOwn, nE = m["nOwn"], m["nE"]; clocks = m["clocks"]; launches = m["launches"]
    solver_fp64_peak = m["fp64_peak"]
    extra = {}
    if a.workload == "laminar" and not a.no_extra_workloads:
        solver.close()
        # the other two device workloads of BASELINE.json in short form, so that the driver's records carry them (full lines: --workload rans / source)
        Kx = max(3, min(K, 5))
        mr = measure_flow(pk, par, chem, mesh, plan, True, rank, world, local, dev, Kx, 3, full=False)
        r_ms, _e, r_units = reduce_flow(mr, world, dev)
        r_kt = mr["solver"].KernelTime(0, 3)
        mr["solver"].close()
        extra["rans"] = {"metric": METRIC, "value": r_units / (r_ms * 1e-3), "unit": "updates/s", "ms_per_step": r_ms, "steps": Kx, "n_gpus": world,
                         "workload": WL[True], "turbulence": mr.get("turbulence"),
                         "edge_flux_visc_ms": r_kt, "sm_iterations_histogram": mr.get("sm_iterations_histogram"), "gpu_launches": mr["launches"]}
        torch.cuda.empty_cache()
        npts = 64_000_000 if a.nx >= 1000 else 1_000_000
        s_ms, s_ok = measure_source(pk, chem, rank, world, local, dev, npts, Kx)
        tt = torch.tensor([s_ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        extra["source"] = {"metric": "source-point updates/s", "value": world * npts / (float(tt[0]) * 1e-3), "unit": "updates/s", "ms_per_step": float(tt[0]),
                           "points_per_gpu": npts, "steps": Kx, "n_gpus": world, "finite": s_ok,
                           "workload": "isolated Arrhenius + PaSR source term (9 species, 2 reactions) over independent points resident in HBM"}
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    hbm_src = "MEASURED_PEAKS.json hbm_gbs (measured)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
    # Dominant kernel: the viscous Stefan-Maxwell edge flux.  SURVEY.md 8(d) bounds it by the FP64 pipe, not by HBM.  Two operation
    # counts per edge are carried: the reference operation sequence (SURVEY.md's estimate, FMA = 2) and the FP64 operations the kernel
    # EXECUTES, counted by ncu (smsp__sass_thread_inst_executed_op_dadd/dmul/dfma_pred_on of the committed capture of this kernel at
    # this mesh size, profiles/r2_traffic.json); `achieved` uses the executed count when the capture matches the mesh.  The FP64 roof is
    # measured here, on this device, by a DFMA loop (rx_fp64_peak; MEASURED_PEAKS.json has no FP64 entry).  The HBM view of the same kernel
    # and the convective AUSM pass (HBM-bound) are carried next to it.
    fp64_peak = solver_fp64_peak
    flops_ref = 1.7e4 if rans else 1.5e4
    bytes_visc = 24.0 + (400.0 + 208.0) / (nE / float(nOwn))
    bytes_conv = 24.0 + (128.0 + 104.0) / (nE / float(nOwn))
    t_visc, t_conv = kt["edge_flux_visc"] * 1e-3, kt["edge_flux_conv"] * 1e-3
    traffic, flops_exec, prof_src, inst_exec, pipe_pct = None, None, None, None, None
    for f in ("r2_traffic.json", "r1_traffic.json"):
        try:   # dram bytes and executed FP64 operations of one launch from the committed ncu capture of this kernel at this workload size
            pr = json.load(open(os.path.join(ROOT, "profiles", f)))
            if pr.get("edges") == nE:
                q = pr.get("k_edge_visc_pair_rans" if rans else "k_edge_visc_pair")
                if isinstance(q, dict):
                    traffic, flops_exec = q.get("dram_bytes"), q.get("fp64_flops_per_edge")
                    inst_exec, pipe_pct = q.get("fp64_inst_per_edge"), q.get("fp64_pipe_pct_of_peak")
                else:
                    traffic = q
                prof_src = "profiles/" + f
                break
        except Exception:
            pass
    flops_per_edge = flops_exec or flops_ref
    ach_tf = flops_per_edge * nE / t_visc / 1e12
    ach_hbm_visc = bytes_visc * nE / t_visc / 1e9
    ach_hbm_conv = bytes_conv * nE / t_conv / 1e9
    line = {
        "metric": METRIC, "value": units / (ms_dev * 1e-3), "unit": "updates/s", "n_gpus": world, "steps": K, "warmup": max(W, 3),
        "ms_per_step": ms_dev, "higher_is_better": True, "scaling": "strong" if a.strong else "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload, "mesh_points_total": int(nxg * a.ny), "mesh_build_s_rank0": round(t_mesh, 1), "points_per_gpu": nOwn, "edges_per_gpu": nE, "cells_per_gpu": int(2 * (nxg - 1) * (a.ny - 1) / world) if a.strong else 2 * (a.nx - 1) * (a.ny - 1), "rk_stages": len(RK_ALPHA),
                   "cache": "inputs larger than L2 (state ~%.1f GB per GPU vs 126 MB L2); no flush needed" % (m["nPoint"] * 8 * 160 / 1e9),
                   "timed_state": "every timed step is the same iteration from the same state: the state after the warm-up iterations, restored device-to-device "
                                  "(rx_state_snapshot) before each step, outside the CUDA events; ms_per_step = mean of the per-step event times",
                   "partition": ("k-way graph partition (rx_partition_kway) with RCM numbering" if a.partition == "graph" else "strips in x") + ", one halo layer; halo U and gradients by ncclSend/ncclRecv inside the library (rx_halo_exchange), norms by one ncclAllReduce" if world > 1 else "single partition",
                   "host_affinity": numa_note},
        "e2e": {"value": units / (ms_e2e * 1e-3), "unit": "updates/s", "ms_per_step": ms_e2e,
                "h2d_bytes_per_step": int(m["nPoint"] * m["nVar"] * 8), "d2h_bytes_per_step": int(m["nPoint"] * m["nVar"] * 8 + m["nVar"] * 8),
                "what": "same iteration from the same state as `value`: conserved state H2D from pinned host memory (rx_set_solution), rx_iterate, residual norms and the UPDATED conserved state D2H to pinned host memory (rx_get_solution)"},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "step_ms_min_max": [m["ms_step_min"], m["ms_step_max"]],
        "turbulence": m.get("turbulence"),
        "sm_iterations_histogram": m.get("sm_iterations_histogram"),
        "roofline": {"kernel": "k_edge_visc_pair<9,%s> (Stefan-Maxwell/BiCGSTAB viscous edge flux%s, two lanes per edge)" % ("true" if rans else "false", " + SST closure (pivoted QR)" if rans else ""),
                     "bound": "fp64", "achieved": ach_tf, "peak": fp64_peak, "unit": "TFLOP/s", "frac": ach_tf / fp64_peak if fp64_peak else None,
                     "traffic": traffic, "traffic_source": prof_src,
                     "peak_source": "DFMA loop measured in this run (rx_fp64_peak); MEASURED_PEAKS.json has no FP64 entry",
                     "flops_per_edge": flops_per_edge, "flops_per_edge_source": ("executed FP64 operations counted by ncu, " + str(prof_src)) if flops_exec else "reference operation sequence (SURVEY.md 8(d) estimate)",
                     "flops_per_edge_reference_sequence": flops_ref, "frac_reference_sequence": flops_ref * nE / t_visc / 1e12 / fp64_peak if fp64_peak else None,
                     "bytes_per_edge": bytes_visc,
                     # the same kernel in FP64 INSTRUCTION slots: -fmad=false (parity with the x86-64 reference) leaves two thirds of the
                     # FP64 instructions as plain DADD / DMUL, which occupy a pipe slot for one FLOP; slots/s against the DFMA slot rate
                     # (peak / 2) is what the pipe sees (ncu: sm__pipe_fp64_cycles_active of the committed capture next to it)
                     "fp64_inst_per_edge": inst_exec,
                     "pipe_slot_frac": (inst_exec * nE / t_visc / (fp64_peak * 1e12 / 2.0)) if (inst_exec and fp64_peak) else None,
                     "ncu_fp64_pipe_pct_of_peak": pipe_pct,
                     "hbm_view": {"achieved": ach_hbm_visc, "peak": hbm, "unit": "GB/s", "frac": ach_hbm_visc / hbm, "peak_source": hbm_src},
                     "conv_kernel": {"kernel": CONV_KERNEL_NAME, "bound": "hbm", "achieved": ach_hbm_conv, "peak": hbm,
                                     "unit": "GB/s", "frac": ach_hbm_conv / hbm, "bytes_per_edge": bytes_conv, "peak_source": hbm_src}},
        "kernel_ms": kt,
    }
    if extra:
        line["workloads"] = extra
    if world == 1 and not a.no_cpu_baseline and not rans:
        try:
            r = run_reference(2, 1, procs=min(physical_cores(), REF_PROCS))
            if r:
                line["cpu_baseline"] = {"value": r["value"], "unit": "updates/s", "cores": r["cores"], "kind": "reference", "sample": r["sample"],
                                        "physical_cores_on_host": physical_cores(), "rss_mb_per_process": r.get("rss_mb"), "anchor_4m_cells": ANCHOR_4M}
        except Exception as e:   # the baseline is a reported number, never a reason to lose the GPU line
            line["cpu_baseline"] = {"value": None, "unit": "updates/s", "cores": 0, "kind": "reference", "sample": "failed: %s" % str(e)[:200]}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
