"""Worker of tests/test_gpu_nccl.py (launched with torch.distributed.run, one rank per GPU): a partitioned run whose halo exchange
and global norms happen INSIDE the library over NCCL (rx_comm_init / rx_halo_exchange / rx_iterate) must equal the single-partition run."""
import importlib
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pk = importlib.import_module("development-of-a-turbulent-numerical-solver-for-reactive-flows-in-su2_b200")
par = importlib.import_module(pk.__name__ + ".parallel")


def turb_mode(rank, world, local, dev, chem):
    """RANS with the SST transport step ON THE DEVICE on every partition (rx_turb_iterate with a communicator: halo exchanges of
    grad(k, omega), of x* between the LU-SGS sweeps, of the preconditioned vector, of (k, omega) after the update; dot products and
    residual norms all-reduced).  Every rank also runs the whole mesh alone as the reference.
      (1) linear system solved to 1e-13 (60 FGMRES iterations allowed): partitioned and single runs solve the SAME global system,
          so (k, omega, mu_t) and the flow state after 3 coupled iterations must agree to 1e-9 — this checks assembly, matrix-vector
          product, dot products and every exchange except the preconditioner's;
      (2) the production setting (5 iterations, tolerance 1e-6): the partition-local LU-SGS is a different (block-Jacobi-coupled)
          preconditioner, as in the reference's MPI build, so the iterates differ at the level of the linear residual: agreement of
          the update to 5 % of its own size, same iteration counts +- 2."""
    opts = pk.default_opts(rans=True, cfl=0.1)
    nx, ny = 61, 23
    g = pk.channel_mesh(nx, ny)
    plan = pk.partition_graph_rank(g, world, rank, rcm=True)
    m = plan["mesh"]
    glob = np.asarray(plan["glob"])
    Ug, Tg, tg = pk.synthetic_state(g, chem, seed=3, rans=True)
    alphas = [0.66667, 0.66667, 1.0]
    ok = True
    report = []
    for label, lin_iter, lin_tol, tol in (("tight", 60, 1e-13, 1e-9), ("production", 5, 1e-6, None)):
        ref = pk.CReactiveNSSolverGPU(g, chem, opts, device=local)
        ref.SetSolution(Ug); ref.SetTemperature(Tg); ref.SetTurbulence(tg)
        ref.ComputeWall_Distance()
        dist_g = ref.GetTurbulence()[:, 2].copy()
        ref.TurbPostprocessing(np.ascontiguousarray(tg[:, 1:3]))
        t0 = ref.GetTurbulence()
        s = pk.CReactiveNSSolverGPU(m, chem, opts, device=local)
        s.SetSolution(Ug[glob]); s.SetTemperature(Tg[glob]); s.SetTurbulence(tg[glob])
        par.library_comm_init(s, plan, dev)
        # wall distance of a partition: the wall nodes of all partitions gathered inside the library (rx_compute_wall_distance with a
        # communicator) — the same distances, bit for bit, as the whole mesh gives; owned and halo points alike
        s.ComputeWall_Distance()
        wd_bad = int(np.sum(s.GetTurbulence()[:, 2] != dist_g[glob]))
        tw = torch.tensor([float(wd_bad)], dtype=torch.float64, device=dev)
        dist.all_reduce(tw, op=dist.ReduceOp.MAX)
        if float(tw[0]) != 0.0:
            ok = False
            report.append("%s: partitioned wall distance differs at %d points" % (label, int(tw[0])))
        par.distributed_turb_postprocess(s, None, np.ascontiguousarray(tg[glob, 1:3]))
        its_p, its_r = [], []
        for it in range(3):
            ref.Iterate(it + 1, alphas); rr, nr = ref.TurbIterate(0.135, 2.0e3, lin_iter=lin_iter, lin_tol=lin_tol)
            s.Iterate(it + 1, alphas); rp, npp = s.TurbIterate(0.135, 2.0e3, lin_iter=lin_iter, lin_tol=lin_tol)
            its_p.append(npp); its_r.append(nr)
        n = s.nPointDomain
        tr, tp = ref.GetTurbulence(), s.GetTurbulence()
        Ur, Up = ref.solution(), s.solution()
        own = glob[:n]
        cols = [0, 1, 6]   # k, omega, mu_t
        scale = np.abs(tr[:, cols]).max(axis=0)
        err_t = (np.abs(tp[:n][:, cols] - tr[own][:, cols]).max(axis=0) / scale).max()
        err_h = (np.abs(tp[n:][:, cols] - tr[glob[n:]][:, cols]).max(axis=0) / scale).max() if len(glob) > n else 0.0   # halo copies after the exchange
        err_u = (np.abs(Up[:n] - Ur[own]).max(axis=0) / np.abs(Ur).max(axis=0)).max()
        upd = (np.abs(tr[:, cols] - t0[:, cols]).max(axis=0) / scale).max()
        rms_err = float(np.abs(rp / rr - 1.0).max())
        t = torch.tensor([err_t, err_h, err_u, rms_err], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        err_t, err_h, err_u, rms_err = (float(x) for x in t.cpu())
        if tol is not None:
            good = max(err_t, err_h, err_u) <= tol and rms_err <= 1e-9
        else:
            good = max(err_t, err_h) <= 0.05 * upd and all(abs(a - b) <= 2 for a, b in zip(its_p, its_r))
        ok = ok and good
        report.append("%s: turb %.2e halo %.2e flow %.2e rms %.2e update %.2e iters part %s single %s" % (label, err_t, err_h, err_u, rms_err, upd, its_p, its_r))
        s.close(); ref.close()
    if rank == 0:
        print("NCCL_WORKER mode turb ranks %d | %s | %s" % (world, " | ".join(report), "OK" if ok else "FAIL"))


def unsteady_mode(rank, world, local, dev, chem):
    """UNSTEADY_SIMULATION= TIME_STEPPING on partitions: the global minimum time step is the minimum over ALL partitions (one ncclAllReduce
    with ncclMin inside rx_time_step, SetTime_Step :5264-5269), so two iterations of a partitioned run equal the single-partition run."""
    opts = dict(pk.default_opts(rans=False, cfl=0.1), timeStepping=1, unstCfl=0.1, deltaUnstTime=1e-7)
    nx, ny = 61, 23
    g = pk.channel_mesh(nx, ny)
    m, plan = pk.channel_strip_mesh(nx, ny, rank, world)
    glob = plan["glob"][1] * nx + plan["glob"][0]
    Ug, Tg, tg = pk.synthetic_state(g, chem, seed=3, rans=False)
    s = pk.CReactiveNSSolverGPU(m, chem, opts, device=local)
    s.SetSolution(Ug[glob]); s.SetTemperature(Tg[glob]); s.SetTurbulence(tg[glob])
    par.library_comm_init(s, plan, dev)
    ref = pk.CReactiveNSSolverGPU(g, chem, opts, device=local)
    ref.SetSolution(Ug); ref.SetTemperature(Tg); ref.SetTurbulence(tg)
    alphas = [0.66667, 0.66667, 1.0]
    for it in range(2):
        s.Iterate(it + 1, alphas); ref.Iterate(it + 1, alphas)
    n = s.nPointDomain
    dt_p, dt_r = s.dt()[:n, 0], ref.dt()[:, 0]
    same_dt = bool(np.all(dt_p == dt_p[0]) and np.all(dt_r == dt_r[0]) and dt_p[0] == dt_r[0])
    Ur = ref.solution()[glob[:n]]
    err = float((np.abs(s.solution()[:n] - Ur).max(axis=0) / np.abs(Ur).max(axis=0)).max())
    t = torch.tensor([0.0 if same_dt else 1.0, err], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    s.close(); ref.close()
    if rank == 0:
        ok = float(t[0]) == 0.0 and float(t[1]) <= 1e-12
        print("NCCL_WORKER mode unsteady ranks %d dt_equal %s worst %.3e %s" % (world, float(t[0]) == 0.0, float(t[1]), "OK" if ok else "FAIL"))


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    chem = pk.load_mechanism()
    mode = sys.argv[1] if len(sys.argv) > 1 else "strips"
    if mode in ("turb", "unsteady"):
        (turb_mode if mode == "turb" else unsteady_mode)(rank, world, local, dev, chem)
        dist.destroy_process_group()
        return
    worst = 0.0
    for rans in (False, True):
        opts = pk.default_opts(rans=rans, cfl=0.1)
        nx, ny = 61, 23
        g = pk.channel_mesh(nx, ny)
        if mode == "graph":
            plan = pk.partition_graph_rank(g, world, rank, rcm=True)   # deterministic: every rank computes the same owner array and builds its own part
            m = plan["mesh"]
            glob = np.asarray(plan["glob"])
        else:
            m, plan = pk.channel_strip_mesh(nx, ny, rank, world)
            glob = plan["glob"][1] * nx + plan["glob"][0]
        Ug, Tg, tg = pk.synthetic_state(g, chem, seed=3, rans=rans)
        s = pk.CReactiveNSSolverGPU(m, chem, opts, device=local)
        s.SetSolution(Ug[glob]); s.SetTemperature(Tg[glob]); s.SetTurbulence(tg[glob])
        par.library_comm_init(s, plan, dev)
        alphas = [0.66667, 0.66667, 1.0]
        for it in range(2):
            rms = s.Iterate(it + 1, alphas)
        U = s.solution()[:s.nPointDomain]
        # rank 0 gathers the owned points and compares with its own single-partition run
        sizes = [torch.zeros(1, dtype=torch.int64, device=dev) for _ in range(world)]
        dist.all_gather(sizes, torch.tensor([s.nPointDomain], dtype=torch.int64, device=dev))
        nmax = int(max(x.item() for x in sizes))
        pad = torch.zeros((nmax, s.nVar + 1), dtype=torch.float64, device=dev)
        pad[:s.nPointDomain, :s.nVar] = torch.as_tensor(U, device=dev)
        pad[:s.nPointDomain, s.nVar] = torch.as_tensor(glob[:s.nPointDomain].astype(np.float64), device=dev)
        allp = [torch.zeros_like(pad) for _ in range(world)]
        dist.all_gather(allp, pad)
        s.close()
        if rank == 0:
            ref = pk.CReactiveNSSolverGPU(g, chem, opts, device=local)
            ref.SetSolution(Ug); ref.SetTemperature(Tg); ref.SetTurbulence(tg)
            for it in range(2):
                rms_ref = ref.Iterate(it + 1, alphas)
            Ur = ref.solution()
            ref.close()
            got = np.zeros_like(Ur)
            for q in range(world):
                a = allp[q][:int(sizes[q].item())].cpu().numpy()
                got[a[:, -1].astype(np.int64)] = a[:, :-1]
            sp = list(range(4, Ur.shape[1]))
            for grp in ([0] + sp, [1, 2], [3]):
                worst = max(worst, np.abs(got[:, grp] - Ur[:, grp]).max() / np.abs(Ur[:, grp]).max())
            live = rms_ref > 1e-30
            worst = max(worst, np.abs(rms[live] / rms_ref[live] - 1.0).max())
    if rank == 0:
        print("NCCL_WORKER mode %s ranks %d worst %.3e %s" % (mode, world, worst, "OK" if worst <= 1e-12 else "FAIL"))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
