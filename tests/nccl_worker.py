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


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    chem = pk.load_mechanism()
    mode = sys.argv[1] if len(sys.argv) > 1 else "strips"
    worst = 0.0
    for rans in (False, True):
        opts = pk.default_opts(rans=rans, cfl=0.1)
        nx, ny = 61, 23
        g = pk.channel_mesh(nx, ny)
        if mode == "graph":
            plan = pk.partition_graph(g, world, rcm=True)[rank]   # deterministic: every rank computes the same partition
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
