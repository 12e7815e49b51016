"""N > 1 host logic on CPU: two (and three) processes over gloo exchange halo values of strip partitions with the
package's HaloExchanger (the same object bench.py drives over NCCL) and reduce the residual norms."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _field(gi, gj, v):
    """A value that identifies (global column, global row, variable)."""
    return 1000.0 * gi + gj + 0.001 * v


def _worker(rank, world, port, nxg, ny, out):
    sys.path.insert(0, HERE)
    from rxcase import load_package
    pk = load_package()
    par = __import__(pk.__name__ + ".parallel", fromlist=["x"])
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mesh, plan = pk.channel_strip_mesh(nxg, ny, rank, world)
        gi, gj = plan["glob"]
        nP, nOwn = mesh["nPoint"], mesh["nPointDomain"]
        fields = {}
        for what, nv in ((par.WHAT_SOLUTION, 13), (par.WHAT_GRADIENT, 26), (par.WHAT_LIMITER, 4), (par.WHAT_TURB_GRADIENT, 4)):
            f = torch.full((nv, nP), float("nan"), dtype=torch.float64)
            for v in range(nv):
                f[v, :nOwn] = torch.from_numpy(_field(gi[:nOwn], gj[:nOwn], v + 100 * what))
            fields[what] = f
        ex = par.HaloExchanger(par.TensorHaloState(fields), plan, torch.device("cpu"))
        nb = {}
        for what in fields:
            nb[what] = ex.exchange(what)
        ok = True
        for what, f in fields.items():
            nv = f.shape[0]
            want = np.stack([_field(gi, gj, v + 100 * what) for v in range(nv)])
            ok = ok and bool(np.array_equal(f.numpy(), want))
            ok = ok and nb[what] == 8 * nv * (nP - nOwn)
        # residual norm of the whole job from per-rank sums of squares (what distributed_iterate reduces)
        local = torch.tensor([float(nOwn), float((gi[:nOwn] ** 2).sum())], dtype=torch.float64)
        dist.all_reduce(local)
        ok = ok and int(local[0]) == nxg * ny
        ok = ok and float(local[1]) == float(sum(i * i for i in range(nxg)) * ny)
        out[rank] = 1 if ok else 0
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_halo_exchange_gloo(world):
    port = _free_port()
    ctx = mp.get_context("spawn")
    out = ctx.Array("i", [-1] * world)
    procs = [ctx.Process(target=_worker, args=(r, world, port, 17, 6, out)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
        assert p.exitcode == 0
    assert list(out) == [1] * world
