"""Partitioned == single partition on the GPU: the strip partitions of a channel, each in its own solver context on cuda:0
and exchanging halos through the C ABI pack/unpack kernels (the path bench.py --gpus N drives over NCCL), reproduce the
single-partition iteration on the owned points."""
import numpy as np
import pytest

import os

from rxcase import GOLDEN, Case, chem_arrays, load_package, mesh_dict, opts_dict

pytestmark = pytest.mark.gpu

RK = [0.66667, 0.66667, 1.0]


class LocalExchanger:
    """HaloExchanger's message pattern between contexts of one process: pack on the sender, unpack on the receiver."""

    def __init__(self, solvers, plans):
        import torch
        self.torch = torch
        self.solvers, self.plans = solvers, plans
        dev = torch.device("cuda", 0)
        self.send = [{q: torch.as_tensor(np.asarray(ids), dtype=torch.int64, device=dev) for q, ids in p["send"].items()} for p in plans]
        self.recv = [{q: torch.as_tensor(np.asarray(ids), dtype=torch.int64, device=dev) for q, ids in p["recv"].items()} for p in plans]
        self.dev = dev

    def exchange(self, what):
        torch = self.torch
        for r, s in enumerate(self.solvers):
            for q, ids in self.send[r].items():
                w = s.HaloWidth(what)
                buf = torch.empty(w * ids.numel(), dtype=torch.float64, device=self.dev)
                s.HaloPack(what, ids.numel(), ids.data_ptr(), buf.data_ptr())
                rid = self.recv[q][r]
                assert rid.numel() == ids.numel()
                self.solvers[q].HaloUnpack(what, rid.numel(), rid.data_ptr(), buf.data_ptr())
        torch.cuda.synchronize()


@pytest.mark.parametrize("nranks", [2, 3])
def test_partitioned_iteration_equals_single(nranks):
    pk = load_package()
    par = __import__(pk.__name__ + ".parallel", fromlist=["x"])
    nxg, ny = 48, 20
    chem = pk.load_mechanism()
    opts = pk.default_opts(rans=False, cfl=0.1)
    g = pk.channel_mesh(nxg, ny)
    Ug, Tg, tg = pk.synthetic_state(g, chem, seed=5)
    ref = pk.CReactiveNSSolverGPU(g, chem, opts)
    ref.SetSolution(Ug); ref.SetTemperature(Tg); ref.SetTurbulence(tg)
    solvers, plans, globs = [], [], []
    for r in range(nranks):
        m, plan = pk.channel_strip_mesh(nxg, ny, r, nranks)
        U, T, tb = pk.synthetic_state(m, chem, seed=5, glob_xy=plan["bbox"])
        glob = plan["glob"][1] * nxg + plan["glob"][0]
        np.testing.assert_array_equal(U, Ug[glob])           # the synthetic state is a function of the coordinates
        s = pk.CReactiveNSSolverGPU(m, chem, opts)
        s.SetSolution(U); s.SetTemperature(T); s.SetTurbulence(tb)
        solvers.append(s); plans.append(plan); globs.append(glob)
    ex = LocalExchanger(solvers, plans)
    for it in (1, 2):
        rms_ref = ref.Iterate(it, RK, want_rms=True)
        # the reference's order: every rank runs a stage, then the exchange (lock step)
        for rk, a in enumerate(RK):
            for s in solvers:
                s.Preprocessing(it, False, count_nonphysical=False)
            ex.exchange(par.WHAT_GRADIENT)
            rms = []
            for s in solvers:
                if rk == 0:
                    s.Set_OldSolution(); s.SetTime_Step()
                s.Residual_All(); s.BC_All()
                rms.append(s.ExplicitRK_Iteration(a, want_rms=True))
            ex.exchange(par.WHAT_SOLUTION)
        for s in solvers:
            s.Preprocessing(it, True, count_nonphysical=False)
        ex.exchange(par.WHAT_GRADIENT)
        sumsq = sum(np.asarray(r) ** 2 * s.nPoint for r, s in zip(rms, solvers))   # a context normalises by its local point count
        rms_all = np.sqrt(sumsq / sum(s.nPointDomain for s in solvers))
        Ur = ref.GetSolution()
        for s, glob in zip(solvers, globs):
            n = s.nPointDomain
            U = s.GetSolution()[:n]
            want = Ur[glob[:n]]
            for grp in ([0] + list(range(4, 13)), [1, 2], [3]):
                err = np.abs(U[:, grp] - want[:, grp]).max() / np.abs(want[:, grp]).max()
                assert err < 1e-12, (it, grp, err)
        np.testing.assert_allclose(rms_all, rms_ref, rtol=1e-9)


@pytest.mark.parametrize("nparts,rcm", [(2, False), (5, True)])
def test_graph_partitioned_rans_iteration_equals_single(nparts, rcm):
    """The shipped UNSTRUCTURED jet-reactor mesh (RANS + PaSR, every BC kind of the case) split by the library's k-way graph
    partitioner (rx_partition_kway, the ParMETIS stand-in), with and without RCM renumbering of the parts: two explicit
    iterations with halo exchange reproduce the single-partition run on the owned points."""
    pk = load_package()
    par = __import__(pk.__name__ + ".parallel", fromlist=["x"])
    case = Case(os.path.join(GOLDEN, "jet_rans_perturbed.npz"))
    g, chem, opts = mesh_dict(case), chem_arrays(case), opts_dict(case)
    it0 = 1
    Ug, Tg, tg = case.stage(it0, 0, "U0"), case.stage(it0, 0, "prev_V")[:, 0], case.stage(it0, 0, "turb")
    alphas = case.meta["cfg"]["rk_alpha"]
    ref = pk.CReactiveNSSolverGPU(g, chem, opts)
    ref.SetSolution(Ug); ref.SetTemperature(Tg); ref.SetTurbulence(tg)
    owner, cut = pk.graph_owner(g, nparts)
    parts = pk.partition_mesh(g, owner, rcm=rcm)
    solvers = []
    for p in parts:
        s = pk.CReactiveNSSolverGPU(p["mesh"], chem, opts)
        s.SetSolution(Ug[p["glob"]]); s.SetTemperature(Tg[p["glob"]]); s.SetTurbulence(tg[p["glob"]])
        solvers.append(s)
    ex = LocalExchanger(solvers, parts)
    for it in (it0, it0 + 1):
        ref.Iterate(it, alphas, want_rms=False)
        for rk, a in enumerate(alphas):
            for s in solvers:
                s.Preprocessing(it, False, count_nonphysical=False)
            ex.exchange(par.WHAT_GRADIENT)
            for s in solvers:
                if rk == 0:
                    s.Set_OldSolution(); s.SetTime_Step()
                s.Residual_All(); s.BC_All()
                s.ExplicitRK_Iteration(a, want_rms=False)
            ex.exchange(par.WHAT_SOLUTION)
        for s in solvers:
            s.Preprocessing(it, True, count_nonphysical=False)
        ex.exchange(par.WHAT_GRADIENT)
        Ur = ref.GetSolution()
        for s, p in zip(solvers, parts):
            n = s.nPointDomain
            U = s.GetSolution()[:n]
            want = Ur[p["glob"][:n]]
            for grp in ([0] + list(range(4, 13)), [1, 2], [3]):
                err = np.abs(U[:, grp] - want[:, grp]).max() / np.abs(want[:, grp]).max()
                assert err < 1e-12, (it, grp, err)
