"""One-process-per-GPU plumbing: halo exchange of conserved variables and gradients between mesh partitions.

Replaces CReactiveEulerSolver::Set_MPI_Solution (solver_direct_reactive.cpp:1530-1636) and Set_MPI_Primitive_Gradient
(:1877-1993): per neighbour one SoA-packed message buf[iVar*nVertex + iVertex] — the reference's own packing — sent with
torch.distributed (NCCL over NVLink on GPUs, gloo in the CPU tests).  The pack/unpack kernels live in the C ABI
(rx_halo_pack / rx_halo_unpack); halo points of one neighbour are contiguous, so the unpack is a straight scatter.
The path has no other collective: residual norms are reduced once per iteration (a 13-double all-reduce).
"""
import torch
import torch.distributed as dist

WHAT_SOLUTION, WHAT_GRADIENT, WHAT_LIMITER, WHAT_TURB_GRADIENT = 0, 1, 2, 3   # 3: grad k, grad omega of TurbPostprocessing (stage 1 -> exchange -> stage 2)


class HaloExchanger:
    """state must provide halo_width(what), halo_pack(what, ids, buf), halo_unpack(what, ids, buf) on torch tensors."""

    def __init__(self, state, plan, device):
        self.state = state
        self.device = device
        self.send = {int(q): torch.as_tensor(ids, dtype=torch.int64, device=device) for q, ids in plan["send"].items()}
        self.recv = {int(q): torch.as_tensor(ids, dtype=torch.int64, device=device) for q, ids in plan["recv"].items()}
        self.bufs = {}

    def _buf(self, what, q, kind, n):
        key = (what, q, kind)
        if key not in self.bufs:
            self.bufs[key] = torch.empty(self.state.halo_width(what) * n, dtype=torch.float64, device=self.device)
        return self.bufs[key]

    def exchange(self, what):
        if not self.send and not self.recv:
            return 0
        ops = []
        for q in sorted(self.send):
            sb = self._buf(what, q, "s", self.send[q].numel())
            self.state.halo_pack(what, self.send[q], sb)
            ops.append(dist.P2POp(dist.isend, sb, q))
        for q in sorted(self.recv):
            rb = self._buf(what, q, "r", self.recv[q].numel())
            ops.append(dist.P2POp(dist.irecv, rb, q))
        for r in dist.batch_isend_irecv(ops):
            r.wait()
        nbytes = 0
        for q in sorted(self.recv):
            rb = self._buf(what, q, "r", self.recv[q].numel())
            self.state.halo_unpack(what, self.recv[q], rb)
            nbytes += rb.numel() * 8
        return nbytes


class GpuHaloState:
    """Adapter of CReactiveNSSolverGPU to the HaloExchanger interface (device pointers straight into the C ABI)."""

    def __init__(self, solver):
        self.s = solver

    def halo_width(self, what):
        return self.s.HaloWidth(what)

    def halo_pack(self, what, ids, buf):
        self.s.HaloPack(what, ids.numel(), ids.data_ptr(), buf.data_ptr())

    def halo_unpack(self, what, ids, buf):
        self.s.HaloUnpack(what, ids.numel(), ids.data_ptr(), buf.data_ptr())


class TensorHaloState:
    """CPU stand-in with the same packing (buf[v*n + k] = field[v][ids[k]]) for the gloo tests of the exchange logic."""

    def __init__(self, fields):
        self.fields = fields   # {what: tensor [nv, nPoint]}

    def halo_width(self, what):
        return self.fields[what].shape[0]

    def halo_pack(self, what, ids, buf):
        buf.copy_(self.fields[what][:, ids].reshape(-1))

    def halo_unpack(self, what, ids, buf):
        self.fields[what][:, ids] = buf.view(self.fields[what].shape[0], -1)


def library_comm_init(solver, plan, device):
    """Hand the halo exchange to the library (rx_comm_init): rank 0 draws the NCCL unique id through the library, torch.distributed
    (already initialised by the launcher) only carries its 128 bytes to the other ranks — the role MPI_Bcast has in an SU2 build.
    From here on rx_iterate runs a whole partitioned iteration, exchanges included, without the host in the loop."""
    from .rx_solver import comm_unique_id
    rank, world = dist.get_rank(), dist.get_world_size()
    t = torch.zeros(128, dtype=torch.uint8, device=device)
    if rank == 0:
        t.copy_(torch.frombuffer(bytearray(comm_unique_id()), dtype=torch.uint8))
    dist.broadcast(t, 0)
    solver.CommInit(bytes(t.cpu().numpy().tobytes()), rank, world, plan)


def global_rms(rms_local, n_local, n_owned, device):
    """Residual norms of the whole job from the per-partition ones (SetResidual_RMS, solver_structure.cpp:172-187: sqrt of
    the squares summed over all owned points / global owned point count): one all-reduce of nVar + 1 doubles.
    A context normalises its own norms by its local point count `n_local` (the serial rule), which is undone here."""
    import numpy as np
    t = torch.empty(len(rms_local) + 1, dtype=torch.float64)
    t[:-1] = torch.from_numpy(np.asarray(rms_local, dtype=np.float64) ** 2 * float(n_local))
    t[-1] = float(n_owned)
    t = t.to(device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t)
    t = t.cpu()
    return torch.sqrt(t[:-1] / t[-1]).numpy()


def distributed_iterate(solver, ex, ext_iter, rk_alpha, last_rms=False):
    """One MultiGrid_Cycle on a partition, with the reference's exchange points: gradients after every Preprocessing
    (Set_MPI_Primitive_Gradient), conserved variables after every update (Set_MPI_Solution)."""
    if getattr(solver, "has_comm", False):   # exchanges and the global norms inside the library (rx_comm_init)
        return solver.Iterate(ext_iter, rk_alpha, want_rms=last_rms)
    rms = None
    for rk, a in enumerate(rk_alpha):
        solver.Preprocessing(ext_iter, False, count_nonphysical=False)
        ex.exchange(WHAT_GRADIENT)
        if solver.opts.get("secondOrder") and solver.opts.get("limiter"):
            ex.exchange(WHAT_LIMITER)   # Set_MPI_Primitive_Limiter (:1756-1870)
        if rk == 0:
            solver.Set_OldSolution()
            solver.SetTime_Step()
        solver.Residual_All()
        solver.BC_All()
        rms = solver.ExplicitRK_Iteration(a, want_rms=(last_rms and rk == len(rk_alpha) - 1))
        ex.exchange(WHAT_SOLUTION)
    solver.Preprocessing(ext_iter, True, count_nonphysical=False)
    ex.exchange(WHAT_GRADIENT)
    if last_rms and rms is not None:
        rms = global_rms(rms, solver.nPoint, solver.nPointDomain, ex.device)
    return rms


def distributed_turb_postprocess(solver, ex, k_omega, sigma_k=0.85):
    """CTurbSSTSolver::Postprocessing on a partition: gradients of (k, omega) on the owned points, their halo values from the
    owners (Set_MPI_Solution_Gradient, solver_direct_turbulent.cpp:227), then blending functions and eddy viscosity everywhere."""
    solver.TurbPostprocessing(k_omega, sigma_k=sigma_k, stage=1)
    if getattr(solver, "has_comm", False):
        solver.HaloExchange(WHAT_TURB_GRADIENT)
    else:
        ex.exchange(WHAT_TURB_GRADIENT)
    solver.TurbPostprocessing(None, sigma_k=sigma_k, stage=2)
