"""rx_residual_rms_static (include/rx_cuda.h): the allocation-free residual norms over a static __device__ scratch array,
checked through the C ABI against the existing entry point (rx_update_explicit's norms) on calls in a row.  The entry point
has no counterpart in the reference.  Collected last on purpose (file name)."""
import ctypes as C
import os

import numpy as np
import pytest

from rxcase import Case, GOLDEN, gpu_solver, load_package

TOL = 1e-12   # the tolerance of the parity tests (tests/test_gpu_parity.py)


def _dev_residual(torch, R):
    """LinSysRes[nPoint][nVar] (host, AoS) -> device SoA [nVar][ld] with a padded leading dimension."""
    nP, nVar = R.shape
    ld = (nP + 15) // 16 * 16
    t = torch.full((nVar, ld), 7.0e300, dtype=torch.float64, device="cuda")   # padding must never be read
    t[:, :nP] = torch.from_numpy(np.ascontiguousarray(R.T)).cuda()
    return t, ld


@pytest.mark.gpu
def test_static_scratch_rms_matches_update_explicit_on_calls_in_a_row():
    import torch
    pk = load_package()
    case = Case(os.path.join(GOLDEN, "channel_lam_rk.npz"))
    it, alphas = 1, case.meta["cfg"]["rk_alpha"]
    s = gpu_solver(case)
    s.set_solution(case.stage(it, 0, "U0")); s.set_temperature(case.stage(it, 0, "prev_V")[:, 0]); s.set_turbulence(case.stage(it, 0, "turb"))
    nP, nD, nVar = s.nPoint, s.nPointDomain, s.nVar
    warmed = False
    worst = 0.0
    out = torch.zeros(nVar, dtype=torch.float64, device="cuda")
    for rk, a in enumerate(alphas):   # one call of the new entry point per RK stage: three calls in a row on the shared scratch
        s.preprocess(it, False)
        if rk == 0:
            s.set_old_solution(); s.time_step()
        s.upwind_residual(); s.viscous_residual(); s.source_residual(); s.boundary_conditions()
        R = s.residual()
        Rt, ld = _dev_residual(torch, R)
        ref = s.explicit_update(a)                      # existing entry point: norms of the same LinSysRes
        torch.cuda.synchronize()
        if not warmed:   # the first call takes the symbol address (and lazily loads the kernel's code)
            pk.residual_rms_static(Rt.data_ptr(), nD, ld, nVar, nP, out.data_ptr())
            torch.cuda.synchronize()
            warmed = True
        out.fill_(-1.0)
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        pk.residual_rms_static(Rt.data_ptr(), nD, ld, nVar, nP, out.data_ptr())
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] == free0, "the no-alloc entry point changed the device's free memory"
        got = out.cpu().numpy()
        rel = np.abs(got - ref) / np.abs(ref)
        worst = max(worst, rel.max())
        assert rel.max() <= TOL, (rk, got, ref)
    # a smaller call after the larger ones must not see their stale partial sums; a one-point call; a non-default stream
    n2 = 200
    sub = torch.from_numpy(np.ascontiguousarray(R.T[:, :n2])).cuda().contiguous()
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        pk.residual_rms_static(sub.data_ptr(), n2, n2, nVar, n2, out.data_ptr(), stream=st.cuda_stream)
    st.synchronize()
    want = np.sqrt((R[:n2] ** 2).sum(axis=0) / n2)
    got = out.cpu().numpy()
    assert np.abs(got - np.maximum(want, 1e-32)).max() <= TOL * np.abs(want).max()
    pk.residual_rms_static(sub.data_ptr(), 1, n2, nVar, 1, out.data_ptr())
    torch.cuda.synchronize()
    assert np.abs(out.cpu().numpy() - np.maximum(np.abs(R[0]), 1e-32)).max() <= TOL * np.abs(R[0]).max()
    print("rx_residual_rms_static vs rx_update_explicit: max rel diff %.3e over %d calls" % (worst, len(alphas)))
    s.close()


@pytest.mark.gpu
def test_static_scratch_rms_rejects_sizes_beyond_the_maximum():
    import torch
    pk = load_package()
    rs = __import__("importlib").import_module(pk.__name__ + ".rx_solver")
    buf = torch.zeros(64, dtype=torch.float64, device="cuda")
    with pytest.raises(pk.RxError) as e:   # rejected before any device memory is touched
        pk.residual_rms_static(buf.data_ptr(), rs.STATIC_RMS_MAX_POINTS + 1, rs.STATIC_RMS_MAX_POINTS + 1, 13, 1, buf.data_ptr())
    assert "exceed the static scratch" in str(e.value) and str(rs.STATIC_RMS_MAX_POINTS) in str(e.value)
    with pytest.raises(pk.RxError) as e:
        pk.residual_rms_static(buf.data_ptr(), 4, 4, rs.STATIC_RMS_MAX_VARS + 1, 4, buf.data_ptr())
    assert "exceed the static scratch" in str(e.value)
    with pytest.raises(pk.RxError):
        pk.residual_rms_static(buf.data_ptr(), 8, 4, 2, 8, buf.data_ptr())   # ld < nPointDomain


def test_header_limits_match_python_mirror():
    """CPU: the limits stated in include/rx_cuda.h are the ones the Python mirror quotes."""
    import re
    from rxcase import ROOT
    pk = load_package()
    rs = __import__("importlib").import_module(pk.__name__ + ".rx_solver")
    txt = open(os.path.join(ROOT, "include", "rx_cuda.h")).read()
    assert int(re.search(r"#define RX_STATIC_RMS_MAX_POINTS (\d+)", txt).group(1)) == rs.STATIC_RMS_MAX_POINTS
    assert int(re.search(r"#define RX_STATIC_RMS_MAX_VARS (\d+)", txt).group(1)) == rs.STATIC_RMS_MAX_VARS
