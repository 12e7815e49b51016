"""Processing order of the viscous edge kernel (csrc/rx_order.cuh): within every tile of consecutive edges the order is a stable
counting sort by the key the kernel's last launch left, and the iteration does not depend on it — with and without ordering, same bits."""
import os

import numpy as np
import pytest

from rxcase import load_package

pytestmark = pytest.mark.gpu

RK = [0.66667, 0.66667, 1.0]
ORD_TILE = 2048   # csrc/rx_order.cuh: edges are re-ordered within tiles of this many consecutive edges


def _run(pk, rans, order, iters=3, nx=97, ny=61):
    chem = pk.load_mechanism()
    opts = pk.default_opts(rans=rans, cfl=0.1)
    mesh = pk.channel_mesh(nx, ny)
    U0, T0, turb = pk.synthetic_state(mesh, chem, seed=77, rans=rans)
    old = os.environ.get("RX_VISC_ORDER")
    os.environ["RX_VISC_ORDER"] = "1" if order else "0"   # read by rx_create; unset: on for laminar, off for RANS
    try:
        s = pk.CReactiveNSSolverGPU(mesh, chem, opts)
    finally:
        if old is None:
            del os.environ["RX_VISC_ORDER"]
        else:
            os.environ["RX_VISC_ORDER"] = old
    s.SetSolution(U0); s.SetTemperature(T0); s.SetTurbulence(turb)
    rms, orders = [], []
    for it in range(iters):
        rms.append(np.array(s.Iterate(it + 1, RK)))
        orders.append(s.EdgeOrder())
    return s.GetSolution(), rms, orders


@pytest.mark.parametrize("rans", [False, True])
def test_order_is_stable_sort_of_previous_keys_and_results_do_not_depend_on_it(rans):
    pk = load_package()
    Ua, rmsA, ordA = _run(pk, rans, order=True)
    Ub, rmsB, ordB = _run(pk, rans, order=False)
    # same bits
    np.testing.assert_array_equal(Ua, Ub)
    for a, b in zip(rmsA, rmsB):
        np.testing.assert_array_equal(a, b)
    nE = ordA[0][0].shape[0]
    # without ordering: identity, but the keys are recorded all the same and agree
    for (oa, ka), (ob, kb) in zip(ordA, ordB):
        np.testing.assert_array_equal(ob, np.arange(nE, dtype=np.int32))
        np.testing.assert_array_equal(ka, kb)
    # iteration 1 ran in edge order (no keys yet); the order of iteration k+1 is the tile-wise stable sort by the keys iteration k left
    np.testing.assert_array_equal(ordA[0][0], np.arange(nE, dtype=np.int32))
    for k in range(1, len(ordA)):
        keys = np.minimum(ordA[k - 1][1], 39)
        expect = np.concatenate([t0 + np.argsort(keys[t0:t0 + ORD_TILE], kind="stable") for t0 in range(0, nE, ORD_TILE)]).astype(np.int32)
        np.testing.assert_array_equal(ordA[k][0], expect)
    keys = ordA[-1][1]
    assert keys.max() <= 18 + (19 if rans else 0)
    assert (keys > 0).any()   # some edge did iterate
