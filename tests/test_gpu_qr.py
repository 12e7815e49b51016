"""The device's two forms of Eigen::ColPivHouseholderQR<MatrixXd>::solve (the SST closure's molar -> mass gradient solve,
numerics_direct_reactive.cpp:743) against the oracle's restatement — which oracle/eigen_qr_check.cpp pins, bit for bit, against the
real Eigen 3.3.7 headers the reference vendors (tests/test_oracle_golden.py::test_qr_restatement_equals_eigen_bit_for_bit).

Bar: every entry of every solution IDENTICAL (same bits; +0 and -0 taken as equal), for the serial routine (boundary vertices) and for
the two-lanes-per-system routine of the viscous edge kernel, on the closure's operator for random compositions, dense random
matrices, rank-deficient matrices, zero rows / columns and the zero matrix (pivot threshold, tau = 0 reflectors)."""
import numpy as np
import pytest

from rxcase import load_package
from test_oracle_golden import oracle_qr, qr_test_systems

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n", [1, 63, 20000])
def test_device_qr_equals_eigen_order_oracle(n):
    pk = load_package()
    A, b = qr_test_systems(n, seed=11 + n)
    want = oracle_qr(A, b)
    Acm = np.ascontiguousarray(np.transpose(A, (0, 2, 1))).reshape(n, 81)
    xs, xp = pk.qr_eval(Acm, b)
    for name, got in (("serial", xs), ("pair", xp)):
        bad = (got != want) & ~(np.isnan(got) & np.isnan(want))
        assert not bad.any(), "%s: %d of %d entries differ; first system %d" % (name, bad.sum(), bad.size, int(np.argwhere(bad)[0][0]))
