"""What the C++ subclasses accept and what they refuse, checked WITHOUT a GPU: oracle/_ref/ref_harness_gpu — the reference's
driver with CReactiveNSSolverGPU as flow solver — is started with no CUDA device visible.  Every guard of the constructor runs before rx_create,
so a cfg the subclass accepts ends with rx_create's "no CUDA device" message, and a cfg it refuses ends with the refusal's own message."""
import os
import re
import subprocess

import pytest

from rxcase import ROOT

HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_harness_gpu")
CASES = os.path.join(ROOT, "oracle", "_ref", "cases")


def _end_message(case, cfg="case.cfg", extra_env=None):
    d = os.path.join(CASES, case)
    env = dict(os.environ, SU2_RX_BACKEND="cuda", CUDA_VISIBLE_DEVICES="", **(extra_env or {}))
    import tempfile
    log = subprocess.run([HARNESS, cfg, "1", os.path.join(tempfile.mkdtemp(prefix="rx_guard_"), "dump"), "99", "1"], cwd=d, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                         timeout=600)
    m = re.findall(r"what\(\):\s*(.*)", log.stdout)
    return m[-1] if m else log.stdout[-500:]


def _need():
    if not (os.path.exists(HARNESS) and os.path.isdir(CASES)):
        pytest.skip("oracle/_ref is not in this snapshot (built by oracle/build_ref.sh where the reference tree is present)")


def test_every_case_cfg_passes_the_guards():
    _need()
    names = sorted(n for n in os.listdir(CASES) if os.path.exists(os.path.join(CASES, n, "case.cfg")))
    assert len(names) >= 20
    for n in names:
        msg = _end_message(n)
        assert "rx_create: no CUDA device" in msg, "%s: %s" % (n, msg)


def _derived(case, tmp_path, repl):
    src = open(os.path.join(CASES, case, "case.cfg")).read()
    for key, val in repl.items():
        src, k = re.subn(r"(?m)^%s\s*=.*$" % re.escape(key), "%s= %s" % (key, val), src)
        if k == 0:
            src += "\n%s= %s\n" % (key, val)
    p = str(tmp_path / ("%s_%s.cfg" % (case, "_".join(repl))))
    open(p, "w").write(src)
    return p


@pytest.mark.parametrize("case,repl,expect", [
    ("channel_lam_walls", {"TIME_DISCRE_FLOW": "EULER_IMPLICIT"}, "EULER_IMPLICIT is not on the CUDA path"),
    ("channel_lam_walls", {"VISCOUS_LIMITER_FLOW": "YES"}, "VISCOUS_LIMITER_FLOW needs SPATIAL_ORDER_FLOW= 2ND_ORDER_LIMITER"),
    ("channel_lam_walls", {"REF_DIMENSIONALIZATION": "REFERENCE"}, "REF_DIMENSIONALIZATION other than DIMENSIONAL"),
    ("channel_lam_muscl", {"SLOPE_LIMITER_FLOW": "SHARP_EDGES"}, "only the VENKATAKRISHNAN and BARTH_JESPERSEN limiters"),
])
def test_unsupported_switches_are_refused_with_their_own_message(case, repl, expect, tmp_path):
    _need()
    msg = _end_message(case, _derived(case, tmp_path, repl))
    assert expect in msg, msg
