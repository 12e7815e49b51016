#!/usr/bin/env python
"""TEST INFRASTRUCTURE — golden vectors of the stand-alone source term from the reference's own ReactingModelLibrary
(oracle/_ref/ref_source_harness, built by oracle/build_ref.sh), on the seeded states of the source microbenchmark
(package synthetic_points: T 300-2500 K, 30 % pure oxidiser, 10 % pure fuel) — SURVEY.md 7.1 step 0(i).

Writes tests/golden/source_points.npz: inputs (T, rho, Y, omegaTurb, vol) and res_laminar / res_pasr [n][Ns].
Runs only in the build container (needs the case data copied from /root/reference by make_golden.py prepare)."""
import importlib
import os
import struct
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
pk = importlib.import_module("development-of-a-turbulent-numerical-solver-for-reactive-flows-in-su2_b200")
HARNESS = os.path.join(HERE, "_ref", "ref_source_harness")
CASE = os.path.join(HERE, "_ref", "cases", "channel_lam_rk")   # any case directory: Mixture/ Chemistry/ Thermo/ Transp/ + test_chem_first.txt
N, SEED, CMU, LB = 4096, 2024, 0.09, 0.2


def run(rans, T, rho, Y, omT, vol):
    with tempfile.TemporaryDirectory() as d:
        fi, fo = os.path.join(d, "in.bin"), os.path.join(d, "out.bin")
        with open(fi, "wb") as f:
            f.write(struct.pack("<qqdd", len(T), int(rans), CMU, LB))
            for a in (T, rho, Y, omT, vol):
                f.write(np.ascontiguousarray(a, dtype=np.float64).tobytes())
        out = subprocess.run([HARNESS, "test_chem_first.txt", fi, fo], cwd=CASE, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if out.returncode != 0:
            raise RuntimeError(out.stdout[-2000:])
        print(out.stdout.strip().splitlines()[-1])
        return np.fromfile(fo, dtype=np.float64).reshape(len(T), -1)


if __name__ == "__main__":
    chem = pk.load_mechanism()
    T, rho, Y, omT, vol = pk.synthetic_points(chem, N, seed=SEED)
    # a slice of hot, fully mixed states on top (the regime the PaSR limiter and both reactions are active in)
    T[:256] = np.linspace(1800.0, 2500.0, 256)
    arrs = dict(T=T, rho=rho, Y=Y, omegaTurb=omT, vol=vol, cMu=np.float64(CMU), pasrLb=np.float64(LB),
                res_laminar=run(False, T, rho, Y, omT, vol), res_pasr=run(True, T, rho, Y, omT, vol))
    path = os.path.join(ROOT, "tests", "golden", "source_points.npz")
    np.savez_compressed(path, **arrs)
    print("wrote %s (%.2f MB); |res| max laminar %.3e pasr %.3e" % (path, os.path.getsize(path) / 1e6, np.abs(arrs["res_laminar"]).max(), np.abs(arrs["res_pasr"]).max()))
