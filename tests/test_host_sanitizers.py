"""The host-only C++ of the library (csrc/host: rx_geometry.cpp, rx_partition.cpp, rx_su2_mesh.cpp) under AddressSanitizer and
UndefinedBehaviorSanitizer: built from the sources with g++ -fsanitize=address,undefined next to tests/host_sanitizer_driver.cpp and run on
jittered triangulations, on well-formed .su2 files (the shipped-layout jet mesh, a mixed triangle / quadrilateral mesh, a file without a final
newline) and on malformed ones (truncated, empty, missing).  Any sanitizer report fails the run."""
import importlib
import os
import shutil
import subprocess

import pytest

from rxcase import ROOT, load_package
from test_geometry import _jet_su2, _mixed_su2

HOST = os.path.join(ROOT, "development-of-a-turbulent-numerical-solver-for-reactive-flows-in-su2_b200", "csrc", "host")


def test_host_code_is_clean_under_address_and_ub_sanitizers(tmp_path):
    if shutil.which("g++") is None:
        pytest.skip("no g++")
    exe = str(tmp_path / "host_san")
    srcs = [os.path.join(ROOT, "tests", "host_sanitizer_driver.cpp")] + [os.path.join(HOST, f) for f in ("rx_geometry.cpp", "rx_partition.cpp", "rx_su2_mesh.cpp")]
    cc = subprocess.run(["g++", "-O1", "-g", "-std=c++17", "-fsanitize=address,undefined", "-fno-sanitize-recover=undefined", "-I", os.path.join(ROOT, "include")] + srcs +
                        ["-ldl", "-o", exe], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if cc.returncode != 0 and "sanitize" in cc.stdout and ("cannot find" in cc.stdout or "unrecognized" in cc.stdout):
        pytest.skip("this g++ has no sanitizer runtime")
    assert cc.returncode == 0, cc.stdout[-3000:]
    pk = load_package()
    geo = importlib.import_module(pk.__name__ + ".geometry")
    jet, mixed = str(tmp_path / "jet.su2"), str(tmp_path / "mixed.su2")
    geo.write_su2(jet, _jet_su2()); geo.write_su2(mixed, _mixed_su2(geo))
    txt = open(mixed).read()
    files = {"cut.su2": txt[:len(txt) // 3], "nonl.su2": txt.rstrip("\n"), "empty.su2": ""}
    for name, body in files.items():
        open(str(tmp_path / name), "w").write(body)
    args = [jet, mixed] + [str(tmp_path / n) for n in files] + [str(tmp_path / "absent.su2")]
    env = dict(os.environ, RX_PARTITIONER="native", ASAN_OPTIONS="detect_leaks=1:abort_on_error=0")
    run = subprocess.run([exe] + args, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env, timeout=600)
    out = run.stdout
    assert run.returncode == 0 and "DONE 0" in out, out[-4000:]
    assert "ERROR: AddressSanitizer" not in out and "runtime error" not in out and "LeakSanitizer" not in out, out[-4000:]
    lines = [l for l in out.splitlines() if l.startswith("su2 ")]
    ok = [l for l in lines if " rc 0 " in l]
    assert len(lines) == 12 and len(ok) == 6, out        # jet, mixed and the file without a final newline read; the other three are refused
    assert "ends inside" in out and "no points or no elements" in out and "cannot open" in out
