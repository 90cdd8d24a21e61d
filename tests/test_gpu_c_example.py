"""The pure-C boundary on a device: examples/c1_lorentz.c (crossinterpolate2 + sum + evaluate for BASELINE config C1, written
against include/tci_b200.h alone, compiled with gcc -std=c99) is built, RUN, and its printed rank / error / sum / value are
compared with the oracle's run of the same configuration — with both driver engines."""
import os
import re
import subprocess

import numpy as np
import pytest

from oracle import oracle as O

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("engine", ["device", "percall"])
def test_c_example_runs_and_matches_the_oracle(tmp_path, engine):
    import tci_b200
    exe = str(tmp_path / "c1_lorentz")
    libdir = os.path.dirname(tci_b200.SO_PATH)
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I" + os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "examples", "c1_lorentz.c"), "-L" + libdir, "-ltci_b200", "-Wl,-rpath," + libdir, "-o", exe])
    r = subprocess.run([exe, engine], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    m = re.search(r"iterations (\d+), rank (\d+), error ([0-9.e+-]+)", r.stdout)
    s = re.search(r"sum over 10\^8 grid points = ([0-9.e+-]+)", r.stdout)
    v = re.search(r"tt\(1\.\.8\) = ([0-9.e+-]+)", r.stdout)
    assert m and s and v, r.stdout
    ld = [10] * 8
    g = O.Func(O.F_LORENTZ, [1.0])
    o, oranks, oerrors = O.crossinterpolate2(g, ld, tolerance=1e-8)
    assert int(m.group(1)) == len(oranks) and int(m.group(2)) == int(oranks[-1])
    assert float(m.group(3)) == pytest.approx(float(oerrors[-1]), rel=1e-3)          # printed with 4 digits
    assert float(s.group(1)) == pytest.approx(O.tt_sum(o.sitetensors()), rel=1e-10)
    assert float(v.group(1)) == pytest.approx(O.tt_evaluate(o.sitetensors(), [[1, 2, 3, 4, 5, 6, 7, 8]])[0], rel=1e-9)
    assert float(v.group(1)) == pytest.approx(1.0 / 205.0, abs=1e-8)
