"""CPU-side checks of the drop-in boundary: libtci_b200.so loads without a GPU and exports every function that
include/tci_b200.h declares; the product package never imports the oracle; compute calls fail loudly without a device."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "tci_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(tci_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    import tci_b200
    lib = ctypes.CDLL(tci_b200.SO_PATH)
    names = _declared()
    assert len(names) >= 20
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing


def test_no_cpu_fallback_without_device():
    import torch
    import tci_b200
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(tci_b200.TCIError, match="no CPU fallback"):
        tci_b200.Handle(0)
    with pytest.raises(tci_b200.TCIError):
        tci_b200.rrlu([[1.0, 2.0], [3.0, 4.0]])


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "tensorcrossinterpolation.jl_b200")
    for dp, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".h", ".cpp")):
                txt = open(os.path.join(dp, fn)).read()
                assert "oracle" not in txt.replace("the oracle", "").lower() or fn == "never", (fn, "mentions oracle")


def test_host_index_helpers():
    import numpy as np
    import tci_b200
    from tci_b200.tensorci2 import union_rows, pushunique
    I = np.array([[1, 2], [3, 4]], dtype=np.int32)
    k = tci_b200.kronecker_left(I, 3)                    # tensorci2.jl:512-517: Iset fastest
    assert k.tolist() == [[1, 2, 1], [3, 4, 1], [1, 2, 2], [3, 4, 2], [1, 2, 3], [3, 4, 3]]
    k = tci_b200.kronecker_right(2, I)                   # tensorci2.jl:524-529: local index fastest
    assert k.tolist() == [[1, 1, 2], [2, 1, 2], [1, 3, 4], [2, 3, 4]]
    u = union_rows(I, np.array([[3, 4], [5, 6], [1, 2], [5, 6]], dtype=np.int32))
    assert u.tolist() == [[1, 2], [3, 4], [5, 6]]
    assert pushunique(I, [1, 2]).shape == (2, 2) and pushunique(I, [9, 9]).shape == (3, 2)
    assert tci_b200.convergencecriterion([1, 2], [1e-2, 1e-5], [0, 0], 1e-4, 4, 3) is False          # test_tensorci2.jl:504-554
    assert tci_b200.convergencecriterion([1, 2, 2, 2], [1e-2, 1e-5, 1e-5, 1e-5], [0, 0, 0, 0], 1e-4, 4, 3) is True
    assert tci_b200.convergencecriterion([1, 2, 2, 2], [1e-2, 1e-5, 1e-5, 1e-5], [0, 1, 0, 0], 1e-4, 4, 3) is False
    assert tci_b200.convergencecriterion([4, 4, 4], [1.0, 1.0, 1.0], [0, 0, 0], 1e-4, 4, 3) is True
    assert tci_b200.forwardsweep("forward", 2) and tci_b200.forwardsweep("backandforth", 1)          # sweepstrategies.jl:41-50
    assert not tci_b200.forwardsweep("backandforth", 2)


def test_header_is_plain_c_and_example_fails_loudly_without_a_device(tmp_path):
    """include/tci_b200.h compiles as C99 (-pedantic) — the boundary is a C ABI, not a C++ one — and examples/c1_lorentz.c (the
    whole of crossinterpolate2 + sum + evaluate from plain C) links against the library; without a CUDA device it stops at
    tci_create with the "no CPU fallback" message."""
    import subprocess
    import torch
    import tci_b200
    exe = str(tmp_path / "c1_lorentz")
    libdir = os.path.dirname(tci_b200.SO_PATH)
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I" + os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "examples", "c1_lorentz.c"), "-L" + libdir, "-ltci_b200", "-Wl,-rpath," + libdir, "-o", exe])
    if torch.cuda.is_available():
        pytest.skip("a GPU is present: tests/test_gpu_c_example.py runs the example")
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 2 and "no CPU fallback" in r.stderr
