"""GPU parity of the compiled sweep driver (csrc/tci2_driver.cpp behind `engine="native"` / native.NativeTCI2) on the real
kernels: against the oracle (pivot sets, ranks, errors identical; tensors to 1e-10 like the Python-driven path) and
against the Python-driven path itself (same kernels in the same order: everything identical).  Its bookkeeping alone is
also checked on the CPU in tests/test_native_driver_mock.py."""
import ctypes as C

import numpy as np
import pytest

from oracle import oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def tci():
    import tci_b200
    return tci_b200


def _same_sets(t, o):
    for b in range(1, len(t) + 1):
        np.testing.assert_array_equal(t.Iset[b - 1], o.Iset(b), err_msg=f"Iset[{b}]")
        np.testing.assert_array_equal(t.Jset[b - 1], o.Jset(b), err_msg=f"Jset[{b}]")


def _close(a, r, tol):
    assert a.shape == r.shape, (a.shape, r.shape)
    if r.size:
        assert np.max(np.abs(a - r)) <= tol * max(np.max(np.abs(r)), 1e-300), np.max(np.abs(a - r))


@pytest.mark.parametrize("n,d,tol", [(5, 10, 1e-8), (8, 10, 1e-8), (6, 4, 1e-10)])
def test_native_matches_oracle(tci, n, d, tol):
    """C1 family through ONE C call (tci_tci2_optimize): the parity bar of tests/test_gpu_tci2.py."""
    f, g = tci.LorentzFunction([d] * n), O.Func(O.F_LORENTZ, [1.0])
    t, ranks, errors = tci.crossinterpolate2(f, [d] * n, tolerance=tol, nsearchglobalpivot=0, maxiter=200, engine="native")
    o, oranks, oerrors = O.crossinterpolate2(g, [d] * n, tolerance=tol, maxiter=200)
    assert list(ranks) == list(oranks)
    np.testing.assert_array_equal(errors, oerrors)
    _same_sets(t, o)
    for a, r in zip(t.sitetensors, o.sitetensors()):
        _close(a, r, 1e-10)
    assert t.maxsamplevalue == o.maxsamplevalue
    np.testing.assert_array_equal(t.bonderrors, o.bonderrors)
    np.testing.assert_array_equal(t.pivoterrors, o.pivoterrors)
    pts = np.random.default_rng(0).integers(1, d + 1, size=(300, n))
    ref = np.array([g(p) for p in pts])
    assert np.max(np.abs(t.evaluate_batch(pts) - ref)) < 100 * tol * t.maxsamplevalue
    assert t.sum() == pytest.approx(O.tt_sum(o.sitetensors()), rel=1e-10)
    assert t.stats["nevals"] == o.stats()["nevals"]


@pytest.mark.parametrize("kw", [dict(sweepstrategy="forward", maxiter=6), dict(maxbonddim=5), dict(strictlynested=True, maxiter=5),
                                dict(normalizeerror=False)], ids=["forward", "maxbonddim", "nested", "unnormalized"])
def test_native_identical_to_python_engine(tci, kw):
    n = 6
    f = tci.LorentzFunction([8] * n)
    a, ra, ea = tci.crossinterpolate2(f, [8] * n, tolerance=1e-7, nsearchglobalpivot=0, engine="native", **kw)
    b, rb, eb = tci.crossinterpolate2(f, [8] * n, tolerance=1e-7, nsearchglobalpivot=0, engine="python", **kw)
    np.testing.assert_array_equal(ra, rb)
    np.testing.assert_array_equal(ea, eb)
    for s in range(n):
        np.testing.assert_array_equal(a.Iset[s], b.Iset[s])
        np.testing.assert_array_equal(a.Jset[s], b.Jset[s])
    for x, y in zip(a.sitetensors, b.sitetensors):
        _close(x, y, 1e-13)
    np.testing.assert_array_equal(a.pivoterrors, b.pivoterrors)
    assert a.maxsamplevalue == b.maxsamplevalue


def test_native_hash_function_hbm_regime(tci):
    """Full-rank function up to maxbonddim 64 (Pi 1280 x 1280: the persistent-kernel regime of the rrLU), pivots vs oracle."""
    ld = [20] * 6
    f, g = tci.HashFunction(ld, 1.0, 1.0), O.Func(O.F_HASH, [1.0, 1.0])
    kw = dict(tolerance=1e-10, maxbonddim=64, maxiter=1)
    t, ranks, errors = tci.crossinterpolate2(f, ld, nsearchglobalpivot=0, engine="native", **kw)
    o, oranks, oerrors = O.crossinterpolate2(g, ld, **kw)
    assert list(ranks) == list(oranks)
    _same_sets(t, o)


def test_native_stepwise_reference_case(tci):
    """test/test_tensorci2.jl:247-273 call by call on a NativeTCI2: updatepivots! per bond (linkdims 2), one global pivot +
    makecanonical (linkdims 3), then sweeps; state compared with the oracle after every stage."""
    from tci_b200.native import NativeTCI2
    ld = [10] * 5
    f, g = tci.LorentzFunction(ld), O.Func(O.F_LORENTZ, [1.0])
    h = f.handle
    nat = NativeTCI2(tci.lib(), h.h, ld)
    nat.addglobalpivots([[1] * 5])
    nat.init_maxsample(f._f, [[1] * 5])
    o = O.TCI2(g, ld, [[1] * 5])

    def same(cores, tol=1e-10):
        for b in range(1, 6):
            np.testing.assert_array_equal(nat.Iset(b), o.Iset(b))
            np.testing.assert_array_equal(nat.Jset(b), o.Jset(b))
        be, pe, mx = nat.errors()
        np.testing.assert_array_equal(be, o.bonderrors)
        np.testing.assert_array_equal(pe, o.pivoterrors)
        assert mx == o.maxsamplevalue
        if cores:
            import torch
            from tci_b200.parallel import _RawCuda
            for b in range(1, 6):
                ptr, shape = nat.core(b)
                got = torch.as_tensor(_RawCuda(ptr, int(np.prod(shape))), device=f"cuda:{h.device}").cpu().numpy().reshape(shape, order="F")
                _close(got, o.sitetensor(b), tol)

    for b in range(1, 5):
        nat.updatepivots(f._f, b, True, reltol=1e-8, maxbonddim=2)
        o.updatepivots(b, True, reltol=1e-8, maxbonddim=2)
    assert nat.linkdims() == [2] * 4
    same(False)
    nat.addglobalpivots([[2, 9, 10, 5, 7]])
    nat.makecanonical(f._f, reltol=1e-8)
    o.addglobalpivots1sitesweep([[2, 9, 10, 5, 7]], reltol=1e-8)
    assert nat.linkdims() == [3] * 4
    same(True)
    nat.sweep2site(f._f, 3, abstol=1e-9, fillsitetensors=False)
    o.sweep2site(3, abstol=1e-9, fillsitetensors=False)
    same(False)
    assert not nat.sitetensors_available()
    nat.sweep1site(f._f, "backward", abstol=1e-9)
    o.sweep1site("backward", abstol=1e-9)
    same(True)
    nat.fillsitetensors(f._f)
    o.fillsitetensors()
    # T = Pi1 * inv(P) (tensorci2.jl:626) with pivots kept down to 1e-9: cond(P) ~ 1e9, and the device solves through a
    # full-pivot LU where the checker restates LAPACK's partial-pivot one — agreement to cond * eps, not to 1e-10
    same(True, tol=1e-6)
    assert nat.stats()["nbondupdates"] == 4 + 3 * 4
    nat.close()


def test_native_global_pivot_search(tci):
    """Default finder inside the driver (one batch of f and one of the tensor train per iteration): a feature that needs
    v2 = v5 = 4 together is invisible to two-site sweeps started at [1, 1, ...] and found by the line search."""
    ld = [4] * 6
    table = np.full(ld, 0.5)
    table[:, 3, :, :, 3, :] = 5.0
    f = tci.DenseFunction(table)
    t0, _, _ = tci.crossinterpolate2(f, ld, tolerance=1e-10, nsearchglobalpivot=0, engine="native")
    assert t0.rank() == 1
    t, ranks, errors = tci.crossinterpolate2(f, ld, tolerance=1e-10, nsearchglobalpivot=40, maxnglobalpivot=5, seed=7, engine="native")
    assert t.rank() == 2
    pts = np.random.default_rng(2).integers(1, 5, size=(200, 6))
    ref = np.array([table[tuple(p - 1)] for p in pts])
    assert np.max(np.abs(t.evaluate_batch(pts) - ref)) < 1e-9


def test_native_quantics_c4_small(tci):
    """R=16 quantics through the compiled driver; the oracle consumes the device-evaluated values (see test_gpu_baseline_configs)."""
    R = 16
    f, g = tci.QuanticsExpPlusSin([2] * R), O.Func(O.F_QEXPPLUSSIN, [1.0, 1e-3, 1000.0])
    fed = O.Func(O.F_BATCHCB, callback=lambda left, right, M: f(left, right, M))
    t, ranks, errors = tci.crossinterpolate2(f, [2] * R, tolerance=1e-10, nsearchglobalpivot=0, maxbonddim=64, engine="native")
    o, oranks, oerrors = O.crossinterpolate2(fed, [2] * R, tolerance=1e-10, maxbonddim=64)
    assert list(ranks) == list(oranks)
    np.testing.assert_array_equal(errors, oerrors)
    for b in range(1, R + 1):
        np.testing.assert_array_equal(t.Iset[b - 1], o.Iset(b))
        np.testing.assert_array_equal(t.Jset[b - 1], o.Jset(b))
    pts = np.random.default_rng(1).integers(1, 3, size=(500, R))
    ref = np.array([g(p) for p in pts])
    assert np.max(np.abs(t.evaluate_batch(pts) - ref)) < 1e-7


def test_native_errors(tci):
    f = tci.LorentzFunction([3, 3, 3])
    with pytest.raises(tci.TCIError, match="Unknown pivot search strategy"):
        tci.crossinterpolate2(f, [3, 3, 3], engine="native", pivotsearch="castle")
    with pytest.raises(tci.TCIError, match="device functor"):
        tci.crossinterpolate2(lambda v: 1.0, [3, 3, 3], engine="native")
    with pytest.raises(tci.TCIError, match="convergence criterion is not reachable"):
        tci.crossinterpolate2(f, [3, 3, 3], engine="native", tolerance=0.0)
    with pytest.raises(tci.TCIError, match="maxsamplevalue is zero!"):
        tci.crossinterpolate2(tci.DenseFunction(np.zeros((3, 3))), [3, 3], engine="native")
    with pytest.raises(tci.TCIError, match="Unknown engine"):
        tci.crossinterpolate2(f, [3, 3, 3], engine="fortran")
    assert C.sizeof(C.c_void_p) == 8


def test_device_max_accumulator(tci):
    """tci_batcheval_acc_f64 / tci_maxacc_set / tci_maxacc_get: the maximum of several batches merged on the device equals
    what tci_batcheval_f64 returns batch by batch (updatemaxsample!, tensorci2.jl:636-638), NaN included."""
    import torch
    from tci_b200._lib import p_i32
    L, ld = tci.lib(), np.asarray([7] * 4, dtype=np.int32)
    f = tci.LorentzFunction(list(ld), 3.0)
    h = f.handle
    dev = f"cuda:{h.device}"
    acc = torch.zeros(2, dtype=torch.float64, device=dev)
    h.check(L.tci_maxacc_set(h.h, C.c_void_p(acc.data_ptr()), C.c_double(1e-3)))
    ref = 1e-3
    rng = np.random.default_rng(4)
    for _ in range(3):
        left = np.ascontiguousarray(rng.integers(2, 8, size=(5, 2)), dtype=np.int32)
        right = np.ascontiguousarray(rng.integers(2, 8, size=(6, 2)), dtype=np.int32)
        a = torch.empty(30, dtype=torch.float64, device=dev)
        b = torch.empty(30, dtype=torch.float64, device=dev)
        h.check(L.tci_batcheval_acc_f64(h.h, f._f, p_i32(ld), 4, p_i32(left), C.c_int64(5), 2, p_i32(right), C.c_int64(6), 2,
                                        C.c_void_p(a.data_ptr()), C.c_int64(0), C.c_int64(0), C.c_void_p(acc.data_ptr())))
        mx = C.c_double(ref)
        h.check(L.tci_batcheval_f64(h.h, f._f, p_i32(ld), 4, p_i32(left), C.c_int64(5), 2, p_i32(right), C.c_int64(6), 2,
                                    C.c_void_p(b.data_ptr()), C.c_int64(0), C.c_int64(0), C.byref(mx)))
        ref = mx.value
        assert torch.equal(a, b)
    got = C.c_double()
    h.check(L.tci_maxacc_get(h.h, C.c_void_p(acc.data_ptr()), C.byref(got)))
    assert got.value == ref and ref > 1e-3
    h.check(L.tci_maxacc_set(h.h, C.c_void_p(acc.data_ptr()), C.c_double(float("nan"))))
    h.check(L.tci_maxacc_get(h.h, C.c_void_p(acc.data_ptr()), C.byref(got)))
    assert np.isnan(got.value)
    table = np.ones((3, 3)); table[1, 2] = np.nan
    g = tci.DenseFunction(table)
    h.check(L.tci_maxacc_set(h.h, C.c_void_p(acc.data_ptr()), C.c_double(0.5)))
    out = torch.empty(9, dtype=torch.float64, device=dev)
    l3 = np.asarray([[1], [2], [3]], dtype=np.int32)
    ld2 = np.asarray([3, 3], dtype=np.int32)
    h.check(L.tci_batcheval_acc_f64(h.h, g._f, p_i32(ld2), 2, p_i32(l3), C.c_int64(3), 1, p_i32(l3), C.c_int64(3), 1,
                                    C.c_void_p(out.data_ptr()), C.c_int64(0), C.c_int64(0), C.c_void_p(acc.data_ptr())))
    h.check(L.tci_maxacc_get(h.h, C.c_void_p(acc.data_ptr()), C.byref(got)))
    assert np.isnan(got.value)


def test_optfirstpivot_native(tci):                        # util.jl:260-291; device functors take the one-call path
    rng = np.random.default_rng(8)
    ld = [5, 3, 7, 2, 6, 4]
    f, g = tci.HashFunction(ld, 1.0, 0.3), O.Func(O.F_HASH, [1.0, 0.3])
    for _ in range(4):
        first = [int(rng.integers(1, d + 1)) for d in ld]
        assert tci.optfirstpivot(f, ld, first) == list(O.optfirstpivot(g, ld, first))
    assert tci.optfirstpivot(tci.SumFunction([3, 4, 5, 6]), [3, 4, 5, 6], [1, 2, 1, 2]) == [3, 4, 5, 6]


def test_engine_auto_selection(tci):
    """engine="auto" (the default): the compiled driver for device functors with the stock keywords, the Python host otherwise."""
    from tci_b200.tensorci2 import _native_eligible
    f = tci.LorentzFunction([4] * 4)
    assert _native_eligible(f, dict(tolerance=1e-8, nsearchglobalpivot=0))
    assert _native_eligible(f, dict(tolerance=1e-8, maxbonddim=7, handle=None))
    assert _native_eligible(f, dict(pivotsearch="rook"))
    assert not _native_eligible(f, dict(rng=np.random.default_rng(0)))
    assert not _native_eligible(f, dict(verbosity=1))
    assert not _native_eligible(f, dict(bondorder="oddeven"))
    assert not _native_eligible(lambda v: 1.0, {})
    assert not _native_eligible(tci.HostFunctionAdapter(lambda v: 1.0, [4] * 4), {})
    assert not _native_eligible(tci.CachedFunction(lambda v: 1.0, [4] * 4), {})
    t, _, _ = tci.crossinterpolate2(f, [4] * 4, tolerance=1e-8, nsearchglobalpivot=0)
    assert "nbondupdates" in t.stats                        # ran in the compiled driver
    t, _, _ = tci.crossinterpolate2(f, [4] * 4, tolerance=1e-8, nsearchglobalpivot=0, engine="python")
    assert "nbondupdates" not in t.stats


@pytest.mark.parametrize("n,d,kw,seed", [(6, 8, dict(tolerance=1e-8), 1), (6, 6, dict(tolerance=1e-12, maxbonddim=5), 7),
                                         (5, 5, dict(tolerance=1e-9, strictlynested=True), 11)])
def test_native_rook_search(tci, n, d, kw, seed):
    """pivotsearch=:rook inside the compiled driver (arrlu over strips, matrixlu.jl:500-569): same seeded stream as the checker
    and as the Python host, so all three sample the same strips — index sets, errors and evaluation counts identical."""
    f, g = tci.LorentzFunction([d] * n), O.Func(O.F_LORENTZ, [1.0])
    t, ranks, errors = tci.crossinterpolate2(f, [d] * n, nsearchglobalpivot=0, pivotsearch="rook", seed=seed, engine="native", **kw)
    o, oranks, oerrors = O.crossinterpolate2(g, [d] * n, pivotsearch="rook", seed=seed, **kw)
    assert list(ranks) == list(oranks)
    np.testing.assert_array_equal(errors, oerrors)
    _same_sets(t, o)
    assert t.stats["nevals"] == o.stats()["nevals"]
    for a, r in zip(t.sitetensors, o.sitetensors()):
        _close(a, r, 1e-10)
    p, pranks, perrors = tci.crossinterpolate2(f, [d] * n, nsearchglobalpivot=0, pivotsearch="rook", seed=seed, engine="python", **kw)
    np.testing.assert_array_equal(ranks, pranks)
    for s in range(n):
        np.testing.assert_array_equal(t.Iset[s], p.Iset[s])
        np.testing.assert_array_equal(t.Jset[s], p.Jset[s])
