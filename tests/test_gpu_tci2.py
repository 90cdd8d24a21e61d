"""GPU parity of the TensorCI2 sweeps (host logic in Python + K1/K2/K3/K4 on the device) against the oracle.
Bar (north star): pivot index sets and ranks bit-exact, tensor entries within 1e-12 relative, errors equal."""
import itertools

import numpy as np
import pytest

from oracle import oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def tci():
    import tci_b200
    return tci_b200


def _same_sets(t, o):
    n = len(t)
    for b in range(1, n + 1):
        np.testing.assert_array_equal(t.Iset[b - 1], o.Iset(b), err_msg=f"Iset[{b}]")
        np.testing.assert_array_equal(t.Jset[b - 1], o.Jset(b), err_msg=f"Jset[{b}]")


def _same_tensors(t, o, tol=1e-12):
    for b, (a, r) in enumerate(zip(t.sitetensors, o.sitetensors())):
        assert a.shape == r.shape, (b, a.shape, r.shape)
        if r.size:
            assert np.max(np.abs(a - r)) <= tol * max(np.max(np.abs(r)), 1e-300), (b, np.max(np.abs(a - r)))


def test_kronecker_order(tci):                          # test/test_tensorci2.jl:9-25
    ms = np.array([list(range(1, 6))] * 5, dtype=np.int32)
    c = tci.kronecker_left(ms, 4)
    assert c.shape == (20, 6) and np.all(c[:, :5] == np.arange(1, 6)) and list(c[:, 5]) == [1] * 5 + [2] * 5 + [3] * 5 + [4] * 5
    d = tci.kronecker_right(4, ms)
    assert d.shape == (20, 6) and np.all(d[:, 1:] == np.arange(1, 6)) and list(d[:, 0]) == [1, 2, 3, 4] * 5


def test_pivoterrors_diagonal(tci):                     # test/test_tensorci2.jl:27-39 (exact)
    diags = [1.0, 1e-5, 0.0]
    f = tci.DenseFunction(np.diag(diags))
    t, ranks, errors = tci.crossinterpolate2(f, [3, 3], [[1, 1]], tolerance=1e-8, nsearchglobalpivot=0)
    assert list(t.pivoterrors) == diags
    g = lambda x: diags[x[0] - 1] if x[0] == x[1] else 0.0      # plain Python callable through the adapter
    t2, _, _ = tci.crossinterpolate2(g, [3, 3], [[1, 1]], tolerance=1e-8, nsearchglobalpivot=0)
    assert list(t2.pivoterrors) == diags


def test_trivial_and_errors(tci):                       # test/test_tensorci2.jl "trivial MPS", "TCI2 errors"
    f = tci.SumFunction([2] * 5, 0.5)
    t = tci.TensorCI2([2] * 5)
    assert len(t) == 5 and t.linkdims() == [0] * 4
    t = tci.TensorCI2([2] * 5, f, [[1] * 5])
    assert t.rank() == 1 and t.linkdims() == [1] * 4
    lor = tci.LorentzFunction([2] * 5)
    with pytest.raises(tci.TCIError):
        tci.crossinterpolate2(lor, [2] * 5, tolerance=1e-9, pivottolerance=1e-2)
    with pytest.raises(tci.TCIError):
        tci.crossinterpolate2(lor, [2] * 5, tolerance=0.0)
    with pytest.raises(tci.TCIError, match="not batch evaluatable"):
        tci.crossinterpolate2(lambda x: 1.0, [2] * 10, checkbatchevaluatable=True)


def test_lorentz_5site_steps(tci):                      # test/test_tensorci2.jl:247-273 + oracle parity at each step
    n = 5
    f, g = tci.LorentzFunction([10] * n), O.Func(O.F_LORENTZ, [1.0])
    t, o = tci.TensorCI2([10] * n, f), O.TCI2(g, [10] * n)
    assert t.linkdims() == [1] * (n - 1)
    for p in range(1, n):
        t.updatepivots(p, f, True, reltol=1e-8, maxbonddim=2)
        o.updatepivots(p, True, reltol=1e-8, maxbonddim=2)
    assert t.linkdims() == [2] * (n - 1) and t.rank() == 2
    _same_sets(t, o)
    t.addglobalpivots1sitesweep(f, [[2, 9, 10, 5, 7]], reltol=1e-12)
    o.addglobalpivots1sitesweep([[2, 9, 10, 5, 7]], reltol=1e-12)
    assert t.linkdims() == [3] * (n - 1) and t.rank() == 3
    _same_sets(t, o)
    _same_tensors(t, o)
    assert t.maxsamplevalue == o.maxsamplevalue
    np.testing.assert_array_equal(t.pivoterrors, o.pivoterrors)


@pytest.mark.parametrize("n,d,tol", [(5, 10, 1e-8), (5, 10, 1e-12), (8, 10, 1e-8), (6, 4, 1e-10)])
def test_crossinterpolate2_matches_oracle(tci, n, d, tol):
    """C1 family: README Lorentzian; index sets, ranks, errors identical; tensors to 1e-12; values within tolerance."""
    f, g = tci.LorentzFunction([d] * n), O.Func(O.F_LORENTZ, [1.0])
    t, ranks, errors = tci.crossinterpolate2(f, [d] * n, tolerance=tol, nsearchglobalpivot=0, maxiter=200)
    o, oranks, oerrors = O.crossinterpolate2(g, [d] * n, tolerance=tol, maxiter=200)
    assert list(ranks) == list(oranks)
    np.testing.assert_array_equal(errors, oerrors)
    _same_sets(t, o)
    _same_tensors(t, o, 1e-10)
    assert t.maxsamplevalue == o.maxsamplevalue
    assert t.pivoterror() <= max(2 * tol, 2e-12)
    rng = np.random.default_rng(0)
    pts = rng.integers(1, d + 1, size=(300, n))
    vals = t.evaluate_batch(pts)
    ref = np.array([g(p) for p in pts])
    assert np.max(np.abs(vals - ref)) < 100 * tol * t.maxsamplevalue
    assert t.sum() == pytest.approx(O.tt_sum(o.sitetensors()), rel=1e-10)


def test_forward_strategy_and_maxbonddim(tci):
    n = 6
    f, g = tci.LorentzFunction([8] * n), O.Func(O.F_LORENTZ, [1.0])
    for kw in (dict(sweepstrategy="forward", maxiter=6), dict(maxbonddim=5), dict(strictlynested=True, maxiter=5),
               dict(normalizeerror=False)):
        t, ranks, errors = tci.crossinterpolate2(f, [8] * n, tolerance=1e-7, nsearchglobalpivot=0, **kw)
        o, oranks, oerrors = O.crossinterpolate2(g, [8] * n, tolerance=1e-7, **kw)
        assert list(ranks) == list(oranks), kw
        np.testing.assert_array_equal(errors, oerrors)
        _same_sets(t, o)


def test_quantics_exp(tci):                             # test/test_tensorci2.jl:55-102
    R, abstol = 8, 1e-4
    f, g = tci.QuanticsExp2([2] * R, 1.0, 0.0, 0.0), O.Func(O.F_QEXP2, [1.0, 0.0, 0.0])
    piv = [[1] * R, [1] + [2] * (R - 1)]
    for strictlynested in (False, True):
        t, _, _ = tci.crossinterpolate2(f, [2] * R, piv, tolerance=abstol, maxbonddim=1, maxiter=2, normalizeerror=False,
                                        nsearchglobalpivot=0, strictlynested=strictlynested)
        assert t.linkdims() == [1] * (R - 1)
        for x in (0.1, 0.3, 0.6, 0.9):
            i = int(np.floor(x * 2 ** R))
            bits = [((i >> (R - 1 - k)) & 1) + 1 for k in range(R)]
            assert abs(t.evaluate(bits) - np.exp(-i / 2 ** R)) < abstol


def test_quantics_oscillatory_c4_small(tci):
    """C4 family at R=16: transcendental functor (libdevice vs libm differ in the last ulp), so the oracle is fed
    the device-evaluated values and every decision is compared bit for bit; the values against libm separately."""
    R = 16
    f, g = tci.QuanticsExpPlusSin([2] * R), O.Func(O.F_QEXPPLUSSIN, [1.0, 1e-3, 1000.0])
    fed = O.Func(O.F_BATCHCB, callback=lambda left, right, M: f(left, right, M))
    t, ranks, errors = tci.crossinterpolate2(f, [2] * R, tolerance=1e-10, nsearchglobalpivot=0, maxbonddim=64, engine="python")
    o, oranks, oerrors = O.crossinterpolate2(fed, [2] * R, tolerance=1e-10, maxbonddim=64)
    assert list(ranks) == list(oranks)
    np.testing.assert_array_equal(errors, oerrors)
    _same_sets(t, o)
    _same_tensors(t, o, 1e-9)
    rng = np.random.default_rng(1)
    pts = rng.integers(1, 3, size=(500, R))
    ref = np.array([g(p) for p in pts])
    assert np.max(np.abs(f.evaluate_points(pts) - ref)) <= 1e-12 * np.max(np.abs(ref))
    assert np.max(np.abs(t.evaluate_batch(pts) - ref)) < 1e-7


def test_global_pivot_finder_runs(tci):                 # default nsearchglobalpivot path (own RNG; quality only)
    n = 6
    f = tci.LorentzFunction([10] * n)
    t, ranks, errors = tci.crossinterpolate2(f, [10] * n, tolerance=1e-8, rng=np.random.default_rng(3))
    assert errors[-1] < 1e-8
    g = O.Func(O.F_LORENTZ, [1.0])
    pts = np.random.default_rng(0).integers(1, 11, size=(300, n))
    assert np.max(np.abs(t.evaluate_batch(pts) - np.array([g(p) for p in pts]))) < 1e-6


def test_reconstruct_from_sets(tci):                    # test/test_tensorci2.jl:461-475
    M = np.random.default_rng(1234).random((10, 10))
    f = tci.DenseFunction(M)
    t, _, _ = tci.crossinterpolate2(f, [10, 10], maxbonddim=5, nsearchglobalpivot=0)
    t2 = tci.TensorCI2([10, 10], f, Iset=t.Iset, Jset=t.Jset)
    assert t2.maxsamplevalue == t.maxsamplevalue or t2.maxsamplevalue <= t.maxsamplevalue
    assert all(np.array_equal(a, b) for a, b in zip(t2.Iset, t.Iset))


def test_sweep0site_matches_oracle(tci):                # sweep0site! / rmbadpivots! tensorci2.jl:559-589
    """Same over-resolved state in the product and the oracle (index sets identical after crossinterpolate2), then bad-pivot
    removal bond by bond: the kept index sets must stay identical."""
    ld = [6] * 5
    g = O.Func(O.F_LORENTZ, [1.0])
    o, _, _ = O.crossinterpolate2(g, ld, tolerance=1e-12, maxiter=8)
    f = tci.LorentzFunction(ld)
    t, _, _ = tci.crossinterpolate2(f, ld, tolerance=1e-12, maxiter=8, nsearchglobalpivot=0, engine="python")
    for b in range(1, 6):
        np.testing.assert_array_equal(t.Iset[b - 1], o.Iset(b))
        np.testing.assert_array_equal(t.Jset[b - 1], o.Jset(b))
    before = list(t.linkdims())
    for b in range(1, 5):
        o.sweep0site(b, reltol=1e-6)
        (t.sweep0site if b % 2 else t.rmbadpivots)(f, b, reltol=1e-6)
        np.testing.assert_array_equal(t.Iset[b], o.Iset(b + 1))
        np.testing.assert_array_equal(t.Jset[b - 1], o.Jset(b))
    assert sum(t.linkdims()) < sum(before)
    assert t.maxsamplevalue == o.maxsamplevalue
    with pytest.raises(tci.TCIError):
        t.sweep0site(f, 5)


def test_threaded_batch_evaluator_equals_serial_adapter(tci):      # ThreadedBatchEvaluator batcheval.jl:247-308
    """A host callable through the threaded adapter: every batch equals the serial adapter's and the oracle's bit for bit
    (M = 0, 1, 2 centre sites, empty sets), exceptions propagate, and a whole crossinterpolate2 run reproduces the oracle."""
    import math
    import threading
    ld = [3, 4, 2, 5, 3]
    seen = set()

    def fun(v):
        seen.add(threading.get_ident())
        return math.cos(0.3 * sum((k + 1) * x for k, x in enumerate(v))) + 0.01 * v[0] * v[-1]
    ser = tci.HostFunctionAdapter(fun, ld)
    thr = tci.ThreadedBatchEvaluator(fun, ld, nthreads=4)
    g = O.Func(O.F_CALLBACK, callback=fun)
    rng = np.random.default_rng(11)
    for nl, M in [(2, 0), (2, 1), (1, 2), (3, 1)]:
        nr = len(ld) - nl - M
        left = np.stack([rng.integers(1, ld[p] + 1, size=7) for p in range(nl)], axis=1)
        right = np.stack([rng.integers(1, ld[len(ld) - nr + p] + 1, size=5) for p in range(nr)], axis=1)
        a = tci.to_numpy(thr.evaluate_batch(left, right, M)[0])
        b = tci.to_numpy(ser.evaluate_batch(left, right, M)[0])
        np.testing.assert_array_equal(a, b)
        np.testing.assert_array_equal(a, O.batchevaluate(g, ld, left, right, M))
    assert len(seen) > 1                                             # the pool really ran the callable on several threads
    assert thr.evaluate_batch(np.zeros((0, 2), dtype=np.int64), np.ones((3, 3), dtype=np.int64), 0)[0].numel() == 0
    assert thr([1, 2, 1, 3, 2]) == fun([1, 2, 1, 3, 2])

    def bad(v):
        raise ValueError("boom")
    with pytest.raises(ValueError, match="boom"):
        tci.ThreadedBatchEvaluator(bad, ld, nthreads=2).evaluate_batch(np.ones((2, 2), dtype=np.int64), np.ones((2, 3), dtype=np.int64), 0)
    t, ranks, errors = tci.crossinterpolate2(thr, ld, tolerance=1e-10, nsearchglobalpivot=0)
    o, oranks, oerrors = O.crossinterpolate2(g, ld, tolerance=1e-10)
    assert list(ranks) == list(oranks)
    np.testing.assert_array_equal(errors, oerrors)
    for b in range(1, len(ld) + 1):
        np.testing.assert_array_equal(t.Iset[b - 1], o.Iset(b))
        np.testing.assert_array_equal(t.Jset[b - 1], o.Jset(b))
