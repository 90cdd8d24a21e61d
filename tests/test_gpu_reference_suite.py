"""The remaining test sets of the reference's own suite for this path, run on the GPU path (and against the oracle where it
restates the same call): test/test_tensorci2.jl (checkbatchevaluatable, custom global pivot finder, small maxbonddim,
insert_global_pivots, globalsearch, crossinterpolate2_ttcache, convergencecriterion), test/test_matrixlu.jl (tiny values,
transpose, solve), test/test_cachedtensortrain.jl (TTCache), test/test_globalsearch.jl (estimatetrueerror),
test/test_util.jl (optfirstpivot, pushunique, randomsubset)."""
import itertools

import numpy as np
import pytest

from oracle import oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def tci():
    import tci_b200
    return tci_b200


def _qx(bits):                                          # QuanticsGrids DiscretizedGrid{1}(R, 0, 1): x = sum (b_i - 1) 2^-i
    return sum((b - 1) * 2.0 ** -(i + 1) for i, b in enumerate(bits))


def _bits(x, R):
    k = int(x * 2 ** R)
    return [((k >> (R - 1 - i)) & 1) + 1 for i in range(R)]


def test_checkbatchevaluatable(tci):                    # test/test_tensorci2.jl:41-53
    with pytest.raises(tci.TCIError, match="not batch evaluatable"):
        tci.crossinterpolate2(lambda x: 1.0, [2] * 10, [[1] * 10], checkbatchevaluatable=True)


class CustomGlobalPivotFinder:                          # test/test_tensorci2.jl:105-118
    def __init__(self, npivots):
        self.npivots = npivots

    def __call__(self, input, f, abstol, verbosity=0, rng=None):
        rng = rng or np.random.default_rng(1240)
        return [[int(rng.integers(1, d + 1)) for d in input.localdims] for _ in range(self.npivots)]


def test_custom_global_pivot_finder(tci):               # test/test_tensorci2.jl:120-166
    R, abstol = 8, 1e-4
    f = tci.QuanticsExp2([2] * R, 1.0, 0.0, 0.0)        # exp(-x)
    piv = [[1] * R, [1] + [2] * (R - 1)]
    t, ranks, errors = tci.crossinterpolate2(f, [2] * R, piv, tolerance=abstol, maxbonddim=1, maxiter=2, normalizeerror=False,
                                             globalpivotfinder=CustomGlobalPivotFinder(10), rng=np.random.default_rng(1240))
    assert all(d == 1 for d in t.linkdims())
    tt = tci.TensorTrain(t)
    for x in [0.1, 0.3, 0.6, 0.9]:
        idx = _bits(x, R)
        assert abs(t.evaluate(idx) - np.exp(-_qx(idx))) < abstol
        assert abs(tt.evaluate(idx) - np.exp(-_qx(idx))) < abstol


def test_small_maxbonddim(tci):                         # test/test_tensorci2.jl:168-213
    R = 8
    f = tci.QuanticsExp2([2] * R, 1.0, 1e-4, 2.0)       # exp(-x) + 1e-4 exp(-2x)
    piv = [[1] * R, [1] + [2] * (R - 1)]
    t, _, _ = tci.crossinterpolate2(f, [2] * R, piv, tolerance=1e-10, maxbonddim=1, maxiter=10, normalizeerror=False,
                                    nsearchglobalpivot=10, rng=np.random.default_rng(1240))
    assert all(d == 1 for d in t.linkdims())
    tt = tci.TensorTrain(t)
    for x in [0.1, 0.3, 0.6, 0.9]:
        idx = _bits(x, R)
        ref = np.exp(-_qx(idx)) + 1e-4 * np.exp(-2 * _qx(idx))
        assert abs(t.evaluate(idx) - ref) < 1e-4 and abs(tt.evaluate(idx) - ref) < 1e-4


@pytest.mark.parametrize("pivotsearch", ["full", "rook"])
@pytest.mark.parametrize("seed", [1, 2])
def test_insert_global_pivots_peaks(tci, pivotsearch, seed):       # test/test_tensorci2.jl:343-393 (R = 20, 100 peaks there;
    R, abstol = 14, 1e-4                                            #  smaller here: the host callback is evaluated point by point)
    delta = 10.0 / 2 ** R
    rng = np.random.default_rng(seed)
    rindex = [[int(v) for v in rng.integers(1, 3, size=R)] for _ in range(30)]
    rpoint = np.array([_qx(r) for r in rindex])

    def f(bits):
        x = _qx(bits)
        return np.exp(-10 * x) + 2 * abstol * float(np.count_nonzero(np.abs(x - rpoint) < delta))

    t, _, _ = tci.crossinterpolate2(f, [2] * R, [[1] * R], tolerance=abstol, maxbonddim=1000, maxiter=20, normalizeerror=False,
                                    pivotsearch=pivotsearch, nsearchglobalpivot=0, seed=seed)
    t.addglobalpivots2sitesweep(f, rindex, tolerance=abstol, normalizeerror=False, maxbonddim=1000, pivotsearch=pivotsearch,
                                ntry=1 if pivotsearch == "full" else 10)
    bad = sum(abs(t.evaluate(r) - f(r)) > abstol for r in rindex)
    assert bad == 0


def test_insert_global_pivots_two_points(tci):          # test/test_tensorci2.jl:395-431, parity with the oracle
    R, abstol = 12, 1e-4
    fn = lambda q: 1.0 if (all(v == 1 for v in q) or all(v == 2 for v in q)) else 0.0
    t, _, _ = tci.crossinterpolate2(fn, [2] * R, [[1] * R], tolerance=abstol, maxbonddim=1000, maxiter=20, normalizeerror=False,
                                    nsearchglobalpivot=0)
    o, _, _ = O.crossinterpolate2(O.Func(O.F_CALLBACK, callback=fn), [2] * R, [[1] * R], tolerance=abstol, maxbonddim=1000,
                                  maxiter=20, normalizeerror=False)
    r = [2] * R
    left = t.addglobalpivots2sitesweep(fn, [r], tolerance=abstol, normalizeerror=False, maxbonddim=1000)
    oleft = o.addglobalpivots2sitesweep([r], tolerance=abstol, normalizeerror=False, maxbonddim=1000)
    assert left == oleft == 0
    assert t.evaluate(r) == pytest.approx(1.0, rel=1e-12)
    for b in range(R):
        np.testing.assert_array_equal(t.Iset[b], o.Iset(b + 1))
        np.testing.assert_array_equal(t.Jset[b], o.Jset(b + 1))
    with pytest.raises(tci.TCIError, match="one index per leg"):
        t.addglobalpivots2sitesweep(fn, [[1, 2]])


def test_globalsearch_nasty_function(tci):              # test/test_tensorci2.jl:433-458
    n = 10
    f = tci.QuanticsExpSin([2] * n, 10.0, 100.0, 1.1)
    rng = np.random.default_rng(1234)
    first = tci.optfirstpivot(f, [2] * n, [int(rng.integers(1, 3)) for _ in range(n)])
    t, ranks, errors = tci.crossinterpolate2(f, [2] * n, [first], tolerance=1e-12, maxbonddim=100, maxiter=100,
                                             nsearchglobalpivot=10, rng=rng)
    assert errors[-1] < 1e-10


def test_crossinterpolate2_ttcache(tci):                # test/test_tensorci2.jl:477-502
    rng = np.random.default_rng(5)
    bonds, localdims = [1, 2, 3, 2, 1], [2, 3, 3, 2]
    cores = [rng.random((bonds[n], localdims[n], bonds[n + 1])) for n in range(4)]
    tt = tci.TensorTrain(cores)
    ttc = tci.TTCache(tt)
    t2, ranks, errors = tci.crossinterpolate2(ttc, localdims, tolerance=1e-10, maxbonddim=10, nsearchglobalpivot=0)
    rec = tci.TensorTrain(t2)
    for idx in itertools.product(*[range(1, d + 1) for d in localdims]):
        assert rec.evaluate(list(idx)) == pytest.approx(tt.evaluate(list(idx)), rel=1e-9, abs=1e-12)


def test_ttcache_batch_and_multileg(tci):               # test/test_cachedtensortrain.jl:8-60, test/test_tensortrain.jl:82-136
    rng = np.random.default_rng(2)
    bonds, localdims = [1, 2, 3, 2, 1], [2, 3, 3, 2]
    cores = [rng.random((bonds[n], localdims[n], bonds[n + 1])) for n in range(4)]
    A = tci.TTCache(cores)
    ref = O.tt_evaluate
    for nleft in range(5):
        for nright in range(5 - nleft):
            ncent = 4 - nleft - nright
            left = [list(p) for p in itertools.product(*[range(1, d + 1) for d in localdims[:nleft]])]
            right = [list(p) for p in itertools.product(*[range(1, d + 1) for d in localdims[4 - nright:]])]
            res = A(left, right, ncent)
            assert res.shape == (len(left),) + tuple(localdims[nleft:nleft + ncent]) + (len(right),)
            for il, l in enumerate(left):
                for ir, r in enumerate(right):
                    for c in itertools.product(*[range(1, d + 1) for d in localdims[nleft:nleft + ncent]]):
                        want = ref(cores, [l + list(c) + r])[0]
                        assert res[(il,) + tuple(v - 1 for v in c) + (ir,)] == pytest.approx(want, rel=1e-13)
    ml = [rng.random((bonds[n], 2, 2, bonds[n + 1])) for n in range(4)]          # multi-leg sites are fused (:35-60)
    B = tci.TTCache(ml)
    assert B.sitedims() == [[2, 2]] * 4 and len(B) == 4
    flat = [c.reshape(c.shape[0], 4, c.shape[-1], order="F") for c in ml]
    assert B([1, 2, 3, 4]) == pytest.approx(O.tt_evaluate(flat, [[1, 2, 3, 4]])[0], rel=1e-13)
    with pytest.raises(tci.TCIError, match="do not match"):
        tci.TTCache(ml, sitedims=[[2, 3]] * 4)


def test_estimatetrueerror(tci):                        # test/test_globalsearch.jl:7-37
    R, abstol = 20, 1e-4
    f = tci.QuanticsExpPlusSin([2] * R, 1.0, 1e-3, 1000.0)
    piv = [[1] * R, [1] + [2] * (R - 1)]
    t, _, _ = tci.crossinterpolate2(f, [2] * R, piv, tolerance=abstol, maxbonddim=1, normalizeerror=False, nsearchglobalpivot=0)
    found = tci.estimatetrueerror(tci.TensorTrain(t), f, rng=np.random.default_rng(1240))
    errs = [e for _, e in found]
    assert len(found) > 0
    for p, e in found:
        assert abs(f(p) - t.evaluate(p)) == pytest.approx(e, rel=1e-9, abs=1e-15)
    assert all(a >= b for a, b in zip(errs[:-1], errs[1:]))
    with pytest.raises(tci.TCIError, match="No search"):
        tci.estimatetrueerror(tci.TensorTrain(t), f, nsearch=0)


def test_convergencecriterion(tci):                     # test/test_tensorci2.jl:504-554
    cc = tci.convergencecriterion
    assert cc([1, 2], [1e-2, 1e-5], [0, 0], 1e-4, 4, 3) is False
    assert cc([1, 2, 2, 2], [1e-2, 1e-5, 1e-5, 1e-5], [0, 0, 0, 0], 1e-4, 4, 3) is True
    assert cc([1, 2, 2, 2], [1e-2, 1e-2, 1e-5, 1e-5], [0, 0, 0, 0], 1e-4, 4, 3) is False
    assert cc([1, 2, 2, 2], [1e-2, 1e-2, 1e-2, 1e-2], [0, 0, 0, 0], 1e-4, 2, 3) is True
    assert cc([1, 2, 2, 2], [1e-2, 1e-2, 1e-2, 1e-2], [0, 1, 1, 1], 1e-4, 2, 3) is True


# ---- test/test_matrixlu.jl ----------------------------------------------------------------------------------------------
def test_rrlu_tiny_values(tci):                         # :197-211
    A = 1e-13 * np.array([[0.585383, 0.124568, 0.352426, 0.573507], [0.865875, 0.600153, 0.727443, 0.902388],
                          [0.913477, 0.954081, 0.116965, 0.817], [0.985918, 0.516114, 0.600366, 0.0200085]])
    lu = tci.rrlu(A, abstol=1e-3)
    ref = O.rrlu(A, abstol=1e-3)
    assert lu.npivots() == 1 == ref.npivot
    assert len(lu.pivoterrors()) > 0 and lu.lastpivoterror() > 0 and lu.size() == (4, 4)
    assert np.max(np.abs(lu.left() @ lu.right() - A)) < 1e-3
    np.testing.assert_array_equal(lu.pivoterrors(), ref.pivoterrors())


def test_rrlu_transpose(tci):                           # :214-222
    rng = np.random.default_rng(1234)
    A = rng.random((5, 10))
    tlu = tci.rrlu(A).transpose()
    assert np.allclose(tlu.left() @ tlu.right(), A.T, rtol=1e-13, atol=1e-14)
    assert tlu.leftorthogonal is False and tlu.size() == (10, 5)


def test_rrlu_solve(tci):                               # :224-238 and the disabled 1000 x 1000 case :240-253
    rng = np.random.default_rng(1234)
    A = np.tril(rng.random((5, 5))) @ np.triu(rng.random((5, 5)))
    b = rng.random((5, 2))
    lua = tci.rrlu(A)
    assert np.allclose(lua.left() @ lua.right(), A, rtol=1e-12)
    assert np.allclose(A @ lua.solve(b), b, rtol=1e-9, atol=1e-10)
    N = 1000
    Q1, Q2 = np.linalg.qr(rng.random((N, N)))[0], np.linalg.qr(rng.random((N, N)))[0]
    Abig, bbig = Q1 @ Q2, rng.random((N, 40))
    x = tci.rrlu(Abig).solve(bbig)
    assert np.allclose(Abig @ x, bbig, rtol=1e-9, atol=1e-9)
    with pytest.raises(tci.TCIError, match="square"):
        tci.rrlu(rng.random((4, 6))).solve(rng.random((4, 1)))
    with pytest.raises(tci.TCIError, match="rank-deficient"):
        tci.rrlu(np.ones((4, 4))).solve(rng.random((4, 1)))


# ---- test/test_util.jl ----------------------------------------------------------------------------------------------------
def test_util_functions(tci):
    f = lambda v: float(sum(v))                          # optfirstpivot :28-36: the maximum of sum(v) is the all-d corner
    assert tci.optfirstpivot(f, [3, 4, 5, 6]) == [3, 4, 5, 6]
    assert tci.optfirstpivot(tci.SumFunction([3, 4, 5, 6]), [3, 4, 5, 6], [1, 2, 1, 2]) == [3, 4, 5, 6]
    from tci_b200.tensorci2 import pushunique
    tab = np.zeros((0, 2), dtype=np.int32)               # pushunique! :38-48
    for row in ([1, 2], [1, 2], [2, 1], [1, 2]):
        tab = pushunique(tab, row)
    assert tab.tolist() == [[1, 2], [2, 1]]
    sub = [3]                                            # randomsubset / pushrandomsubset! :61-75
    tci.pushrandomsubset(sub, 10, 4, tci.Rng(3))
    assert len(sub) == 5 and len(set(sub)) == 5 and all(1 <= v <= 10 for v in sub)
    sub2 = [1, 2]
    tci.pushrandomsubset(sub2, 3, 5, tci.Rng(0))         # n is clipped to what is left
    assert sorted(sub2) == [1, 2, 3]
