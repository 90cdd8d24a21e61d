"""Pins the oracle's batch evaluation and TensorCI2 sweeps against the reference's tests
(test/test_batcheval.jl, test/test_tensorci2.jl, test/test_tensortrain.jl)."""
import itertools

import numpy as np
import pytest

from oracle import oracle as O


def test_batcheval_M1_M2():                             # test/test_batcheval.jl:13-38
    f = O.Func(O.F_SUM, [1.0])
    ld = [2] * 5
    left, right = [[1, 1]] * 100, [[1, 1]] * 100
    res = O.batchevaluate(f, ld, left, right, 1)
    ref = np.array([[[sum(l) + c + sum(r) for r in right] for c in (1, 2)] for l in left], dtype=float)
    np.testing.assert_array_equal(res, ref)
    left = [[1]] * 100
    res = O.batchevaluate(f, ld, left, right, 2)
    assert res.shape == (100, 2, 2, 100)
    for c, cp in itertools.product((1, 2), (1, 2)):
        assert res[3, c - 1, cp - 1, 7] == 1 + c + cp + 2


def test_batcheval_shape_and_order():                   # test/test_batcheval.jl:40-69, batcheval.jl:157-171
    rng = np.random.default_rng(5)
    ld = [3, 2, 4, 3, 2, 3]
    T = rng.random(ld)
    f = O.Func(O.F_DENSE, table=T)
    left = [[1, 2], [3, 1], [2, 2]]
    right = [[1, 3], [2, 1]]
    res = O.batchevaluate(f, ld, left, right, 2)
    assert res.shape == (3, 4, 3, 2)
    for i, l in enumerate(left):
        for c in range(4):
            for cp in range(3):
                for j, r in enumerate(right):
                    assert res[i, c, cp, j] == T[l[0] - 1, l[1] - 1, c, cp, r[0] - 1, r[1] - 1]
    assert O.batchevaluate(f, ld, [], right, 2).shape == (0, 0, 0, 0)


def test_pivoterrors_diagonal():                        # test/test_tensorci2.jl:27-39 (exact equality)
    diags = [1.0, 1e-5, 0.0]
    f = O.Func(O.F_DENSE, table=np.diag(diags))
    tci, ranks, errors = O.crossinterpolate2(f, [3, 3], [[1, 1]], tolerance=1e-8)
    assert list(tci.pivoterrors) == diags


def test_callback_matches_dense():
    diags = [1.0, 1e-5, 0.0]
    f = O.Func(O.F_CALLBACK, callback=lambda x: diags[x[0] - 1] if x[0] == x[1] else 0.0)
    tci, _, _ = O.crossinterpolate2(f, [3, 3], [[1, 1]], tolerance=1e-8)
    assert list(tci.pivoterrors) == diags


def test_trivial_mps():                                 # test/test_tensorci2.jl "trivial MPS"
    f = O.Func(O.F_SUM, [0.5])
    tci = O.TCI2(f, [2] * 5, [])
    assert tci.linkdims() == [0] * 4
    tci = O.TCI2(f, [2] * 5, [[1] * 5])
    assert tci.linkdims() == [1] * 4 and tci.rank() == 1


def test_lorentz_5site_linkdims():                      # test/test_tensorci2.jl:247-273
    n = 5
    f = O.Func(O.F_LORENTZ, [1.0])
    tci = O.TCI2(f, [10] * n)
    assert tci.linkdims() == [1] * (n - 1)
    assert len(tci.Iset(1)) == 1 and len(tci.Jset(n)) == 1
    for p in range(1, n):
        tci.updatepivots(p, True, reltol=1e-8, maxbonddim=2)
    assert tci.linkdims() == [2] * (n - 1) and tci.rank() == 2
    tci.addglobalpivots1sitesweep([[2, 9, 10, 5, 7]], reltol=1e-12)
    assert tci.linkdims() == [3] * (n - 1) and tci.rank() == 3
    assert len(tci.Iset(1)) == 1 and len(tci.Jset(n)) == 1


def test_lorentz_5site_tolerance():                     # test/test_tensorci2.jl:297-339
    n = 5
    f = O.Func(O.F_LORENTZ, [1.0])
    tci3, ranks, errors = O.crossinterpolate2(f, [10] * n, [[1] * n], tolerance=1e-12, maxiter=200)
    assert tci3.pivoterror() <= 2e-12
    assert tci3.rank() <= 200
    cores = tci3.sitetensors()
    for v in itertools.product(range(1, 4), repeat=n):
        val = tci3.evaluate(v)
        assert val == pytest.approx(f(v), rel=1e-8)
        assert val == pytest.approx(O.tt_evaluate(cores, [v])[0], rel=1e-12)
    piv = [[1, 1, 1, 1, 1], [10, 8, 10, 4, 4], [5, 4, 8, 9, 3], [7, 7, 10, 5, 9], [7, 7, 10, 5, 9]]
    tci4, _, _ = O.crossinterpolate2(f, [10] * n, piv, tolerance=1e-12, maxiter=200)
    assert tci4.pivoterror() <= 2e-12 and tci4.rank() <= 200


def test_forward_vs_backandforth_rank():                # test/test_tensorci2.jl:275-296
    n = 5
    f = O.Func(O.F_LORENTZ, [1.0])
    tci = O.TCI2(f, [10] * n)
    for _ in range(20):
        for p in range(1, n):
            tci.updatepivots(p, True, reltol=1e-8)
    tci2, _, _ = O.crossinterpolate2(f, [10] * n, [[1] * n], tolerance=1e-8, maxiter=8, sweepstrategy="forward")
    assert tci.rank() == tci2.rank()


def test_quantics_exp_bond1():                          # test/test_tensorci2.jl:55-102 (nsearchglobalpivot=0 cases)
    R, abstol = 8, 1e-4
    f = O.Func(O.F_QEXP2, [1.0, 0.0, 0.0])
    for strictlynested in (False, True):
        tci, _, _ = O.crossinterpolate2(f, [2] * R, [[1] * R, [1] + [2] * (R - 1)], tolerance=abstol, maxbonddim=1,
                                        maxiter=2, normalizeerror=False, strictlynested=strictlynested)
        assert tci.linkdims() == [1] * (R - 1)
        for x in (0.1, 0.3, 0.6, 0.9):
            i = int(np.floor(x * 2 ** R))               # grid index of x = i / 2^R
            bits = [((i >> (R - 1 - k)) & 1) + 1 for k in range(R)]
            assert abs(tci.evaluate(bits) - f(bits)) < abstol
            assert abs(f(bits) - np.exp(-i / 2 ** R)) < 1e-15


def test_reconstruct_from_sets():                       # test/test_tensorci2.jl:461-475 (maxbonddim run)
    rng = np.random.default_rng(1234)
    M = rng.random((10, 10))
    f = O.Func(O.F_DENSE, table=M)
    tci, ranks, _ = O.crossinterpolate2(f, [10, 10], maxbonddim=5)
    assert tci.rank() == 5
    mx = max(abs(M[i[0] - 1, j[0] - 1]) for i in tci.Iset(2) for j in tci.Jset(1))
    assert tci.maxsamplevalue >= mx


def test_tt_sum_bruteforce():                           # test/test_tensortrain.jl:10-32
    f = O.Func(O.F_LORENTZ, [1.0])
    tci, _, _ = O.crossinterpolate2(f, [6] * 4, tolerance=1e-12)
    ref = sum(f(v) for v in itertools.product(range(1, 7), repeat=4))
    assert O.tt_sum(tci.sitetensors()) == pytest.approx(ref, rel=1e-9)


def test_convergence_and_errors_c1():                   # BASELINE config C1 invariants
    f = O.Func(O.F_LORENTZ, [1.0])
    tci, ranks, errors = O.crossinterpolate2(f, [10] * 8, tolerance=1e-8)
    assert errors[-1] < 1e-8
    rng = np.random.default_rng(0)
    pts = rng.integers(1, 11, size=(500, 8))
    assert max(abs(tci.evaluate(p) - f(p)) for p in pts) < 1e-7 * tci.maxsamplevalue


# ---- rook pivot search (matrixlu.jl:500-569, tensorci2.jl:863-905) ---------------------------------------------
A4 = np.array([[0.711002, 0.724557, 0.789335, 0.382373], [0.910429, 0.726781, 0.719957, 0.486302],
               [0.632716, 0.39967, 0.571809, 0.0803125], [0.885709, 0.531645, 0.569399, 0.481214]])


@pytest.mark.parametrize("leftorthogonal", [True, False])
def test_arrlu_reference_matrix(leftorthogonal):        # test/test_matrixlu.jl:71-86
    for seed in range(4):
        lu, _ = O.arrlu(A4, [1], [1], leftorthogonal=leftorthogonal, seed=seed)
        assert lu.L.shape[0] == 4 and lu.U.shape[1] == 4
        unit, other = (lu.L, lu.U) if leftorthogonal else (lu.U.T, lu.L.T)
        assert np.all(np.tril(lu.L) == lu.L) and np.all(np.triu(lu.U) == lu.U)
        assert np.all(np.diag(unit) == 1.0) and other.shape[0] == 4
        assert np.allclose(lu.left() @ lu.right(), A4, rtol=1e-13, atol=0)


def test_arrlu_lowrank_samples_few_elements():
    rng = np.random.default_rng(3)
    B = rng.standard_normal((300, 7)) @ rng.standard_normal((7, 400))
    lu, nev = O.arrlu(B, [1], [1], reltol=1e-10, seed=11)
    assert lu.npivot == 7
    assert nev < 0.35 * B.size                           # the point of the rook search: most of the matrix is never sampled
    assert np.allclose(lu.left() @ lu.right(), B, rtol=0, atol=1e-10 * np.abs(B).max())
    assert sorted(lu.rowpermutation) == list(range(1, 301)) and sorted(lu.colpermutation) == list(range(1, 401))


def test_lorentz_5site_linkdims_rook():                 # test/test_tensorci2.jl:247-273, pivotsearch=:rook
    n = 5
    f = O.Func(O.F_LORENTZ, [1.0])
    tci = O.TCI2(f, [10] * n)
    tci.set_pivotsearch("rook", seed=7)
    for p in range(1, n):
        tci.updatepivots(p, True, reltol=1e-8, maxbonddim=2)
    assert tci.linkdims() == [2] * (n - 1) and tci.rank() == 2
    tci.addglobalpivots1sitesweep([[2, 9, 10, 5, 7]], reltol=1e-12)
    assert tci.linkdims() == [3] * (n - 1) and tci.rank() == 3
    for _ in range(4, 21):
        for p in range(1, n):
            tci.updatepivots(p, True, reltol=1e-8)
    assert len(tci.Iset(1)) == 1 and len(tci.Jset(n)) == 1


def test_lorentz_5site_tolerance_rook():                # test/test_tensorci2.jl:297-339, pivotsearch=:rook
    n = 5
    f = O.Func(O.F_LORENTZ, [1.0])
    tci3, ranks, errors = O.crossinterpolate2(f, [10] * n, [[1] * n], tolerance=1e-12, maxiter=200, pivotsearch="rook", seed=1)
    assert tci3.pivoterror() <= 2e-12 and tci3.rank() <= 200
    for v in itertools.product(range(1, 4), repeat=n):
        assert tci3.evaluate(v) == pytest.approx(f(v), rel=1e-8)


# ---- integrate (src/integration.jl, test/test_integration.jl) ---------------------------------------------------------
POLY = [0.23637074801483304, 0.20661524945577847, 0.1850826417895819, 0.8433788714289417, 0.5801482873508491,
        0.20339438932656262, 0.21593267492457668, 0.8052490409622802, 0.7189346124875339, 0.9400806688257749,
        0.355210845205325, 0.5251561513473092, 0.6819965273401778, 0.9221987248861162, 0.04166444723413998]


def _poly(x):
    return sum(c * x ** i for i, c in enumerate(POLY))


def _polyint(x):
    return sum(c * x ** (i + 1) / (i + 1) for i, c in enumerate(POLY))


def test_integrate_polynomials():                       # test/test_integration.jl:5-29
    N = 5
    f = lambda x: float(np.prod([_poly(v) for v in x]))
    assert O.integrate(f, [0.0] * N, [1.0] * N) == pytest.approx(_polyint(1.0) ** N, rel=1e-8)
    rng = np.random.default_rng(1234)
    a, b = rng.random(N), rng.random(N)
    exact = float(np.prod([_polyint(v) - _polyint(u) for u, v in zip(a, b)]))
    assert O.integrate(f, a, b) == pytest.approx(exact, rel=1e-8)


def test_kronrod15_rule_is_exact_to_degree_22():
    x, w = O.kronrod15()
    assert np.all(np.diff(x) > 0) and abs(w.sum() - 2.0) < 1e-15
    for k in range(0, 23, 2):
        assert abs(np.sum(w * x ** k) - 2.0 / (k + 1)) < 1e-14


def test_insert_global_pivots_oracle():                 # test/test_tensorci2.jl:395-431
    R, abstol = 12, 1e-4
    f = O.Func(O.F_CALLBACK, callback=lambda q: 1.0 if (all(v == 1 for v in q) or all(v == 2 for v in q)) else 0.0)
    tci, _, _ = O.crossinterpolate2(f, [2] * R, [[1] * R], tolerance=abstol, maxbonddim=1000, maxiter=20, normalizeerror=False)
    r = [2] * R
    tci.addglobalpivots2sitesweep([r], tolerance=abstol, normalizeerror=False, maxbonddim=1000)
    assert tci.evaluate(r) == pytest.approx(1.0, rel=1e-12)
    assert tci.evaluate([1] * R) == pytest.approx(1.0, rel=1e-12)


def test_sweep0site_removes_bad_pivots():               # sweep0site! / rmbadpivots! tensorci2.jl:559-589
    """An over-resolved TCI (tolerance 1e-12) has pivot matrices with diagonal entries down to 1e-12 of the first; sweep0site!
    with reltol 1e-6 keeps exactly the leading pivots of rrlu(P) whose |U_ii / U_11| exceeds it.  Checked against the same
    steps written with the oracle's rrlu on a P evaluated point by point."""
    ld = [6] * 5
    g = O.Func(O.F_LORENTZ, [1.0])
    o, _, _ = O.crossinterpolate2(g, ld, tolerance=1e-12, maxiter=8)
    before = o.linkdims()
    for b in range(1, 5):
        I, J = o.Iset(b + 1), o.Jset(b)
        P = np.array([[g(list(i) + list(j)) for j in J] for i in I])
        lu = O.rrlu(P, reltol=1e-6)
        d = np.abs(np.diag(lu.U)[:lu.npivot])
        nd = int(np.sum((d > 0.0) & (d / d[0] > 1e-6)))
        o.rmbadpivots(b, reltol=1e-6)
        np.testing.assert_array_equal(o.Iset(b + 1), I[lu.rowindices()[:nd] - 1])
        np.testing.assert_array_equal(o.Jset(b), J[lu.colindices()[:nd] - 1])
    after = o.linkdims()
    assert all(a <= b for a, b in zip(after, before)) and sum(after) < sum(before)
    with pytest.raises(O.OracleError):
        o.sweep0site(5)
