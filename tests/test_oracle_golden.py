"""Pins the CPU oracle against the reference's own known-answer tests.
Each test names the reference test it restates."""
import numpy as np
import pytest

from oracle import oracle as O


def test_argmax_finder(literals):                       # test/test_matrixlu.jl:7-29
    A = literals["test_matrixlu.jl:8:A"]
    m, n = A.shape
    allr, allc = list(range(1, m + 1)), list(range(1, n + 1))
    am = np.unravel_index(np.argmax(A), A.shape)
    assert O.submatrixargmax(A, [3], [5]) == (3, 5)
    assert O.submatrixargmax(A, allr, allc) == (am[0] + 1, am[1] + 1)
    assert O.submatrixargmax(A, [1], allc) == (1, int(np.argmax(A[0, :])) + 1)
    assert O.submatrixargmax(A, allr, [1]) == (int(np.argmax(A[:, 0])) + 1, 1)
    k = min(m, n)
    assert O.submatrixargmax(A, list(range(k, m + 1)), list(range(k, n + 1))) == (k, k)


def test_argmax_throws():                               # test/test_matrixlu.jl:31-37
    A = np.random.default_rng(0).random((10, 10))
    with pytest.raises(O.OracleError, match="rows must not be empty"):
        O.submatrixargmax(A, [], [1])
    with pytest.raises(O.OracleError, match="cols must not be empty"):
        O.submatrixargmax(A, [3], [])


def test_argmax_tie_rule():                             # matrixlu.jl:70-85: smallest column, then smallest row
    A = np.array([[1.0, -2.0, 2.0], [2.0, 2.0, -2.0], [-2.0, 1.0, 2.0]])
    assert O.submatrixargmax(A, [1, 2, 3], [1, 2, 3], mode=1) == (2, 1)
    B = A.copy()
    B[0, 0] = np.nan
    assert O.submatrixargmax(B, [1, 2, 3], [1, 2, 3], mode=1) == (2, 1)     # NaN never wins


def test_rrlu_4x4(literals):                            # test/test_matrixlu.jl:54-69
    A = literals["test_matrixlu.jl:55:A"]
    lu = O.rrlu(A)
    L, U = lu.left(permute=False), lu.right(permute=False)
    assert L.shape == (4, 4) and U.shape == (4, 4)
    assert np.all(L == np.tril(L)) and np.all(np.diag(L) == 1.0)
    assert np.all(U == np.triu(U))
    np.testing.assert_allclose(lu.left() @ lu.right(), A, rtol=1e-13, atol=1e-15)


def test_rrlu_rank1():                                  # test/test_matrixlu.jl:88-97
    A = np.zeros((3, 3))
    A[0, 0] = 1.0
    assert O.rrlu(A).npivot == 1


def test_rrlu_8x6(literals):                            # test/test_matrixlu.jl:99-139
    A = literals["test_matrixlu.jl:100:A"]
    lu = O.rrlu(A, maxrank=4)
    assert len(lu.rowindices()) == 4 and len(lu.colindices()) == 4
    L, U = lu.left(permute=False), lu.right(permute=False)
    assert L.shape == (8, 4) and np.all(L == np.tril(L))
    assert U.shape == (4, 6) and np.all(U == np.triu(U))
    A2 = np.hstack([A, A + 1e-3 * np.random.default_rng(1).random((8, 6))])
    lu = O.rrlu(A2, reltol=1e-2)
    assert len(lu.rowindices()) < 8 and len(lu.colindices()) < 12
    assert np.max(np.abs(lu.left() @ lu.right() - A2)) < 1e-2


def test_rrlu_exact_lowrank(literals):                  # test/test_matrixlu.jl:141-165
    A = literals["test_matrixlu.jl:142:p"] @ literals["test_matrixlu.jl:154:q"]
    lu = O.rrlu(A)
    assert lu.npivot == 3
    np.testing.assert_allclose(lu.left() @ lu.right(), A, rtol=1e-12, atol=1e-14)


def test_rrlu_identity_errors():                        # test/test_matrixlu.jl:167-175
    lu = O.rrlu(np.eye(2))
    assert list(lu.pivoterrors()) == [1.0, 1.0, 0.0]
    assert lu.error == 0.0


def test_rrlu_lastpivoterror(literals):                 # test/test_matrixlu.jl:177-195
    A = literals["test_matrixlu.jl:178:A"]
    lu = O.rrlu(A, maxrank=2)
    assert len(lu.pivoterrors()) == 3 and lu.error > 0
    assert O.rrlu(A, abstol=0.5).error < 0.5
    assert O.rrlu(A, abstol=0.0).error == 0.0


def test_rrlu_tiny_values(literals):                    # test/test_matrixlu.jl:197-211
    A = literals["test_matrixlu.jl:198:A"]
    lu = O.rrlu(A, abstol=1e-3)
    assert lu.npivot == 1 and lu.error > 0
    assert np.max(np.abs(lu.left() @ lu.right() - A)) < 1e-3


def test_luci_8x6(literals):                            # test/test_matrixluci.jl:6-46
    A = literals["test_matrixluci.jl:7:A"]
    lu = O.rrlu(A, maxrank=4)
    I, J = lu.rowindices() - 1, lu.colindices() - 1
    P = A[np.ix_(I, J)]
    for lo in (True, False):
        lu = O.rrlu(A, maxrank=4, leftorthogonal=lo)
        I, J = lu.rowindices() - 1, lu.colindices() - 1
        P = A[np.ix_(I, J)]
        left, right = lu.luci_factors()
        assert left.shape == (8, 4) and right.shape == (4, 6)
        if lo:      # colstimespivotinv = A[:, J] P^-1 (MatrixCI.leftmatrix), rowmatrix = A[I, :]
            np.testing.assert_allclose(left, A[:, J] @ np.linalg.inv(P), rtol=1e-10, atol=1e-12)
            np.testing.assert_allclose(right, A[I, :], rtol=1e-12, atol=1e-14)
        else:       # colmatrix = A[:, J], pivotinvtimesrows = P^-1 A[I, :] (MatrixCI.rightmatrix)
            np.testing.assert_allclose(left, A[:, J], rtol=1e-12, atol=1e-14)
            np.testing.assert_allclose(right, np.linalg.inv(P) @ A[I, :], rtol=1e-10, atol=1e-12)
        np.testing.assert_allclose(left @ right, A[:, J] @ np.linalg.inv(P) @ A[I, :], rtol=1e-10, atol=1e-12)


def test_luci_exact_lowrank(literals):                  # test/test_matrixluci.jl:48-74
    A = literals["test_matrixluci.jl:49:p"] @ literals["test_matrixluci.jl:61:q"]
    lu = O.rrlu(A)
    assert lu.npivot == 3
    left, right = lu.luci_factors()
    np.testing.assert_allclose(left @ right, A, rtol=1e-11, atol=1e-13)
    P = A[np.ix_(lu.rowindices() - 1, lu.colindices() - 1)]
    assert np.linalg.cond(P) < 1e12


def test_solve_right():                                 # tensorci2.jl:626
    rng = np.random.default_rng(3)
    P, B = rng.random((6, 6)) + np.eye(6), rng.random((9, 6))
    np.testing.assert_allclose(O.solve_right(P, B) @ P, B, rtol=1e-12, atol=1e-13)


def test_threaded_rrlu_is_the_same_loop():
    """optimizerrlu_mt (the loop of matrixlu.jl:346-369 with the column ranges spread over host threads) against the scalar
    loop: permutations, pivot errors, L, U bit for bit — full rank, low rank, ties, maxrank and abstol stops, both orientations."""
    rng = np.random.default_rng(5)
    mats = [rng.standard_normal((97, 113)), rng.standard_normal((150, 20)) @ rng.standard_normal((20, 90)),
            np.ones((40, 33)), np.round(rng.standard_normal((64, 64)) * 2) / 2, np.eye(17)]
    for A in mats:
        for kw in (dict(), dict(maxrank=9), dict(abstol=0.5), dict(leftorthogonal=False), dict(reltol=1e-3)):
            a = O.rrlu(A, **kw)
            for nt in (2, 5):
                b = O.rrlu(A, threads=nt, **kw)
                assert a.npivot == b.npivot and (a.error == b.error or (np.isnan(a.error) and np.isnan(b.error)))
                np.testing.assert_array_equal(a.rowpermutation, b.rowpermutation)
                np.testing.assert_array_equal(a.colpermutation, b.colpermutation)
                np.testing.assert_array_equal(a.L, b.L)
                np.testing.assert_array_equal(a.U, b.U)
                np.testing.assert_array_equal(a.pivoterrors(), b.pivoterrors())


def test_batch_callback_functor_matches_builtin():
    """F_BATCHCB (the BatchEvaluator protocol, batcheval.jl:196-214): an oracle run fed with externally produced batches is the
    same run as the built-in functor's — the mechanism the GPU tests use to hand the oracle device-evaluated values."""
    ld = [2] * 12
    f0 = O.Func(O.F_QEXPSIN, [10.0, 100.0, 1.1])
    calls = []

    def cb(left, right, M):
        calls.append((left.shape, right.shape, M))
        return O.batchevaluate(f0, ld, [list(r) for r in left], [list(r) for r in right], M)

    fb = O.Func(O.F_BATCHCB, callback=cb)
    piv = [[2, 1] * 6]
    assert fb(piv[0]) == f0(piv[0])
    a, ra, ea = O.crossinterpolate2(f0, ld, piv, tolerance=1e-10, nsearchglobalpivot=5, seed=3)
    b, rb, eb = O.crossinterpolate2(fb, ld, piv, tolerance=1e-10, nsearchglobalpivot=5, seed=3)
    np.testing.assert_array_equal(ra, rb)
    np.testing.assert_array_equal(ea, eb)
    for s in range(1, 13):
        np.testing.assert_array_equal(a.Iset(s), b.Iset(s))
        np.testing.assert_array_equal(a.Jset(s), b.Jset(s))
        np.testing.assert_array_equal(a.sitetensor(s), b.sitetensor(s))
    assert any(M == 1 for _, _, M in calls) and any(M == 0 for _, _, M in calls)
