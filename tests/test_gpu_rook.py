"""GPU parity of the rook pivot search (pivotsearch=:rook): arrlu (matrixlu.jl:500-569) and the :rook branch of
updatepivots! (tensorci2.jl:863-905) against the oracle.  Product and oracle share the seeded splitmix64 stream that
stands in for Julia's global RNG, so permutations, pivot counts and index sets must be IDENTICAL; L/U and site tensors
are compared to 1e-12 relative (the completion solves run blocked on the device)."""
import itertools

import numpy as np
import pytest

from oracle import oracle as O

pytestmark = pytest.mark.gpu

A4 = np.array([[0.711002, 0.724557, 0.789335, 0.382373], [0.910429, 0.726781, 0.719957, 0.486302],
               [0.632716, 0.39967, 0.571809, 0.0803125], [0.885709, 0.531645, 0.569399, 0.481214]])


@pytest.fixture(scope="module")
def tci():
    import tci_b200
    return tci_b200


def _close(a, b, tol=1e-12):
    assert a.shape == b.shape
    if b.size:
        assert np.max(np.abs(a - b)) <= tol * max(np.max(np.abs(b)), 1e-300), np.max(np.abs(a - b))


def _check_lu(tci, A, I0, J0, seed, **kw):
    lu = tci.arrlu(lambda i, j: A[i - 1, j - 1], A.shape, I0, J0, rng=tci.Rng(seed), **kw)
    ref, _ = O.arrlu(A, I0, J0, seed=seed, **kw)
    assert lu.npivot == ref.npivot
    np.testing.assert_array_equal(lu.rowpermutation, ref.rowpermutation)
    np.testing.assert_array_equal(lu.colpermutation, ref.colpermutation)
    np.testing.assert_array_equal(lu.pivoterrors(), ref.pivoterrors())
    assert lu.error == ref.error
    _close(lu.L.cpu().numpy(), ref.L)
    _close(lu.U.cpu().numpy(), ref.U)
    return lu, ref


@pytest.mark.parametrize("lo", [True, False])
def test_arrlu_reference_matrix(tci, lo):               # test/test_matrixlu.jl:71-86
    for seed in range(4):
        lu, _ = _check_lu(tci, A4, [1], [1], seed, leftorthogonal=lo)
        assert lu.size() == (4, 4)
        L, U = lu.left(permute=False).cpu().numpy(), lu.right(permute=False).cpu().numpy()
        assert np.all(np.tril(L) == L) and np.all(np.triu(U) == U)
        assert np.all(np.diag(L if lo else U) == 1.0)
        assert np.allclose(lu.left().cpu().numpy() @ lu.right().cpu().numpy(), A4, rtol=1e-13, atol=0)


@pytest.mark.parametrize("shape,rank", [((300, 400), 7), ((513, 130), 12), ((64, 700), 5)])
@pytest.mark.parametrize("lo", [True, False])
def test_arrlu_lowrank(tci, shape, rank, lo):
    rng = np.random.default_rng(shape[0] + rank)
    B = rng.standard_normal((shape[0], rank)) @ rng.standard_normal((rank, shape[1]))
    lu, ref = _check_lu(tci, B, [1], [1], 11, reltol=1e-10, leftorthogonal=lo)
    assert lu.npivot == rank
    assert np.allclose(lu.left().cpu().numpy() @ lu.right().cpu().numpy(), B, rtol=0, atol=1e-9 * np.abs(B).max())


def test_arrlu_maxrank_and_initial_sets(tci):
    rng = np.random.default_rng(5)
    B = rng.standard_normal((120, 90))
    lu, _ = _check_lu(tci, B, [3, 17, 40], [2, 9, 33], 3, maxrank=8)
    assert lu.npivot == 8 and lu.L.shape == (120, 8) and lu.U.shape == (8, 90)


def test_rrlu_function_dispatch(tci):                   # matrixlu.jl:599-617
    f = lambda i, j: A4[i - 1, j - 1]
    full = tci.rrlu_function(f, (4, 4), pivotsearch="full")
    ref = O.rrlu(A4)
    np.testing.assert_array_equal(full.rowpermutation, ref.rowpermutation)
    rook = tci.rrlu_function(f, (4, 4), [1], [1], pivotsearch="rook")
    assert rook.npivot == 4
    with pytest.raises(tci.TCIError, match="Unknown pivot search strategy"):
        tci.rrlu_function(f, (4, 4), pivotsearch="nope")


def _same_sets(t, o):
    for b in range(1, len(t) + 1):
        np.testing.assert_array_equal(t.Iset[b - 1], o.Iset(b), err_msg=f"Iset[{b}]")
        np.testing.assert_array_equal(t.Jset[b - 1], o.Jset(b), err_msg=f"Jset[{b}]")


def test_lorentz_5site_steps_rook(tci):                 # test/test_tensorci2.jl:247-280, pivotsearch=:rook
    n = 5
    f, g = tci.LorentzFunction([10] * n), O.Func(O.F_LORENTZ, [1.0])
    t, o = tci.TensorCI2([10] * n, f), O.TCI2(g, [10] * n)
    t.rng = tci.Rng(7)
    o.set_pivotsearch("rook", seed=7)
    for p in range(1, n):
        t.updatepivots(p, f, True, reltol=1e-8, maxbonddim=2, pivotsearch="rook")
        o.updatepivots(p, True, reltol=1e-8, maxbonddim=2)
    assert t.linkdims() == [2] * (n - 1) and t.rank() == 2
    _same_sets(t, o)
    t.addglobalpivots1sitesweep(f, [[2, 9, 10, 5, 7]], reltol=1e-12)
    o.addglobalpivots1sitesweep([[2, 9, 10, 5, 7]], reltol=1e-12)
    assert t.linkdims() == [3] * (n - 1)
    for _ in range(4, 21):
        for p in range(1, n):
            t.updatepivots(p, f, True, reltol=1e-8, pivotsearch="rook")
            o.updatepivots(p, True, reltol=1e-8)
    _same_sets(t, o)
    assert t.maxsamplevalue == o.maxsamplevalue
    np.testing.assert_array_equal(t.pivoterrors, o.pivoterrors)
    assert t.Iset[0].shape[0] == 1 and t.Jset[n - 1].shape[0] == 1


@pytest.mark.parametrize("n,d,tol,strategy", [(5, 10, 1e-8, "forward"), (5, 10, 1e-12, "backandforth"), (7, 6, 1e-10, "backandforth")])
def test_crossinterpolate2_rook_matches_oracle(tci, n, d, tol, strategy):      # test/test_tensorci2.jl:282-339
    f, g = tci.LorentzFunction([d] * n), O.Func(O.F_LORENTZ, [1.0])
    t, ranks, errors = tci.crossinterpolate2(f, [d] * n, tolerance=tol, nsearchglobalpivot=0, maxiter=200, sweepstrategy=strategy,
                                             pivotsearch="rook", seed=21)
    o, oranks, oerrors = O.crossinterpolate2(g, [d] * n, tolerance=tol, maxiter=200, sweepstrategy=strategy, pivotsearch="rook", seed=21)
    assert list(ranks) == list(oranks)
    np.testing.assert_array_equal(np.asarray(errors), oerrors)
    _same_sets(t, o)
    for a, r in zip(t.sitetensors, o.sitetensors()):
        _close(a, r, 1e-11)
    assert t.pivoterror() <= 2 * tol
    for v in itertools.product(range(1, 4), repeat=n):
        assert t.evaluate(list(v)) == pytest.approx(g(v), rel=max(1e-8, 100 * tol))


def test_quantics_rook_agrees_with_full(tci):
    """C3 family (quantics, long chain of dimension-2 legs): the rook search reaches the same tolerance as the full search."""
    R = 20
    f = tci.QuanticsExpPlusSin([2] * R, 1.0, 0.5, 30.0)
    t, _, _ = tci.crossinterpolate2(f, [2] * R, tolerance=1e-9, nsearchglobalpivot=0, maxiter=30, pivotsearch="rook", seed=2)
    t2, _, _ = tci.crossinterpolate2(f, [2] * R, tolerance=1e-9, nsearchglobalpivot=0, maxiter=30)
    assert t.pivoterror() <= 2e-9 * t.maxsamplevalue
    rng = np.random.default_rng(0)
    for v in rng.integers(1, 3, size=(50, R)):
        assert t.evaluate(list(v)) == pytest.approx(t2.evaluate(list(v)), abs=1e-6)
