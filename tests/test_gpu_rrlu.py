"""GPU parity of the rrLU path (through the C ABI) against the CPU oracle: pivots, permutations, L and U
must be BIT-EXACT (same IEEE operations in the same order, ties included)."""
import numpy as np
import pytest

from oracle import oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def tci():
    import tci_b200
    return tci_b200


def _same(lu, ref):
    assert lu.npivot == ref.npivot
    np.testing.assert_array_equal(lu.rowpermutation, ref.rowpermutation)
    np.testing.assert_array_equal(lu.colpermutation, ref.colpermutation)
    np.testing.assert_array_equal(lu.L, ref.L)
    np.testing.assert_array_equal(lu.U, ref.U)
    np.testing.assert_array_equal(lu.pivoterrors(), ref.pivoterrors())
    assert lu.error == ref.error or (np.isnan(lu.error) and np.isnan(ref.error))


def test_golden_literals(tci, literals):
    for key, A in literals.items():
        if A.shape[0] < 2:
            continue
        for lo in (True, False):
            _same(tci.rrlu(A, leftorthogonal=lo), O.rrlu(A, leftorthogonal=lo))
    A = literals["test_matrixlu.jl:100:A"]
    _same(tci.rrlu(A, maxrank=4), O.rrlu(A, maxrank=4))
    A = literals["test_matrixlu.jl:178:A"]
    for kw in (dict(maxrank=2), dict(abstol=0.5), dict(abstol=0.0)):
        _same(tci.rrlu(A, **kw), O.rrlu(A, **kw))
    A = literals["test_matrixlu.jl:198:A"]
    lu = tci.rrlu(A, abstol=1e-3)
    assert lu.npivot == 1 and lu.error > 0
    lu = tci.rrlu(np.eye(2))
    assert list(lu.pivoterrors()) == [1.0, 1.0, 0.0] and lu.lastpivoterror() == 0.0
    A = np.zeros((3, 3)); A[0, 0] = 1.0
    assert tci.rrlu(A).npivot == 1


@pytest.mark.parametrize("shape", [(1, 1), (1, 7), (9, 1), (33, 65), (64, 64), (34, 34), (4096, 1), (1, 4096), (3000, 8), (8, 3100),
                                   (257, 130), (513, 515), (700, 300), (1025, 1024)])
@pytest.mark.parametrize("lo", [True, False])
def test_random_shapes(tci, shape, lo):
    rng = np.random.default_rng(shape[0] * 1000 + shape[1])
    A = rng.standard_normal(shape)
    kw = dict(maxrank=min(64, *shape), leftorthogonal=lo)
    lu, ref = tci.rrlu(A, **kw), O.rrlu(A, **kw)
    _same(lu, ref)
    if lu.npivot == min(shape):
        np.testing.assert_allclose(lu.left() @ lu.right(), A, rtol=1e-9, atol=1e-9)


def test_lowrank_stops(tci):
    rng = np.random.default_rng(7)
    A = rng.standard_normal((300, 12)) @ rng.standard_normal((12, 280))
    lu, ref = tci.rrlu(A, reltol=1e-10), O.rrlu(A, reltol=1e-10)
    assert lu.npivot == 12
    _same(lu, ref)
    np.testing.assert_allclose(lu.left() @ lu.right(), A, rtol=1e-8, atol=1e-10)


def test_ties_follow_reference_order(tci):
    """A Lorentzian Pi matrix has many exactly equal entries (duplicate rows/columns): the tie rule
    (smallest column, then smallest row, in swapped coordinates) decides the pivots."""
    v = np.arange(1, 11)
    rows = [(a, b) for b in v for a in v]
    cols = [(a, b) for b in v for a in v]
    a2 = np.array([r[0] ** 2 + r[1] ** 2 for r in rows], dtype=np.int64)
    b2 = np.array([c[0] ** 2 + c[1] ** 2 for c in cols], dtype=np.int64)
    A = 1.0 / (1 + a2[:, None] + b2[None, :]).astype(np.float64)
    for lo in (True, False):
        kw = dict(reltol=1e-14, abstol=1e-9, leftorthogonal=lo)
        _same(tci.rrlu(A, **kw), O.rrlu(A, **kw))
    B = np.ones((40, 50))
    _same(tci.rrlu(B), O.rrlu(B))
    Z = np.zeros((17, 9))      # the reference always takes a first pivot (matrixlu.jl:363): 0/0 -> NaN -> error
    with pytest.raises(O.OracleError, match="contains NaNs"):
        O.rrlu(Z)
    with pytest.raises(tci.TCIError, match="contains NaNs"):
        tci.rrlu(Z)


def test_nan_raises_like_reference(tci):
    A = np.random.default_rng(0).standard_normal((20, 20))
    A[3, 4] = np.nan
    with pytest.raises(O.OracleError, match="contains NaNs"):
        O.rrlu(A)
    with pytest.raises(tci.TCIError, match="contains NaNs"):
        tci.rrlu(A)


def test_argmax(tci, literals):
    A = literals["test_matrixlu.jl:8:A"]
    m, n = A.shape
    for k in (1, 2, 5, min(m, n)):
        ref = O.submatrixargmax(A, list(range(k, m + 1)), list(range(k, n + 1)), mode=1)
        assert tci.submatrixargmax(A, k) == ref
    with pytest.raises(tci.TCIError, match="rows must not be empty"):
        tci.submatrixargmax(np.ones((10, 10)), 100)


def test_device_entry_point_and_odd_lda(tci):
    import torch
    from tci_b200.matrixlu import _rrlu_device
    h = tci.default_handle()
    rng = np.random.default_rng(11)
    for (m, n, lda) in [(100, 90, 100), (101, 90, 101), (64, 200, 67), (300, 300, 304)]:
        A = rng.standard_normal((m, n))
        buf = np.zeros((n, lda))
        buf[:, :m] = A.T
        dA = torch.from_numpy(buf).cuda()
        lu = _rrlu_device(h, dA, m, n, lda, 50, 1e-14, 0.0, True)
        ref = O.rrlu(A, maxrank=50)
        assert lu.npivot == ref.npivot
        np.testing.assert_array_equal(lu.rowpermutation, ref.rowpermutation)
        np.testing.assert_array_equal(lu.colpermutation, ref.colpermutation)
        np.testing.assert_array_equal(lu.L.cpu().numpy(), ref.L)
        np.testing.assert_array_equal(lu.U.cpu().numpy(), ref.U)


def test_large_size_properties(tci):
    """Full-size property check (BASELINE config C2 shape family at 2048): L*U reproduces A on the pivot
    rows/cols and the residual is below the last pivot."""
    rng = np.random.default_rng(20260001)
    N, R = 2048, 96
    sig = 2.0 ** (-np.arange(R) / 4.0)
    A = (rng.standard_normal((N, R)) * sig) @ rng.standard_normal((R, N))
    lu = tci.rrlu(A, maxrank=R)
    assert lu.npivot == R
    ref = O.rrlu(A, maxrank=R)
    _same(lu, ref)
    resid = np.max(np.abs(lu.left() @ lu.right() - A))
    assert resid < 1e-6 * np.max(np.abs(A))
