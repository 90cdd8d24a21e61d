"""GPU parity of K4a (DMMA GEMM, triangular solves, MatrixLUCI factors, Pi1 * inv(P)) and K4b (TT sum /
evaluate) against the oracle / numpy in FP64.  Tolerance: 1e-12 relative to the largest entry (north star)."""
import itertools

import numpy as np
import pytest

from oracle import oracle as O

pytestmark = pytest.mark.gpu
RTOL = 1e-12


@pytest.fixture(scope="module")
def tci():
    import tci_b200
    return tci_b200


def _close(a, b, tol=RTOL):
    scale = max(np.max(np.abs(b)), 1e-300)
    assert a.shape == b.shape
    assert np.max(np.abs(a - b)) <= tol * scale, np.max(np.abs(a - b)) / scale


@pytest.mark.parametrize("shape", [(1, 1, 1), (7, 5, 3), (64, 64, 16), (65, 130, 33), (300, 17, 257), (512, 256, 512)])
@pytest.mark.parametrize("ta,tb", [(False, False), (True, False), (False, True), (True, True)])
def test_gemm(tci, shape, ta, tb):
    M, N, K = shape
    rng = np.random.default_rng(M + N + K)
    A = rng.standard_normal((K, M) if ta else (M, K))
    B = rng.standard_normal((N, K) if tb else (K, N))
    h = tci.default_handle()
    dev = f"cuda:{h.device}"
    Cm = tci.gemm(h, tci.Mat.from_numpy(A, dev, pad=4), tci.Mat.from_numpy(B, dev), transa=ta, transb=tb)
    ref = (A.T if ta else A) @ (B.T if tb else B)
    _close(Cm.numpy(), ref, 1e-13 * max(K, 1))


def test_gemm_accumulate(tci):
    rng = np.random.default_rng(0)
    h = tci.default_handle()
    dev = f"cuda:{h.device}"
    A, B, C0 = rng.standard_normal((50, 20)), rng.standard_normal((20, 30)), rng.standard_normal((50, 30))
    Cm = tci.Mat.from_numpy(C0, dev)
    tci.gemm(h, tci.Mat.from_numpy(A, dev), tci.Mat.from_numpy(B, dev), C_out=Cm, alpha=-2.0, beta=0.5)
    _close(Cm.numpy(), -2.0 * A @ B + 0.5 * C0)


@pytest.mark.parametrize("lo", [True, False])
@pytest.mark.parametrize("shape,r", [((8, 6), 4), ((100, 90), 33), ((300, 520), 64), ((257, 129), 129), ((40, 40), 40)])
def test_luci_factors(tci, shape, r, lo):
    rng = np.random.default_rng(shape[0] + r)
    A = rng.standard_normal((shape[0], r)) @ rng.standard_normal((r, shape[1])) + 1e-3 * rng.standard_normal(shape)
    luci = tci.MatrixLUCI(A, maxrank=r, leftorthogonal=lo)
    ref = O.rrlu(A, maxrank=r, leftorthogonal=lo)
    assert luci.npivots() == ref.npivot
    np.testing.assert_array_equal(luci.rowindices(), ref.rowindices())
    np.testing.assert_array_equal(luci.colindices(), ref.colindices())
    left, right = ref.luci_factors()
    _close(luci.left(), left, 1e-11)
    _close(luci.right(), right, 1e-11)
    I, J = ref.rowindices() - 1, ref.colindices() - 1
    P = A[np.ix_(I, J)]
    _close(luci.colmatrix(), A[:, J], 1e-11)
    _close(luci.rowmatrix(), A[I, :], 1e-11)
    if np.linalg.cond(P) < 1e6:
        _close(luci.colstimespivotinv(), A[:, J] @ np.linalg.inv(P), 1e-9)
        _close(luci.pivotinvtimesrows(), np.linalg.inv(P) @ A[I, :], 1e-9)


def test_luci_golden(tci, literals):                     # test/test_matrixluci.jl:6-46
    A = literals["test_matrixluci.jl:7:A"]
    luci = tci.MatrixLUCI(A, maxrank=4)
    assert luci.size() == (8, 6) and len(luci.rowindices()) == 4 and len(luci.colindices()) == 4
    I, J = luci.rowindices() - 1, luci.colindices() - 1
    P = A[np.ix_(I, J)]
    _close(luci.colstimespivotinv(), A[:, J] @ np.linalg.inv(P), 1e-11)
    _close(luci.pivotinvtimesrows(), np.linalg.inv(P) @ A[I, :], 1e-11)
    _close(luci.left() @ luci.right(), A[:, J] @ np.linalg.inv(P) @ A[I, :], 1e-11)
    A3 = literals["test_matrixluci.jl:49:p"] @ literals["test_matrixluci.jl:61:q"]
    luci = tci.MatrixLUCI(A3)
    assert luci.npivots() == 3
    _close(luci.left() @ luci.right(), A3, 1e-11)


@pytest.mark.parametrize("q,r", [(1, 1), (30, 7), (129, 33), (1000, 40), (200, 64), (300, 65), (5120, 256)])
def test_solve_right(tci, q, r):                         # tensorci2.jl:626
    rng = np.random.default_rng(q + r)
    P = rng.standard_normal((r, r)) + 3.0 * np.eye(r)
    B = rng.standard_normal((q, r))
    h = tci.default_handle()
    dev = f"cuda:{h.device}"
    X = tci.solve_right(h, tci.Mat.from_numpy(B, dev), tci.Mat.from_numpy(P, dev)).numpy()
    ref = O.solve_right(P, B)
    _close(X, ref, 1e-12 * np.linalg.cond(P))
    _close(X @ P, B, 1e-12 * np.linalg.cond(P))


def test_solve_right_small_kernel_cases(tci):
    """The one-launch path for pivot matrices up to 64 x 64: padded leading dimensions, a badly scaled P (full pivoting),
    exact singularity reported like the general path, and agreement with that path (TCI_SOLVE_SMALL is read once per process,
    so the general path is reached through r = 65 above and through the factor identities here)."""
    h = tci.default_handle()
    dev = f"cuda:{h.device}"
    rng = np.random.default_rng(11)
    q, r = 77, 20
    P = rng.standard_normal((r, r)) * np.logspace(0, -9, r)[None, :]          # columns scaled over nine decades
    B = rng.standard_normal((q, r))
    X = tci.solve_right(h, tci.Mat.from_numpy(B, dev, pad=16), tci.Mat.from_numpy(P, dev, pad=16)).numpy()
    resid = np.abs(X @ P - B)                                                 # backward error: small against |X| |P| whatever cond(P) is
    assert np.max(resid / (np.abs(X) @ np.abs(P))) < 1e-13 * r
    Ps = P.copy(); Ps[:, 3] = 0.0                                             # a zero column stays exactly zero: singular
    with pytest.raises(tci.TCIError, match="singular"):
        tci.solve_right(h, tci.Mat.from_numpy(B, dev), tci.Mat.from_numpy(Ps, dev))
    Pn = P.copy(); Pn[2, 2] = np.nan
    try:                                                                      # NaN input must not hang or crash; either outcome is
        tci.solve_right(h, tci.Mat.from_numpy(B, dev), tci.Mat.from_numpy(Pn, dev))   # an error or a NaN-carrying result
    except tci.TCIError:
        pass


def test_tt_sum_and_evaluate(tci):                       # test/test_tensortrain.jl:10-32
    rng = np.random.default_rng(6)
    bonds = [1, 3, 7, 5, 2, 1]
    dims = [6, 4, 5, 3, 6]
    cores = [rng.standard_normal((bonds[i], dims[i], bonds[i + 1])) for i in range(5)]
    tt = tci.TensorTrain(cores)
    full = cores[0].reshape(dims[0], -1)
    for c in cores[1:]:
        full = (full @ c.reshape(c.shape[0], -1)).reshape(-1, c.shape[-1])
    full = full.reshape(dims)
    assert tt.sum() == pytest.approx(full.sum(), rel=1e-12, abs=1e-12 * np.abs(full).sum())
    assert tt.sum() == pytest.approx(O.tt_sum(cores), rel=1e-12, abs=1e-12 * np.abs(full).sum())
    pts = [[int(rng.integers(1, d + 1)) for d in dims] for _ in range(300)]
    got = tt.evaluate_batch(pts)
    ref = O.tt_evaluate(cores, pts)
    assert np.max(np.abs(got - ref)) <= 1e-12 * np.max(np.abs(full))
    for p in pts[:5]:
        assert tt(p) == pytest.approx(full[tuple(v - 1 for v in p)], abs=1e-12 * np.max(np.abs(full)))
    assert tt.linkdims() == bonds[1:-1] and tt.rank() == 7
    _close(tt.fulltensor(), full, 1e-12)
    with pytest.raises(tci.TCIError, match="you have to provide 5 indices"):
        tt.evaluate([1, 1])


def test_tt_large_bond(tci):
    rng = np.random.default_rng(8)
    chi, d, n = 200, 16, 6
    bonds = [1] + [chi] * (n - 1) + [1]
    cores = [rng.standard_normal((bonds[i], d, bonds[i + 1])) / np.sqrt(chi * d) for i in range(n)]
    tt = tci.TensorTrain(cores)
    ref = O.tt_sum(cores)
    scale = np.prod([np.abs(c).sum(axis=1).max() for c in cores])
    assert abs(tt.sum() - ref) <= 1e-12 * max(scale, abs(ref))
    pts = rng.integers(1, d + 1, size=(64, n))
    got, refv = tt.evaluate_batch(pts), O.tt_evaluate(cores, pts)
    assert np.max(np.abs(got - refv)) <= 1e-12 * np.max(np.abs(refv)) * chi
