"""GPU parity of the device SVD (csrc/svd.cu, one-sided Jacobi) and of the tensor-train arithmetic built on it
(_factorize(:SVD), compress!(:SVD), add/subtract/norm2, multiply/divide/reverse, SVD-default contract_naive/zipup)
against the numpy oracle (LAPACK gesdd, like the reference's LinearAlgebra.svd).  Singular values agree to 1e-12
relative to the largest; factors are compared through their product and their orthonormality (signs are free)."""
import numpy as np
import pytest

from oracle import contraction_oracle as CO

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def tci():
    import tci_b200
    return tci_b200


def _svd(tci, A):
    h = tci.default_handle()
    from tci_b200.dense import svd
    U, S, Vt = svd(h, tci.Mat.from_numpy(A, "cuda:0"))
    return U.numpy(), S, Vt.numpy()


@pytest.mark.parametrize("shape", [(1, 1), (7, 1), (1, 9), (5, 5), (33, 17), (17, 33), (200, 64), (64, 200), (513, 130), (300, 301)])
def test_svd_random(tci, shape):
    rng = np.random.default_rng(shape[0] * 1000 + shape[1])
    A = rng.standard_normal(shape)
    U, S, Vt = _svd(tci, A)
    k = min(shape)
    Sref = np.linalg.svd(A, compute_uv=False)
    assert U.shape == (shape[0], k) and Vt.shape == (k, shape[1]) and S.shape == (k,)
    assert np.all(np.diff(S) <= 0)
    assert np.max(np.abs(S - Sref)) <= 1e-12 * Sref[0]
    assert np.max(np.abs(U @ np.diag(S) @ Vt - A)) <= 1e-12 * Sref[0]
    assert np.max(np.abs(U.T @ U - np.eye(k))) <= 1e-12 and np.max(np.abs(Vt @ Vt.T - np.eye(k))) <= 1e-12


def test_svd_graded_and_rank_deficient(tci):
    rng = np.random.default_rng(2)
    sig = 10.0 ** -np.arange(0, 16, 2.0)
    A = np.linalg.qr(rng.standard_normal((60, 8)))[0] @ np.diag(sig) @ np.linalg.qr(rng.standard_normal((40, 8)))[0].T
    U, S, Vt = _svd(tci, A)
    Sref = np.linalg.svd(A, compute_uv=False)               # A itself carries 1e-16 absolute rounding: compare with LAPACK on the same A
    assert np.max(np.abs(S - Sref)) <= 1e-13 and np.max(np.abs(S[:5] - sig[:5]) / sig[:5]) < 1e-6
    assert np.max(np.abs(U @ np.diag(S) @ Vt - A)) <= 1e-14
    Z = np.zeros((6, 4))
    U, S, Vt = _svd(tci, Z)
    assert np.all(S == 0) and np.all(U == 0)


@pytest.mark.parametrize("lo", [True, False])
def test_factorize_svd_matches_oracle(tci, lo):            # tensortrain.jl:244-262
    rng = np.random.default_rng(5)
    A = rng.standard_normal((30, 8)) @ np.diag([1, 1e-1, 1e-2, 1e-3, 1e-4, 1e-5, 1e-6, 1e-7]) @ rng.standard_normal((8, 20))
    for tol, mb, ne in [(0.0, None, True), (1e-4, None, True), (0.0, 3, True), (1e-3, None, False), (1e-6, 5, True)]:
        l, r, k = tci._factorize(tci.Mat.from_numpy(A, "cuda:0"), "SVD", tol, mb, leftorthogonal=lo, normalizeerror=ne)
        ol, orr, ok = CO._factorize(A, "SVD", tol, mb, leftorthogonal=lo, normalizeerror=ne)
        assert k == ok and l.numpy().shape == ol.shape and r.numpy().shape == orr.shape
        assert np.max(np.abs(l.numpy() @ r.numpy() - ol @ orr)) <= 1e-12 * np.abs(A).max()
        iso = l.numpy().T @ l.numpy() if lo else r.numpy() @ r.numpy().T
        assert np.max(np.abs(iso - np.eye(k))) <= 1e-10


def _tt(rng, linkdims, localdims):
    return [rng.standard_normal((linkdims[n], localdims[n], linkdims[n + 1])) for n in range(len(localdims))]


def _close(a, b, tol=1e-12):
    assert a.shape == b.shape and np.max(np.abs(a - b)) <= tol * max(1.0, np.abs(b).max()), np.max(np.abs(a - b))


def test_add_multiply_divide_subtract(tci):                # test/test_tensortrain.jl:166-199
    rng = np.random.default_rng(10)
    c1, c2 = _tt(rng, [1, 2, 3, 1], [2, 2, 2]), _tt(rng, [1, 2, 3, 1], [2, 2, 2])
    t1, t2 = tci.TensorTrain(c1), tci.TensorTrain(c2)
    f1, f2 = CO.fulltensor(c1), CO.fulltensor(c2)
    _close(tci.add(t1, t2).fulltensor(), f1 + f2)
    _close((t1 + t2).fulltensor(), CO.fulltensor(CO.add(c1, c2)))
    assert (t1 + t2).linkdims() == [c.shape[-1] for c in CO.add(c1, c2)[:-1]]
    mul = 1.6 * t1
    _close(mul.fulltensor(), 1.6 * f1)
    _close((t1 * 1.6).fulltensor(), 1.6 * f1)
    div = mul / 3.2
    _close(div.fulltensor(), f1 / 2.0)
    _close((t1 - div).fulltensor(), f1 / 2.0)
    _close(t1.fulltensor(), f1)                             # multiply/divide copy, the operand is untouched
    with pytest.raises(tci.TCIError, match="different length"):
        tci.add(t1, tci.TensorTrain(c1[:-1]))
    ml = [rng.standard_normal((l, 2, 2, r)) for l, r in [(1, 2), (2, 3), (3, 1)]]
    tm = tci.TensorTrain(ml)
    with pytest.raises(tci.TCIError, match="same indices"):
        tci.add(t1, tm)
    got = (tm + tm)
    for idx in [((1, 1), (2, 1), (1, 2)), ((2, 2), (1, 2), (2, 1))]:
        assert got.evaluate(idx) == pytest.approx(2 * tm.evaluate(idx), rel=1e-12)


def test_norm_and_reverse(tci):                            # test/test_tensortrain.jl:201-214, tensortrain.jl:452-456
    rng = np.random.default_rng(3)
    c = _tt(rng, [1, 3, 4, 2, 1], [2, 3, 2, 3])
    t = tci.TensorTrain(c)
    assert tci.norm2(t) == pytest.approx(CO.norm2(c), rel=1e-12)
    assert tci.norm(t) == pytest.approx(np.sqrt(CO.norm2(c)), rel=1e-12)
    rev = t.reverse()
    for a, b in zip(rev.sitetensors, CO.reverse(c)):
        assert np.array_equal(a, b)
    assert tci.norm2(rev) == pytest.approx(CO.norm2(c), rel=1e-12)


def test_compress_svd(tci):                                # test/test_tensortrain.jl:216-239
    rng = np.random.default_rng(10)
    N, chi, tol = 10, 10, 0.1
    bond = [1] + [chi] * (N - 1) + [1]
    cores = [rng.standard_normal((bond[n], 2, bond[n + 1])) for n in range(N)]
    tt = tci.TensorTrain(cores)
    comp = tt.copy()
    comp.compress("SVD", tolerance=tol)
    ocomp = CO.compress(cores, "SVD", tolerance=tol)
    assert comp.linkdims() == [c.shape[-1] for c in ocomp[:-1]]
    assert np.sqrt(tci.norm2(tt - comp) / tci.norm2(tt)) < np.sqrt(N) * tol
    _close(comp.fulltensor(), CO.fulltensor(ocomp), 1e-11)
    comp2 = tt.copy()
    comp2.compress("SVD", tolerance=tci.norm(tt) * tol, normalizeerror=False)
    assert np.sqrt(tci.norm2(tt - comp2) / tci.norm2(tt)) < np.sqrt(N) * tol
    assert comp2.linkdims() == [c.shape[-1] for c in CO.compress(cores, "SVD", tolerance=np.sqrt(CO.norm2(cores)) * tol, normalizeerror=False)[:-1]]


def test_contract_defaults_use_svd(tci):                   # contraction.jl:616-637, :751-788, test/test_contraction.jl:73-106, :192-202
    rng = np.random.default_rng(4)
    bonds = [1, 2, 3, 2, 1]
    a = [rng.standard_normal((bonds[n], 2, 3, bonds[n + 1])) for n in range(4)]
    b = [rng.standard_normal((bonds[n], 3, 2, bonds[n + 1])) for n in range(4)]
    A, B = tci.TensorTrain(a), tci.TensorTrain(b)
    dense = CO.tomat(a) @ CO.tomat(b)
    zp = tci.contract(A, B, algorithm="zipup")
    _close(CO.tomat(zp.sitetensors), dense, 1e-11)
    assert zp.linkdims() == [c.shape[-1] for c in CO.contract_zipup(a, b)[:-1]]
    nv = tci.contract(A, B, algorithm="naive", tolerance=1e-12)
    _close(CO.tomat(nv.sitetensors), dense, 1e-11)
    assert nv.linkdims() == [c.shape[-1] for c in CO.contract_naive(a, b, tolerance=1e-12)[:-1]]
    tr = tci.contract_zipup(A, B, maxbonddim=2)
    _close(CO.tomat(tr.sitetensors), CO.tomat(CO.contract_zipup(a, b, maxbonddim=2)), 1e-10)
