"""The numpy restatement of src/contraction.jl, pinned by the identities the reference's own tests assert
(test/test_contraction.jl): `_contract` against an explicit permute + matmul, MPO x MPO against the dense matrix
product, projector slices of batchevaluate, MPO x MPS, zip-up and LU/CI compression."""
import itertools

import numpy as np
import pytest

from oracle import contraction_oracle as CO


def _mpo(rng, bonds, d1, d2):
    return [rng.standard_normal((bonds[n], d1[n], d2[n], bonds[n + 1])) for n in range(len(d1))]


def _testdata(seed=0):                                     # test/test_contraction.jl:32-50 (Float64 instead of ComplexF64)
    rng = np.random.default_rng(seed)
    bonds = [1, 2, 3, 2, 1]
    return _mpo(rng, bonds, [2] * 4, [3] * 4), _mpo(rng, bonds, [3] * 4, [2] * 4)


def test_contract_identity():                              # test/test_contraction.jl:25-30
    rng = np.random.default_rng(1)
    a, b = rng.standard_normal((2, 3, 4)), rng.standard_normal((2, 5, 4))
    ab = CO._contract(a, b, (1, 3), (1, 3))
    ref = np.einsum("ijk,ilk->jl", a, b)
    assert ab.shape == (3, 5) and np.allclose(ab, ref, rtol=1e-13, atol=1e-13)


def test_evaluate_and_batchevaluate_match_dense():         # test/test_contraction.jl:73-106, :108-153
    a, b = _testdata()
    dense = CO.tomat(a) @ CO.tomat(b)                      # (2^4) x (2^4)
    obj = CO.Contraction(a, b)
    for idx in itertools.product(*[range(1, 5)] * 4):      # fused index i + 2 (j - 1) per site
        ij = [obj._unfuse(n, v) for n, v in enumerate(idx)]
        r = sum((p[0] - 1) * 2 ** n for n, p in enumerate(ij))
        c = sum((p[1] - 1) * 2 ** n for n, p in enumerate(ij))
        assert abs(obj.evaluate(list(idx)) - dense[r, c]) < 1e-12
    full = obj.batchevaluate([[1], [3]], [[2], [4]], 2)
    assert full.shape == (2, 4, 4, 2)
    for (li, l), (ri, r_) in itertools.product(enumerate([1, 3]), enumerate([2, 4])):
        for c1, c2 in itertools.product(range(1, 5), range(1, 5)):
            assert abs(full[li, c1 - 1, c2 - 1, ri] - obj.evaluate([l, c1, c2, r_])) < 1e-12
    ref = obj.batchevaluate([[1]], [[1]], 2).reshape((1, 2, 2, 2, 2, 1), order="F")
    res = obj.batchevaluate([[1]], [[1]], 2, [[0, 0], [1, 0]])
    assert np.allclose(res.reshape(-1, order="F"), ref[:, :, :, 0:1, :, :].reshape(-1, order="F"), atol=1e-13)
    res = obj.batchevaluate([[1]], [[1]], 2, [[0, 1], [1, 0]])
    assert np.allclose(res.reshape(-1, order="F"), ref[:, :, 0:1, 0:1, :, :].reshape(-1, order="F"), atol=1e-13)


def test_elementwise_f():
    a, b = _testdata(2)
    plain, doubled = CO.Contraction(a, b), CO.Contraction(a, b, f=lambda x: 2 * x)
    assert abs(doubled.evaluate([1, 2, 3, 4]) - 2 * plain.evaluate([1, 2, 3, 4])) < 1e-13
    assert np.allclose(doubled.batchevaluate([[1]], [[2]], 2), 2 * plain.batchevaluate([[1]], [[2]], 2))


@pytest.mark.parametrize("method", ["LU", "CI"])
def test_naive_zipup_compress(method):                     # test/test_contraction.jl:73-106, :192-202; tensortrain.jl:302-348
    a, b = _testdata(3)
    dense = CO.tomat(a) @ CO.tomat(b)
    naive = CO.contract_naive(a, b)
    assert [c.shape for c in naive] == [(1, 2, 2, 4), (4, 2, 2, 9), (9, 2, 2, 4), (4, 2, 2, 1)]
    assert np.allclose(CO.tomat(naive), dense, atol=1e-12)
    comp = CO.compress(naive, method, tolerance=1e-13)
    assert np.allclose(CO.tomat(comp), dense, atol=1e-10) and max(c.shape[-1] for c in comp) <= 9
    zp = CO.contract_zipup(a, b, tolerance=1e-13, method=method)
    assert np.allclose(CO.tomat(zp), dense, atol=1e-10)
    trunc = CO.contract_zipup(a, b, tolerance=1e-13, method=method, maxbonddim=2)
    assert max(c.shape[-1] for c in trunc) <= 2


def test_mpo_mps():                                        # test/test_contraction.jl:155-189
    rng = np.random.default_rng(4)
    bonds = [1, 2, 3, 2, 1]
    a = _mpo(rng, bonds, [3] * 4, [3] * 4)
    v = [rng.standard_normal((bonds[n], 3, 1, bonds[n + 1])) for n in range(4)]      # MPS as an MPO with a unit leg
    res = CO.contract_naive(a, v)
    assert np.allclose(CO.tomat(res), CO.tomat(a) @ CO.tomat(v), atol=1e-12)


# ---- SVD factorisation and tensor-train arithmetic (tensortrain.jl:244-262, abstracttensortrain.jl:467-648) -----------
def _tt(rng, linkdims, localdims):
    return [rng.standard_normal((linkdims[n], localdims[n], linkdims[n + 1])) for n in range(len(localdims))]


def test_add_multiply_subtract():                          # test/test_tensortrain.jl:166-199
    rng = np.random.default_rng(10)
    t1, t2 = _tt(rng, [1, 2, 3, 1], [2, 2, 2]), _tt(rng, [1, 2, 3, 1], [2, 2, 2])
    f1, f2 = CO.fulltensor(t1), CO.fulltensor(t2)
    assert np.allclose(CO.fulltensor(CO.add(t1, t2)), f1 + f2, rtol=1e-12, atol=1e-13)
    half = [c.copy() for c in t1]
    half[-1] = half[-1] * 1.6 / 3.2
    assert np.allclose(CO.fulltensor(CO.subtract(t1, half)), f1 / 2.0, rtol=1e-12, atol=1e-13)
    with pytest.raises(ValueError):
        CO.add(t1, t1[:-1])
    ml = [rng.standard_normal((l, 2, 2, r)) for l, r in [(1, 2), (2, 3), (3, 1)]]
    assert np.allclose(CO.fulltensor(CO.add(ml, ml)), 2 * CO.fulltensor(ml), rtol=1e-12, atol=1e-13)


def test_norm():                                           # test/test_tensortrain.jl:201-214
    rng = np.random.default_rng(3)
    t = _tt(rng, [1, 3, 4, 2, 1], [2, 3, 2, 3])
    full = CO.fulltensor(t)
    assert CO.norm2(t) == pytest.approx(float(np.sum(full ** 2)), rel=1e-13)
    assert CO.norm2(CO.reverse(t)) == pytest.approx(CO.norm2(t), rel=1e-13)


def test_compress_svd():                                   # test/test_tensortrain.jl:216-239
    rng = np.random.default_rng(10)
    N, chi, tol = 10, 10, 0.1
    bond = [1] + [chi] * (N - 1) + [1]
    tt = [rng.standard_normal((bond[n], 2, bond[n + 1])) for n in range(N)]
    comp = CO.compress(tt, "SVD", tolerance=tol)
    assert np.sqrt(CO.norm2(CO.subtract(tt, comp)) / CO.norm2(tt)) < np.sqrt(N) * tol
    comp2 = CO.compress(tt, "SVD", tolerance=np.sqrt(CO.norm2(tt)) * tol, normalizeerror=False)
    assert np.sqrt(CO.norm2(CO.subtract(tt, comp2)) / CO.norm2(tt)) < np.sqrt(N) * tol
    assert max(c.shape[-1] for c in comp) < chi


def test_factorize_svd_truncation_rule():                  # tensortrain.jl:246-251
    rng = np.random.default_rng(2)
    A = rng.standard_normal((30, 8)) @ np.diag([1, 1e-1, 1e-2, 1e-3, 1e-4, 1e-5, 1e-6, 1e-7]) @ rng.standard_normal((8, 20))
    S = np.linalg.svd(A, compute_uv=False)
    l, r, k = CO._factorize(A, "SVD", 0.0, None)
    assert k == 20 and np.allclose(l @ r, A, atol=1e-13)
    l, r, k = CO._factorize(A, "SVD", 1e-4, None, leftorthogonal=True)
    tail = np.array([np.sum(S[n + 1:] ** 2) for n in range(len(S))]) / np.sum(S ** 2)
    assert k == int(np.nonzero(tail < 1e-8)[0][0]) + 1
    assert np.allclose(l.T @ l, np.eye(k), atol=1e-13)
    l, r, k = CO._factorize(A, "SVD", 0.0, 3)
    assert k == 3 and np.allclose(r @ r.T, np.eye(3), atol=1e-13)


def test_zipup_and_naive_default_to_svd():                 # contraction.jl:616-637, :751-788
    a, b = _testdata(4)
    dense = CO.tomat(a) @ CO.tomat(b)
    assert np.allclose(CO.tomat(CO.contract_zipup(a, b)), dense, rtol=1e-10, atol=1e-10)
    assert np.allclose(CO.tomat(CO.contract_naive(a, b, tolerance=1e-12)), dense, rtol=1e-10, atol=1e-10)
