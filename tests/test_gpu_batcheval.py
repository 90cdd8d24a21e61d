"""GPU parity of K1 (batched evaluation + Pi assembly) against the oracle's restatement of
src/batcheval.jl:131-175.  Integer-exact functors (Lorentzian, sum, dense table) must be bit-exact;
transcendental ones (CUDA libdevice vs libm) within 1e-12 relative to the largest value."""
import itertools

import numpy as np
import pytest

from oracle import oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def tci():
    import tci_b200
    return tci_b200


def _sets(rng, localdims, nl, nr, nleft, nright):
    N = len(localdims)
    left = [[int(rng.integers(1, localdims[p] + 1)) for p in range(nl)] for _ in range(nleft)]
    right = [[int(rng.integers(1, localdims[N - nr + p] + 1)) for p in range(nr)] for _ in range(nright)]
    return left, right


def test_reference_batcheval_cases(tci):                # test/test_batcheval.jl:13-49
    ld = [2] * 5
    f, g = tci.SumFunction(ld), O.Func(O.F_SUM, [1.0])
    left, right = [[1, 1]] * 100, [[1, 1]] * 100
    np.testing.assert_array_equal(f(left, right, 1), O.batchevaluate(g, ld, left, right, 1))
    left = [[1]] * 100
    res = f(left, right, 2)
    assert res.shape == (100, 2, 2, 100)
    np.testing.assert_array_equal(res, O.batchevaluate(g, ld, left, right, 2))
    f3 = tci.SumFunction([3, 3, 3])
    assert f3([[1], [2]], [[1], [2]], 1).shape == (2, 3, 2)
    assert f3([], [[1]], 2).shape == (0, 0, 0, 0)       # batcheval.jl:140-142
    with pytest.raises(tci.TCIError, match="Invalid number of central indices"):   # tensorci2.jl:492
        f3([[1]], [[1]], 2)


@pytest.mark.parametrize("M", [0, 1, 2])
def test_lorentz_bit_exact(tci, M):
    rng = np.random.default_rng(M)
    ld = [10, 7, 5, 10, 3, 10, 8]
    nl, nr = 2, len(ld) - 2 - M
    left, right = _sets(rng, ld, nl, nr, 37, 53)
    f, g = tci.LorentzFunction(ld, 1.0), O.Func(O.F_LORENTZ, [1.0])
    np.testing.assert_array_equal(f(left, right, M), O.batchevaluate(g, ld, left, right, M))
    assert f([1] * len(ld)) == g([1] * len(ld))


def test_dense_and_ordering(tci):                       # centre index order: first index fastest (batcheval.jl:160)
    rng = np.random.default_rng(5)
    ld = [3, 2, 4, 3, 2, 3]
    T = rng.random(ld)
    f, g = tci.DenseFunction(T), O.Func(O.F_DENSE, table=T)
    left, right = [[1, 2], [3, 1], [2, 2]], [[1, 3], [2, 1]]
    res = f(left, right, 2)
    assert res.shape == (3, 4, 3, 2)
    np.testing.assert_array_equal(res, O.batchevaluate(g, ld, left, right, 2))
    for (i, l), c, cp, (j, r) in itertools.product(enumerate(left), range(4), range(3), enumerate(right)):
        assert res[i, c, cp, j] == T[l[0] - 1, l[1] - 1, c, cp, r[0] - 1, r[1] - 1]


@pytest.mark.parametrize("cls,kind,params", [("QuanticsExpSin", O.F_QEXPSIN, [10.0, 100.0, 1.1]),
                                             ("QuanticsExpPlusSin", O.F_QEXPPLUSSIN, [1.0, 1e-3, 1000.0]),
                                             ("QuanticsExp2", O.F_QEXP2, [1.0, 1e-4, 2.0])])
def test_quantics_functions(tci, cls, kind, params):
    R = 40
    ld = [2] * R
    rng = np.random.default_rng(9)
    left, right = _sets(rng, ld, 17, R - 19, 64, 48)
    f, g = getattr(tci, cls)(ld, *params), O.Func(kind, params)
    got, ref = f(left, right, 2), O.batchevaluate(g, ld, left, right, 2)
    assert np.max(np.abs(got - ref)) <= 1e-12 * np.max(np.abs(ref))


def test_maxabs_and_strides(tci):
    import torch
    rng = np.random.default_rng(2)
    ld = [10] * 6
    left, right = _sets(rng, ld, 2, 2, 33, 21)
    f, g = tci.LorentzFunction(ld, -2.5), O.Func(O.F_LORENTZ, [-2.5])
    ref = O.batchevaluate(g, ld, left, right, 2)
    rows = 33 * 100
    ldp = (rows + 15) // 16 * 16
    out = torch.full((21, ldp), 7.0, dtype=torch.float64, device="cuda")
    _, mx = f.evaluate_batch(left, right, 2, out=out, stride_c=33, stride_j=ldp, maxabs=0.0)
    got = out.cpu().numpy()
    np.testing.assert_array_equal(got[:, :rows].T, ref.reshape(rows, 21, order="F"))
    assert np.all(got[:, rows:] == 7.0)                 # padding untouched
    assert mx == np.max(np.abs(ref))
    _, mx2 = f.evaluate_batch(left, right, 2, maxabs=100.0)
    assert mx2 == 100.0


def test_user_functor_plugin(tci):
    src = r"""
    __device__ double tci_user_f(const int* idx, int n, const double* params, int nparams) {
        long long s = 0;
        for (int i = 0; i < n; ++i) s += (long long)idx[i] * idx[i];
        return params[0] / (double)(1 + s);
    }
    """
    rng = np.random.default_rng(4)
    ld = [10] * 8
    left, right = _sets(rng, ld, 3, 3, 50, 60)
    f, g = tci.UserFunction(src, ld, params=[1.0]), O.Func(O.F_LORENTZ, [1.0])
    np.testing.assert_array_equal(f(left, right, 2), O.batchevaluate(g, ld, left, right, 2))
    with pytest.raises(tci.TCIError, match="failed to compile"):
        tci.UserFunction("__device__ double tci_user_f(const int* idx, int n) { return oops; }", ld)


def test_full_size_pi_c3(tci):
    """C3-sized Pi (5120 x 5120 = 26.2 M evaluations): spot-check entries and the maximum."""
    rng = np.random.default_rng(3)
    ld = [20] * 20
    I = rng.integers(1, 21, size=(256, 9))
    J = rng.integers(1, 21, size=(256, 9))
    left = np.concatenate([np.repeat(I, 1, axis=0)] * 1)
    Ic = np.array([list(i) + [s] for s in range(1, 21) for i in I], dtype=np.int32)
    Jc = np.array([[s] + list(j) for j in J for s in range(1, 21)], dtype=np.int32)
    f, g = tci.LorentzFunction(ld, 1.0), O.Func(O.F_LORENTZ, [1.0])
    t, mx = f.evaluate_batch(Ic, Jc, 0, maxabs=0.0)
    A = tci.to_numpy(t)
    assert A.shape == (5120, 5120)
    for _ in range(200):
        i, j = int(rng.integers(5120)), int(rng.integers(5120))
        assert A[i, j] == g(list(Ic[i]) + list(Jc[j]))
    assert mx == A.max()


@pytest.mark.gpu
def test_hash_function_bit_exact(tci):
    """The synthetic full-rank benchmark functor: integer hash, so GPU and oracle agree bit for bit."""
    ld = [5, 4, 3, 6, 2]
    f, g = tci.HashFunction(ld, 0.75, 1.5), O.Func(O.F_HASH, [0.75, 1.5])
    rng = np.random.default_rng(3)
    left = [[int(rng.integers(1, ld[0] + 1)), int(rng.integers(1, ld[1] + 1))] for _ in range(7)]
    right = [[int(rng.integers(1, ld[3] + 1)), int(rng.integers(1, ld[4] + 1))] for _ in range(5)]
    got = f(left, right, 1)
    ref = O.batchevaluate(g, ld, left, right, 1)
    assert got.shape == ref.shape == (7, 3, 5) and np.array_equal(got, ref)
    assert np.abs(ref).max() <= 0.75 + 1.5 and len(np.unique(ref)) > 50
