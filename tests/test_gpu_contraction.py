"""GPU parity of the contraction stack (K4c: strided copies + DMMA GEMMs + rrLU) against the numpy oracle of
src/contraction.jl — FP64, tolerance 1e-12 relative to the largest entry (north star)."""
import itertools

import numpy as np
import pytest

from oracle import contraction_oracle as CO

pytestmark = pytest.mark.gpu
RTOL = 1e-12


@pytest.fixture(scope="module")
def tci():
    import tci_b200
    return tci_b200


def _mpo(rng, bonds, d1, d2):
    return [rng.standard_normal((bonds[n], d1[n], d2[n], bonds[n + 1])) for n in range(len(d1))]


def _close(a, b, tol=RTOL):
    scale = max(1.0, float(np.max(np.abs(b))))
    assert a.shape == b.shape, (a.shape, b.shape)
    assert float(np.max(np.abs(a - b))) <= tol * scale, float(np.max(np.abs(a - b))) / scale


def test_strided_copy_and_contract(tci):                   # contraction.jl:193-215, test/test_contraction.jl:25-30
    rng = np.random.default_rng(0)
    dev = "cuda:0"
    x = rng.standard_normal((5, 33, 4, 40))
    t = tci.DTensor.from_numpy(x, dev)
    for perm in [(0, 1, 2, 3), (3, 1, 2, 0), (1, 0, 3, 2), (2, 3, 0, 1), (1, 3, 0, 2)]:
        assert np.array_equal(t.permute(perm).numpy(), np.transpose(x, perm))
    assert np.array_equal(t.select(1, 7).permute((2, 0, 1)).numpy(), np.transpose(x[:, 7], (2, 0, 1)))
    assert np.array_equal(t.keep(2, 3).numpy(), x[:, :, 3:4, :])
    a, b = rng.standard_normal((2, 3, 4)), rng.standard_normal((2, 5, 4))
    ab = tci._contract(tci.DTensor.from_numpy(a, dev), tci.DTensor.from_numpy(b, dev), (1, 3), (1, 3))
    _close(ab.numpy(), CO._contract(a, b, (1, 3), (1, 3)))
    a, b = rng.standard_normal((7, 20, 3, 9)), rng.standard_normal((3, 20, 11))
    for ia, ib in [((2,), (2,)), ((3, 2), (1, 2)), ((2, 3), (2, 1))]:
        got = tci._contract(tci.DTensor.from_numpy(a, dev), tci.DTensor.from_numpy(b, dev), ia, ib)
        _close(got.numpy(), CO._contract(a, b, ia, ib))
    with pytest.raises(tci.TCIError):
        tci._contract(tci.DTensor.from_numpy(a, dev), tci.DTensor.from_numpy(b, dev), (1, 1), (1, 2))


def test_contraction_evaluate_and_batch(tci):              # test/test_contraction.jl:73-153
    rng = np.random.default_rng(1)
    bonds = [1, 2, 3, 2, 1]
    a, b = _mpo(rng, bonds, [2] * 4, [3] * 4), _mpo(rng, bonds, [3] * 4, [2] * 4)
    ref = CO.Contraction(a, b)
    obj = tci.Contraction(tci.TensorTrain(a), tci.TensorTrain(b))
    assert obj.sitedims == ref.sitedims
    for idx in [(1, 1, 1, 1), (4, 3, 2, 1), (2, 4, 4, 3)]:
        assert abs(obj(list(idx)) - ref.evaluate(list(idx))) < 1e-12
    assert abs(obj([[2, 1], [1, 2], [2, 2], [1, 1]]) - ref.evaluate([(2, 1), (1, 2), (2, 2), (1, 1)])) < 1e-12
    _close(obj([[1], [3]], [[2], [4]], 2), ref.batchevaluate([[1], [3]], [[2], [4]], 2))
    _close(obj([[1, 2], [3, 1], [4, 4]], [[2]], 1), ref.batchevaluate([[1, 2], [3, 1], [4, 4]], [[2]], 1))
    for proj in [[[0, 0], [1, 0]], [[0, 0], [1, 1]], [[0, 1], [1, 0]]]:
        _close(obj.batchevaluate([[1]], [[1]], 2, proj).numpy(), ref.batchevaluate([[1]], [[1]], 2, proj))
    dbl = tci.Contraction(tci.TensorTrain(a), tci.TensorTrain(b), f=tci.ElementwiseMap("affine", 2.0))
    _close(dbl([[1]], [[2]], 2), 2 * ref.batchevaluate([[1]], [[2]], 2))
    assert abs(dbl([1, 2, 3, 4]) - 2 * ref.evaluate([1, 2, 3, 4])) < 1e-12


@pytest.mark.parametrize("method", ["LU", "CI"])
def test_naive_zipup_compress(tci, method):                # test/test_contraction.jl:73-106, :192-202
    rng = np.random.default_rng(2)
    bonds = [1, 2, 3, 2, 1]
    a, b = _mpo(rng, bonds, [2] * 4, [3] * 4), _mpo(rng, bonds, [3] * 4, [2] * 4)
    A, B = tci.TensorTrain(a), tci.TensorTrain(b)
    dense = CO.tomat(a) @ CO.tomat(b)
    naive = tci.contract_naive(A, B)
    assert naive.sitedims() == [[2, 2]] * 4
    for got, want in zip(naive.sitetensors, CO.contract_naive(a, b)):
        _close(got, want)
    _close(CO.tomat(tci.contract(A, B, algorithm="naive", method=method).sitetensors), dense, 1e-9)
    zp = tci.contract_zipup(A, B, tolerance=1e-13, method=method)
    ozp = CO.contract_zipup(a, b, tolerance=1e-13, method=method)
    assert [c.shape for c in zp.sitetensors] == [c.shape for c in ozp]
    _close(CO.tomat(zp.sitetensors), dense, 1e-10)
    for got, want in zip(zp.sitetensors, ozp):             # same pivots (tie-free random data) -> same tensors
        _close(got, want, 1e-9)
    naive.compress(method, tolerance=1e-13)
    _close(CO.tomat(naive.sitetensors), dense, 1e-10)
    trunc = tci.contract(A, B, algorithm="zipup", method=method, maxbonddim=2)
    assert max(trunc.linkdims()) <= 2
    with pytest.raises(tci.TCIError):
        tci.contract(A, B, algorithm="naive", f=tci.ElementwiseMap("affine", 2.0))
    with pytest.raises(tci.TCIError):
        tci.contract(A, B, algorithm="nope")


def test_contract_tci_and_mps(tci):                        # test/test_contraction.jl:73-106 (:TCI), :155-189
    rng = np.random.default_rng(3)
    bonds = [1, 2, 3, 2, 1]
    a, b = _mpo(rng, bonds, [2] * 4, [3] * 4), _mpo(rng, bonds, [3] * 4, [2] * 4)
    A, B = tci.TensorTrain(a), tci.TensorTrain(b)
    dense = CO.tomat(a) @ CO.tomat(b)
    ab = tci.contract(A, B, algorithm="TCI", tolerance=1e-12, rng=np.random.default_rng(5))
    assert ab.sitedims() == [[2, 2]] * 4
    _close(CO.tomat(ab.sitetensors), dense, 1e-9)
    ab2 = tci.contract(A, B, algorithm="TCI", tolerance=1e-12, f=tci.ElementwiseMap("affine", 2.0), rng=np.random.default_rng(5))
    _close(CO.tomat(ab2.sitetensors), 2 * dense, 1e-9)
    v = [rng.standard_normal((bonds[n], 3, bonds[n + 1])) for n in range(4)]
    a3 = _mpo(rng, bonds, [3] * 4, [3] * 4)
    av = tci.contract(tci.TensorTrain(a3), tci.TensorTrain(v), algorithm="naive")
    assert av.sitedims() == [[3]] * 4
    vm = [x.reshape(x.shape[0], 3, 1, x.shape[2]) for x in v]
    want = (CO.tomat(a3) @ CO.tomat(vm))[:, 0]
    got = CO.tomat([c.reshape(c.shape[0], 3, 1, c.shape[2]) for c in av.sitetensors])[:, 0]
    _close(got, want, 1e-11)
    va = tci.contract(tci.TensorTrain(v), tci.TensorTrain(a3), algorithm="zipup", method="LU")
    vrow = [x.reshape(x.shape[0], 1, 3, x.shape[2]) for x in v]
    want = (CO.tomat(vrow) @ CO.tomat(a3))[0, :]
    got = CO.tomat([c.reshape(c.shape[0], 1, 3, c.shape[2]) for c in va.sitetensors])[0, :]
    _close(got, want, 1e-10)


def test_zipup_medium(tci):
    """A size where the GEMMs are many tiles and the rrLU truncates: zip-up with maxbonddim vs the oracle."""
    rng = np.random.default_rng(4)
    N, chi = 5, 12
    bonds = [1] + [chi] * (N - 1) + [1]
    a = [x / np.sqrt(3 * chi) for x in _mpo(rng, bonds, [3] * N, [3] * N)]
    b = [x / np.sqrt(3 * chi) for x in _mpo(rng, bonds, [3] * N, [3] * N)]
    zp = tci.contract_zipup(tci.TensorTrain(a), tci.TensorTrain(b), tolerance=1e-12, method="LU", maxbonddim=20)
    ozp = CO.contract_zipup(a, b, tolerance=1e-12, method="LU", maxbonddim=20)
    assert zp.linkdims() == [c.shape[-1] for c in ozp[:-1]]
    _close(CO.tomat(zp.sitetensors), CO.tomat(ozp), 1e-9)
    assert abs(zp.sum() - float(np.sum(CO.tomat(ozp)))) < 1e-9 * max(1.0, abs(float(np.sum(CO.tomat(ozp)))))
