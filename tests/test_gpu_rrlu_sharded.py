"""Column-sharded rrLU (tci_rrlu_dist_*): bit-exact parity with the oracle's rrlu (and hence with the single-GPU
kernels) for any number of ranks — here with virtual ranks on one GPU; the NCCL exchange is covered by test_gpu_parallel.py."""
import numpy as np
import pytest

from oracle import oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def tci():
    import tci_b200
    return tci_b200


def _check(tci, A, nranks, **kw):
    from tci_b200.parallel import assemble_sharded_U, rrlu_sharded_virtual
    outs, blocks = rrlu_sharded_virtual(A, nranks, **kw)
    ref = O.rrlu(A, **{k: v for k, v in kw.items()})
    for o in outs:
        assert o["npivot"] == ref.npivot
        np.testing.assert_array_equal(o["rowperm"], ref.rowpermutation)
        np.testing.assert_array_equal(o["colperm"], ref.colpermutation)
        np.testing.assert_array_equal(o["pivoterrors"], ref.pivoterrors())
        assert o["error"] == ref.error
        np.testing.assert_array_equal(o["L"].numpy()[:, :ref.npivot], ref.L)
    U = assemble_sharded_U(outs, A.shape[1])
    np.testing.assert_array_equal(U, ref.U)
    return outs


@pytest.mark.parametrize("nranks", [1, 2, 3, 4])
@pytest.mark.parametrize("lo", [True, False])
def test_random_matrices(tci, nranks, lo):
    rng = np.random.default_rng(7 + nranks)
    for shape in [(33, 65), (64, 64), (257, 130), (130, 257), (9, 4)]:
        _check(tci, rng.standard_normal(shape), nranks, leftorthogonal=lo)


@pytest.mark.parametrize("nranks", [2, 4])
def test_lowrank_stop_and_maxrank(tci, nranks):
    rng = np.random.default_rng(1)
    B = rng.standard_normal((300, 12)) @ rng.standard_normal((12, 200))
    outs = _check(tci, B, nranks, reltol=1e-10)
    assert outs[0]["npivot"] == 12
    _check(tci, rng.standard_normal((200, 180)), nranks, maxrank=17)
    _check(tci, B, nranks, abstol=1e-3, reltol=0.0)


@pytest.mark.parametrize("nranks", [2, 3])
def test_ties_follow_reference_order(tci, nranks):
    """Equal |values| all over: the pivot order is decided by the (position, row) tie rule across rank boundaries."""
    A = np.ones((12, 15))
    A[::2, ::3] = -1.0
    A[5, 7] = 1.0
    _check(tci, A, nranks)
    rng = np.random.default_rng(3)
    S = rng.integers(-2, 3, size=(40, 36)).astype(float)     # many exact ties, exact arithmetic
    _check(tci, S, nranks, maxrank=10)


def test_more_ranks_than_columns(tci):
    rng = np.random.default_rng(5)
    _check(tci, rng.standard_normal((20, 3)), 4)
