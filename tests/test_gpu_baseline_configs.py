"""BASELINE.json's configurations C1-C5 AT THEIR STATED SIZES on the GPU against the oracle (oracle/tci_oracle.cpp, the CPU
restatement of the reference; no Julia exists on either box, so "identical to the reference" below always means identical
to that restatement).  Bar (north star): pivot index sets and ranks bit-exact, tensor entries within 1e-12 relative, errors
equal.

  C1  8-d Lorentzian, d = 10, tol 1e-8 — with the reference's DEFAULT kwargs (nsearchglobalpivot = 5) on a shared seeded stream
  C2  rrlu of the 8192 x 8192 rank-512 bench matrix, maxrank 512: permutations, pivot errors, L, U bit for bit (all 512 pivots)
  C3  20-d Lorentzian, d = 20, maxbonddim 256, tol 1e-10; C3H the same shape with the full-rank hash functor (bounded sample)
  C4  quantics R = 40, both functions of SURVEY 8d.  Transcendental functors: CUDA libdevice and libm differ in the last ulp,
      so the oracle is FED the device-evaluated values (hard part 1) — every decision it takes is then comparable
      bit for bit — and the values themselves are compared with libm separately at 1e-12.
  C5  MPO x MPO zip-up (:LU), legs 4 x 4, bond 200: a 3-site sample of the 30-site trains against the numpy oracle."""
import os

import numpy as np
import pytest

from oracle import oracle as O

pytestmark = pytest.mark.gpu

ENGINES = ["python", "native", "device"]


@pytest.fixture(scope="module")
def tci():
    import tci_b200
    return tci_b200


def _engines(tci):
    return [e for e in ENGINES if e in getattr(tci, "ENGINES", ("python", "native"))]


def _same_sets(t, o):
    for b in range(1, len(t) + 1):
        np.testing.assert_array_equal(t.Iset[b - 1], o.Iset(b), err_msg=f"Iset[{b}]")
        np.testing.assert_array_equal(t.Jset[b - 1], o.Jset(b), err_msg=f"Jset[{b}]")


def _same_tensors(t, o, tol):
    for b, (a, r) in enumerate(zip(t.sitetensors, o.sitetensors())):
        assert a.shape == r.shape, (b, a.shape, r.shape)
        if r.size:
            assert np.max(np.abs(a - r)) <= tol * max(np.max(np.abs(r)), 1e-300), (b, np.max(np.abs(a - r)))


def _device_fed(f):
    """Oracle functor whose values ARE the device functor's (BatchEvaluator protocol, batcheval.jl:196-214)."""
    return O.Func(O.F_BATCHCB, callback=lambda left, right, M: f(left, right, M))


# ---- C1 --------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("seed", [1, 2026])
def test_c1_default_kwargs_global_pivots(tci, seed):
    """README call (README.md:21-29) with the reference's default kwargs: nsearchglobalpivot = 5, maxnglobalpivot = 5,
    tolmarginglobalsearch = 10 (tensorci2.jl:1032-1034).  The finder's starts come from the seeded stream shared with the
    oracle (globalpivotfinder.jl:219 draws from Julia's RNG)."""
    ld = [10] * 8
    g = O.Func(O.F_LORENTZ, [1.0])
    o, oranks, oerrors = O.crossinterpolate2(g, ld, tolerance=1e-8, nsearchglobalpivot=5, seed=seed)
    for engine in _engines(tci):
        f = tci.LorentzFunction(ld)
        t, ranks, errors = tci.crossinterpolate2(f, ld, tolerance=1e-8, seed=seed, engine=engine)     # defaults otherwise
        assert list(ranks) == list(oranks), engine
        np.testing.assert_array_equal(errors, oerrors)
        _same_sets(t, o)
        _same_tensors(t, o, 1e-10)


def test_c1_global_pivots_found_identically(tci):
    """A function whose feature only the global search finds: the found pivots (hence index sets, ranks, errors) equal the
    oracle's for every engine."""
    ld = [4] * 6
    table = np.full(ld, 0.5)
    table[:, 3, :, :, 3, :] = 5.0
    table[2, :, 1, :, :, 0] += 0.25
    g = O.Func(O.F_DENSE, table=table)
    o, oranks, oerrors = O.crossinterpolate2(g, ld, tolerance=1e-10, nsearchglobalpivot=40, maxnglobalpivot=5, seed=7)
    assert o.rank() > 1
    for engine in _engines(tci):
        f = tci.DenseFunction(table)
        t, ranks, errors = tci.crossinterpolate2(f, ld, tolerance=1e-10, nsearchglobalpivot=40, maxnglobalpivot=5, seed=7, engine=engine)
        assert list(ranks) == list(oranks), engine
        _same_sets(t, o)
        np.testing.assert_allclose(errors, oerrors, rtol=1e-9, atol=1e-15)


# ---- C2 --------------------------------------------------------------------------------------------------------------
def test_c2_full_size_bit_exact(tci):
    """rrlu(A; maxrank=512) of the 8192 x 8192 rank-512 bench matrix (bench.py's C2): all 512 pivots against the oracle's
    loop (thread-parallel execution of the same loop; checked against the scalar loop in tests/test_oracle_golden.py)."""
    import torch
    import bench
    n, r = 8192, 512
    X, Y, sig = bench.make_factors(n, r)
    dev = "cuda:0"
    A = ((torch.from_numpy(X * sig).to(dev)) @ torch.from_numpy(Y).to(dev).T).cpu().numpy()
    A = np.asfortranarray(A)
    lu = tci.rrlu(A, maxrank=r)
    ref = O.rrlu(A, maxrank=r, threads=min(O.max_threads(), os.cpu_count() or 1))
    assert lu.npivot == ref.npivot == r
    np.testing.assert_array_equal(lu.rowpermutation, ref.rowpermutation)
    np.testing.assert_array_equal(lu.colpermutation, ref.colpermutation)
    np.testing.assert_array_equal(lu.pivoterrors(), ref.pivoterrors())
    np.testing.assert_array_equal(lu.L, ref.L)
    np.testing.assert_array_equal(lu.U, ref.U)


# ---- C3 / C3H --------------------------------------------------------------------------------------------------------
def test_c3_full_size(tci):
    ld = [20] * 20
    kw = dict(tolerance=1e-10, maxbonddim=256)
    g = O.Func(O.F_LORENTZ, [1.0])
    o, oranks, oerrors = O.crossinterpolate2(g, ld, **kw)
    for engine in _engines(tci):
        f = tci.LorentzFunction(ld)
        t, ranks, errors = tci.crossinterpolate2(f, ld, nsearchglobalpivot=0, engine=engine, **kw)
        assert list(ranks) == list(oranks), engine
        np.testing.assert_array_equal(errors, oerrors)
        _same_sets(t, o)
        _same_tensors(t, o, 1e-9)
        assert t.maxsamplevalue == o.maxsamplevalue
        np.testing.assert_array_equal(t.pivoterrors, o.pivoterrors)


def test_c3h_sample(tci):
    """The shape of C3 with the synthetic full-rank functor (every bond saturates maxbonddim): maxbonddim 64, one iteration
    (the bounded sample bench.py's sweep object times on the CPU too)."""
    ld = [20] * 10
    kw = dict(tolerance=1e-10, maxbonddim=64, maxiter=1)
    g = O.Func(O.F_HASH, [1.0, 1.0])
    o, oranks, oerrors = O.crossinterpolate2(g, ld, **kw)
    for engine in _engines(tci):
        f = tci.HashFunction(ld, 1.0, 1.0)
        t, ranks, errors = tci.crossinterpolate2(f, ld, nsearchglobalpivot=0, engine=engine, **kw)
        assert list(ranks) == list(oranks) == [64], engine
        np.testing.assert_array_equal(errors, oerrors)
        _same_sets(t, o)


# ---- C4 --------------------------------------------------------------------------------------------------------------
C4 = [("expsin", "QuanticsExpSin", O.F_QEXPSIN, [10.0, 100.0, 1.1]),            # test/test_tensorci2.jl:437
      ("expplussin", "QuanticsExpPlusSin", O.F_QEXPPLUSSIN, [1.0, 1e-3, 1000.0])]   # test/test_globalsearch.jl:14


@pytest.mark.parametrize("name,cls,kind,params", C4, ids=[c[0] for c in C4])
def test_c4_full_size_device_fed_oracle(tci, name, cls, kind, params):
    R = 40
    ld = [2] * R
    kw = dict(tolerance=1e-10, maxbonddim=512)
    f = getattr(tci, cls)(ld)
    fed = _device_fed(f)
    first = tci.optfirstpivot(f, ld)
    assert list(first) == list(O.optfirstpivot(fed, ld))
    o, oranks, oerrors = O.crossinterpolate2(fed, ld, [list(first)], **kw)
    for engine in _engines(tci):
        t, ranks, errors = tci.crossinterpolate2(f, ld, [list(first)], nsearchglobalpivot=0, engine=engine, **kw)
        assert list(ranks) == list(oranks), engine
        np.testing.assert_array_equal(errors, oerrors)
        _same_sets(t, o)
        _same_tensors(t, o, 1e-9)
        assert t.maxsamplevalue == o.maxsamplevalue
    # the function values themselves: libdevice against libm
    rng = np.random.default_rng(40)
    pts = rng.integers(1, 3, size=(4000, R)).astype(np.int32)
    got = f.evaluate_points(pts)
    g = O.Func(kind, params)
    ref = np.array([g(p) for p in pts])
    assert np.max(np.abs(got - ref)) <= 1e-12 * np.max(np.abs(ref))
    # and the interpolation error on random points within the requested tolerance (normalised like the reference's error)
    assert np.max(np.abs(t.evaluate_batch(pts) - ref)) <= 100 * 1e-10 * t.maxsamplevalue


# ---- C5 --------------------------------------------------------------------------------------------------------------
def test_c5_zipup_three_site_sample(tci):
    """contract_zipup(A, B; method=:LU, maxbonddim=200, tolerance=1e-12) (contraction.jl:751-788) on the first three sites of
    the C5 trains (legs 4 x 4, bond 200; closed with a random boundary vector so the last bond is 1): link dimensions equal,
    site tensors to 1e-12 relative, sum equal."""
    from oracle import contraction_oracle as CO
    from benchmarks import contract_bench as CB
    a, b = CB.make(30, 200, 4)
    k = 3
    rng = np.random.default_rng(1)
    asub = a[:k - 1] + [np.einsum("abcd,d->abc", a[k - 1], rng.standard_normal(a[k - 1].shape[3]))[..., None]]
    bsub = b[:k - 1] + [np.einsum("abcd,d->abc", b[k - 1], rng.standard_normal(b[k - 1].shape[3]))[..., None]]
    ref = CO.contract_zipup(asub, bsub, tolerance=1e-12, method="LU", maxbonddim=200)
    tt = tci.contract_zipup(tci.TensorTrain(asub), tci.TensorTrain(bsub), tolerance=1e-12, method="LU", maxbonddim=200)
    assert tt.linkdims() == [c.shape[-1] for c in ref[:-1]]
    for x, y in zip(tt.sitetensors, ref):
        assert x.shape == y.shape
        assert np.max(np.abs(x - y)) <= 1e-12 * np.max(np.abs(y))
    assert tt.sum() == pytest.approx(O.tt_sum([c.reshape(c.shape[0], -1, c.shape[-1], order="F") for c in ref]), rel=1e-11)
