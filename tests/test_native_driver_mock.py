"""CPU check of the compiled sweep driver's BOOKKEEPING (csrc/tci2_driver.cpp): the unchanged driver source is linked
against tests/native_mock/mock_cabi.cpp — host-memory stand-ins of the few C-ABI entry points it calls, backed by the
oracle's arithmetic — and driven through the product's own ctypes front end (native.py).  Index sets, ranks, errors
and site tensors must then be IDENTICAL to the oracle's TensorCI2 (same arithmetic underneath, so any difference is
a bookkeeping bug: union/kronecker order, history extras, error merging, sweep order, convergence).
The GPU parity tests of the same driver on the real kernels are in tests/test_gpu_native_driver.py."""
import ctypes as C
import importlib.util
import itertools
import os
import subprocess

import numpy as np
import pytest

from oracle import oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "tensorcrossinterpolation.jl_b200")


def _load(name):
    """Load one host-side module of the package without importing the package (no CUDA library needed)."""
    import tci_b200                                          # noqa: F401  (importable without a GPU)
    return importlib.import_module(f"tci_b200.{name}")


@pytest.fixture(scope="module")
def mock():
    O.build()
    out = os.path.join(ROOT, "tests", "_build")
    os.makedirs(out, exist_ok=True)
    so = os.path.join(out, "libtci2_mock.so")
    srcs = [os.path.join(ROOT, "tests", "native_mock", "mock_cabi.cpp"), os.path.join(PKG, "csrc", "tci2_driver.cpp")]
    if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wall", "-fPIC", "-shared", "-o", so, *srcs,
                               "-L" + os.path.join(ROOT, "oracle", "_build"), "-ltci_oracle",
                               "-Wl,-rpath," + os.path.join(ROOT, "oracle", "_build")])
    L = C.CDLL(so)
    h = C.c_void_p()
    assert L.mock_create(C.byref(h)) == 0
    yield L, h
    L.mock_destroy(h)


def _native(mock, localdims, f, pivots):
    L, h = mock
    N = _load("native").NativeTCI2
    t = N(L, h, localdims)
    t.addglobalpivots(pivots)
    t.init_maxsample(f.h, pivots)
    return t


def _same_state(t, o, cores=True):
    n = len(t)
    for b in range(1, n + 1):
        np.testing.assert_array_equal(t.Iset(b), o.Iset(b))
        np.testing.assert_array_equal(t.Jset(b), o.Jset(b))
    be, pe, mx = t.errors()
    np.testing.assert_array_equal(be, o.bonderrors)
    np.testing.assert_array_equal(pe, o.pivoterrors)
    assert mx == o.maxsamplevalue
    if cores:
        assert t.sitetensors_available()
        for b in range(1, n + 1):
            ptr, shape = t.core(b)
            size = int(np.prod(shape))
            got = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_double)), shape=(size,)).reshape(shape, order="F")
            np.testing.assert_array_equal(got, o.sitetensor(b))


CASES = [
    ("lorentz5", O.F_LORENTZ, [1.0], [10] * 5, dict(tolerance=1e-8)),
    ("lorentz8_c1", O.F_LORENTZ, [1.0], [10] * 8, dict(tolerance=1e-8)),
    ("lorentz_maxbond", O.F_LORENTZ, [1.0], [6] * 6, dict(tolerance=1e-12, maxbonddim=5)),
    ("lorentz_forward", O.F_LORENTZ, [1.0], [5] * 6, dict(tolerance=1e-9, sweepstrategy="forward")),
    ("lorentz_nested", O.F_LORENTZ, [1.0], [5] * 6, dict(tolerance=1e-9, strictlynested=True)),
    ("quantics_c4_small", O.F_QEXPSIN, [10.0, 100.0, 1.1], [2] * 16, dict(tolerance=1e-10, maxbonddim=64)),
    ("hash_fullrank", O.F_HASH, [1.0, 0.0], [4] * 6, dict(tolerance=1e-10, maxbonddim=12, maxiter=4)),
    ("mixed_dims", O.F_LORENTZ, [2.5], [3, 7, 2, 5, 4], dict(tolerance=1e-10, normalizeerror=False)),
]


@pytest.mark.parametrize("name,kind,params,localdims,kw", CASES, ids=[c[0] for c in CASES])
def test_optimize_identical_to_oracle(mock, name, kind, params, localdims, kw):
    f = O.Func(kind, params)
    piv = [[1] * len(localdims)] if kind != O.F_QEXPSIN else [[2, 1] * (len(localdims) // 2)]   # f(x = 0) = 0 there
    o, oranks, oerrors = O.crossinterpolate2(f, localdims, piv, **kw)
    t = _native(mock, localdims, f, piv)
    ranks, errors = t.optimize(f.h, nsearchglobalpivot=0, **kw)
    np.testing.assert_array_equal(ranks, oranks)
    np.testing.assert_array_equal(errors, oerrors)
    _same_state(t, o)
    assert t.stats()["nevals"] > 0


@pytest.mark.parametrize("limit", [1 << 30, 400, 0])
@pytest.mark.parametrize("name,kind,params,localdims,kw", CASES, ids=[c[0] for c in CASES])
def test_device_engine_bookkeeping_identical_to_oracle(mock, name, kind, params, localdims, kw, limit):
    """engine = device: the driver packs the index sets (and the history extras) into one arena per half-sweep and unpacks the
    new sets / pivot errors (tci_dsweep_halfsweep_f64; here the arena protocol restated on the host).  `limit` makes bonds
    with more Pi entries "not fit", so the per-bond fall-back and the continuation behind it are exercised (0: every bond
    falls back)."""
    L, h = mock
    f = O.Func(kind, params)
    piv = [[1] * len(localdims)] if kind != O.F_QEXPSIN else [[2, 1] * (len(localdims) // 2)]
    o, oranks, oerrors = O.crossinterpolate2(f, localdims, piv, **kw)
    t = _native(mock, localdims, f, piv)
    t.set_engine("device")
    L.mock_set_dsweep_limit(h, C.c_int64(limit))
    try:
        ranks, errors = t.optimize(f.h, nsearchglobalpivot=0, **kw)
    finally:
        L.mock_set_dsweep_limit(h, C.c_int64(1 << 30))
    np.testing.assert_array_equal(ranks, oranks)
    np.testing.assert_array_equal(errors, oerrors)
    _same_state(t, o)
    assert t.stats()["nevals"] == o.stats()["nevals"]


def test_stepwise_calls_identical_to_oracle(mock):
    """test/test_tensorci2.jl:247-273 driven call by call: updatepivots! per bond, addglobalpivots + makecanonical,
    sweep2site without the final fill, sweep1site in both directions, fillsitetensors."""
    localdims = [10] * 5
    f = O.Func(O.F_LORENTZ, [1.0])
    piv = [[1] * 5]
    o = O.TCI2(f, localdims, piv)
    t = _native(mock, localdims, f, piv)
    for b in range(1, 5):
        o.updatepivots(b, True, reltol=1e-8, maxbonddim=2)
        t.updatepivots(f.h, b, True, reltol=1e-8, maxbonddim=2)
    assert t.linkdims() == [2] * 4
    _same_state(t, o, cores=False)
    gp = [[2, 9, 10, 5, 7]]
    o.addglobalpivots1sitesweep(gp, reltol=1e-8)
    t.addglobalpivots(gp)
    t.makecanonical(f.h, reltol=1e-8)
    assert t.linkdims() == [3] * 4
    _same_state(t, o)
    o.sweep2site(3, iter1=1, abstol=1e-9, fillsitetensors=False)
    t.sweep2site(f.h, 3, iter1=1, abstol=1e-9, fillsitetensors=False)
    _same_state(t, o, cores=False)
    assert not t.sitetensors_available()
    o.sweep1site("backward", abstol=1e-9, updatetensors=True)
    t.sweep1site(f.h, "backward", abstol=1e-9, updatetensors=True)
    _same_state(t, o)
    o.sweep1site("forward", abstol=1e-9, updatetensors=False)
    t.sweep1site(f.h, "forward", abstol=1e-9, updatetensors=False)
    _same_state(t, o, cores=False)
    o.fillsitetensors()
    t.fillsitetensors(f.h)
    _same_state(t, o)


def test_pivoterrors_diagonal_exact(mock):                 # test/test_tensorci2.jl:27-39
    diags = [1.0, 1e-5, 0.0]
    f = O.Func(O.F_DENSE, table=np.diag(diags))
    t = _native(mock, [3, 3], f, [[1, 1]])
    t.optimize(f.h, tolerance=1e-8, nsearchglobalpivot=0)
    assert list(t.errors()[1]) == diags


def test_global_pivot_search_finds_hidden_feature(mock):
    """The line search of the default finder (globalpivotfinder.jl:205-264) on a function whose feature the two-site
    sweeps started at [1, 1, ...] cannot see: with the search the interpolation is exact on the feature, without it is not."""
    localdims = [4] * 6
    table = np.ones(localdims) * 0.5
    table[:, 3, :, :, 3, :] = 5.0                          # needs v2 = v5 = 4 together: no two-site Pi built from [1, 1, ...] holds it
    f = O.Func(O.F_DENSE, table=table)
    Nat = _load("native")
    t0 = _native(mock, localdims, f, [[1] * 6])
    t0.optimize(f.h, tolerance=1e-10, nsearchglobalpivot=0)
    assert t0.rank() == 1
    t = _native(mock, localdims, f, [[1] * 6])
    t.seed(7)
    ranks, errors = t.optimize(f.h, tolerance=1e-10, nsearchglobalpivot=40, maxnglobalpivot=5)
    assert t.rank() == 2 and ranks[-1] == 2
    cores = [np.ctypeslib.as_array(C.cast(t.core(b)[0], C.POINTER(C.c_double)),
                                   shape=(int(np.prod(t.core(b)[1])),)).reshape(t.core(b)[1], order="F") for b in range(1, 7)]
    v = np.ones((1, 1))
    for c, i in zip(cores, [1, 4, 1, 1, 4, 1]):
        v = v @ c[:, i - 1, :]
    assert abs(v[0, 0] - 5.0) < 1e-9
    assert Nat.STRATEGY["backandforth"] == 1


GLOBAL = [
    ("hidden_feature", 7, [4] * 6, dict(tolerance=1e-10, nsearchglobalpivot=40, maxnglobalpivot=5)),
    ("hidden_feature_defaults", 3, [4] * 6, dict(tolerance=1e-10, nsearchglobalpivot=5, maxnglobalpivot=5)),
    ("hidden_feature_margin", 11, [4] * 6, dict(tolerance=1e-6, nsearchglobalpivot=20, maxnglobalpivot=3, tolmarginglobalsearch=2.0)),
]


@pytest.mark.parametrize("name,seed,localdims,kw", GLOBAL, ids=[c[0] for c in GLOBAL])
def test_global_pivot_finder_identical_to_oracle(mock, name, seed, localdims, kw):
    """DefaultGlobalPivotFinder (globalpivotfinder.jl:205-264) in the driver against the checker's restatement on the shared
    seeded stream: the same starts, the same found pivots, hence identical index sets, ranks, errors and tensors."""
    table = np.ones(localdims) * 0.5
    table[:, 3, :, :, 3, :] = 5.0
    table[2, :, 1, :, :, 0] += 0.25
    f = O.Func(O.F_DENSE, table=table)
    piv = [[1] * len(localdims)]
    o, oranks, oerrors = O.crossinterpolate2(f, localdims, piv, seed=seed, **kw)
    t = _native(mock, localdims, f, piv)
    t.seed(seed)
    ranks, errors = t.optimize(f.h, **kw)
    np.testing.assert_array_equal(ranks, oranks)
    np.testing.assert_array_equal(errors, oerrors)
    _same_state(t, o)
    o0, r0, _ = O.crossinterpolate2(f, localdims, piv, tolerance=kw["tolerance"])
    assert o.rank() >= o0.rank()


def test_errors_mirror_the_reference(mock):
    L, h = mock
    Nat = _load("native")
    TCIError = _load("_lib").TCIError
    with pytest.raises(TCIError, match="at least 2 elements"):
        Nat.NativeTCI2(L, h, [3])
    f = O.Func(O.F_DENSE, table=np.zeros((3, 3)))
    t = Nat.NativeTCI2(L, h, [3, 3])
    t.addglobalpivots([[1, 1]])
    with pytest.raises(TCIError, match="maxsamplevalue is zero!"):
        t.init_maxsample(f.h, [[1, 1]])
    with pytest.raises(TCIError, match="one index per leg"):
        t.addglobalpivots([[1, 1, 1]])
    g = O.Func(O.F_LORENTZ, [1.0])
    t = _native(mock, [3, 3, 3], g, [[1, 1, 1]])
    with pytest.raises(TCIError, match="convergence criterion is not reachable"):
        t.optimize(g.h, tolerance=0.0)
    with pytest.raises(TCIError, match="nsearchglobalpivot < maxnglobalpivot"):
        t.optimize(g.h, tolerance=1e-8, nsearchglobalpivot=2, maxnglobalpivot=5)
    with pytest.raises(TCIError, match="site tensors are not available"):
        t.core(1)
    t2 = _native(mock, [3, 3, 3], g, [[1, 1, 1]])
    t2.set_indexset(0, 2, [[1], [2]])                       # |Iset[2]| = 2 but |Jset[1]| = 1: not square
    with pytest.raises(TCIError, match="Pivot matrix at bond 1 is not square!"):
        t2.fillsitetensors(g.h)


def test_optfirstpivot_identical_to_oracle(mock):          # util.jl:260-291 through tci_optfirstpivot_f64
    from tci_b200._lib import p_i32
    L, h = mock
    rng = np.random.default_rng(3)
    for kind, params, ld in [(O.F_HASH, [1.0, 0.3], [5, 3, 7, 2, 6, 4]), (O.F_QEXPSIN, [10.0, 100.0, 1.1], [2] * 20),
                             (O.F_SUM, [1.0], [3, 4, 5, 6])]:
        f = O.Func(kind, params)
        for _ in range(3):
            first = [int(rng.integers(1, d + 1)) for d in ld]
            pv = np.asarray(first, dtype=np.int32)
            lda = np.asarray(ld, dtype=np.int32)
            assert L.tci_optfirstpivot_f64(h, f.h, p_i32(lda), len(ld), p_i32(pv), 1000) == 0
            assert pv.tolist() == list(O.optfirstpivot(f, ld, first))
    assert pv.tolist() == [3, 4, 5, 6]                     # test/test_util.jl:28-36: the maximum of sum(v) is the all-d corner


EDGE = [
    ("two_sites", O.F_LORENTZ, [1.0], [7, 5], dict(tolerance=1e-10)),
    ("unit_dims", O.F_LORENTZ, [1.0], [1, 4, 1, 3, 1], dict(tolerance=1e-10)),
    ("maxiter_1", O.F_LORENTZ, [1.0], [4] * 5, dict(tolerance=1e-12, maxiter=1)),
    ("ncheck_1", O.F_LORENTZ, [1.0], [4] * 5, dict(tolerance=1e-6, ncheckhistory=1)),
    ("maxbond_1", O.F_LORENTZ, [1.0], [4] * 5, dict(tolerance=1e-12, maxbonddim=1, maxiter=3)),
    ("noglobalcheck", O.F_HASH, [1.0, 0.5], [3] * 5, dict(tolerance=1e-9, checkconvglobalpivot=False, maxiter=6)),
]


@pytest.mark.parametrize("name,kind,params,localdims,kw", EDGE, ids=[c[0] for c in EDGE])
def test_edge_configurations_identical_to_oracle(mock, name, kind, params, localdims, kw):
    f = O.Func(kind, params)
    piv = [[1] * len(localdims)]
    o, oranks, oerrors = O.crossinterpolate2(f, localdims, piv, **kw)
    t = _native(mock, localdims, f, piv)
    ranks, errors = t.optimize(f.h, nsearchglobalpivot=0, **kw)
    np.testing.assert_array_equal(ranks, oranks)
    np.testing.assert_array_equal(errors, oerrors)
    _same_state(t, o)


def test_several_initial_pivots_and_state_transfer(mock):
    """Two initial pivots (test/test_tensorci2.jl:55-102 style), then the state is copied into a fresh object through
    set_indexset / set_maxsamplevalue and one more sweep gives the same result on both."""
    localdims = [2] * 8
    f = O.Func(O.F_QEXP2, [1.0, 0.5, 3.0])
    piv = [[1] * 8, [1] + [2] * 7]
    o = O.TCI2(f, localdims, piv)
    t = _native(mock, localdims, f, piv)
    o.sweep2site(2, abstol=1e-9)
    t.sweep2site(f.h, 2, abstol=1e-9)
    _same_state(t, o)
    L, h = mock
    u = _load("native").NativeTCI2(L, h, localdims)
    for b in range(1, 9):
        u.set_indexset(0, b, t.Iset(b))
        u.set_indexset(1, b, t.Jset(b))
    u.set_maxsamplevalue(t.errors()[2])
    assert not u.sitetensors_available()
    # a fresh object has no sweep history: compare with strictly nested sweeps, where the history is not used
    t.sweep2site(f.h, 2, iter1=3, abstol=1e-9, strictlynested=True)
    u.sweep2site(f.h, 2, iter1=3, abstol=1e-9, strictlynested=True)
    for b in range(1, 9):
        np.testing.assert_array_equal(t.Iset(b), u.Iset(b))
        np.testing.assert_array_equal(t.Jset(b), u.Jset(b))
    np.testing.assert_array_equal(t.errors()[0], u.errors()[0])
    assert t.errors()[2] == u.errors()[2]


def test_nan_function_reports_like_the_reference(mock):
    """A NaN in Pi: the rrLU's "lu.L contains NaNs" / "lu.U contains NaNs" (matrixlu.jl:376-381) comes through the driver."""
    table = np.ones((3, 3, 3)); table[0, 1, 0] = np.nan       # in the first pivot's row of the first Pi
    f = O.Func(O.F_DENSE, table=table)
    TCIError = _load("_lib").TCIError
    t = _native(mock, [3, 3, 3], f, [[1, 1, 1]])
    with pytest.raises(TCIError, match="contains NaNs"):
        t.optimize(f.h, tolerance=1e-8, nsearchglobalpivot=0)


ROOK = [
    ("lorentz6", O.F_LORENTZ, [1.0], [8] * 6, dict(tolerance=1e-8), 1),
    ("lorentz_maxbond", O.F_LORENTZ, [1.0], [6] * 6, dict(tolerance=1e-12, maxbonddim=5), 7),
    ("hash_fullrank", O.F_HASH, [1.0, 0.5], [4] * 6, dict(tolerance=1e-10, maxbonddim=12, maxiter=3), 3),
    ("nested", O.F_LORENTZ, [1.0], [5] * 5, dict(tolerance=1e-9, strictlynested=True), 11),
    ("forward", O.F_LORENTZ, [1.0], [5] * 5, dict(tolerance=1e-9, sweepstrategy="forward", maxiter=6), 5),
]


@pytest.mark.parametrize("name,kind,params,localdims,kw,seed", ROOK, ids=[c[0] for c in ROOK])
def test_rook_search_identical_to_oracle(mock, name, kind, params, localdims, kw, seed):
    """pivotsearch=:rook (arrlu over SubMatrix strips, matrixlu.jl:500-569) in the driver against the checker's restatement with
    the same seeded stream: the same strips are sampled, so index sets, errors, evaluation counts and tensors are identical."""
    f = O.Func(kind, params)
    piv = [[1] * len(localdims)]
    o, oranks, oerrors = O.crossinterpolate2(f, localdims, piv, pivotsearch="rook", seed=seed, **kw)
    t = _native(mock, localdims, f, piv)
    t.set_pivotsearch("rook", seed)
    ranks, errors = t.optimize(f.h, nsearchglobalpivot=0, **kw)
    np.testing.assert_array_equal(ranks, oranks)
    np.testing.assert_array_equal(errors, oerrors)
    _same_state(t, o)
    assert t.stats()["nevals"] == o.stats()["nevals"]


def test_last_bond_tensors_survive_a_sweep_without_fill(mock):
    """sweep2site!(...; fillsitetensors=false) with strictly nested sets: every updatepivots! invalidates the tensors and sets
    those of its own two sites (tensorci2.jl:917-920), so exactly the last bond's pair is available afterwards — the driver
    skips forming the others, which nobody can observe."""
    localdims = [4] * 5
    f = O.Func(O.F_LORENTZ, [1.0])
    t = _native(mock, localdims, f, [[1] * 5])
    t.sweep2site(f.h, 1, abstol=1e-9, strictlynested=True, fillsitetensors=False)      # forward: last bond is 4 (sites 4, 5)
    TCIError = _load("_lib").TCIError
    for site in (1, 2, 3):
        with pytest.raises(TCIError, match="not available"):
            t.core(site)
    o = O.TCI2(f, localdims, [[1] * 5])
    o.sweep2site(1, abstol=1e-9, strictlynested=True, fillsitetensors=False)
    for site in (4, 5):
        ptr, shape = t.core(site)
        got = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_double)), shape=(int(np.prod(shape)),)).reshape(shape, order="F")
        np.testing.assert_array_equal(got, o.sitetensor(site))


@pytest.mark.parametrize("seed", range(24))
def test_random_configurations_identical_to_oracle(mock, seed):
    """Randomised sweep over the driver's control flow: chain length, mixed local dimensions, function family, tolerance,
    maxbonddim, sweep strategy, nesting, normalisation, history length, pivot search — everything identical to the checker."""
    rng = np.random.default_rng(1000 + seed)
    n = int(rng.integers(2, 8))
    localdims = [int(rng.integers(1, 6)) for _ in range(n)]
    if all(d == 1 for d in localdims):
        localdims[0] = 3
    kind, params = [(O.F_LORENTZ, [1.0]), (O.F_HASH, [1.0, 0.7]), (O.F_SUM, [0.25])][int(rng.integers(0, 3))]
    kw = dict(tolerance=float(10.0 ** -rng.integers(3, 13)), maxiter=int(rng.integers(1, 7)),
              sweepstrategy=["backandforth", "forward"][int(rng.integers(0, 2))], strictlynested=bool(rng.integers(0, 2)),
              normalizeerror=bool(rng.integers(0, 2)), ncheckhistory=int(rng.integers(1, 4)))
    if rng.integers(0, 2):
        kw["maxbonddim"] = int(rng.integers(1, 9))
    rook = bool(rng.integers(0, 2))
    piv = [[int(rng.integers(1, d + 1)) for d in localdims]]
    f = O.Func(kind, params)
    okw = dict(pivotsearch="rook", seed=seed) if rook else {}
    o, oranks, oerrors = O.crossinterpolate2(f, localdims, piv, **okw, **kw)
    t = _native(mock, localdims, f, piv)
    if rook:
        t.set_pivotsearch("rook", seed)
    ranks, errors = t.optimize(f.h, nsearchglobalpivot=0, **kw)
    np.testing.assert_array_equal(ranks, oranks)
    np.testing.assert_array_equal(errors, oerrors)
    _same_state(t, o)
    assert t.stats()["nevals"] == o.stats()["nevals"]


def test_evaluate_and_sum_on_the_object(mock):             # abstracttensortrain.jl:328-342, :428-441; test/test_tensortrain.jl:10-32
    localdims = [6] * 4
    f = O.Func(O.F_LORENTZ, [2.0])
    t = _native(mock, localdims, f, [[1] * 4])
    TCIError = _load("_lib").TCIError
    with pytest.raises(TCIError, match="not available"):
        t.sum()
    t.optimize(f.h, tolerance=1e-12, nsearchglobalpivot=0)
    pts = np.asarray(list(itertools.product(range(1, 7), repeat=4)), dtype=np.int32)
    ref = np.array([f(p) for p in pts])
    vals = t.evaluate(pts)
    assert np.max(np.abs(vals - ref)) < 1e-10
    assert t.sum() == pytest.approx(ref.sum(), rel=1e-10)
    o, _, _ = O.crossinterpolate2(f, localdims, [[1] * 4], tolerance=1e-12)
    np.testing.assert_array_equal(vals, O.tt_evaluate(o.sitetensors(), pts))
    assert t.sum() == O.tt_sum(o.sitetensors())


def test_bulk_export_matches_per_site_accessors(mock):
    """tci_tci2_export_indexsets / tci_tci2_export_cores (the whole result in two calls) against the per-site accessors."""
    localdims = [3, 5, 2, 4, 3]
    f = O.Func(O.F_HASH, [1.0, 0.5])
    t = _native(mock, localdims, f, [[1] * 5])
    t.optimize(f.h, tolerance=1e-9, nsearchglobalpivot=0, maxiter=4)
    I, J = t.export_indexsets()
    for b in range(1, 6):
        np.testing.assert_array_equal(I[b - 1], t.Iset(b))
        np.testing.assert_array_equal(J[b - 1], t.Jset(b))
    keep = {}

    def alloc(n):                                           # "device" memory of the mock is host memory
        keep["buf"] = np.zeros(n)
        return keep["buf"].ctypes.data, keep["buf"]
    flat, shapes, offs = t.export_cores(alloc)
    for b in range(1, 6):
        ptr, shape = t.core(b)
        assert tuple(int(v) for v in shapes[b - 1]) == shape
        size = int(np.prod(shape))
        got = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_double)), shape=(size,))
        np.testing.assert_array_equal(flat[int(offs[b - 1]):int(offs[b - 1]) + size], got)
