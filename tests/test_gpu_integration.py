"""GPU tests of `integrate` (src/integration.jl) and `CachedFunction` (src/cachedfunction.jl): the reference's own tests
(test/test_integration.jl incl. its stored 10-d value, test/test_cachedfunction.jl) plus parity against the oracle."""
import numpy as np
import pytest

from oracle import oracle as O

pytestmark = pytest.mark.gpu

POLY = [0.23637074801483304, 0.20661524945577847, 0.1850826417895819, 0.8433788714289417, 0.5801482873508491,
        0.20339438932656262, 0.21593267492457668, 0.8052490409622802, 0.7189346124875339, 0.9400806688257749,
        0.355210845205325, 0.5251561513473092, 0.6819965273401778, 0.9221987248861162, 0.04166444723413998]


@pytest.fixture(scope="module")
def tci():
    import tci_b200
    return tci_b200


def _poly(x):
    return sum(c * x ** i for i, c in enumerate(POLY))


def _polyint(x):
    return sum(c * x ** (i + 1) / (i + 1) for i, c in enumerate(POLY))


def test_kronrod_matches_quadpack(tci):
    x, w, gw = tci.kronrod(7)
    xr, wr = O.kronrod15()
    assert np.max(np.abs(x - xr)) < 5e-16 and np.max(np.abs(w - wr)) < 5e-16
    assert abs(gw.sum() - 2.0) < 1e-14 and np.count_nonzero(gw) == 7
    x21, w21, _ = tci.kronrod(10)                        # QUADPACK qk21: first node, centre weight
    assert abs(x21[0] + 0.995657163025808080735527280689003) < 5e-16
    assert abs(w21[10] - 0.149445554002916905664936468389821) < 5e-16


def test_integrate_polynomials_host_function(tci):      # test/test_integration.jl:5-29
    N = 5
    f = lambda x: float(np.prod([_poly(v) for v in x]))
    assert tci.integrate(f, [0.0] * N, [1.0] * N) == pytest.approx(_polyint(1.0) ** N, rel=1e-8)
    rng = np.random.default_rng(1234)
    a, b = rng.random(N), rng.random(N)
    exact = float(np.prod([_polyint(v) - _polyint(u) for u, v in zip(a, b)]))
    assert tci.integrate(f, a, b) == pytest.approx(exact, rel=1e-8)
    ref = O.integrate(f, a, b)
    assert tci.integrate(f, a, b, nsearchglobalpivot=0) == pytest.approx(ref, rel=1e-12)


INTEGRAND_10D = r"""
__device__ double tci_integrand(const double* x, int n, const double* params, int nparams) {
    double s = 0.0, s2 = 0.0;
    for (int i = 0; i < n; ++i) { s += x[i]; s2 += x[i] * x[i]; }
    return 1000.0 * cos(10.0 * s2) * exp(-(s * s * s * s) / 1000.0);
}
"""


def test_integrate_10d_reference_value(tci):            # test/test_integration.jl:31-41: Iref = -5.4960415218049, |I15 - Iref| < 1e-3
    I15 = tci.integrate(INTEGRAND_10D, [-1.0] * 10, [1.0] * 10, GKorder=15, tolerance=1e-8)
    assert abs(I15 - (-5.4960415218049)) < 1e-3


def test_integrate_device_functor_matches_oracle(tci):
    src = r"""
    __device__ double tci_integrand(const double* x, int n, const double* p, int np_) {
        double s = 0.0; for (int i = 0; i < n; ++i) s += x[i] * x[i];
        return p[0] / (1.0 + s);
    }"""
    a, b = [-1.0, 0.0, 0.5, -2.0], [1.0, 2.0, 1.5, 0.0]
    got = tci.integrate(src, a, b, params=[2.5], nsearchglobalpivot=0, tolerance=1e-10)
    ref = O.integrate(lambda x: 2.5 / (1.0 + sum(v * v for v in x)), a, b, tolerance=1e-10)
    assert got == pytest.approx(ref, rel=1e-10)


def test_integrate_argument_errors(tci):
    with pytest.raises(tci.TCIError, match="must be odd"):
        tci.integrate(lambda x: 1.0, [0.0], [1.0], GKorder=14)
    with pytest.raises(tci.TCIError, match="same dimensionality"):
        tci.integrate(lambda x: 1.0, [0.0, 0.0], [1.0])


# ---- CachedFunction (test/test_cachedfunction.jl) ----------------------------------------------------------------------
def test_cache_point_form(tci):                         # :47-58
    f = lambda x: 2 * (x[0] - 1) + (x[1] - 1)
    cf = tci.CachedFunction(f, [4, 2])
    assert cf.f is f
    for i in range(1, 5):
        for j in range(1, 3):
            assert cf([i, j]) == f([i, j])
            assert cf._key([i, j]) in cf.cachedata()
            assert cf([i, j]) == f([i, j])
    assert cf.ncacheddata() == 8


def test_cache_batch_form(tci):                         # :80-91
    localdims = [2, 2, 2, 2, 2]
    calls = []

    def g(x):
        calls.append(tuple(x))
        return float(sum(x))

    cf = tci.CachedFunction(g, localdims)
    left, right = [[1, 1]] * 100, [[1, 1]] * 100
    res = tci._batchevaluate_dispatch(cf, localdims, left, right, 1)
    ref = np.array([[[sum(l + [c] + r) for r in right] for c in (1, 2)] for l in left], dtype=float)
    assert res.shape == (100, 2, 100) and np.array_equal(res, ref)
    assert len(calls) == 2                               # 20000 requested points, 2 distinct: everything else came from the cache
    res2 = tci._batchevaluate_dispatch(cf, localdims, [[1, 2], [2, 1]], [[2, 2]], 1)
    assert np.array_equal(res2[:, :, 0], np.array([[8.0, 9.0], [8.0, 9.0]])) and len(calls) == 6


def test_cache_many_bits_and_key_check(tci):            # :93-101, :132-141
    N = 256
    cf = tci.CachedFunction(lambda x: 1.0, [2] * N)
    x = [1] * N
    assert cf(x) == 1.0 and cf._key(x) == 0
    assert cf._key([2] * N) == 2 ** N - 1
    with pytest.raises(tci.TCIError):
        cf([1] * (2 * N))


def test_tci_through_cached_function_equals_uncached(tci):
    n = 5
    evals = []

    def lor(x):
        evals.append(1)
        return 1.0 / (1.0 + sum(v * v for v in x))

    cf = tci.CachedFunction(lor, [6] * n)
    t, ranks, errors = tci.crossinterpolate2(cf, [6] * n, tolerance=1e-8, nsearchglobalpivot=0)
    t2, ranks2, errors2 = tci.crossinterpolate2(tci.LorentzFunction([6] * n), [6] * n, tolerance=1e-8, nsearchglobalpivot=0)
    assert list(ranks) == list(ranks2)
    for b in range(n):
        np.testing.assert_array_equal(t.Iset[b], t2.Iset[b])
    assert len(evals) == cf.ncacheddata() <= 6 ** n      # every distinct point evaluated once
    assert t.stats["nevals"] > len(evals)
