"""Device-resident half-sweeps (csrc/sweep.cu, engine="device"): every bond of a half-sweep in ONE launch.  Index sets, ranks,
errors must be IDENTICAL to the oracle's (and hence to the per-bond engines); both the single-CTA kernel and the 8-CTA
cluster kernel (forced through TCI_DSWEEP_CLUSTER) run every case."""
import os

import numpy as np
import pytest

from oracle import oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def tci():
    import tci_b200
    return tci_b200


@pytest.fixture(params=["1", "8"], ids=["one_cta", "cluster8"])
def variant(request):
    old = os.environ.get("TCI_DSWEEP_CLUSTER")
    os.environ["TCI_DSWEEP_CLUSTER"] = request.param
    yield request.param
    if old is None:
        os.environ.pop("TCI_DSWEEP_CLUSTER", None)
    else:
        os.environ["TCI_DSWEEP_CLUSTER"] = old


def _same(t, o, ranks, oranks, errors, oerrors):
    assert list(ranks) == list(oranks)
    np.testing.assert_array_equal(errors, oerrors)
    for b in range(1, len(t) + 1):
        np.testing.assert_array_equal(t.Iset[b - 1], o.Iset(b), err_msg=f"Iset[{b}]")
        np.testing.assert_array_equal(t.Jset[b - 1], o.Jset(b), err_msg=f"Jset[{b}]")
    assert t.maxsamplevalue == o.maxsamplevalue
    np.testing.assert_array_equal(t.pivoterrors, o.pivoterrors)
    np.testing.assert_array_equal(t.bonderrors, o.bonderrors)


CASES = [
    ("lorentz5", "LorentzFunction", (), O.F_LORENTZ, [1.0], [10] * 5, dict(tolerance=1e-8)),
    ("lorentz8_c1", "LorentzFunction", (), O.F_LORENTZ, [1.0], [10] * 8, dict(tolerance=1e-8)),
    ("lorentz_tight", "LorentzFunction", (), O.F_LORENTZ, [1.0], [10] * 5, dict(tolerance=1e-12)),
    ("lorentz_maxbond", "LorentzFunction", (), O.F_LORENTZ, [1.0], [6] * 6, dict(tolerance=1e-12, maxbonddim=5)),
    ("lorentz_forward", "LorentzFunction", (), O.F_LORENTZ, [1.0], [5] * 6, dict(tolerance=1e-9, sweepstrategy="forward")),
    ("lorentz_nested", "LorentzFunction", (), O.F_LORENTZ, [1.0], [5] * 6, dict(tolerance=1e-9, strictlynested=True)),
    ("lorentz_nonorm", "LorentzFunction", (2.5,), O.F_LORENTZ, [2.5], [3, 7, 2, 5, 4], dict(tolerance=1e-10, normalizeerror=False)),
    ("hash_fullrank", "HashFunction", (1.0, 0.0), O.F_HASH, [1.0, 0.0], [4] * 6, dict(tolerance=1e-10, maxbonddim=12, maxiter=4)),
    ("hash_mid", "HashFunction", (1.0, 1.0), O.F_HASH, [1.0, 1.0], [8] * 6, dict(tolerance=1e-10, maxbonddim=20, maxiter=3)),
    ("sum_rank2", "SumFunction", (0.25,), O.F_SUM, [0.25], [3, 4, 5, 6], dict(tolerance=1e-12)),
    ("two_sites", "LorentzFunction", (), O.F_LORENTZ, [1.0], [7, 5], dict(tolerance=1e-10)),
    ("unit_dims", "LorentzFunction", (), O.F_LORENTZ, [1.0], [1, 4, 1, 3, 1], dict(tolerance=1e-10)),
    ("c3_shape_short", "LorentzFunction", (), O.F_LORENTZ, [1.0], [20] * 6, dict(tolerance=1e-10, maxbonddim=256)),
]


@pytest.mark.parametrize("name,cls,cargs,kind,params,ld,kw", CASES, ids=[c[0] for c in CASES])
def test_device_engine_identical_to_oracle(tci, variant, name, cls, cargs, kind, params, ld, kw):
    f = getattr(tci, cls)(ld, *cargs)
    g = O.Func(kind, params)
    o, oranks, oerrors = O.crossinterpolate2(g, ld, **kw)
    t, ranks, errors = tci.crossinterpolate2(f, ld, nsearchglobalpivot=0, engine="device", **kw)
    _same(t, o, ranks, oranks, errors, oerrors)
    for a, r in zip(t.sitetensors, o.sitetensors()):
        assert a.shape == r.shape
        assert np.max(np.abs(a - r)) <= 1e-9 * max(np.max(np.abs(r)), 1e-300)


def test_device_engine_dense_functor_and_ties(tci, variant):
    """A table with many exactly equal entries: the tie rule of submatrixargmax (smallest column, then smallest row, in
    swapped coordinates; matrixlu.jl:46-87) decides most pivots."""
    rng = np.random.default_rng(3)
    table = np.round(rng.standard_normal((4, 5, 3, 4, 5)) * 2) / 2
    table[0, 0, 0, 0, 0] = 3.0
    g = O.Func(O.F_DENSE, table=table)
    f = tci.DenseFunction(table)
    ld = list(table.shape)
    for kw in (dict(tolerance=1e-10), dict(tolerance=1e-10, maxbonddim=3), dict(tolerance=1e-3, strictlynested=True)):
        o, oranks, oerrors = O.crossinterpolate2(g, ld, **kw)
        t, ranks, errors = tci.crossinterpolate2(f, ld, nsearchglobalpivot=0, engine="device", **kw)
        _same(t, o, ranks, oranks, errors, oerrors)


def test_device_engine_nan_reports_like_the_reference(tci, variant):
    table = np.ones((3, 3, 3)); table[0, 1, 0] = np.nan
    f = tci.DenseFunction(table)
    with pytest.raises(tci.TCIError, match="contains NaNs"):
        tci.crossinterpolate2(f, [3, 3, 3], tolerance=1e-8, nsearchglobalpivot=0, engine="device")


def test_device_engine_global_pivots(tci, variant):
    ld = [4] * 6
    table = np.full(ld, 0.5)
    table[:, 3, :, :, 3, :] = 5.0
    table[2, :, 1, :, :, 0] += 0.25
    g = O.Func(O.F_DENSE, table=table)
    o, oranks, oerrors = O.crossinterpolate2(g, ld, tolerance=1e-10, nsearchglobalpivot=40, maxnglobalpivot=5, seed=7)
    t, ranks, errors = tci.crossinterpolate2(tci.DenseFunction(table), ld, tolerance=1e-10, nsearchglobalpivot=40, maxnglobalpivot=5,
                                             seed=7, engine="device")
    assert list(ranks) == list(oranks)
    for b in range(1, 7):
        np.testing.assert_array_equal(t.Iset[b - 1], o.Iset(b))
        np.testing.assert_array_equal(t.Jset[b - 1], o.Jset(b))


def test_device_engine_uses_few_launches(tci):
    """The point of the engine: the quantics R = 40 run (BASELINE C4) takes tens of launches for its two-site sweeps instead of
    thousands."""
    ld = [2] * 40
    f = tci.QuanticsExpSin(ld)
    first = tci.optfirstpivot(f, ld)
    h = tci.default_handle()
    n = {}
    for engine in ("native", "device"):
        l0 = h.launch_count()
        t, ranks, errors = tci.crossinterpolate2(f, ld, [list(first)], tolerance=1e-10, maxbonddim=512, nsearchglobalpivot=0, engine=engine)
        n[engine] = h.launch_count() - l0
    assert n["device"] < n["native"] / 2
