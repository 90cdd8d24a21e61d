"""tci_batcheval_noalloc_f64 (include/tci_b200.h): the batch evaluation with its scratch in a file-scope `static __device__` array
instead of the handle's cudaMalloc'ed workspaces.  No counterpart in the reference (added
for callers that cannot call cudaMalloc).  Checked through the C-ABI against tci_batcheval_acc_f64 on the same inputs — values
and the merged maximum BIT for bit, the bar of tests/test_gpu_native_driver.py::test_device_max_accumulator — on several calls
in a row that share the array (large after small after large: the array keeps the previous call's contents), plus the size
limit and that device memory does not move.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

LIMIT = 4 << 20                                            # TCI_BATCHEVAL_NOALLOC_SCRATCH_BYTES


@pytest.fixture(scope="module")
def tci():
    import tci_b200
    return tci_b200


def _call(L, name, h, f, ld, left, right, out, acc):
    from tci_b200._lib import p_i32
    nl = left.shape[1] if left.ndim == 2 else 0
    nr = right.shape[1] if right.ndim == 2 else 0
    return getattr(L, name)(h.h, f._f, p_i32(ld), C.c_int(len(ld)), p_i32(left), C.c_int64(left.shape[0]), C.c_int(nl),
                            p_i32(right), C.c_int64(right.shape[0]), C.c_int(nr), C.c_void_p(out.data_ptr()), C.c_int64(0), C.c_int64(0),
                            C.c_void_p(acc.data_ptr()))


def _sets(rng, ld, nl, nr, nleft, nright):
    left = np.ascontiguousarray(np.stack([rng.integers(1, ld[p] + 1, size=nleft) for p in range(nl)], axis=1), dtype=np.int32)
    right = np.ascontiguousarray(np.stack([rng.integers(1, ld[len(ld) - nr + p] + 1, size=nright) for p in range(nr)], axis=1), dtype=np.int32)
    return left, right


def _cases(tci):
    """(functor, localdims, nl, nr, nleft, nright): small (plain assembly kernel), bulk (vectorised kernel with per-pair states),
    a transcendental functor, a table functor, then the bulk shape again with other rows after the small ones."""
    ld8 = [10] * 8
    q = [2] * 24
    table = np.random.default_rng(5).standard_normal((4, 3, 5, 2, 3))
    return [(tci.LorentzFunction(ld8), ld8, 3, 3, 7, 9),
            (tci.LorentzFunction(ld8), ld8, 3, 3, 700, 40),                  # Pi 7000 x 400: bulk
            (tci.QuanticsExpSin(q), q, 11, 11, 600, 37),                     # bulk, libdevice transcendental
            (tci.DenseFunction(table), [4, 3, 5, 2, 3], 2, 1, 12, 3),
            (tci.LorentzFunction(ld8, 3.0), ld8, 2, 4, 5, 6),
            (tci.HashFunction(ld8, 1.0, 1.0), ld8, 3, 3, 1030, 33)]


def test_noalloc_matches_allocating_entry_point_on_calls_in_a_row(tci):
    import torch
    L = tci.lib()
    rng = np.random.default_rng(20260002)
    cases = _cases(tci)
    h = cases[0][0].handle
    dev = f"cuda:{h.device}"
    work = []
    for f, ld, nl, nr, nleft, nright in cases:                                 # every buffer exists before the first call
        assert f.handle is h or f.handle.device == h.device
        ld = np.asarray(ld, dtype=np.int32)
        left, right = _sets(rng, ld, nl, nr, nleft, nright)
        n = nleft * int(np.prod(ld[nl:len(ld) - nr])) * nright
        work.append((f, ld, left, right, torch.full((n,), -7.0, dtype=torch.float64, device=dev),
                     torch.full((n,), -9.0, dtype=torch.float64, device=dev)))
    acc_a = torch.zeros(2, dtype=torch.float64, device=dev)
    acc_b = torch.zeros(2, dtype=torch.float64, device=dev)
    for acc in (acc_a, acc_b):
        f0 = work[0][0]
        f0.handle.check(L.tci_maxacc_set(f0.handle.h, C.c_void_p(acc.data_ptr()), C.c_double(1e-9)))
    import gc
    for rounds in range(2):                                                     # the whole list twice: 12 calls share the array
        gc.collect()                                                            # objects of earlier tests release their device memory now, not mid-loop
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info(h.device)[0]
        for k, (f, ld, left, right, a, b) in enumerate(work):
            f.handle.check(_call(L, "tci_batcheval_noalloc_f64", f.handle, f, ld, left, right, a, acc_a))
            torch.cuda.synchronize()
            if rounds == 1:                                                     # round 0 loads the kernels' code lazily; round 1 must not take memory
                assert torch.cuda.mem_get_info(h.device)[0] >= free0, "free device memory shrank across an allocation-free call"
        for k, (f, ld, left, right, a, b) in enumerate(work):
            f.handle.check(_call(L, "tci_batcheval_acc_f64", f.handle, f, ld, left, right, b, acc_b))
            torch.cuda.synchronize()
            assert torch.equal(a, b), (rounds, k)
            a.fill_(-7.0); b.fill_(-9.0)
    got_a, got_b = C.c_double(), C.c_double()
    h.check(L.tci_maxacc_get(h.h, C.c_void_p(acc_a.data_ptr()), C.byref(got_a)))
    h.check(L.tci_maxacc_get(h.h, C.c_void_p(acc_b.data_ptr()), C.byref(got_b)))
    assert got_a.value == got_b.value and got_a.value > 1e-9


def test_noalloc_back_to_back_without_host_sync(tci):
    """Two calls enqueued back to back on one stream (no synchronisation between them): stream order keeps the second call's
    uploads behind the first call's kernels, so both results are right."""
    import torch
    L = tci.lib()
    rng = np.random.default_rng(3)
    ld = np.asarray([10] * 8, dtype=np.int32)
    f = tci.LorentzFunction(list(ld))
    dev = f"cuda:{f.handle.device}"
    l1, r1 = _sets(rng, ld, 3, 3, 900, 25)
    l2, r2 = _sets(rng, ld, 3, 3, 640, 31)
    a1 = torch.empty(900 * 100 * 25, dtype=torch.float64, device=dev); b1 = torch.empty_like(a1)
    a2 = torch.empty(640 * 100 * 31, dtype=torch.float64, device=dev); b2 = torch.empty_like(a2)
    acc = torch.zeros(2, dtype=torch.float64, device=dev)
    f.handle.check(_call(L, "tci_batcheval_noalloc_f64", f.handle, f, ld, l1, r1, a1, acc))
    f.handle.check(_call(L, "tci_batcheval_noalloc_f64", f.handle, f, ld, l2, r2, a2, acc))
    f.handle.check(_call(L, "tci_batcheval_acc_f64", f.handle, f, ld, l1, r1, b1, acc))
    f.handle.check(_call(L, "tci_batcheval_acc_f64", f.handle, f, ld, l2, r2, b2, acc))
    torch.cuda.synchronize()
    assert torch.equal(a1, b1) and torch.equal(a2, b2)
    pts = np.concatenate([l1[:5], np.tile([[2, 3]], (5, 1)), r1[:5]], axis=1)          # spot check against the formula
    ref = 1.0 / (1.0 + np.sum(pts.astype(np.float64) ** 2, axis=1))
    got = np.array([a1.view(25, 10, 10, 900)[j, 2, 1, i].item() for i, j in zip(range(5), range(5))])
    np.testing.assert_allclose(got, ref, rtol=1e-14)


def test_noalloc_rejects_sizes_beyond_the_static_array(tci):
    import torch
    L = tci.lib()
    ld = np.asarray([10] * 8, dtype=np.int32)
    f = tci.LorentzFunction(list(ld))
    dev = f"cuda:{f.handle.device}"
    nleft = 200000                                                            # 4 * 4 * nleft + 16 * nleft = 6.4 MB > 4 MiB
    assert 16 * nleft + 16 * nleft > LIMIT
    left = np.ones((nleft, 4), dtype=np.int32)
    right = np.ones((2, 4), dtype=np.int32)
    out = torch.zeros(16, dtype=torch.float64, device=dev)
    acc = torch.zeros(2, dtype=torch.float64, device=dev)
    with pytest.raises(tci.TCIError, match="static array holds 4194304"):
        f.handle.check(_call(L, "tci_batcheval_noalloc_f64", f.handle, f, ld, left, right, out, acc))
    torch.cuda.synchronize()
    assert float(out.abs().max()) == 0.0                                       # nothing was launched
    with pytest.raises(tci.TCIError, match="NULL accumulator"):
        from tci_b200._lib import p_i32
        f.handle.check(L.tci_batcheval_noalloc_f64(f.handle.h, f._f, p_i32(ld), C.c_int(8), p_i32(right), C.c_int64(2), C.c_int(4),
                                                   p_i32(right), C.c_int64(2), C.c_int(4), C.c_void_p(out.data_ptr()), C.c_int64(0),
                                                   C.c_int64(0), C.c_void_p(0)))
    # the largest request that fits still works (index tables + states only; no room for per-pair states -> plain kernel)
    nleft = (LIMIT - 4096) // 32
    left = np.ascontiguousarray(np.random.default_rng(1).integers(1, 11, size=(nleft, 4)), dtype=np.int32)
    a = torch.empty(nleft * 2, dtype=torch.float64, device=dev); b = torch.empty_like(a)
    f.handle.check(_call(L, "tci_batcheval_noalloc_f64", f.handle, f, ld, left, right, a, acc))
    f.handle.check(_call(L, "tci_batcheval_acc_f64", f.handle, f, ld, left, right, b, acc))
    torch.cuda.synchronize()
    assert torch.equal(a, b)
