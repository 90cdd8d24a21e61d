#!/usr/bin/env python3
"""Per-call host wall time of the C-ABI entry points in the latency-bound regime (Pi of a few tens of rows): what one bond
of a low-rank sweep costs, call by call.  Run under ncu with --metrics gpu__time_duration.sum for the kernel durations."""
import ctypes as C
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import tci_b200 as T
from tci_b200._lib import p_f64, p_i32, p_i64

h = T.default_handle()
L = T.lib()
dev = f"cuda:{h.device}"
REPS = int(os.environ.get("REPS", 300))


def timeit(fn, reps=REPS):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e6


rng = np.random.default_rng(0)
for m, n, r in [(34, 34, 17), (64, 64, 32), (120, 120, 12), (360, 360, 18)]:
    A0 = rng.standard_normal((m, r)) @ rng.standard_normal((r, n))
    ld = (m + 15) // 16 * 16
    src = T.Mat.from_numpy(A0, dev, pad=16)
    work = torch.empty_like(src.buf)
    Lm, Um = T.Mat.empty(m, r, dev), T.Mat.empty(r, n, dev)
    rp, cp, pe = np.zeros(m, dtype=np.int64), np.zeros(n, dtype=np.int64), np.zeros(r + 1)
    npv, err = C.c_int64(), C.c_double()

    def rrlu():
        work.copy_(src.buf)
        h.check(L.tci_rrlu_inplace_f64(h.h, C.c_void_p(work.data_ptr()), C.c_int64(m), C.c_int64(n), C.c_int64(ld), C.c_int64(r),
                                       C.c_double(1e-14), C.c_double(0.0), 1, p_i64(rp), p_i64(cp), C.byref(npv), C.byref(err),
                                       Lm.ptr, C.c_int64(Lm.ld), Um.ptr, C.c_int64(Um.ld), p_f64(pe)))

    def copy_only():
        work.copy_(src.buf)
        torch.cuda.synchronize()

    print(f"rrlu {m}x{n} r={r}: {timeit(rrlu):7.1f} us per call (copy + sync alone {timeit(copy_only):5.1f} us), npivot {npv.value}", flush=True)

for q, r in [(34, 17), (240, 12), (360, 18)]:
    P = T.Mat.from_numpy(rng.standard_normal((r, r)) + 3 * np.eye(r), dev, pad=16)
    B = T.Mat.from_numpy(rng.standard_normal((q, r)), dev, pad=16)
    X = T.Mat.empty(q, r, dev)

    def solve():
        h.check(L.tci_solve_right_f64(h.h, B.ptr, C.c_int64(q), C.c_int64(B.ld), P.ptr, C.c_int64(r), C.c_int64(P.ld), X.ptr, C.c_int64(X.ld)))

    print(f"solve_right {q}x{r}: {timeit(solve):7.1f} us per call", flush=True)

ldims = np.asarray([2] * 40, dtype=np.int32)
f = T.QuanticsExpSin(list(ldims))
left = np.ascontiguousarray(rng.integers(1, 3, size=(34, 20)), dtype=np.int32)
right = np.ascontiguousarray(rng.integers(1, 3, size=(34, 20)), dtype=np.int32)
out = torch.empty(34 * 48, dtype=torch.float64, device=dev)
acc = torch.zeros(2, dtype=torch.float64, device=dev)
mx = C.c_double(0.0)


def ev_sync():
    h.check(L.tci_batcheval_f64(h.h, f._f, p_i32(ldims), 40, p_i32(left), C.c_int64(34), 20, p_i32(right), C.c_int64(34), 20,
                                C.c_void_p(out.data_ptr()), C.c_int64(34), C.c_int64(48), C.byref(mx)))


def ev_acc():
    h.check(L.tci_batcheval_acc_f64(h.h, f._f, p_i32(ldims), 40, p_i32(left), C.c_int64(34), 20, p_i32(right), C.c_int64(34), 20,
                                    C.c_void_p(out.data_ptr()), C.c_int64(34), C.c_int64(48), C.c_void_p(acc.data_ptr())))


print(f"batcheval 34x34 with host maximum: {timeit(ev_sync):6.1f} us; with device accumulator (no sync): {timeit(ev_acc):6.1f} us", flush=True)


def sync_only():
    h.synchronize()


print(f"empty stream synchronize: {timeit(sync_only):5.1f} us", flush=True)
