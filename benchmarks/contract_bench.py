#!/usr/bin/env python3
"""BASELINE config C5: MPO x MPO contraction, 30 sites, legs 4 x 4, bond 200, FP64.

    python benchmarks/contract_bench.py [--sites 30] [--chi 200] [--cpu-sites 3]

GPU: contract_zipup(A, B; method=:LU, maxbonddim=chi, tolerance=1e-12) + sum of the result, through the public host API;
device-synchronised wall time, achieved FP64 TFLOP/s of the two GEMMs per site (2*M*N*K each).  CPU (numpy oracle: OpenBLAS
matmul on all cores + the single-threaded rrLU restatement) on a bounded sample: the first `--cpu-sites` sites of the same
trains; the GPU is timed on that sample too."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def make(n, chi, d, seed=20260005):
    rng = np.random.default_rng(seed)
    bonds = [1] + [chi] * (n - 1) + [1]
    mk = lambda: [rng.standard_normal((bonds[k], d, d, bonds[k + 1])) / np.sqrt(d * chi) for k in range(n)]
    return mk(), mk()


def zipup_flops(n, chi, d):
    bonds = [1] + [chi] * (n - 1) + [1]
    tot, lc = 0.0, 1
    for k in range(n):
        la, lb, la2, lb2 = bonds[k], bonds[k], bonds[k + 1], bonds[k + 1]
        tot += 2.0 * (lc * lb) * la * (d * d * la2)              # R x A
        tot += 2.0 * (lc * d * la2) * (lb * d) * (d * lb2)       # (RA) x B
        lc = min(chi, lc * d * d, la2 * lb2)
    return tot


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sites", type=int, default=30)
    ap.add_argument("--chi", type=int, default=200)
    ap.add_argument("--d", type=int, default=4)
    ap.add_argument("--cpu-sites", type=int, default=3)
    ap.add_argument("--tci", action="store_true", help="also time contract_TCI(maxbonddim=chi)")
    ap.add_argument("--tci-maxiter", type=int, default=2)
    args = ap.parse_args()
    import torch
    import tci_b200 as T
    a, b = make(args.sites, args.chi, args.d)
    A, B = T.TensorTrain(a), T.TensorTrain(b)
    h = T.default_handle()

    def gpu_run(A, B):
        best = None
        for _ in range(2):
            torch.cuda.synchronize(); l0 = h.launch_count(); t0 = time.perf_counter()
            tt = T.contract_zipup(A, B, tolerance=1e-12, method="LU", maxbonddim=args.chi)
            s = tt.sum()
            torch.cuda.synchronize(); dt = time.perf_counter() - t0
            if best is None or dt < best[0]:
                best = (dt, tt, s, h.launch_count() - l0)
        return best

    dt, tt, s, nl = gpu_run(A, B)
    fl = zipup_flops(args.sites, args.chi, args.d)
    # ---- per-phase DEVICE time of one more run: the two GEMMs per site (_contract) and the factorisation (_factorize -> rrLU),
    #      bracketed with CUDA events on the stream the library launches on (torch's current stream)
    from tci_b200 import contraction as CM
    ev = {"gemm": [], "factorize": []}
    orig_c, orig_f = CM._contract, CM._factorize

    def timed(kind, fn):
        def w(*a, **k):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); r = fn(*a, **k); e1.record()
            ev[kind].append((e0, e1))
            return r
        return w
    CM._contract, CM._factorize = timed("gemm", orig_c), timed("factorize", orig_f)
    try:
        torch.cuda.synchronize(); t0 = time.perf_counter()
        T.contract_zipup(A, B, tolerance=1e-12, method="LU", maxbonddim=args.chi)
        torch.cuda.synchronize(); dt_inst = time.perf_counter() - t0
    finally:
        CM._contract, CM._factorize = orig_c, orig_f
    t_gemm = sum(a.elapsed_time(b) for a, b in ev["gemm"]) * 1e-3
    t_fact = sum(a.elapsed_time(b) for a, b in ev["factorize"]) * 1e-3
    # algorithmic Schur-update bytes of the factorisations (16 B per updated element): site n factorises a
    # (lc * d * d) x (chi * chi) matrix with maxrank chi
    bonds = [1] + [args.chi] * (args.sites - 1) + [1]
    lc, upd = 1, 0.0
    for k in range(args.sites - 1):
        m_, n_ = lc * args.d * args.d, bonds[k + 1] * bonds[k + 1]
        r_ = min(args.chi, m_, n_)
        upd += sum(16.0 * (m_ - q) * (n_ - q) for q in range(1, r_ + 1))
        lc = r_
    peak_hbm = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    torch.cuda.synchronize(); t0 = time.perf_counter(); ssum = tt.sum(); torch.cuda.synchronize(); t_sum = time.perf_counter() - t0
    line = {"config": f"C5: MPO x MPO zip-up (:LU), {args.sites} sites, legs {args.d}x{args.d}, bond {args.chi}, maxbonddim={args.chi}, FP64",
            "gpu": {"wall_s": dt, "gemm_flop": fl, "gemm_tflops_over_wall": fl / dt / 1e12, "sum": s, "linkdims_max": max(tt.linkdims()),
                    "launches": nl,
                    "phases": {"instrumented_wall_s": dt_inst, "gemm_device_s": t_gemm, "gemm_tflops": fl / t_gemm / 1e12,
                               "gemm_frac_of_dgemm_peak_35.3": fl / t_gemm / 1e12 / 35.3,
                               "factorize_device_s": t_fact, "rrlu_update_GB": upd / 1e9, "rrlu_update_GBps": upd / t_fact / 1e9,
                               "rrlu_frac_of_hbm_peak": upd / t_fact / 1e9 / peak_hbm, "hbm_peak_GBps": peak_hbm,
                               "note": "factorize = rrLU of the (lc d^2) x chi^2 matrix incl. its swap / extraction kernels and the "
                                       "un-permutation of L and U; GB/s counts the Schur-update bytes only"},
                    "sum_wall_s": t_sum}}
    if args.tci:
        torch.cuda.synchronize(); t0 = time.perf_counter()
        tt2 = T.contract_TCI(A, B, tolerance=1e-10, maxbonddim=args.chi, maxiter=args.tci_maxiter, rng=np.random.default_rng(5), nsearchglobalpivot=0)
        torch.cuda.synchronize(); dt2 = time.perf_counter() - t0
        line["gpu"]["contract_TCI"] = {"wall_s": dt2, "maxbonddim": args.chi, "maxiter": args.tci_maxiter, "linkdims_max": max(tt2.linkdims()),
                                       "sum": tt2.sum(), "zipup_sum": ssum}
    if args.cpu_sites > 0:
        from oracle import contraction_oracle as CO
        k = args.cpu_sites
        asub, bsub = a[:k], b[:k]
        # close the sample trains with a random boundary vector so that the last bond is 1
        rng = np.random.default_rng(1)
        asub = asub[:-1] + [np.einsum("abcd,d->abc", asub[-1], rng.standard_normal(asub[-1].shape[3]))[..., None]]
        bsub = bsub[:-1] + [np.einsum("abcd,d->abc", bsub[-1], rng.standard_normal(bsub[-1].shape[3]))[..., None]]
        t0 = time.perf_counter()
        ref = CO.contract_zipup(asub, bsub, tolerance=1e-12, method="LU", maxbonddim=args.chi)
        cdt = time.perf_counter() - t0
        gdt, gtt, gs, _ = gpu_run(T.TensorTrain(asub), T.TensorTrain(bsub))
        got = gtt.sitetensors
        err = max(float(np.max(np.abs(x - y))) / max(1e-300, float(np.max(np.abs(y)))) for x, y in zip(got, ref))
        line["cpu"] = {"wall_s": cdt, "sample": f"first {k} sites of the same trains (closed with a random boundary vector)",
                       "cores": os.cpu_count(), "kind": "port (oracle/contraction_oracle.py: numpy matmul + oracle rrLU, 1 thread)"}
        line["gpu_same_as_cpu"] = {"wall_s": gdt}
        line["parity"] = {"max_rel_diff_site_tensors": err, "linkdims_equal": gtt.linkdims() == [c.shape[-1] for c in ref[:-1]]}
        line["speedup_wall"] = cdt / gdt
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
