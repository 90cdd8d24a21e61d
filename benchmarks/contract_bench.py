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
    line = {"config": f"C5: MPO x MPO zip-up (:LU), {args.sites} sites, legs {args.d}x{args.d}, bond {args.chi}, maxbonddim={args.chi}, FP64",
            "gpu": {"wall_s": dt, "gemm_flop": fl, "gemm_tflops_over_wall": fl / dt / 1e12, "sum": s, "linkdims_max": max(tt.linkdims()),
                    "launches": nl}}
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
