#!/usr/bin/env python3
"""TCI2 sweep time and f-evals/s on BASELINE configs C1, C3, C4, GPU path vs the CPU oracle.

    python benchmarks/sweep_bench.py [--configs c1,c3,c4] [--cpu] [--c3-cpu-maxbonddim 64] [--engine python|native|both]

One JSON line per config.  `gpu` = crossinterpolate2 through the public host API (tci_b200), wall time with a device
synchronize on both sides, plus the per-phase split the reference prints at verbosity > 2 (Pi fill vs LU,
src/tensorci2.jl:864-867).  `cpu` (with --cpu) = the oracle port of the reference algorithm on the host cores for the
same call; for C3 the CPU leg is bounded by a smaller maxbonddim (stated in the line) and the GPU runs that reduced
configuration too (`gpu_same_as_cpu`), so the two are comparable.  Parity of ranks / pivot sets between the two legs is
checked and reported (`parity`).  --engine native runs the same call with the compiled sweep driver (engine="native":
csrc/tci2_driver.cpp, one C call for the whole optimize!); --engine both reports it as `gpu_native` next to `gpu`, with
`native_identical_to_python` = pivot sets of the two runs equal."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


ENGINE = "python"


def run_gpu(make_f, localdims, **kw):
    if ENGINE == "both":
        return run_gpu_engine(make_f, localdims, "python", **kw)
    return run_gpu_engine(make_f, localdims, ENGINE, **kw)


def add_native(line, make_f, localdims, gt, **kw):
    """--engine both: the native-driver run of the same call beside the Python-driven one."""
    if ENGINE != "both":
        return
    g, nt = run_gpu_engine(make_f, localdims, "native", **kw)
    line["gpu_native"] = g
    line["native_identical_to_python"] = bool(all(np.array_equal(gt.Iset[b], nt.Iset[b]) and np.array_equal(gt.Jset[b], nt.Jset[b])
                                                  for b in range(len(gt.localdims))))
    line["native_speedup_over_python"] = line["gpu"]["wall_s"] / g["wall_s"]
    if "cpu" in line:
        line["native_speedup_over_cpu"] = line["cpu"]["wall_s"] / g["wall_s"]


def run_gpu_engine(make_f, localdims, engine, **kw):
    import torch
    import tci_b200 as T
    f = make_f(T)
    kw = dict(kw, engine=engine)
    h = T.default_handle()
    best = None
    for rep in range(kw.pop("reps", 2)):                 # first rep warms the allocator / NVRTC / workspaces
        torch.cuda.synchronize()
        l0 = h.launch_count()
        t0 = time.perf_counter()
        if kw.get("optfirst"):
            kw2 = {k: v for k, v in kw.items() if k != "optfirst"}
            tci, ranks, errors = T.crossinterpolate2(f, localdims, [T.optfirstpivot(f, localdims)], nsearchglobalpivot=0, **kw2)
        else:
            tci, ranks, errors = T.crossinterpolate2(f, localdims, nsearchglobalpivot=0, **kw)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        res = {"wall_s": dt, "ranks": [int(r) for r in ranks], "final_error": float(errors[-1]), "niter": len(ranks),
               "nevals": int(tci.stats["nevals"]), "evals_per_s": tci.stats["nevals"] / dt,
               "t_eval_s": tci.stats["t_eval"], "t_lu_s": tci.stats["t_lu"], "t_factors_s": tci.stats["t_factors"],
               "t_host_s": tci.stats["t_host"], "launches": h.launch_count() - l0,
               "half_sweeps": 2 * len(ranks), "s_per_half_sweep": dt / (2 * len(ranks)), "linkdims": tci.linkdims()}
        if best is None or dt < best[0]["wall_s"]:
            best = (res, tci)
    return best


def run_cpu(kind, params, localdims, **kw):
    from oracle import oracle as O
    f = O.Func(kind, params)
    t0 = time.perf_counter()
    piv = [O.optfirstpivot(f, localdims)] if kw.pop("optfirst", False) else None
    tci, ranks, errors = O.crossinterpolate2(f, localdims, piv, **kw)
    dt = time.perf_counter() - t0
    st = tci.stats()
    return {"wall_s": dt, "ranks": [int(r) for r in ranks], "final_error": float(errors[-1]), "niter": len(ranks),
            "nevals": int(st["nevals"]), "evals_per_s": st["nevals"] / dt, "t_eval_s": st["t_eval"], "t_lu_s": st["t_lu"],
            "cores": 1, "kind": "port (oracle/tci_oracle.cpp)", "linkdims": tci.linkdims()}, tci


def parity(gt, ot):
    same = all(np.array_equal(gt.Iset[b], ot.Iset(b + 1)) and np.array_equal(gt.Jset[b], ot.Jset(b + 1))
               for b in range(len(gt.localdims)))
    return {"pivot_sets_identical": bool(same)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="c1,c3,c3h,c4")
    ap.add_argument("--cpu", action="store_true")
    ap.add_argument("--c3-cpu-maxbonddim", type=int, default=64)
    ap.add_argument("--c3-maxiter", type=int, default=20)
    ap.add_argument("--c3h-maxiter", type=int, default=2)
    ap.add_argument("--engine", default="python", choices=["python", "native", "both"])
    args = ap.parse_args()
    global ENGINE
    ENGINE = args.engine
    from oracle import oracle as O
    for c in args.configs.split(","):
        if c == "c1":
            ld, kw = [10] * 8, dict(tolerance=1e-8)
            g, gt = run_gpu(lambda T: T.LorentzFunction(ld), ld, **kw)
            line = {"config": "C1: 8-d Lorentzian, localdims=fill(10,8), tolerance=1e-8, nsearchglobalpivot=0", "gpu": g}
            if args.cpu:
                o, ot = run_cpu(O.F_LORENTZ, [1.0], ld, **kw)
                line["cpu"] = o; line["parity"] = parity(gt, ot); line["speedup_wall"] = o["wall_s"] / g["wall_s"]
            add_native(line, lambda T: T.LorentzFunction(ld), ld, gt, **kw)
        elif c == "c3":
            ld, kw = [20] * 20, dict(tolerance=1e-10, maxbonddim=256, maxiter=args.c3_maxiter)
            g, gt = run_gpu(lambda T: T.LorentzFunction(ld), ld, reps=1, **kw)
            line = {"config": f"C3: 20-d Lorentzian, localdims=fill(20,20), maxbonddim=256, tolerance=1e-10, maxiter={args.c3_maxiter}", "gpu": g}
            if args.cpu:
                o, ot = run_cpu(O.F_LORENTZ, [1.0], ld, **kw)
                line["cpu"] = o; line["parity"] = parity(gt, ot); line["speedup_wall"] = o["wall_s"] / g["wall_s"]
            add_native(line, lambda T: T.LorentzFunction(ld), ld, gt, reps=2, **kw)
        elif c == "c3h":
            # the Lorentzian of C3 converges at rank 18, far below maxbonddim=256; this variant (same shape, synthetic
            # full-rank function) is the one that reaches the 5120 x 5120 Pi matrices SURVEY.md 8a sized C3 for
            ld, kw = [20] * 10, dict(tolerance=1e-10, maxbonddim=256, maxiter=args.c3h_maxiter)
            g, gt = run_gpu(lambda T: T.HashFunction(ld, 1.0, 1.0), ld, reps=2, **kw)
            line = {"config": f"C3H: full-rank hash function, localdims=fill(20,10), maxbonddim=256, maxiter={args.c3h_maxiter}", "gpu": g}
            if args.cpu:
                kw2 = dict(tolerance=1e-10, maxbonddim=args.c3_cpu_maxbonddim, maxiter=1)
                o, ot = run_cpu(O.F_HASH, [1.0, 1.0], ld, **kw2)
                g2, gt2 = run_gpu(lambda T: T.HashFunction(ld, 1.0, 1.0), ld, **kw2)
                line["cpu"] = o; line["cpu"]["sample"] = f"same function with maxbonddim={args.c3_cpu_maxbonddim}, maxiter=1"
                line["gpu_same_as_cpu"] = g2; line["parity"] = parity(gt2, ot); line["speedup_wall"] = o["wall_s"] / g2["wall_s"]
            if ENGINE == "both":
                cpu = line.pop("cpu", None)                      # the CPU leg ran a reduced configuration: no native/CPU ratio here
                add_native(line, lambda T: T.HashFunction(ld, 1.0, 1.0), ld, gt, reps=2, **kw)
                if cpu is not None:
                    line["cpu"] = cpu
        elif c == "c4":
            ld, kw = [2] * 40, dict(tolerance=1e-10, maxbonddim=512, optfirst=True)
            g, gt = run_gpu(lambda T: T.QuanticsExpSin(ld), ld, **kw)
            line = {"config": "C4: quantics R=40, f=exp(-10x) sin(2 pi 100 x^1.1), maxbonddim=512, tolerance=1e-10", "gpu": g}
            if args.cpu:
                o, ot = run_cpu(O.F_QEXPSIN, [10.0, 100.0, 1.1], ld, **dict(kw))
                line["cpu"] = o; line["speedup_wall"] = o["wall_s"] / g["wall_s"]
                line["parity"] = {"ranks_equal": o["ranks"] == g["ranks"], "note": "libdevice vs libm differ in the last ulp: pivots not bit-comparable"}
            add_native(line, lambda T: T.QuanticsExpSin(ld), ld, gt, **kw)
        else:
            continue
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
