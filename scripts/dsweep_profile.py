#!/usr/bin/env python3
"""Per-config wall time and driver statistics of the device engine (csrc/sweep.cu) on C1 / C3 / C4; run it under
`ncu --metrics gpu__time_duration.sum` for the per-launch kernel durations."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import tci_b200 as T

which = sys.argv[1] if len(sys.argv) > 1 else "c1,c3,c4"
engine = sys.argv[2] if len(sys.argv) > 2 else "device"
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
h = T.default_handle()
for c in which.split(","):
    if c == "c1":
        ld, kw, mk, piv = [10] * 8, dict(tolerance=1e-8), lambda: T.LorentzFunction(ld), None
    elif c == "c3":
        ld, kw, mk, piv = [20] * 20, dict(tolerance=1e-10, maxbonddim=256), lambda: T.LorentzFunction(ld), None
    elif c == "c4":
        ld, kw, mk = [2] * 40, dict(tolerance=1e-10, maxbonddim=512), lambda: T.QuanticsExpSin(ld)
        piv = [list(T.optfirstpivot(mk(), ld))]
    else:
        ld, kw, mk, piv = [20] * 10, dict(tolerance=1e-10, maxbonddim=256, maxiter=2), lambda: T.HashFunction(ld, 1.0, 1.0), None
    for rep in range(reps):
        f = mk()
        torch.cuda.synchronize()
        l0, t0 = h.launch_count(), time.perf_counter()
        tci, ranks, errors = T.crossinterpolate2(f, ld, piv, nsearchglobalpivot=0, engine=engine, **kw)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        st = {k: (round(v * 1e3, 3) if k.startswith("t_") else v) for k, v in tci.stats.items()}
        print(json.dumps({"config": c, "engine": engine, "rep": rep, "wall_ms": round(dt * 1e3, 3), "launches": h.launch_count() - l0,
                          "ranks": [int(r) for r in ranks], "stats_ms": st}), flush=True)
