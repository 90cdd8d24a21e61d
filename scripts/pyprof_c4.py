#!/usr/bin/env python3
"""Host-side profile of one crossinterpolate2 call on C4 (engine="device"): where the wall time outside the kernels goes."""
import cProfile
import os
import pstats
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import tci_b200 as T

ld = [2] * 40
kw = dict(tolerance=1e-10, maxbonddim=512, nsearchglobalpivot=0, engine="device")
f = T.QuanticsExpSin(ld)
piv = [list(T.optfirstpivot(f, ld))]
for _ in range(3):
    T.crossinterpolate2(f, ld, piv, **kw)
torch.cuda.synchronize()
ts = []
for _ in range(5):
    t0 = time.perf_counter()
    tci, ranks, errors = T.crossinterpolate2(f, ld, piv, **kw)
    torch.cuda.synchronize()
    ts.append(1e3 * (time.perf_counter() - t0))
print("wall ms:", [round(t, 3) for t in ts], "stats:", {k: (round(1e3 * v, 3) if k.startswith("t_") else v) for k, v in tci.stats.items()})
pr = cProfile.Profile()
pr.enable()
for _ in range(5):
    T.crossinterpolate2(f, ld, piv, **kw)
torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
