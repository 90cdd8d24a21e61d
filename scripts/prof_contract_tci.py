#!/usr/bin/env python3
"""Where contract_TCI spends its time at C5 size (30 sites, legs 4 x 4, bond 200, maxbonddim 200, one iteration): device time
inside _contract (GEMM), host profile (cProfile) and the sweep driver's own counters."""
import cProfile
import os
import pstats
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import tci_b200 as T
from benchmarks import contract_bench as CB
from tci_b200 import contraction as CM
from tci_b200 import tensorci2 as T2

sites, chi, d = int(sys.argv[1]) if len(sys.argv) > 1 else 30, int(sys.argv[2]) if len(sys.argv) > 2 else 200, 4
a, b = CB.make(sites, chi, d)
A, B = T.TensorTrain(a), T.TensorTrain(b)
ev = []
orig = CM._contract


def timed(*x, **k):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); r = orig(*x, **k); e1.record(); ev.append((e0, e1)); return r


CM._contract = timed
captured = {}
orig_ci2 = T2.crossinterpolate2


def ci2(*x, **k):
    r = orig_ci2(*x, **k); captured["tci"] = r[0]; return r


T2.crossinterpolate2 = ci2
pr = cProfile.Profile()
torch.cuda.synchronize(); t0 = time.perf_counter()
pr.enable()
tt = T.contract_TCI(A, B, tolerance=1e-10, maxbonddim=chi, maxiter=1, rng=np.random.default_rng(5), nsearchglobalpivot=0)
torch.cuda.synchronize()
pr.disable()
dt = time.perf_counter() - t0
tg = sum(x.elapsed_time(y) for x, y in ev) * 1e-3
print(f"wall {dt:.2f} s; device time inside _contract {tg:.2f} s over {len(ev)} calls; linkdims max {max(tt.linkdims())}")
if "tci" in captured:
    print("stats:", {k: (round(v, 3) if isinstance(v, float) else v) for k, v in captured["tci"].stats.items()})
pstats.Stats(pr).sort_stats("tottime").print_stats(22)
