#!/usr/bin/env python3
"""Host-side profile (rank 0) of the column-sharded C3H sweep; run under torchrun with N >= 2."""
import cProfile
import os
import pstats
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
import tci_b200 as T

rank, local = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
h = T.Handle(local)
comm = T.Comm()
ld = [20] * 10
kw = dict(tolerance=1e-10, maxbonddim=256, maxiter=2, nsearchglobalpivot=0, handle=h)
f = T.HashFunction(ld, 1.0, 1.0, handle=h)
for _ in range(2):
    T.crossinterpolate2(f, ld, comm=comm, **kw)
torch.cuda.synchronize(); dist.barrier()
pr = cProfile.Profile()
t0 = time.perf_counter()
pr.enable()
tci, _, _ = T.crossinterpolate2(f, ld, comm=comm, **kw)
torch.cuda.synchronize()
pr.disable()
dt = time.perf_counter() - t0
if rank == 0:
    print(f"wall {dt*1e3:.1f} ms; stats", {k: (round(1e3 * v, 1) if k.startswith("t_") else v) for k, v in tci.stats.items()})
    pstats.Stats(pr).sort_stats("tottime").print_stats(25)
dist.destroy_process_group()
