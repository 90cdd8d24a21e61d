#!/usr/bin/env python3
"""Multi-GPU measurements; launch with
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P benchmarks/multigpu_bench.py

(a) one Pi fill of the C3 worst-case shape (5120 x 5120, 26.2 M evaluations, 210 MB) with the full-rank hash functor and
    with an arithmetic-heavy NVRTC plugin functor: local (every rank evaluates everything), sharded + NCCL all-gather,
    sharded + fused peer-store kernel.  Device time, max over ranks.
(b) C3H sweep (full-rank function, localdims=fill(20,10), maxbonddim=256, 2 iterations = 4 half-sweeps): sequential bonds on
    every rank (replicas) vs bondorder='oddeven' with the bonds spread over the ranks.  Wall time, max over ranks."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

HEAVY = """
__device__ double tci_user_f(const int* idx, int n, const double* p, int np) {
    double s = 0.0;
    for (int i = 0; i < n; ++i) s += sin(p[0] * idx[i] * (i + 1)) * exp(-p[1] * idx[i]);
    return s / (1.0 + s * s);
}
"""


def main():
    import torch
    import torch.distributed as dist
    import tci_b200 as T
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    h = T.Handle(local)
    comm = T.Comm()
    dev = f"cuda:{local}"
    ld = [20] * 10
    rng = np.random.default_rng(0)
    nI = nJ = 5120
    left = rng.integers(1, 21, size=(nI, 5)).astype(np.int32)
    right = rng.integers(1, 21, size=(nJ, 5)).astype(np.int32)

    def timed(fn, reps=5):
        fn(); fn()
        ts = []
        for _ in range(reps):
            comm.barrier(); torch.cuda.synchronize()
            t0 = time.perf_counter()
            fn()
            torch.cuda.synchronize()
            ts.append(time.perf_counter() - t0)
        t = torch.tensor([min(ts)], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    out = {}
    for name, f in [("hash", T.HashFunction(ld, 1.0, 1.0, handle=h)), ("heavy_plugin", T.UserFunction(HEAVY, ld, [0.37, 0.05], handle=h))]:
        buf = torch.empty(nI * nJ, dtype=torch.float64, device=dev)
        t_local = timed(lambda: f.evaluate_batch(left, right, 0, out=buf, stride_c=nI, stride_j=nI, maxabs=0.0))
        sh = T.ShardedBatchEvaluator(f, comm)
        t_gather = timed(lambda: sh.evaluate_batch(left, right, 0, out=buf, stride_c=nI, stride_j=nI, maxabs=0.0))
        fu = T.ShardedBatchEvaluator(f, comm, fused=True)

        def fused():
            m = fu.alloc_matrix(nI, nJ, 16)
            o = m.buf if m is not None else buf
            fu.evaluate_batch(left, right, 0, out=o, stride_c=nI, stride_j=nI, maxabs=0.0)
        t_fused = timed(fused)
        ref = f.evaluate_batch(left, right, 0, out=torch.empty_like(buf), stride_c=nI, stride_j=nI)[0]
        m = fu.alloc_matrix(nI, nJ, 16)
        o = m.buf if m is not None else buf
        fu.evaluate_batch(left, right, 0, out=o, stride_c=nI, stride_j=nI, maxabs=0.0)
        same = bool(torch.equal(o.reshape(-1)[: nI * nJ], ref.reshape(-1)[: nI * nJ]))
        out[name] = {"evals": nI * nJ, "local_ms": 1e3 * t_local, "sharded_allgather_ms": 1e3 * t_gather, "sharded_fused_ms": 1e3 * t_fused,
                     "fused_equals_local": same, "bytes_over_nvlink_per_rank": 8 * nI * nJ * (world - 1) / world,
                     "fused_nvlink_GBps_per_rank": 8 * nI * nJ * (world - 1) / world / t_fused / 1e9}

    f = T.HashFunction(ld, 1.0, 1.0, handle=h)
    kw = dict(tolerance=1e-10, maxbonddim=256, maxiter=2, nsearchglobalpivot=0, handle=h)

    def sweep(**extra):
        T.crossinterpolate2(f, ld, **kw, **extra)
    t_seq = timed(lambda: sweep(), reps=2)
    t_oe1 = timed(lambda: sweep(bondorder="oddeven"), reps=2)
    t_oe = timed(lambda: sweep(bondorder="oddeven", comm=comm), reps=2)
    # sequential bond order (the reference's) with the columns of every large Pi spread over the ranks: K1 evaluates only
    # the local column block, the sharded rrLU factorises it in place; must reproduce the one-GPU pivots exactly
    ref_tci = T.crossinterpolate2(f, ld, **kw)[0]
    sh_tci = T.crossinterpolate2(f, ld, **kw, comm=comm)[0]
    same = all(np.array_equal(a, b) for a, b in zip(ref_tci.Iset, sh_tci.Iset)) and \
        all(np.array_equal(a, b) for a, b in zip(ref_tci.Jset, sh_tci.Jset))
    t_cs = timed(lambda: sweep(comm=comm), reps=2)
    out["c3h_sweep"] = {"sequential_replicated_s": t_seq, "oddeven_one_gpu_s": t_oe1, "oddeven_bonds_over_ranks_s": t_oe,
                        "speedup_vs_sequential": t_seq / t_oe, "sequential_columns_over_ranks_s": t_cs,
                        "columns_speedup_vs_sequential": t_seq / t_cs, "columns_pivots_identical": bool(same),
                        "columns_sharded_bonds": int(sh_tci.stats.get("nsharded", 0))}
    if os.environ.get("C3H512", "1") != "0":             # the same sweep at maxbonddim=512: Pi up to 10240 x 10240
        kw5 = dict(kw, maxbonddim=512, maxiter=2)
        t1 = timed(lambda: T.crossinterpolate2(f, ld, **kw5), reps=1)
        t8 = timed(lambda: T.crossinterpolate2(f, ld, **kw5, comm=comm), reps=1)
        out["c3h512_sweep"] = {"sequential_replicated_s": t1, "sequential_columns_over_ranks_s": t8, "speedup": t1 / t8}
    if rank == 0:
        print(json.dumps({"n_gpus": world, **out}), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
