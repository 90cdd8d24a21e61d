#!/usr/bin/env python3
"""Column-sharded rrLU of the C2 workload (8192 x 8192, rank 512, maxrank 512) over N GPUs; launch with
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P benchmarks/rrlu_sharded_bench.py [--size 8192]

Every rank builds the same matrix from the C2 seed, keeps its column block, and runs `rrlu_sharded` (one all-gather of the
ranks' candidate records per pivot step).  Device time of the whole factorisation, max over ranks; rank 0 also runs the
single-GPU rrlu of the full matrix and checks that permutations, pivot errors and L agree bit for bit."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    import tci_b200 as T
    from tci_b200.parallel import partition, rrlu_sharded
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=8192)
    ap.add_argument("--rank", type=int, default=512)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--exchange", default="nccl")
    args = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    h = T.Handle(local)
    comm = T.Comm()
    dev = f"cuda:{local}"
    N, R = args.size, args.rank
    g = torch.Generator(device=dev); g.manual_seed(20260001)
    X = torch.randn((N, R), generator=g, device=dev, dtype=torch.float64)
    Y = torch.randn((N, R), generator=g, device=dev, dtype=torch.float64)
    sig = torch.pow(2.0, -torch.arange(R, device=dev, dtype=torch.float64) / 16.0)
    A = (X * sig) @ Y.T                                   # same on every rank (same seed, same device type)
    s, e = partition(N, world)[rank]
    # Mat storage: (cols, ld) row-major == column-major m x cols
    def local_block():
        return T.Mat(A[:, s:e].T.contiguous(), N, e - s)
    from tci_b200.parallel import PeerBuffers, ShardedRRLU
    res = {"workload": f"C2-like rrlu {N}x{N}, rank {R}, maxrank {R}", "n_gpus": world,
           "algorithmic_update_GB": sum(16.0 * (N - k) * (N - k) for k in range(1, R + 1)) / 1e9}
    peers = None
    if world > 1:
        probe = ShardedRRLU(local_block(), N, s, rank, world, R, handle=h)
        peers = PeerBuffers(h, comm, probe.mailbox_doubles(), nslots=1)
        probe.close()
    out = None
    for mode in (["nccl", "fused"] if world > 1 else ["nccl"]):
        times = []
        for rep in range(args.reps + 1):
            blk = local_block()
            comm.barrier(); torch.cuda.synchronize()
            t0 = time.perf_counter()
            out = rrlu_sharded(blk, N, s, comm if world > 1 else None, maxrank=R, handle=h, check_every=0, fused=(mode == "fused"), peers=peers)
            torch.cuda.synchronize()
            if rep:
                times.append(time.perf_counter() - t0)
        t = torch.tensor([min(times)], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        res[f"{mode}_ms"] = float(t.item()) * 1e3
        res[f"{mode}_GBps_aggregate"] = res["algorithmic_update_GB"] / float(t.item())
    res["npivot"] = int(out["npivot"])
    res["exchange"] = {"nccl": "NCCL all_gather_into_tensor issued from the host between the step kernels",
                       "fused": "two rounds over peer-mapped mailboxes inside the step kernels (32-byte candidate headers from the update kernel, then the winner's column from the owner's swap kernel); two launches per step, no host round trip"}
    if rank == 0:
        full = T.Mat(A.T.contiguous(), N, N)
        torch.cuda.synchronize(); t0 = time.perf_counter()
        ref = T.rrlu_device(full, maxrank=R, handle=h)
        torch.cuda.synchronize()
        res["single_gpu_ms_incl_extract"] = (time.perf_counter() - t0) * 1e3
        same = (np.array_equal(ref.rowpermutation, out["rowperm"]) and np.array_equal(ref.colpermutation, out["colperm"])
                and np.array_equal(ref.pivoterrors(), out["pivoterrors"]) and bool(torch.equal(ref.L[:, :R], out["L"].view()[:, :R])))
        res["bit_identical_to_single_gpu"] = bool(same)
        print(json.dumps(res), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
