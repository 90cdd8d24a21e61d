"""World-size-2 tests of the multi-GPU host logic on CPU (gloo): ragged record exchange, column-block all-gather, the
sharded batch evaluator (with the ORACLE as the stand-in evaluator — test infrastructure) and the odd/even
parallel-bond pass.  The CUDA path itself is covered by tests/test_gpu_parallel.py."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class OracleEvaluator:
    """CPU stand-in with the evaluate_batch protocol of tci_b200.BatchEvaluator, backed by the oracle."""

    def __init__(self, localdims):
        from oracle import oracle as O
        self.O, self.localdims, self.handle = O, np.asarray(localdims, dtype=np.int32), None
        self.g = O.Func(O.F_LORENTZ, [1.0])

    def evaluate_batch(self, left, right, M, out=None, stride_c=0, stride_j=0, maxabs=None):
        ref = self.O.batchevaluate(self.g, [int(d) for d in self.localdims], [list(r) for r in left], [list(r) for r in right], M)
        nI, nJ = len(left), len(right)
        ncent = ref.size // (nI * nJ)
        sc = stride_c if stride_c > 0 else nI
        sj = stride_j if stride_j > 0 else sc * ncent
        vals = torch.from_numpy(np.ascontiguousarray(ref.reshape(nI, ncent, nJ, order="F").transpose(2, 1, 0)))
        if out is None:
            out = torch.zeros(nJ * sj, dtype=torch.float64)
        torch.as_strided(out.reshape(-1), (nJ, ncent, nI), (sj, sc, 1)).copy_(vals)
        if maxabs is not None:
            maxabs = max(abs(maxabs), float(np.abs(ref).max()))
        return out, maxabs


class FakeTCI:
    """Deterministic stand-in for TensorCI2: updatepivots derives the new sets from exactly the sets the real one reads."""

    def __init__(self, n):
        self.Iset = [np.full((2, b), b + 1, dtype=np.int32) for b in range(n)]
        self.Jset = [np.full((2, n - b - 1), 10 * (b + 1), dtype=np.int32) for b in range(n)]
        self.maxsamplevalue, self.bonderrors, self.pivoterrors = 1.0, np.zeros(n - 1), np.zeros(0)

    def updateerrors(self, b0, errs):
        self.bonderrors[b0] = errs[-1]

    def update(self, b):
        b0 = b - 1
        seed = int(self.Iset[b0].sum() * 7 + self.Jset[b0 + 1].sum() * 13 + b) % 1000
        r = 1 + seed % 3
        self.Iset[b0 + 1] = np.full((r, b0 + 1), seed, dtype=np.int32)
        self.Jset[b0] = np.full((r, len(self.Iset) - b0 - 1), seed + 1, dtype=np.int32)
        errs = np.array([seed / 7.0, seed / 11.0])
        self.updateerrors(b0, errs)
        self.maxsamplevalue = max(self.maxsamplevalue, float(seed))
        return errs


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import tci_b200
        from tci_b200.parallel import Comm, ShardedBatchEvaluator, bond_passes, partition, run_pass
        comm = Comm()
        res = {}
        # 1. ragged records
        recs = [(10 * rank + k, [np.arange(6 * (k + 1), dtype=np.int32).reshape(-1, 3) + rank, np.zeros((0, 4), dtype=np.int32)],
                 np.array([rank + 0.5, k])) for k in range(rank + 1)]
        got = comm.allgather_records(recs)
        res["keys"] = sorted(got)
        res["rec_ok"] = bool(np.array_equal(got[10][0][0], np.arange(6, dtype=np.int32).reshape(-1, 3) + 1) and got[11][0][1].shape == (0, 4)
                             and np.allclose(got[11][1], [1.5, 1.0]))
        res["max"] = comm.allreduce_max(float(rank) + 2.0)
        res["nanmax"] = comm.allreduce_max(float("nan") if rank == 1 else 1.0)
        # 2. sharded evaluation == unsharded evaluation (padded leading dimension, M = 0 and M = 1)
        ld = [4, 3, 5, 2]
        f = OracleEvaluator(ld)
        sh = ShardedBatchEvaluator(f, comm)
        left, right = [[1], [3], [4]], [[1, 1], [2, 2], [5, 1], [3, 2], [4, 1]]
        full, mfull = f.evaluate_batch(left, right, 1, out=torch.zeros(5 * 16), stride_c=3, stride_j=16, maxabs=0.0)
        mine, mmine = sh.evaluate_batch(left, right, 1, out=torch.zeros(5 * 16), stride_c=3, stride_j=16, maxabs=0.0)
        res["shard_eq"] = bool(torch.equal(full, mine)) and mfull == mmine
        full0, _ = f.evaluate_batch([[1, 2], [4, 3]], [[1, 1], [2, 2], [5, 1]], 0)
        mine0, _ = sh.evaluate_batch([[1, 2], [4, 3]], [[1, 1], [2, 2], [5, 1]], 0)
        res["shard_eq0"] = bool(torch.equal(full0.reshape(-1), mine0.reshape(-1)))
        # 3. odd/even pass over the ranks == the same passes on one process
        n = 7
        par, ser = FakeTCI(n), FakeTCI(n)
        for fwd in (True, False):
            for bonds in bond_passes(n - 1, fwd, "oddeven"):
                run_pass(par, bonds, par.update, comm)
                run_pass(ser, bonds, ser.update, None)
        res["sets_eq"] = all(np.array_equal(a, b) for a, b in zip(par.Iset + par.Jset, ser.Iset + ser.Jset))
        res["err_eq"] = bool(np.array_equal(par.bonderrors, ser.bonderrors)) and par.maxsamplevalue == ser.maxsamplevalue
        res["partition"] = partition(7, 3)
        q.put((rank, res))
    finally:
        dist.destroy_process_group()


def test_world_size_2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = dict(q.get(timeout=180) for _ in range(2))
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    for r in range(2):
        res = out[r]
        assert res["keys"] == [0, 10, 11] and res["rec_ok"]
        assert res["max"] == 3.0 and np.isnan(res["nanmax"])
        assert res["shard_eq"] and res["shard_eq0"]
        assert res["sets_eq"] and res["err_eq"]
        assert res["partition"] == [(0, 3), (3, 5), (5, 7)]


def test_bond_passes():
    sys.path.insert(0, ROOT)
    from tci_b200.parallel import bond_passes
    assert bond_passes(5, True, "sequential") == [[1, 2, 3, 4, 5]]
    assert bond_passes(5, False, "sequential") == [[5, 4, 3, 2, 1]]
    assert bond_passes(6, True, "oddeven") == [[1, 3, 5], [2, 4, 6]]
    assert bond_passes(6, False, "oddeven") == [[5, 3, 1], [6, 4, 2]]
    with pytest.raises(ValueError):
        bond_passes(3, True, "zigzag")
