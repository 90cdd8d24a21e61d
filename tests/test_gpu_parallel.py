"""Parallel-bond sweep (bondorder='oddeven') and the sharded function-evaluation batch on the GPU.

Single GPU: the odd/even schedule must reproduce the oracle's restatement of the same schedule bit for bit.
Two or more GPUs (skipped otherwise): one process per GPU over NCCL must give exactly the single-process result."""
import os
import sys

import numpy as np
import pytest

from oracle import oracle as O

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def tci():
    import tci_b200
    return tci_b200


@pytest.mark.parametrize("n,d,tol", [(6, 10, 1e-8), (7, 4, 1e-10)])
def test_oddeven_matches_oracle(tci, n, d, tol):
    f = tci.LorentzFunction([d] * n)
    t, ranks, errors = tci.crossinterpolate2(f, [d] * n, tolerance=tol, nsearchglobalpivot=0, bondorder="oddeven")
    o, oranks, oerrors = O.crossinterpolate2(O.Func(O.F_LORENTZ, [1.0]), [d] * n, tolerance=tol, bondorder="oddeven")
    assert list(ranks) == list(oranks)
    assert np.allclose(errors, oerrors, rtol=1e-10, atol=0)
    for b in range(n):
        assert np.array_equal(t.Iset[b], o.Iset(b + 1)) and np.array_equal(t.Jset[b], o.Jset(b + 1))
    seq, sranks, _ = tci.crossinterpolate2(f, [d] * n, tolerance=tol, nsearchglobalpivot=0)
    rng = np.random.default_rng(0)
    pts = rng.integers(1, d + 1, size=(200, n))
    assert np.max(np.abs(t.evaluate_batch(pts) - seq.evaluate_batch(pts))) < 50 * tol     # same tolerance envelope


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device(f"cuda:{rank}"))
    try:
        import tci_b200 as T
        h = T.Handle(rank)
        comm = T.Comm()
        n, d = 7, 6
        f = T.LorentzFunction([d] * n, handle=h)
        t, ranks, errors = T.crossinterpolate2(f, [d] * n, tolerance=1e-9, nsearchglobalpivot=0, bondorder="oddeven", comm=comm, handle=h)
        sf = T.ShardedBatchEvaluator(f, comm)
        t2, ranks2, errors2 = T.crossinterpolate2(sf, [d] * n, tolerance=1e-9, nsearchglobalpivot=0, handle=h)   # sharded K1, sequential bonds
        ff = T.ShardedBatchEvaluator(f, comm, fused=True)      # peer-store kernel: compute + all-gather in one launch
        t3, ranks3, errors3 = T.crossinterpolate2(ff, [d] * n, tolerance=1e-9, nsearchglobalpivot=0, handle=h)
        fused_used = ff.peers is not None
        # sequential bonds, columns of every Pi with >= 64 entries spread over the ranks (sharded K1 + sharded rrLU)
        t1, ranks1, errors1 = T.crossinterpolate2(f, [d] * n, tolerance=1e-9, nsearchglobalpivot=0, handle=h)
        os.environ["TCI_SHARD_MIN_ELEMS"] = "64"
        t4, ranks4, errors4 = T.crossinterpolate2(f, [d] * n, tolerance=1e-9, nsearchglobalpivot=0, comm=comm, handle=h)
        os.environ.pop("TCI_SHARD_MIN_ELEMS")
        cols = {"ranks4": [int(r) for r in ranks4], "errors4": [float(e) for e in errors4], "nsharded": int(t4.stats.get("nsharded", 0)),
                "I4": [s.tolist() for s in t4.Iset], "J4": [s.tolist() for s in t4.Jset],
                "same_as_one_gpu": bool(all(np.array_equal(a, b) for a, b in zip(t1.Iset, t4.Iset))
                                        and all(np.array_equal(a, b) for a, b in zip(t1.Jset, t4.Jset))
                                        and list(errors1) == list(errors4))}
        rng = np.random.default_rng(0)
        pts = rng.integers(1, d + 1, size=(50, n))
        pts4 = np.random.default_rng(1).integers(1, d + 1, size=(50, n))
        cols["vals4"] = t4.evaluate_batch(pts4).tolist()
        cols["vals1"] = t1.evaluate_batch(pts4).tolist()
        q.put((rank, {**cols, "ranks3": [int(r) for r in ranks3], "I3": [s.tolist() for s in t3.Iset], "vals3": t3.evaluate_batch(pts).tolist(),
                      "fused_used": fused_used,"ranks": [int(r) for r in ranks], "errors": [float(e) for e in errors],
                      "I": [s.tolist() for s in t.Iset], "J": [s.tolist() for s in t.Jset], "vals": t.evaluate_batch(pts).tolist(),
                      "ranks2": [int(r) for r in ranks2], "I2": [s.tolist() for s in t2.Iset], "vals2": t2.evaluate_batch(pts).tolist()}))
    finally:
        dist.destroy_process_group()


def test_multi_gpu_matches_single_process(tci):
    import torch
    world = min(torch.cuda.device_count(), 4)
    if world < 2:
        pytest.skip("needs at least 2 GPUs")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = dict(q.get(timeout=600) for _ in range(world))
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    n, d = 7, 6
    o, oranks, oerrors = O.crossinterpolate2(O.Func(O.F_LORENTZ, [1.0]), [d] * n, tolerance=1e-9, bondorder="oddeven")
    s, sranks, serrors = O.crossinterpolate2(O.Func(O.F_LORENTZ, [1.0]), [d] * n, tolerance=1e-9)
    for r in range(world):
        res = out[r]
        assert res["ranks"] == list(oranks) and np.allclose(res["errors"], oerrors, rtol=1e-10)
        for b in range(n):
            tab = lambda rows, ln: np.asarray(rows, dtype=np.int32).reshape(len(rows), ln)
            assert np.array_equal(tab(res["I"][b], b), o.Iset(b + 1))
            assert np.array_equal(tab(res["J"][b], n - b - 1), o.Jset(b + 1))
            assert np.array_equal(tab(res["I2"][b], b), s.Iset(b + 1))
        assert res["ranks2"] == list(sranks) and res["ranks3"] == list(sranks) and res["fused_used"]
        assert res["I3"] == res["I2"] and res["vals3"] == res["vals2"]
        assert res["nsharded"] > 0 and res["same_as_one_gpu"] and res["ranks4"] == list(sranks)
        assert np.allclose(res["errors4"], serrors, rtol=1e-10) and res["I4"] == res["I2"]
        assert np.allclose(res["vals4"], res["vals1"], rtol=1e-12, atol=0) and res["vals4"] == out[0]["vals4"]
        assert res["vals"] == out[0]["vals"] and res["vals2"] == out[0]["vals2"]      # replicas agree bit for bit


def _sharded_worker(rank, world, port, q):
    import os
    import numpy as np
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", device_id=torch.device(f"cuda:{rank}"))
    import tci_b200 as T
    from tci_b200.parallel import partition, rrlu_sharded
    h = T.Handle(rank)
    comm = T.Comm()
    rng = np.random.default_rng(11)
    A = rng.standard_normal((300, 12)) @ rng.standard_normal((12, 260)) + 1e-9 * rng.standard_normal((300, 260))
    s, e = partition(260, world)[rank]
    res = {}
    for mode in ("nccl", "fused"):
        blk = T.Mat.from_numpy(A[:, s:e], f"cuda:{rank}", pad=2)
        out = rrlu_sharded(blk, 260, s, comm, maxrank=40, reltol=1e-12, handle=h, fused=(mode == "fused"), check_every=8)
        res[mode] = (out["rowperm"], out["colperm"], out["npivot"], out["pivoterrors"], out["L"].numpy(), out["locpos"], out["Uloc"].numpy())
    q.put((rank, res))
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_rrlu_two_gpus_nccl_and_fused(tci):
    """Column-sharded rrLU over 2 processes / 2 GPUs, both exchange modes, against the oracle (bit-exact)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_sharded_worker, args=(r, 2, 29631, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=240) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
    rng = np.random.default_rng(11)
    A = rng.standard_normal((300, 12)) @ rng.standard_normal((12, 260)) + 1e-9 * rng.standard_normal((300, 260))
    ref = O.rrlu(A, maxrank=40, reltol=1e-12)
    for mode in ("nccl", "fused"):
        U = np.zeros((ref.npivot, 260))
        for r in range(2):
            rowperm, colperm, npiv, pe, L, locpos, Uloc = got[r][mode]
            assert npiv == ref.npivot
            np.testing.assert_array_equal(rowperm, ref.rowpermutation)
            np.testing.assert_array_equal(colperm, ref.colpermutation)
            np.testing.assert_array_equal(pe, ref.pivoterrors())
            np.testing.assert_array_equal(L[:, :npiv], ref.L)
            U[:, locpos] = Uloc[:npiv]
        np.testing.assert_array_equal(U, ref.U)
