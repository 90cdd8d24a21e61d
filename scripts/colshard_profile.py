"""Per-phase split and cProfile (rank 0) of the C3H sweep with the columns of Pi spread over the ranks; launch with torchrun."""
import cProfile, pstats, io, os, sys, time
sys.path.insert(0, '.')
import numpy as np
import torch, torch.distributed as dist
import tci_b200 as T
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
h = T.Handle(local); comm = T.Comm()
ld = [20] * 10
f = T.HashFunction(ld, 1.0, 1.0, handle=h)
kw = dict(tolerance=1e-10, maxbonddim=int(os.environ.get("MAXBOND", 256)), maxiter=2, nsearchglobalpivot=0, handle=h)
for mode in ("one", "cols"):
    extra = {} if mode == "one" else {"comm": comm}
    T.crossinterpolate2(f, ld, **kw, **extra)
    for rep in range(2):
        comm.barrier(); torch.cuda.synchronize(); t0 = time.perf_counter()
        tci, r, e = T.crossinterpolate2(f, ld, **kw, **extra); torch.cuda.synchronize()
        if rank == 0:
            print(mode, "wall", round(time.perf_counter() - t0, 4), {k: (round(v, 4) if isinstance(v, float) else v) for k, v in tci.stats.items()}, flush=True)
pr = cProfile.Profile(); pr.enable()
T.crossinterpolate2(f, ld, **kw, comm=comm); torch.cuda.synchronize()
pr.disable()
if rank == 0:
    s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("tottime").print_stats(25); print(s.getvalue()[:6000])
dist.destroy_process_group()
