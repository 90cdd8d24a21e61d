"""Where does a pivot step of the column-sharded rrLU spend its time?  TCI_DIST_TRACE=1 torchrun ... scripts/dist_trace.py [size rank]"""
import os, sys, time
sys.path.insert(0, '.')
os.environ["TCI_DIST_TRACE"] = "1"
import numpy as np, torch, torch.distributed as dist
import tci_b200 as T
from tci_b200.parallel import partition, ShardedRRLU, PeerBuffers
from tci_b200.dense import Mat
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
h = T.Handle(local); comm = T.Comm()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 5120
r = int(sys.argv[2]) if len(sys.argv) > 2 else 256
g = torch.Generator(device=f"cuda:{local}"); g.manual_seed(7)
X = torch.randn(n, r, generator=g, device=f"cuda:{local}", dtype=torch.float64)
Y = torch.randn(r, n, generator=g, device=f"cuda:{local}", dtype=torch.float64)
A = (X @ Y)
s0, e0 = partition(n, world)[rank]
for rep in range(2):
    Al = Mat.empty(n, e0 - s0, f"cuda:{local}", pad=16)
    Al.buf[: e0 - s0, :n].copy_(A[:, s0:e0].T)
    s = ShardedRRLU(Al, n, s0, rank, world, r, 1e-14, 0.0, True, h)
    peers = PeerBuffers(h, comm, s.mailbox_doubles(), nslots=1)
    s.set_mailboxes(peers.ptrs); comm.barrier()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    s.run(); out = s.finish(); dt = time.perf_counter() - t0
    tr = s.trace(r).astype(np.int64)
    s.close(); peers.close()
lo, hi = r // 4, 3 * r // 4
names = ["swap start", "headers in", "column out/in", "update start", "block0 done", "all tiles done", "header released"]
res = {"rank": rank, "ms": 1e3 * dt, "step_us": float((tr[hi, 0] - tr[lo, 0]) / (hi - lo) / 1e3)}
prev = tr[lo:hi, 6] * 0
# per-step deltas: previous step's header release -> this step's stamps
base = tr[lo - 1:hi - 1, 6]
for i, nm in enumerate(names):
    d = (tr[lo:hi, i] - base) / 1e3
    res[nm] = [round(float(np.median(d)), 2), round(float(np.percentile(d, 90)), 2)]
outs = [None] * world
dist.all_gather_object(outs, res)
if rank == 0:
    print(f"size {n} rank {r} gpus {world}: microseconds after the previous step's header release [median, p90]")
    for o in outs[:3] + outs[-1:]:
        print(o)
dist.destroy_process_group()
