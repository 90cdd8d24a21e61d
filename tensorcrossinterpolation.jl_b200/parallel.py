"""Multi-GPU layer: one process per GPU, `torch.distributed` (NCCL over NVLink on the box, gloo in the
CPU tests).  Only what shards naturally is sharded:

  * the function-evaluation batch (the reference's own parallel axis, src/batcheval.jl:300): `ShardedBatchEvaluator`
    splits the column multi-indices of a Pi matrix over the ranks; with peer-mapped buffers the assembly kernel stores
    every value straight into all ranks' Pi buffers (compute + all-gather in one kernel, csrc/batcheval.cu); without
    them the column blocks are exchanged with an all-gather;
  * the bonds of a half-sweep (NOT in the reference, src/tensorci2.jl:1226-1250 is sequential): `bondorder="oddeven"`
    updates all odd bonds, then all even bonds; inside one pass the bonds read (Iset[b], Jset[b+1]) and write
    (Iset[b+1], Jset[b]) disjoint index sets, so they run concurrently on different GPUs and the new pivot sets are
    exchanged afterwards.  The result is identical to running the same passes on one GPU (and to the oracle's
    restatement of the schedule), whatever the number of ranks;
  * `fillsitetensors!` (src/globalsearch.jl:202-208): sites are independent, dealt round-robin, tensors broadcast.

  * the rrLU itself (an experiment next to the single-GPU default): `rrlu_sharded` deals the COLUMNS of the matrix to the
    ranks; one candidate record per rank and pivot step is exchanged (NCCL all-gather, or peer stores into mailboxes
    fused into the update kernel); bit-identical to the single-GPU factorisation.

`sum`, `evaluate` and zip-up do not shard (replicas only)."""
import numpy as np


def partition(n, parts):
    """Contiguous, balanced split of range(n) into `parts` blocks: list of (start, stop)."""
    base, rem = divmod(int(n), int(parts))
    out, s = [], 0
    for p in range(parts):
        e = s + base + (1 if p < rem else 0)
        out.append((s, e))
        s = e
    return out


def owner_of(items, size):
    """Round-robin assignment: item k of the list belongs to rank k % size."""
    return {it: k % size for k, it in enumerate(items)}


class Comm:
    """Thin wrapper over a torch.distributed process group for the three exchange patterns the path needs."""

    def __init__(self, group=None, device=None):
        import torch.distributed as dist
        self.dist, self.group = dist, group
        self.rank, self.size = dist.get_rank(group), dist.get_world_size(group)
        self.backend = dist.get_backend(group)
        if device is None:
            import torch
            device = f"cuda:{torch.cuda.current_device()}" if self.backend == "nccl" else "cpu"
        self.device = device

    # ---- ragged all-gather of a flat numpy vector
    def _allgather_flat(self, vec, dtype):
        import torch
        t = torch.from_numpy(np.ascontiguousarray(vec)).to(self.device)
        n = torch.tensor([t.numel()], dtype=torch.int64, device=self.device)
        sizes = [torch.zeros(1, dtype=torch.int64, device=self.device) for _ in range(self.size)]
        self.dist.all_gather(sizes, n, group=self.group)
        sizes = [int(s.item()) for s in sizes]
        mx = max(max(sizes), 1)
        pad = torch.zeros(mx, dtype=dtype, device=self.device)
        pad[: t.numel()] = t
        outs = [torch.zeros(mx, dtype=dtype, device=self.device) for _ in range(self.size)]
        self.dist.all_gather(outs, pad, group=self.group)
        return [o[:s].cpu().numpy() for o, s in zip(outs, sizes)]

    def allgather_records(self, records):
        """records: list of (key:int, tables:list of int32 2-D arrays, floats:1-D float64 array) owned by this rank.
        Returns the union over all ranks as a dict key -> (tables, floats)."""
        import torch
        ints, flts = [len(records)], []
        for key, tabs, fl in records:
            fl = np.asarray(fl, dtype=np.float64).reshape(-1)
            ints += [int(key), len(tabs), fl.size]
            for t in tabs:
                t = np.asarray(t, dtype=np.int32)
                ints += [t.shape[0], t.shape[1]]
                ints += t.reshape(-1).tolist()
            flts.append(fl)
        iv = np.asarray(ints, dtype=np.int32)
        fv = np.concatenate(flts) if flts else np.zeros(0)
        gi = self._allgather_flat(iv, torch.int32)
        gf = self._allgather_flat(fv, torch.float64)
        out = {}
        for iv, fv in zip(gi, gf):
            pos, fpos = 1, 0
            for _ in range(int(iv[0])):
                key, ntab, nfl = int(iv[pos]), int(iv[pos + 1]), int(iv[pos + 2])
                pos += 3
                tabs = []
                for _ in range(ntab):
                    r, c = int(iv[pos]), int(iv[pos + 1])
                    pos += 2
                    tabs.append(np.asarray(iv[pos:pos + r * c], dtype=np.int32).reshape(r, c))
                    pos += r * c
                out[key] = (tabs, np.asarray(fv[fpos:fpos + nfl], dtype=np.float64))
                fpos += nfl
        return out

    def allreduce_max(self, x):
        """NaN-propagating max of one float over the ranks (Julia's max, util.jl:34-43)."""
        import torch
        t = torch.tensor([np.inf if np.isnan(x) else float(x), 1.0 if np.isnan(x) else 0.0], dtype=torch.float64, device=self.device)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX, group=self.group)
        return float("nan") if t[1].item() > 0 else float(t[0].item())

    def broadcast(self, tensor, src):
        self.dist.broadcast(tensor, src=src, group=self.group)
        return tensor

    def allgather_column_blocks(self, out2d, blocks):
        """out2d: torch (ncols, ld) matrix storage (column j = row j); every rank has filled rows blocks[rank]; after the
        call every rank has all rows.  Plain all-gather fallback of the fused peer-store kernel."""
        import torch
        mx = max(e - s for s, e in blocks)
        if mx == 0:
            return
        ld = out2d.shape[1]
        s, e = blocks[self.rank]
        mine = torch.zeros((mx, ld), dtype=out2d.dtype, device=out2d.device)
        mine[: e - s] = out2d[s:e]
        outs = [torch.empty_like(mine) for _ in range(self.size)]
        self.dist.all_gather(outs, mine, group=self.group)
        for r, (s2, e2) in enumerate(blocks):
            if r != self.rank and e2 > s2:
                out2d[s2:e2] = outs[r][: e2 - s2]

    def barrier(self):
        self.dist.barrier(group=self.group)


def bond_passes(nbonds, forward, bondorder):
    """Bond visiting order of one half-sweep as a list of passes (lists of 1-based bond numbers).  'sequential' is the
    reference's single pass (tensorci2.jl:1226-1250); 'oddeven' is two passes whose bonds are mutually independent."""
    order = list(range(1, nbonds + 1)) if forward else list(range(nbonds, 0, -1))
    if bondorder == "sequential":
        return [order]
    if bondorder == "oddeven":
        return [[b for b in order if b % 2 == 1], [b for b in order if b % 2 == 0]]
    raise ValueError(f"Unknown bond order {bondorder}: choose 'sequential' or 'oddeven'.")


def run_pass(tci, bonds, update, comm):
    """One pass of mutually independent bond updates.  `update(b)` runs updatepivots! on bond b and returns its pivot
    errors; with a communicator the bonds are dealt round-robin over the ranks and the results exchanged."""
    if comm is None or comm.size == 1:
        for b in bonds:
            update(b)
        return
    own = owner_of(bonds, comm.size)
    records = []
    for b in bonds:
        if own[b] == comm.rank:
            errs = update(b)
            records.append((b, [tci.Iset[b], tci.Jset[b - 1]], np.asarray(errs, dtype=np.float64)))
    got = comm.allgather_records(records)
    for b in bonds:
        if own[b] != comm.rank:
            tabs, errs = got[b]
            tci.Iset[b], tci.Jset[b - 1] = tabs[0], tabs[1]
            tci.updateerrors(b - 1, errs)
    tci.maxsamplevalue = comm.allreduce_max(tci.maxsamplevalue)


from .batcheval import BatchEvaluator          # noqa: E402  (no CUDA needed to import)


class _RawCuda:
    """__cuda_array_interface__ carrier so torch can wrap memory that was allocated with tci_malloc."""

    def __init__(self, ptr, nelem):
        self.__cuda_array_interface__ = {"shape": (int(nelem),), "typestr": "<f8", "data": (int(ptr), False), "version": 3}


class PeerBuffers:
    """`nslots` equal slots of FP64 storage on every rank, allocated with tci_malloc and mapped into every peer process
    with CUDA IPC (tci_ipc_export / tci_ipc_open): ptrs[r] is rank r's buffer as seen from this process."""

    def __init__(self, handle, comm, slot_elems, nslots=3):
        import ctypes as C
        import torch
        from ._lib import lib
        self.handle, self.comm, self.slot_elems, self.nslots = handle, comm, int(slot_elems), nslots
        nbytes = 8 * self.slot_elems * nslots
        own = C.c_void_p()
        handle.check(lib().tci_malloc(handle.h, C.byref(own), C.c_size_t(nbytes)))
        self.own = own.value
        hb = (C.c_ubyte * 64)()
        handle.check(lib().tci_ipc_export(handle.h, own, hb))
        mine = torch.tensor(list(hb), dtype=torch.uint8).to(comm.device)
        alls = [torch.zeros(64, dtype=torch.uint8, device=comm.device) for _ in range(comm.size)]
        comm.dist.all_gather(alls, mine, group=comm.group)
        self.ptrs = []
        for r in range(comm.size):
            if r == comm.rank:
                self.ptrs.append(self.own)
                continue
            raw = (C.c_ubyte * 64)(*alls[r].cpu().tolist())
            p = C.c_void_p()
            handle.check(lib().tci_ipc_open(handle.h, raw, C.byref(p)))
            self.ptrs.append(p.value)
        self._keep = _RawCuda(self.own, self.slot_elems * nslots)
        self.tensor = torch.as_tensor(self._keep, device=f"cuda:{handle.device}")
        self.next_slot = 0

    def take(self, nelem):
        """Next slot (round-robin) as a flat torch view of `nelem` doubles, plus its element offset in the buffer."""
        k = self.next_slot
        self.next_slot = (k + 1) % self.nslots
        off = k * self.slot_elems
        return self.tensor[off:off + int(nelem)], off

    def close(self):
        import ctypes as C
        from ._lib import lib
        self.comm.barrier()
        for r, p in enumerate(self.ptrs):
            if r != self.comm.rank:
                lib().tci_ipc_close(self.handle.h, C.c_void_p(p))
        self.tensor = None
        lib().tci_free(self.handle.h, C.c_void_p(self.own))
        self.ptrs = []


class ShardedBatchEvaluator(BatchEvaluator):
    """Wraps a BatchEvaluator: the right (column) index set of every batch is split over the ranks of `comm`; each rank
    evaluates its column block into the shared-layout output and the blocks are exchanged, so every rank ends up with the
    whole array (what the replicated rrLU needs).  Exposes the same `evaluate_batch`/call protocol as the wrapped object."""

    def __init__(self, f, comm, fused=False):
        """fused=True: outputs live in peer-mapped buffers (`alloc_matrix`) and the assembly kernel stores every value
        into all ranks' buffers (tci_batcheval_multi_f64) — compute and all-gather are one kernel.  fused=False: each rank
        fills its block and the blocks are all-gathered (NCCL / gloo)."""
        import ctypes
        self.fused, self.peers = bool(fused), None
        self.f, self.comm = f, comm                        # (the parent's __init__ would create a device handle)
        self.localdims = np.asarray(f.localdims, dtype=np.int32)
        self.handle = getattr(f, "handle", None)
        self._f = ctypes.c_void_p()

    def __call__(self, *args):
        if len(args) == 1:
            return self.f(*args)
        left, right, M = args
        from .batcheval import to_numpy
        t, _ = self.evaluate_batch(left, right, M)
        return to_numpy(t)

    def evaluate_points(self, pts):
        return self.f.evaluate_points(pts)

    # ---- fused mode: peer-mapped output storage ------------------------------------------------------------------------
    def alloc_matrix(self, rows, cols, pad=16):
        """Pi storage for TensorCI2._fill: a `Mat` in the next peer-mapped slot (three slots: at most two fills are alive
        at any time — Pi1 and P of setsitetensor! — and a rank can run at most one fill ahead of its peers)."""
        from .dense import Mat
        if not self.fused or self.comm.size == 1:
            return None
        ld = max((int(rows) + pad - 1) // pad * pad, 1)
        need = ld * max(int(cols), 1)
        if self.peers is None or self.peers.slot_elems < need:        # collective: every rank sees the same sizes
            if self.peers is not None:
                self.peers.close()
            self.peers = PeerBuffers(self.handle, self.comm, max(need, 1 << 20))
        flat, off = self.peers.take(need)
        m = Mat(flat.reshape(max(int(cols), 1), ld), rows, cols)
        m.transient, m.peer_offset = True, off
        return m

    def _fused_batch(self, lt, rt, M, out, sc, sj, maxabs):
        import ctypes as C
        from ._lib import lib, p_i32
        h, comm = self.handle, self.comm
        off = (out.data_ptr() - self.peers.own) // 8
        blocks = partition(rt.shape[0], comm.size)
        s, e = blocks[comm.rank]
        mx = C.c_double(0.0 if maxabs is None else float(maxabs))
        if e > s:
            order = [comm.rank] + [r for r in range(comm.size) if r != comm.rank]
            outs = (C.c_void_p * comm.size)(*[self.peers.ptrs[r] + 8 * (off + s * sj) for r in order])
            sub = np.ascontiguousarray(rt[s:e])
            h.check(lib().tci_batcheval_multi_f64(h.h, self.f._f, p_i32(self.f.localdims), len(self.f.localdims), p_i32(lt),
                                                  C.c_int64(lt.shape[0]), lt.shape[1], p_i32(sub), C.c_int64(e - s), rt.shape[1],
                                                  outs, comm.size, C.c_int64(sc), C.c_int64(sj), C.byref(mx)))
        h.synchronize()                                            # my peer stores are complete ...
        return comm.allreduce_max(mx.value)                        # ... and so are everyone else's (collective = barrier)

    def evaluate_batch(self, left, right, M, out=None, stride_c=0, stride_j=0, maxabs=None):
        import torch
        from .batcheval import as_table
        lt, rt = as_table(left), as_table(right)
        nI, nJ = lt.shape[0], rt.shape[0]
        if nI * nJ == 0 or self.comm.size == 1:
            return self.f.evaluate_batch(lt, rt, M, out=out, stride_c=stride_c, stride_j=stride_j, maxabs=maxabs)
        nl = lt.shape[1]
        cd = [int(d) for d in self.localdims[nl:nl + M]]
        ncent = int(np.prod(cd)) if cd else 1
        sc = stride_c if stride_c > 0 else nI
        sj = stride_j if stride_j > 0 else sc * ncent
        if out is None:
            dev = self.comm.device if self.handle is None else f"cuda:{self.handle.device}"
            out = torch.empty((nJ,) + tuple(reversed(cd)) + (nI,), dtype=torch.float64, device=dev)
        if (self.fused and self.peers is not None and getattr(self.f, "_f", None)
                and self.peers.own <= out.data_ptr() < self.peers.own + 8 * self.peers.slot_elems * self.peers.nslots):
            mx = self._fused_batch(lt, rt, M, out, sc, sj, maxabs)
            return out, (mx if maxabs is not None else None)
        flat = out.reshape(-1)
        blocks = partition(nJ, self.comm.size)
        s, e = blocks[self.comm.rank]
        mx = maxabs
        if e > s:
            sub = flat[s * sj:]                                    # column block s..e starts at element s * stride_j
            _, mx = self.f.evaluate_batch(lt, rt[s:e], M, out=sub, stride_c=sc, stride_j=sj, maxabs=maxabs)
        rows = torch.as_strided(flat, (nJ, sj), (sj, 1)) if flat.numel() >= nJ * sj else None
        if rows is None:                                           # last column shorter than stride_j: gather on a padded copy
            pad = torch.zeros(nJ * sj, dtype=flat.dtype, device=flat.device)
            pad[: flat.numel()] = flat
            self.comm.allgather_column_blocks(pad.reshape(nJ, sj), blocks)
            flat.copy_(pad[: flat.numel()])
        else:
            self.comm.allgather_column_blocks(rows, blocks)
        if maxabs is not None:
            mx = self.comm.allreduce_max(mx)
        return out, mx


# ---- column-sharded full-pivot rrLU (SURVEY.md 8e row "rrLU": one exchange per pivot step) ----------------------------------
class ShardedRRLU:
    """One rank's part of a column-sharded rrlu (csrc/rrlu.cu, tci_rrlu_dist_*).  `A` is this rank's column block as a
    `dense.Mat` (m x nloc, overwritten), `col0` its first global column."""

    def __init__(self, A, n, col0, rank, nranks, maxrank=None, reltol=1e-14, abstol=0.0, leftorthogonal=True, handle=None):
        import ctypes as C
        import torch
        from . import _lib
        from ._lib import lib
        self.h = handle or _lib.default_handle()
        self.A, self.m, self.n, self.nloc, self.col0 = A, A.rows, int(n), A.cols, int(col0)
        self.rank, self.nranks, self.leftorthogonal = rank, nranks, bool(leftorthogonal)
        mn = min(self.m, self.n)
        self.maxrank = mn if maxrank is None or maxrank <= 0 or maxrank > mn else int(maxrank)
        self._d = C.c_void_p()
        self.h.check(lib().tci_rrlu_dist_create(self.h.h, C.c_int64(self.m), C.c_int64(self.n), C.c_int64(self.nloc), C.c_int64(col0),
                                                int(rank), int(nranks), C.c_int64(self.maxrank), C.c_double(reltol), C.c_double(abstol),
                                                int(self.leftorthogonal), C.byref(self._d)))
        lib().tci_rrlu_dist_record_doubles.restype = C.c_int64
        self.reclen = int(lib().tci_rrlu_dist_record_doubles(self._d))
        self.record = torch.zeros(self.reclen, dtype=torch.float64, device=A.buf.device)

    def _aptr(self):
        import ctypes as C
        return C.c_void_p(self.A.buf.data_ptr())

    def search(self):
        import ctypes as C
        from ._lib import lib
        self.h.check(lib().tci_rrlu_dist_search(self.h.h, self._d, self._aptr(), C.c_int64(self.A.ld), C.c_void_p(self.record.data_ptr())))

    def step(self, gathered):
        import ctypes as C
        from ._lib import lib
        self.h.check(lib().tci_rrlu_dist_step(self.h.h, self._d, self._aptr(), C.c_int64(self.A.ld), C.c_void_p(gathered.data_ptr()),
                                              C.c_void_p(self.record.data_ptr())))

    def mailbox_doubles(self):
        import ctypes as C
        from ._lib import lib
        lib().tci_rrlu_dist_mailbox_bytes.restype = C.c_int64
        return (int(lib().tci_rrlu_dist_mailbox_bytes(self._d)) + 7) // 8

    def set_mailboxes(self, ptrs):
        import ctypes as C
        from ._lib import lib
        arr = (C.c_void_p * len(ptrs))(*[C.c_void_p(p) for p in ptrs])
        self.h.check(lib().tci_rrlu_dist_set_mailboxes(self.h.h, self._d, arr))

    def run(self):
        """Enqueue the whole factorisation (fused peer-store exchange); returns without synchronising."""
        import ctypes as C
        from ._lib import lib
        self.h.check(lib().tci_rrlu_dist_run(self.h.h, self._d, self._aptr(), C.c_int64(self.A.ld)))

    def trace(self, nsteps):
        """Device timestamps of the step kernels (TCI_DIST_TRACE=1 at creation): (nsteps, 8) uint64 ns, see tci_b200.h."""
        import ctypes as C
        from ._lib import lib
        out = np.zeros((int(nsteps), 8), dtype=np.uint64)
        self.h.check(lib().tci_rrlu_dist_trace(self.h.h, self._d, out.ctypes.data_as(C.c_void_p), C.c_int64(int(nsteps))))
        return out

    def status(self):
        import ctypes as C
        from ._lib import lib
        np_, stop = C.c_int64(), C.c_int()
        self.h.check(lib().tci_rrlu_dist_status(self.h.h, self._d, C.byref(np_), C.byref(stop)))
        return np_.value, bool(stop.value)

    def finish(self):
        """-> dict(rowperm, colperm (1-based, full length), npivot, error, pivoterrors, L (Mat m x r), Uloc (Mat r x nloc), locpos)."""
        import ctypes as C
        from ._lib import lib, p_f64, p_i64
        from .dense import Mat
        rowperm, colperm = np.zeros(self.m, dtype=np.int64), np.zeros(self.n, dtype=np.int64)
        locpos = np.zeros(max(self.nloc, 1), dtype=np.int64)
        pe = np.zeros(self.maxrank + 1)
        npiv, err = C.c_int64(), C.c_double()
        dev = self.A.buf.device
        L = Mat.empty(self.m, max(self.maxrank, 1), dev)
        U = Mat.empty(max(self.maxrank, 1), self.nloc, dev)
        self.h.check(lib().tci_rrlu_dist_finish(self.h.h, self._d, self._aptr(), C.c_int64(self.A.ld), p_i64(rowperm), p_i64(colperm),
                                                C.byref(npiv), C.byref(err), L.ptr, C.c_int64(L.ld), U.ptr, C.c_int64(U.ld),
                                                p_i64(locpos), p_f64(pe)))
        r = npiv.value
        L.cols, U.rows = r, r
        return {"rowperm": rowperm, "colperm": colperm, "npivot": r, "error": err.value, "pivoterrors": pe[: r + 1].copy(),
                "L": L, "Uloc": U, "locpos": locpos[: self.nloc]}

    def close(self):
        from ._lib import lib
        if self._d:
            lib().tci_rrlu_dist_destroy(self.h.h, self._d)
            self._d = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def rrlu_sharded(A_local, n, col0, comm=None, maxrank=None, reltol=1e-14, abstol=0.0, leftorthogonal=True, handle=None,
                 check_every=32, fused=False, peers=None):
    """rrlu!(A; ...) (matrixlu.jl:420-430) with the columns of A spread over the ranks of `comm` (NCCL).  Per pivot step the
    ranks exchange their best candidates: fused=False -> one record per rank (header + the candidate's column,
    8 x (m + 4) bytes) through an NCCL all-gather issued from the host between the step kernels; fused=True -> two rounds
    over peer-mapped mailboxes inside the step kernels (32-byte headers stored by the update kernel, then only the
    winner's column, stored by the owner's swap kernel; `peers`: a PeerBuffers with >= mailbox_doubles() elements per
    slot or a MailboxPool, created once and reused) and the whole factorisation is enqueued without any host round trip.  Returns this rank's `ShardedRRLU.finish()` dict; rowperm /
    colperm / npivot / pivoterrors / L are the same on every rank and bit-identical to the single-GPU rrlu."""
    import torch
    size = 1 if comm is None else comm.size
    rank = 0 if comm is None else comm.rank
    s = ShardedRRLU(A_local, n, col0, rank, size, maxrank, reltol, abstol, leftorthogonal, handle)
    if fused and size > 1:
        if hasattr(peers, "get"):                        # a MailboxPool: grows collectively when this matrix needs more
            peers = peers.get(s.mailbox_doubles())
        own_peers = peers is None
        if own_peers:
            peers = PeerBuffers(s.h, comm, s.mailbox_doubles(), nslots=1)
        if peers.slot_elems * peers.nslots < s.mailbox_doubles():
            raise ValueError("rrlu_sharded: the peer buffers are smaller than one mailbox")
        s.set_mailboxes(peers.ptrs)                      # zeroes the own mailbox
        comm.barrier()                                   # nobody stores into a mailbox that is still being cleared
        s.run()
        out = s.finish()
        s.close()
        if own_peers:
            peers.close()
        return out
    gathered = torch.empty(size * s.reclen, dtype=torch.float64, device=A_local.buf.device)
    s.search()
    for it in range(s.maxrank):
        if size == 1:
            gathered = s.record
        else:
            comm.dist.all_gather_into_tensor(gathered, s.record, group=comm.group)
        s.step(gathered)
        if check_every and (it + 1) % check_every == 0 and it + 1 < s.maxrank:
            if s.status()[1]:                            # every rank takes the same decision: the loop lengths agree
                break
    out = s.finish()
    s.close()
    return out


class MailboxPool:
    """The peer-mapped mailboxes of the fused sharded rrLU, created once per run and re-created (collectively: every rank
    sees the same matrix sizes) only when a larger matrix arrives."""

    def __init__(self, handle, comm):
        self.handle, self.comm, self.peers = handle, comm, None

    def get(self, need):
        if self.peers is None or self.peers.slot_elems < need:
            if self.peers is not None:
                self.peers.close()
            self.peers = PeerBuffers(self.handle, self.comm, max(int(need), 1 << 16), nslots=1)
        return self.peers

    def close(self):
        if self.peers is not None:
            self.peers.close()
            self.peers = None


def gather_sharded_U(out, n, blocks, comm):
    """The full U (npivot x n, swapped column order) as a device `Mat` on every rank, from the per-rank pieces of
    `rrlu_sharded`: one all-gather of the (padded) local blocks, each column tagged with its global position, then a
    device scatter of the columns to their positions."""
    import torch
    from .dense import Mat
    r, Uloc, locpos = out["npivot"], out["Uloc"], out["locpos"]
    dev = Uloc.buf.device
    mx = max(e - s for s, e in blocks)
    nloc = Uloc.cols
    mine = torch.zeros((mx, r + 1), dtype=torch.float64, device=dev)
    if nloc:
        mine[:nloc, :r] = Uloc.buf[:nloc, :r]
        mine[:nloc, r] = torch.from_numpy(np.asarray(locpos, dtype=np.float64)).to(dev)
    allb = torch.empty((comm.size * mx, r + 1), dtype=torch.float64, device=dev)
    comm.dist.all_gather_into_tensor(allb, mine, group=comm.group)
    keep = np.concatenate([np.arange(q * mx, q * mx + (e - s)) for q, (s, e) in enumerate(blocks)])
    if len(keep) != n:
        raise ValueError("gather_sharded_U: the column blocks do not cover the matrix")
    rows = allb[torch.from_numpy(keep).to(dev)] if len(keep) != comm.size * mx else allb
    U = Mat.empty(max(r, 1), n, dev)
    U.rows = r
    U.buf[:n, :r].index_copy_(0, rows[:, r].to(torch.int64), rows[:, :r].contiguous())
    return U


def rrlu_sharded_virtual(A, nranks, maxrank=None, reltol=1e-14, abstol=0.0, leftorthogonal=True, handle=None):
    """The sharded factorisation with `nranks` VIRTUAL ranks on one GPU (lock-step, records concatenated by device copies):
    the algorithm and kernels of `rrlu_sharded` without a process group — used by the single-GPU parity tests.
    A: numpy matrix.  Returns (list of per-rank finish() dicts, column blocks)."""
    import torch
    from . import _lib
    from .dense import Mat
    h = handle or _lib.default_handle()
    dev = f"cuda:{h.device}"
    A = np.asarray(A, dtype=np.float64)
    m, n = A.shape
    blocks = partition(n, nranks)
    parts = [ShardedRRLU(Mat.from_numpy(A[:, s:e], dev, pad=2), n, s, r, nranks, maxrank, reltol, abstol, leftorthogonal, h)
             for r, (s, e) in enumerate(blocks)]
    reclen = parts[0].reclen
    gathered = torch.empty(nranks * reclen, dtype=torch.float64, device=dev)
    for p in parts:
        p.search()
    for it in range(parts[0].maxrank):
        for r, p in enumerate(parts):
            gathered[r * reclen:(r + 1) * reclen].copy_(p.record)
        for p in parts:
            p.step(gathered)
    outs = [p.finish() for p in parts]
    for p in parts:
        p.close()
    return outs, blocks


def assemble_sharded_U(outs, n):
    """Full U (npivot x n, swapped column order) from the per-rank pieces of `rrlu_sharded` (host, numpy)."""
    r = outs[0]["npivot"]
    U = np.zeros((r, n))
    for o in outs:
        if o["Uloc"].cols:
            U[:, o["locpos"]] = o["Uloc"].numpy()[:r]
    return U
