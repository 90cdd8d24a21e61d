"""Host-side mirror of src/tensorci2.jl: TensorCI2 state, 1-/2-site sweeps, `optimize!`, `crossinterpolate2`.

This file is bookkeeping only (index sets, sweep order, convergence control) — what the north star keeps in
the host language.  Every number is produced by the CUDA library: K1 batch evaluation (batcheval.py), K2/K3
rrLU (matrixlu.py), K4a factors/solves (dense.py), K4b evaluate/sum (tensortrain.py).  Index sets are int32
tables (count x length), 1-based values like Julia.  Bond/site arguments of public methods are 1-based.
"""
import os
import time

import numpy as np

from . import _lib
from ._lib import TCIError
from .batcheval import BatchEvaluator, as_table
from .dense import Mat, solve_right
from .matrixlu import INT_MAX
from .matrixluci import MatrixLUCI
from .tensortrain import AbstractTensorTrain, DeviceCore


# ---- index-set helpers (tensorci2.jl:512-529, util.jl:94-98, Base.union) --------------------------------
def kronecker_left(Iset, d):
    """[is..., j] for is in Iset (fastest), j in 1:d — tensorci2.jl:512-517."""
    n = Iset.shape[0]
    return np.concatenate([np.tile(Iset, (d, 1)), np.repeat(np.arange(1, d + 1, dtype=np.int32), n)[:, None]], axis=1)


def kronecker_right(d, Jset):
    """[i, js...] for i in 1:d (fastest), js in Jset — tensorci2.jl:524-529."""
    n = Jset.shape[0]
    return np.concatenate([np.tile(np.arange(1, d + 1, dtype=np.int32), n)[:, None], np.repeat(Jset, d, axis=0)], axis=1)


def _keys(tab):
    return [r.tobytes() for r in np.ascontiguousarray(tab, dtype=np.int32)]


def _first_seen_unique(tab):
    """Rows of `tab` with duplicates dropped, first occurrence kept, original order (Base.union / unique semantics)."""
    n, ln = tab.shape
    if n <= 1 or ln == 0:
        return tab[: min(n, 1)] if ln == 0 else tab
    t64 = np.ascontiguousarray(tab, dtype=np.int64)
    span = t64.max(axis=0) + 1                              # per-column radix: a mixed-radix key identifies a row
    if float(np.prod(span.astype(np.float64))) < 2.0 ** 62:
        w = np.concatenate([[1], np.cumprod(span[:-1])])
        _, first = np.unique(t64 @ w, return_index=True)   # integer sort: far cheaper than sorting the raw rows
    else:
        flat = np.ascontiguousarray(tab, dtype=np.int32).view(np.dtype((np.void, 4 * ln))).reshape(-1)
        _, first = np.unique(flat, return_index=True)
    if len(first) == n:
        return tab
    return tab[np.sort(first)]


def union_rows(a, b):
    """Base.union on vectors of multi-indices: first-seen order, duplicates dropped."""
    if b is None or b.shape[0] == 0:
        return _first_seen_unique(a)
    return _first_seen_unique(np.concatenate([a, b.reshape(-1, a.shape[1])], axis=0))


def _positions(sub, tab):
    """1-based positions in `tab` of the rows of `sub` that occur in it, in `sub` order
    (findfirst(isequal(i), Icombined) filtered for nothing — tensorci2.jl:866-867)."""
    where = {k: i + 1 for i, k in reversed(list(enumerate(_keys(tab))))}
    return [where[k] for k in _keys(sub) if k in where]


def pushunique(tab, row):
    row = np.asarray(row, dtype=np.int32).reshape(1, -1)
    if tab.shape[0] and any(k == row.tobytes() for k in _keys(tab)):
        return tab
    return np.concatenate([tab, row], axis=0)


def _jl_max(a, b):
    if np.isnan(a) or np.isnan(b):
        return float("nan")
    return a if a > b else b


def forwardsweep(sweepstrategy, iteration):                # sweepstrategies.jl:41-50
    return sweepstrategy == "forward" or (sweepstrategy == "backandforth" and iteration % 2 == 1)


class HostFunctionAdapter(BatchEvaluator):
    """BatchEvaluatorAdapter (batcheval.jl:32-107) for an arbitrary Python callable f(indexset): the USER's function
    runs where the user wrote it (the host); the values are uploaded and everything downstream runs on the GPU."""

    def __init__(self, f, localdims, handle=None):
        super().__init__(localdims, handle)
        self.f = f

    def _host_values(self, lt, rt, cents, vals):
        """The triple loop of _batchevaluate_dispatch (batcheval.jl:157-171): one call of the user's function per element."""
        nleft, nright = lt.shape[0], rt.shape[0]
        for j in range(nright):
            for c, cidx in enumerate(cents):
                mid = list(reversed(cidx))
                for i in range(nleft):
                    vals[j, c, i] = self.f(list(lt[i]) + mid + list(rt[j]))

    def evaluate_batch(self, left, right, M, out=None, stride_c=0, stride_j=0, maxabs=None):
        import itertools
        import torch
        lt, rt = as_table(left), as_table(right)
        nleft, nright = lt.shape[0], rt.shape[0]
        dev = f"cuda:{self.handle.device}"
        if nleft * nright == 0:
            return torch.zeros((0,) * (M + 2), dtype=torch.float64, device=dev), maxabs
        nl = lt.shape[1]
        cd = [int(d) for d in self.localdims[nl:nl + M]]
        ncent = int(np.prod(cd)) if cd else 1
        vals = np.empty((nright, ncent, nleft))
        cents = list(itertools.product(*[range(1, d + 1) for d in reversed(cd)]))
        self._host_values(lt, rt, cents, vals)
        sc = stride_c if stride_c > 0 else nleft
        sj = stride_j if stride_j > 0 else sc * ncent
        if out is None:
            out = torch.empty((nright,) + tuple(reversed(cd)) + (nleft,), dtype=torch.float64, device=dev)
        flat = out.reshape(-1)
        host = torch.from_numpy(vals)
        view = torch.as_strided(flat, (nright, ncent, nleft), (sj, sc, 1))
        view.copy_(host)
        if maxabs is not None:
            m = float(np.max(np.abs(vals))) if not np.isnan(vals).any() else float("nan")
            maxabs = _jl_max(abs(maxabs), m)
        return out, maxabs

    def __call__(self, *args):
        if len(args) == 1:
            return float(self.f([int(v) for v in args[0]]))
        return super().__call__(*args)


class ThreadedBatchEvaluator(HostFunctionAdapter):
    """ThreadedBatchEvaluator (batcheval.jl:247-308): the same element-by-element evaluation of a host callable, with the
    elements of a batch spread over a pool of host threads (the reference's `Threads.@threads` over `elements`, :300).  Like the
    reference's it only threads INSIDE one batch call.  Python threads run concurrently only while the callable releases the
    GIL (numpy / ctypes / I/O-bound functions), which is the case this class is for; an expensive pure function belongs in a
    device functor (`UserFunction`, tci_functor_from_source)."""

    def __init__(self, f, localdims, nthreads=None, handle=None):
        super().__init__(f, localdims, handle)
        import os
        self.nthreads = int(nthreads) if nthreads else max(1, len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1))

    def _host_values(self, lt, rt, cents, vals):
        from concurrent.futures import ThreadPoolExecutor
        nleft, nright, ncent = lt.shape[0], rt.shape[0], len(cents)
        mids = [list(reversed(c)) for c in cents]
        lrows, rrows = [list(r) for r in lt], [list(r) for r in rt]
        total = nright * ncent * nleft

        def work(lo, hi):                                    # a contiguous stretch of the (j, c, i) element list
            for e in range(lo, hi):
                i, rem = e % nleft, e // nleft
                c, j = rem % ncent, rem // ncent
                vals[j, c, i] = self.f(lrows[i] + mids[c] + rrows[j])
        nt = max(1, min(self.nthreads, total))
        step = (total + nt - 1) // nt
        with ThreadPoolExecutor(max_workers=nt) as pool:
            for fut in [pool.submit(work, lo, min(total, lo + step)) for lo in range(0, total, step)]:
                fut.result()                                 # re-raises the callable's exception, like @threads does


def makebatchevaluatable(f, localdims, handle=None):       # batcheval.jl:57
    return f if isinstance(f, BatchEvaluator) else HostFunctionAdapter(f, localdims, handle)


class TensorCI2(AbstractTensorTrain):
    """mutable struct TensorCI2{Float64} (tensorci2.jl:50-93)."""

    def __init__(self, localdims, f=None, initialpivots=None, Iset=None, Jset=None, handle=None):
        localdims = [int(d) for d in localdims]
        if len(localdims) <= 1:
            raise TCIError(_lib.TCI_ERR_INVALID, "localdims should have at least 2 elements!")
        self.handle = handle or getattr(f, "handle", None) or _lib.default_handle()
        self.device = f"cuda:{self.handle.device}"
        n = len(localdims)
        self.localdims = localdims
        self.Iset = [np.zeros((0, b), dtype=np.int32) for b in range(n)]
        self.Jset = [np.zeros((0, n - b - 1), dtype=np.int32) for b in range(n)]
        self._T = [None] * n                                # device cores; None = invalidated
        self.pivoterrors = np.zeros(0)
        self.bonderrors = np.zeros(n - 1)
        self.maxsamplevalue = 0.0
        self.Iset_history, self.Jset_history = [], []
        self.stats = {"nevals": 0, "t_eval": 0.0, "t_lu": 0.0, "t_factors": 0.0, "t_host": 0.0}
        self._pool = {}                                     # recycled Pi buffers (see _pool_mat)
        self._colcomm, self._mailpool = None, None          # column-sharded bonds (sweep2site with a communicator)
        self.shard_min_elems = int(os.environ.get("TCI_SHARD_MIN_ELEMS", 1 << 20))   # smaller Pi stay replicated
        from .matrixlu import Rng
        self.rng = Rng(0)                                   # stream behind the rook search's random rows/columns (seed: `seed=` of crossinterpolate2)
        if f is not None and Iset is not None:              # tensorci2.jl:123-137
            self.Iset = [as_table(s, b) for b, s in enumerate(Iset)]
            self.Jset = [as_table(s, n - b - 1) for b, s in enumerate(Jset)]
            piv = reconstractglobalpivotsfromijset(localdims, self.Iset, self.Jset)
            self._init_maxsample(f, piv)
        elif f is not None:                                 # tensorci2.jl:105-116
            if initialpivots is None:
                initialpivots = [[1] * n]
            self.addglobalpivots(initialpivots)
            self._init_maxsample(f, initialpivots)

    def _init_maxsample(self, f, pivots):
        f = makebatchevaluatable(f, self.localdims, self.handle)
        m = -np.inf
        for p in pivots:
            m = _jl_max(m, abs(f([int(v) for v in p])))
        self.maxsamplevalue = m
        if not abs(self.maxsamplevalue) > 0.0:
            raise TCIError(_lib.TCI_ERR_INVALID, "maxsamplevalue is zero!")
        self.invalidatesitetensors()

    # ---- AbstractTensorTrain plumbing
    def _dev_cores(self):
        if not self.issitetensorsavailable():
            raise TCIError(_lib.TCI_ERR_INVALID, "site tensors are not available (call fillsitetensors! or a sweep first)")
        return self._T

    def __len__(self):
        return len(self.localdims)

    @property
    def sitetensors(self):
        n = len(self)
        return [self._T[b].numpy() if self._T[b] is not None else np.zeros((0, self.localdims[b], 0)) for b in range(n)]

    def linkdims(self):                                     # tensorci2.jl:189-191
        return [self.Iset[b + 1].shape[0] for b in range(len(self) - 1)]

    def rank(self):
        return max(self.linkdims())

    def invalidatesitetensors(self):                        # tensorci2.jl:206-211
        self._T = [None] * len(self)

    def issitetensorsavailable(self):                       # tensorci2.jl:218-220
        return all(t is not None and t.chil * t.chir != 0 for t in self._T)

    # ---- errors (tensorci2.jl:231-289)
    def updateerrors(self, b0, errors):
        self.bonderrors[b0] = errors[-1]
        k = max(len(self.pivoterrors), len(errors))
        a = np.zeros(k); a[: len(self.pivoterrors)] = self.pivoterrors
        e = np.zeros(k); e[: len(errors)] = errors
        self.pivoterrors = np.maximum(a, e)                 # propagates NaN like Julia's max (padzero + max., :252-260)

    def flushpivoterror(self):
        self.pivoterrors = np.zeros(0)

    def pivoterror(self):                                   # tensorci2.jl:277-279, :243-245
        m = -np.inf
        for x in self.bonderrors:
            m = _jl_max(m, x)
        return m

    # ---- pivots
    def addglobalpivots(self, pivots):                      # tensorci2.jl:335-357
        n = len(self)
        pivots = [[int(v) for v in p] for p in pivots]
        if any(len(p) != n for p in pivots):
            raise TCIError(_lib.TCI_ERR_INVALID, "Please specify a pivot as one index per leg of the MPS.")
        for p in pivots:
            for b in range(n):
                self.Iset[b] = pushunique(self.Iset[b], p[:b])
                self.Jset[b] = pushunique(self.Jset[b], p[b + 1:])
        if pivots:
            self.invalidatesitetensors()

    # ---- K1 call sites -----------------------------------------------------------------------------------
    def _pool_mat(self, slot, rows, cols, pad=16):
        """Pi matrices are consumed (factorised in place) right after they are filled: recycle one buffer per slot instead
        of a fresh allocation per bond."""
        import torch
        ld = max((int(rows) + pad - 1) // pad * pad, 1)
        need = max(int(cols), 1) * ld
        buf = self._pool.get(slot)
        if buf is None or buf.numel() < need:
            buf = torch.empty(need, dtype=torch.float64, device=self.device)
            self._pool[slot] = buf
        return Mat(buf[:need].view(max(int(cols), 1), ld), rows, cols)

    def _fill(self, f, I, J, M, rows, cols, slot=None):
        """filltensor (tensorci2.jl:475-497) + updatemaxsample! (:636-638) into a device matrix rows x cols
        (column-major reshape of the (|I|, d..., |J|) array, exactly the reference's `reshape`).  `slot`: recycle the
        buffer of that pool slot (for matrices that do not outlive the next fill of the same slot)."""
        t0 = time.perf_counter()
        P = f.alloc_matrix(rows, cols, 16) if hasattr(f, "alloc_matrix") else None     # peer-mapped storage (sharded batch)
        if P is None:
            P = self._pool_mat(slot, rows, cols) if slot is not None else Mat.empty(rows, cols, self.device, pad=16)
        nI, nJ = I.shape[0], J.shape[0]
        if nI * nJ:
            if M == 0:
                sc, sj = nI, P.ld
            elif rows == nI:                                # rows = I, cols = (c, J)
                sc, sj = P.ld, P.ld * (cols // nJ)
            else:                                           # rows = (I, c), cols = J
                sc, sj = nI, P.ld
            _, self.maxsamplevalue = f.evaluate_batch(I, J, M, out=P.buf, stride_c=sc, stride_j=sj,
                                                      maxabs=self.maxsamplevalue)
            self.stats["nevals"] += rows * cols
        self.stats["t_eval"] += time.perf_counter() - t0
        return P

    def _set_core(self, b0, mat):
        """setsitetensor!(tci, b, T) (tensorci2.jl:536-545): reshape to (|Iset[b]|, d_b, |Jset[b]|)."""
        chil, d, chir = self.Iset[b0].shape[0], self.localdims[b0], self.Jset[b0].shape[0]
        if mat.rows * mat.cols != chil * d * chir:
            raise TCIError(_lib.TCI_ERR_INVALID, "setsitetensor!: size mismatch")
        if getattr(mat, "transient", False):                # lives in a recycled peer-mapped slot: keep a private copy
            keep = Mat.empty(mat.rows, mat.cols, self.device)
            keep.buf[: mat.cols, : mat.rows].copy_(mat.buf[: mat.cols, : mat.rows])
            mat = keep
        if mat.rows != chil * d:                            # r x (d*chi) matrix: same memory, other 2-D view
            if mat.ld != mat.rows:
                mat = Mat.from_numpy(mat.numpy().reshape(chil * d, chir, order="F"), self.device)
            else:
                mat = Mat(mat.buf.reshape(-1)[: chil * d * chir].reshape(chir, chil * d), chil * d, chir)
        self._T[b0] = DeviceCore(mat, chil, (d,), chir)

    def updatepivots(self, b, f, leftorthogonal, reltol=1e-14, abstol=0.0, maxbonddim=INT_MAX, sweepdirection="forward",
                     pivotsearch="full", verbosity=0, extraIset=None, extraJset=None):
        """updatepivots! (tensorci2.jl:825-930): pivotsearch=:full fills the whole Pi matrix; :rook samples strips of it
        through a SubMatrix (:764-804) and never holds Pi."""
        if pivotsearch not in ("full", "rook"):
            raise TCIError(_lib.TCI_ERR_INVALID, f"Unknown pivot search strategy {pivotsearch}. Choose from :rook, :full.")
        f = makebatchevaluatable(f, self.localdims, self.handle)
        b0 = b - 1
        self.invalidatesitetensors()
        th = time.perf_counter()
        noextra = (extraIset is None or len(extraIset) == 0) and (extraJset is None or len(extraJset) == 0)
        Ic = union_rows(kronecker_left(self.Iset[b0], self.localdims[b0]), None if extraIset is None else as_table(extraIset, b0 + 1))
        Jc = union_rows(kronecker_right(self.localdims[b0 + 1], self.Jset[b0 + 1]),
                        None if extraJset is None else as_table(extraJset, len(self) - b0 - 1))
        self.stats["t_host"] += time.perf_counter() - th
        luci = None
        if pivotsearch == "rook":                           # :863-905
            I0 = _positions(self.Iset[b0 + 1], Ic)          # :866-867, current pivots that survive in the combined sets
            J0 = _positions(self.Jset[b0], Jc)

            def pif(irows, icols):                          # SubMatrix (:764-804): strips of Pi on demand, on the device
                return self._fill(f, Ic[irows - 1], Jc[icols - 1], 0, len(irows), len(icols), slot=0)

            t0 = time.perf_counter()
            te = self.stats["t_eval"]
            luci = MatrixLUCI.from_function(pif, (Ic.shape[0], Jc.shape[0]), I0, J0, reltol=reltol, abstol=abstol,
                                            maxrank=maxbonddim, leftorthogonal=leftorthogonal, pivotsearch="rook",
                                            usebatcheval=True, rng=self.rng, handle=self.handle)
            self.stats["t_lu"] += time.perf_counter() - t0 - (self.stats["t_eval"] - te)
            if luci.npivots() == 0:                         # :886-899 fall back to the full search
                luci = None
        cc = self._colcomm
        if luci is None and cc is not None and Jc.shape[0] >= 2 * cc.size \
                and Ic.shape[0] * Jc.shape[0] >= self.shard_min_elems:
            luci = self._luci_colsharded(f, Ic, Jc, cc, reltol, abstol, maxbonddim, leftorthogonal)
            t0 = time.perf_counter()
        if luci is None:
            Pi = self._fill(f, Ic, Jc, 0, Ic.shape[0], Jc.shape[0], slot=0)
            t0 = time.perf_counter()
            luci = MatrixLUCI(Pi, reltol=reltol, abstol=abstol, maxrank=maxbonddim, leftorthogonal=leftorthogonal,
                              handle=self.handle)
            self.stats["t_lu"] += time.perf_counter() - t0
            t0 = time.perf_counter()
        if verbosity > 2:
            print(f"    Computing Pi ({Ic.shape[0]} x {Jc.shape[0]}) at bond {b}: LU: {time.perf_counter() - t0} sec")
        self.Iset[b0 + 1] = Ic[luci.rowindices() - 1]
        self.Jset[b0] = Jc[luci.colindices() - 1]
        if noextra:
            t0 = time.perf_counter()
            self._set_core(b0, luci.left_mat())
            self._set_core(b0 + 1, luci.right_mat())
            self.stats["t_factors"] += time.perf_counter() - t0
        errs = luci.pivoterrors()
        self.updateerrors(b0, errs)
        self.last_luci = luci
        return errs

    def _luci_colsharded(self, f, Ic, Jc, comm, reltol, abstol, maxbonddim, leftorthogonal):
        """The full-search branch of updatepivots! with the COLUMNS of Pi spread over the ranks of `comm`:
        every rank evaluates only its column block (K1, no exchange), the blocks are factorised in place by the
        column-sharded rrLU (one candidate record per rank and pivot step through the peer mailboxes) and only the
        r x n factor U is gathered.  Pi is never assembled on one GPU; pivots, L, U are bit-identical to the one-GPU rrLU."""
        from .matrixlu import rrLU
        from .parallel import MailboxPool, gather_sharded_U, partition, rrlu_sharded
        nI, nJ = Ic.shape[0], Jc.shape[0]
        blocks = partition(nJ, comm.size)
        s, e = blocks[comm.rank]
        Pl = self._fill(f, Ic, Jc[s:e], 0, nI, e - s, slot=0)
        t0 = time.perf_counter()
        if self._mailpool is None:
            self._mailpool = MailboxPool(self.handle, comm)
        out = rrlu_sharded(Pl, nJ, s, comm, maxrank=maxbonddim, reltol=reltol, abstol=abstol, leftorthogonal=leftorthogonal,
                           handle=self.handle, fused=True, peers=self._mailpool)
        U = gather_sharded_U(out, nJ, blocks, comm)
        L = out["L"]
        lu = rrLU(out["rowperm"], out["colperm"], L.view(), U.view(), leftorthogonal, out["npivot"], out["error"],
                  out["pivoterrors"])
        lu.Lmat, lu.Umat = L, U
        self.stats["t_lu"] += time.perf_counter() - t0
        self.stats["nsharded"] = self.stats.get("nsharded", 0) + 1
        return MatrixLUCI.from_rrlu(lu, nI, nJ, self.handle)

    def close_parallel(self):
        """Release the peer-mapped mailboxes (collective: every rank calls it)."""
        if self._mailpool is not None:
            self._mailpool.close()
            self._mailpool = None

    def setsitetensor(self, f, b, leftorthogonal=True):
        """setsitetensor!(tci, f, b) (tensorci2.jl:599-629)."""
        if not leftorthogonal:
            raise TCIError(_lib.TCI_ERR_INVALID, "leftorthogonal==false is not supported!")
        f = makebatchevaluatable(f, self.localdims, self.handle)
        b0 = b - 1
        nI, nJ, d = self.Iset[b0].shape[0], self.Jset[b0].shape[0], self.localdims[b0]
        Pi1 = self._fill(f, self.Iset[b0], self.Jset[b0], 1, nI * d, nJ, slot=None if b0 == len(self) - 1 else 1)
        if b0 == len(self) - 1:
            self._set_core(b0, Pi1)
            return
        nI1 = self.Iset[b0 + 1].shape[0]
        P = self._fill(f, self.Iset[b0 + 1], self.Jset[b0], 0, nI1, nJ, slot=2)
        if nI1 != nJ:
            raise TCIError(_lib.TCI_ERR_NOT_SQUARE, f"Pivot matrix at bond {b} is not square!")
        t0 = time.perf_counter()
        T = solve_right(self.handle, Pi1, P)
        self.stats["t_factors"] += time.perf_counter() - t0
        self._T[b0] = DeviceCore(T, nI, (d,), nI1)

    def fillsitetensors(self, f, comm=None):               # globalsearch.jl:202-208
        n = len(self)
        if comm is None or comm.size == 1:
            for b in range(1, n + 1):
                self.setsitetensor(f, b)
            return
        # sites are independent (each reads only the index sets): deal them over the ranks, broadcast the tensors
        import torch
        for b in range(1, n + 1):
            if (b - 1) % comm.size == comm.rank:
                self.setsitetensor(f, b)
        for b0 in range(n):
            chil, d, chir = self.Iset[b0].shape[0], self.localdims[b0], self.Jset[b0].shape[0]
            src = b0 % comm.size
            if src == comm.rank:
                m = self._T[b0].mat
                dense = m.buf[: m.cols, : m.rows].contiguous()
            else:
                dense = torch.empty((chir, chil * d), dtype=torch.float64, device=self.device)
            comm.broadcast(dense, src)
            if src != comm.rank:
                self._T[b0] = DeviceCore(Mat(dense, chil * d, chir), chil, (d,), chir)
        self.maxsamplevalue = comm.allreduce_max(self.maxsamplevalue)

    def sweep1site(self, f, sweepdirection="forward", reltol=1e-14, abstol=0.0, maxbonddim=INT_MAX, updatetensors=True):
        """sweep1site! (tensorci2.jl:659-725)."""
        import torch
        f = makebatchevaluatable(f, self.localdims, self.handle)
        self.flushpivoterror()
        self.invalidatesitetensors()
        if sweepdirection not in ("forward", "backward"):
            raise TCIError(_lib.TCI_ERR_INVALID, f"Unknown sweep direction {sweepdirection}: choose between :forward, :backward.")
        fwd = sweepdirection == "forward"
        n = len(self)
        for b0 in (range(0, n - 1) if fwd else range(n - 1, 0, -1)):
            d = self.localdims[b0]
            Is = kronecker_left(self.Iset[b0], d) if fwd else self.Iset[b0]
            Js = self.Jset[b0] if fwd else kronecker_right(d, self.Jset[b0])
            Pi = self._fill(f, self.Iset[b0], self.Jset[b0], 1, Is.shape[0], Js.shape[0], slot=0)
            t0 = time.perf_counter()
            luci = MatrixLUCI(Pi, reltol=reltol, abstol=abstol, maxrank=maxbonddim, leftorthogonal=fwd, handle=self.handle)
            self.stats["t_lu"] += time.perf_counter() - t0
            if fwd:
                self.Iset[b0 + 1] = Is[luci.rowindices() - 1]
                self.Jset[b0] = Js[luci.colindices() - 1]
            else:
                self.Iset[b0] = Is[luci.rowindices() - 1]
                self.Jset[b0 - 1] = Js[luci.colindices() - 1]
            if updatetensors:
                t0 = time.perf_counter()
                self._set_core(b0, luci.left_mat() if fwd else luci.right_mat())
                if bool(torch.isnan(self._T[b0].mat.view()).any()):
                    raise TCIError(_lib.TCI_ERR_NAN_TENSOR, f"Error: NaN in tensor T[{b0 + 1}]")
                self.stats["t_factors"] += time.perf_counter() - t0
            self.updateerrors(b0 if fwd else b0 - 1, luci.pivoterrors())
        if updatetensors:
            last = n - 1 if fwd else 0
            rows = self.Iset[last].shape[0] if fwd else self.localdims[0]
            cols = self.localdims[last] if fwd else self.Jset[0].shape[0]
            self._set_core(last, self._fill(f, self.Iset[last], self.Jset[last], 1, rows, cols))

    def sweep0site(self, f, b, reltol=1e-14, abstol=0.0):
        """sweep0site! (tensorci2.jl:559-587): remove bad pivots at bond b (1-based) — fill the pivot matrix
        P = f(Iset[b+1], Jset[b]), factorise it (MatrixLUCI, leftorthogonal), keep as many leading pivots as there are diagonal
        entries of U above both thresholds (:580 counts over the whole diagonal)."""
        f = makebatchevaluatable(f, self.localdims, self.handle)
        if not 1 <= b < len(self):
            raise TCIError(_lib.TCI_ERR_INVALID, f"sweep0site!: bond {b} out of range 1:{len(self) - 1}")
        self.invalidatesitetensors()
        I, J = self.Iset[b], self.Jset[b - 1]
        P = self._fill(f, I, J, 0, I.shape[0], J.shape[0], slot=0)
        luci = MatrixLUCI(P, reltol=reltol, abstol=abstol, leftorthogonal=True, handle=self.handle)
        diag = np.asarray(luci.pivoterrors(), dtype=np.float64)[:luci.npivots()]       # |U[i, i]| (leftorthogonal: U carries the pivots)
        ndiag = int(np.count_nonzero((diag > abstol) & (diag / diag[0] > reltol))) if diag.size else 0
        self.Iset[b] = I[luci.rowindices()[:ndiag] - 1]
        self.Jset[b - 1] = J[luci.colindices()[:ndiag] - 1]

    rmbadpivots = sweep0site                                                        # tensorci2.jl:589

    def makecanonical(self, f, reltol=1e-14, abstol=0.0, maxbonddim=INT_MAX):      # tensorci2.jl:738-749
        self.sweep1site(f, "forward", reltol=0.0, abstol=0.0, maxbonddim=INT_MAX, updatetensors=False)
        self.sweep1site(f, "backward", reltol=reltol, abstol=abstol, maxbonddim=maxbonddim, updatetensors=False)
        self.sweep1site(f, "forward", reltol=reltol, abstol=abstol, maxbonddim=maxbonddim, updatetensors=True)

    def addglobalpivots1sitesweep(self, f, pivots, reltol=1e-14, abstol=0.0, maxbonddim=INT_MAX):   # :367-377
        self.addglobalpivots(pivots)
        self.makecanonical(f, reltol=reltol, abstol=abstol, maxbonddim=maxbonddim)

    def addglobalpivots2sitesweep(self, f, pivots, tolerance=1e-8, normalizeerror=True, maxbonddim=INT_MAX, pivotsearch="full",
                                  verbosity=0, ntry=10, strictlynested=False):
        """addglobalpivots2sitesweep! (tensorci2.jl:407-453): add the pivots, run two 2-site half-sweeps, repeat with the
        pivots that are still not interpolated within tolerance.  Returns how many remain."""
        pivots = [[int(v) for v in p] for p in pivots]
        if any(len(p) != len(self) for p in pivots):
            raise TCIError(_lib.TCI_ERR_INVALID, "Please specify a pivot as one index per leg of the MPS.")
        f = makebatchevaluatable(f, self.localdims, self.handle)
        pivots_ = pivots
        for _ in range(ntry):
            errornormalization = self.maxsamplevalue if normalizeerror else 1.0
            abstol = tolerance * errornormalization
            self.addglobalpivots(pivots_)
            self.sweep2site(f, 2, abstol=abstol, maxbonddim=maxbonddim, pivotsearch=pivotsearch, strictlynested=strictlynested,
                            verbosity=verbosity)
            if pivots:
                pts = np.asarray(pivots, dtype=np.int32)
                err = np.abs(self.evaluate_batch(pts) - _point_values(f, pts))      # one K4b + one K1 call for all pivots
                newpivots = [p for p, e in zip(pivots, err) if e > abstol]
            else:
                newpivots = []
            if verbosity > 0:
                print(f"Trying to add {len(pivots_)} global pivots, {len(newpivots)} still remain.")
            if len(newpivots) == 0 or set(map(tuple, newpivots)) == set(map(tuple, pivots_)):
                return len(newpivots)
            pivots_ = newpivots
        return len(pivots_)

    def sweep2site(self, f, niter, iter1=1, abstol=1e-8, maxbonddim=INT_MAX, sweepstrategy="backandforth",
                   pivotsearch="full", verbosity=0, strictlynested=False, fillsitetensors=True, bondorder="sequential",
                   comm=None):
        """sweep2site! (tensorci2.jl:1195-1258).  bondorder='oddeven' (+ optional `comm`) is the parallel-bond variant
        (parallel.py); the default is the reference's sequential order, and with a communicator of several ranks that order
        is kept: every rank visits every bond and each Pi of at least `shard_min_elems` entries is evaluated and
        factorised column-sharded (`_luci_colsharded`), the smaller ones redundantly on every rank."""
        from .parallel import bond_passes, run_pass
        f = makebatchevaluatable(f, self.localdims, self.handle)
        colsharded = comm is not None and comm.size > 1 and bondorder == "sequential"
        if colsharded and pivotsearch != "full":
            raise TCIError(_lib.TCI_ERR_INVALID, "with several ranks the sequential bond order shards the columns of Pi, "
                           "which needs pivotsearch=:full")
        self._colcomm = comm if colsharded else None       # every rank visits every bond; large Pi are column-sharded
        self.invalidatesitetensors()
        n = len(self)
        for it in range(iter1, iter1 + niter):
            extraI = [None] * n
            extraJ = [None] * n
            if not strictlynested and self.Iset_history:
                extraI, extraJ = self.Iset_history[-1], self.Jset_history[-1]
            self.Iset_history.append([s.copy() for s in self.Iset])
            self.Jset_history.append([s.copy() for s in self.Jset])
            self.flushpivoterror()
            fwd = forwardsweep(sweepstrategy, it)

            def update(b):
                return self.updatepivots(b, f, fwd, abstol=abstol, maxbonddim=maxbonddim,
                                         sweepdirection="forward" if fwd else "backward", pivotsearch=pivotsearch,
                                         verbosity=verbosity, extraIset=extraI[b], extraJset=extraJ[b - 1])

            for bonds in bond_passes(n - 1, fwd, bondorder):
                run_pass(self, bonds, update, None if colsharded else comm)
        if colsharded:
            self._colcomm = None
            self.maxsamplevalue = comm.allreduce_max(self.maxsamplevalue)
        if fillsitetensors:
            self.fillsitetensors(f, comm)


def reconstractglobalpivotsfromijset(localdims, Isets, Jsets):     # tensorci2.jl:303-320
    piv, seen = [], set()
    for i in range(len(Isets)):
        for I in Isets[i]:
            for J in Jsets[i]:
                for j in range(1, localdims[i] + 1):
                    p = tuple(int(v) for v in I) + (j,) + tuple(int(v) for v in J)
                    if p not in seen:
                        seen.add(p); piv.append(list(p))
    return piv


ENGINES = ("python", "native", "device")


def convergencecriterion(ranks, errors, nglobalpivots, tolerance, maxbonddim, ncheckhistory, checkconvglobalpivot=True):
    """tensorci2.jl:947-966."""
    if len(errors) < ncheckhistory:
        return False
    lastranks = list(ranks)[-ncheckhistory:]
    lastng = list(nglobalpivots)[-ncheckhistory:]
    return ((all(e < tolerance for e in list(errors)[-ncheckhistory:])
             and (all(g == 0 for g in lastng) if checkconvglobalpivot else True)
             and min(lastranks) == lastranks[-1])
            or all(r >= maxbonddim for r in lastranks))


class GlobalPivotSearchInput:
    """globalpivotfinder.jl:33-68."""

    def __init__(self, tci):
        from .tensortrain import TensorTrain
        self.localdims = list(tci.localdims)
        self.current_tt = TensorTrain(tci)
        self.maxsamplevalue = tci.maxsamplevalue
        self.Iset, self.Jset = tci.Iset, tci.Jset


class DefaultGlobalPivotFinder:
    """DefaultGlobalPivotFinder (globalpivotfinder.jl:145-265): random starts, one greedy line search per site on
    |f - tt|.  The host keeps the RNG and the greedy control flow; all the |f - tt| values of one start
    (sum_p d_p points) are evaluated in ONE K1 batch and ONE K4b batch instead of point by point.
    (numpy's RNG replaces Julia's Xoshiro: global pivots are not comparable with a Julia run — SURVEY.md 7.)"""

    def __init__(self, nsearch=5, maxnglobalpivot=5, tolmarginglobalsearch=10.0):
        self.nsearch, self.maxnglobalpivot, self.tolmarginglobalsearch = nsearch, maxnglobalpivot, tolmarginglobalsearch

    def __call__(self, input, f, abstol, verbosity=0, rng=None):
        rng = rng or np.random.default_rng()
        L = len(input.localdims)
        found = []
        if self.nsearch <= 0:
            return found
        if hasattr(rng, "randint") and hasattr(rng, "next"):       # matrixlu.Rng: the seeded stream the compiled driver and the CPU checker use
            starts = [[1 + int(rng.randint(input.localdims[p])) for p in range(L)] for _ in range(self.nsearch)]
        else:
            starts = [[int(rng.integers(1, input.localdims[p] + 1)) for p in range(L)] for _ in range(self.nsearch)]
        pts = []
        for point in starts:                                # globalpivotfinder.jl:223-246: vary one site at a time
            for p in range(L):
                for v in range(1, input.localdims[p] + 1):
                    q = list(point); q[p] = v
                    pts.append(q)
        pts = np.asarray(pts, dtype=np.int32)
        fv = np.array([f(list(q)) for q in pts]) if not isinstance(f, BatchEvaluator) else _point_values(f, pts)
        tv = input.current_tt.evaluate_batch(pts)
        err = np.abs(fv - tv)
        per = sum(input.localdims)
        for s, point in enumerate(starts):
            e = err[s * per:(s + 1) * per]
            best_error, best_point = 0.0, list(point)
            for k in range(per):                            # strict '>' keeps the first maximum, as the reference loop does
                if e[k] > best_error:
                    best_error, best_point = float(e[k]), [int(v) for v in pts[s * per + k]]
            if best_error > abstol * self.tolmarginglobalsearch:
                found.append(best_point)
        if len(found) > self.maxnglobalpivot:
            found = found[: self.maxnglobalpivot]
        if verbosity > 0:
            print(f"Found {len(found)} global pivots")
        return found


def _point_values(f, pts):
    return f.evaluate_points(pts)


def optimize(tci, f, tolerance=None, pivottolerance=None, maxbonddim=INT_MAX, maxiter=20, sweepstrategy="backandforth",
             pivotsearch="full", verbosity=0, loginterval=10, normalizeerror=True, ncheckhistory=3, globalpivotfinder=None,
             maxnglobalpivot=5, nsearchglobalpivot=5, tolmarginglobalsearch=10.0, strictlynested=False,
             checkbatchevaluatable=False, checkconvglobalpivot=True, rng=None, bondorder="sequential", comm=None):
    """optimize! (tensorci2.jl:1018-1172).  Returns (ranks, errors ./ errornormalization)."""
    import warnings
    errors, ranks, nglobalpivots = [], [], []
    if checkbatchevaluatable and not isinstance(f, BatchEvaluator):
        raise TCIError(_lib.TCI_ERR_INVALID, "Function `f` is not batch evaluatable")
    if 0 < nsearchglobalpivot < maxnglobalpivot:
        raise TCIError(_lib.TCI_ERR_INVALID, "nsearchglobalpivot < maxnglobalpivot!")
    if pivottolerance is not None:
        if tolerance is not None and tolerance != pivottolerance:
            raise TCIError(_lib.TCI_ERR_INVALID, "Got different values for pivottolerance and tolerance in optimize!(TCI2). "
                           "For TCI2, both of these options have the same meaning. Please assign only `tolerance`.")
        warnings.warn("The option `pivottolerance` of `optimize!(tci::TensorCI2, f)` is deprecated. Please update your code "
                      "to use `tolerance`, as `pivottolerance` will be removed in the future.", DeprecationWarning)
        tol = pivottolerance
    elif tolerance is not None:
        tol = tolerance
    else:
        tol = 1e-8
    if maxbonddim >= INT_MAX and tol <= 0:
        raise TCIError(_lib.TCI_ERR_INVALID, "Specify either tolerance > 0 or some maxbonddim; otherwise, the convergence "
                       "criterion is not reachable!")
    f = makebatchevaluatable(f, tci.localdims, tci.handle)
    finder = globalpivotfinder or DefaultGlobalPivotFinder(nsearch=nsearchglobalpivot, maxnglobalpivot=maxnglobalpivot,
                                                           tolmarginglobalsearch=tolmarginglobalsearch)
    tstart = time.perf_counter()
    for it in range(1, maxiter + 1):
        errornormalization = tci.maxsamplevalue if normalizeerror else 1.0
        abstol = tol * errornormalization
        if verbosity > 1:
            print(f"  Walltime {time.perf_counter() - tstart} sec: starting 2site sweep")
        tci.sweep2site(f, 2, iter1=1, abstol=abstol, maxbonddim=maxbonddim, pivotsearch=pivotsearch,
                       strictlynested=strictlynested, verbosity=verbosity, sweepstrategy=sweepstrategy, fillsitetensors=True,
                       bondorder=bondorder, comm=comm)
        errors.append(tci.pivoterror())
        globalpivots = finder(GlobalPivotSearchInput(tci), f, abstol, verbosity=verbosity, rng=rng) \
            if getattr(finder, "nsearch", 1) != 0 else []
        tci.addglobalpivots(globalpivots)
        nglobalpivots.append(len(globalpivots))
        ranks.append(tci.rank())
        if verbosity > 0 and it % loginterval == 0:
            print(f"iteration = {it}, rank = {ranks[-1]}, error= {errors[-1]}, maxsamplevalue= {tci.maxsamplevalue}, "
                  f"nglobalpivot={len(globalpivots)}")
        if convergencecriterion(ranks, errors, nglobalpivots, abstol, maxbonddim, ncheckhistory,
                                checkconvglobalpivot=checkconvglobalpivot):
            break
    errornormalization = tci.maxsamplevalue if normalizeerror else 1.0
    abstol = tol * errornormalization
    tci.sweep1site(f, abstol=abstol, maxbonddim=maxbonddim)
    if comm is not None:
        tci.close_parallel()
    _sanitycheck(tci)
    return np.asarray(ranks, dtype=np.int64), np.asarray(errors) / errornormalization


def optfirstpivot(f, localdims, firstpivot=None, maxsweep=1000):
    """optfirstpivot (util.jl:260-291): greedy one-site-at-a-time search for a large |f|.  All d_i candidates of a site go
    through ONE K1 call; taking the first maximum that beats the current value is exactly the reference's sequential
    strict-`>` scan (the other coordinates are fixed while site i is scanned)."""
    localdims = [int(d) for d in localdims]
    n = len(localdims)
    f = makebatchevaluatable(f, localdims)
    pivot = [1] * n if firstpivot is None else [int(v) for v in firstpivot]
    if isinstance(f, BatchEvaluator) and getattr(f, "_f", None):            # device functor: the same scan as one C call
        import ctypes as C
        ld, pv = np.asarray(localdims, dtype=np.int32), np.asarray(pivot, dtype=np.int32)
        f.handle.check(_lib.lib().tci_optfirstpivot_f64(f.handle.h, f._f, _lib.p_i32(ld), n, _lib.p_i32(pv), C.c_int(int(maxsweep))))
        return [int(v) for v in pv]
    valf = abs(f(pivot))
    for _ in range(maxsweep):
        valf_prev = valf
        for i in range(n):
            pts = np.tile(np.asarray(pivot, dtype=np.int32), (localdims[i], 1))
            pts[:, i] = np.arange(1, localdims[i] + 1)
            vals = np.abs(_point_values(f, pts)) if isinstance(f, BatchEvaluator) and not isinstance(f, HostFunctionAdapter) \
                else np.array([abs(f(list(q))) for q in pts])
            for d in range(localdims[i]):
                if vals[d] > valf:
                    valf = float(vals[d]); pivot[i] = d + 1
        if valf_prev == valf:
            break
    return pivot


def _sanitycheck(tci):                                      # globalsearch.jl:226-233
    for b in range(len(tci) - 1):
        if tci.Iset[b + 1].shape[0] != tci.Jset[b].shape[0]:
            raise TCIError(_lib.TCI_ERR_NOT_SQUARE, f"Pivot matrix at bond {b + 1} is not square!")
    return True


_NATIVE_KW = {"tolerance", "maxbonddim", "maxiter", "sweepstrategy", "normalizeerror", "ncheckhistory", "nsearchglobalpivot",
              "maxnglobalpivot", "tolmarginglobalsearch", "strictlynested", "checkconvglobalpivot", "pivotsearch", "verbosity",
              "loginterval", "checkbatchevaluatable"}


def _native_eligible(f, kwargs):
    """engine="auto": does the compiled driver cover this call?  Device functor evaluated by the stock batch path (wrappers
    that override evaluate_batch — sharded, cached, contraction, host adapters — stay on the Python host), only keywords the
    driver implements, nothing to print."""
    return (isinstance(f, BatchEvaluator) and bool(getattr(f, "_f", None))
            and type(f).evaluate_batch is BatchEvaluator.evaluate_batch
            and set(kwargs) <= (_NATIVE_KW | {"handle"})
            and kwargs.get("pivotsearch", "full") in ("full", "rook")
            and not kwargs.get("verbosity", 0)
            and not kwargs.get("checkbatchevaluatable", False))


def _crossinterpolate2_native(f, localdims, initialpivots, seed, handle, kwargs, device=False):
    """crossinterpolate2 with the whole optimize! loop run by the compiled sweep driver (csrc/tci2_driver.cpp, native.py): one
    C call instead of an interpreted host loop over bonds.  Same kernels, same order of operations; the result is handed
    back as an ordinary TensorCI2 (index sets on the host, site tensors in torch-owned device memory)."""
    import torch
    from .native import NativeTCI2
    from .parallel import _RawCuda
    unknown = set(kwargs) - _NATIVE_KW
    if unknown:
        raise TCIError(_lib.TCI_ERR_INVALID, f"engine='native' does not take {sorted(unknown)} (use engine='python')")
    pivotsearch = kwargs.pop("pivotsearch", "full")
    if pivotsearch not in ("full", "rook"):
        raise TCIError(_lib.TCI_ERR_INVALID, f"Unknown pivot search strategy {pivotsearch}. Choose from :rook, :full.")
    if not isinstance(f, BatchEvaluator) or not getattr(f, "_f", None):
        raise TCIError(_lib.TCI_ERR_INVALID, "engine='native' needs a device functor (a BatchEvaluator built on tci_functor_*); "
                       "host callables run with engine='python'")
    for k in ("verbosity", "loginterval", "checkbatchevaluatable"):
        kwargs.pop(k, None)
    h = handle or f.handle
    h.use_torch_stream()
    localdims = [int(d) for d in localdims]
    n = len(localdims)
    if initialpivots is None:
        initialpivots = [[1] * n]
    nat = NativeTCI2(_lib.lib(), h.h, localdims)
    if device:
        nat.set_engine("device")
    if seed is not None:
        nat.seed(seed)
    nat.set_pivotsearch(pivotsearch, 0 if seed is None else seed)        # same default stream as TensorCI2.rng = Rng(0)
    nat.addglobalpivots(initialpivots)
    nat.init_maxsample(f._f, initialpivots)
    if kwargs.get("tolerance") is None:
        kwargs["tolerance"] = 1e-8
    ranks, errors = nat.optimize(f._f, **kwargs)
    tci = TensorCI2(localdims, handle=h)
    tci.Iset, tci.Jset = nat.export_indexsets()
    tci.bonderrors, tci.pivoterrors, tci.maxsamplevalue = nat.errors()
    dev = f"cuda:{h.device}"

    def alloc(ndoubles):                                    # one torch-owned buffer for all site tensors (the driver's die with `nat`)
        buf = torch.empty(int(ndoubles), dtype=torch.float64, device=dev)
        return buf.data_ptr(), buf
    flat, shapes, offs = nat.export_cores(alloc)
    for b in range(n):
        chil, d, chir = (int(v) for v in shapes[b])
        o = int(offs[b])
        tci._T[b] = DeviceCore(Mat(flat[o:o + chil * d * chir].view(chir, chil * d), chil * d, chir), chil, (d,), chir)
    st = nat.stats()
    tci.stats.update({k: st[k] for k in ("nevals", "t_eval", "t_lu", "t_factors", "t_host")})
    tci.stats["nbondupdates"] = st["nbondupdates"]
    nat.close()
    return tci, ranks, errors


def crossinterpolate2(f, localdims, initialpivots=None, **kwargs):
    """crossinterpolate2(Float64, f, localdims, initialpivots; kwargs...) — tensorci2.jl:1313-1323.
    `engine` (not in the reference): "native" hands the optimize! loop to the compiled sweep driver (fewer microseconds of host
    work per bond, identical kernels and results), "device" is that driver with the bond loop of every half-sweep running
    inside ONE kernel launch wherever the Pi matrices fit shared memory (csrc/sweep.cu), "python" runs the loop in this module;
    the default "auto" takes the compiled driver whenever it covers the call (device functor, one GPU, no Python-side finder /
    rng / verbosity)."""
    seed = kwargs.pop("seed", None)                         # not in the reference (it uses Julia's global RNG): starts the two
                                                            # splitmix64 streams (rook columns, global-pivot starts) for every engine;
                                                            # without it the Python engine's finder draws from an unseeded numpy RNG
    engine = kwargs.pop("engine", "auto")
    if engine == "auto":                                    # the compiled driver whenever it covers the call, else this module;
        engine = "python"                                   # with the bond loop on the device for the built-in functors
        if _native_eligible(f, kwargs):
            h = kwargs.get("handle") or f.handle
            engine = "device" if (kwargs.get("pivotsearch", "full") == "full" and _lib.lib().tci_dsweep_supported(h.h, f._f)) else "native"
    if engine in ("native", "device"):
        return _crossinterpolate2_native(f, localdims, initialpivots, seed, kwargs.pop("handle", None), kwargs, device=engine == "device")
    if engine != "python":
        raise TCIError(_lib.TCI_ERR_INVALID, f"Unknown engine {engine}: choose 'python', 'native' or 'device'.")
    tci = TensorCI2(localdims, f, initialpivots, handle=kwargs.pop("handle", None))
    if seed is not None:
        from .matrixlu import Rng
        tci.rng = Rng(seed)
        if kwargs.get("rng") is None:                       # `seed` means the same for every engine: it also starts the stream of
            kwargs["rng"] = Rng(seed)                       # the global pivot finder's random starts (tci_tci2_seed in the driver)
    ranks, errors = optimize(tci, f, **kwargs)
    return tci, ranks, errors
