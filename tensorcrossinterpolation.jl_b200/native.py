"""ctypes front end of the compiled sweep driver (csrc/tci2_driver.cpp, `tci_tci2_*` in include/tci_b200.h).

`NativeTCI2` is the `TensorCI2{Float64}` state held by the library: index sets on the host, site tensors in device
memory, and the reference's sweeps (`sweep2site!`, `sweep1site!`, `optimize!`, ... — src/tensorci2.jl) as single C
calls, so that a half-sweep over small Pi matrices does not pay an interpreter's bookkeeping per bond.  Every number
still comes from the same kernels as the Python-driven path (`tensorci2.py`).  The class takes the ctypes library
and the raw handle explicitly: the product passes libtci_b200.so; the CPU test of the driver's bookkeeping passes the
same driver source linked against a host-memory stand-in of the kernels (tests/native_mock/)."""
import ctypes as C

import numpy as np

from ._lib import TCIError, p_f64, p_i32, p_i64

STRATEGY = {"forward": 0, "backandforth": 1, "backward": 2}
_UNBOUNDED = np.iinfo(np.int64).max


def _mbd(maxbonddim):
    """typemax(Int) -> 0 (the C side's 'unbounded')."""
    return 0 if maxbonddim is None or maxbonddim >= _UNBOUNDED else int(maxbonddim)


def _table(rows, length):
    a = np.asarray(rows, dtype=np.int32)
    n = a.shape[0] if a.ndim == 2 else len(rows)
    return np.ascontiguousarray(a.reshape(n, length)), n


class NativeTCI2:
    def __init__(self, L, handle_ptr, localdims):
        self.L, self.localdims = L, np.ascontiguousarray(localdims, dtype=np.int32)
        L.tci_tci2_last_error.restype = C.c_char_p
        L.tci_tci2_last_error.argtypes = [C.c_void_p]
        self._t = C.c_void_p()
        if len(self.localdims) <= 1:
            raise TCIError(-1, "localdims should have at least 2 elements!")
        rc = L.tci_tci2_create(handle_ptr, p_i32(self.localdims), len(self.localdims), C.byref(self._t))
        if rc != 0:
            raise TCIError(rc, "tci_tci2_create failed")

    def __len__(self):
        return len(self.localdims)

    def close(self):
        if getattr(self, "_t", None):
            self.L.tci_tci2_destroy(self._t)
            self._t = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            raise TCIError(rc, self.L.tci_tci2_last_error(self._t).decode())

    # ---- state
    def seed(self, seed):
        self._check(self.L.tci_tci2_seed(self._t, C.c_uint64(int(seed) & ((1 << 64) - 1))))

    def set_engine(self, engine):
        """'percall' (one C-ABI call per bond phase) or 'device' (half-sweeps whose Pi fit shared memory as ONE launch each,
        csrc/sweep.cu); identical index sets, ranks and errors."""
        self._check(self.L.tci_tci2_set_engine(self._t, {"percall": 0, "device": 1}[engine]))

    def set_pivotsearch(self, pivotsearch, seed=0):
        """'full' or 'rook' (tensorci2.jl:845-914); `seed` starts the stream behind pushrandomsubset!."""
        if pivotsearch not in ("full", "rook"):
            raise TCIError(-1, f"Unknown pivot search strategy {pivotsearch}. Choose from :rook, :full.")
        self._check(self.L.tci_tci2_set_pivotsearch(self._t, int(pivotsearch == "rook"), C.c_uint64(int(seed) & ((1 << 64) - 1))))

    def addglobalpivots(self, pivots):
        if any(len(p) != len(self) for p in pivots):
            raise TCIError(-1, "Please specify a pivot as one index per leg of the MPS.")
        tab, n = _table(pivots, len(self)) if len(pivots) else (np.zeros((0, len(self)), dtype=np.int32), 0)
        self._check(self.L.tci_tci2_addglobalpivots(self._t, p_i32(tab), C.c_int64(n)))

    def init_maxsample(self, fptr, pivots):
        tab, n = _table(pivots, len(self))
        self._check(self.L.tci_tci2_init_maxsample(self._t, fptr, p_i32(tab), C.c_int64(n)))

    def set_indexset(self, which, site, rows):
        length = site - 1 if which == 0 else len(self) - site
        tab, n = _table(rows, length)
        self._check(self.L.tci_tci2_set_indexset(self._t, which, site, p_i32(tab) if tab.size else None, C.c_int64(n)))

    def indexset(self, which, site):
        n = C.c_int64()
        self._check(self.L.tci_tci2_indexset(self._t, which, site, None, C.byref(n)))
        length = site - 1 if which == 0 else len(self) - site
        out = np.zeros((n.value, length), dtype=np.int32)
        if out.size:
            self._check(self.L.tci_tci2_indexset(self._t, which, site, p_i32(out), C.byref(n)))
        return out

    def Iset(self, site):
        return self.indexset(0, site)

    def Jset(self, site):
        return self.indexset(1, site)

    def linkdims(self):
        return [self.indexset(0, b + 1).shape[0] for b in range(1, len(self))]

    def rank(self):
        return max(self.linkdims())

    def set_maxsamplevalue(self, v):
        self._check(self.L.tci_tci2_set_maxsamplevalue(self._t, C.c_double(v)))

    def errors(self):
        """(bonderrors, pivoterrors, maxsamplevalue)"""
        n, mx = C.c_int64(), C.c_double()
        be = np.zeros(len(self) - 1)
        self._check(self.L.tci_tci2_errors(self._t, p_f64(be), None, C.byref(n), C.byref(mx)))
        pe = np.zeros(max(n.value, 1))
        self._check(self.L.tci_tci2_errors(self._t, None, p_f64(pe), C.byref(n), None))
        return be, pe[: n.value].copy(), mx.value

    def sitetensors_available(self):
        return bool(self.L.tci_tci2_sitetensors_available(self._t))

    def core(self, site):
        """(device pointer, (chi_l, d, chi_r)) of site tensor `site`; valid until the next mutating call."""
        ptr = C.c_void_p()
        shape = np.zeros(3, dtype=np.int64)
        self._check(self.L.tci_tci2_core(self._t, site, C.byref(ptr), p_i64(shape)))
        return ptr.value, tuple(int(v) for v in shape)

    def export_indexsets(self):
        """(Iset, Jset): lists of (count, length) int32 tables for every site, fetched with two C calls."""
        n = len(self)
        counts, tot = np.zeros(2 * n, dtype=np.int64), C.c_int64()
        self._check(self.L.tci_tci2_export_indexsets(self._t, None, C.c_int64(0), p_i64(counts), C.byref(tot)))
        buf = np.zeros(max(tot.value, 1), dtype=np.int32)
        self._check(self.L.tci_tci2_export_indexsets(self._t, p_i32(buf), C.c_int64(buf.size), p_i64(counts), C.byref(tot)))
        I, J, off = [], [], 0
        for b in range(n):
            li, lj = b, n - b - 1
            ci, cj = int(counts[2 * b]), int(counts[2 * b + 1])
            I.append(buf[off:off + ci * li].reshape(ci, li).copy()); off += ci * li
            J.append(buf[off:off + cj * lj].reshape(cj, lj).copy()); off += cj * lj
        return I, J

    def export_cores(self, alloc):
        """All site tensors in ONE device buffer: `alloc(ndoubles)` returns (device pointer, owner); -> (owner, shapes, offsets)."""
        n = len(self)
        shapes, offs, tot = np.zeros(3 * n, dtype=np.int64), np.zeros(n, dtype=np.int64), C.c_int64()
        self._check(self.L.tci_tci2_export_cores(self._t, None, C.c_int64(0), p_i64(shapes), p_i64(offs), C.byref(tot)))
        ptr, owner = alloc(max(tot.value, 1))
        self._check(self.L.tci_tci2_export_cores(self._t, C.c_void_p(ptr), C.c_int64(max(tot.value, 1)), p_i64(shapes), p_i64(offs), C.byref(tot)))
        return owner, shapes.reshape(n, 3), offs

    def evaluate(self, points):
        """evaluate(tci, indexset) at many multi-indices (npoints x nsites, 1-based): numpy vector."""
        tab, n = _table(points, len(self))
        out = np.zeros(n)
        self._check(self.L.tci_tci2_evaluate(self._t, p_i32(tab), C.c_int64(n), p_f64(out)))
        return out

    def sum(self):
        out = C.c_double()
        self._check(self.L.tci_tci2_sum(self._t, C.byref(out)))
        return out.value

    def stats(self):
        counts, secs = np.zeros(2, dtype=np.int64), np.zeros(4)
        self._check(self.L.tci_tci2_stats(self._t, p_i64(counts), p_f64(secs)))
        return {"nevals": int(counts[0]), "nbondupdates": int(counts[1]), "t_eval": float(secs[0]), "t_lu": float(secs[1]),
                "t_factors": float(secs[2]), "t_host": float(secs[3])}

    # ---- sweeps (src/tensorci2.jl)
    def updatepivots(self, fptr, b, leftorthogonal, reltol=1e-14, abstol=0.0, maxbonddim=_UNBOUNDED, extraIset=None, extraJset=None):
        eI, nI = _table(extraIset, b) if extraIset is not None and len(extraIset) else (None, 0)
        eJ, nJ = _table(extraJset, len(self) - b) if extraJset is not None and len(extraJset) else (None, 0)
        self._check(self.L.tci_tci2_updatepivots(self._t, fptr, int(b), int(bool(leftorthogonal)), C.c_double(reltol), C.c_double(abstol),
                                                 C.c_int64(_mbd(maxbonddim)), p_i32(eI) if nI else None, C.c_int64(nI),
                                                 p_i32(eJ) if nJ else None, C.c_int64(nJ)))

    def sweep2site(self, fptr, niter, iter1=1, abstol=1e-8, maxbonddim=_UNBOUNDED, sweepstrategy="backandforth", strictlynested=False,
                   fillsitetensors=True):
        self._check(self.L.tci_tci2_sweep2site(self._t, fptr, int(niter), int(iter1), C.c_double(abstol), C.c_int64(_mbd(maxbonddim)),
                                               STRATEGY[sweepstrategy], int(bool(strictlynested)), int(bool(fillsitetensors))))

    def sweep1site(self, fptr, sweepdirection="forward", reltol=1e-14, abstol=0.0, maxbonddim=_UNBOUNDED, updatetensors=True):
        if sweepdirection not in ("forward", "backward"):
            raise TCIError(-1, f"Unknown sweep direction {sweepdirection}: choose between :forward, :backward.")
        self._check(self.L.tci_tci2_sweep1site(self._t, fptr, int(sweepdirection == "forward"), C.c_double(reltol), C.c_double(abstol),
                                               C.c_int64(_mbd(maxbonddim)), int(bool(updatetensors))))

    def makecanonical(self, fptr, reltol=1e-14, abstol=0.0, maxbonddim=_UNBOUNDED):
        self._check(self.L.tci_tci2_makecanonical(self._t, fptr, C.c_double(reltol), C.c_double(abstol), C.c_int64(_mbd(maxbonddim))))

    def fillsitetensors(self, fptr):
        self._check(self.L.tci_tci2_fillsitetensors(self._t, fptr))

    def optimize(self, fptr, tolerance=1e-8, maxbonddim=_UNBOUNDED, maxiter=20, sweepstrategy="backandforth", normalizeerror=True,
                 ncheckhistory=3, nsearchglobalpivot=5, maxnglobalpivot=5, tolmarginglobalsearch=10.0, strictlynested=False,
                 checkconvglobalpivot=True):
        ranks = np.zeros(max(maxiter, 1), dtype=np.int64)
        errors = np.zeros(max(maxiter, 1))
        niter = C.c_int()
        self._check(self.L.tci_tci2_optimize(self._t, fptr, C.c_double(tolerance), C.c_int64(_mbd(maxbonddim)), int(maxiter),
                                             STRATEGY[sweepstrategy], int(bool(normalizeerror)), int(ncheckhistory), int(nsearchglobalpivot),
                                             int(maxnglobalpivot), C.c_double(tolmarginglobalsearch), int(bool(strictlynested)),
                                             int(bool(checkconvglobalpivot)), p_i64(ranks), p_f64(errors), C.byref(niter)))
        return ranks[: niter.value].copy(), errors[: niter.value].copy()
