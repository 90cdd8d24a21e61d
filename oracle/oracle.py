"""ctypes front end of the CPU oracle (oracle/tci_oracle.cpp).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline
legs.  The product package never imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libtci_oracle.so")

F_CALLBACK, F_LORENTZ, F_QEXPSIN, F_QEXPPLUSSIN, F_SUM, F_QEXP2, F_DENSE, F_HASH, F_BATCHCB = range(9)
INT64_MAX = np.iinfo(np.int64).max

CALLBACK = C.CFUNCTYPE(C.c_double, C.POINTER(C.c_int32), C.c_int32, C.c_void_p)
BATCH_CALLBACK = C.CFUNCTYPE(None, C.POINTER(C.c_int32), C.c_int64, C.c_int32, C.POINTER(C.c_int32), C.c_int64, C.c_int32,
                             C.c_int32, C.POINTER(C.c_double), C.c_void_p)


def build(force=False):
    src = os.path.join(_HERE, "tci_oracle.cpp")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE], stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = C.CDLL(_SO)
        _lib.orc_last_error.restype = C.c_char_p
        _lib.orc_func_create.restype = C.c_void_p
        _lib.orc_func_eval.restype = C.c_double
        _lib.orc_tci_create.restype = C.c_void_p
        _lib.orc_tci_setsize.restype = C.c_int64
        _lib.orc_tci_maxsamplevalue.restype = C.c_double
        _lib.orc_tci_npivoterrors.restype = C.c_int64
    return _lib


class OracleError(RuntimeError):
    pass


_MSG = {-1: "lu.L contains NaNs", -2: "lu.U contains NaNs", -3: "Error: NaN in tensor",
        -4: "Pivot matrix is not square!"}


def _check(st):
    if st == 0:
        return
    if st in _MSG:
        raise OracleError(_MSG[st])
    raise OracleError(lib().orc_last_error().decode())


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def _f64(a):
    return np.asfortranarray(np.asarray(a, dtype=np.float64))


def submatrixargmax(A, rows, cols, mode=0):
    """rows/cols: 1-based index lists; returns 1-based (row, col).  matrixlu.jl:46-87"""
    A = _f64(A)
    r = np.asarray(rows, dtype=np.int64) - 1
    c = np.asarray(cols, dtype=np.int64) - 1
    outr, outc = C.c_int64(), C.c_int64()
    _check(lib().orc_submatrixargmax(_p(A, C.c_double), C.c_int64(A.shape[0]), mode, _p(r, C.c_int64), C.c_int64(len(r)),
                                     _p(c, C.c_int64), C.c_int64(len(c)), C.byref(outr), C.byref(outc)))
    return outr.value + 1, outc.value + 1


class RRLU:
    """Mirror of rrLU{Float64} (matrixlu.jl:200-207) with the accessors :716-811."""

    def __init__(self, rowperm, colperm, L, U, leftorthogonal, npivot, error, pivoterrors):
        self.rowpermutation, self.colpermutation = rowperm, colperm
        self.L, self.U = L, U
        self.leftorthogonal, self.npivot, self.error = leftorthogonal, npivot, error
        self._pivoterrors = pivoterrors

    def left(self, permute=True):
        if not permute:
            return self.L
        out = np.empty_like(self.L)
        out[self.rowpermutation - 1, :] = self.L
        return out

    def right(self, permute=True):
        if not permute:
            return self.U
        out = np.empty_like(self.U)
        out[:, self.colpermutation - 1] = self.U
        return out

    def rowindices(self):
        return self.rowpermutation[: self.npivot]

    def colindices(self):
        return self.colpermutation[: self.npivot]

    def pivoterrors(self):
        return self._pivoterrors

    def luci_factors(self):
        m, n, r = self.L.shape[0], self.U.shape[1], self.npivot
        left = np.zeros((m, r), order="F")
        right = np.zeros((r, n), order="F")
        L, U = _f64(self.L), _f64(self.U)
        _check(lib().orc_luci_factors(C.c_int64(m), C.c_int64(n), C.c_int64(r), int(self.leftorthogonal),
                                      _p(self.rowpermutation, C.c_int64), _p(self.colpermutation, C.c_int64),
                                      _p(L, C.c_double), C.c_int64(max(m, 1)), _p(U, C.c_double), C.c_int64(max(r, 1)),
                                      _p(left, C.c_double), _p(right, C.c_double)))
        return left, right


def max_threads():
    return int(lib().orc_max_threads())


def rrlu(A, maxrank=INT64_MAX, reltol=1e-14, abstol=0.0, leftorthogonal=True, threads=1):
    """rrlu (matrixlu.jl:455-463).  threads > 1 runs the same loop with the column ranges of the update and of the search
    spread over host threads; every value and every decision is unchanged (see optimizerrlu_mt)."""
    lib().orc_rrlu_set_threads(int(threads))
    try:
        return _rrlu(A, maxrank, reltol, abstol, leftorthogonal)
    finally:
        lib().orc_rrlu_set_threads(1)


def _rrlu(A, maxrank, reltol, abstol, leftorthogonal):
    A = _f64(A)
    m, n = A.shape
    rmax = int(min(maxrank, m, n))
    rowperm = np.zeros(m, dtype=np.int64)
    colperm = np.zeros(n, dtype=np.int64)
    L = np.zeros((m, max(rmax, 1)), order="F")
    U = np.zeros((max(rmax, 1), n), order="F")
    pe = np.zeros(rmax + 1)
    npivot, err = C.c_int64(), C.c_double()
    _check(lib().orc_rrlu(_p(A, C.c_double), C.c_int64(m), C.c_int64(n), C.c_int64(m), C.c_int64(int(min(maxrank, INT64_MAX))),
                          C.c_double(reltol), C.c_double(abstol), int(leftorthogonal), _p(rowperm, C.c_int64),
                          _p(colperm, C.c_int64), C.byref(npivot), C.byref(err), _p(L, C.c_double), C.c_int64(m),
                          _p(U, C.c_double), C.c_int64(max(rmax, 1)), _p(pe, C.c_double)))
    r = npivot.value
    return RRLU(rowperm, colperm, np.asfortranarray(L[:, :r]), np.asfortranarray(U[:r, :]), leftorthogonal, r,
                err.value, pe[: r + 1].copy())


def arrlu(A, I0=(), J0=(), maxrank=INT64_MAX, reltol=1e-14, abstol=0.0, leftorthogonal=True, numrookiter=5, seed=0):
    """arrlu(Float64, (i, j) -> A[i, j], size(A), I0, J0; ...) — matrixlu.jl:500-569 (rook pivot search).  The random
    column/row draws come from the oracle's own seeded stream (see tci_oracle.cpp `Rng`), not Julia's RNG.
    Returns (RRLU, number of matrix elements sampled)."""
    A = _f64(A)
    m, n = A.shape
    rmax = int(min(maxrank, m, n))
    rowperm = np.zeros(m, dtype=np.int64)
    colperm = np.zeros(n, dtype=np.int64)
    L = np.zeros((m, max(rmax, 1)), order="F")
    U = np.zeros((max(rmax, 1), n), order="F")
    pe = np.zeros(rmax + 1)
    npivot, err, nev = C.c_int64(), C.c_double(), C.c_int64()
    i0 = np.asarray(list(I0), dtype=np.int64)
    j0 = np.asarray(list(J0), dtype=np.int64)
    _check(lib().orc_arrlu(_p(A, C.c_double), C.c_int64(m), C.c_int64(n), C.c_int64(m), _p(i0, C.c_int64), C.c_int64(len(i0)),
                           _p(j0, C.c_int64), C.c_int64(len(j0)), C.c_int64(int(min(maxrank, INT64_MAX))), C.c_double(reltol),
                           C.c_double(abstol), int(leftorthogonal), int(numrookiter), C.c_uint64(seed), _p(rowperm, C.c_int64),
                           _p(colperm, C.c_int64), C.byref(npivot), C.byref(err), _p(L, C.c_double), C.c_int64(m),
                           _p(U, C.c_double), C.c_int64(max(rmax, 1)), _p(pe, C.c_double), C.byref(nev)))
    r = npivot.value
    return RRLU(rowperm, colperm, np.asfortranarray(L[:, :r]), np.asfortranarray(U[:r, :]), leftorthogonal, r,
                err.value, pe[: r + 1].copy()), nev.value


def solve_right(P, B):
    P, B = _f64(P), _f64(B)
    X = np.zeros(B.shape, order="F")
    _check(lib().orc_solve_right(_p(P, C.c_double), C.c_int64(P.shape[0]), _p(B, C.c_double), C.c_int64(B.shape[0]),
                                 _p(X, C.c_double)))
    return X


class Func:
    """Target function handle: a built-in kind, a dense table, a Python callback (1-based indices), or (F_BATCHCB) a batch
    evaluator `callback(left, right, M) -> array of shape (len(left), d..., len(right))` with left/right as int32 tables —
    the BatchEvaluator protocol of the reference (batcheval.jl:196-214)."""

    def __init__(self, kind, params=(), table=None, callback=None):
        self.kind = kind
        p = np.asarray(params, dtype=np.float64)
        dims = np.zeros(0, dtype=np.int32)
        tab = np.zeros(0)
        if kind == F_DENSE:
            t = np.asarray(table, dtype=np.float64)
            dims = np.asarray(t.shape, dtype=np.int32)
            tab = np.ascontiguousarray(t.reshape(-1, order="F"))
        self._cb = None
        cb = C.cast(None, CALLBACK)
        if kind == F_CALLBACK:
            def _tramp(idx, n, _user):
                return float(callback([idx[i] for i in range(n)]))
            self._cb = CALLBACK(_tramp)
            cb = self._cb
        if kind == F_BATCHCB:
            def _btramp(left, nleft, nl, right, nright, nr, M, out, _user):
                lt = np.ctypeslib.as_array(left, shape=(nleft, nl)).copy() if nl > 0 else np.zeros((nleft, 0), dtype=np.int32)
                rt = np.ctypeslib.as_array(right, shape=(nright, nr)).copy() if nr > 0 else np.zeros((nright, 0), dtype=np.int32)
                vals = np.asarray(callback(lt, rt, int(M)), dtype=np.float64).reshape(-1, order="F")
                np.ctypeslib.as_array(out, shape=(vals.size,))[:] = vals
            self._cb = BATCH_CALLBACK(_btramp)
            cb = C.cast(self._cb, CALLBACK)
        self.h = C.c_void_p(lib().orc_func_create(kind, _p(p, C.c_double), len(p), _p(dims, C.c_int32), len(dims),
                                                  _p(tab, C.c_double), cb, None))
        if not self.h:
            raise OracleError(lib().orc_last_error().decode())

    def __call__(self, idx):
        v = np.asarray(idx, dtype=np.int32)
        return lib().orc_func_eval(self.h, _p(v, C.c_int32), len(v))

    def __del__(self):
        try:
            lib().orc_func_destroy(self.h)
        except Exception:
            pass


def _table(sets, length):
    a = np.asarray(sets, dtype=np.int32).reshape(len(sets), length)
    return np.ascontiguousarray(a)


def batchevaluate(f, localdims, left, right, M):
    """_batchevaluate_dispatch (batcheval.jl:131-175): returns array of shape (|left|, d..., |right|)."""
    ld = np.asarray(localdims, dtype=np.int32)
    if len(left) == 0 or len(right) == 0:
        return np.zeros((0,) * (M + 2))
    nl, nr = len(left[0]), len(right[0])
    lt, rt = _table(left, nl), _table(right, nr)
    shape = (len(left),) + tuple(int(d) for d in ld[nl:nl + M]) + (len(right),)
    out = np.zeros(int(np.prod(shape)))
    _check(lib().orc_batcheval(f.h, _p(ld, C.c_int32), len(ld), _p(lt, C.c_int32), C.c_int64(len(left)), nl,
                               _p(rt, C.c_int32), C.c_int64(len(right)), nr, M, _p(out, C.c_double)))
    return out.reshape(shape, order="F")


class TCI2:
    """Handle on the oracle's TensorCI2 (tensorci2.jl:50-93).  Bond/site numbers are 1-based."""

    def __init__(self, f, localdims, initialpivots=None):
        self.f = f
        self.localdims = np.asarray(localdims, dtype=np.int32)
        n = len(self.localdims)
        if initialpivots is None:
            initialpivots = [[1] * n]
        piv = _table(initialpivots, n) if len(initialpivots) else np.zeros((0, n), dtype=np.int32)
        self.h = C.c_void_p(lib().orc_tci_create(f.h, _p(self.localdims, C.c_int32), n, _p(piv, C.c_int32),
                                                 C.c_int64(len(initialpivots))))
        if not self.h:
            raise OracleError(lib().orc_last_error().decode())

    def __del__(self):
        try:
            lib().orc_tci_destroy(self.h)
        except Exception:
            pass

    def __len__(self):
        return len(self.localdims)

    def set_bondorder(self, order):
        """'sequential' (the reference) or 'oddeven' (the product's parallel-bond schedule restated: all odd bonds, then
        all even bonds; within one pass the bonds read and write disjoint index sets, so the order inside a pass is
        immaterial and a concurrent execution gives the same result)."""
        lib().orc_tci_set_bondorder(self.h, {"sequential": 0, "oddeven": 1}[order])

    def set_pivotsearch(self, pivotsearch, seed=0):
        """'full' or 'rook' (tensorci2.jl:845-914); `seed` starts the stream behind pushrandomsubset!."""
        lib().orc_tci_set_pivotsearch(self.h, {"full": 0, "rook": 1}[pivotsearch], C.c_uint64(seed))

    def optimize(self, tolerance=1e-8, maxbonddim=INT64_MAX, maxiter=20, sweepstrategy="backandforth",
                 normalizeerror=True, ncheckhistory=3, strictlynested=False, checkconvglobalpivot=True,
                 nsearchglobalpivot=0, maxnglobalpivot=5, tolmarginglobalsearch=10.0, globalseed=0):
        """optimize! (tensorci2.jl:1018-1172).  nsearchglobalpivot > 0 runs the DefaultGlobalPivotFinder with its random
        starts drawn from the splitmix64 stream seeded with `globalseed` (the stream the product's `seed=` uses)."""
        ranks = np.zeros(maxiter, dtype=np.int64)
        errors = np.zeros(maxiter)
        niter = C.c_int()
        _check(lib().orc_tci_optimize_gp(self.h, C.c_double(tolerance), C.c_int64(int(maxbonddim)), maxiter,
                                         0 if sweepstrategy == "forward" else 1, int(normalizeerror), ncheckhistory,
                                         int(strictlynested), int(checkconvglobalpivot), int(nsearchglobalpivot),
                                         int(maxnglobalpivot), C.c_double(tolmarginglobalsearch), C.c_uint64(int(globalseed)),
                                         _p(ranks, C.c_int64), _p(errors, C.c_double), C.byref(niter)))
        return ranks[: niter.value].copy(), errors[: niter.value].copy()

    def updatepivots(self, b, leftorthogonal, reltol=1e-14, abstol=0.0, maxbonddim=INT64_MAX):
        _check(lib().orc_tci_updatepivots(self.h, b, int(leftorthogonal), C.c_double(reltol), C.c_double(abstol),
                                          C.c_int64(int(maxbonddim))))

    def sweep2site(self, niter, iter1=1, abstol=1e-8, maxbonddim=INT64_MAX, sweepstrategy="backandforth",
                   strictlynested=False, fillsitetensors=True):
        _check(lib().orc_tci_sweep2site(self.h, niter, iter1, C.c_double(abstol), C.c_int64(int(maxbonddim)),
                                        0 if sweepstrategy == "forward" else 1, int(strictlynested), int(fillsitetensors)))

    def sweep1site(self, direction="forward", reltol=1e-14, abstol=0.0, maxbonddim=INT64_MAX, updatetensors=True):
        _check(lib().orc_tci_sweep1site(self.h, int(direction == "forward"), C.c_double(reltol), C.c_double(abstol),
                                        C.c_int64(int(maxbonddim)), int(updatetensors)))

    def sweep0site(self, b, reltol=1e-14, abstol=0.0):                       # sweep0site! tensorci2.jl:559-587, b 1-based
        _check(lib().orc_tci_sweep0site(self.h, int(b), C.c_double(reltol), C.c_double(abstol)))

    rmbadpivots = sweep0site                                                 # tensorci2.jl:589

    def makecanonical(self, reltol=1e-14, abstol=0.0, maxbonddim=INT64_MAX):
        _check(lib().orc_tci_makecanonical(self.h, C.c_double(reltol), C.c_double(abstol), C.c_int64(int(maxbonddim))))

    def addglobalpivots(self, pivots):
        piv = _table(pivots, len(self))
        _check(lib().orc_tci_addglobalpivots(self.h, _p(piv, C.c_int32), C.c_int64(len(pivots))))

    def addglobalpivots1sitesweep(self, pivots, reltol=1e-14, abstol=0.0, maxbonddim=INT64_MAX):
        self.addglobalpivots(pivots)                      # tensorci2.jl:367-377
        self.makecanonical(reltol, abstol, maxbonddim)

    def addglobalpivots2sitesweep(self, pivots, tolerance=1e-8, normalizeerror=True, maxbonddim=INT64_MAX, ntry=10,
                                  strictlynested=False):
        """addglobalpivots2sitesweep! (tensorci2.jl:407-453)."""
        pivots = [[int(v) for v in p] for p in pivots]
        if any(len(p) != len(self) for p in pivots):
            raise OracleError("Please specify a pivot as one index per leg of the MPS.")
        pivots_ = pivots
        for _ in range(ntry):
            errnorm = self.maxsamplevalue if normalizeerror else 1.0
            abstol = tolerance * errnorm
            self.addglobalpivots(pivots_)
            self.sweep2site(2, abstol=abstol, maxbonddim=maxbonddim, strictlynested=strictlynested)
            newpivots = [p for p in pivots if abs(self.evaluate(p) - self.f(p)) > abstol]
            if len(newpivots) == 0 or set(map(tuple, newpivots)) == set(map(tuple, pivots_)):
                return len(newpivots)
            pivots_ = newpivots
        return len(pivots_)

    def fillsitetensors(self):
        _check(lib().orc_tci_fillsitetensors(self.h))

    def _set(self, which, b):
        ln = C.c_int()
        cnt = lib().orc_tci_setsize(self.h, which, b, C.byref(ln))
        out = np.zeros((cnt, ln.value), dtype=np.int32)
        if out.size:
            lib().orc_tci_getset(self.h, which, b, _p(out, C.c_int32))
        return out

    def Iset(self, b):
        return self._set(0, b)

    def Jset(self, b):
        return self._set(1, b)

    def linkdims(self):
        return [len(self.Iset(b + 1)) for b in range(1, len(self))]

    def rank(self):
        return max(self.linkdims())

    def sitetensor(self, b):
        shp = np.zeros(3, dtype=np.int64)
        lib().orc_tci_sitetensor_shape(self.h, b, _p(shp, C.c_int64))
        out = np.zeros(int(np.prod(shp)))
        if out.size:
            lib().orc_tci_sitetensor(self.h, b, _p(out, C.c_double))
        return out.reshape(tuple(int(s) for s in shp), order="F")

    def sitetensors(self):
        return [self.sitetensor(b) for b in range(1, len(self) + 1)]

    @property
    def maxsamplevalue(self):
        return lib().orc_tci_maxsamplevalue(self.h)

    @property
    def pivoterrors(self):
        out = np.zeros(lib().orc_tci_npivoterrors(self.h))
        if out.size:
            lib().orc_tci_pivoterrors(self.h, _p(out, C.c_double))
        return out

    @property
    def bonderrors(self):
        out = np.zeros(len(self) - 1)
        lib().orc_tci_bonderrors(self.h, _p(out, C.c_double))
        return out

    def pivoterror(self):
        return float(np.max(self.bonderrors))

    def evaluate(self, idx):
        v = np.asarray(idx, dtype=np.int32)
        out = C.c_double()
        _check(lib().orc_tci_evaluate(self.h, _p(v, C.c_int32), C.byref(out)))
        return out.value

    def stats(self):
        ne, te, tl = C.c_int64(), C.c_double(), C.c_double()
        lib().orc_tci_stats(self.h, C.byref(ne), C.byref(te), C.byref(tl))
        return {"nevals": ne.value, "t_eval": te.value, "t_lu": tl.value}


def crossinterpolate2(f, localdims, initialpivots=None, bondorder="sequential", pivotsearch="full", seed=0, **kw):
    """crossinterpolate2 (tensorci2.jl:1313-1323); nsearchglobalpivot defaults to 0 here (the reference's default is 5).
    `seed` starts both random streams (rook columns and global-pivot starts), like the product's `seed=`."""
    tci = TCI2(f, localdims, initialpivots)
    tci.set_bondorder(bondorder)
    tci.set_pivotsearch(pivotsearch, seed)
    kw.setdefault("globalseed", seed)
    ranks, errors = tci.optimize(**kw)
    return tci, ranks, errors


def optfirstpivot(f, localdims, firstpivot=None, maxsweep=1000):
    """util.jl:260-291, literally (sequential strict-> scan)."""
    n = len(localdims)
    pivot = [1] * n if firstpivot is None else [int(v) for v in firstpivot]
    valf = abs(f(pivot))
    for _ in range(maxsweep):
        valf_prev = valf
        for i in range(n):
            for d in range(1, int(localdims[i]) + 1):
                bak = pivot[i]
                pivot[i] = d
                newval = abs(f(pivot))
                if newval > valf:
                    valf = newval
                else:
                    pivot[i] = bak
        if valf_prev == valf:
            break
    return pivot


def _flatten_tt(cores):
    shapes = np.asarray([[c.shape[0], int(np.prod(c.shape[1:-1])), c.shape[-1]] for c in cores], dtype=np.int64)
    flat = np.concatenate([np.asarray(c, dtype=np.float64).reshape(-1, order="F") for c in cores])
    return flat, shapes


def tt_sum(cores):
    flat, shapes = _flatten_tt(cores)
    out = C.c_double()
    _check(lib().orc_tt_sum(_p(flat, C.c_double), _p(shapes, C.c_int64), len(cores), C.byref(out)))
    return out.value


def tt_evaluate(cores, points):
    flat, shapes = _flatten_tt(cores)
    pts = np.ascontiguousarray(np.asarray(points, dtype=np.int32).reshape(-1, len(cores)))
    out = np.zeros(len(pts))
    _check(lib().orc_tt_evaluate(_p(flat, C.c_double), _p(shapes, C.c_int64), len(cores), _p(pts, C.c_int32),
                                 C.c_int64(len(pts)), _p(out, C.c_double)))
    return out


# ---- integrate (src/integration.jl:68-161) -------------------------------------------------------------------------
# QuadGK.kronrod(7, -1, +1) (integration.jl:63; QuadGK.jl is a registry dependency, not in the reference tree): the 15-point
# Gauss–Kronrod rule.  The constants below are the published QUADPACK qk15 values (positive half, centre last).
_QK15_X = [0.991455371120812639206854697526329, 0.949107912342758524526189684047851, 0.864864423359769072789712788640926,
           0.741531185599394439863864773280788, 0.586087235467691130294144838258730, 0.405845151377397166906606412076961,
           0.207784955007898467600689403773245, 0.0]
_QK15_W = [0.022935322010529224963732008058970, 0.063092092629978553290700663189204, 0.104790010322250183839876322541518,
           0.140653259715525918745189590510238, 0.169004726639267902826583426598550, 0.190350578064785409913256402421014,
           0.204432940075298892414161999234649, 0.209482141084727828012999174891714]


def kronrod15():
    x = np.array([-v for v in _QK15_X] + _QK15_X[-2::-1])
    w = np.array(_QK15_W + _QK15_W[-2::-1])
    return x, w


def integrate(f, a, b, **kw):
    """integrate(Float64, f, a, b; GKorder=15, kwargs...) with nsearchglobalpivot=0 (the oracle has no global search)."""
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    if len(a) != len(b):
        raise OracleError(f"Integral bounds must have the same dimensionality, but got {len(a)} lower bounds and {len(b)} upper bounds.")
    x1, w1 = kronrod15()
    N = len(a)
    nodes = (b - a)[:, None] * (x1[None, :] + 1.0) / 2.0 + a[:, None]
    weights = (b - a)[:, None] * w1[None, :] / 2.0
    normalization = 15.0 ** N

    def F(indices):
        x = [float(nodes[n, i - 1]) for n, i in enumerate(indices)]
        w = 1.0
        for n, i in enumerate(indices):
            w = w * float(weights[n, i - 1])
        return w * f(x) * normalization

    tci, ranks, errors = crossinterpolate2(Func(F_CALLBACK, callback=F), [15] * N, **kw)
    return tt_sum(tci.sitetensors()) / normalization
