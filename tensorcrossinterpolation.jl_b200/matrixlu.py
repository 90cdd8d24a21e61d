"""Host-side mirror of src/matrixlu.jl and src/matrixluci.jl: `rrlu`, `rrlu_inplace` (rrlu!), the
`rrLU` result type with its accessors, and `submatrixargmax`.  All arithmetic runs in the CUDA
library (csrc/rrlu.cu); this file only marshals arguments and mirrors names/defaults/errors."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import TCIError, lib, p_f64, p_i64

INT_MAX = np.iinfo(np.int64).max


class rrLU:
    """rrLU{Float64} (matrixlu.jl:200-207).  rowpermutation/colpermutation are 1-based like Julia.
    L and U are numpy (host) or torch.cuda (device) matrices depending on the entry point used."""

    def __init__(self, rowpermutation, colpermutation, L, U, leftorthogonal, npivot, error, pivoterrors):
        self.rowpermutation = rowpermutation
        self.colpermutation = colpermutation
        self.L, self.U = L, U
        self.leftorthogonal = bool(leftorthogonal)
        self.npivot = int(npivot)
        self.error = float(error)
        self._pivoterrors = pivoterrors

    def size(self):                                       # matrixlu.jl:685-687
        return (self.L.shape[0], self.U.shape[1])

    def left(self, permute=True):                         # matrixlu.jl:716-724
        if not permute:
            return self.L
        out = _empty_like(self.L)
        out[_index(self.rowpermutation - 1, self.L), :] = self.L
        return out

    def right(self, permute=True):                        # matrixlu.jl:738-746
        if not permute:
            return self.U
        out = _empty_like(self.U)
        out[:, _index(self.colpermutation - 1, self.U)] = self.U
        return out

    def diag(self):                                       # matrixlu.jl:756-762
        M = self.U if self.leftorthogonal else self.L
        return np.array([float(M[i, i]) for i in range(self.npivot)])

    def rowindices(self):                                 # matrixlu.jl:769-771
        return self.rowpermutation[: self.npivot].copy()

    def colindices(self):                                 # matrixlu.jl:778-780
        return self.colpermutation[: self.npivot].copy()

    def npivots(self):                                    # matrixlu.jl:787-789
        return self.npivot

    def pivoterrors(self):                                # matrixlu.jl:799-801
        return self._pivoterrors.copy()

    def lastpivoterror(self):                             # matrixlu.jl:811-813
        return self.error

    def transpose(self):                                  # Base.transpose(A::rrLU) — matrixlu.jl:918-922
        tr = (lambda M: np.asfortranarray(M.T)) if isinstance(self.L, np.ndarray) else (lambda M: M.T)
        return rrLU(self.colpermutation, self.rowpermutation, tr(self.U), tr(self.L), not self.leftorthogonal, self.npivot,
                    self.error, self._pivoterrors)

    def solve(self, b, handle=None):                      # Base.:\(A::rrLU, b) — matrixlu.jl:891-901
        m, n = self.size()
        if m != n:
            raise TCIError(_lib.TCI_ERR_INVALID, "Matrix must be square.")
        if self.npivot != m:
            raise TCIError(_lib.TCI_ERR_INVALID, "rank-deficient matrix is not supportred!")
        b = np.asarray(b, dtype=np.float64)
        x_perm = solve(_host(self.L), _host(self.U), b[self.rowpermutation - 1, :], handle=handle)
        x = np.empty_like(x_perm)
        x[self.colpermutation - 1, :] = x_perm
        return x


def _host(M):
    return M if isinstance(M, np.ndarray) else M.cpu().numpy()


def solve(L, U, b, handle=None):
    """solve(L, U, b) — matrixlu.jl:839-868: x with (L * U) x = b by forward and backward substitution; here two device
    triangular solves (csrc/dense.cu) on host matrices."""
    from .dense import Mat, trsm_right
    h = handle or _lib.default_handle()
    L, U, b = np.asarray(L, dtype=np.float64), np.asarray(U, dtype=np.float64), np.asarray(b, dtype=np.float64)
    n2 = L.shape[1]
    if U.shape[0] != n2 or U.shape[1] != n2 or L.shape[0] < n2 or b.shape[0] < n2:
        raise TCIError(_lib.TCI_ERR_INVALID, "solve: L and U must be square factors of the same size")
    dev = f"cuda:{h.device}"
    Lm, Um, Bm = Mat.from_numpy(L[:n2], dev), Mat.from_numpy(U, dev), Mat.from_numpy(b[:n2], dev)
    trsm_right(h, Lm, n2, True, False, Bm, left=True)      # L y = b
    trsm_right(h, Um, n2, False, False, Bm, left=True)     # U x = y
    return Bm.numpy()


def _empty_like(M):
    if isinstance(M, np.ndarray):
        return np.empty_like(M)
    import torch
    return torch.empty_like(M)


def _index(idx, M):
    if isinstance(M, np.ndarray):
        return idx
    import torch
    return torch.as_tensor(idx, device=M.device)


def _maxrank(maxrank, m, n):
    return int(min(maxrank, m, n)) if maxrank is not None else int(min(m, n))


def rrlu(A, maxrank=INT_MAX, reltol=1e-14, abstol=0.0, leftorthogonal=True, handle=None):
    """rrlu(A; maxrank, reltol, abstol, leftorthogonal) — matrixlu.jl:455-463.  A: host matrix (numpy);
    it is not modified.  Returns an rrLU with host L/U."""
    h = handle or _lib.default_handle()
    A = np.asarray(A, dtype=np.float64)
    if A.ndim != 2:
        raise TCIError(_lib.TCI_ERR_INVALID, "rrlu: A must be a matrix")
    if not A.flags.f_contiguous:
        A = np.asfortranarray(A)
    m, n = A.shape
    mr = _maxrank(maxrank, m, n)
    rowperm = np.zeros(m, dtype=np.int64)
    colperm = np.zeros(n, dtype=np.int64)
    L = np.zeros((m, max(mr, 1)), order="F")
    U = np.zeros((max(mr, 1), n), order="F")
    pe = np.zeros(mr + 1)
    npivot, err = C.c_int64(), C.c_double()
    h.check(lib().tci_rrlu_host_f64(h.h, p_f64(A), C.c_int64(m), C.c_int64(n), C.c_int64(max(m, 1)), C.c_int64(mr),
                                    C.c_double(reltol), C.c_double(abstol), int(bool(leftorthogonal)), p_i64(rowperm),
                                    p_i64(colperm), C.byref(npivot), C.byref(err), p_f64(L), C.c_int64(max(m, 1)),
                                    p_f64(U), C.c_int64(max(mr, 1)), p_f64(pe)))
    r = npivot.value
    return rrLU(rowperm, colperm, np.asfortranarray(L[:, :r]), np.asfortranarray(U[:r, :]), leftorthogonal, r, err.value,
                pe[: r + 1].copy())


def _rrlu_device(h, At, m, n, lda, maxrank, reltol, abstol, leftorthogonal, extract=True):
    """rrlu! on a device matrix stored in torch tensor `At` of shape (n, lda) (column-major m x n, ld lda)."""
    import torch
    from .dense import Mat
    assert At.is_cuda and At.dtype == torch.float64 and At.is_contiguous()
    mr = _maxrank(maxrank, m, n)
    rowperm = np.zeros(m, dtype=np.int64)
    colperm = np.zeros(n, dtype=np.int64)
    pe = np.zeros(mr + 1)
    npivot, err = C.c_int64(), C.c_double()
    Lm = Mat.empty(m, max(mr, 1), At.device) if extract else None
    Um = Mat.empty(max(mr, 1), n, At.device) if extract else None
    h.check(lib().tci_rrlu_inplace_f64(h.h, C.c_void_p(At.data_ptr()), C.c_int64(m), C.c_int64(n), C.c_int64(lda),
                                       C.c_int64(mr), C.c_double(reltol), C.c_double(abstol), int(bool(leftorthogonal)),
                                       p_i64(rowperm), p_i64(colperm), C.byref(npivot), C.byref(err),
                                       Lm.ptr if extract else None, C.c_int64(Lm.ld if extract else 1),
                                       Um.ptr if extract else None, C.c_int64(Um.ld if extract else 1), p_f64(pe)))
    r = npivot.value
    L = U = None
    if extract:
        Lm.cols = r
        Um.rows = r
        L, U = Lm.view(), Um.view()                       # logical m x r and r x n views (column-major storage)
    lu = rrLU(rowperm, colperm, L, U, leftorthogonal, r, err.value, pe[: r + 1].copy())
    lu.Lmat, lu.Umat = Lm, Um
    return lu


def rrlu_device(A, maxrank=INT_MAX, reltol=1e-14, abstol=0.0, leftorthogonal=True, handle=None, extract=True):
    """rrlu!(A; ...) on a `dense.Mat` (device, overwritten)."""
    h = handle or _lib.default_handle()
    return _rrlu_device(h, A.buf, A.rows, A.cols, A.ld, maxrank, reltol, abstol, leftorthogonal, extract)


def submatrixargmax(A, startindex, handle=None):
    """submatrixargmax(abs2, A, startindex) — matrixlu.jl:133-135; A host matrix; 1-based result."""
    import torch
    h = handle or _lib.default_handle()
    A = np.asfortranarray(np.asarray(A, dtype=np.float64))
    m, n = A.shape
    if startindex > m or startindex < 1:
        raise TCIError(_lib.TCI_ERR_INVALID, "rows must not be empty")
    if startindex > n:
        raise TCIError(_lib.TCI_ERR_INVALID, "cols must not be empty")
    dA = torch.from_numpy(np.ascontiguousarray(A.T)).to(f"cuda:{h.device}")
    r, c = C.c_int64(), C.c_int64()
    h.check(lib().tci_submatrixargmax_abs2_f64(h.h, C.c_void_p(dA.data_ptr()), C.c_int64(m), C.c_int64(n), C.c_int64(m),
                                               C.c_int64(startindex), C.byref(r), C.byref(c)))
    return r.value, c.value


def rrlu_last_timing(handle=None):
    h = handle or _lib.default_handle()
    ms = np.zeros(3)
    steps = C.c_int64()
    h.check(lib().tci_rrlu_last_timing(h.h, p_f64(ms), C.byref(steps)))
    return {"loop_ms": float(ms[0]), "extract_ms": float(ms[2]), "steps": steps.value}


# ---- rook pivot search: arrlu (matrixlu.jl:500-569) -------------------------------------------------------------------
class Rng:
    """Counter-based splitmix64 stream behind `pushrandomsubset!` (util.jl:173-219).  The reference draws from Julia's
    global RNG, whose stream has no portable restatement; the CPU checker under tests/ restates this same generator, so
    both pick identical random rows/columns for a given seed."""
    MASK = (1 << 64) - 1

    def __init__(self, seed=0):
        self.s = int(seed) & self.MASK

    def next(self):
        self.s = (self.s + 0x9E3779B97F4A7C15) & self.MASK
        z = self.s
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & self.MASK
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & self.MASK
        return z ^ (z >> 31)

    def randint(self, n):                                  # 0-based stand-in for rand(1:n)
        return self.next() % int(n)


def pushrandomsubset(subset, setsize, n, rng):
    """pushrandomsubset!(subset, 1:setsize, n) — util.jl:214-219 with randomsubset :173-191 (draw, then deleteat!)."""
    taken = np.zeros(setsize + 1, dtype=bool)
    taken[np.asarray(subset, dtype=np.int64)] = True
    c = [int(v) for v in np.nonzero(~taken[1:])[0] + 1]
    for _ in range(min(int(n), len(c))):
        subset.append(c.pop(rng.randint(len(c))))


def _as_device_mat(x, h):
    from .dense import Mat
    if isinstance(x, Mat):
        return x
    return Mat.from_numpy(np.asarray(x, dtype=np.float64), f"cuda:{h.device}", pad=16)


def arrlu(f, matrixsize, I0=(), J0=(), maxrank=INT_MAX, reltol=1e-14, abstol=0.0, leftorthogonal=True, numrookiter=5,
          usebatcheval=False, rng=None, handle=None):
    """arrlu(Float64, f, matrixsize, I0, J0; ...) — matrixlu.jl:500-569.  `f(i, j)` (1-based scalars; usebatcheval=False: the
    user's function runs on the host and the sampled strips are uploaded) or `f(rows, cols)` returning the sub-matrix as a
    `dense.Mat`/array (usebatcheval=True; the TCI path passes a device evaluator).  Every strip is factorised on the GPU
    by the same rrLU kernels as the full search; the final L/U completion (:558-573) is two device triangular solves."""
    import torch
    from .dense import Mat, trsm_right
    h = handle or _lib.default_handle()
    rng = rng or Rng(0)
    m, n = int(matrixsize[0]), int(matrixsize[1])
    if usebatcheval:
        batchf = lambda r, c: _as_device_mat(f(r, c), h)
    else:
        batchf = lambda r, c: _as_device_mat(np.array([[f(int(i), int(j)) for j in c] for i in r], dtype=np.float64).reshape(len(r), len(c)), h)
    rowperm = np.arange(1, m + 1, dtype=np.int64)
    colperm = np.arange(1, n + 1, dtype=np.int64)
    I0, J0 = [int(v) for v in I0], [int(v) for v in J0]
    islowrank = False
    maxrank = int(min(maxrank, m, n))
    lus = None
    while True:
        if leftorthogonal:                                 # :521-525
            pushrandomsubset(J0, n, max(1, len(J0)), rng)
        else:
            pushrandomsubset(I0, m, max(1, len(I0)), rng)
        for rookiter in range(1, numrookiter + 1):
            colmove = ((rookiter % 2 == 0) == bool(leftorthogonal))                 # :529
            sub = batchf(np.asarray(I0, dtype=np.int64), colperm.copy()) if colmove else batchf(rowperm.copy(), np.asarray(J0, dtype=np.int64))
            kr, kc = sub.rows, sub.cols
            lus = rrlu_device(sub, maxrank=maxrank, reltol=reltol, abstol=abstol, leftorthogonal=leftorthogonal, handle=h)
            # _optimizerrlu! swaps lu.rowpermutation/colpermutation at the POSITIONS of the strip (:254-275, :539-540)
            rowperm[:kr] = rowperm[:kr][lus.rowpermutation - 1]
            colperm[:kc] = colperm[:kc][lus.colpermutation - 1]
            r = lus.npivot
            islowrank = islowrank or r < min(kr, kc)                                 # :541
            ri, ci = [int(v) for v in rowperm[:r]], [int(v) for v in colperm[:r]]
            if ri == I0 and ci == J0:                                                # :544-546
                break
            I0, J0 = ri, ci
        if islowrank or len(I0) >= maxrank:                                          # :552-554
            break
    r = lus.npivot
    dev = lus.Lmat.buf.device
    Lm, Um = lus.Lmat, lus.Umat
    if Lm.rows < m:                                                                  # :558-564
        i0 = np.asarray(I0, dtype=np.int64)
        I2 = np.setdiff1d(np.arange(1, m + 1, dtype=np.int64), i0)                   # ascending, like setdiff(1:m, I0)
        rowperm = np.concatenate([i0, I2])
        L2 = batchf(I2, np.asarray(J0, dtype=np.int64))
        trsm_right(h, Um, r, False, False, L2)                                       # cols2Lmatrix! :627-648
        full = Mat.empty(m, max(r, 1), dev)
        full.buf[:r, :r].copy_(Lm.buf[:r, :r])
        full.buf[:r, r:m].copy_(L2.buf[:r, : m - r])
        full.cols = r
        Lm = full
    if Um.cols < n:                                                                  # :567-573
        j0 = np.asarray(J0, dtype=np.int64)
        J2 = np.setdiff1d(np.arange(1, n + 1, dtype=np.int64), j0)
        colperm = np.concatenate([j0, J2])
        U2 = batchf(np.asarray(I0, dtype=np.int64), J2)
        trsm_right(h, Lm, r, True, False, U2, left=True)                             # rows2Umatrix! :655-674
        full = Mat.empty(max(r, 1), n, dev)
        full.buf[:r, :r].copy_(Um.buf[:r, :r])
        full.buf[r:n, :r].copy_(U2.buf[: n - r, :r])
        full.rows = r
        Um = full
    lu = rrLU(rowperm, colperm, Lm.view(), Um.view(), leftorthogonal, r, lus.error, lus.pivoterrors())
    lu.Lmat, lu.Umat = Lm, Um
    return lu


def rrlu_function(f, matrixsize, I0=(), J0=(), pivotsearch="full", usebatcheval=False, handle=None, **kw):
    """rrlu(Float64, f, matrixsize, I0, J0; pivotsearch, kwargs...) — matrixlu.jl:599-617."""
    h = handle or _lib.default_handle()
    if pivotsearch == "rook":
        return arrlu(f, matrixsize, I0, J0, usebatcheval=usebatcheval, handle=h, **kw)
    if pivotsearch == "full":
        m, n = int(matrixsize[0]), int(matrixsize[1])
        kw.pop("numrookiter", None), kw.pop("rng", None)
        rows, cols = np.arange(1, m + 1, dtype=np.int64), np.arange(1, n + 1, dtype=np.int64)
        A = f(rows, cols) if usebatcheval else np.array([[f(int(i), int(j)) for j in cols] for i in rows], dtype=np.float64).reshape(m, n)
        return rrlu_device(_as_device_mat(A, h), handle=h, **kw)
    raise TCIError(_lib.TCI_ERR_INVALID, f"Unknown pivot search strategy {pivotsearch}. Choose between :rook and :full.")
