"""Host-side mirror of src/batcheval.jl and the BatchEvaluator protocol (src/cachedtensortrain.jl:31,
docs/src/index.md:181-237).  A `BatchEvaluator` here owns a device functor; calling it with
(leftindexset, rightindexset, M) evaluates the whole batch on the GPU (csrc/batcheval.cu) and returns
a torch CUDA tensor whose memory is the column-major Array{Float64,M+2} the reference returns."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import TCIError, lib, p_f64, p_i32

KIND_LORENTZ, KIND_QEXPSIN, KIND_QEXPPLUSSIN, KIND_SUM, KIND_QEXP2, KIND_HASH = 1, 2, 3, 4, 5, 7


def as_table(indexset, length=None):
    """list of multi-indices (or an (n, len) array) -> C-contiguous int32 (n, len) table."""
    if isinstance(indexset, np.ndarray) and indexset.ndim == 2:
        return np.ascontiguousarray(indexset, dtype=np.int32)
    n = len(indexset)
    if n == 0:
        return np.zeros((0, length or 0), dtype=np.int32)
    ln = len(indexset[0])
    for x in indexset:
        if len(x) != ln:
            raise TCIError(_lib.TCI_ERR_INVALID, "all multi-indices of a set must have the same length")
    return np.ascontiguousarray(np.asarray(indexset, dtype=np.int32).reshape(n, ln))


class BatchEvaluator:
    """abstract type BatchEvaluator{Float64} (cachedtensortrain.jl:31): point form `f(indexset)` and batch form
    `f(leftindexset, rightindexset, M)`."""

    def __init__(self, localdims, handle=None):
        self.localdims = np.asarray(localdims, dtype=np.int32)
        self.handle = handle or _lib.default_handle()
        self._f = C.c_void_p()

    # -- device functor lifetime
    def _destroy(self):
        if getattr(self, "_f", None):
            lib().tci_functor_destroy(self.handle.h, self._f)
            self._f = C.c_void_p()

    def __del__(self):
        try:
            self._destroy()
        except Exception:
            pass

    # -- batch form: returns a torch CUDA tensor `t` with t.shape == (nright, *reversed(centre dims), nleft) whose
    #    memory is the column-major (nleft, centre..., nright) array; use `to_numpy(t)` for the Julia-shaped array.
    def evaluate_batch(self, left, right, M, out=None, stride_c=0, stride_j=0, maxabs=None):
        import torch
        h = self.handle
        N = len(self.localdims)
        lt, rt = as_table(left), as_table(right)
        nleft, nright = lt.shape[0], rt.shape[0]
        if nleft * nright == 0:                                                     # batcheval.jl:140-142
            return torch.zeros((0,) * (M + 2), dtype=torch.float64, device=f"cuda:{h.device}"), maxabs
        nl, nr = lt.shape[1], rt.shape[1]
        if N - nl - nr != M:
            raise TCIError(_lib.TCI_ERR_INVALID, "Invalid number of central indices")   # tensorci2.jl:492
        cd = [int(d) for d in self.localdims[nl:nl + M]]
        if out is None:
            out = torch.empty((nright,) + tuple(reversed(cd)) + (nleft,), dtype=torch.float64, device=f"cuda:{h.device}")
        mx = C.c_double(0.0 if maxabs is None else float(maxabs))
        h.check(lib().tci_batcheval_f64(h.h, self._f, p_i32(self.localdims), N, p_i32(lt), C.c_int64(nleft), nl,
                                        p_i32(rt), C.c_int64(nright), nr, C.c_void_p(out.data_ptr()),
                                        C.c_int64(stride_c), C.c_int64(stride_j), C.byref(mx) if maxabs is not None else None))
        return out, (mx.value if maxabs is not None else None)

    def evaluate_points(self, pts):
        """f at arbitrary full-length multi-indices (npoints x N) in ONE K1 call: the points are the left set, the right
        set is the single empty multi-index.  Returns a numpy vector."""
        t, _ = self.evaluate_batch(np.ascontiguousarray(pts, dtype=np.int32), np.zeros((1, 0), dtype=np.int32), 0)
        return t.reshape(-1).cpu().numpy()

    def __call__(self, *args):
        if len(args) == 1:                                                          # point form f(indexset)
            idx = [int(v) for v in args[0]]
            if len(idx) != len(self.localdims):
                raise TCIError(_lib.TCI_ERR_INVALID, "wrong number of indices")
            t, _ = self.evaluate_batch([idx[:1]], [idx[1:]], 0)
            return float(t.cpu().reshape(-1)[0])
        left, right, M = args
        t, _ = self.evaluate_batch(left, right, M)
        return to_numpy(t)


def to_numpy(t):
    """torch tensor in 'reversed-shape row-major' storage -> numpy array with the reference's shape (column-major)."""
    a = t.cpu().numpy()
    return a.transpose(tuple(reversed(range(a.ndim))))


class BuiltinFunction(BatchEvaluator):
    def __init__(self, kind, params, localdims, handle=None):
        super().__init__(localdims, handle)
        p = np.asarray(params, dtype=np.float64)
        self.kind, self.params = kind, p
        self.handle.check(lib().tci_functor_builtin(self.handle.h, int(kind), p_f64(p), len(p), C.byref(self._f)))


class LorentzFunction(BuiltinFunction):
    """f(v) = coeff / (1 + sum v_i^2)  (README.md:21-29)."""

    def __init__(self, localdims, coeff=1.0, handle=None):
        super().__init__(KIND_LORENTZ, [coeff], localdims, handle)


class SumFunction(BuiltinFunction):
    """f(v) = scale * sum v_i  (test/test_batcheval.jl:18)."""

    def __init__(self, localdims, scale=1.0, handle=None):
        super().__init__(KIND_SUM, [scale], localdims, handle)


class QuanticsExpSin(BuiltinFunction):
    """exp(-a x) sin(2 pi w x^p) on the quantics grid x = sum (b_i-1) 2^-i  (test/test_tensorci2.jl:437)."""

    def __init__(self, localdims, a=10.0, w=100.0, p=1.1, handle=None):
        super().__init__(KIND_QEXPSIN, [a, w, p], localdims, handle)


class QuanticsExpPlusSin(BuiltinFunction):
    """exp(-a x) + c sin(w x)  (test/test_globalsearch.jl:14)."""

    def __init__(self, localdims, a=1.0, c=1e-3, w=1000.0, handle=None):
        super().__init__(KIND_QEXPPLUSSIN, [a, c, w], localdims, handle)


class QuanticsExp2(BuiltinFunction):
    """exp(-a x) + c exp(-b x)  (test/test_tensorci2.jl:63, :146)."""

    def __init__(self, localdims, a=1.0, c=0.0, b=0.0, handle=None):
        super().__init__(KIND_QEXP2, [a, c, b], localdims, handle)


class HashFunction(BuiltinFunction):
    """amp * u(v) + lor / (1 + sum v_i^2) with u an integer hash of v in [-1, 1): synthetic full-rank benchmark function."""

    def __init__(self, localdims, amp=1.0, lor=0.0, handle=None):
        super().__init__(KIND_HASH, [amp, lor], localdims, handle)


class DenseFunction(BatchEvaluator):
    """f(v) = T[v_1, ..., v_N] for a dense table T (any numpy array; its shape gives localdims)."""

    def __init__(self, table, handle=None):
        t = np.asarray(table, dtype=np.float64)
        super().__init__(t.shape, handle)
        flat = np.ascontiguousarray(t.reshape(-1, order="F"))
        self.handle.check(lib().tci_functor_dense(self.handle.h, p_f64(flat), p_i32(self.localdims), t.ndim, C.byref(self._f)))


class UserFunction(BatchEvaluator):
    """Device functor plugin: CUDA source defining
    `__device__ double tci_user_f(const int* idx, int n, const double* params, int nparams)`; NVRTC -> sm_100a."""

    def __init__(self, cuda_source, localdims, params=(), handle=None):
        super().__init__(localdims, handle)
        p = np.asarray(params, dtype=np.float64)
        self.handle.check(lib().tci_functor_from_source(self.handle.h, cuda_source.encode(), p_f64(p) if len(p) else None,
                                                        len(p), C.byref(self._f)))


def _batchevaluate_dispatch(f, localdims, leftindexset, rightindexset, M):
    """batcheval.jl:196-214 (BatchEvaluator method): numpy array of shape (|left|, d..., |right|)."""
    return f(leftindexset, rightindexset, M)


def filltensor(f, localdims, Iset, Jset, M):
    """filltensor (tensorci2.jl:475-497)."""
    return _batchevaluate_dispatch(f, localdims, Iset, Jset, M)
