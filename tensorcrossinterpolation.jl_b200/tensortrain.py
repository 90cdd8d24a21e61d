"""Host-side mirror of src/tensortrain.jl / src/abstracttensortrain.jl: a TensorTrain whose cores live on the
device; `sum` and `evaluate` run in csrc/tt.cu."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import TCIError, lib, p_f64, p_i32, p_i64
from .dense import Mat


class DeviceCore:
    """One core (chi_l, *sitedims, chi_r), column-major, stored as a Mat of (chi_l*prod(sitedims)) x chi_r."""

    def __init__(self, mat, chil, sitedims, chir):
        self.mat, self.chil, self.sitedims, self.chir = mat, int(chil), tuple(int(d) for d in sitedims), int(chir)

    @property
    def d(self):
        return int(np.prod(self.sitedims)) if self.sitedims else 1

    @property
    def shape(self):
        return (self.chil,) + self.sitedims + (self.chir,)

    def numpy(self):
        if self.chil * self.d * self.chir == 0:
            return np.zeros(self.shape)
        return self.mat.numpy().reshape(self.shape, order="F")

    @staticmethod
    def from_numpy(a, device):
        a = np.asarray(a, dtype=np.float64)
        chil, chir = a.shape[0], a.shape[-1]
        m = Mat.from_numpy(a.reshape(-1, chir, order="F"), device)
        return DeviceCore(m, chil, a.shape[1:-1], chir)


class AbstractTensorTrain:
    """abstract type AbstractTensorTrain (abstracttensortrain.jl:67) over device cores `self._cores`."""

    handle = None

    def _dev_cores(self):
        raise NotImplementedError

    def __len__(self):                                     # abstracttensortrain.jl:232-234
        return len(self._dev_cores())

    def sitetensor(self, i):                               # :220-222 (1-based like the reference)
        return self._dev_cores()[i - 1].numpy()

    @property
    def sitetensors(self):                                 # :204-206
        return [c.numpy() for c in self._dev_cores()]

    def linkdims(self):                                    # :106-110
        return [c.chil for c in self._dev_cores()[1:]]

    def linkdim(self, i):                                  # :127-129
        return self._dev_cores()[i].chil

    def sitedims(self):                                    # :151-155
        return [list(c.sitedims) for c in self._dev_cores()]

    def rank(self):                                        # :188-190
        return max(self.linkdims())

    def _core_args(self):
        cores = self._dev_cores()
        n = len(cores)
        ptrs = (C.c_void_p * n)(*[c.mat.buf.data_ptr() for c in cores])
        shapes = np.asarray([[c.chil, c.d, c.chir] for c in cores], dtype=np.int64).reshape(-1)
        lds = np.asarray([c.mat.ld for c in cores], dtype=np.int64)
        return ptrs, shapes, lds, n

    def evaluate_batch(self, points):
        """evaluate at many multi-indices at once: points (npoints, nsites) 1-based; returns numpy (npoints,)."""
        h = self.handle or _lib.default_handle()
        ptrs, shapes, lds, n = self._core_args()
        pts = np.ascontiguousarray(np.asarray(points, dtype=np.int32).reshape(-1, n))
        out = np.zeros(len(pts))
        h.check(lib().tci_tt_evaluate_f64(h.h, ptrs, p_i64(shapes), p_i64(lds), n, p_i32(pts), C.c_int64(len(pts)), p_f64(out)))
        return out

    def evaluate(self, indexset):                          # :328-342 and the multi-leg form :371-388
        n = len(self)
        if len(indexset) != n:
            raise TCIError(_lib.TCI_ERR_INVALID, f"To evaluate a tt of length {n}, you have to provide {n} indices, "
                                                 f"but there were {len(indexset)}.")
        flat = []
        for c, i in zip(self._dev_cores(), indexset):
            if np.ndim(i) == 0:
                flat.append(int(i))
            else:                                          # multi-leg index -> fused column-major index
                if len(i) != len(c.sitedims):
                    raise TCIError(_lib.TCI_ERR_INVALID, f"The index set {list(i)} does not have the correct length "
                                                         f"for the tensor of size {c.shape}.")
                lin, st = 0, 1
                for v, d in zip(i, c.sitedims):
                    lin += (int(v) - 1) * st
                    st *= d
                flat.append(lin + 1)
        return float(self.evaluate_batch([flat])[0])

    def __call__(self, indexset):                          # :400-402
        return self.evaluate(indexset)

    def sum(self):                                         # :428-441
        h = self.handle or _lib.default_handle()
        ptrs, shapes, lds, n = self._core_args()
        out = C.c_double()
        h.check(lib().tci_tt_sum_f64(h.h, ptrs, p_i64(shapes), p_i64(lds), n, C.byref(out)))
        return out.value


class TensorTrain(AbstractTensorTrain):
    """TensorTrain{Float64,N} (tensortrain.jl:58-79).  Construct from a list of numpy cores (chi_l, d..., chi_r),
    from device cores, or from another tensor train (e.g. a TensorCI2, tensortrain.jl:145-147)."""

    def __init__(self, sitetensors, handle=None):
        self.handle = handle or getattr(sitetensors, "handle", None) or _lib.default_handle()
        dev = f"cuda:{self.handle.device}"
        if isinstance(sitetensors, AbstractTensorTrain):
            cores = list(sitetensors._dev_cores())
        else:
            cores = [c if isinstance(c, DeviceCore) else DeviceCore.from_numpy(c, dev) for c in sitetensors]
        for i in range(len(cores) - 1):                    # tensortrain.jl:68-75
            if cores[i].chir != cores[i + 1].chil:
                raise TCIError(_lib.TCI_ERR_INVALID, f"The tensors at {i + 1} and {i + 2} must have consistent dimensions "
                                                     "for a tensor train.")
        self._cores = cores

    def _dev_cores(self):
        return self._cores

    def compress(self, method="LU", tolerance=1e-12, maxbonddim=None, normalizeerror=True):   # compress! tensortrain.jl:302
        return compress(self, method, tolerance, maxbonddim, normalizeerror)

    def fulltensor(self):                                  # tensortrain.jl:580-600 (small trains only; host reshape)
        from .dense import gemm
        h = self.handle
        cores = self._cores
        res = Mat(cores[0].mat.buf, cores[0].chil * cores[0].d, cores[0].chir)
        res.ld = cores[0].mat.ld
        left = cores[0].chil * cores[0].d
        for c in cores[1:]:
            nxt = c.mat.numpy().reshape(c.chil, c.d * c.chir, order="F")
            nm = Mat.from_numpy(nxt, res.buf.device)
            prod = gemm(h, res, nm)
            left *= c.d
            res = Mat.from_numpy(prod.numpy().reshape(left, c.chir, order="F"), res.buf.device)
        dims = [d for c in cores for d in c.sitedims]
        return res.numpy().reshape(dims, order="F")


def tensortrain(tci):                                      # tensortrain.jl:187-189
    return TensorTrain(tci)


# ---- factorisation / compression (tensortrain.jl:219-348) -------------------------------------------------------------
def _unpermute(h, M, perm, by_rows):
    """left(lu; permute=true) / right(lu; permute=true) (matrixlu.jl:716-746) on the device."""
    out = Mat.empty(M.rows, M.cols, M.buf.device)
    pm = np.ascontiguousarray(perm, dtype=np.int64)
    h.check(lib().tci_unpermute_f64(h.h, M.ptr, C.c_int64(M.ld), out.ptr, C.c_int64(out.ld), p_i64(pm), C.c_int64(M.rows),
                                    C.c_int64(M.cols), 1 if by_rows else 0))
    return out


def _factorize(A, method, tolerance, maxbonddim, leftorthogonal=False, normalizeerror=True, handle=None):
    """_factorize(A, method; tolerance, maxbonddim, leftorthogonal, normalizeerror) — tensortrain.jl:219-272.
    A: device `Mat` (consumed).  Returns (left Mat m x r, right Mat r x n, r).  method 'LU' or 'CI'; 'SVD' (LAPACK gesdd in
    the reference) is outside this build's scope."""
    from .matrixlu import INT_MAX, rrlu_device
    from .matrixluci import MatrixLUCI
    h = handle or _lib.default_handle()
    reltol, abstol = 1e-14, 0.0
    if normalizeerror:
        reltol = tolerance
    else:
        abstol = tolerance
    maxrank = INT_MAX if maxbonddim is None else maxbonddim
    if method == "LU":
        lu = rrlu_device(A, maxrank=maxrank, reltol=reltol, abstol=abstol, leftorthogonal=leftorthogonal, handle=h)
        r = lu.npivot
        return _unpermute(h, lu.Lmat, lu.rowpermutation, True), _unpermute(h, lu.Umat, lu.colpermutation, False), r
    if method == "CI":
        luci = MatrixLUCI(A, maxrank=maxrank, reltol=reltol, abstol=abstol, leftorthogonal=leftorthogonal, handle=h)
        return luci.left_mat(), luci.right_mat(), luci.npivots()
    if method == "SVD":
        raise TCIError(_lib.TCI_ERR_INVALID, "method=:SVD is not implemented in this build (no GPU SVD on the TCI2 hot path); use :LU or :CI")
    raise TCIError(_lib.TCI_ERR_INVALID, "Not implemented yet.")


def _dense_copy(h, M):
    """Mat with ld == rows (a fresh copy if the source is padded) so that it can be re-viewed with another 2-D shape."""
    if M.ld == M.rows or M.cols <= 1:
        return M
    out = Mat.empty(M.rows, M.cols, M.buf.device)
    out.buf.copy_(M.buf[: M.cols, : M.rows])
    return out


def compress(tt, method="LU", tolerance=1e-12, maxbonddim=None, normalizeerror=True):
    """compress!(tt, method; tolerance, maxbonddim, normalizeerror) — tensortrain.jl:302-348 (in place)."""
    from .dense import gemm
    h = tt.handle
    cores = tt._cores
    n = len(cores)
    for ell in range(n - 1):                               # left-to-right, no truncation
        c = cores[ell]
        A = _dense_copy(h, c.mat)
        A = Mat(A.buf.clone(), A.rows, A.cols) if A is c.mat else A
        left, right, r = _factorize(A, method, 0.0, None, leftorthogonal=True, handle=h)
        cores[ell] = DeviceCore(left, c.chil, c.sitedims, r)
        nx = cores[ell + 1]
        nm = _dense_copy(h, nx.mat)
        nxt = Mat(nm.buf.reshape(-1)[: nx.chil * nx.d * nx.chir].reshape(nx.d * nx.chir, nx.chil), nx.chil, nx.d * nx.chir)
        prod = gemm(h, right, nxt)                         # (r x chil) * (chil x d*chir)
        cores[ell + 1] = DeviceCore(Mat(prod.buf.reshape(-1)[: r * nx.d * nx.chir].reshape(nx.chir, r * nx.d), r * nx.d, nx.chir),
                                    r, nx.sitedims, nx.chir)
    for ell in range(n - 1, 0, -1):                        # right-to-left, truncating
        c = cores[ell]
        cm = _dense_copy(h, c.mat)
        A = Mat(cm.buf.reshape(-1)[: c.chil * c.d * c.chir].clone().reshape(c.d * c.chir, c.chil), c.chil, c.d * c.chir)
        left, right, r = _factorize(A, method, tolerance, maxbonddim, leftorthogonal=False, normalizeerror=normalizeerror, handle=h)
        rd = _dense_copy(h, right)
        cores[ell] = DeviceCore(Mat(rd.buf.reshape(-1)[: r * c.d * c.chir].reshape(c.chir, r * c.d), r * c.d, c.chir), r, c.sitedims, c.chir)
        pv = cores[ell - 1]
        prod = gemm(h, _dense_copy(h, pv.mat), left)       # (chil*d x chir) * (chir x r)
        cores[ell - 1] = DeviceCore(prod, pv.chil, pv.sitedims, r)
    return None
