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

    def copy(self):                                        # deepcopy(tt)
        return TensorTrain([DeviceCore(Mat(c.mat.buf.clone(), c.mat.rows, c.mat.cols), c.chil, c.sitedims, c.chir) for c in self._cores],
                           handle=self.handle)

    def multiply_(self, a):                                # multiply!(tt, a) — tensortrain.jl:359-368: scales the LAST core
        c = self._cores[-1]
        self.handle.check(lib().tci_elementwise_f64(self.handle.h, 1, C.c_double(float(a)), C.c_double(0.0), c.mat.ptr,
                                                    C.c_int64(c.mat.buf.numel())))

    def divide_(self, a):                                  # divide!(tt, a) — tensortrain.jl:412-415
        c = self._cores[-1]                                # true division, like `./ a`
        self.handle.check(lib().tci_elementwise_f64(self.handle.h, 5, C.c_double(float(a)), C.c_double(0.0), c.mat.ptr,
                                                    C.c_int64(c.mat.buf.numel())))

    def multiply(self, a):                                 # :375-384
        t = self.copy(); t.multiply_(a); return t

    def divide(self, a):                                   # :422-426
        t = self.copy(); t.divide_(a); return t

    def __mul__(self, a):                                  # :399-405
        return self.multiply(a)

    __rmul__ = __mul__

    def __truediv__(self, a):                              # :433-435
        return self.divide(a)

    def __add__(self, other):                              # abstracttensortrain.jl:578-580
        return add(self, other)

    def __sub__(self, other):                              # :590-592
        return subtract(self, other)

    def reverse(self):                                     # Base.reverse — tensortrain.jl:452-456
        out = []
        for c in reversed(self._cores):                    # permutedims(T, (N, 2:N-1..., 1)): swap the two bond legs (device copy)
            m = _dense_copy(self.handle, c.mat)
            t = m.buf.reshape(-1)[: c.chil * c.d * c.chir].reshape(c.chir, c.d, c.chil).permute(2, 1, 0).contiguous()
            out.append(DeviceCore(Mat(t.reshape(c.chil, c.d * c.chir), c.chir * c.d, c.chil), c.chir, c.sitedims, c.chil))
        return TensorTrain(out, handle=self.handle)

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


def _addtttensor(A, B, factorA=1.0, factorB=1.0, lefttensor=False, righttensor=False, handle=None):
    """_addtttensor (abstracttensortrain.jl:467-485): block-diagonal stacking [factorA*A 0; 0 factorB*B] of two device cores
    (first/last cores stack along one bond only).  Data movement (torch slice copies) + the affine element-wise kernel."""
    import torch
    h = handle or _lib.default_handle()
    if len(A.sitedims) != len(B.sitedims):
        raise TCIError(_lib.TCI_ERR_INVALID, "Elementwise addition only works if both tensors have the same indices, but A and B "
                       f"have different numbers ({len(A.sitedims) + 2} and {len(B.sitedims) + 2}) of indices.")
    if tuple(A.sitedims) != tuple(B.sitedims):
        raise TCIError(_lib.TCI_ERR_INVALID, "Elementwise addition only works if both tensors have the same local dimensions.")
    off1 = 0 if lefttensor else A.chil
    off3 = 0 if righttensor else A.chir
    cl, cr, d = off1 + B.chil, off3 + B.chir, A.d
    out = torch.zeros((cr, d, cl), dtype=torch.float64, device=A.mat.buf.device)      # memory of the (cl, d, cr) column-major core

    def block(core):
        return _dense_copy(h, core.mat).buf.reshape(-1)[: core.chil * core.d * core.chir].reshape(core.chir, core.d, core.chil)

    def scaled(core, factor):
        t = block(core).clone()
        if factor != 1.0:
            h.check(lib().tci_elementwise_f64(h.h, 1, C.c_double(factor), C.c_double(0.0), C.c_void_p(t.data_ptr()), C.c_int64(t.numel())))
        return t
    out[: A.chir, :, : A.chil].copy_(scaled(A, factorA))             # :482
    out[off3:, :, off1:].copy_(scaled(B, factorB))                   # :483
    return DeviceCore(Mat(out.reshape(cr, d * cl), cl * d, cr), cl, A.sitedims, cr)


def add(lhs, rhs, factorlhs=1.0, factorrhs=1.0, tolerance=0.0, maxbonddim=None):
    """add(lhs, rhs; factorlhs, factorrhs, tolerance, maxbonddim) — abstracttensortrain.jl:524-545 (then compress!(:SVD))."""
    if len(lhs) != len(rhs):
        raise TCIError(_lib.TCI_ERR_INVALID, f"Two tensor trains with different length ({len(lhs)} and {len(rhs)}) cannot be added elementwise.")
    L = len(lhs)
    h = lhs.handle or _lib.default_handle()
    a, b = lhs._dev_cores(), rhs._dev_cores()
    tt = TensorTrain([_addtttensor(a[ell], b[ell], factorA=factorlhs if ell == L - 1 else 1.0,
                                   factorB=factorrhs if ell == L - 1 else 1.0, lefttensor=ell == 0, righttensor=ell == L - 1,
                                   handle=h) for ell in range(L)], handle=h)
    compress(tt, "SVD", tolerance=tolerance, maxbonddim=maxbonddim)
    return tt


def subtract(lhs, rhs, tolerance=0.0, maxbonddim=None):     # abstracttensortrain.jl:560-565
    return add(lhs, rhs, factorrhs=-1.0, tolerance=tolerance, maxbonddim=maxbonddim)


def norm2(tt):
    """LA.norm2(tt) — abstracttensortrain.jl:625-634: sum of |f|^2 over all indices.  The reference multiplies the
    (chi^2 x chi^2) transfer matrices; here the same contraction runs as a left-to-right environment
    E <- sum_s T_s^T E T_s (two GEMMs per site, chi^3 d work instead of chi^4 d)."""
    from .dense import gemm
    import torch
    h = tt.handle or _lib.default_handle()
    cores = tt._dev_cores()
    dev = cores[0].mat.buf.device
    E = Mat(torch.ones((1, 1), dtype=torch.float64, device=dev), 1, 1)
    if cores[0].chil != 1 or cores[-1].chir != 1:
        raise TCIError(_lib.TCI_ERR_INVALID, "norm2: the tensor train must have boundary bond dimensions 1")
    for c in cores:
        T = _dense_copy(h, c.mat)                          # (chil*d) x chir
        Tw = Mat(T.buf.reshape(-1)[: c.chil * c.d * c.chir].reshape(c.d * c.chir, c.chil), c.chil, c.d * c.chir)   # chil x (d*chir)
        X = gemm(h, E, Tw)                                 # E (chil x chil) * T (chil x d*chir)
        Xm = Mat(X.buf.reshape(-1)[: c.chil * c.d * c.chir].reshape(c.chir, c.chil * c.d), c.chil * c.d, c.chir)  # (chil*d) x chir
        E = gemm(h, T, Xm, transa=True)                    # T^T (chir x chil*d) * X
    return float(E.numpy()[0, 0])


def norm(tt):                                              # abstracttensortrain.jl:646-648
    return float(np.sqrt(norm2(tt)))


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
    A: device `Mat` (consumed).  Returns (left Mat m x r, right Mat r x n, r).  method 'LU', 'CI' or 'SVD' (one-sided Jacobi
    on the device, csrc/svd.cu, where the reference calls LAPACK gesdd)."""
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
    if method == "SVD":                                    # tensortrain.jl:244-262
        from .dense import diag_scale, svd
        U, S, Vt = svd(h, A)
        S2 = S * S
        err = np.array([np.sum(S2[i + 1:]) for i in range(len(S))])          # err[n] = sum(S[n+1:end].^2)
        tot = np.sum(S2)
        with np.errstate(divide="ignore", invalid="ignore"):
            nerr = err / tot
        def first(mask):                                   # replacenothing(findfirst(...), length)
            w = np.nonzero(mask)[0]
            return int(w[0]) + 1 if len(w) else len(S)
        trunci = min(first(err < abstol ** 2), first(nerr < reltol ** 2), maxrank)
        left = Mat(U.buf[:max(trunci, 1)], U.rows, trunci)
        right = Vt if trunci == Vt.rows else Mat(Vt.buf[:, :max(trunci, 1)].contiguous(), trunci, Vt.cols)   # Vt[1:trunci, :]
        if leftorthogonal:
            diag_scale(h, right, S[:trunci], True)
        else:
            diag_scale(h, left, S[:trunci], False)
        return left, right, trunci
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
