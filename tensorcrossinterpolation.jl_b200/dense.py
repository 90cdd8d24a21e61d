"""Thin wrappers over the K4a entry points (csrc/dense.cu) on torch CUDA tensors.

Storage convention used throughout the host code: a logical column-major matrix M (rows x cols, leading
dimension ld) lives in a torch tensor `t` of shape (cols, ld) — `t.T[:rows]` is M.  `Mat` bundles that."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import lib, p_i64


class Mat:
    """Column-major device matrix: rows x cols with leading dimension ld, backed by torch tensor (cols, ld)."""

    def __init__(self, buf, rows, cols):
        self.buf, self.rows, self.cols = buf, int(rows), int(cols)
        self.ld = int(buf.shape[1]) if buf.dim() == 2 else max(int(rows), 1)

    @staticmethod
    def empty(rows, cols, device, pad=1):
        import torch
        ld = max((int(rows) + pad - 1) // pad * pad, 1)
        return Mat(torch.empty((max(int(cols), 1), ld), dtype=torch.float64, device=device), rows, cols)

    @staticmethod
    def from_numpy(a, device, pad=1):
        import torch
        a = np.asarray(a, dtype=np.float64)
        m = Mat.empty(a.shape[0], a.shape[1], device, pad)
        if a.size:
            m.buf[: a.shape[1], : a.shape[0]].copy_(torch.from_numpy(np.ascontiguousarray(a.T)))
        return m

    @property
    def ptr(self):
        return C.c_void_p(self.buf.data_ptr())

    def view(self):
        """torch view with logical shape (rows, cols)."""
        return self.buf[: self.cols, : self.rows].T

    def numpy(self):
        return np.asfortranarray(self.view().cpu().numpy())


def gemm(h, A, B, C_out=None, transa=False, transb=False, alpha=1.0, beta=0.0):
    """C = alpha * op(A) * op(B) + beta * C on `Mat`s."""
    M = A.cols if transa else A.rows
    K = A.rows if transa else A.cols
    K2 = B.cols if transb else B.rows
    N = B.rows if transb else B.cols
    if K != K2:
        raise _lib.TCIError(_lib.TCI_ERR_INVALID, "gemm: inner dimensions do not match")
    if C_out is None:
        C_out = Mat.empty(M, N, A.buf.device)
    sam, sak = (A.ld, 1) if transa else (1, A.ld)
    sbk, sbn = (B.ld, 1) if transb else (1, B.ld)
    h.check(lib().tci_gemm_f64(h.h, C.c_int64(M), C.c_int64(N), C.c_int64(K), C.c_double(alpha), A.ptr, C.c_int64(sam),
                               C.c_int64(sak), B.ptr, C.c_int64(sbk), C.c_int64(sbn), C.c_double(beta), C_out.ptr,
                               C.c_int64(1), C.c_int64(C_out.ld), C.c_int64(1), C.c_int64(0), C.c_int64(0), C.c_int64(0)))
    return C_out


def luci_factors(h, m, n, r, leftorthogonal, rowperm, colperm, L, U, want_left=True, want_right=True):
    """left(luci), right(luci) as `Mat`s (matrixluci.jl:256-283) from rrLU outputs L (m x r), U (r x n)."""
    dev = L.buf.device
    left = Mat.empty(m, r, dev) if want_left else None
    right = Mat.empty(r, n, dev) if want_right else None
    rp = np.ascontiguousarray(rowperm, dtype=np.int64)
    cp = np.ascontiguousarray(colperm, dtype=np.int64)
    h.check(lib().tci_luci_factors_f64(h.h, C.c_int64(m), C.c_int64(n), C.c_int64(r), int(bool(leftorthogonal)), p_i64(rp),
                                       p_i64(cp), L.ptr, C.c_int64(L.ld), U.ptr, C.c_int64(U.ld),
                                       left.ptr if want_left else None, C.c_int64(left.ld if want_left else 1),
                                       right.ptr if want_right else None, C.c_int64(right.ld if want_right else 1)))
    return left, right


def solve_right(h, B, P):
    """X = B * inv(P) (tensorci2.jl:626)."""
    if P.rows != P.cols or B.cols != P.rows:
        raise _lib.TCIError(_lib.TCI_ERR_NOT_SQUARE, "Pivot matrix is not square!")
    X = Mat.empty(B.rows, B.cols, B.buf.device)
    h.check(lib().tci_solve_right_f64(h.h, B.ptr, C.c_int64(B.rows), C.c_int64(B.ld), P.ptr, C.c_int64(P.rows),
                                      C.c_int64(P.ld), X.ptr, C.c_int64(X.ld)))
    return X


def trsm_right(h, T, r, lower, unit, B, left=False):
    """In place on B: X * T = B (left=False, B is q x r) or T * X = B (left=True, B is r x q; the right solve on transposed
    strides) with the leading r x r triangle of `T` — tci_trsm_right_f64.  Serves cols2Lmatrix!/rows2Umatrix!
    (matrixlu.jl:627-674) of the rook search."""
    if r == 0 or B.rows * B.cols == 0:
        return B
    if left:
        h.check(lib().tci_trsm_right_f64(h.h, T.ptr, C.c_int64(T.ld), C.c_int64(1), C.c_int64(r), int(not lower), int(bool(unit)),
                                         B.ptr, C.c_int64(B.ld), C.c_int64(1), C.c_int64(B.cols)))
    else:
        h.check(lib().tci_trsm_right_f64(h.h, T.ptr, C.c_int64(1), C.c_int64(T.ld), C.c_int64(r), int(bool(lower)), int(bool(unit)),
                                         B.ptr, C.c_int64(1), C.c_int64(B.ld), C.c_int64(B.rows)))
    return B


def svd(h, A):
    """Thin SVD of the device matrix A (consumed): (U Mat m x k, S numpy[k] descending, Vt Mat k x n) — tci_svd_f64,
    replaces LinearAlgebra.svd(A) (tensortrain.jl:244)."""
    m, n = A.rows, A.cols
    k = min(m, n)
    dev = A.buf.device
    U, Vt = Mat.empty(m, max(k, 1), dev), Mat.empty(max(k, 1), n, dev)
    S = np.zeros(max(k, 1))
    sweeps = C.c_int()
    h.check(lib().tci_svd_f64(h.h, A.ptr, C.c_int64(m), C.c_int64(n), C.c_int64(A.ld), U.ptr, C.c_int64(U.ld),
                              S.ctypes.data_as(C.POINTER(C.c_double)), Vt.ptr, C.c_int64(Vt.ld), C.byref(sweeps)))
    U.cols, Vt.rows = k, k
    return U, S[:k], Vt


def diag_scale(h, M, s, by_rows):
    """M[i, j] *= s[i] (by_rows) or s[j], in place — tci_diag_scale_f64."""
    s = np.ascontiguousarray(s, dtype=np.float64)
    h.check(lib().tci_diag_scale_f64(h.h, M.ptr, C.c_int64(M.rows), C.c_int64(M.cols), C.c_int64(M.ld),
                                     s.ctypes.data_as(C.POINTER(C.c_double)), 1 if by_rows else 0))
    return M
