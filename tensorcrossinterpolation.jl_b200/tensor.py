"""Device tensors for the contraction code: an N-d column-major VIEW (offset + element strides) over a flat torch CUDA
buffer, and `_contract` (src/contraction.jl:193-215).  `permutedims`, slicing and `reshape` are view operations here; a
view is only materialised (csrc/tensor.cu: tci_strided_copy_f64) when its index groups cannot be expressed as the two
element strides the GEMM kernel (csrc/dense.cu: tci_gemm_f64) accepts."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import TCIError, lib, p_i64


def _colmajor_strides(shape):
    st, acc = [], 1
    for s in shape:
        st.append(acc)
        acc *= int(s)
    return tuple(st)


class DTensor:
    def __init__(self, buf, shape, strides=None, offset=0):
        self.buf = buf                                   # flat torch.float64 CUDA tensor
        self.shape = tuple(int(s) for s in shape)
        self.strides = tuple(int(s) for s in strides) if strides is not None else _colmajor_strides(self.shape)
        self.offset = int(offset)

    # ---- construction
    @staticmethod
    def empty(shape, device):
        import torch
        n = int(np.prod(shape)) if len(shape) else 1
        return DTensor(torch.empty(max(n, 1), dtype=torch.float64, device=device), shape)

    @staticmethod
    def from_numpy(a, device):
        import torch
        a = np.asarray(a, dtype=np.float64)
        flat = np.ascontiguousarray(a.reshape(-1, order="F"))
        buf = torch.from_numpy(flat).to(device) if flat.size else torch.empty(1, dtype=torch.float64, device=device)
        return DTensor(buf, a.shape)

    @staticmethod
    def ones(shape, device):
        import torch
        n = int(np.prod(shape))
        return DTensor(torch.ones(n, dtype=torch.float64, device=device), shape)

    # ---- properties
    @property
    def ndim(self):
        return len(self.shape)

    @property
    def size(self):
        return int(np.prod(self.shape)) if self.shape else 1

    @property
    def device(self):
        return self.buf.device

    @property
    def ptr(self):
        return C.c_void_p(self.buf.data_ptr() + 8 * self.offset)

    def is_dense(self):
        return self.strides == _colmajor_strides(self.shape) or self.size <= 1

    # ---- views (no data movement)
    def permute(self, perm):
        """permutedims(a, perm) as a view; perm is 0-based."""
        return DTensor(self.buf, [self.shape[p] for p in perm], [self.strides[p] for p in perm], self.offset)

    def select(self, dim, idx0):
        """a[.., idx, ..] with dimension `dim` dropped (0-based)."""
        if not 0 <= idx0 < self.shape[dim]:
            raise TCIError(_lib.TCI_ERR_INVALID, f"index {idx0 + 1} out of range 1..{self.shape[dim]}")
        shp = self.shape[:dim] + self.shape[dim + 1:]
        st = self.strides[:dim] + self.strides[dim + 1:]
        return DTensor(self.buf, shp, st, self.offset + idx0 * self.strides[dim])

    def keep(self, dim, idx0):
        """a[.., idx:idx, ..]: dimension kept with extent 1."""
        shp = self.shape[:dim] + (1,) + self.shape[dim + 1:]
        return DTensor(self.buf, shp, self.strides, self.offset + idx0 * self.strides[dim])

    def reshape(self, shape, handle=None):
        shape = tuple(int(s) for s in shape)
        if int(np.prod(shape)) != self.size:
            raise TCIError(_lib.TCI_ERR_INVALID, f"reshape: cannot reshape {self.shape} to {shape}")
        t = self if self.is_dense() else self.contiguous(handle)
        return DTensor(t.buf, shape, None, t.offset)

    # ---- materialisation
    def contiguous(self, handle=None):
        if self.is_dense():
            return self
        h = handle or _lib.default_handle()
        out = DTensor.empty(self.shape, self.device)
        if self.size:
            shp = np.asarray(self.shape, dtype=np.int64)
            st = np.asarray(self.strides, dtype=np.int64)
            h.check(lib().tci_strided_copy_f64(h.h, C.c_void_p(self.buf.data_ptr()), C.c_int64(self.offset), p_i64(shp), p_i64(st),
                                               len(self.shape), out.ptr))
        return out

    def numpy(self, handle=None):
        t = self.contiguous(handle)
        if t.size == 0:
            return np.zeros(t.shape)
        flat = t.buf[t.offset:t.offset + t.size].cpu().numpy()
        return flat.reshape(t.shape, order="F")


def _merge(shape, strides):
    """(extent, stride) of an index group seen as ONE index, or None if the group is not uniformly strided."""
    dims = [(s, t) for s, t in zip(shape, strides) if s != 1]
    if not dims:
        return 1, 1
    for (s0, t0), (_, t1) in zip(dims[:-1], dims[1:]):
        if t1 != t0 * s0:
            return None
    return int(np.prod([s for s, _ in dims])), dims[0][1]


def _contract(a, b, idx_a, idx_b, handle=None):
    """_contract(a, b, idx_a, idx_b) — contraction.jl:193-215 (idx are 1-based like the reference).  Result: dense DTensor
    of shape (rest dims of a..., rest dims of b...)."""
    h = handle or _lib.default_handle()
    ia, ib = [int(i) - 1 for i in idx_a], [int(i) - 1 for i in idx_b]
    if len(ia) != len(ib):
        raise TCIError(_lib.TCI_ERR_INVALID, "length(idx_a) != length(idx_b)")
    if len(set(ia)) != len(ia):
        raise TCIError(_lib.TCI_ERR_INVALID, "idx_a contains duplicate elements")
    if len(set(ib)) != len(ib):
        raise TCIError(_lib.TCI_ERR_INVALID, "idx_b contains duplicate elements")
    if not all(0 <= i < a.ndim for i in ia):
        raise TCIError(_lib.TCI_ERR_INVALID, "idx_a contains elements out of range")
    if not all(0 <= i < b.ndim for i in ib):
        raise TCIError(_lib.TCI_ERR_INVALID, "idx_b contains elements out of range")
    ra = [d for d in range(a.ndim) if d not in ia]
    rb = [d for d in range(b.ndim) if d not in ib]
    if [a.shape[i] for i in ia] != [b.shape[i] for i in ib]:
        raise TCIError(_lib.TCI_ERR_INVALID, "DimensionMismatch: contracted extents differ")
    ap, bp = a.permute(ra + ia), b.permute(ib + rb)
    nra, nk = len(ra), len(ia)

    def groups(t, split):
        g1 = _merge(t.shape[:split], t.strides[:split])
        g2 = _merge(t.shape[split:], t.strides[split:])
        if g1 is None or g2 is None:
            t = t.contiguous(h)
            g1 = _merge(t.shape[:split], t.strides[:split])
            g2 = _merge(t.shape[split:], t.strides[split:])
        return t, g1, g2

    ap, (M, sam), (K, sak) = groups(ap, nra)
    bp, (K2, sbk), (N, sbn) = groups(bp, nk)
    assert K == K2
    out_shape = tuple(a.shape[d] for d in ra) + tuple(b.shape[d] for d in rb)
    out = DTensor.empty(out_shape, a.device)
    if M * N:
        h.check(lib().tci_gemm_f64(h.h, C.c_int64(M), C.c_int64(N), C.c_int64(K), C.c_double(1.0), ap.ptr, C.c_int64(sam),
                                   C.c_int64(sak), bp.ptr, C.c_int64(sbk), C.c_int64(sbn), C.c_double(0.0), out.ptr,
                                   C.c_int64(1), C.c_int64(M), C.c_int64(1), C.c_int64(0), C.c_int64(0), C.c_int64(0)))
    return out
