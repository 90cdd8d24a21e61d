"""Host-side mirror of src/matrixluci.jl: MatrixLUCI wraps an rrLU and forms the cross-interpolation
factors on the device (csrc/dense.cu)."""
import numpy as np

from . import _lib
from .dense import Mat, gemm, luci_factors
from .matrixlu import INT_MAX, rrlu_device


class MatrixLUCI:
    """MatrixLUCI(A; maxrank, reltol, abstol, leftorthogonal) — matrixluci.jl:32-57.  `A` is a host matrix
    (numpy; copied to the device, not modified) or a `Mat` (device; consumed in place like rrlu!)."""

    def __init__(self, A, maxrank=INT_MAX, reltol=1e-14, abstol=0.0, leftorthogonal=True, handle=None):
        self.handle = handle or _lib.default_handle()
        if not isinstance(A, Mat):
            A = Mat.from_numpy(np.asarray(A, dtype=np.float64), f"cuda:{self.handle.device}", pad=16)
        self.m, self.n = A.rows, A.cols
        self.lu = rrlu_device(A, maxrank=maxrank, reltol=reltol, abstol=abstol, leftorthogonal=leftorthogonal,
                              handle=self.handle)
        self._left = self._right = None

    @classmethod
    def from_function(cls, f, matrixsize, I0=(), J0=(), handle=None, **kw):
        """MatrixLUCI(Float64, f, matrixsize, I0, J0; kwargs...) — matrixluci.jl:75-84: never holds the full matrix;
        kwargs go to matrixlu.rrlu_function (pivotsearch="rook" runs the rook search arrlu)."""
        from .matrixlu import rrlu_function
        self = cls.__new__(cls)
        self.handle = handle or _lib.default_handle()
        self.m, self.n = int(matrixsize[0]), int(matrixsize[1])
        self.lu = rrlu_function(f, matrixsize, I0, J0, handle=self.handle, **kw)
        self._left = self._right = None
        return self

    @classmethod
    def from_rrlu(cls, lu, m, n, handle=None):
        """MatrixLUCI around an rrLU that was computed elsewhere (the column-sharded factorisation): `lu` carries device
        factors `Lmat` (m x r) and `Umat` (r x n)."""
        self = cls.__new__(cls)
        self.handle = handle or _lib.default_handle()
        self.m, self.n = int(m), int(n)
        self.lu = lu
        self._left = self._right = None
        return self

    def size(self):                                        # matrixluci.jl:94-96
        return (self.m, self.n)

    def npivots(self):                                     # matrixluci.jl:121-123
        return self.lu.npivot

    def rowindices(self):                                  # matrixluci.jl:133-135
        return self.lu.rowindices()

    def colindices(self):                                  # matrixluci.jl:145-147
        return self.lu.colindices()

    def pivoterrors(self):                                 # matrixluci.jl:294-296
        return self.lu.pivoterrors()

    def lastpivoterror(self):                              # matrixluci.jl:309-311
        return self.lu.lastpivoterror()

    def _factors(self, leftorthogonal, want_left, want_right):
        lu = self.lu
        return luci_factors(self.handle, self.m, self.n, lu.npivot, leftorthogonal, lu.rowpermutation, lu.colpermutation,
                            lu.Lmat, lu.Umat, want_left, want_right)

    def left_mat(self):                                    # matrixluci.jl:256-262, device result
        if self._left is None:
            self._left = self._factors(self.lu.leftorthogonal, True, False)[0]
        return self._left

    def right_mat(self):                                   # matrixluci.jl:277-283, device result
        if self._right is None:
            self._right = self._factors(self.lu.leftorthogonal, False, True)[1]
        return self._right

    def left(self):
        return self.left_mat().numpy()

    def right(self):
        return self.right_mat().numpy()

    # the four named factor forms (matrixluci.jl:161-241), independent of lu.leftorthogonal
    def colstimespivotinv(self):
        return self._factors(True, True, False)[0].numpy()

    def rowmatrix(self):
        return self._factors(True, False, True)[1].numpy()

    def colmatrix(self):
        return self._factors(False, True, False)[0].numpy()

    def pivotinvtimesrows(self):
        return self._factors(False, False, True)[1].numpy()
