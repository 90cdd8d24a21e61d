"""Host-side mirror of src/cachedtensortrain.jl and the diagnostics of src/globalsearch.jl: `TTCache` presents a tensor train
as a `BatchEvaluator` (so that TCI can run ON a tensor train: recompression, `crossinterpolate2(ttcache, ...)`), and
`estimatetrueerror` / `_floatingzone` search for the largest |f - tt|.

The reference memoises left/right environment vectors in dictionaries (cachedtensortrain.jl:165-243) because every point
costs a chain of host mat-vecs.  On the device the whole batch goes through the batched chain kernel (K4b, csrc/tt.cu) in
one launch, which is cheaper than any per-point lookup, so there is no cache to keep."""
import itertools

import numpy as np

from . import _lib
from ._lib import TCIError
from .batcheval import BatchEvaluator, as_table
from .tensortrain import AbstractTensorTrain, DeviceCore, TensorTrain


class TTCache(BatchEvaluator, AbstractTensorTrain):
    """TTCache{Float64}(sitetensors[, sitedims]) / TTCache(tt) — cachedtensortrain.jl:63-123."""

    _CHUNK = 1 << 22                                       # points per K4b call

    def __init__(self, tt, sitedims=None, handle=None):
        if isinstance(tt, AbstractTensorTrain):
            handle = handle or tt.handle
            cores = list(tt._dev_cores())
        else:
            handle = handle or _lib.default_handle()
            cores = [c if isinstance(c, DeviceCore) else DeviceCore.from_numpy(c, f"cuda:{handle.device}") for c in tt]
        if sitedims is None:
            sitedims = [list(c.sitedims) for c in cores]
        if len(sitedims) != len(cores):
            raise TCIError(_lib.TCI_ERR_INVALID, "The number of site tensors and site dimensions must be the same.")
        for n, (c, sd) in enumerate(zip(cores, sitedims)):
            if int(np.prod(sd)) != c.d:
                raise TCIError(_lib.TCI_ERR_INVALID, f"Site dimensions do not match the site tensor dimensions at {n + 1}.")
        BatchEvaluator.__init__(self, [c.d for c in cores], handle)
        self._cores = [DeviceCore(c.mat, c.chil, (c.d,), c.chir) for c in cores]     # fused legs, like the reference's reshape
        self._sitedims = [list(sd) for sd in sitedims]

    def _dev_cores(self):
        return self._cores

    def sitedims(self):                                    # :132-134
        return self._sitedims

    def __len__(self):                                     # :141-143
        return len(self._cores)

    def evaluate(self, indexset):                          # :261-263
        return AbstractTensorTrain.evaluate(self, indexset)

    def evaluate_points(self, pts):
        return AbstractTensorTrain.evaluate_batch(self, np.asarray(pts, dtype=np.int32))

    def __call__(self, *args):                             # :271-273 and the batch form :290-323
        if len(args) == 1:
            return self.evaluate(args[0])
        return BatchEvaluator.__call__(self, *args)

    def evaluate_batch(self, left, right=None, M=None, out=None, stride_c=0, stride_j=0, maxabs=None):
        """Two roles: AbstractTensorTrain.evaluate_batch(points) and the BatchEvaluator batch form (left, right, M)."""
        if right is None:
            return AbstractTensorTrain.evaluate_batch(self, left)
        import torch
        lt, rt = as_table(left), as_table(right)
        nleft, nright = lt.shape[0], rt.shape[0]
        dev = f"cuda:{self.handle.device}"
        if nleft * nright == 0:
            return torch.zeros((0,) * (M + 2), dtype=torch.float64, device=dev), maxabs
        nl, nr = lt.shape[1], rt.shape[1]
        if len(self) - nl - nr != M:
            raise TCIError(_lib.TCI_ERR_INVALID, "Invalid number of central indices")
        cd = [int(d) for d in self.localdims[nl:nl + M]]
        ncent = int(np.prod(cd)) if cd else 1
        cents = np.asarray([tuple(reversed(c)) for c in itertools.product(*[range(1, d + 1) for d in reversed(cd)])],
                           dtype=np.int32).reshape(ncent, M)
        vals = np.empty((nright, ncent, nleft))
        per_j = ncent * nleft
        rows_per_chunk = max(1, self._CHUNK // per_j)
        for j0 in range(0, nright, rows_per_chunk):
            j1 = min(nright, j0 + rows_per_chunk)
            pts = np.empty((j1 - j0, ncent, nleft, len(self)), dtype=np.int32)
            pts[..., :nl] = lt[None, None, :, :]
            pts[..., nl:nl + M] = cents[None, :, None, :]
            pts[..., nl + M:] = rt[j0:j1, None, None, :]
            vals[j0:j1] = AbstractTensorTrain.evaluate_batch(self, pts.reshape(-1, len(self))).reshape(j1 - j0, ncent, nleft)
        sc = stride_c if stride_c > 0 else nleft
        sj = stride_j if stride_j > 0 else sc * ncent
        if out is None:
            out = torch.empty((nright,) + tuple(reversed(cd)) + (nleft,), dtype=torch.float64, device=dev)
        torch.as_strided(out.reshape(-1), (nright, ncent, nleft), (sj, sc, 1)).copy_(torch.from_numpy(vals))
        if maxabs is not None:
            m = float(np.max(np.abs(vals))) if not np.isnan(vals).any() else float("nan")
            a = abs(maxabs)
            maxabs = float("nan") if (np.isnan(a) or np.isnan(m)) else max(a, m)
        return out, maxabs


def _floatingzone(ttcache, f, earlystoptol=float("inf"), nsweeps=None, initp=None, rng=None):
    """_floatingzone (globalsearch.jl:119-186): coordinate ascent on |f - tt| from `initp`; every site scan is one batch of
    f and one of the tensor train."""
    from .tensorci2 import _point_values, makebatchevaluatable
    if nsweeps is not None and nsweeps <= 0:
        raise TCIError(_lib.TCI_ERR_INVALID, "nsweeps should be positive!")
    localdims = [int(sd[0]) for sd in ttcache.sitedims()]
    n = len(ttcache)
    f = makebatchevaluatable(f, localdims, ttcache.handle)
    rng = rng or np.random.default_rng()
    pivot = [int(rng.integers(1, d + 1)) for d in localdims] if initp is None else [int(v) for v in initp]
    maxerror = abs(f(pivot) - ttcache(pivot))
    isweep = 0
    while nsweeps is None or isweep < nsweeps:
        isweep += 1
        prev = maxerror
        for ipos in range(n):
            pts = np.tile(np.asarray(pivot, dtype=np.int32), (localdims[ipos], 1))
            pts[:, ipos] = np.arange(1, localdims[ipos] + 1)
            err = np.abs(_point_values(f, pts) - ttcache.evaluate_points(pts))       # filltensor(..., Val(1)) for f and for tt
            pivot[ipos] = int(np.argmax(err)) + 1
            maxerror = max(float(np.max(err)), maxerror)
        if maxerror == prev or maxerror > earlystoptol:
            break
    return pivot, maxerror


def estimatetrueerror(tt, f, nsearch=100, initialpoints=None, rng=None):
    """estimatetrueerror(tt, f; nsearch, initialpoints) — globalsearch.jl:52-73: (point, error) pairs found by floating-zone
    searches, sorted by decreasing error, duplicates removed."""
    if nsearch <= 0 and initialpoints is None:
        raise TCIError(_lib.TCI_ERR_INVALID, "No search is performed")
    if nsearch < 0:
        raise TCIError(_lib.TCI_ERR_INVALID, "nsearch must be non-negative")
    rng = rng or np.random.default_rng()
    if nsearch > 0 and initialpoints is None:
        initialpoints = [[int(rng.integers(1, sd[0] + 1)) for sd in tt.sitedims()] for _ in range(nsearch)]
    ttcache = tt if isinstance(tt, TTCache) else TTCache(tt)
    found = [_floatingzone(ttcache, f, initp=list(p), rng=rng) for p in initialpoints]
    order = sorted(range(len(found)), key=lambda i: -found[i][1])                    # sortperm(..., rev=true) is stable
    out, seen = [], set()
    for i in order:
        key = (tuple(found[i][0]), found[i][1])
        if key not in seen:
            seen.add(key)
            out.append((list(found[i][0]), found[i][1]))
    return out
