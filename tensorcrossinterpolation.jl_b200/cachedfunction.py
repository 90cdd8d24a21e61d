"""Host-side mirror of src/cachedfunction.jl: `CachedFunction` memoises an expensive HOST function (a Python callable or
any BatchEvaluator whose values cost more than a dictionary lookup) keyed by the mixed-radix integer of the multi-index
(cachedfunction.jl:197-199).  Device functors need no cache — re-evaluating them in the K1 kernel is cheaper than any
lookup — so this class exists for the host-function case of the reference API: the batch form evaluates only the
points that are not cached yet, uploads the assembled (|left|, centre..., |right|) block once, and everything downstream
runs on the GPU as usual."""
import itertools

import numpy as np

from . import _lib
from ._lib import TCIError
from .batcheval import BatchEvaluator, as_table


class CachedFunction(BatchEvaluator):
    """CachedFunction{Float64,K}(f, localdims[, cache]) — cachedfunction.jl:53-113.  Python integers are unbounded, so the
    key type parameter K (UInt32 … UInt512 / BigInt) and its overflow check (:74-84) have no counterpart."""

    def __init__(self, f, localdims, cache=None, handle=None):
        super().__init__(localdims, handle)
        self.f = f
        self.cache = {} if cache is None else cache
        self.coeffs = [1] * len(self.localdims)                       # :63-66 coeffs[n] = prod(localdims[1:n-1])
        for n in range(1, len(self.localdims)):
            self.coeffs[n] = self.coeffs[n - 1] * int(self.localdims[n - 1])
        total = self.coeffs[-1] * int(self.localdims[-1])
        self._np_coeffs = np.asarray(self.coeffs, dtype=np.int64) if total < 2 ** 62 else None

    def cachedata(self):                                              # :135-137
        return self.cache

    cacheddata = cachedata

    def ncacheddata(self):                                            # :151-153
        return len(self.cache)

    def _key(self, indexset):                                         # :197-199
        if len(indexset) != len(self.localdims):
            raise TCIError(_lib.TCI_ERR_INVALID, f"Invalid length of the indexset: {len(indexset)} != {len(self.localdims)}")
        return sum((int(v) - 1) * c for v, c in zip(indexset, self.coeffs))

    key = _key

    def haskey(self, indexset):                                       # :206-208
        return self._key(indexset) in self.cache

    def clearcache(self):                                             # :304-307
        self.cache.clear()

    def _call_f(self, idx):
        return float(self.f(idx))

    def __call__(self, *args):
        if len(args) == 1:                                            # :231-237  get!(cache, key) do f(indexset) end
            idx = [int(v) for v in args[0]]
            k = self._key(idx)
            if k not in self.cache:
                self.cache[k] = self._call_f(idx)
            return self.cache[k]
        return super().__call__(*args)

    def evaluate_points(self, pts):
        return np.array([self(list(p)) for p in np.asarray(pts)])

    def evaluate_batch(self, left, right, M, out=None, stride_c=0, stride_j=0, maxabs=None):
        """Batch form (:255-289): cached values are reused, the missing ones are computed (through the wrapped function's
        own batch form when it has one) and stored; the block is uploaded into `out` with the K1 strides."""
        import torch
        lt, rt = as_table(left), as_table(right)
        nleft, nright = lt.shape[0], rt.shape[0]
        dev = f"cuda:{self.handle.device}"
        if nleft * nright == 0:
            return torch.zeros((0,) * (M + 2), dtype=torch.float64, device=dev), maxabs
        nl, nr = lt.shape[1], rt.shape[1]
        if len(self.localdims) - nl - nr != M:
            raise TCIError(_lib.TCI_ERR_INVALID, "Invalid number of central indices")
        cd = [int(d) for d in self.localdims[nl:nl + M]]
        ncent = int(np.prod(cd)) if cd else 1
        cents = np.asarray([tuple(reversed(c)) for c in itertools.product(*[range(1, d + 1) for d in reversed(cd)])],
                           dtype=np.int64).reshape(ncent, M)            # first centre index fastest (:274)
        vals = np.empty((nright, ncent, nleft))
        if self._np_coeffs is not None:                               # vectorised keys
            co = self._np_coeffs
            kl = (lt.astype(np.int64) - 1) @ co[:nl]
            kc = (cents - 1) @ co[nl:nl + M] if M else np.zeros(1, dtype=np.int64)
            kr = (rt.astype(np.int64) - 1) @ co[nl + M:]
            keys = kr[:, None, None] + kc[None, :, None] + kl[None, None, :]
            flat = keys.reshape(-1)
            out_flat = vals.reshape(-1)
            cache = self.cache
            missing = []
            for p, k in enumerate(flat.tolist()):
                v = cache.get(k)
                if v is None:
                    missing.append(p)
                else:
                    out_flat[p] = v
            for p in missing:
                k = int(flat[p])
                if k in cache:                                        # duplicate index rows inside one batch
                    out_flat[p] = cache[k]
                    continue
                j, rem = divmod(p, ncent * nleft)
                c, i = divmod(rem, nleft)
                v = self._call_f([int(x) for x in lt[i]] + [int(x) for x in cents[c]] + [int(x) for x in rt[j]])
                cache[k] = v
                out_flat[p] = v
        else:
            for j in range(nright):
                for c in range(ncent):
                    for i in range(nleft):
                        vals[j, c, i] = self([int(x) for x in lt[i]] + [int(x) for x in cents[c]] + [int(x) for x in rt[j]])
        sc = stride_c if stride_c > 0 else nleft
        sj = stride_j if stride_j > 0 else sc * ncent
        if out is None:
            out = torch.empty((nright,) + tuple(reversed(cd)) + (nleft,), dtype=torch.float64, device=dev)
        view = torch.as_strided(out.reshape(-1), (nright, ncent, nleft), (sj, sc, 1))
        view.copy_(torch.from_numpy(vals))
        if maxabs is not None:
            m = float(np.max(np.abs(vals))) if not np.isnan(vals).any() else float("nan")
            a = abs(maxabs)
            maxabs = float("nan") if (np.isnan(a) or np.isnan(m)) else max(a, m)
        return out, maxabs
