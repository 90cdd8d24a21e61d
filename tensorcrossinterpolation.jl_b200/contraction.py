"""Host-side mirror of src/contraction.jl: `Contraction` (MPO x MPO as a lazily evaluated BatchEvaluator with memoised
left/right environments), `contract_naive`, `contract_zipup`, `contract_TCI`, `contract`.

As in the reference the environments are memoised in host dictionaries keyed by the index prefix/suffix; the matrices
themselves live on the device and every product is a DMMA GEMM on strided views (tensor._contract)."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import TCIError, lib
from .batcheval import BatchEvaluator, as_table
from .dense import Mat
from .tensor import DTensor, _contract
from .tensortrain import DeviceCore, TensorTrain, _factorize


class ElementwiseMap:
    """The optional `f` of Contraction (contraction.jl:64): applied to every contracted element on the device
    (tci_elementwise_f64).  kinds: 'affine' a*x+b, 'abs2', 'abs', 'exp' exp(a*x)."""
    KINDS = {"affine": 1, "abs2": 2, "abs": 3, "exp": 4}

    def __init__(self, kind, p0=1.0, p1=0.0):
        if kind not in self.KINDS:
            raise TCIError(_lib.TCI_ERR_INVALID, f"unknown elementwise map {kind}")
        self.kind, self.p0, self.p1 = kind, float(p0), float(p1)

    def __call__(self, x):
        if self.kind == "affine":
            return self.p0 * x + self.p1
        if self.kind == "abs2":
            return x * x
        if self.kind == "abs":
            return abs(x)
        return float(np.exp(self.p0 * x))

    def apply(self, h, t):
        h.check(lib().tci_elementwise_f64(h.h, self.KINDS[self.kind], C.c_double(self.p0), C.c_double(self.p1), t.ptr,
                                          C.c_int64(t.size)))


def _core_tensor(core):
    """DeviceCore -> DTensor view (chi_l, legs..., chi_r) (no copy; honours a padded leading dimension)."""
    shape = (core.chil,) + core.sitedims + (core.chir,)
    st, acc = [], 1
    for s in shape[:-1]:
        st.append(acc)
        acc *= s
    st.append(core.mat.ld)
    return DTensor(core.mat.buf.reshape(-1), shape, st, 0)


def _tensor_core(t, handle):
    """dense DTensor (chi_l, legs..., chi_r) -> DeviceCore."""
    t = t.contiguous(handle)
    chil, chir = t.shape[0], t.shape[-1]
    rows = t.size // max(chir, 1)
    buf = t.buf[t.offset:t.offset + t.size].reshape(chir, rows)
    return DeviceCore(Mat(buf, rows, chir), chil, t.shape[1:-1], chir)


class Contraction(BatchEvaluator):
    """struct Contraction{Float64} <: BatchEvaluator (contraction.jl:60-152)."""

    def __init__(self, a, b, f=None, handle=None):
        if len(a) != len(b):
            raise TCIError(_lib.TCI_ERR_INVALID, "Tensor trains must have the same length.")
        self.a = [_core_tensor(c) for c in a._dev_cores()]
        self.b = [_core_tensor(c) for c in b._dev_cores()]
        for n, (x, y) in enumerate(zip(self.a, self.b)):
            if x.ndim != 4 or y.ndim != 4:
                raise TCIError(_lib.TCI_ERR_INVALID, "Contraction needs tensor trains with two site legs (MPOs)")
            if x.shape[2] != y.shape[1]:
                raise TCIError(_lib.TCI_ERR_INVALID, f"Tensor trains must share the identical index at n={n + 1}!")
        self.sitedims = [[x.shape[1], y.shape[2]] for x, y in zip(self.a, self.b)]
        super().__init__([s[0] * s[1] for s in self.sitedims], handle or a.handle)
        self.f = f
        self.leftcache, self.rightcache = {}, {}
        self.dev = f"cuda:{self.handle.device}"
        self._one = DTensor.ones((1, 1), self.dev)
        # right environments are built with the bond legs exchanged (contraction.jl:343-344): views, no copies
        self.ar = [t.permute((3, 1, 2, 0)) for t in self.a]
        self.br = [t.permute((3, 1, 2, 0)) for t in self.b]

    def __len__(self):
        return len(self.a)

    # ---- index helpers (contraction.jl:222-234)
    def _unfuse(self, n0, idx):
        d1 = self.sitedims[n0][0]
        q, r = divmod(int(idx) - 1, d1)
        return (r + 1, q + 1)

    def _fuse(self, n0, ij):
        return ij[0] + self.sitedims[n0][0] * (ij[1] - 1)

    # ---- environments
    def _extend_cache(self, old, a_ell, b_ell, i, j):                   # contraction.jl:253-259
        tmp1 = _contract(old, a_ell.select(1, i - 1), (1,), (1,), self.handle)
        return _contract(tmp1, b_ell.select(2, j - 1), (1, 2), (1, 2), self.handle)

    def evaluateleft(self, indexset):                                   # contraction.jl:279-310
        indexset = tuple((int(i), int(j)) for i, j in indexset)
        if len(indexset) >= len(self):
            raise TCIError(_lib.TCI_ERR_INVALID, f"Invalid indexset: {list(indexset)}")
        if len(indexset) == 0:
            return self._one
        # iterative form of the reference's recursion: extend from the longest cached prefix
        start = len(indexset)
        while start > 1 and indexset[:start] not in self.leftcache:
            start -= 1
        if start == 1 and indexset[:1] not in self.leftcache:
            i, j = indexset[0]
            a1 = self.a[0].select(0, 0).select(0, i - 1)                # (s2, chi_a')
            b1 = self.b[0].select(0, 0).select(1, j - 1)                # (s2, chi_b')
            self.leftcache[indexset[:1]] = _contract(a1, b1, (1,), (1,), self.handle)   # transpose(a) * b
        env = self.leftcache[indexset[:start]]
        for ell in range(start, len(indexset)):
            i, j = indexset[ell]
            env = self._extend_cache(env, self.a[ell], self.b[ell], i, j)
            self.leftcache[indexset[:ell + 1]] = env
        return env

    def evaluateright(self, indexset):                                  # contraction.jl:323-354
        indexset = tuple((int(i), int(j)) for i, j in indexset)
        N = len(self)
        if len(indexset) >= N:
            raise TCIError(_lib.TCI_ERR_INVALID, f"Invalid indexset: {list(indexset)}")
        if len(indexset) == 0:
            return self._one
        L = len(indexset)
        start = L                                                       # length of the longest cached suffix
        while start > 1 and indexset[L - start:] not in self.rightcache:
            start -= 1
        if start == 1 and indexset[L - 1:] not in self.rightcache:
            i, j = indexset[-1]
            aN = self.a[-1].select(3, 0).select(1, i - 1)               # (chi_a, s2)
            bN = self.b[-1].select(3, 0).select(2, j - 1)               # (chi_b, s2)
            self.rightcache[indexset[L - 1:]] = _contract(aN, bN, (2,), (2,), self.handle)   # a * transpose(b)
        env = self.rightcache[indexset[L - start:]]
        for cnt in range(start + 1, L + 1):                             # suffix of length cnt
            ell = N - cnt                                               # 0-based site of its first index
            i, j = indexset[L - cnt]
            env = self._extend_cache(env, self.ar[ell], self.br[ell], i, j)
            self.rightcache[indexset[L - cnt:]] = env
        return env

    # ---- batched environments: what the per-row recursion above computes, for a whole index set at once
    def _fill_cache(self, rows, left):
        """Memoise the environment of every row (tuples of (i, j) pairs of equal length) level by level: the prefixes (left) /
        suffixes (right) of one length that are still missing are grouped by their new index pair, the parents' matrices of a
        group are stacked, and the group is extended by TWO contractions (`_extend_cache` for all its rows at once) instead of
        two per row.  A TCI sweep asks for |Iset| = chi rows per bond whose parents were the previous bond's rows, so this is
        2 d^2 GEMMs per bond instead of 2 chi (contraction.jl:279-354 is a recursion per index set; same values)."""
        import torch
        if not rows or len(rows[0]) == 0:
            return
        cache = self.leftcache if left else self.rightcache
        L, N, h = len(rows[0]), len(self), self.handle
        first = 1                                                       # the deepest level that is already complete (in a sweep: L - 1)
        for lev in range(L, 0, -1):
            if all((r[:lev] if left else r[L - lev:]) in cache for r in rows):
                first = lev + 1
                break
        for lev in range(first, L + 1):
            missing = [k for k in {(r[:lev] if left else r[L - lev:]) for r in rows} if k not in cache]
            if not missing:
                continue
            if lev == 1:
                for k in missing:
                    (self.evaluateleft if left else self.evaluateright)(k)
                continue
            ell = lev - 1 if left else N - lev                          # 0-based site of the index that is added
            a_t, b_t = (self.a[ell], self.b[ell]) if left else (self.ar[ell], self.br[ell])
            groups = {}
            for k in missing:
                groups.setdefault(k[-1] if left else k[0], []).append(k)
            for (i, j), ks in groups.items():
                parents = [cache[k[:-1] if left else k[1:]] for k in ks]
                ca, cb = parents[0].shape
                flat = torch.stack([q.buf[q.offset:q.offset + ca * cb] for q in parents]).reshape(-1)     # (ca, cb, n), n slowest
                stack = DTensor(flat, (ca, cb, len(ks)))
                tmp = _contract(stack, a_t.select(1, i - 1), (1,), (1,), h)                  # (cb, n, s2, ca')
                new = _contract(tmp, b_t.select(2, j - 1), (1, 3), (1, 2), h)                # (n, ca', cb')
                newp = new.permute((1, 2, 0)).contiguous(h)                                  # (ca', cb', n): one dense matrix per row
                for q, k in enumerate(ks):
                    cache[k] = newp.select(2, q)

    # ---- point evaluation (contraction.jl:365-406)
    def evaluate(self, indexset):
        if len(indexset) != len(self):
            raise TCIError(_lib.TCI_ERR_INVALID, f"Length mismatch: {len(self)} != {len(indexset)}")
        if np.ndim(indexset[0]) == 0:
            indexset = [self._unfuse(n, v) for n, v in enumerate(indexset)]
        mid = len(self) // 2
        left = self.evaluateleft(indexset[:mid])
        right = self.evaluateright(indexset[mid:])
        n = left.size
        res = _contract(left.reshape((1, n), self.handle), right.reshape((n, 1), self.handle), (2,), (1,), self.handle)   # sum(left .* right)
        val = float(res.numpy()[0, 0])
        return self.f(val) if self.f is not None else val

    def __call__(self, *args):
        if len(args) == 1:
            idx = args[0]
            if np.ndim(idx[0]) != 0:                                    # vector of [i, j] pairs -> fused (contraction.jl:455)
                idx = [self._fuse(n, p) for n, p in enumerate(idx)]
            return self.evaluate(list(idx))
        left, right, M = args
        return self.batchevaluate(left, right, M).numpy(self.handle)

    # ---- batch evaluation (contraction.jl:483-575)
    def batchevaluate(self, leftindexset, rightindexset, M, projector=None):
        h = self.handle
        N = len(self)
        lt, rt = as_table(leftindexset), as_table(rightindexset)
        nI, nJ = lt.shape[0], rt.shape[0]
        Nl, Nr = lt.shape[1], rt.shape[1]
        s_, e_ = Nl, N - Nr                                             # centre sites s_ .. e_-1 (0-based)
        if e_ - s_ != M:
            raise TCIError(_lib.TCI_ERR_INVALID, "Invalid number of central indices")
        if projector is None:
            projector = [[0, 0] for _ in range(s_, e_)]
        if len(projector) != M:
            raise TCIError(_lib.TCI_ERR_INVALID, f"Length mismatch: length of projector (={len(projector)}) must be {M}")
        for n in range(s_, e_):
            p = projector[n - s_]
            if len(p) != 2:
                raise TCIError(_lib.TCI_ERR_INVALID, f"Invalid projector at {n + 1}: {list(p)}, the length must be 2")
            if not all(0 <= int(x) <= d for x, d in zip(p, self.sitedims[n])):
                raise TCIError(_lib.TCI_ERR_INVALID, f"Invalid projector: {list(p)}")
        chi_a = [1] + [t.shape[3] for t in self.a]                      # linkdims with the boundary 1s
        chi_b = [1] + [t.shape[3] for t in self.b]
        ca, cb = chi_a[s_], chi_b[s_]
        # environment stacks: each memoised matrix is copied into one slot of (chi_a, chi_b, count); the left stack is then
        # VIEWED as (count, chi_a, chi_b), which is what the reference materialises (contraction.jl:516-521)
        import torch
        lrows = [tuple(self._unfuse(n, v) for n, v in enumerate(row)) for row in lt]
        rrows = [tuple(self._unfuse(N - Nr + n, v) for n, v in enumerate(row)) for row in rt]
        self._fill_cache(lrows, True)
        self._fill_cache(rrows, False)
        ra, rb = chi_a[e_], chi_b[e_]

        def stacked(rows, getter, r0, r1):                              # (r0, r1, count): one memoised matrix per slot
            envs = [getter(r).contiguous(h) for r in rows]
            return DTensor(torch.stack([e.buf[e.offset:e.offset + r0 * r1] for e in envs]).reshape(-1), (r0, r1, len(rows)))
        lstack = stacked(lrows, self.evaluateleft, ca, cb)
        rstack = stacked(rrows, self.evaluateright, ra, rb)
        leftobj = DTensor(lstack.buf, (nI, ca, cb, 1), (ca * cb, 1, ca, 1))
        sizes = []
        for n in range(s_, e_):
            pa, pb = int(projector[n - s_][0]), int(projector[n - s_][1])
            a_n = self.a[n] if pa == 0 else self.a[n].keep(1, pa - 1)
            b_n = self.b[n] if pb == 0 else self.b[n].keep(2, pb - 1)
            sizes.append(a_n.shape[1] * b_n.shape[2])
            tmp1 = _contract(leftobj, a_n, (2,), (1,), h)               # (I, chi_b, S, s1, s2, chi_a')
            tmp2 = _contract(tmp1, b_n, (2, 5), (1, 2), h)              # (I, S, s1, chi_a', s3, chi_b')
            tmp3 = tmp2.permute((0, 3, 5, 1, 2, 4))                     # (I, chi_a', chi_b', S, s1, s3)
            leftobj = tmp3.reshape(tmp3.shape[:3] + (tmp3.shape[3] * tmp3.shape[4] * tmp3.shape[5],), h)
        res = _contract(leftobj, rstack, (2, 3), (1, 2), h)             # (I, S, J)
        if self.f is not None:
            self.f.apply(h, res)
        return res.reshape((nI,) + tuple(sizes) + (nJ,), h)

    def evaluate_points(self, pts):
        """f at full-length multi-indices.  Points that differ in ONE site only (the scan of optfirstpivot, util.jl:260-291: all
        d candidates of a site with the others fixed) are one batch with a single centre site — one left environment, one right
        environment, one contraction — instead of d point evaluations with a device round trip each."""
        pts = np.asarray(pts)
        if pts.ndim == 2 and pts.shape[0] > 1:
            varying = np.nonzero((pts != pts[0]).any(axis=0))[0]
            if varying.size == 1:
                c = int(varying[0])
                vals = self.batchevaluate(pts[:1, :c], pts[:1, c + 1:], 1).numpy(self.handle).reshape(-1)
                return vals[pts[:, c].astype(np.int64) - 1]
        return np.array([self.evaluate([int(v) for v in p]) for p in pts])

    # ---- the K1-compatible interface TensorCI2 uses (batcheval.BatchEvaluator.evaluate_batch)
    def evaluate_batch(self, left, right, M, out=None, stride_c=0, stride_j=0, maxabs=None):
        import torch
        h = self.handle
        lt, rt = as_table(left), as_table(right)
        nI, nJ = lt.shape[0], rt.shape[0]
        if nI * nJ == 0:
            return torch.zeros((0,) * (M + 2), dtype=torch.float64, device=self.dev), maxabs
        res = self.batchevaluate(lt, rt, M)                             # dense (I, c..., J)
        ncent = res.size // (nI * nJ)
        flat = res.buf[res.offset:res.offset + res.size]
        if out is None:
            cd = res.shape[1:-1]
            out = flat.reshape((nJ,) + tuple(reversed(cd)) + (nI,))
        else:
            sc = stride_c if stride_c > 0 else nI
            sj = stride_j if stride_j > 0 else sc * ncent
            view = torch.as_strided(out.reshape(-1), (nJ, ncent, nI), (sj, sc, 1))
            view.copy_(flat.reshape(nJ, ncent, nI))
        if maxabs is not None:
            mx = C.c_double(float(maxabs))
            h.check(lib().tci_maxabs_f64(h.h, res.ptr, C.c_int64(res.size), C.c_int64(1), C.c_int64(res.size), C.byref(mx)))
            maxabs = mx.value
        return out, maxabs


# ---- algorithms -----------------------------------------------------------------------------------------------------
def _contractsitetensors(a, b, h):                                      # contraction.jl:591-602
    ab = _contract(a, b, (3,), (2,), h)                                 # (la, s1, la', lb, s3, lb')
    abp = ab.permute((0, 3, 1, 4, 2, 5))
    return abp.reshape((a.shape[0] * b.shape[0], a.shape[1], b.shape[2], a.shape[3] * b.shape[3]), h)


def contract_naive(a, b=None, tolerance=0.0, maxbonddim=None, method="SVD"):
    """contract_naive (contraction.jl:616-637): site-wise product, then compress!(tt, :SVD; tolerance, maxbonddim) when a
    tolerance or maxbonddim is given (`method` 'LU'/'CI' selects the rrLU compression instead; not a reference option)."""
    obj = a if isinstance(a, Contraction) else Contraction(a, b)
    if obj.f is not None:
        raise TCIError(_lib.TCI_ERR_INVALID, "Naive contraction implementation cannot contract matrix product with a function. "
                                             "Use algorithm=:TCI instead.")
    h = obj.handle
    tt = TensorTrain([_tensor_core(_contractsitetensors(x, y, h), h) for x, y in zip(obj.a, obj.b)], handle=h)
    if tolerance > 0 or maxbonddim is not None:
        tt.compress(method, tolerance=tolerance, maxbonddim=maxbonddim)
    return tt


def contract_zipup(A, B, tolerance=1e-12, method="SVD", maxbonddim=None):
    """contract_zipup (contraction.jl:751-788); method 'SVD' (default, like the reference) | 'LU' | 'CI'."""
    if len(A) != len(B):
        raise TCIError(_lib.TCI_ERR_INVALID, "Cannot contract tensor trains with different length.")
    h = A.handle
    dev = f"cuda:{h.device}"
    a = [_core_tensor(c) for c in A._dev_cores()]
    b = [_core_tensor(c) for c in B._dev_cores()]
    R = DTensor.ones((1, 1, 1), dev)
    cores = []
    N = len(a)
    for n in range(N):
        RA = _contract(R, a[n], (2,), (1,), h)                          # (lc, lb, s1, s2, la')
        Cn = _contract(RA, b[n], (2, 4), (1, 2), h).permute((0, 1, 3, 2, 4))   # (lc, s1, s3, la', lb')
        shp = Cn.shape
        if n == N - 1:
            cores.append(_tensor_core(Cn.reshape(shp[:3] + (1,), h), h))
            break
        Cm = Cn.reshape((shp[0] * shp[1] * shp[2], shp[3] * shp[4]), h)
        Am = Mat(Cm.buf[Cm.offset:Cm.offset + Cm.size].reshape(Cm.shape[1], Cm.shape[0]), Cm.shape[0], Cm.shape[1])
        left, right, r = _factorize(Am, method, tolerance, maxbonddim, handle=h)
        cores.append(DeviceCore(left, shp[0], shp[1:3], r))
        rd = right if right.ld == right.rows else None
        if rd is None:
            rd = Mat.empty(right.rows, right.cols, right.buf.device)
            rd.buf.copy_(right.buf[: right.cols, : right.rows])
        R = DTensor(rd.buf.reshape(-1), (r, shp[3], shp[4]))
    return TensorTrain(cores, handle=h)


def _findinitialpivots(f, localdims, nmaxpivots, rng=None):             # contraction.jl:666-677 (numpy RNG, not Xoshiro)
    from .tensorci2 import optfirstpivot
    rng = rng or np.random.default_rng()
    pivots = []
    for _ in range(nmaxpivots):
        p = [int(rng.integers(1, d + 1)) for d in localdims]
        p = optfirstpivot(f, localdims, p)
        if abs(f(p)) == 0.0:
            continue
        pivots.append(p)
    return pivots


def contract_TCI(A, B, initialpivots=10, f=None, rng=None, **kwargs):
    """contract_TCI (contraction.jl:692-732)."""
    from .tensorci2 import crossinterpolate2
    if len(A) != len(B):
        raise TCIError(_lib.TCI_ERR_INVALID, "Cannot contract tensor trains with different length.")
    if not all(sa[1] == sb[0] for sa, sb in zip(A.sitedims(), B.sitedims())):
        raise TCIError(_lib.TCI_ERR_INVALID, "Cannot contract tensor trains with non-matching site dimensions.")
    mp = Contraction(A, B, f=f)
    localdims = [int(d) for d in mp.localdims]
    if isinstance(initialpivots, int):
        initialpivots = _findinitialpivots(mp, localdims, initialpivots, rng)
        if not initialpivots:
            raise TCIError(_lib.TCI_ERR_INVALID, "No initial pivots found.")
    tci, ranks, errors = crossinterpolate2(mp, localdims, initialpivots, **kwargs)
    cores = []
    for c, sd in zip(tci._dev_cores(), mp.sitedims):
        cores.append(DeviceCore(c.mat, c.chil, tuple(sd), c.chir))      # _reshape_splitsites: same memory, two legs
    return TensorTrain(cores, handle=mp.handle)


def _as_mpo(tt, which):
    """TensorTrain{3} -> TensorTrain{4} with a unit leg (contraction.jl:875, :889)."""
    cores = []
    for c in tt._dev_cores():
        sd = (1,) + c.sitedims if which == "left" else c.sitedims + (1,)
        cores.append(DeviceCore(c.mat, c.chil, sd, c.chir))
    return TensorTrain(cores, handle=tt.handle)


def contract(A, B, algorithm="TCI", tolerance=1e-12, maxbonddim=None, f=None, **kwargs):
    """contract (contraction.jl:832-891), MPO x MPO, MPS x MPO and MPO x MPS."""
    na = len(A._dev_cores()[0].sitedims)
    nb = len(B._dev_cores()[0].sitedims)
    if na == 1 and nb == 2:
        tt = contract(_as_mpo(A, "left"), B, algorithm, tolerance, maxbonddim, f, **kwargs)
        return TensorTrain([DeviceCore(c.mat, c.chil, (c.d,), c.chir) for c in tt._dev_cores()], handle=tt.handle)
    if na == 2 and nb == 1:
        tt = contract(A, _as_mpo(B, "right"), algorithm, tolerance, maxbonddim, f, **kwargs)
        return TensorTrain([DeviceCore(c.mat, c.chil, (c.d,), c.chir) for c in tt._dev_cores()], handle=tt.handle)
    if algorithm == "TCI":
        kw = dict(kwargs)
        if maxbonddim is not None:
            kw["maxbonddim"] = maxbonddim
        return contract_TCI(A, B, tolerance=tolerance, f=f, **kw)
    if algorithm == "naive":
        if f is not None:
            raise TCIError(_lib.TCI_ERR_INVALID, "Naive contraction implementation cannot contract matrix product with a function. "
                                                 "Use algorithm=:TCI instead.")
        return contract_naive(A, B, tolerance=tolerance, maxbonddim=maxbonddim, **kwargs)
    if algorithm == "zipup":
        if f is not None:
            raise TCIError(_lib.TCI_ERR_INVALID, "Zipup contraction implementation cannot contract matrix product with a function. "
                                                 "Use algorithm=:TCI instead.")
        return contract_zipup(A, B, tolerance=tolerance, maxbonddim=maxbonddim, **kwargs)
    raise TCIError(_lib.TCI_ERR_INVALID, f"Unknown algorithm {algorithm}.")
