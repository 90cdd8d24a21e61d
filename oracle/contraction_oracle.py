"""CPU restatement (numpy) of the reference's MPO contraction code — src/contraction.jl and the LU/CI branches of
src/tensortrain.jl:219-348.

TEST INFRASTRUCTURE ONLY (imported by tests/ and the CPU legs of benchmarks/); the product never imports it.
Parity status: pinned by the reference's own test identities (test/test_contraction.jl:25-30 `_contract` vs explicit
permute+matmul, :73-106 MPO x MPO == dense matrix product for :TCI / :naive, :108-153 projector slices, :155-189
MPO x MPS, :192-202 zip-up), which tests/test_oracle_contraction.py re-runs on seeded inputs.  The reference's tests use
random ComplexF64 tensors and compare against dense products, so there are no literal golden vectors for this part.

Tensors are numpy arrays indexed exactly like the Julia arrays (shape (chi_l, legs..., chi_r)); index VALUES are 1-based."""
import numpy as np

from . import oracle as O


def _contract(a, b, idx_a, idx_b):
    """contraction.jl:193-215: permutedims -> reshape -> one matrix product -> reshape (Julia column-major reshapes)."""
    ia, ib = [i - 1 for i in idx_a], [i - 1 for i in idx_b]
    assert len(ia) == len(ib) and len(set(ia)) == len(ia) and len(set(ib)) == len(ib)
    ra = [d for d in range(a.ndim) if d not in ia]
    rb = [d for d in range(b.ndim) if d not in ib]
    sa, sb = a.shape, b.shape
    amat = np.transpose(a, ra + ia).reshape((int(np.prod([sa[d] for d in ra])), int(np.prod([sa[d] for d in ia]))), order="F")
    bmat = np.transpose(b, ib + rb).reshape((int(np.prod([sb[d] for d in ib])), int(np.prod([sb[d] for d in rb]))), order="F")
    return (amat @ bmat).reshape(tuple(sa[d] for d in ra) + tuple(sb[d] for d in rb), order="F")


class Contraction:
    """contraction.jl:60-152, evaluate :365-406, batchevaluate :483-575 (dict-memoised environments, :279-354)."""

    def __init__(self, a, b, f=None):
        if len(a) != len(b):
            raise ValueError("Tensor trains must have the same length.")
        for n in range(len(a)):
            if a[n].shape[2] != b[n].shape[1]:
                raise RuntimeError(f"Tensor trains must share the identical index at n={n + 1}!")
        self.a, self.b, self.f = a, b, f
        self.sitedims = [[x.shape[1], y.shape[2]] for x, y in zip(a, b)]
        self.localdims = [s[0] * s[1] for s in self.sitedims]
        self.leftcache, self.rightcache = {}, {}

    def __len__(self):
        return len(self.a)

    def _unfuse(self, n0, idx):                                          # :222-224
        q, r = divmod(idx - 1, self.sitedims[n0][0])
        return (r + 1, q + 1)

    @staticmethod
    def _extend_cache(old, a_ell, b_ell, i, j):                          # :253-259
        tmp1 = _contract(old, a_ell[:, i - 1, :, :], (1,), (1,))
        return _contract(tmp1, b_ell[:, :, j - 1, :], (1, 2), (1, 2))

    def evaluateleft(self, indexset):                                    # :279-310
        indexset = tuple(indexset)
        if len(indexset) >= len(self):
            raise RuntimeError("Invalid indexset")
        if len(indexset) == 0:
            return np.ones((1, 1))
        ell = len(indexset)
        if ell == 1:
            i, j = indexset[0]
            return self.a[0][0, i - 1, :, :].T @ self.b[0][0, :, j - 1, :]
        if indexset not in self.leftcache:
            i, j = indexset[-1]
            self.leftcache[indexset] = self._extend_cache(self.evaluateleft(indexset[:-1]), self.a[ell - 1], self.b[ell - 1], i, j)
        return self.leftcache[indexset]

    def evaluateright(self, indexset):                                   # :323-354
        indexset = tuple(indexset)
        N = len(self)
        if len(indexset) >= N:
            raise RuntimeError("Invalid indexset")
        if len(indexset) == 0:
            return np.ones((1, 1))
        if len(indexset) == 1:
            i, j = indexset[0]
            return self.a[-1][:, i - 1, :, 0] @ self.b[-1][:, :, j - 1, 0].T
        ell = N - len(indexset)                                          # 0-based site
        if indexset not in self.rightcache:
            i, j = indexset[0]
            self.rightcache[indexset] = self._extend_cache(self.evaluateright(indexset[1:]),
                                                           np.transpose(self.a[ell], (3, 1, 2, 0)),
                                                           np.transpose(self.b[ell], (3, 1, 2, 0)), i, j)
        return self.rightcache[indexset]

    def evaluate(self, indexset):                                        # :365-406
        if len(indexset) != len(self):
            raise RuntimeError("Length mismatch")
        if np.ndim(indexset[0]) == 0:
            indexset = [self._unfuse(n, int(v)) for n, v in enumerate(indexset)]
        indexset = [tuple(int(x) for x in p) for p in indexset]
        mid = len(self) // 2
        res = float(np.sum(self.evaluateleft(indexset[:mid]) * self.evaluateright(indexset[mid:])))
        return self.f(res) if self.f is not None else res

    def __call__(self, indexset):
        return self.evaluate(indexset)

    def batchevaluate(self, leftindexset, rightindexset, M, projector=None):     # :483-575
        N = len(self)
        Nr = len(rightindexset[0])
        s_, e_ = len(leftindexset[0]), N - Nr
        if projector is None:
            projector = [[0, 0] for _ in range(s_, e_)]
        assert len(projector) == M
        lu = [[self._unfuse(n, int(v)) for n, v in enumerate(idx)] for idx in leftindexset]
        ru = [[self._unfuse(N - Nr + n, int(v)) for n, v in enumerate(idx)] for idx in rightindexset]
        la = [1] + [t.shape[3] for t in self.a]
        lb = [1] + [t.shape[3] for t in self.b]
        left_ = np.zeros((len(leftindexset), la[s_], lb[s_]))
        for i, idx in enumerate(lu):
            left_[i, :, :] = self.evaluateleft(idx)
        right_ = np.zeros((la[e_], lb[e_], len(rightindexset)))
        for i, idx in enumerate(ru):
            right_[:, :, i] = self.evaluateright(idx)
        leftobj = left_.reshape(left_.shape + (1,), order="F")
        sizes = []
        for n in range(s_, e_):
            pa, pb = projector[n - s_]
            a_n = self.a[n] if pa == 0 else self.a[n][:, pa - 1:pa, :, :]
            b_n = self.b[n] if pb == 0 else self.b[n][:, :, pb - 1:pb, :]
            sizes.append(a_n.shape[1] * b_n.shape[2])
            tmp1 = _contract(leftobj, a_n, (2,), (1,))
            tmp2 = _contract(tmp1, b_n, (2, 5), (1, 2))
            tmp3 = np.transpose(tmp2, (0, 3, 5, 1, 2, 4))
            leftobj = tmp3.reshape(tmp3.shape[:3] + (-1,), order="F")
        res = _contract(leftobj, right_, (2, 3), (1, 2))
        if self.f is not None:
            res = np.vectorize(self.f)(res)
        return res.reshape((len(leftindexset),) + tuple(sizes) + (len(rightindexset),), order="F")


def _factorize(A, method, tolerance, maxbonddim, leftorthogonal=False, normalizeerror=True):     # tensortrain.jl:219-239
    reltol, abstol = 1e-14, 0.0
    if normalizeerror:
        reltol = tolerance
    else:
        abstol = tolerance
    mr = O.INT64_MAX if maxbonddim is None else maxbonddim
    if method in ("LU", "CI"):
        lu = O.rrlu(np.asfortranarray(A), maxrank=mr, reltol=reltol, abstol=abstol, leftorthogonal=leftorthogonal)
        if method == "LU":
            return lu.left(), lu.right(), lu.npivot
        left, right = lu.luci_factors()
        return left, right, lu.npivot
    if method == "SVD":                                                  # tensortrain.jl:244-262 (LinearAlgebra.svd = LAPACK gesdd)
        U, S, Vt = np.linalg.svd(np.asarray(A, dtype=np.float64), full_matrices=False)
        err = np.array([np.sum(S[n + 1:] ** 2) for n in range(len(S))])
        with np.errstate(divide="ignore", invalid="ignore"):
            nerr = err / np.sum(S ** 2)

        def first(mask):                                                 # replacenothing(findfirst(...), length(err))
            w = np.nonzero(mask)[0]
            return int(w[0]) + 1 if len(w) else len(S)
        trunci = min(first(err < abstol ** 2), first(nerr < reltol ** 2), mr)
        if leftorthogonal:
            return U[:, :trunci], np.diag(S[:trunci]) @ Vt[:trunci, :], trunci
        return U[:, :trunci] @ np.diag(S[:trunci]), Vt[:trunci, :], trunci
    raise NotImplementedError(method)


def _addtttensor(A, B, factorA=1.0, factorB=1.0, lefttensor=False, righttensor=False):   # abstracttensortrain.jl:467-485
    if A.ndim != B.ndim:
        raise ValueError("Elementwise addition only works if both tensors have the same indices")
    off1 = 0 if lefttensor else A.shape[0]
    off3 = 0 if righttensor else A.shape[-1]
    C = np.zeros((off1 + B.shape[0],) + A.shape[1:-1] + (off3 + B.shape[-1],))
    C[: A.shape[0], ..., : A.shape[-1]] = factorA * A
    C[off1:, ..., off3:] = factorB * B
    return C


def add(lhs, rhs, factorlhs=1.0, factorrhs=1.0, tolerance=0.0, maxbonddim=None):          # abstracttensortrain.jl:524-545
    if len(lhs) != len(rhs):
        raise ValueError("Two tensor trains with different length cannot be added elementwise.")
    L = len(lhs)
    tt = [_addtttensor(lhs[e], rhs[e], factorA=factorlhs if e == L - 1 else 1.0, factorB=factorrhs if e == L - 1 else 1.0,
                       lefttensor=e == 0, righttensor=e == L - 1) for e in range(L)]
    return compress(tt, "SVD", tolerance=tolerance, maxbonddim=maxbonddim)


def subtract(lhs, rhs, tolerance=0.0, maxbonddim=None):                                   # :560-565
    return add(lhs, rhs, factorrhs=-1.0, tolerance=tolerance, maxbonddim=maxbonddim)


def norm2(tt):                                                                            # :625-634, literally
    def _f(t):
        t3 = t.reshape((t.shape[0], -1, t.shape[-1]), order="F")
        tct = np.transpose(_contract(t3, t3, (2,), (2,)), (0, 2, 1, 3))
        return tct.reshape((tct.shape[0] * tct.shape[1], tct.shape[2] * tct.shape[3]), order="F")
    acc = _f(tt[0])
    for t in tt[1:]:
        acc = acc @ _f(t)
    assert acc.shape == (1, 1)
    return float(acc[0, 0])


def reverse(tt):                                                                          # tensortrain.jl:452-456
    return [np.transpose(T, (T.ndim - 1,) + tuple(range(1, T.ndim - 1)) + (0,)) for T in reversed(tt)]


def fulltensor(tt):
    res = tt[0]
    for t in tt[1:]:
        res = np.tensordot(res, t, axes=([res.ndim - 1], [0]))
    return res.reshape(res.shape[1:-1])


def compress(cores, method="LU", tolerance=1e-12, maxbonddim=None, normalizeerror=True):          # tensortrain.jl:302-348
    cores = [np.array(c, order="F") for c in cores]
    n = len(cores)
    for ell in range(n - 1):
        shl = cores[ell].shape
        left, right, r = _factorize(cores[ell].reshape((-1, shl[-1]), order="F"), method, 0.0, None, leftorthogonal=True)
        cores[ell] = left.reshape(shl[:-1] + (r,), order="F")
        shr = cores[ell + 1].shape
        cores[ell + 1] = (right @ cores[ell + 1].reshape((shr[0], -1), order="F")).reshape((r,) + shr[1:], order="F")
    for ell in range(n - 1, 0, -1):
        shr = cores[ell].shape
        left, right, r = _factorize(cores[ell].reshape((shr[0], -1), order="F"), method, tolerance, maxbonddim,
                                    leftorthogonal=False, normalizeerror=normalizeerror)
        cores[ell] = right.reshape((r,) + shr[1:], order="F")
        shl = cores[ell - 1].shape
        cores[ell - 1] = (cores[ell - 1].reshape((-1, shl[-1]), order="F") @ left).reshape(shl[:-1] + (r,), order="F")
    return cores


def _contractsitetensors(a, b):                                          # contraction.jl:591-602
    ab = _contract(a, b, (3,), (2,))
    abp = np.transpose(ab, (0, 3, 1, 4, 2, 5))
    return abp.reshape((a.shape[0] * b.shape[0], a.shape[1], b.shape[2], a.shape[3] * b.shape[3]), order="F")


def contract_naive(a, b, tolerance=0.0, maxbonddim=None, method="SVD"):  # contraction.jl:616-637 (the reference compresses with :SVD)
    tt = [_contractsitetensors(x, y) for x, y in zip(a, b)]
    if tolerance > 0 or maxbonddim is not None:
        tt = compress(tt, method, tolerance=tolerance, maxbonddim=maxbonddim)
    return tt


def contract_zipup(A, B, tolerance=1e-12, method="SVD", maxbonddim=None):  # contraction.jl:751-788
    R = np.ones((1, 1, 1))
    out = []
    N = len(A)
    for n in range(N):
        RA = _contract(R, A[n], (2,), (1,))
        Cn = np.transpose(_contract(RA, B[n], (2, 4), (1, 2)), (0, 1, 3, 2, 4))
        shp = Cn.shape
        if n == N - 1:
            out.append(Cn.reshape(shp[:3] + (1,), order="F"))
            break
        left, right, r = _factorize(Cn.reshape((shp[0] * shp[1] * shp[2], shp[3] * shp[4]), order="F"), method, tolerance, maxbonddim)
        out.append(left.reshape(shp[:3] + (r,), order="F"))
        R = right.reshape((r,) + shp[3:5], order="F")
    return out


def tomat(cores):
    """_tomat of test/test_contraction.jl:5-17: dense (prod d1) x (prod d2) matrix of an MPO, first site fastest."""
    res = cores[0]                                                       # (1, s1, s2, chi)
    res = res.reshape(res.shape[1:], order="F")                          # (s1, s2, chi)
    rows, cols = res.shape[0], res.shape[1]
    acc = np.transpose(res, (0, 1, 2)).reshape((rows, cols, -1), order="F")
    for c in cores[1:]:
        t = np.tensordot(acc, c, axes=([2], [0]))                        # (R, C, s1, s2, chi')
        t = np.transpose(t, (0, 2, 1, 3, 4))                             # (R, s1, C, s2, chi')
        acc = t.reshape((t.shape[0] * t.shape[1], t.shape[2] * t.shape[3], t.shape[4]), order="F")
    return acc[:, :, 0]
