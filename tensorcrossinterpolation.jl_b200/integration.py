"""Host-side mirror of src/integration.jl: `integrate(f, a, b; GKorder, kwargs...)` — tensor-product Gauss–Kronrod
quadrature compressed by TCI2, then `sum`.  A thin caller of the hot path: the weighted integrand is a
device functor when the user supplies CUDA source (every evaluation then runs in the K1 kernel), or a host callable
(evaluated where the user wrote it, values uploaded) otherwise.

`kronrod(n)` restates what the reference obtains from QuadGK.jl (`QuadGK.kronrod(n, -1, +1)`, integration.jl:63; a
dependency that is not vendored in the reference tree): the 2n+1 Kronrod points are the n Gauss–Legendre points plus the
n+1 zeros of the Stieltjes polynomial E_{n+1} (orthogonal to all polynomials of degree <= n with weight P_n), weights
from exactness on P_0..P_2n."""
import numpy as np

from . import _lib
from ._lib import TCIError


def kronrod(n):
    """(x, w, wg): the 2n+1 Gauss–Kronrod nodes on [-1, 1] in ascending order, their weights, and the n embedded Gauss
    weights (zero-padded at the Kronrod-only nodes).  Checked against the QUADPACK G7K15 / G10K21 constants in the tests."""
    from numpy.polynomial import legendre as Lg
    n = int(n)
    if n < 1:
        raise TCIError(_lib.TCI_ERR_INVALID, "kronrod: n must be >= 1")
    xg, wg = Lg.leggauss(n)
    # Stieltjes polynomial E = P_{n+1} + sum_{i<=n, i = n+1 mod 2} c_i P_i with  int P_n E P_k = 0  for k = 0..n
    xq, wq = Lg.leggauss(2 * n + 4)                      # exact for degree 4n+7 >= 3n+1
    V = Lg.legvander(xq, n + 1)                          # V[:, i] = P_i(xq)
    idx = [i for i in range(n + 1) if (i - (n + 1)) % 2 == 0]
    ks = [k for k in range(n + 1) if k % 2 == 1]         # P_n * E is odd: the conditions with even k hold identically
    A = np.array([[np.sum(wq * V[:, n] * V[:, i] * V[:, k]) for i in idx] for k in ks])
    rhs = -np.array([np.sum(wq * V[:, n] * V[:, n + 1] * V[:, k]) for k in ks])
    c = np.zeros(n + 2)
    c[n + 1] = 1.0
    c[idx] = np.linalg.solve(A, rhs)
    xk = np.real(Lg.legroots(c))
    dc = Lg.legder(c)
    for _ in range(4):                                   # Newton polish of the companion-matrix roots
        xk = xk - Lg.legval(xk, c) / Lg.legval(xk, dc)
    x = np.sort(np.concatenate([xg, xk]))
    x = 0.5 * (x - x[::-1])                              # enforce the exact symmetry of the rule
    m = 2 * n + 1
    W = Lg.legvander(x, m - 1).T                         # exact for P_0 .. P_{2n}: interpolatory weights
    b = np.zeros(m)
    b[0] = 2.0
    w = np.linalg.solve(W, b)
    w = 0.5 * (w + w[::-1])
    gw = np.zeros(m)
    for xv, wv in zip(xg, wg):
        gw[int(np.argmin(np.abs(x - xv)))] = wv
    return x, w, gw


_WRAPPER = r"""
__device__ double tci_integrand(const double* x, int n, const double* params, int nparams);
__device__ double tci_user_f(const int* idx, int n, const double* p, int np_) {
    const int K = %(K)d, NU = %(NU)d;
    const double* nodes = p + NU;                  // nodes[site * K + k]
    const double* wts = p + NU + %(N)d * K;
    double x[%(N)d];
    double w = 1.0;
    for (int i = 0; i < n; ++i) { x[i] = nodes[i * K + idx[i] - 1]; w = w * wts[i * K + idx[i] - 1]; }
    return w * tci_integrand(x, n, p, NU) * p[NU + 2 * %(N)d * K];
}
"""


def integrate(f, a, b, GKorder=15, params=(), handle=None, **kwargs):
    """integrate(Float64, f, a, b; GKorder=15, kwargs...) — integration.jl:68-161.

    f: a Python callable f(x) on a list of coordinates (host function), or a string of CUDA source defining
       `__device__ double tci_integrand(const double* x, int n, const double* params, int nparams)` (device functor;
       `params` are passed through).  kwargs go to crossinterpolate2; like the reference the default is
       nsearchglobalpivot=10.  Returns the integral of f over the box [a, b]."""
    from .batcheval import UserFunction
    from .tensorci2 import crossinterpolate2
    a = np.asarray(a, dtype=np.float64).reshape(-1)
    b = np.asarray(b, dtype=np.float64).reshape(-1)
    if iseven(GKorder):
        raise TCIError(_lib.TCI_ERR_INVALID, "Gauss--Kronrod order must be odd, e.g. 15 or 61.")
    if len(a) != len(b):
        raise TCIError(_lib.TCI_ERR_INVALID, "Integral bounds must have the same dimensionality, but got "
                       f"{len(a)} lower bounds and {len(b)} upper bounds.")
    nodes1d, weights1d, _ = kronrod(GKorder // 2)        # :63
    N, K = len(a), len(nodes1d)
    nodes = (b - a)[:, None] * (nodes1d[None, :] + 1.0) / 2.0 + a[:, None]           # :66
    weights = (b - a)[:, None] * weights1d[None, :] / 2.0                            # :68
    normalization = float(GKorder) ** N                                              # :70
    localdims = [K] * N
    kwargs.setdefault("nsearchglobalpivot", 10)
    if isinstance(f, str):
        p = np.concatenate([np.asarray(params, dtype=np.float64).reshape(-1), nodes.reshape(-1), weights.reshape(-1),
                            [normalization]])
        src = f + _WRAPPER % {"K": K, "N": N, "NU": len(np.asarray(params).reshape(-1))}
        F = UserFunction(src, localdims, p, handle=handle)
    else:
        def F(indices):                                  # :86-93
            x = [float(nodes[n, i - 1]) for n, i in enumerate(indices)]
            w = 1.0
            for n, i in enumerate(indices):
                w = w * float(weights[n, i - 1])
            return w * f(x) * normalization
    tci2, ranks, errors = crossinterpolate2(F, localdims, handle=handle, **kwargs)   # :96-102
    return tci2.sum() / normalization                                                # :105


def iseven(k):
    return int(k) % 2 == 0
