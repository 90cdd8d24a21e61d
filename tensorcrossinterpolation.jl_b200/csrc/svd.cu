// svd.cu — thin SVD on the device by one-sided Jacobi (Hestenes) rotations, FP64.
//
// Replaces `LinearAlgebra.svd(A)` in _factorize(A, :SVD) (src/tensortrain.jl:244-262), the factorisation behind
// compress!(tt, :SVD) (:302-348), add/subtract (src/abstracttensortrain.jl:524-545) and the default method of
// contract_naive / contract_zipup (src/contraction.jl:616-637, :751-788).  The reference calls LAPACK gesdd; singular values
// are unique, so they agree to rounding, and the vectors agree up to the sign of each pair (and a rotation inside degenerate
// subspaces) — parity is checked on the singular values and on the reconstructed product.
//
// Method: columns of W (= A if m >= n, else A^T) are orthogonalised pair by pair; one launch handles one round of the
// round-robin tournament (n/2 disjoint column pairs, one CTA per pair: three dot products, then the rotation of the two
// columns of W and of the accumulated V).  HBM-bound: a round reads and writes W and V once, 16 B per element.
// A sweep is n-1 rounds; sweeps repeat until no pair needed a rotation (|a.b| <= tol |a||b|, tol = sqrt(m) eps, the dgesvj
// criterion).  Singular values are the column norms, sorted on the host (k doubles), U/V^T are emitted in that order.
// One sweep (n-1 launches of a few microseconds) is captured once into a CUDA graph and replayed per sweep.
#include "context.h"
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <numeric>
#include <vector>

namespace {

constexpr int JT = 256;

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// round-robin ("circle") pairing: npad even, rounds 0..npad-2, pair index 0..npad/2-1
__device__ __forceinline__ void pair_of(int round, int k, int npad, int& p, int& q) {
    const int M1 = npad - 1;
    if (k == 0) { p = M1; q = round; }
    else { p = (round + k) % M1; q = (round - k + M1) % M1; }
    if (p > q) { const int t = p; p = q; q = t; }
}

__global__ void __launch_bounds__(JT)
jacobi_round_kernel(double* __restrict__ W, long long ldw, int m, double* __restrict__ V, long long ldv, int nv,
                    int n, int npad, int round, double tol, int* __restrict__ rotated) {
    int p, q;
    pair_of(round, blockIdx.x, npad, p, q);
    if (q >= n) return;                                   // padding column
    double* a = W + (long long)p * ldw;
    double* b = W + (long long)q * ldw;
    double al = 0.0, be = 0.0, ga = 0.0;
    for (int i = threadIdx.x; i < m; i += JT) { const double x = a[i], y = b[i]; al += x * x; be += y * y; ga += x * y; }
    __shared__ double red[3][JT / 32];
    __shared__ double cs[2];
    al = warp_sum(al); be = warp_sum(be); ga = warp_sum(ga);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) { red[0][w] = al; red[1][w] = be; red[2][w] = ga; }
    __syncthreads();
    if (w == 0) {
        al = l < JT / 32 ? red[0][l] : 0.0; be = l < JT / 32 ? red[1][l] : 0.0; ga = l < JT / 32 ? red[2][l] : 0.0;
        al = warp_sum(al); be = warp_sum(be); ga = warp_sum(ga);
        if (l == 0) {
            double c = 1.0, s = 0.0;
            if (fabs(ga) > tol * sqrt(al * be) && ga != 0.0) {
                const double zeta = (be - al) / (2.0 * ga);
                const double t = copysign(1.0, zeta) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                c = 1.0 / sqrt(1.0 + t * t);
                s = c * t;
                if (s != 0.0) *rotated = 1;
            }
            cs[0] = c; cs[1] = s;
        }
    }
    __syncthreads();
    const double c = cs[0], s = cs[1];
    if (s == 0.0) return;
    for (int i = threadIdx.x; i < m; i += JT) { const double x = a[i], y = b[i]; a[i] = c * x - s * y; b[i] = s * x + c * y; }
    double* vp = V + (long long)p * ldv;
    double* vq = V + (long long)q * ldv;
    for (int i = threadIdx.x; i < nv; i += JT) { const double x = vp[i], y = vq[i]; vp[i] = c * x - s * y; vq[i] = s * x + c * y; }
}

__global__ void identity_kernel(double* V, long long ldv, int n) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (long long)n * n) return;
    const int i = (int)(t % n), j = (int)(t / n);
    V[(long long)j * ldv + i] = (i == j) ? 1.0 : 0.0;
}

__global__ void __launch_bounds__(JT) col_norm_kernel(const double* __restrict__ W, long long ldw, int m, double* __restrict__ S) {
    const double* a = W + (long long)blockIdx.x * ldw;
    double s = 0.0;
    for (int i = threadIdx.x; i < m; i += JT) { const double x = a[i]; s += x * x; }
    __shared__ double red[JT / 32];
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x < 32) {
        s = threadIdx.x < JT / 32 ? red[threadIdx.x] : 0.0;
        s = warp_sum(s);
        if (threadIdx.x == 0) S[blockIdx.x] = sqrt(s);
    }
}

// out[i, j] = in[i, perm[j]] (/ S[perm[j]])            i < rows, j < k
__global__ void emit_cols_kernel(double* __restrict__ out, long long ldo, const double* __restrict__ in, long long ldi,
                                 int rows, int k, const int* __restrict__ perm, const double* __restrict__ S) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x, j = blockIdx.y;
    if (i >= rows || j >= k) return;
    const int pj = perm[j];
    double v = in[(long long)pj * ldi + i];
    if (S) { const double s = S[pj]; v = s > 0.0 ? v / s : 0.0; }
    out[(long long)j * ldo + i] = v;
}
// out[j, i] = in[i, perm[j]] (/ S[perm[j]])            i < rows, j < k   (the transposed emission)
__global__ void emit_rows_kernel(double* __restrict__ out, long long ldo, const double* __restrict__ in, long long ldi,
                                 int rows, int k, const int* __restrict__ perm, const double* __restrict__ S) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x, j = blockIdx.y;
    if (i >= rows || j >= k) return;
    const int pj = perm[j];
    double v = in[(long long)pj * ldi + i];
    if (S) { const double s = S[pj]; v = s > 0.0 ? v / s : 0.0; }
    out[(long long)i * ldo + j] = v;
}

// out (n x m, ld ldo) = in^T, in (m x n, ld ldi)
__global__ void __launch_bounds__(256) transpose_kernel(const double* __restrict__ in, long long ldi, int m, int n,
                                                        double* __restrict__ out, long long ldo) {
    __shared__ double tile[32][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int i0 = blockIdx.x * 32, j0 = blockIdx.y * 32;
#pragma unroll
    for (int r = 0; r < 32; r += 8) {
        const int i = i0 + tx, j = j0 + ty + r;
        if (i < m && j < n) tile[ty + r][tx] = in[(long long)j * ldi + i];
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < 32; r += 8) {
        const int j = j0 + tx, i = i0 + ty + r;
        if (i < m && j < n) out[(long long)i * ldo + j] = tile[tx][ty + r];
    }
}

__global__ void diag_scale_kernel(double* __restrict__ M, long long ld, int rows, int cols, const double* __restrict__ s, int by_rows) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows) return;
    for (int j = blockIdx.y; j < cols; j += gridDim.y)          // grid.y is capped at 65535: wide matrices loop
        M[(long long)j * ld + i] = M[(long long)j * ld + i] * (by_rows ? s[i] : s[j]);
}

} // namespace

extern "C" {

int tci_svd_f64(tci_handle h, double* d_A, int64_t m64, int64_t n64, int64_t lda, double* d_U, int64_t ldu, double* h_S,
                double* d_Vt, int64_t ldvt, int* h_sweeps) {
    if (!h) return TCI_ERR_INVALID;
    if (m64 < 0 || n64 < 0 || m64 > INT32_MAX || n64 > INT32_MAX) return tci_fail(h, TCI_ERR_INVALID, "tci_svd_f64: bad sizes");
    if (h_sweeps) *h_sweeps = 0;
    const int m = (int)m64, n = (int)n64, k = m < n ? m : n;
    if (k == 0) return TCI_OK;
    if (!d_A || !h_S || lda < m || (d_U && ldu < m) || (d_Vt && ldvt < k)) return tci_fail(h, TCI_ERR_INVALID, "tci_svd_f64: bad arguments");
    if (k > 65535) return tci_fail(h, TCI_ERR_INVALID, "tci_svd_f64: min(m, n) = %d exceeds 65535", k);
    TCI_CUDA(h, cudaSetDevice(h->device));
    const bool tall = m >= n;
    const int wm = tall ? m : n, wn = k;                 // W is wm x wn; rotations accumulate in V (wn x wn)
    int rc = tci_ws_reserve(h, h->ws_a, sizeof(double) * (size_t)wn * wn);
    if (rc) return rc;
    rc = tci_ws_reserve(h, h->ws_c, sizeof(double) * (size_t)wn + sizeof(int) * (size_t)wn + 256);
    if (rc) return rc;
    rc = tci_pinned_reserve(h, sizeof(double) * (size_t)wn + sizeof(int) * (size_t)wn + 64);
    if (rc) return rc;
    double* V = (double*)h->ws_a.ptr;
    double* dS = (double*)h->ws_c.ptr;
    int* dperm = (int*)(dS + wn);
    int* dflag = dperm + wn;
    double* W = d_A; long long ldw = lda;
    // The sweep loop is launch-bound for the bond dimensions TCI produces (n - 1 rounds of a few microseconds each), so one
    // sweep is captured into a CUDA graph and replayed.  The legacy default stream cannot be captured: all of this call's
    // work runs on the handle's own stream, ordered after the caller's stream by an event.
    cudaStream_t caller = h->stream;
    cudaStream_t cs = h->own_stream ? h->own_stream : h->stream;
    if (cs != caller) {
        TCI_CUDA(h, cudaEventRecord(h->ev[3], caller));
        TCI_CUDA(h, cudaStreamWaitEvent(cs, h->ev[3], 0));
    }
    struct StreamGuard { tci_handle h; cudaStream_t s; ~StreamGuard() { h->stream = s; } } guard{h, caller};
    h->stream = cs;
    if (!tall) {
        rc = tci_ws_reserve(h, h->ws_b, sizeof(double) * (size_t)wm * wn);
        if (rc) return rc;
        W = (double*)h->ws_b.ptr; ldw = wm;
        transpose_kernel<<<dim3((m + 31) / 32, (n + 31) / 32), 256, 0, h->stream>>>(d_A, lda, m, n, W, ldw);
        TCI_LAUNCH_CHECK(h);
    }
    identity_kernel<<<(unsigned)(((long long)wn * wn + 255) / 256), 256, 0, h->stream>>>(V, wn, wn);
    TCI_LAUNCH_CHECK(h);
    const int npad = (wn + 1) & ~1;
    const double tol = sqrt((double)wm) * 1.1102230246251565e-16;
    int sweeps = 0, flag = wn > 1 ? 1 : 0;
    const int kMaxSweeps = 60;
    cudaGraphExec_t exec = nullptr;
    static const int use_graph = []() { const char* e = getenv("TCI_SVD_GRAPH"); return e ? atoi(e) : 1; }();
    if (flag && use_graph && npad - 1 >= 8 && cs != nullptr && cs != cudaStreamLegacy) {
        if (cudaStreamBeginCapture(cs, cudaStreamCaptureModeRelaxed) == cudaSuccess) {
            cudaMemsetAsync(dflag, 0, sizeof(int), cs);
            for (int round = 0; round < npad - 1; ++round)
                jacobi_round_kernel<<<npad / 2, JT, 0, cs>>>(W, ldw, wm, V, wn, wn, wn, npad, round, tol, dflag);
            cudaGraph_t g = nullptr;
            if (cudaStreamEndCapture(cs, &g) == cudaSuccess && g) {
                if (cudaGraphInstantiate(&exec, g, 0) != cudaSuccess) exec = nullptr;
                cudaGraphDestroy(g);
            }
        }
        (void)cudaGetLastError();                         // a failed capture falls back to plain launches
    }
    while (flag && sweeps < kMaxSweeps) {
        if (exec) {
            TCI_CUDA(h, cudaGraphLaunch(exec, cs));
            h->launches += npad - 1;
        } else {
            TCI_CUDA(h, cudaMemsetAsync(dflag, 0, sizeof(int), h->stream));
            for (int round = 0; round < npad - 1; ++round) {
                jacobi_round_kernel<<<npad / 2, JT, 0, h->stream>>>(W, ldw, wm, V, wn, wn, wn, npad, round, tol, dflag);
                TCI_LAUNCH_CHECK(h);
            }
        }
        TCI_CUDA(h, cudaMemcpyAsync(h->pinned, dflag, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
        TCI_CUDA(h, cudaStreamSynchronize(h->stream));
        flag = *(int*)h->pinned;
        ++sweeps;
    }
    if (exec) cudaGraphExecDestroy(exec);
    if (h_sweeps) *h_sweeps = sweeps;
    if (flag) return tci_fail(h, TCI_ERR_INVALID, "tci_svd_f64: Jacobi iteration did not converge in %d sweeps", kMaxSweeps);
    col_norm_kernel<<<wn, JT, 0, h->stream>>>(W, ldw, wm, dS);
    TCI_LAUNCH_CHECK(h);
    double* hS = (double*)h->pinned;
    TCI_CUDA(h, cudaMemcpyAsync(hS, dS, sizeof(double) * wn, cudaMemcpyDeviceToHost, h->stream));
    TCI_CUDA(h, cudaStreamSynchronize(h->stream));
    std::vector<int> perm(wn);
    std::iota(perm.begin(), perm.end(), 0);
    std::stable_sort(perm.begin(), perm.end(), [&](int a, int b) { return hS[a] > hS[b]; });
    for (int j = 0; j < wn; ++j) h_S[j] = hS[perm[j]];
    int* hperm = (int*)(hS + wn);
    for (int j = 0; j < wn; ++j) hperm[j] = perm[j];
    TCI_CUDA(h, cudaMemcpyAsync(dperm, hperm, sizeof(int) * wn, cudaMemcpyHostToDevice, h->stream));
    // tall: A = (W/S) S V^T        -> U = W[:, perm]/S,  Vt[j, :] = V[:, perm[j]]
    // wide: A^T = (W/S) S V^T      -> U = V[:, perm],    Vt[j, :] = W[:, perm[j]]/S
    if (d_U) {
        emit_cols_kernel<<<dim3((m + 255) / 256, k), 256, 0, h->stream>>>(d_U, ldu, tall ? W : V, tall ? ldw : wn, m, k, dperm, tall ? dS : nullptr);
        TCI_LAUNCH_CHECK(h);
    }
    if (d_Vt) {
        emit_rows_kernel<<<dim3((n + 255) / 256, k), 256, 0, h->stream>>>(d_Vt, ldvt, tall ? V : W, tall ? wn : ldw, n, k, dperm, tall ? nullptr : dS);
        TCI_LAUNCH_CHECK(h);
    }
    TCI_CUDA(h, cudaStreamSynchronize(h->stream));       // hperm (pinned) must stay valid until the copy has run
    return TCI_OK;
}

int tci_diag_scale_f64(tci_handle h, double* d_M, int64_t rows, int64_t cols, int64_t ld, const double* h_s, int by_rows) {
    if (!h) return TCI_ERR_INVALID;
    if (rows < 0 || cols < 0 || rows > INT32_MAX || cols > INT32_MAX) return tci_fail(h, TCI_ERR_INVALID, "tci_diag_scale_f64: bad sizes");
    if (rows == 0 || cols == 0) return TCI_OK;
    if (!d_M || !h_s || ld < rows) return tci_fail(h, TCI_ERR_INVALID, "tci_diag_scale_f64: bad arguments");
    TCI_CUDA(h, cudaSetDevice(h->device));
    const size_t ns = (size_t)(by_rows ? rows : cols);
    int rc = tci_ws_reserve(h, h->ws_c, sizeof(double) * ns);
    if (rc) return rc;
    std::vector<double> stage(h_s, h_s + ns);             // pageable copy: returns after the data is staged
    TCI_CUDA(h, cudaMemcpyAsync(h->ws_c.ptr, stage.data(), sizeof(double) * ns, cudaMemcpyHostToDevice, h->stream));
    diag_scale_kernel<<<dim3((unsigned)((rows + 255) / 256), (unsigned)(cols > 65535 ? 65535 : cols)), 256, 0, h->stream>>>(d_M, ld, (int)rows, (int)cols,
                                                                                                 (const double*)h->ws_c.ptr, by_rows);
    TCI_LAUNCH_CHECK(h);
    return TCI_OK;
}

} // extern "C"
