// solve_small.h — X = B * inv(P) for small pivot matrices (tensorci2.jl:626), the pieces shared by tci_solve_right_f64's
// one-launch kernel (dense.cu) and the batched site-tensor kernel of the device-resident sweeps (sweep.cu): a CTA
// factorises P in shared memory by full-pivot LU, then every thread carries one row of B through the two substitutions.
// P[rp, cp] = L U  =>  X[:, rp] = (B[:, cp] U^-1) L^-1.
#pragma once

constexpr int kSolveRows = 128;                            // rows of B per CTA = threads per CTA
constexpr int kSolveMaxR = 64;

template <int NT> struct SmallLuScratch { double a[NT / 32]; int e[NT / 32]; int best; };

// S[j * lds + i] = P(i, j) (r x r) becomes L \ U in swapped coordinates; rp / cp = the row / column permutations.  Returns
// true when P is singular (zero or all-NaN trailing block).  All NT threads of the CTA must call it; S must be complete
// (barrier) before the call, and the factors are complete (barrier) after it.
template <int NT>
__device__ __forceinline__ bool small_lu_fullpivot(double* S, const int lds, const int r, int* rp, int* cp, SmallLuScratch<NT>& sc) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int t = tid; t < r; t += NT) { rp[t] = t; cp[t] = t; }
    __syncthreads();
    for (int k = 0; k < r; ++k) {
        // full pivot search over the trailing block; ties -> smallest column, then smallest row (deterministic, so every CTA
        // takes the same decisions); NaN never wins
        const int rem = r - k;
        double ba = -1.0; int be = 0x7fffffff;
        for (int e = tid; e < rem * rem; e += NT) {
            const double a = fabs(S[(k + e / rem) * lds + k + e % rem]);
            if (a > ba) { ba = a; be = e; }
        }
        for (int d = 16; d > 0; d >>= 1) {
            const double oa = __shfl_down_sync(0xffffffffu, ba, d);
            const int oe = __shfl_down_sync(0xffffffffu, be, d);
            if (oa > ba || (oa == ba && oe < be)) { ba = oa; be = oe; }
        }
        if (lane == 0) { sc.a[warp] = ba; sc.e[warp] = be; }
        __syncthreads();
        if (tid == 0) {
            for (int w = 1; w < NT / 32; ++w)
                if (sc.a[w] > ba || (sc.a[w] == ba && sc.e[w] < be)) { ba = sc.a[w]; be = sc.e[w]; }
            sc.best = (ba > 0.0) ? be : -1;                // zero (or all-NaN) trailing block: P is singular
        }
        __syncthreads();
        const int e = sc.best;
        if (e < 0) return true;
        const int pi = k + e % rem, pj = k + e / rem;
        if (pi != k) for (int j = tid; j < r; j += NT) { const double a = S[j * lds + k]; S[j * lds + k] = S[j * lds + pi]; S[j * lds + pi] = a; }
        __syncthreads();
        if (pj != k) for (int i = tid; i < r; i += NT) { const double a = S[k * lds + i]; S[k * lds + i] = S[pj * lds + i]; S[pj * lds + i] = a; }
        if (tid == 0) { int t = rp[k]; rp[k] = rp[pi]; rp[pi] = t; t = cp[k]; cp[k] = cp[pj]; cp[pj] = t; }
        __syncthreads();
        const double piv = S[k * lds + k];
        for (int i = k + 1 + tid; i < r; i += NT) S[k * lds + i] = S[k * lds + i] / piv;
        __syncthreads();
        const int rm = rem - 1;
        for (int t = tid; t < rm * rm; t += NT) {
            const int i = k + 1 + t % rm, j = k + 1 + t / rm;
            S[j * lds + i] = S[j * lds + i] - S[k * lds + i] * S[j * lds + k];
        }
        __syncthreads();
    }
    return false;
}

// W[k * NT + tid], k < r: this thread's row of B[:, cp] on entry, its row of X[:, rp] on return.
template <int NT>
__device__ __forceinline__ void small_solve_row(double* W, const double* S, const int lds, const int r) {
    const int tid = threadIdx.x;
    for (int k = 0; k < r; ++k) {                          // Z U = B[:, cp]   (U(j, k) = S[k * lds + j], j <= k)
        double s = W[k * NT + tid];
        for (int j = 0; j < k; ++j) s = s - W[j * NT + tid] * S[k * lds + j];
        W[k * NT + tid] = s / S[k * lds + k];
    }
    for (int k = r - 2; k >= 0; --k) {                     // Xr L = Z, L unit lower (L(j, k) = S[k * lds + j], j > k)
        double s = W[k * NT + tid];
        for (int j = k + 1; j < r; ++j) s = s - W[j * NT + tid] * S[k * lds + j];
        W[k * NT + tid] = s;
    }
}
