// rrlu.cu — K2 + K3: full-pivot rank-revealing LU on the device.
//
// Restates (does not translate) src/matrixlu.jl:346-396 (_optimizerrlu!), :46-87 (submatrixargmax),
// :295-322 (addpivot!) of the reference for one B200:
//
//   per pivot step k                      reference                          here
//   ------------------------------------  ---------------------------------  ---------------------------------
//   find argmax |A[i,j]|^2, i,j >= k      separate full scan (:359)          fused into the previous step's
//                                                                            update pass (no extra HBM read)
//   swap rows k<->pr, cols k<->pc,        two strided whole-row/col loops    rrlu_swap_kernel: one thread per
//   scale the pivot column (or row)       (:254-275, :303-309)               swapped pair, corners by one thread
//   A[i,j] -= x[i]*y[j]                   scalar double loop (:316-320)      rrlu_update_kernel<...>: 128-bit
//                                                                            coalesced column accesses, pivot row
//                                                                            read from a packed copy (ybuf)
//   stop test                             host (:363)                        last-finishing block of the update
//                                                                            kernel (device-side state, no host
//                                                                            round trip per step)
//
// The matrix keeps the reference's physically swapped layout, so the argmax tie rule (smallest
// column, then smallest row, in swapped coordinates) and the final tril/triu extraction are
// bit-identical to the reference, ties included.  Arithmetic is __dmul_rn/__dsub_rn (no FMA
// contraction; Julia does not contract) and IEEE division.
//
// HBM roofline: 16 B per updated element (one FP64 read + one write); the fused
// search adds no bytes.  DESIGN.md has the per-kernel figures.
#include "context.h"
#include <cmath>
#include <cstdlib>
#include <cstring>

namespace {

constexpr int kThreads = 256;

struct Cand {            // argmax candidate: signed value and 0-based position; i < 0 means "none yet"
    double x;
    int i, j;
};

struct RrluState {
    int npivot;          // pivots placed so far
    int stop;            // 1 once the loop of matrixlu.jl:356-369 has ended
    int pr, pc;          // pending pivot (0-based, swapped coordinates) for step `npivot`
    double pivval;       // A[pr, pc]
    double maxerror;
    double error;        // lu.error
    unsigned int ticket; // blocks of the running update kernel that have finished
    int flags;           // bit 0: NaN in L, bit 1: NaN in U, bit 2: grid barrier timed out
    long long tphase[8]; // persistent kernel: clock64 cycles block 0 spent per phase (diagnostics)
};

__device__ __forceinline__ bool cand_better(const Cand& a, const Cand& b) {
    // true iff a precedes b in the reference's scan: larger abs2, then smaller column, then smaller row
    if (a.i < 0) return false;
    if (b.i < 0) return true;
    const double va = __dmul_rn(a.x, a.x), vb = __dmul_rn(b.x, b.x);
    if (va > vb) return true;
    if (va == vb) return (a.j < b.j) || (a.j == b.j && a.i < b.i);
    return false;
}

__device__ __forceinline__ Cand cand_shfl_down(const Cand& c, int delta) {
    Cand o;
    o.x = __shfl_down_sync(0xffffffffu, c.x, delta);
    o.i = __shfl_down_sync(0xffffffffu, c.i, delta);
    o.j = __shfl_down_sync(0xffffffffu, c.j, delta);
    return o;
}

__device__ __forceinline__ Cand block_reduce(Cand c, Cand* smem /* kThreads/32 entries */) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        Cand o = cand_shfl_down(c, d);
        if (cand_better(o, c)) c = o;
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) smem[warp] = c;
    __syncthreads();
    if (warp == 0) {
        c = (lane < kThreads / 32) ? smem[lane] : Cand{0.0, -1, -1};
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            Cand o = cand_shfl_down(c, d);
            if (cand_better(o, c)) c = o;
        }
    }
    return c;   // valid in thread 0
}

__device__ __forceinline__ double jl_max(double a, double b) {   // Julia's max propagates NaN
    if (isnan(a) || isnan(b)) return nan("");
    return a > b ? a : b;
}

__global__ void rrlu_init_kernel(RrluState* st, int* rowperm, int* colperm, int m, int n, unsigned int* bar) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < m) rowperm[t] = t;
    if (t < n) colperm[t] = t;
    if (t == 0) {
        st->npivot = 0; st->stop = 0; st->pr = 0; st->pc = 0; st->pivval = 0.0;
        st->maxerror = 0.0; st->error = nan(""); st->ticket = 0u; st->flags = 0;
        if (bar) *bar = 0u;
    }
}

// One pass over the trailing block.  UPDATE: apply the rank-1 Schur update of the pivot that the swap
// kernel just placed at position k = st->npivot, and find the next pivot in the same pass.
// !UPDATE: only the search over the whole matrix (first pivot).  The last block to finish reduces the
// per-block candidates and evaluates the stop rule (matrixlu.jl:360-367).
// Tiles (TR rows x COLS columns) are dealt round-robin to a grid sized to exactly one resident wave; odd
// steps walk the tiles backwards so that the ~L2-sized tail written by step k is the first thing
// step k+1 reads (serpentine order: those reads, and the re-writes of the still-dirty lines, never
// reach HBM).
// DIST (column-sharded factorisation): local column j stands at position vpos[j] of the reference's swapped layout, and ties
// are broken by that position instead of the local index.
#define TCI_CONSIDER(q, val, ii, jj)                                                             \
    if ((q) > bv || ((q) == bv && ((jj) < bj || ((jj) == bj && (ii) < bi)))) { bv = (q); bx = (val); bi = (ii); bj = (jj); }
#define TCI_CONSIDER_D(q, val, ii, jj)                                                           \
    if ((q) > bv || ((q) == bv && (col_before<DIST>((jj), bj, vpos) || ((jj) == bj && (ii) < bi)))) { bv = (q); bx = (val); bi = (ii); bj = (jj); }

struct ColOrder { const int* gid; const int* col2pos; };   // local column -> global column id -> position in the swapped order

template <bool DIST> __device__ __forceinline__ bool col_before(int a, int b, const ColOrder& vpos) {
    if (!DIST) return a < b;
    return b < 0 ? true : vpos.col2pos[vpos.gid[a]] < vpos.col2pos[vpos.gid[b]];
}

// Loads that must see data written by OTHER blocks earlier in the same (persistent) kernel go around L1.
template <bool NOL1> __device__ __forceinline__ double2 ld2(const double* p) {
    return NOL1 ? __ldcg(reinterpret_cast<const double2*>(p)) : *reinterpret_cast<const double2*>(p);
}
template <bool NOL1> __device__ __forceinline__ double ld1(const double* p) { return NOL1 ? __ldcg(p) : *p; }

// One pass of this block over its tiles of the trailing block [r0, m) x [c0, n), r0 = c0 = np.  UPDATE: apply the
// rank-1 update of pivot k = np - 1 (scaled pivot column in A[:, k], packed pivot row in ybuf) and track the
// abs2-argmax of the updated values; !UPDATE: search only.
template <bool UPDATE, bool VEC, int COLS, bool NOL1, bool DIST = false>
__device__ __forceinline__ void scan_tiles(double* __restrict__ A, int m, int n, long long lda, const double* __restrict__ ybuf,
                                           const double* __restrict__ xcol /* scaled pivot column, indexed by row (UPDATE only) */,
                                           int r0, int c0, int np, int serpentine, const ColOrder vpos,
                                           double& bv, double& bx, int& bi, int& bj) {
    const int smode = serpentine & 7;
    const bool rev = smode == 2 ? true : (smode == 3 ? !(np & 1) : (smode && (np & 1)));
    const int nct = (n - c0 + COLS - 1) / COLS;
    if (VEC) {
        // Row tiling: start at the 128-byte line that holds row r0 (rows below r0 are masked), and split the
        // height into nrt EQUAL tiles (a ragged last tile would leave the blocks that own it short of work:
        // tiles are dealt round-robin and gridDim.x and nrt share factors).
        constexpr int TR = kThreads * 2;
        const int rbase = r0 & ~15;
        const int h = m - rbase;
        const int nrt = (h + TR - 1) / TR;
        const int rpt = (((h + nrt - 1) / nrt) + 15) & ~15;          // rows per tile, multiple of 16, <= TR
        const int ntiles = nrt * nct;                                // < 2^31: m, n < 2^31 - 1024 and TR * COLS >= 2048
        const int il = threadIdx.x * 2;
        // serpentine & 8: contiguous tile ranges per block (each block streams down its own columns) instead of round-robin
        const bool blocked = (serpentine & 8) != 0;
        const int per = (ntiles + (int)gridDim.x - 1) / (int)gridDim.x;
        const int tb = blocked ? (int)blockIdx.x * per : (int)blockIdx.x;
        const int te = blocked ? min(ntiles, tb + per) : ntiles;
        const int ts = blocked ? 1 : (int)gridDim.x;
        if (il < rpt) {
            for (int t0 = tb; t0 < te; t0 += ts) {
                const int t = rev ? ntiles - 1 - t0 : t0;
                const int rt = t % nrt, ct = t / nrt;
                const int i0 = rbase + rt * rpt + il;
                const int j0 = c0 + ct * COLS;
                if (i0 + 1 < r0 || i0 >= m) continue;               // both rows masked, or past the end
                if (i0 + 1 < m) {
                    const bool v0 = i0 >= r0;                       // false only for the pair that straddles r0
                    double2 x = make_double2(0.0, 0.0);
                    if (UPDATE) x = ld2<NOL1>(xcol + i0);
                    double2 a[COLS];
#pragma unroll
                    for (int c = 0; c < COLS; ++c)
                        if (j0 + c < n) a[c] = ld2<NOL1>(A + (long long)(j0 + c) * lda + i0);
#pragma unroll
                    for (int c = 0; c < COLS; ++c) {
                        const int j = j0 + c;
                        if (j < n) {
                            double2 v = a[c];
                            if (UPDATE) {
                                const double y = ld1<NOL1>(ybuf + j);
                                if (v0) v.x = __dsub_rn(v.x, __dmul_rn(x.x, y));    // row r0-1 is the pivot row: rewritten unchanged
                                v.y = __dsub_rn(v.y, __dmul_rn(x.y, y));
                                *reinterpret_cast<double2*>(A + (long long)j * lda + i0) = v;
                            }
                            const double q0 = v0 ? __dmul_rn(v.x, v.x) : -INFINITY, q1 = __dmul_rn(v.y, v.y);
                            TCI_CONSIDER_D(q0, v.x, i0, j)
                            TCI_CONSIDER_D(q1, v.y, i0 + 1, j)
                        }
                    }
                } else {                                            // i0 == m - 1: odd tail row, scalar
                    const int i = i0;
                    if (i < r0) continue;
                    const double x = UPDATE ? ld1<NOL1>(xcol + i) : 0.0;
#pragma unroll
                    for (int c = 0; c < COLS; ++c) {
                        const int j = j0 + c;
                        if (j < n) {
                            double v = ld1<NOL1>(A + (long long)j * lda + i);
                            if (UPDATE) { v = __dsub_rn(v, __dmul_rn(x, ld1<NOL1>(ybuf + j))); A[(long long)j * lda + i] = v; }
                            const double q = __dmul_rn(v, v);
                            TCI_CONSIDER_D(q, v, i, j)
                        }
                    }
                }
            }
        }
    } else {
        constexpr int TR = kThreads;
        const int nrt = (m - r0 + TR - 1) / TR;
        const long long ntiles = (long long)nrt * nct;
        for (long long t0 = blockIdx.x; t0 < ntiles; t0 += gridDim.x) {
            const long long t = rev ? ntiles - 1 - t0 : t0;
            const int rt = (int)(t % nrt), ct = (int)(t / nrt);
            const int i = r0 + rt * TR + threadIdx.x;
            const int j0 = c0 + ct * COLS;
            if (i >= m) continue;
            const double x = UPDATE ? ld1<NOL1>(xcol + i) : 0.0;
            double a[COLS];
#pragma unroll
            for (int c = 0; c < COLS; ++c)
                if (j0 + c < n) a[c] = ld1<NOL1>(A + (long long)(j0 + c) * lda + i);
#pragma unroll
            for (int c = 0; c < COLS; ++c) {
                const int j = j0 + c;
                if (j < n) {
                    double v = a[c];
                    if (UPDATE) { v = __dsub_rn(v, __dmul_rn(x, ld1<NOL1>(ybuf + j))); A[(long long)j * lda + i] = v; }
                    const double q = __dmul_rn(v, v);
                    TCI_CONSIDER_D(q, v, i, j)
                }
            }
        }
    }
}

template <bool UPDATE, bool VEC, int COLS>
__global__ void __launch_bounds__(kThreads, (COLS <= 4 ? 5 : (COLS == 8 ? 3 : 2)))
rrlu_update_kernel(double* __restrict__ A, int m, int n, long long lda, const double* __restrict__ ybuf,
                   RrluState* st, Cand* parts, int maxrank, double reltol, double abstol, int start0, int serpentine) {
    __shared__ Cand s_c[kThreads / 32];
    __shared__ int s_last;
    cudaGridDependencySynchronize();                    // programmatic dependent launch: block start-up overlaps the previous kernel's tail
    if (st->stop) return;
    const int k = st->npivot;
    const int np = UPDATE ? k + 1 : start0;             // pivots in place once this pass is done
    double bv = -INFINITY, bx = 0.0;
    int bi = -1, bj = -1;
    if (np < maxrank)                                   // otherwise the loop ends; nothing reads the block
        scan_tiles<UPDATE, VEC, COLS, false>(A, m, n, lda, ybuf, A + (long long)k * lda, np, np, np, serpentine, ColOrder{nullptr, nullptr}, bv, bx, bi, bj);

    Cand c = block_reduce(Cand{bx, bi, bj}, s_c);
    if (threadIdx.x == 0) {
        parts[blockIdx.x] = c;
        __threadfence();
        const unsigned int done = atomicAdd(&st->ticket, 1u);
        s_last = (done == gridDim.x - 1);
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    Cand best{0.0, -1, -1};
    for (int b = threadIdx.x; b < (int)gridDim.x; b += kThreads) {
        Cand o = parts[b];
        if (cand_better(o, best)) best = o;
    }
    __syncthreads();
    best = block_reduce(best, s_c);
    if (threadIdx.x == 0) {
        st->npivot = np;
        st->ticket = 0u;
        if (np >= maxrank) {
            st->stop = 1;                                           // while lu.npivot < maxrank   (:356)
        } else {
            if (best.i < 0) { best.i = np; best.j = np; best.x = nan(""); }   // every entry NaN: first(rows), first(cols)
            const double err = fabs(best.x);                        // :360
            st->error = err;
            if ((err < reltol * st->maxerror || err < abstol) && np > 0) {    // :363
                st->stop = 1;
            } else {
                st->maxerror = jl_max(st->maxerror, err);           // :367
                st->pr = best.i; st->pc = best.j; st->pivval = best.x;
            }
        }
    }
}

// swaprow!/swapcol! (:254-275) + the scaling of addpivot! (:303-309) for the pending pivot, plus the
// packed copy of the new pivot row that the update kernel reads.  Thread t < n handles column t of the
// two swapped rows, n <= t < n+m handles row t-n of the two swapped columns, t == n+m the four corners
// and the permutation bookkeeping.
template <bool NOL1>
__device__ __forceinline__ void swap_item(int t, double* __restrict__ A, int m, int n, long long lda, double* __restrict__ ybuf,
                                          int k, int pr, int pc, double piv, int* rowperm, int* colperm, double* pivvals, int leftorth) {
    if (t < n) {
        const int j = t;
        if (j == k || j == pc) return;
        double* pk = A + (long long)j * lda + k;
        double* pp = A + (long long)j * lda + pr;
        const double a = ld1<NOL1>(pk);
        double newk = (pr != k) ? ld1<NOL1>(pp) : a;
        if (!leftorth && j > k) newk = newk / piv;
        *pk = newk;
        if (pr != k) *pp = a;
        if (j > k) ybuf[j] = newk;
    } else if (t < n + m) {
        const int i = t - n;
        if (i == k || i == pr) return;
        double* pk = A + (long long)k * lda + i;
        double* pp = A + (long long)pc * lda + i;
        const double a = ld1<NOL1>(pk);
        double newk = (pc != k) ? ld1<NOL1>(pp) : a;
        if (leftorth && i > k) newk = newk / piv;
        *pk = newk;
        if (pc != k) *pp = a;
    } else if (t == n + m) {
        const double o_kk = ld1<NOL1>(A + (long long)k * lda + k), o_kpc = ld1<NOL1>(A + (long long)pc * lda + k);
        const double o_prk = ld1<NOL1>(A + (long long)k * lda + pr), o_prpc = ld1<NOL1>(A + (long long)pc * lda + pr);
        double n_kk = o_prpc, n_kpc = o_prk, n_prk = o_kpc, n_prpc = o_kk;
        if (leftorth) { if (pr > k) n_prk = n_prk / piv; }
        else          { if (pc > k) n_kpc = n_kpc / piv; }
        // When pr == k or pc == k two of the four corners are the same address; this store order makes
        // the (possibly scaled) canonical value the one that lands last.
        A[(long long)pc * lda + pr] = n_prpc;
        A[(long long)pc * lda + k] = n_kpc;
        A[(long long)k * lda + pr] = n_prk;
        A[(long long)k * lda + k] = n_kk;
        if (pc > k) ybuf[pc] = n_kpc;
        int tmp = rowperm[k]; rowperm[k] = rowperm[pr]; rowperm[pr] = tmp;
        tmp = colperm[k]; colperm[k] = colperm[pc]; colperm[pc] = tmp;
        pivvals[k] = piv;
    }
}

__global__ void rrlu_swap_kernel(double* __restrict__ A, int m, int n, long long lda, double* __restrict__ ybuf,
                                 RrluState* st, int* rowperm, int* colperm, double* pivvals, int leftorth) {
    cudaGridDependencySynchronize();
    if (st->stop) return;
    swap_item<false>(blockIdx.x * blockDim.x + threadIdx.x, A, m, n, lda, ybuf, st->npivot, st->pr, st->pc, st->pivval,
                     rowperm, colperm, pivvals, leftorth);
}

// ---- single-CTA variant: the matrix lives in shared memory for the whole factorisation -------------------------------
// For the Pi matrices of low-rank problems (BASELINE C1/C3/C4: at most a few hundred rows and columns) every pivot step is
// pure latency.  One CTA of 1024 threads keeps A (<= ~200 KB) in shared memory in the reference's physically swapped
// layout and runs search -> swaps -> scale -> update with __syncthreads between the phases: ~1-2 us per pivot instead of
// two kernel launches or two grid barriers.  Same arithmetic (mul_rn/sub_rn, IEEE division) and same tie rule.
constexpr int kSmemThreads = 1024;

template <int NW>
__device__ __forceinline__ Cand block_reduce_all(Cand c, Cand* smem /* 32 entries */) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        Cand o = cand_shfl_down(c, d);
        if (cand_better(o, c)) c = o;
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) smem[warp] = c;
    __syncthreads();
    c = (lane < NW) ? smem[lane] : Cand{0.0, -1, -1};     // every warp reduces the NW partials redundantly: no broadcast step
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        Cand o;
        o.x = __shfl_xor_sync(0xffffffffu, c.x, d); o.i = __shfl_xor_sync(0xffffffffu, c.i, d); o.j = __shfl_xor_sync(0xffffffffu, c.j, d);
        if (cand_better(o, c)) c = o;
    }
    __syncthreads();                                      // smem[] may be rewritten by the next reduction
    return c;                                             // valid in every thread
}

// NT threads: 1024 in general; 256 for matrices of at most 64 x 64 elements' worth of work, where 8 warps cover the columns
// and every __syncthreads is several times cheaper (the quantics configs: Pi <= 34 x 34, ~430 factorisations per run).
template <int NT>
__global__ void __launch_bounds__(NT, 1)
rrlu_smem_kernel(double* __restrict__ A, int m, int n, long long lda, RrluState* st, int* g_rowperm, int* g_colperm,
                 double* pivvals, int maxrank, double reltol, double abstol, int leftorth,
                 double* __restrict__ L, long long ldl, double* __restrict__ U, long long ldu, int extract) {
    extern __shared__ double S[];                         // m x n, column-major, leading dimension lds (odd: conflict-free rows)
    __shared__ Cand s_c[32];
    const int lds = m | 1;
    // the permutations live in shared memory too (after S): thread 0 swaps two entries of each per pivot, and a global
    // round trip there would sit on the critical path of every step
    int* rowperm = (int*)(S + (size_t)lds * n);
    int* colperm = rowperm + m;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int j = warp; j < n; j += NT / 32)
        for (int i = lane; i < m; i += 32) S[j * lds + i] = A[(long long)j * lda + i];
    for (int t = tid; t < m; t += NT) rowperm[t] = t;
    for (int t = tid; t < n; t += NT) colperm[t] = t;
    if (tid == 0) st->flags = 0;                          // this kernel needs no rrlu_init_kernel before it
    __syncthreads();
    int np = 0;
    double maxerror = 0.0, error = nan("");
    double bv = -INFINITY, bx = 0.0; int bi = -1, bj = -1;
    for (int j = warp; j < n; j += NT / 32)
        for (int i = lane; i < m; i += 32) { const double v = S[j * lds + i]; const double q = __dmul_rn(v, v); TCI_CONSIDER(q, v, i, j) }
    for (;;) {
        Cand best = block_reduce_all<NT / 32>(Cand{bx, bi, bj}, s_c);
        if (best.i < 0) { best.i = np; best.j = np; best.x = nan(""); }
        const double err = fabs(best.x);                                        // :360
        error = err;
        if ((err < reltol * maxerror || err < abstol) && np > 0) break;         // :363
        maxerror = jl_max(maxerror, err);                                       // :367
        const int k = np, pr = best.i, pc = best.j;
        const double piv = best.x;
        // swaprow!(k, pr) over all columns, then swapcol!(k, pc) over all rows (:254-275, :297-298)
        if (pr != k) for (int j = tid; j < n; j += NT) { const double a = S[j * lds + k]; S[j * lds + k] = S[j * lds + pr]; S[j * lds + pr] = a; }
        __syncthreads();
        if (pc != k) for (int i = tid; i < m; i += NT) { const double a = S[k * lds + i]; S[k * lds + i] = S[pc * lds + i]; S[pc * lds + i] = a; }
        if (tid == 0) {
            int t = rowperm[k]; rowperm[k] = rowperm[pr]; rowperm[pr] = t;
            t = colperm[k]; colperm[k] = colperm[pc]; colperm[pc] = t;
            pivvals[k] = piv;
        }
        __syncthreads();
        if (leftorth) { for (int i = k + 1 + tid; i < m; i += NT) S[k * lds + i] = S[k * lds + i] / piv; }      // :303-305
        else          { for (int j = k + 1 + tid; j < n; j += NT) S[j * lds + k] = S[j * lds + k] / piv; }      // :307-309
        __syncthreads();
        np = k + 1;
        if (np >= maxrank) break;                                               // :356
        bv = -INFINITY; bx = 0.0; bi = -1; bj = -1;
        for (int j = np + warp; j < n; j += NT / 32) {                               // :316-320 + the next search
            const double y = S[j * lds + k];
            for (int i = np + lane; i < m; i += 32) {
                const double v = __dsub_rn(S[j * lds + i], __dmul_rn(S[k * lds + i], y));
                S[j * lds + i] = v;
                const double q = __dmul_rn(v, v);
                TCI_CONSIDER(q, v, i, j)
            }
        }
    }
    __syncthreads();
    for (int j = warp; j < n; j += NT / 32)
        for (int i = lane; i < m; i += 32) A[(long long)j * lda + i] = S[j * lds + i];
    for (int t = tid; t < m; t += NT) g_rowperm[t] = rowperm[t];
    for (int t = tid; t < n; t += NT) g_colperm[t] = colperm[t];
    if (tid == 0) { st->npivot = np; st->stop = 1; st->error = error; st->maxerror = maxerror; }
    if (extract) {
        // L = tril(A[:, 1:r]), U = triu(A[1:r, :]), NaN flags, unit diagonal (:372-388): what rrlu_extract_kernel does, from
        // shared memory, so the small-matrix path is ONE launch
        const int r = np;
        int flags = 0;
        for (int j = warp; j < r; j += NT / 32)
            for (int i = lane; i < m; i += 32) {
                double v = (i >= j) ? S[j * lds + i] : 0.0;
                if (isnan(v)) flags |= 1;
                if (i == j && leftorth) v = 1.0;
                if (L) L[(long long)j * ldl + i] = v;
            }
        for (int j = warp; j < n; j += NT / 32)
            for (int i = lane; i < r; i += 32) {
                double v = (j >= i) ? S[j * lds + i] : 0.0;
                if (isnan(v)) flags |= 2;
                if (i == j && !leftorth) v = 1.0;
                if (U) U[(long long)j * ldu + i] = v;
            }
        if (flags) atomicOr(&st->flags, flags);
    }
}

// Second generation of the single-CTA kernel: TWO block barriers per pivot instead of six.  The first generation follows the
// reference phase by phase (search | swap rows | swap columns | scale | update), each phase waiting for the previous one
// through __syncthreads, and at ~0.5 us per phase that chain IS the run time of a small factorisation.  Here, once the pivot
// (pr, pc) is known, one short phase copies what the swaps would overwrite — old row k and old column k — and forms the final
// pivot column and row (already swapped, the column or row already divided by the pivot); after ONE barrier a single pass
// writes every element its final value: moved rows/columns come from the copies, the trailing block gets
// final(i, j) = old(sigma(i), tau(j)) - Lk[i] * Uk[j]  (sigma, tau = the two transpositions) and is searched for the next
// pivot on the way.  Operands and operations per element are exactly those of the first generation, so results stay bit-identical.
template <int NT>
__global__ void __launch_bounds__(NT, 1)
rrlu_smem2_kernel(double* __restrict__ A, int m, int n, long long lda, RrluState* st, int* g_rowperm, int* g_colperm,
                  double* pivvals, int maxrank, double reltol, double abstol, int leftorth,
                  double* __restrict__ L, long long ldl, double* __restrict__ U, long long ldu, int extract) {
    extern __shared__ double S[];                         // m x n, column-major, leading dimension lds
    __shared__ Cand s_c[2][32];                           // double-buffered by step parity: no barrier to protect its reuse
    const int lds = m | 1;
    double* Lk = S + (size_t)lds * n;                     // final column k (all rows; rows > k scaled when leftorthogonal)
    double* Uk = Lk + m;                                  // final row k (all columns; columns > k scaled otherwise)
    double* colK = Uk + n;                                // old column k
    double* rowK = colK + m;                              // old row k
    int* rowperm = (int*)(rowK + n);
    int* colperm = rowperm + m;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int NW = NT / 32;
    for (int j = warp; j < n; j += NW)
        for (int i = lane; i < m; i += 32) S[j * lds + i] = A[(long long)j * lda + i];
    for (int t = tid; t < m; t += NT) rowperm[t] = t;
    for (int t = tid; t < n; t += NT) colperm[t] = t;
    if (tid == 0) st->flags = 0;
    __syncthreads();
    int np = 0;
    double maxerror = 0.0, error = nan("");
    double bv = -INFINITY, bx = 0.0; int bi = -1, bj = -1;
    for (int j = warp; j < n; j += NW)
        for (int i = lane; i < m; i += 32) { const double v = S[j * lds + i]; const double q = __dmul_rn(v, v); TCI_CONSIDER(q, v, i, j) }
    for (;;) {
        // ---- reduce the per-thread candidates; every thread ends up with the winner (barrier 1)
        Cand c{bx, bi, bj};
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) { Cand o = cand_shfl_down(c, d); if (cand_better(o, c)) c = o; }
        Cand* sc = s_c[np & 1];
        if (lane == 0) sc[warp] = c;
        __syncthreads();
        Cand best = (lane < NW) ? sc[lane] : Cand{0.0, -1, -1};
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            Cand o;
            o.x = __shfl_xor_sync(0xffffffffu, best.x, d); o.i = __shfl_xor_sync(0xffffffffu, best.i, d); o.j = __shfl_xor_sync(0xffffffffu, best.j, d);
            if (cand_better(o, best)) best = o;
        }
        if (best.i < 0) { best.i = np; best.j = np; best.x = nan(""); }
        const double err = fabs(best.x);                                        // :360
        error = err;
        if ((err < reltol * maxerror || err < abstol) && np > 0) break;         // :363
        maxerror = jl_max(maxerror, err);                                       // :367
        const int k = np, pr = best.i, pc = best.j;
        const double piv = best.x;
        // ---- save what the swaps overwrite, form the final pivot column and row (barrier 2)
        for (int t = tid; t < m + n; t += NT) {
            if (t < m) {
                const int i = t, si = (i == k) ? pr : ((i == pr) ? k : i);      // sigma: rows k <-> pr
                colK[i] = S[k * lds + i];
                const double v = S[pc * lds + si];                              // final column k = old column pc, rows swapped
                Lk[i] = (leftorth && i > k) ? v / piv : v;                      // :303-305
            } else {
                const int j = t - m, tj = (j == k) ? pc : ((j == pc) ? k : j);  // tau: columns k <-> pc
                rowK[j] = S[j * lds + k];
                const double v = S[tj * lds + pr];                              // final row k = old row pr, columns swapped
                Uk[j] = (!leftorth && j > k) ? v / piv : v;                     // :307-309
            }
        }
        if (tid == 0) {
            int t = rowperm[k]; rowperm[k] = rowperm[pr]; rowperm[pr] = t;
            t = colperm[k]; colperm[k] = colperm[pc]; colperm[pc] = t;
            pivvals[k] = piv;
        }
        __syncthreads();
        np = k + 1;
        const bool last = np >= maxrank;                                        // :356: swaps and scaling still happen, the update does not
        // ---- one pass: every element gets its final value
        for (int t = tid; t < m + n; t += NT) {
            if (t < m) {
                const int i = t;
                S[k * lds + i] = Lk[i];                                         // column k, all rows
                if (pc != k && i < k) S[pc * lds + i] = colK[i];                // column pc above the pivot row: old column k
            } else {
                const int j = t - m;
                if (j != k) S[j * lds + k] = Uk[j];                             // row k, all other columns
                if (pr != k && j < k) S[j * lds + pr] = rowK[j];                // row pr left of the pivot column: old row k
            }
        }
        bv = -INFINITY; bx = 0.0; bi = -1; bj = -1;
        for (int j = np + warp; j < n; j += NW) {                               // trailing block: :316-320 + the next search
            const double y = Uk[j];
            const int tj = (j == pc) ? k : j;
            for (int i = np + lane; i < m; i += 32) {
                // value at (i, j) after the swaps: old(sigma(i), tau(j)); only row pr and column pc actually moved
                const double a = (i == pr) ? rowK[tj] : ((j == pc) ? colK[i] : S[j * lds + i]);
                if (last) { S[j * lds + i] = a; continue; }
                const double v = __dsub_rn(a, __dmul_rn(Lk[i], y));
                S[j * lds + i] = v;
                const double q = __dmul_rn(v, v);
                TCI_CONSIDER(q, v, i, j)
            }
        }
        if (last) break;
    }
    __syncthreads();
    for (int j = warp; j < n; j += NW)
        for (int i = lane; i < m; i += 32) A[(long long)j * lda + i] = S[j * lds + i];
    for (int t = tid; t < m; t += NT) g_rowperm[t] = rowperm[t];
    for (int t = tid; t < n; t += NT) g_colperm[t] = colperm[t];
    if (tid == 0) { st->npivot = np; st->stop = 1; st->error = error; st->maxerror = maxerror; }
    if (extract) {
        const int r = np;
        int flags = 0;
        for (int j = warp; j < r; j += NW)
            for (int i = lane; i < m; i += 32) {
                double v = (i >= j) ? S[j * lds + i] : 0.0;
                if (isnan(v)) flags |= 1;
                if (i == j && leftorth) v = 1.0;
                if (L) L[(long long)j * ldl + i] = v;
            }
        for (int j = warp; j < n; j += NW)
            for (int i = lane; i < r; i += 32) {
                double v = (j >= i) ? S[j * lds + i] : 0.0;
                if (isnan(v)) flags |= 2;
                if (i == j && !leftorth) v = 1.0;
                if (U) U[(long long)j * ldu + i] = v;
            }
        if (flags) atomicOr(&st->flags, flags);
    }
}

// ---- persistent variant: the whole loop of _optimizerrlu! (:346-369) in ONE cooperative launch ---------------------
// Per pivot step: [every block reduces the per-block candidates and evaluates the stop rule redundantly] ->
// swap/scale phase spread over the grid -> grid barrier -> update + next search -> grid barrier.  No launch gaps,
// no host round trip; the same scan_tiles/swap_item code as the step kernels, with L1-bypassing loads because
// blocks read what other blocks wrote earlier in the same launch.
__device__ __forceinline__ bool grid_barrier(unsigned int* bar, unsigned int target, int* s_abort) {
    __syncthreads();
    if (threadIdx.x == 0) {
        // release-add without a return value (no round trip), then poll with acquire loads
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" :: "l"(bar) : "memory");
        unsigned int v;
        int ab = 0, spins = 0;
        do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(bar) : "memory");
            if (++spins > (1 << 23)) { ab = 1; break; }              // a lost block must not hang the GPU
        } while (v < target);
        *s_abort = ab;
    }
    __syncthreads();
    return *s_abort == 0;
}

template <bool VEC, int COLS>
__global__ void __launch_bounds__(kThreads, (COLS <= 4 ? 5 : (COLS == 8 ? 3 : 2)))
rrlu_persistent_kernel(double* __restrict__ A, int m, int n, long long lda, double* __restrict__ ybuf, RrluState* st,
                       Cand* parts, int* rowperm, int* colperm, double* pivvals, int maxrank, double reltol, double abstol,
                       int leftorth, int serpentine, unsigned int* bar) {
    __shared__ Cand s_c[kThreads / 32];
    __shared__ Cand s_best;
    __shared__ int s_abort;
    unsigned int phase = 0;
    int np = 0;
    double maxerror = 0.0, error = nan("");
    double bv = -INFINITY, bx = 0.0;
    int bi = -1, bj = -1;
    scan_tiles<false, VEC, COLS, true>(A, m, n, lda, ybuf, A, 0, 0, 0, serpentine, ColOrder{nullptr, nullptr}, bv, bx, bi, bj);
    long long tp[6] = {0, 0, 0, 0, 0, 0}, t0 = clock64(), t1;
#define TCI_TICK(slot) { t1 = clock64(); tp[slot] += t1 - t0; t0 = t1; }
    for (;;) {
        Cand c = block_reduce(Cand{bx, bi, bj}, s_c);
        if (threadIdx.x == 0) parts[blockIdx.x] = c;
        TCI_TICK(0)
        phase += gridDim.x;
        if (!grid_barrier(bar, phase, &s_abort)) { if (threadIdx.x == 0) atomicOr(&st->flags, 4); return; }
        TCI_TICK(1)
        Cand best{0.0, -1, -1};
        for (int b = threadIdx.x; b < (int)gridDim.x; b += kThreads) {
            const double2 raw = __ldcg(reinterpret_cast<const double2*>(parts + b));
            Cand o; o.x = raw.x; o.i = (int)(__double_as_longlong(raw.y) & 0xffffffffLL); o.j = (int)(__double_as_longlong(raw.y) >> 32);
            if (cand_better(o, best)) best = o;
        }
        best = block_reduce(best, s_c);
        if (threadIdx.x == 0) s_best = best;
        __syncthreads();
        best = s_best;
        if (best.i < 0) { best.i = np; best.j = np; best.x = nan(""); }       // every entry NaN: first(rows), first(cols)
        const double err = fabs(best.x);                                        // :360
        error = err;
        if ((err < reltol * maxerror || err < abstol) && np > 0) break;         // :363
        maxerror = jl_max(maxerror, err);                                       // :367
        TCI_TICK(2)
        const int k = np;
        for (int t = blockIdx.x * kThreads + threadIdx.x; t <= n + m; t += gridDim.x * kThreads)
            swap_item<true>(t, A, m, n, lda, ybuf, k, best.i, best.j, best.x, rowperm, colperm, pivvals, leftorth);
        TCI_TICK(3)
        phase += gridDim.x;
        if (!grid_barrier(bar, phase, &s_abort)) { if (threadIdx.x == 0) atomicOr(&st->flags, 4); return; }
        TCI_TICK(4)
        np = k + 1;
        if (np >= maxrank) break;                                               // while lu.npivot < maxrank   (:356)
        bv = -INFINITY; bx = 0.0; bi = -1; bj = -1;
        scan_tiles<true, VEC, COLS, true>(A, m, n, lda, ybuf, A + (long long)k * lda, np, np, np, serpentine, ColOrder{nullptr, nullptr}, bv, bx, bi, bj);
        TCI_TICK(5)
    }
#undef TCI_TICK
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        st->npivot = np; st->stop = 1; st->error = error; st->maxerror = maxerror;
        for (int i = 0; i < 6; ++i) st->tphase[i] = tp[i];
    }
}

// L = tril(A[:, 1:r]), U = triu(A[1:r, :]), NaN flags, unit diagonal (:372-388).
__global__ void rrlu_extract_kernel(const double* __restrict__ A, int m, int n, long long lda, int r_in, int leftorth,
                                    double* __restrict__ L, long long ldl, double* __restrict__ U, long long ldu,
                                    RrluState* st) {
    const int r = r_in >= 0 ? r_in : st->npivot;          // r_in < 0: the host has not read the pivot count back yet
    const long long nL = (long long)m * r, nU = (long long)r * n;
    int flags = 0;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < nL + nU; t += (long long)gridDim.x * blockDim.x) {
        if (t < nL) {
            const int i = (int)(t % m), j = (int)(t / m);
            double v = (i >= j) ? A[(long long)j * lda + i] : 0.0;
            if (isnan(v)) flags |= 1;
            if (i == j && leftorth) v = 1.0;
            if (L) L[(long long)j * ldl + i] = v;
        } else {
            const long long u = t - nL;
            const int i = (int)(u % r), j = (int)(u / r);
            double v = (j >= i) ? A[(long long)j * lda + i] : 0.0;
            if (isnan(v)) flags |= 2;
            if (i == j && !leftorth) v = 1.0;
            if (U) U[(long long)j * ldu + i] = v;
        }
    }
    if (flags) atomicOr(&st->flags, flags);
}

struct SmallWs {                 // carve-up of tci_context::ws_small
    RrluState* st; int* rowperm; int* colperm; double* ybuf; double* pivvals; Cand* parts; unsigned int* bar;
};

int carve(tci_handle h, int m, int n, int grid, SmallWs* w) {
    auto al = [](size_t x) { return (x + 255) & ~(size_t)255; };
    const size_t o_st = 0, o_bar = al(sizeof(RrluState)), o_rp = o_bar + 256, o_cp = o_rp + al(sizeof(int) * m),
                 o_y = o_cp + al(sizeof(int) * n), o_pv = o_y + al(sizeof(double) * n),
                 o_parts = o_pv + al(sizeof(double) * (size_t)(m < n ? m : n)), total = o_parts + al(sizeof(Cand) * grid);
    int rc = tci_ws_reserve(h, h->ws_small, total);
    if (rc) return rc;
    char* b = (char*)h->ws_small.ptr;
    w->st = (RrluState*)(b + o_st); w->bar = (unsigned int*)(b + o_bar); w->rowperm = (int*)(b + o_rp); w->colperm = (int*)(b + o_cp);
    w->ybuf = (double*)(b + o_y); w->pivvals = (double*)(b + o_pv); w->parts = (Cand*)(b + o_parts);
    return TCI_OK;
}

struct UpdateCfg { int cols; int grid; int serpentine; };

template <bool UPDATE, bool VEC, int COLS>
int occupancy_grid(tci_handle h) {
    int nb = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, rrlu_update_kernel<UPDATE, VEC, COLS>, kThreads, 0) != cudaSuccess || nb < 1) nb = 1;
    return nb * h->num_sms;
}

// Tunables can be overridden from the environment for experiments (TCI_RRLU_COLS = 4|8|16,
// TCI_RRLU_BLOCKS_PER_SM, TCI_RRLU_SERPENTINE = 0|1).
UpdateCfg update_cfg(tci_handle h, bool vec) {
    UpdateCfg c; c.cols = 8; c.serpentine = 1;
    if (const char* e = getenv("TCI_RRLU_COLS")) { int v = atoi(e); if (v == 4 || v == 8 || v == 16) c.cols = v; }
    if (const char* e = getenv("TCI_RRLU_SERPENTINE")) c.serpentine = atoi(e);
    if (vec) c.grid = c.cols == 4 ? occupancy_grid<true, true, 4>(h) : c.cols == 8 ? occupancy_grid<true, true, 8>(h) : occupancy_grid<true, true, 16>(h);
    else     c.grid = occupancy_grid<true, false, 8>(h);
    if (const char* e = getenv("TCI_RRLU_BLOCKS_PER_SM")) { int v = atoi(e); if (v > 0 && v * h->num_sms < c.grid) c.grid = v * h->num_sms; }
    return c;
}

// Launch with programmatic stream serialization (PDL): the kernel may be scheduled while its predecessor drains; both
// kernels call cudaGridDependencySynchronize() before they read anything the predecessor wrote.
template <typename... KArgs, typename... Args>
cudaError_t launch_pdl(tci_handle h, void (*kern)(KArgs...), unsigned grid, unsigned block, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(block); cfg.dynamicSmemBytes = 0; cfg.stream = h->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = h->rrlu_pdl ? 1 : 0;
    cfg.attrs = attr; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

// Same, plus an L2 access-policy window: accesses inside [win, win + winbytes) are cached as PERSISTING, everything else as
// streaming.  The step kernels put the window on the part of the trailing block that the current pass touches LAST and the
// next (reversed) pass touches FIRST, so those lines are still in L2 — dirty — when they are read and rewritten again.
template <typename... KArgs, typename... Args>
cudaError_t launch_pdl_win(tci_handle h, void (*kern)(KArgs...), unsigned grid, unsigned block, void* win, size_t winbytes, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(block); cfg.dynamicSmemBytes = 0; cfg.stream = h->stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = h->rrlu_pdl ? 1 : 0;
    int na = 1;
    if (win && winbytes) {
        attr[1].id = cudaLaunchAttributeAccessPolicyWindow;
        attr[1].val.accessPolicyWindow.base_ptr = win;
        attr[1].val.accessPolicyWindow.num_bytes = winbytes;
        attr[1].val.accessPolicyWindow.hitRatio = 1.0f;
        attr[1].val.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
        attr[1].val.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
        na = 2;
    }
    cfg.attrs = attr; cfg.numAttrs = na;
    return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

template <bool UPDATE>
void launch_update(tci_handle h, bool vec, const UpdateCfg& c, double* A, int m, int n, long long lda, const SmallWs& w,
                   int maxrank, double reltol, double abstol, int start0 = 0, void* win = nullptr, size_t winbytes = 0) {
#define TCI_LAUNCH_UPD(V, CC) launch_pdl_win(h, rrlu_update_kernel<UPDATE, V, CC>, (unsigned)c.grid, (unsigned)kThreads, win, winbytes, A, m, n, lda, (const double*)w.ybuf, w.st, w.parts, maxrank, reltol, abstol, start0, c.serpentine)
    if (!vec) TCI_LAUNCH_UPD(false, 8);
    else if (c.cols == 4) TCI_LAUNCH_UPD(true, 4);
    else if (c.cols == 16) TCI_LAUNCH_UPD(true, 16);
    else TCI_LAUNCH_UPD(true, 8);
#undef TCI_LAUNCH_UPD
}

template <bool VEC, int COLS>
int launch_persistent_t(tci_handle h, double* A, int m, int n, long long lda, const SmallWs& w, int maxrank, double reltol,
                        double abstol, int leftorth, int serpentine) {
    int nb = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, rrlu_persistent_kernel<VEC, COLS>, kThreads, 0) != cudaSuccess || nb < 1)
        return tci_fail(h, TCI_ERR_CUDA, "rrlu: occupancy query failed for the persistent kernel");
    long long grid = (long long)nb * h->num_sms;
    // small matrices: no more blocks than there are tiles / swap items at the first step (cheaper grid barrier)
    const long long rows_per_tile = VEC ? 2 * kThreads : kThreads;
    long long want = ((m + rows_per_tile - 1) / rows_per_tile) * (((long long)n + COLS - 1) / COLS);
    const long long swapb = ((long long)m + n + 1 + kThreads - 1) / kThreads;
    if (want < swapb) want = swapb;
    if (grid > want) grid = want;
    if (const char* e = getenv("TCI_RRLU_PGRID")) { long long v = atoll(e); if (v > 0 && v < grid) grid = v; }
    double* ybuf = w.ybuf; RrluState* st = w.st; Cand* parts = w.parts; int* rp = w.rowperm; int* cp = w.colperm;
    double* pv = w.pivvals; unsigned int* bar = w.bar;
    void* args[] = {&A, &m, &n, &lda, &ybuf, &st, &parts, &rp, &cp, &pv, &maxrank, &reltol, &abstol, &leftorth, &serpentine, &bar};
    cudaError_t e = cudaLaunchCooperativeKernel((const void*)rrlu_persistent_kernel<VEC, COLS>, dim3((unsigned)grid), dim3(kThreads), args, 0, h->stream);
    h->launches++;
    if (e != cudaSuccess) return tci_fail(h, TCI_ERR_CUDA, "cudaLaunchCooperativeKernel(rrlu): %s", cudaGetErrorString(e));
    return TCI_OK;
}

int launch_persistent(tci_handle h, bool vec, const UpdateCfg& c, double* A, int m, int n, long long lda, const SmallWs& w,
                      int maxrank, double reltol, double abstol, int leftorth) {
    if (!vec) return launch_persistent_t<false, 8>(h, A, m, n, lda, w, maxrank, reltol, abstol, leftorth, c.serpentine);
    if (c.cols == 4) return launch_persistent_t<true, 4>(h, A, m, n, lda, w, maxrank, reltol, abstol, leftorth, c.serpentine);
    if (c.cols == 16) return launch_persistent_t<true, 16>(h, A, m, n, lda, w, maxrank, reltol, abstol, leftorth, c.serpentine);
    return launch_persistent_t<true, 8>(h, A, m, n, lda, w, maxrank, reltol, abstol, leftorth, c.serpentine);
}


// ---- column-sharded factorisation over several GPUs (SURVEY.md 8e, row "rrLU") ----------------------------------------
// Rank g owns nloc columns of the m x n matrix (all rows).  Per pivot step every rank
//   1. reads the G candidate records of the previous step (its own + the peers', gathered by the host layer's collective or
//      written by the peers into its mailbox), picks the global pivot by the reference's order (largest abs2, then smallest
//      column POSITION in the swapped layout, then smallest row) and applies the stop rule — identically on every rank;
//   2. swaps rows k <-> pr in its columns (and in the L columns gathered so far); the owner of the pivot column also moves
//      that column to the front of its local trailing block (a LOCAL swap: columns never cross ranks — the reference's
//      global column order lives in the pos2col/col2pos maps that every rank updates identically);
//   3. applies the rank-1 update to its local trailing columns with the broadcast pivot column (part of the winner's
//      record) and its own piece of the pivot row, searching the updated values in the same pass; the last block to finish
//      writes this rank's next record: best candidate + that candidate's whole column.
// Arithmetic per element is identical to the single-GPU kernels (same scan_tiles), so the factors are bit-identical.
struct DistState {
    int npivot, stop, pr, owner, pcl, posw;   // pending pivot: row, owning rank, local column there, its position
    int kl, kl_next;                          // local columns already pivoted (the local trailing block starts here)
    int flags;
    unsigned int ticket, ticket2;             // last-block detection of the update kernel / of the owner's swap kernel
    double pivval, maxerror, error;
    double maxerr2[2];                        // mailbox mode: running maximum by step parity (every block reads [k & 1], block 0 writes the other)
};
struct Cand4 { double x; int i, pos, jl; };

__device__ __forceinline__ bool cand4_better(const Cand4& a, const Cand4& b) {
    if (a.i < 0) return false;
    if (b.i < 0) return true;
    const double va = __dmul_rn(a.x, a.x), vb = __dmul_rn(b.x, b.x);
    if (va > vb) return true;
    if (va == vb) return (a.pos < b.pos) || (a.pos == b.pos && a.i < b.i);
    return false;
}
__device__ __forceinline__ Cand4 cand4_shfl_down(const Cand4& c, int d) {
    Cand4 o;
    o.x = __shfl_down_sync(0xffffffffu, c.x, d); o.i = __shfl_down_sync(0xffffffffu, c.i, d);
    o.pos = __shfl_down_sync(0xffffffffu, c.pos, d); o.jl = __shfl_down_sync(0xffffffffu, c.jl, d);
    return o;
}
__device__ __forceinline__ Cand4 block_reduce4(Cand4 c, Cand4* smem) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) { Cand4 o = cand4_shfl_down(c, d); if (cand4_better(o, c)) c = o; }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) smem[warp] = c;
    __syncthreads();
    if (warp == 0) {
        c = (lane < kThreads / 32) ? smem[lane] : Cand4{0.0, -1, -1, -1};
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) { Cand4 o = cand4_shfl_down(c, d); if (cand4_better(o, c)) c = o; }
    }
    return c;   // valid in thread 0
}

// Mailbox exchange over peer-mapped memory (CUDA IPC / NVLink), two rounds per pivot step, with a self-validating packet
// format instead of flags: every 16-byte packet is two 8-byte words {32 data bits | 32-bit sequence number k + 1}, so a
// double travels as one packet and a receiver that sees the expected sequence number in both words has the data — no
// release/acquire fences, no flag round trip, nothing to order (8-byte stores are single transactions on NVLink, and each
// word validates itself).  Rank g's mailbox holds, for each parity of the step counter, one 3-packet candidate HEADER per
// rank (value | row, position | local column) and one pivot COLUMN of m packets.  Round 1: the last block of a rank's
// update kernel stores its header into slot [parity][rank] of EVERY rank's mailbox; every block of the next swap kernel
// polls the G headers of its own mailbox and picks the winner.  Round 2: the winner's owner stores the pivot column into
// every peer's mailbox from the (grid-wide) swap kernel — the threads that swap that column locally hold its entries
// anyway; the peers' builder threads poll their own entry.  Only one column per step crosses NVLink, written by a whole
// grid.  A rank can be at most one step ahead of a peer, so two parities suffice; the mailbox starts zeroed (sequence 0
// never matches).
//   layout (16-byte packets): hdr[2][n][3] | col[2][mpad];  mpad = reclen - 4
constexpr int kMaxPeers = 8;
struct PeerBox { double* box[kMaxPeers]; int n; };       // n == 0: records go to `record` and are gathered by the caller
struct alignas(16) Packet { unsigned long long w0, w1; };

__device__ __forceinline__ Packet* box_hdr(double* box, int n, int parity, int g) {
    return reinterpret_cast<Packet*>(box) + ((long long)parity * n + g) * 3;
}
__device__ __forceinline__ Packet* box_col(double* box, long long reclen, int n, int parity) {
    return reinterpret_cast<Packet*>(box) + 6LL * n + (long long)parity * (reclen - 4);
}
__device__ __forceinline__ void packet_store(Packet* dst, unsigned int lo, unsigned int hi, unsigned int seq) {
    const unsigned long long w0 = (unsigned long long)lo | ((unsigned long long)seq << 32);
    const unsigned long long w1 = (unsigned long long)hi | ((unsigned long long)seq << 32);
    asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(dst), "l"(w0), "l"(w1) : "memory");
}
__device__ __forceinline__ void packet_store_f64(Packet* dst, double v, unsigned int seq) {
    const unsigned long long b = (unsigned long long)__double_as_longlong(v);
    packet_store(dst, (unsigned int)(b & 0xffffffffULL), (unsigned int)(b >> 32), seq);
}
__device__ __forceinline__ bool packet_try_load(const Packet* src, unsigned int seq, unsigned int& lo, unsigned int& hi) {
    unsigned long long w0, w1;
    asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(w0), "=l"(w1) : "l"(src) : "memory");
    lo = (unsigned int)(w0 & 0xffffffffULL); hi = (unsigned int)(w1 & 0xffffffffULL);
    return (unsigned int)(w0 >> 32) == seq && (unsigned int)(w1 >> 32) == seq;
}
__device__ __forceinline__ unsigned long long globaltimer_ns();
// spin until the packet carries `seq` (false after 4 s: the sender died)
__device__ __forceinline__ bool packet_wait(const Packet* src, unsigned int seq, unsigned int& lo, unsigned int& hi) {
    if (packet_try_load(src, seq, lo, hi)) return true;
    const unsigned long long t0 = globaltimer_ns();
    for (;;) {
        if (packet_try_load(src, seq, lo, hi)) return true;
        if (globaltimer_ns() - t0 > 4000000000ULL) return false;
        __nanosleep(50);
    }
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

struct DistArrays {
    DistState* st; int* rowperm; int* pos2col; int* col2pos; int* gid;
    double* ybuf; double* xcol; double* pivvals; double* Lbuf; Cand4* parts;
    long long ldl;
    unsigned long long* trace;               // TCI_DIST_TRACE=1: 8 %globaltimer stamps per step (tci_rrlu_dist_trace), else NULL
};
#define TCI_TRACE(a, step, slot) do { if ((a).trace) (a).trace[(long long)(step) * 8 + (slot)] = globaltimer_ns(); } while (0)

__global__ void rrlu_dist_init_kernel(DistArrays a, int m, int n, int nloc, int col0) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < m) a.rowperm[t] = t;
    if (t < n) { a.pos2col[t] = t; a.col2pos[t] = t; }
    if (t < nloc) a.gid[t] = col0 + t;
    if (t == 0) {
        DistState* st = a.st;
        st->npivot = 0; st->stop = 0; st->pr = 0; st->owner = 0; st->pcl = 0; st->posw = 0; st->kl = 0; st->kl_next = 0;
        st->flags = 0; st->ticket = 0u; st->ticket2 = 0u; st->pivval = 0.0; st->maxerror = 0.0; st->error = nan("");
        st->maxerr2[0] = 0.0; st->maxerr2[1] = 0.0;
    }
}

// Step part 1 (one block): global pivot choice + stop rule (matrixlu.jl:359-367), pivot column broadcast -> xcol / Lbuf.
__global__ void __launch_bounds__(kThreads)
rrlu_dist_select_kernel(DistArrays a, const double* __restrict__ gathered, long long reclen, int nranks, int m,
                        int maxrank, double reltol, double abstol, int leftorth) {
    __shared__ int s_commit;
    cudaGridDependencySynchronize();
    DistState* st = a.st;
    const int k = st->npivot;                               // written only by the update kernel
    __shared__ double s_hdr[kMaxPeers * 2][4];              // the G record headers, fetched in parallel
    if ((int)threadIdx.x < 4 * nranks && nranks <= kMaxPeers * 2)
        s_hdr[threadIdx.x >> 2][threadIdx.x & 3] = __ldcg(gathered + (long long)(threadIdx.x >> 2) * reclen + (threadIdx.x & 3));
    __syncthreads();
    if (threadIdx.x == 0) {
        s_commit = 0;
        if (st->stop) {
            // the loop has ended on an earlier step: nothing to do
        } else if (st->flags & 4) {
            st->stop = 1;                                           // exchange timed out
        } else if (k >= maxrank) {
            st->stop = 1;                                           // while lu.npivot < maxrank (:356)
        } else {
            Cand4 best{0.0, -1, -1, -1}; int owner = -1;
            for (int g = 0; g < nranks; ++g) {
                double hdr[4];
                if (nranks <= kMaxPeers * 2) { hdr[0] = s_hdr[g][0]; hdr[1] = s_hdr[g][1]; hdr[2] = s_hdr[g][2]; hdr[3] = s_hdr[g][3]; }
                else { const double* r = gathered + (long long)g * reclen; hdr[0] = __ldcg(r); hdr[1] = __ldcg(r + 1); hdr[2] = __ldcg(r + 2); hdr[3] = __ldcg(r + 3); }
                Cand4 c{hdr[0], (int)hdr[1], (int)hdr[2], (int)hdr[3]};
                if (cand4_better(c, best)) { best = c; owner = g; }
            }
            if (best.i < 0) {                                       // every remaining entry is NaN: not supported in the sharded path
                st->flags |= 8; st->stop = 1; st->error = nan("");
            } else {
                const double err = fabs(best.x);                    // :360
                st->error = err;
                if ((err < reltol * st->maxerror || err < abstol) && k > 0) {    // :363
                    st->stop = 1;
                } else {
                    st->maxerror = jl_max(st->maxerror, err);       // :367
                    st->pr = best.i; st->owner = owner; st->pcl = best.jl; st->posw = best.pos; st->pivval = best.x;
                    s_commit = 1;
                }
            }
        }
    }
    __syncthreads();
    if (!s_commit) return;
    if (threadIdx.x == 0) {
        const int pr = st->pr;
        int t = a.rowperm[k]; a.rowperm[k] = a.rowperm[pr]; a.rowperm[pr] = t;
        const int P = st->posw, ck = a.pos2col[k], cw = a.pos2col[P];                 // swapcol!(k, P) on the position maps
        a.pos2col[k] = cw; a.pos2col[P] = ck; a.col2pos[cw] = k; a.col2pos[ck] = P;
        a.pivvals[k] = st->pivval;
    }
}

// Step part 2: the row swap in the local columns and in the L columns gathered so far; on the owner also the local column
// swap that brings the pivot column to the front of the trailing block; packed copy of the local piece of the pivot row;
// the pivot column with the row swap applied -> xcol and column k of L.  Mailbox mode (peers.n > 0): round 2 of the
// exchange — the owner's column-swap threads also store the column into the peers' mailboxes and build xcol / L locally,
// the peers wait for the column flag (one lane per warp) before they build theirs.
__global__ void rrlu_dist_swap_kernel(DistArrays a, double* __restrict__ A, int m, int nloc, long long lda, int rank, int leftorth,
                                      const double* __restrict__ gathered, long long reclen, int nranks, PeerBox peers,
                                      int maxrank, double reltol, double abstol) {
    cudaGridDependencySynchronize();
    DistState* st = a.st;
    if (st->stop) return;
    const bool mail = peers.n > 0;
    const int k = st->npivot, kl = st->kl;
    int pr, pcl, owner;
    double piv;
    if (blockIdx.x == 0 && threadIdx.x == 0) TCI_TRACE(a, k, 0);   // swap kernel running
    if (mail) {
        // round 1, receiving side, fused with the pivot choice: EVERY block waits for the G headers of its own mailbox and
        // takes the decision of the select kernel redundantly (pivot, stop rule :359-367); block 0 alone writes the shared
        // state.  The running maximum is double-buffered by step parity so that no block reads what block 0 writes.
        __shared__ unsigned int s_hw[kMaxPeers][3][2];
        const bool lead = blockIdx.x == 0 && threadIdx.x == 0;
        if (k >= maxrank) { if (lead) st->stop = 1; return; }       // while lu.npivot < maxrank (:356)
        double* mybox = peers.box[rank];
        int to = 0;
        if ((int)threadIdx.x < 3 * nranks) {                        // one thread per header packet
            const int g = threadIdx.x / 3, part = threadIdx.x % 3;
            unsigned int lo, hi;
            if (!packet_wait(box_hdr(mybox, nranks, k & 1, g) + part, (unsigned int)(k + 1), lo, hi)) to = 1;   // a peer died
            s_hw[g][part][0] = lo; s_hw[g][part][1] = hi;
        }
        to = __syncthreads_or(to);
        if (to) { if (threadIdx.x == 0) { atomicOr(&st->flags, 4); st->stop = 1; } return; }
        Cand4 best{0.0, -1, -1, -1};
        owner = -1;
        for (int g = 0; g < nranks; ++g) {
            const double x = __longlong_as_double((long long)((unsigned long long)s_hw[g][0][0] | ((unsigned long long)s_hw[g][0][1] << 32)));
            const Cand4 c{x, (int)s_hw[g][1][0], (int)s_hw[g][1][1], (int)s_hw[g][2][0]};
            if (cand4_better(c, best)) { best = c; owner = g; }
        }
        if (best.i < 0) {                                           // every remaining entry is NaN: not supported in the sharded path
            if (lead) { st->flags |= 8; st->stop = 1; st->error = nan(""); }
            return;
        }
        const double err = fabs(best.x);                            // :360
        const double maxerr = st->maxerr2[k & 1];
        if ((err < reltol * maxerr || err < abstol) && k > 0) {     // :363
            if (lead) { st->error = err; st->stop = 1; }
            return;
        }
        pr = best.i; pcl = best.jl; piv = best.x;
        if (lead) {
            TCI_TRACE(a, k, 1);                                     // headers arrived, pivot chosen
            st->error = err;
            st->maxerr2[(k + 1) & 1] = jl_max(maxerr, err);         // :367
            st->pr = pr; st->owner = owner; st->pcl = pcl; st->posw = best.pos; st->pivval = piv;
            int t0 = a.rowperm[k]; a.rowperm[k] = a.rowperm[pr]; a.rowperm[pr] = t0;
            const int P = best.pos, ck = a.pos2col[k], cw = a.pos2col[P];          // swapcol!(k, P) on the position maps
            a.pos2col[k] = cw; a.pos2col[P] = ck; a.col2pos[cw] = k; a.col2pos[ck] = P;
            a.pivvals[k] = piv;
        }
    } else {
        pr = st->pr; pcl = st->pcl; owner = st->owner; piv = st->pivval;
    }
    const bool mine = owner == rank;
    const int klnext = kl + (mine ? 1 : 0);
    const int par = k & 1;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int b_cols = nloc, b_L = nloc + m, b_corner = nloc + m + k, b_build = nloc + m + k + 1;   // first item of each kind
    const bool builder = t >= b_build && t < b_build + m && !(mail && mine);
    const double* col = mail ? nullptr : gathered + (long long)owner * reclen + 4;
    if (t < b_cols) {
        const int j = t;
        if (!(mine && (j == kl || j == pcl))) {                     // the corner item handles these two columns
            double* pk = A + (long long)j * lda + k;
            double* pp = A + (long long)j * lda + pr;
            const double old_k = *pk;
            double newk = (pr != k) ? *pp : old_k;
            if (j >= klnext) {                                      // trailing column: its row-k entry is part of the pivot row
                if (!leftorth) newk = newk / piv;                   // :306-308
                a.ybuf[j] = newk;
            }
            *pk = newk;
            if (pr != k) *pp = old_k;
        }
    } else if (t < b_L) {
        const int i = t - nloc;
        if (mine && i != k && i != pr) {
            if (mail) {
                double* p1 = A + (long long)pcl * lda + i;
                const double v1 = *p1;                              // entry i of the pivot column
                if (pcl != kl) { double* p0 = A + (long long)kl * lda + i; const double v0 = *p0; *p0 = v1; *p1 = v0; }
                for (int g = 0; g < peers.n; ++g)
                    if (g != rank) packet_store_f64(box_col(peers.box[g], reclen, nranks, par) + i, v1, (unsigned int)(k + 1));
                double v = v1;
                if (leftorth && i > k) v = v / piv;                 // :303-305
                a.xcol[i] = v;
                if (i >= k) a.Lbuf[(long long)k * a.ldl + i] = v;
            } else if (pcl != kl) {
                double* p0 = A + (long long)kl * lda + i;
                double* p1 = A + (long long)pcl * lda + i;
                const double v = *p0; *p0 = *p1; *p1 = v;
            }
        }
    } else if (t < b_corner) {
        if (pr != k) {
            const int c = t - b_L;
            double* p0 = a.Lbuf + (long long)c * a.ldl + k;
            double* p1 = a.Lbuf + (long long)c * a.ldl + pr;
            const double v = *p0; *p0 = *p1; *p1 = v;
        }
    } else if (t == b_corner) {
        if (mine) {
            const double o_kk = A[(long long)kl * lda + k], o_kp = A[(long long)pcl * lda + k];
            const double o_rk = A[(long long)kl * lda + pr], o_rp = A[(long long)pcl * lda + pr];
            double n_kk = o_rp, n_kp = o_rk, n_rk = o_kp, n_rp = o_kk;           // new[r][c] = old[rowswap(r)][colswap(c)]
            if (!leftorth && pcl != kl) n_kp = n_kp / piv;                       // column pcl stays in the trailing block
            // coinciding addresses (pr == k or pcl == kl): the canonical value lands last, as in swap_item
            A[(long long)pcl * lda + pr] = n_rp;
            A[(long long)pcl * lda + k] = n_kp;
            A[(long long)kl * lda + pr] = n_rk;
            A[(long long)kl * lda + k] = n_kk;
            if (pcl != kl) a.ybuf[pcl] = n_kp;
            const int g0 = a.gid[kl]; a.gid[kl] = a.gid[pcl]; a.gid[pcl] = g0;
            if (mail) {                                             // rows k and pr of the pivot column (before the swaps)
                for (int g = 0; g < peers.n; ++g)
                    if (g != rank) {
                        Packet* c = box_col(peers.box[g], reclen, nranks, par);
                        packet_store_f64(c + k, o_kp, (unsigned int)(k + 1));
                        packet_store_f64(c + pr, o_rp, (unsigned int)(k + 1));
                    }
                a.xcol[k] = o_rp;                                   // the pivot itself lands in row k
                a.Lbuf[(long long)k * a.ldl + k] = o_rp;
                if (pr != k) {
                    const double v = leftorth ? o_kp / piv : o_kp;
                    a.xcol[pr] = v;
                    a.Lbuf[(long long)k * a.ldl + pr] = v;
                }
            }
        }
        st->kl_next = klnext;
    } else if (builder) {
        // the broadcast pivot column with the row swap applied -> xcol and column k of L
        const int i = t - b_build;
        const int src = (i == k) ? pr : (i == pr ? k : i);
        double v;
        if (mail) {                                                 // round 2, receiving side: poll the own entry
            unsigned int lo, hi;
            if (!packet_wait(box_col(peers.box[rank], reclen, nranks, par) + src, (unsigned int)(k + 1), lo, hi)) {
                atomicOr(&st->flags, 4); st->stop = 1;              // the owner died
            }
            v = __longlong_as_double((long long)((unsigned long long)lo | ((unsigned long long)hi << 32)));
            if (i == 0) TCI_TRACE(a, k, 2);                         // peer: pivot column arriving
        } else {
            v = __ldcg(col + src);                                  // written by a peer: around L1
        }
        if (leftorth && i > k) v = v / piv;                         // :303-305
        a.xcol[i] = v;
        if (i >= k) a.Lbuf[(long long)k * a.ldl + i] = v;
    }
    if (mail && mine && t == b_corner) TCI_TRACE(a, k, 2);         // owner: corner entries stored
}

// Step part 3: local rank-1 update + search; the last block writes this rank's record for the next step.
template <bool UPDATE, bool VEC, int COLS>
__global__ void __launch_bounds__(kThreads, (COLS == 8 ? 3 : 2))
rrlu_dist_update_kernel(DistArrays a, double* __restrict__ A, int m, int nloc, long long lda, int maxrank, int serpentine,
                        double* __restrict__ record, PeerBox peers, int rank, long long reclen) {
    __shared__ Cand4 s_c[kThreads / 32];
    __shared__ int s_last;
    cudaGridDependencySynchronize();
    DistState* st = a.st;
    if (st->stop) return;
    const int k = st->npivot;
    const int np = UPDATE ? k + 1 : 0;
    const int c0 = UPDATE ? st->kl_next : 0;
    double bv = -INFINITY, bx = 0.0;
    int bi = -1, bj = -1;
    const ColOrder ord{a.gid, a.col2pos};
    if (UPDATE && blockIdx.x == 0 && threadIdx.x == 0) TCI_TRACE(a, k, 3);        // update kernel running
    if (np < maxrank && c0 < nloc)
        scan_tiles<UPDATE, VEC, COLS, false, true>(A, m, nloc, lda, a.ybuf, a.xcol, np, c0, np, serpentine, ord, bv, bx, bi, bj);
    if (UPDATE && blockIdx.x == 0 && threadIdx.x == 0) TCI_TRACE(a, k, 4);        // block 0 done with its tiles
    Cand4 c{bx, bi, bj >= 0 ? a.col2pos[a.gid[bj]] : -1, bj};
    c = block_reduce4(c, s_c);
    if (threadIdx.x == 0) {
        a.parts[blockIdx.x] = c;
        __threadfence();
        const unsigned int done = atomicAdd(&st->ticket, 1u);
        s_last = (done == gridDim.x - 1);
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    if (UPDATE && threadIdx.x == 0) TCI_TRACE(a, k, 5);             // last block: every tile done
    Cand4 best{0.0, -1, -1, -1};
    for (int b = threadIdx.x; b < (int)gridDim.x; b += kThreads) {
        Cand4 o = a.parts[b];
        if (cand4_better(o, best)) best = o;
    }
    __syncthreads();
    best = block_reduce4(best, s_c);
    __shared__ int s_jl;
    if (peers.n == 0) {
        if (threadIdx.x == 0) {
            record[0] = best.x; record[1] = (double)best.i; record[2] = (double)best.pos; record[3] = (double)best.jl;
            s_jl = best.i >= 0 ? best.jl : -1;
            st->npivot = np; st->kl = c0; st->ticket = 0u;
        }
        __syncthreads();
        const int jl = s_jl;
        if (jl >= 0) {
            const double* src = A + (long long)jl * lda;
            for (int i0 = threadIdx.x; i0 < m; i0 += 8 * kThreads) {
                double v[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) { const int i = i0 + u * kThreads; v[u] = i < m ? __ldcg(src + i) : 0.0; }
#pragma unroll
                for (int u = 0; u < 8; ++u) { const int i = i0 + u * kThreads; if (i < m) record[4 + i] = v[u]; }
            }
        }
        return;
    }
    // round 1 of the mailbox exchange: the 4-double header goes straight into every rank's mailbox (peer stores over
    // NVLink), each followed by its own release flag; the pivot column follows in round 2 from the owner's swap kernel
    __shared__ Cand4 s_best;
    if (threadIdx.x == 0) {
        s_best = best;
        st->npivot = np; st->kl = c0; st->ticket = 0u;
    }
    __syncthreads();
    if ((int)threadIdx.x < 3 * peers.n && np < maxrank) {           // one thread per (peer, packet); nobody reads a header after the last step
        const Cand4 bb = s_best;
        const int g = threadIdx.x / 3, part = threadIdx.x % 3;
        Packet* r = box_hdr(peers.box[g], peers.n, np & 1, rank) + part;
        const unsigned long long xb = (unsigned long long)__double_as_longlong(bb.x);
        const unsigned int lo = part == 0 ? (unsigned int)(xb & 0xffffffffULL) : (part == 1 ? (unsigned int)bb.i : (unsigned int)bb.jl);
        const unsigned int hi = part == 0 ? (unsigned int)(xb >> 32) : (part == 1 ? (unsigned int)bb.pos : 0u);
        packet_store(r, lo, hi, (unsigned int)(np + 1));
        if (UPDATE && threadIdx.x == 0) TCI_TRACE(a, k, 6);         // header stored
    }
}

// L (m x r): tril of the gathered columns with the reference's diagonal; U_loc (r x nloc): triu of the local columns in
// terms of their POSITIONS (:372-388); locpos[j] = position of local column j.
__global__ void rrlu_dist_extract_kernel(DistArrays a, const double* __restrict__ A, int m, int nloc, long long lda, int r, int leftorth,
                                         double* __restrict__ L, long long ldl, double* __restrict__ U, long long ldu, int* __restrict__ locpos) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long nl = (long long)m * r, nu = (long long)r * nloc;
    if (t < nl) {
        if (!L) return;
        const int i = (int)(t % m), c = (int)(t / m);
        double v = 0.0;
        if (i > c) v = a.Lbuf[(long long)c * a.ldl + i];
        else if (i == c) v = leftorth ? 1.0 : a.Lbuf[(long long)c * a.ldl + c];
        L[(long long)c * ldl + i] = v;
    } else if (t < nl + nu) {
        if (!U) return;
        const long long e = t - nl;
        const int c = (int)(e % r), j = (int)(e / r);
        const int pos = a.col2pos[a.gid[j]];
        double v = 0.0;
        if (c < pos) v = A[(long long)j * lda + c];
        else if (c == pos) v = leftorth ? A[(long long)j * lda + c] : 1.0;
        U[(long long)j * ldu + c] = v;
    } else if (t < nl + nu + nloc) {
        const int j = (int)(t - nl - nu);
        locpos[j] = a.col2pos[a.gid[j]];
    }
}

} // namespace

// Shared driver: factorises d_A in place, leaves perms/pivot values in ws_small, reads the final state.
static int rrlu_run(tci_handle h, double* d_A, int m, int n, long long lda, int mr, double reltol, double abstol,
                    int leftorth, SmallWs* w, RrluState* final_state,
                    double* d_L = nullptr, long long ldl = 0, double* d_U = nullptr, long long ldu = 0, bool may_extract = false) {
    h->rrlu_extracted = 0;
    const bool vec = (lda % 2 == 0) && (((uintptr_t)d_A & 15u) == 0);
    const UpdateCfg cfg = update_cfg(h, vec);
    const int grid = cfg.grid;
    int rc = carve(h, m, n, grid > h->num_sms * 8 ? grid : h->num_sms * 8, w);
    if (rc) return rc;
    {   // read-back staging: either the packed layout or an image of the workspace up to the pivot values (rrlu_finish)
        const size_t packed = sizeof(RrluState) + sizeof(int) * ((size_t)m + n) + sizeof(double) * (size_t)(mr + 1) + 32;
        const size_t image = (size_t)((char*)w->pivvals - (char*)w->st) + sizeof(double) * (size_t)(mr + 1);
        rc = tci_pinned_reserve(h, packed > image ? packed : image);
        if (rc) return rc;
    }
    const int mx = m > n ? m : n;
    RrluState* hst = (RrluState*)h->pinned;
    // One cooperative launch for the whole loop when the matrix is L2-resident (launch/latency-bound regime);
    // step kernels when it streams from HBM (their launch gaps overlap, a grid barrier's do not).
    bool persistent = (double)m * n * sizeof(double) <= 48e6;
    if (const char* e = getenv("TCI_RRLU_MODE")) persistent = strcmp(e, "persistent") == 0 ? true : (strcmp(e, "steps") == 0 ? false : persistent);
    if (h->rrlu_profile) persistent = false;                // per-launch profiling needs the step kernels
    if (const char* e = getenv("TCI_RRLU_PDL")) h->rrlu_pdl = atoi(e) ? 1 : 0;
    if (const char* e = getenv("TCI_RRLU_LATE_STATE")) h->rrlu_late_state = atoi(e) ? 1 : 0;
    const size_t smem_need = sizeof(double) * (size_t)(m | 1) * n + sizeof(int) * (((size_t)m + n + 1) & ~(size_t)1);   // matrix + permutations
    bool in_smem = persistent && smem_need <= 200 * 1024;
    if (const char* e = getenv("TCI_RRLU_SMEM")) in_smem = in_smem && atoi(e) != 0;
    if (!in_smem) {                                         // the single-CTA kernel initialises its own state
        rrlu_init_kernel<<<(mx + 255) / 256, 256, 0, h->stream>>>(w->st, w->rowperm, w->colperm, m, n, w->bar);
        TCI_LAUNCH_CHECK(h);
    }
    TCI_CUDA(h, cudaEventRecord(h->ev[0], h->stream));
    if (persistent) {
        if (in_smem) {
            if (smem_need > 48 * 1024 && !(h->rrlu_smem_optin & 1)) {
                TCI_CUDA(h, cudaFuncSetAttribute(rrlu_smem_kernel<kSmemThreads>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
                h->rrlu_smem_optin |= 1;
            }
            static const bool tiny_off = getenv("TCI_RRLU_TINY") && atoi(getenv("TCI_RRLU_TINY")) == 0;
            const int extract = may_extract ? 1 : 0;       // L/U straight from shared memory: no separate extraction launch
            // second-generation kernel (two barriers per pivot) when its four extra vectors fit as well
            const size_t smem2 = smem_need + sizeof(double) * 2 * ((size_t)m + n);
            static const bool gen2_off = getenv("TCI_RRLU_SMEM2") && atoi(getenv("TCI_RRLU_SMEM2")) == 0;
            if (!gen2_off && smem2 <= 200 * 1024) {
                const bool tiny = (long long)m * n <= 4096 && smem2 <= 48 * 1024 && !tiny_off;
                if (!tiny && smem2 > 48 * 1024 && !(h->rrlu_smem_optin & 2)) {
                    TCI_CUDA(h, cudaFuncSetAttribute(rrlu_smem2_kernel<kSmemThreads>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
                    h->rrlu_smem_optin |= 2;
                }
                if (tiny)
                    rrlu_smem2_kernel<256><<<1, 256, smem2, h->stream>>>(d_A, m, n, lda, w->st, w->rowperm, w->colperm, w->pivvals, mr, reltol, abstol, leftorth, d_L, ldl, d_U, ldu, extract);
                else
                    rrlu_smem2_kernel<kSmemThreads><<<1, kSmemThreads, smem2, h->stream>>>(d_A, m, n, lda, w->st, w->rowperm, w->colperm, w->pivvals, mr, reltol, abstol, leftorth, d_L, ldl, d_U, ldu, extract);
            } else
            if ((long long)m * n <= 4096 && smem_need <= 48 * 1024 && !tiny_off)
                rrlu_smem_kernel<256><<<1, 256, smem_need, h->stream>>>(d_A, m, n, lda, w->st, w->rowperm, w->colperm, w->pivvals, mr, reltol, abstol, leftorth, d_L, ldl, d_U, ldu, extract);
            else
                rrlu_smem_kernel<kSmemThreads><<<1, kSmemThreads, smem_need, h->stream>>>(d_A, m, n, lda, w->st, w->rowperm, w->colperm, w->pivvals, mr, reltol, abstol, leftorth, d_L, ldl, d_U, ldu, extract);
            TCI_LAUNCH_CHECK(h);
            h->rrlu_extracted = extract;
        } else {
            rc = launch_persistent(h, vec, cfg, d_A, m, n, lda, *w, mr, reltol, abstol, leftorth);
            if (rc) return rc;
        }
        TCI_CUDA(h, cudaEventRecord(h->ev[1], h->stream));
        if (in_smem && h->rrlu_late_state) {
            // Single-CTA kernel: nothing to check before the extraction (no grid barrier that could time out), so the
            // state is read back once, together with the permutations, by rrlu_finish — one host synchronisation per
            // factorisation instead of two (the small-Pi regime is bound by exactly these round trips).
            final_state->npivot = -1; final_state->flags = 0;
            h->rrlu_ms[1] = 0.0;
            return TCI_OK;
        }
        TCI_CUDA(h, cudaMemcpyAsync(hst, w->st, sizeof(RrluState), cudaMemcpyDeviceToHost, h->stream));
        TCI_CUDA(h, cudaStreamSynchronize(h->stream));
        if (hst->flags & 4) return tci_fail(h, TCI_ERR_CUDA, "rrlu: grid barrier timed out in the persistent kernel");
        *final_state = *hst;
        h->rrlu_steps = hst->npivot;
        h->rrlu_ms[1] = 0.0;
        if (getenv("TCI_RRLU_TPHASE")) {
            fprintf(stderr, "rrlu persistent %dx%d steps=%d cycles/step:", m, n, hst->npivot);
            for (int i = 0; i < 6; ++i) fprintf(stderr, " %s=%.0f", (const char*[]){"publish", "barrierA", "reduce", "swap", "barrierB", "scan"}[i], (double)hst->tphase[i] / (hst->npivot ? hst->npivot : 1));
            fprintf(stderr, "\n");
        }
        return TCI_OK;
    }
    // Persisting-L2 window (see launch_pdl_win): sized to the device's persisting set-aside; TCI_RRLU_L2PERSIST = MB (0 off).
    if (h->l2_window_bytes < 0) {
        h->l2_window_bytes = 0;
        long long want_mb = 16;                                // measured best on B200: +1.6 % (8192^2) / +2.9 % (5120^2); 32 MB and more lose
        if (const char* e = getenv("TCI_RRLU_L2PERSIST")) want_mb = atoll(e);
        int maxpersist = 0, maxwin = 0;
        cudaDeviceGetAttribute(&maxpersist, cudaDevAttrMaxPersistingL2CacheSize, h->device);
        cudaDeviceGetAttribute(&maxwin, cudaDevAttrMaxAccessPolicyWindowSize, h->device);
        if (want_mb != 0 && maxpersist > 0 && maxwin > 0) {
            long long bytes = want_mb < 0 ? (long long)maxpersist : want_mb * (1LL << 20);
            if (bytes > maxpersist) bytes = maxpersist;
            if (cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, (size_t)bytes) == cudaSuccess)
                h->l2_window_bytes = bytes < maxwin ? bytes : maxwin;
            if (getenv("TCI_RRLU_L2PERSIST_VERBOSE"))
                fprintf(stderr, "rrlu: persisting L2 max %d B, max window %d B, using %lld B\n", maxpersist, maxwin, h->l2_window_bytes);
        }
        (void)cudaGetLastError();
    }
    const bool use_win = h->l2_window_bytes > 0 && (cfg.serpentine & 7) == 1 && (double)m * n * 8.0 > 2.0 * (double)h->l2_window_bytes;
    auto window_of = [&](int np, void** base, size_t* bytes) {   // the tail of the pass that ends with np pivots in place
        *base = nullptr; *bytes = 0;
        if (!use_win || np >= n) return;
        long long wcols = h->l2_window_bytes / (long long)(lda * 8);
        if (wcols > n - np) wcols = n - np;
        if (wcols <= 0) return;
        const bool rev = (np & 1) != 0;                          // same rule as scan_tiles (serpentine mode 1)
        const long long c_first = rev ? np : n - wcols;          // forward passes end on the highest columns, reversed ones on the lowest
        *base = d_A + c_first * lda;
        *bytes = (size_t)(wcols * lda * 8);
    };
    {
        void* wb; size_t wn; window_of(0, &wb, &wn);
        launch_update<false>(h, vec, cfg, d_A, m, n, lda, *w, mr, reltol, abstol, 0, wb, wn);
    }
    TCI_LAUNCH_CHECK(h);
    const int swap_blocks = (m + n + 1 + 255) / 256;
    int launched = 0, chunk = 8;
    h->rrlu_steps = 0;
    if (h->rrlu_profile)
        while ((int)h->ev_pool.size() < 2 * mr) { cudaEvent_t e; TCI_CUDA(h, cudaEventCreate(&e)); h->ev_pool.push_back(e); }
    while (launched < mr) {
        const int nb = (mr - launched < chunk) ? mr - launched : chunk;
        for (int s = 0; s < nb; ++s) {
            launch_pdl(h, rrlu_swap_kernel, (unsigned)swap_blocks, 256u, d_A, m, n, lda, w->ybuf, w->st, w->rowperm, w->colperm, w->pivvals, leftorth);
            TCI_LAUNCH_CHECK(h);
            if (h->rrlu_profile) TCI_CUDA(h, cudaEventRecord(h->ev_pool[2 * (launched + s)], h->stream));
            {
                void* wb; size_t wn; window_of(launched + s + 1, &wb, &wn);
                launch_update<true>(h, vec, cfg, d_A, m, n, lda, *w, mr, reltol, abstol, 0, wb, wn);
            }
            TCI_LAUNCH_CHECK(h);
            if (h->rrlu_profile) TCI_CUDA(h, cudaEventRecord(h->ev_pool[2 * (launched + s) + 1], h->stream));
        }
        launched += nb;
        if (chunk < 64) chunk *= 2;
        TCI_CUDA(h, cudaMemcpyAsync(hst, w->st, sizeof(RrluState), cudaMemcpyDeviceToHost, h->stream));
        TCI_CUDA(h, cudaStreamSynchronize(h->stream));
        if (hst->stop) break;
    }
    if (launched >= mr && !hst->stop) {        // cannot happen: the last update's tail always stops at maxrank
        return tci_fail(h, TCI_ERR_CUDA, "rrlu: device loop did not terminate (npivot=%d, maxrank=%d)", hst->npivot, mr);
    }
    TCI_CUDA(h, cudaEventRecord(h->ev[1], h->stream));
    if (use_win) (void)cudaCtxResetPersistingL2Cache();    // hand the set-aside lines back
    *final_state = *hst;
    h->rrlu_steps = hst->npivot;
    h->rrlu_ms[1] = 0.0;
    if (h->rrlu_profile) {
        TCI_CUDA(h, cudaStreamSynchronize(h->stream));
        double tot = 0.0;
        h->rrlu_launch_ms.assign(launched, 0.f);
        for (int s = 0; s < launched; ++s) { float ms = 0.f; cudaEventElapsedTime(&ms, h->ev_pool[2 * s], h->ev_pool[2 * s + 1]); tot += ms; h->rrlu_launch_ms[s] = ms; }
        h->rrlu_ms[1] = tot;
    }
    return TCI_OK;
}

static int rrlu_finish(tci_handle h, const double* d_A, int m, int n, long long lda, int leftorth, const SmallWs& w,
                       RrluState* fs, int mr, int64_t* h_rowperm, int64_t* h_colperm, int64_t* h_npivot, double* h_error,
                       double* d_L, long long ldl, double* d_U, long long ldu, double* h_pivoterrors) {
    const bool late = fs->npivot < 0;                       // rrlu_run left the state on the device (see there)
    int r = late ? mr : fs->npivot;                         // late: an upper bound, the kernel reads the count itself
    if (r > 0 && !h->rrlu_extracted) {
        const long long work = (long long)m * r + (long long)r * n;
        long long blocks = (work + 255) / 256;
        if (blocks > h->num_sms * 8) blocks = h->num_sms * 8;
        rrlu_extract_kernel<<<(int)blocks, 256, 0, h->stream>>>(d_A, m, n, lda, late ? -1 : r, leftorth, d_L, ldl, d_U, ldu, w.st);
        TCI_LAUNCH_CHECK(h);
    }
    TCI_CUDA(h, cudaEventRecord(h->ev[2], h->stream));
    char* pin = (char*)h->pinned;
    RrluState* hst = (RrluState*)pin;
    int* hrp = (int*)(pin + sizeof(RrluState));
    int* hcp = hrp + m;
    double* hpv = (double*)(hcp + n + ((m + n) & 1));      // keep 8-byte alignment
    const size_t image = (size_t)((const char*)w.pivvals - (const char*)w.st) + sizeof(double) * (size_t)(r > 0 ? r : 0);
    if (image <= 64 * 1024) {
        // small problem: state, permutations and pivot values are one contiguous stretch of the workspace (carve) —
        // one copy instead of four
        TCI_CUDA(h, cudaMemcpyAsync(pin, w.st, image, cudaMemcpyDeviceToHost, h->stream));
        hrp = (int*)(pin + ((const char*)w.rowperm - (const char*)w.st));
        hcp = (int*)(pin + ((const char*)w.colperm - (const char*)w.st));
        hpv = (double*)(pin + ((const char*)w.pivvals - (const char*)w.st));
    } else {
        TCI_CUDA(h, cudaMemcpyAsync(hst, w.st, sizeof(RrluState), cudaMemcpyDeviceToHost, h->stream));
        TCI_CUDA(h, cudaMemcpyAsync(hrp, w.rowperm, sizeof(int) * m, cudaMemcpyDeviceToHost, h->stream));
        TCI_CUDA(h, cudaMemcpyAsync(hcp, w.colperm, sizeof(int) * n, cudaMemcpyDeviceToHost, h->stream));
        if (r > 0) TCI_CUDA(h, cudaMemcpyAsync(hpv, w.pivvals, sizeof(double) * r, cudaMemcpyDeviceToHost, h->stream));
    }
    TCI_CUDA(h, cudaStreamSynchronize(h->stream));
    if (late) { r = hst->npivot; *fs = *hst; h->rrlu_steps = r; }
    float ms = 0.f;
    cudaEventElapsedTime(&ms, h->ev[0], h->ev[1]); h->rrlu_ms[0] = ms;
    cudaEventElapsedTime(&ms, h->ev[1], h->ev[2]); h->rrlu_ms[2] = ms;
    double err = hst->error;
    if (r >= (m < n ? m : n)) err = 0.0;                    // matrixlu.jl:391-393
    if (h_rowperm) for (int i = 0; i < m; ++i) h_rowperm[i] = (int64_t)hrp[i] + 1;
    if (h_colperm) for (int j = 0; j < n; ++j) h_colperm[j] = (int64_t)hcp[j] + 1;
    if (h_npivot) *h_npivot = r;
    if (h_error) *h_error = err;
    if (h_pivoterrors) { for (int i = 0; i < r; ++i) h_pivoterrors[i] = fabs(hpv[i]); h_pivoterrors[r] = err; }
    fs->flags = hst->flags;
    if (hst->flags & 1) return tci_fail(h, TCI_ERR_NAN_L, "lu.L contains NaNs");
    if (hst->flags & 2) return tci_fail(h, TCI_ERR_NAN_U, "lu.U contains NaNs");
    return TCI_OK;
}

static int check_dims(tci_handle h, int64_t m, int64_t n, int64_t lda, const char* who) {
    if (!h) return TCI_ERR_INVALID;
    if (m < 0 || n < 0 || m > INT32_MAX - 1024 || n > INT32_MAX - 1024) return tci_fail(h, TCI_ERR_INVALID, "%s: bad matrix size %lld x %lld", who, (long long)m, (long long)n);
    if (lda < (m > 1 ? m : 1)) return tci_fail(h, TCI_ERR_INVALID, "%s: lda (%lld) < m (%lld)", who, (long long)lda, (long long)m);
    return TCI_OK;
}

extern "C" {

int tci_rrlu_inplace_f64(tci_handle h, double* d_A, int64_t m, int64_t n, int64_t lda, int64_t maxrank, double reltol,
                         double abstol, int leftorthogonal, int64_t* h_rowperm, int64_t* h_colperm, int64_t* h_npivot,
                         double* h_error, double* d_L, int64_t ldl, double* d_U, int64_t ldu, double* h_pivoterrors) {
    int rc = check_dims(h, m, n, lda, "tci_rrlu_inplace_f64");
    if (rc) return rc;
    TCI_CUDA(h, cudaSetDevice(h->device));
    int64_t mn = m < n ? m : n;
    int64_t mr = (maxrank <= 0 || maxrank > mn) ? mn : maxrank;
    if (mn == 0) {                                          // empty matrix: zero pivots, error 0 (:353, :391)
        if (h_rowperm) for (int64_t i = 0; i < m; ++i) h_rowperm[i] = i + 1;
        if (h_colperm) for (int64_t j = 0; j < n; ++j) h_colperm[j] = j + 1;
        if (h_npivot) *h_npivot = 0;
        if (h_error) *h_error = 0.0;
        if (h_pivoterrors) h_pivoterrors[0] = 0.0;
        return TCI_OK;
    }
    if (!d_A) return tci_fail(h, TCI_ERR_INVALID, "tci_rrlu_inplace_f64: d_A is NULL");
    if (d_L && ldl < m) return tci_fail(h, TCI_ERR_INVALID, "tci_rrlu_inplace_f64: ldl < m");
    if (d_U && ldu < mr) return tci_fail(h, TCI_ERR_INVALID, "tci_rrlu_inplace_f64: ldu < min(maxrank, m, n)");
    SmallWs w; RrluState fs;
    rc = rrlu_run(h, d_A, (int)m, (int)n, lda, (int)mr, reltol, abstol, leftorthogonal ? 1 : 0, &w, &fs, d_L, ldl, d_U, ldu, true);
    if (rc) return rc;
    return rrlu_finish(h, d_A, (int)m, (int)n, lda, leftorthogonal ? 1 : 0, w, &fs, (int)mr, h_rowperm, h_colperm, h_npivot, h_error,
                       d_L, ldl, d_U, ldu, h_pivoterrors);
}

int tci_rrlu_host_f64(tci_handle h, const double* h_A, int64_t m, int64_t n, int64_t lda, int64_t maxrank, double reltol,
                      double abstol, int leftorthogonal, int64_t* h_rowperm, int64_t* h_colperm, int64_t* h_npivot,
                      double* h_error, double* h_L, int64_t ldl, double* h_U, int64_t ldu, double* h_pivoterrors) {
    int rc = check_dims(h, m, n, lda, "tci_rrlu_host_f64");
    if (rc) return rc;
    TCI_CUDA(h, cudaSetDevice(h->device));
    int64_t mn = m < n ? m : n;
    int64_t mr = (maxrank <= 0 || maxrank > mn) ? mn : maxrank;
    if (mn == 0) return tci_rrlu_inplace_f64(h, nullptr, m, n, lda, maxrank, reltol, abstol, leftorthogonal, h_rowperm,
                                             h_colperm, h_npivot, h_error, nullptr, 0, nullptr, 0, h_pivoterrors);
    if (!h_A) return tci_fail(h, TCI_ERR_INVALID, "tci_rrlu_host_f64: h_A is NULL");
    if (h_L && ldl < m) return tci_fail(h, TCI_ERR_INVALID, "tci_rrlu_host_f64: ldl < m");
    if (h_U && ldu < 1) return tci_fail(h, TCI_ERR_INVALID, "tci_rrlu_host_f64: ldu < 1");
    const int64_t ldw = (m + 15) & ~(int64_t)15;           // 128-byte aligned columns for the working copy (rrlu copies A, :462)
    rc = tci_ws_reserve(h, h->ws_matrix, sizeof(double) * (size_t)ldw * n);
    if (rc) return rc;
    double* dA = (double*)h->ws_matrix.ptr;
    TCI_CUDA(h, cudaMemcpy2DAsync(dA, sizeof(double) * ldw, h_A, sizeof(double) * lda, sizeof(double) * m, n,
                                  cudaMemcpyHostToDevice, h->stream));
    double *dL = nullptr, *dU = nullptr;
    if (h_L || h_U) {
        rc = tci_ws_reserve(h, h->ws_lu, sizeof(double) * ((size_t)m * mr + (size_t)mr * n));
        if (rc) return rc;
        dL = (double*)h->ws_lu.ptr; dU = dL + (size_t)m * mr;
    }
    SmallWs w; RrluState fs;
    rc = rrlu_run(h, dA, (int)m, (int)n, ldw, (int)mr, reltol, abstol, leftorthogonal ? 1 : 0, &w, &fs, h_L ? dL : nullptr, m,
                  h_U ? dU : nullptr, mr, true);
    if (rc) return rc;
    rc = rrlu_finish(h, dA, (int)m, (int)n, ldw, leftorthogonal ? 1 : 0, w, &fs, (int)mr, h_rowperm, h_colperm, h_npivot, h_error,
                     h_L ? dL : nullptr, m, h_U ? dU : nullptr, mr, h_pivoterrors);
    if (rc) return rc;
    const int64_t r = fs.npivot;
    if (r > 0 && h_L)
        TCI_CUDA(h, cudaMemcpy2DAsync(h_L, sizeof(double) * ldl, dL, sizeof(double) * m, sizeof(double) * m, r, cudaMemcpyDeviceToHost, h->stream));
    if (r > 0 && h_U) {
        if (ldu < r) return tci_fail(h, TCI_ERR_INVALID, "tci_rrlu_host_f64: ldu (%lld) < npivot (%lld)", (long long)ldu, (long long)r);
        TCI_CUDA(h, cudaMemcpy2DAsync(h_U, sizeof(double) * ldu, dU, sizeof(double) * mr, sizeof(double) * r, n, cudaMemcpyDeviceToHost, h->stream));
    }
    TCI_CUDA(h, cudaStreamSynchronize(h->stream));
    return TCI_OK;
}

int tci_submatrixargmax_abs2_f64(tci_handle h, const double* d_A, int64_t m, int64_t n, int64_t lda, int64_t k,
                                 int64_t* h_row, int64_t* h_col) {
    int rc = check_dims(h, m, n, lda, "tci_submatrixargmax_abs2_f64");
    if (rc) return rc;
    if (k < 1 || k > m) return tci_fail(h, TCI_ERR_INVALID, "rows must not be empty");      // matrixlu.jl:58
    if (k > n) return tci_fail(h, TCI_ERR_INVALID, "cols must not be empty");               // matrixlu.jl:59
    TCI_CUDA(h, cudaSetDevice(h->device));
    const bool vec = (lda % 2 == 0) && (((uintptr_t)d_A & 15u) == 0);
    const UpdateCfg cfg = update_cfg(h, vec);
    SmallWs w;
    rc = carve(h, (int)m, (int)n, cfg.grid, &w);
    if (rc) return rc;
    rc = tci_pinned_reserve(h, sizeof(RrluState));
    if (rc) return rc;
    rrlu_init_kernel<<<1, 32, 0, h->stream>>>(w.st, w.rowperm, w.colperm, 0, 0, w.bar);
    TCI_LAUNCH_CHECK(h);
    launch_update<false>(h, vec, cfg, const_cast<double*>(d_A), (int)m, (int)n, lda, w, INT32_MAX, 0.0, 0.0, (int)(k - 1));
    TCI_LAUNCH_CHECK(h);
    RrluState* hst = (RrluState*)h->pinned;
    TCI_CUDA(h, cudaMemcpyAsync(hst, w.st, sizeof(RrluState), cudaMemcpyDeviceToHost, h->stream));
    TCI_CUDA(h, cudaStreamSynchronize(h->stream));
    *h_row = hst->pr + 1; *h_col = hst->pc + 1;
    return TCI_OK;
}

int tci_rrlu_set_profiling(tci_handle h, int on) {
    if (!h) return TCI_ERR_INVALID;
    h->rrlu_profile = on ? 1 : 0;
    return TCI_OK;
}

int64_t tci_rrlu_last_launch_ms(tci_handle h, float* h_ms, int64_t cap) {
    if (!h) return 0;
    const int64_t n = (int64_t)h->rrlu_launch_ms.size();
    if (h_ms) for (int64_t i = 0; i < n && i < cap; ++i) h_ms[i] = h->rrlu_launch_ms[i];
    return n;
}

int tci_rrlu_last_timing(tci_handle h, double* h_ms3, int64_t* h_steps) {
    if (!h) return TCI_ERR_INVALID;
    if (h_ms3) for (int i = 0; i < 3; ++i) h_ms3[i] = h->rrlu_ms[i];
    if (h_steps) *h_steps = h->rrlu_steps;
    return TCI_OK;
}

} // extern "C"

// ---- column-sharded rrLU: host side ------------------------------------------------------------------------------------
struct tci_rrlu_dist_s {
    int m, n, nloc, col0, rank, nranks, maxrank, leftorth, vec, grid;
    double reltol, abstol;
    long long reclen;
    void* block;                 // one allocation behind all the arrays
    size_t block_bytes;
    DistArrays a;
    PeerBox peers;               // n == 0 until tci_rrlu_dist_set_mailboxes
};

template <bool UPDATE>
static void launch_dist_update(tci_handle h, tci_rrlu_dist_s* d, double* A, long long lda, double* record) {
    PeerBox pb = d->peers;
    if (record) pb.n = 0;                                // explicit record buffer: the caller gathers
    if (d->vec) launch_pdl(h, rrlu_dist_update_kernel<UPDATE, true, 8>, (unsigned)d->grid, (unsigned)kThreads, d->a, A, d->m, d->nloc, lda, d->maxrank, 1, record, pb, d->rank, d->reclen);
    else        launch_pdl(h, rrlu_dist_update_kernel<UPDATE, false, 8>, (unsigned)d->grid, (unsigned)kThreads, d->a, A, d->m, d->nloc, lda, d->maxrank, 1, record, pb, d->rank, d->reclen);
}

extern "C" {

int tci_rrlu_dist_create(tci_handle h, int64_t m, int64_t n, int64_t nloc, int64_t col0, int rank, int nranks, int64_t maxrank,
                         double reltol, double abstol, int leftorthogonal, tci_rrlu_dist* out) {
    if (!h || !out) return TCI_ERR_INVALID;
    if (m < 1 || n < 1 || nloc < 0 || col0 < 0 || col0 + nloc > n || nranks < 1 || rank < 0 || rank >= nranks || m > (1 << 30) || n > (1 << 30))
        return tci_fail(h, TCI_ERR_INVALID, "tci_rrlu_dist_create: bad arguments");
    TCI_CUDA(h, cudaSetDevice(h->device));
    tci_rrlu_dist_s* d = new tci_rrlu_dist_s();
    d->m = (int)m; d->n = (int)n; d->nloc = (int)nloc; d->col0 = (int)col0; d->rank = rank; d->nranks = nranks;
    const int64_t mn = m < n ? m : n;
    d->maxrank = (int)((maxrank <= 0 || maxrank > mn) ? mn : maxrank);
    d->leftorth = leftorthogonal ? 1 : 0; d->reltol = reltol; d->abstol = abstol;
    d->reclen = 4 + ((m + 1) & ~(int64_t)1);
    d->vec = 0; d->peers.n = 0;
    int nb = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, rrlu_dist_update_kernel<true, true, 8>, kThreads, 0) != cudaSuccess || nb < 1) nb = 1;
    d->grid = nb * h->num_sms;
    auto al = [](size_t x) { return (x + 255) & ~(size_t)255; };
    const size_t o_st = 0, o_rp = al(sizeof(DistState)), o_p2c = o_rp + al(sizeof(int) * m), o_c2p = o_p2c + al(sizeof(int) * n),
                 o_gid = o_c2p + al(sizeof(int) * n), o_y = o_gid + al(sizeof(int) * (nloc + 1)), o_x = o_y + al(sizeof(double) * (nloc + 1)),
                 o_pv = o_x + al(sizeof(double) * (m + 2)), o_parts = o_pv + al(sizeof(double) * d->maxrank),
                 o_L = o_parts + al(sizeof(Cand4) * d->grid), o_tr = o_L + al(sizeof(double) * (size_t)m * d->maxrank);
    const char* tr_env = getenv("TCI_DIST_TRACE");
    const bool tracing = tr_env && atoi(tr_env) > 0;
    const size_t total = o_tr + (tracing ? al(sizeof(unsigned long long) * 8 * ((size_t)d->maxrank + 1)) : 0);
    // a sweep creates one of these per bond: take the block the previous one parked in the handle when it is large enough
    // (cudaMalloc / cudaFree cost milliseconds each once peer mappings exist)
    if (h->dist_parked.ptr && h->dist_parked.bytes >= total) {
        d->block = h->dist_parked.ptr; d->block_bytes = h->dist_parked.bytes;
        h->dist_parked.ptr = nullptr; h->dist_parked.bytes = 0;
    } else {
        if (cudaMalloc(&d->block, total) != cudaSuccess) { delete d; return tci_fail(h, TCI_ERR_NOMEM, "tci_rrlu_dist_create: cudaMalloc of %zu bytes failed", total); }
        d->block_bytes = total;
    }
    char* b = (char*)d->block;
    d->a.st = (DistState*)(b + o_st); d->a.rowperm = (int*)(b + o_rp); d->a.pos2col = (int*)(b + o_p2c); d->a.col2pos = (int*)(b + o_c2p);
    d->a.gid = (int*)(b + o_gid); d->a.ybuf = (double*)(b + o_y); d->a.xcol = (double*)(b + o_x); d->a.pivvals = (double*)(b + o_pv);
    d->a.parts = (Cand4*)(b + o_parts); d->a.Lbuf = (double*)(b + o_L); d->a.ldl = m;
    d->a.trace = tracing ? (unsigned long long*)(b + o_tr) : nullptr;
    if (tracing) cudaMemsetAsync(d->a.trace, 0, sizeof(unsigned long long) * 8 * ((size_t)d->maxrank + 1), h->stream);
    const int64_t mx = m > n ? m : n;
    rrlu_dist_init_kernel<<<(unsigned)((mx + 255) / 256), 256, 0, h->stream>>>(d->a, d->m, d->n, d->nloc, d->col0);
    TCI_LAUNCH_CHECK(h);
    *out = d;
    return TCI_OK;
}

int64_t tci_rrlu_dist_record_doubles(tci_rrlu_dist d) { return d ? d->reclen : 0; }

int tci_rrlu_dist_search(tci_handle h, tci_rrlu_dist d, double* d_A, int64_t lda, double* d_record) {
    if (!h || !d || !d_record || (d->nloc > 0 && (!d_A || lda < d->m))) return tci_fail(h, TCI_ERR_INVALID, "tci_rrlu_dist_search: bad arguments");
    TCI_CUDA(h, cudaSetDevice(h->device));
    d->vec = (lda % 2 == 0) && (((uintptr_t)d_A & 15) == 0) ? 1 : 0;
    launch_dist_update<false>(h, d, d_A, lda, d_record);
    TCI_LAUNCH_CHECK(h);
    return TCI_OK;
}

int tci_rrlu_dist_step(tci_handle h, tci_rrlu_dist d, double* d_A, int64_t lda, const double* d_gathered, double* d_record) {
    if (!h || !d || !d_gathered || !d_record) return tci_fail(h, TCI_ERR_INVALID, "tci_rrlu_dist_step: bad arguments");
    TCI_CUDA(h, cudaSetDevice(h->device));
    rrlu_dist_select_kernel<<<1, kThreads, 0, h->stream>>>(d->a, d_gathered, d->reclen, d->nranks, d->m, d->maxrank, d->reltol, d->abstol, d->leftorth);
    TCI_LAUNCH_CHECK(h);
    const long long items = (long long)d->nloc + 2LL * d->m + d->maxrank + 1;
    PeerBox nobox = d->peers; nobox.n = 0;
    launch_pdl(h, rrlu_dist_swap_kernel, (unsigned)((items + 255) / 256), 256u, d->a, d_A, d->m, d->nloc, (long long)lda, d->rank, d->leftorth,
               d_gathered, d->reclen, d->nranks, nobox, d->maxrank, d->reltol, d->abstol);
    TCI_LAUNCH_CHECK(h);
    launch_dist_update<true>(h, d, d_A, lda, d_record);
    TCI_LAUNCH_CHECK(h);
    return TCI_OK;
}

int64_t tci_rrlu_dist_mailbox_bytes(tci_rrlu_dist d) {
    if (!d) return 0;
    return (int64_t)((6LL * d->nranks + 2LL * (d->reclen - 4)) * 16 + 256);    // 16-byte packets: hdr[2][n][3] | col[2][mpad]
}

int tci_rrlu_dist_set_mailboxes(tci_handle h, tci_rrlu_dist d, void* const* h_boxes) {
    if (!h || !d || !h_boxes) return TCI_ERR_INVALID;
    if (d->nranks > kMaxPeers) return tci_fail(h, TCI_ERR_INVALID, "tci_rrlu_dist_set_mailboxes: at most %d ranks", kMaxPeers);
    for (int g = 0; g < d->nranks; ++g) {
        if (!h_boxes[g]) return tci_fail(h, TCI_ERR_INVALID, "tci_rrlu_dist_set_mailboxes: NULL mailbox");
        d->peers.box[g] = (double*)h_boxes[g];
    }
    d->peers.n = d->nranks;
    TCI_CUDA(h, cudaSetDevice(h->device));
    TCI_CUDA(h, cudaMemsetAsync(h_boxes[d->rank], 0, (size_t)tci_rrlu_dist_mailbox_bytes(d), h->stream));   // own flags start at 0
    TCI_CUDA(h, cudaStreamSynchronize(h->stream));
    return TCI_OK;
}

int tci_rrlu_dist_run(tci_handle h, tci_rrlu_dist d, double* d_A, int64_t lda) {
    if (!h || !d || (d->nloc > 0 && (!d_A || lda < d->m))) return tci_fail(h, TCI_ERR_INVALID, "tci_rrlu_dist_run: bad arguments");
    if (d->peers.n != d->nranks) return tci_fail(h, TCI_ERR_INVALID, "tci_rrlu_dist_run: call tci_rrlu_dist_set_mailboxes first");
    TCI_CUDA(h, cudaSetDevice(h->device));
    d->vec = (lda % 2 == 0) && (((uintptr_t)d_A & 15) == 0) ? 1 : 0;
    launch_dist_update<false>(h, d, d_A, lda, nullptr);
    TCI_LAUNCH_CHECK(h);
    const long long items = (long long)d->nloc + 2LL * d->m + d->maxrank + 1;
    for (int s = 0; s < d->maxrank; ++s) {
        launch_pdl(h, rrlu_dist_swap_kernel, (unsigned)((items + 255) / 256), 256u, d->a, d_A, d->m, d->nloc, (long long)lda, d->rank, d->leftorth,
                   (const double*)nullptr, d->reclen, d->nranks, d->peers, d->maxrank, d->reltol, d->abstol);
        TCI_LAUNCH_CHECK(h);
        launch_dist_update<true>(h, d, d_A, lda, nullptr);
        TCI_LAUNCH_CHECK(h);
    }
    return TCI_OK;
}

int tci_rrlu_dist_status(tci_handle h, tci_rrlu_dist d, int64_t* h_npivot, int* h_stop) {
    if (!h || !d) return TCI_ERR_INVALID;
    int rc = tci_pinned_reserve(h, sizeof(DistState)); if (rc) return rc;
    TCI_CUDA(h, cudaMemcpyAsync(h->pinned, d->a.st, sizeof(DistState), cudaMemcpyDeviceToHost, h->stream));
    TCI_CUDA(h, cudaStreamSynchronize(h->stream));
    const DistState* st = (const DistState*)h->pinned;
    if (h_npivot) *h_npivot = st->npivot;
    if (h_stop) *h_stop = st->stop || st->npivot >= d->maxrank;
    if (st->flags & 8) return tci_fail(h, TCI_ERR_NAN_L, "sharded rrlu: every remaining entry is NaN");
    return TCI_OK;
}

int tci_rrlu_dist_trace(tci_handle h, tci_rrlu_dist d, uint64_t* h_stamps, int64_t nsteps) {
    if (!h || !d || !h_stamps) return TCI_ERR_INVALID;
    if (!d->a.trace) return tci_fail(h, TCI_ERR_INVALID, "tci_rrlu_dist_trace: create the factorisation with TCI_DIST_TRACE=1");
    if (nsteps > d->maxrank) nsteps = d->maxrank;
    TCI_CUDA(h, cudaSetDevice(h->device));
    TCI_CUDA(h, cudaStreamSynchronize(h->stream));
    TCI_CUDA(h, cudaMemcpy(h_stamps, d->a.trace, sizeof(uint64_t) * 8 * (size_t)nsteps, cudaMemcpyDeviceToHost));
    return TCI_OK;
}

int tci_rrlu_dist_finish(tci_handle h, tci_rrlu_dist d, const double* d_A, int64_t lda, int64_t* h_rowperm, int64_t* h_colperm,
                         int64_t* h_npivot, double* h_error, double* d_L, int64_t ldl, double* d_Uloc, int64_t ldu,
                         int64_t* h_locpos, double* h_pivoterrors) {
    if (!h || !d || !h_rowperm || !h_colperm || !h_npivot || !h_error) return tci_fail(h, TCI_ERR_INVALID, "tci_rrlu_dist_finish: bad arguments");
    TCI_CUDA(h, cudaSetDevice(h->device));
    const int m = d->m, n = d->n, nloc = d->nloc;
    const size_t pin = sizeof(DistState) + sizeof(int) * ((size_t)m + n + nloc) + sizeof(double) * d->maxrank + 256;
    int rc = tci_pinned_reserve(h, pin); if (rc) return rc;
    TCI_CUDA(h, cudaMemcpyAsync(h->pinned, d->a.st, sizeof(DistState), cudaMemcpyDeviceToHost, h->stream));
    TCI_CUDA(h, cudaStreamSynchronize(h->stream));
    DistState st = *(const DistState*)h->pinned;
    const int r = st.npivot;
    if (r > 0 && ((d_L && ldl < m) || (d_Uloc && ldu < r))) return tci_fail(h, TCI_ERR_INVALID, "tci_rrlu_dist_finish: leading dimensions too small");
    rc = tci_ws_reserve(h, h->ws_c, sizeof(int) * (size_t)(nloc + 1)); if (rc) return rc;
    int* d_locpos = (int*)h->ws_c.ptr;
    const long long items = (long long)m * r + (long long)r * nloc + nloc;
    if (items > 0) {
        rrlu_dist_extract_kernel<<<(unsigned)((items + 255) / 256), 256, 0, h->stream>>>(d->a, d_A, m, nloc, lda, r, d->leftorth, d_L, ldl, d_Uloc, ldu, d_locpos);
        TCI_LAUNCH_CHECK(h);
    }
    int* p_rp = (int*)((char*)h->pinned + 256); int* p_cp = p_rp + m; int* p_lp = p_cp + n; double* p_pv = (double*)(p_lp + nloc + ((m + n + nloc) & 1));
    TCI_CUDA(h, cudaMemcpyAsync(p_rp, d->a.rowperm, sizeof(int) * m, cudaMemcpyDeviceToHost, h->stream));
    TCI_CUDA(h, cudaMemcpyAsync(p_cp, d->a.pos2col, sizeof(int) * n, cudaMemcpyDeviceToHost, h->stream));
    if (nloc > 0) TCI_CUDA(h, cudaMemcpyAsync(p_lp, d_locpos, sizeof(int) * nloc, cudaMemcpyDeviceToHost, h->stream));
    if (r > 0) TCI_CUDA(h, cudaMemcpyAsync(p_pv, d->a.pivvals, sizeof(double) * r, cudaMemcpyDeviceToHost, h->stream));
    TCI_CUDA(h, cudaStreamSynchronize(h->stream));
    for (int i = 0; i < m; ++i) h_rowperm[i] = p_rp[i] + 1;
    for (int j = 0; j < n; ++j) h_colperm[j] = p_cp[j] + 1;
    if (h_locpos) for (int j = 0; j < nloc; ++j) h_locpos[j] = p_lp[j];
    *h_npivot = r;
    *h_error = (r >= (m < n ? m : n)) ? 0.0 : st.error;                 // :391-393
    if (h_pivoterrors) { for (int c = 0; c < r; ++c) h_pivoterrors[c] = fabs(p_pv[c]); h_pivoterrors[r] = *h_error; }
    if (st.flags & 8) return tci_fail(h, TCI_ERR_NAN_L, "sharded rrlu: every remaining entry is NaN");
    if (st.flags & 4) return tci_fail(h, TCI_ERR_CUDA, "sharded rrlu: a peer's record did not arrive within 4 s");
    return TCI_OK;
}

int tci_rrlu_dist_destroy(tci_handle h, tci_rrlu_dist d) {
    if (!d) return TCI_OK;
    if (h) cudaSetDevice(h->device);
    if (h && d->block_bytes > h->dist_parked.bytes) {    // park the larger block (stream order protects its reuse)
        if (h->dist_parked.ptr) { cudaStreamSynchronize(h->stream); cudaFree(h->dist_parked.ptr); }
        h->dist_parked.ptr = d->block; h->dist_parked.bytes = d->block_bytes;
    } else {
        if (h) cudaStreamSynchronize(h->stream);
        cudaFree(d->block);
    }
    delete d;
    return TCI_OK;
}

} // extern "C"
