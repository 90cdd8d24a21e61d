// dense.cu — K4a: FP64 dense algebra for site-tensor formation.
//
//   tci_gemm_f64          strided/batched DGEMM on the FP64 tensor pipe (DMMA.8x8x4 via mma.sync m8n8k4.f64 —
//                         tcgen05 has no FP64 kind, this is the FP64 tensor path sm_100a exposes)
//   tci_trsm_right_f64    X * T = B for triangular T (blocked: 32x32 diagonal-block inverses + GEMMs)
//   tci_luci_factors_f64  left(luci) / right(luci)   src/matrixluci.jl:161-283
//   tci_solve_right_f64   Pi1 * inv(P)               src/tensorci2.jl:626
//
// Operands are addressed with explicit element strides (row stride, column stride), so transposes and
// the reference's `permutedims`+`reshape` views (src/contraction.jl:211-212) cost nothing.
#include "context.h"
#include "solve_small.h"

static inline unsigned ycap(long long cols) { return (unsigned)(cols < 1 ? 1 : (cols > 65535 ? 65535 : cols)); }   // gridDim.y limit; the kernels loop
#include <cmath>
#include <cstdlib>
#include <vector>

namespace {

constexpr int BM = 64, BN = 64, BK = 16, LDS = 68;      // LDS = 4 mod 16: conflict-free 64-bit fragment loads
constexpr int GEMM_THREADS = 256;

struct GemmArgs {
    int M, N, K;
    double alpha, beta;
    const double* A; long long sam, sak, bsa;           // A(m,k) = A[m*sam + k*sak + z*bsa]
    const double* B; long long sbk, sbn, bsb;           // B(k,n)
    double* C; long long scm, scn, bsc;                 // C(m,n)
};

__device__ __forceinline__ void dmma884(double& d0, double& d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

// Shared-memory layout of an operand tile follows the operand's contiguous direction so that BOTH the coalesced global
// load -> shared store and the DMMA fragment load are bank-conflict free:
//   row index (m or n) contiguous in global memory: tile[k][row], leading dimension LDS = 68  (68 mod 16 = 4)
//   k contiguous in global memory:                  tile[row][k], leading dimension LDK = 20  (20 mod 16 = 4)
// (64-bit accesses are served per half-warp; 16 addresses are conflict-free iff distinct mod 16 in 8-byte units.)
constexpr int LDK = 20;
constexpr int TILE_DOUBLES = (BK * LDS > BM * LDK) ? BK * LDS : BM * LDK;

template <bool KFAST> __device__ __forceinline__ int tile_off(int row, int k) { return KFAST ? row * LDK + k : k * LDS + row; }

template <bool A_KFAST, bool B_KFAST>
__global__ void __launch_bounds__(GEMM_THREADS)
gemm_dmma_kernel(GemmArgs g) {
    __shared__ double As[2][TILE_DOUBLES];
    __shared__ double Bs[2][TILE_DOUBLES];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int wm = warp & 1, wn = warp >> 1;            // warp tile: 32 rows x 16 cols
    const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
    const double* A = g.A + (long long)blockIdx.z * g.bsa;
    const double* B = g.B + (long long)blockIdx.z * g.bsb;
    double* C = g.C + (long long)blockIdx.z * g.bsc;

    double acc[4][2][2];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 2; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    double ra[4], rb[4];
    auto gload = [&](int k0) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            int m, k;
            if (!A_KFAST) { m = t & 63; k = (t >> 6) + 4 * i; } else { k = t & 15; m = (t >> 4) + 16 * i; }
            const int gm = m0 + m, gk = k0 + k;
            ra[i] = (gm < g.M && gk < g.K) ? A[(long long)gm * g.sam + (long long)gk * g.sak] : 0.0;
            int n, kb;
            if (B_KFAST) { kb = t & 15; n = (t >> 4) + 16 * i; } else { n = t & 63; kb = (t >> 6) + 4 * i; }
            const int gn = n0 + n, gkb = k0 + kb;
            rb[i] = (gn < g.N && gkb < g.K) ? B[(long long)gkb * g.sbk + (long long)gn * g.sbn] : 0.0;
        }
    };
    auto sstore = [&](int buf) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            int m, k;
            if (!A_KFAST) { m = t & 63; k = (t >> 6) + 4 * i; } else { k = t & 15; m = (t >> 4) + 16 * i; }
            As[buf][tile_off<A_KFAST>(m, k)] = ra[i];
            int n, kb;
            if (B_KFAST) { kb = t & 15; n = (t >> 4) + 16 * i; } else { n = t & 63; kb = (t >> 6) + 4 * i; }
            Bs[buf][tile_off<B_KFAST>(n, kb)] = rb[i];
        }
    };

    const int nk = (g.K + BK - 1) / BK;
    gload(0);
    sstore(0);
    __syncthreads();
    for (int kt = 0; kt < nk; ++kt) {
        const int buf = kt & 1;
        if (kt + 1 < nk) gload((kt + 1) * BK);
#pragma unroll
        for (int kk = 0; kk < BK; kk += 4) {
            double a[4], b[2];
#pragma unroll
            for (int i = 0; i < 4; ++i) a[i] = As[buf][tile_off<A_KFAST>(wm * 32 + i * 8 + (lane >> 2), kk + (lane & 3))];
#pragma unroll
            for (int j = 0; j < 2; ++j) b[j] = Bs[buf][tile_off<B_KFAST>(wn * 16 + j * 8 + (lane >> 2), kk + (lane & 3))];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 2; ++j) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
        }
        if (kt + 1 < nk) sstore(buf ^ 1);
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 2; ++j)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int gm = m0 + wm * 32 + i * 8 + (lane >> 2);
                const int gn = n0 + wn * 16 + j * 8 + (lane & 3) * 2 + e;
                if (gm < g.M && gn < g.N) {
                    double* p = C + (long long)gm * g.scm + (long long)gn * g.scn;
                    double v = g.alpha * acc[i][j][e];
                    if (g.beta != 0.0) v += g.beta * *p;
                    *p = v;
                }
            }
}

// ---- pipelined variant for large products: 128 x 64 x 16 CTA tile, 8 warps of 32 x 32, 3-stage cp.async ring -----------
// Same fragment layout, same k order (so bit-identical results to the kernel above); the operand tiles are fetched by
// 8-byte cp.async (any element strides; out-of-range elements are zero-filled with src-size 0) two k-tiles ahead of the
// DMMAs, with one __syncthreads per k-tile.
constexpr int PM = 128, PN = 64, PSTAGES = 3;
constexpr int LDSA = 132;                                // 132 mod 16 = 4: conflict-free like LDS = 68
constexpr int PA_DOUBLES = (BK * LDSA > PM * LDK) ? BK * LDSA : PM * LDK;     // 2560
constexpr int PB_DOUBLES = (BK * LDS > PN * LDK) ? BK * LDS : PN * LDK;       // 1280
constexpr int PIPE_SMEM_BYTES = PSTAGES * (PA_DOUBLES + PB_DOUBLES) * 8;      // 92160

template <bool KFAST> __device__ __forceinline__ int tile_off_a(int row, int k) { return KFAST ? row * LDK + k : k * LDSA + row; }

__device__ __forceinline__ void cp_async8(double* smem, const double* gptr, bool valid) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    const int sz = valid ? 8 : 0;
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(s), "l"(gptr), "r"(sz) : "memory");
}

template <bool A_KFAST, bool B_KFAST>
__global__ void __launch_bounds__(GEMM_THREADS, 2)
gemm_dmma_pipe_kernel(GemmArgs g) {
    extern __shared__ __align__(16) double psm[];
    double* As = psm;                                    // [PSTAGES][PA_DOUBLES]
    double* Bs = psm + PSTAGES * PA_DOUBLES;             // [PSTAGES][PB_DOUBLES]
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int wm = warp & 3, wn = warp >> 2;             // warp tile: 32 rows x 32 cols
    const int m0 = blockIdx.x * PM, n0 = blockIdx.y * PN;
    const double* A = g.A + (long long)blockIdx.z * g.bsa;
    const double* B = g.B + (long long)blockIdx.z * g.bsb;
    double* C = g.C + (long long)blockIdx.z * g.bsc;

    double acc[4][4][2];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    // Per-thread load plan, hoisted out of the k loop: element i of this thread's share of a tile sits at
    // base + i * step (global) and off0 + i * sstep (shared); only the k offset changes from tile to tile.
    const int am = A_KFAST ? (t >> 4) : (t & 127), ak = A_KFAST ? (t & 15) : (t >> 7);      // i = 0 element (row, k)
    const int bn = B_KFAST ? (t >> 4) : (t & 63), bk = B_KFAST ? (t & 15) : (t >> 6);
    const long long a_step = A_KFAST ? 16 * g.sam : 2 * g.sak;          // A_KFAST: row += 16 per i; else k += 2 per i
    const long long b_step = B_KFAST ? 16 * g.sbn : 4 * g.sbk;
    constexpr int a_sstep = A_KFAST ? 16 * LDK : 2 * LDSA;
    constexpr int b_sstep = B_KFAST ? 16 * LDK : 4 * LDS;
    const int a_soff = tile_off_a<A_KFAST>(am, ak), b_soff = tile_off<B_KFAST>(bn, bk);
    const double* a_base = A + (long long)(m0 + am) * g.sam + (long long)ak * g.sak;
    const double* b_base = B + (long long)(n0 + bn) * g.sbn + (long long)bk * g.sbk;
    unsigned a_rows = 0, b_rows = 0;                     // bit i: the row of element i is inside the matrix
#pragma unroll
    for (int i = 0; i < 8; ++i) a_rows |= ((m0 + am + (A_KFAST ? 16 * i : 0)) < g.M ? 1u : 0u) << i;
#pragma unroll
    for (int i = 0; i < 4; ++i) b_rows |= ((n0 + bn + (B_KFAST ? 16 * i : 0)) < g.N ? 1u : 0u) << i;
    const long long a_kstep = (long long)BK * g.sak, b_kstep = (long long)BK * g.sbk;

    const unsigned a_saddr = (unsigned)__cvta_generic_to_shared(As + a_soff);
    const unsigned b_saddr = (unsigned)__cvta_generic_to_shared(Bs + b_soff);
    const bool interior = (m0 + PM <= g.M) && (n0 + PN <= g.N);       // CTA tile entirely inside C: no row checks
    const double* a_next = a_base;                       // running pointers: the next k-tile to be loaded
    const double* b_next = b_base;

    auto load_stage = [&](int stage, int kt_) {
        const int krem = g.K - kt_ * BK;                 // k indices < krem are valid in this tile
        if (interior && krem >= BK) {                    // the common case: plain copies, two adds per element
            const double* ap = a_next;
            const double* bp = b_next;
            const unsigned sa = a_saddr + stage * (PA_DOUBLES * 8), sb = b_saddr + stage * (PB_DOUBLES * 8);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(sa + i * (a_sstep * 8)), "l"(ap) : "memory");
                ap += a_step;
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(sb + i * (b_sstep * 8)), "l"(bp) : "memory");
                bp += b_step;
            }
        } else {
            double* as = As + stage * PA_DOUBLES + a_soff;
            double* bs = Bs + stage * PB_DOUBLES + b_soff;
#pragma unroll
            for (int i = 0; i < 8; ++i) {                // A tile: 128 x 16
                const bool ok = ((a_rows >> i) & 1u) && (ak + (A_KFAST ? 0 : 2 * i)) < krem;
                cp_async8(as + i * a_sstep, ok ? a_next + i * a_step : A, ok);
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {                // B tile: 16 x 64
                const bool ok = ((b_rows >> i) & 1u) && (bk + (B_KFAST ? 0 : 4 * i)) < krem;
                cp_async8(bs + i * b_sstep, ok ? b_next + i * b_step : B, ok);
            }
        }
        a_next += a_kstep;
        b_next += b_kstep;
    };

    const int nk = (g.K + BK - 1) / BK;
#pragma unroll
    for (int s = 0; s < PSTAGES - 1; ++s) {
        if (s < nk) load_stage(s, s);
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    for (int kt = 0; kt < nk; ++kt) {
        asm volatile("cp.async.wait_group %0;" ::"n"(PSTAGES - 2) : "memory");
        __syncthreads();                                 // stage kt has landed for everyone; stage kt-1 is free again
        const int nxt = kt + PSTAGES - 1;
        if (nxt < nk) load_stage(nxt % PSTAGES, nxt);
        asm volatile("cp.async.commit_group;" ::: "memory");
        const double* as = As + (kt % PSTAGES) * PA_DOUBLES;
        const double* bs = Bs + (kt % PSTAGES) * PB_DOUBLES;
#pragma unroll
        for (int kk = 0; kk < BK; kk += 4) {
            double a[4], b[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) a[i] = as[tile_off_a<A_KFAST>(wm * 32 + i * 8 + (lane >> 2), kk + (lane & 3))];
#pragma unroll
            for (int j = 0; j < 4; ++j) b[j] = bs[tile_off<B_KFAST>(wn * 32 + j * 8 + (lane >> 2), kk + (lane & 3))];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
        }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int gm = m0 + wm * 32 + i * 8 + (lane >> 2);
                const int gn = n0 + wn * 32 + j * 8 + (lane & 3) * 2 + e;
                if (gm < g.M && gn < g.N) {
                    double* p = C + (long long)gm * g.scm + (long long)gn * g.scn;
                    double v = g.alpha * acc[i][j][e];
                    if (g.beta != 0.0) v += g.beta * *p;
                    *p = v;
                }
            }
}

template <bool AK, bool BKF>
int launch_pipe(tci_handle h, const GemmArgs& g, int batch) {
    static bool optin[8] = {false, false, false, false, false, false, false, false};   // per device
    const int dev = h->device & 7;
    if (!optin[dev]) {
        TCI_CUDA(h, cudaFuncSetAttribute(gemm_dmma_pipe_kernel<AK, BKF>, cudaFuncAttributeMaxDynamicSharedMemorySize, PIPE_SMEM_BYTES));
        optin[dev] = true;
    }
    dim3 grid((g.M + PM - 1) / PM, (g.N + PN - 1) / PN, batch);
    gemm_dmma_pipe_kernel<AK, BKF><<<grid, GEMM_THREADS, PIPE_SMEM_BYTES, h->stream>>>(g);
    TCI_LAUNCH_CHECK(h);
    return TCI_OK;
}

__global__ void copy_strided_kernel(const double* __restrict__ src, long long ssr, long long ssc, double* __restrict__ dst,
                                        long long dsr, long long dsc, int rows, int cols) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows) return;
    for (int j = blockIdx.y; j < cols; j += gridDim.y)       // grid.y is capped at 65535: wide matrices loop
        dst[(long long)i * dsr + (long long)j * dsc] = src[(long long)i * ssr + (long long)j * ssc];
}

int launch_gemm(tci_handle h, const GemmArgs& g, int batch) {
    if (g.M <= 0 || g.N <= 0 || batch <= 0) return TCI_OK;
    dim3 grid((g.M + BM - 1) / BM, (g.N + BN - 1) / BN, batch);
    const bool a_kfast = (g.sak == 1) && (g.sam != 1), b_kfast = (g.sbk == 1);
    static const int pipe_min = []() { const char* e = getenv("TCI_GEMM_PIPE_MIN"); return e ? atoi(e) : 128; }();
    if (pipe_min >= 0 && g.M >= pipe_min && g.N >= 64 && g.K >= 64 && (long long)((g.M + PM - 1) / PM) * ((g.N + PN - 1) / PN) * batch >= 2LL * h->num_sms) {   // at least one full wave of 2 CTAs per SM
        if (a_kfast) return b_kfast ? launch_pipe<true, true>(h, g, batch) : launch_pipe<true, false>(h, g, batch);
        return b_kfast ? launch_pipe<false, true>(h, g, batch) : launch_pipe<false, false>(h, g, batch);
    }
    if (a_kfast) { if (b_kfast) gemm_dmma_kernel<true, true><<<grid, GEMM_THREADS, 0, h->stream>>>(g); else gemm_dmma_kernel<true, false><<<grid, GEMM_THREADS, 0, h->stream>>>(g); }
    else         { if (b_kfast) gemm_dmma_kernel<false, true><<<grid, GEMM_THREADS, 0, h->stream>>>(g); else gemm_dmma_kernel<false, false><<<grid, GEMM_THREADS, 0, h->stream>>>(g); }
    TCI_LAUNCH_CHECK(h);
    return TCI_OK;
}

// Inverse of each NB x NB diagonal block of a triangular matrix T (element (i,j) at T[i*sr + j*sc]), one
// CUDA block per diagonal block.  inv is stored dense column-major NB x NB per block (zero outside the
// triangle / beyond r).  lower != 0: lower triangular; unit != 0: implicit unit diagonal.
constexpr int NB = 32;
__global__ void diag_inverse_kernel(const double* __restrict__ T, long long sr, long long sc, int r, int lower, int unit,
                                    double* __restrict__ inv) {
    __shared__ double S[NB][NB + 1];
    __shared__ double X[NB][NB + 1];
    const int b = blockIdx.x, o = b * NB, nb = min(NB, r - o);
    for (int e = threadIdx.x; e < NB * NB; e += blockDim.x) {
        const int i = e % NB, j = e / NB;
        double v = 0.0;
        if (i < nb && j < nb) {
            const bool tri = lower ? (i >= j) : (i <= j);
            if (tri) v = (i == j && unit) ? 1.0 : T[(long long)(o + i) * sr + (long long)(o + j) * sc];
        } else if (i == j) v = 1.0;
        S[i][j] = v; X[i][j] = 0.0;
    }
    __syncthreads();
    const int j = threadIdx.x;                           // column j of the inverse: S * x = e_j
    if (j < NB) {
        if (lower) {
            for (int i = j; i < NB; ++i) {
                double s = (i == j) ? 1.0 : 0.0;
                for (int k = j; k < i; ++k) s -= S[i][k] * X[k][j];
                X[i][j] = s / S[i][i];
            }
        } else {
            for (int i = j; i >= 0; --i) {
                double s = (i == j) ? 1.0 : 0.0;
                for (int k = i + 1; k <= j; ++k) s -= S[i][k] * X[k][j];
                X[i][j] = s / S[i][i];
            }
        }
    }
    __syncthreads();
    for (int e = threadIdx.x; e < NB * NB; e += blockDim.x) {
        const int i = e % NB, jj = e / NB;
        inv[(long long)b * NB * NB + (long long)jj * NB + i] = (i < nb && jj < nb) ? X[i][jj] : 0.0;
    }
}

// X * T = B in place on B (q x r, element (i,j) at B[i*sbr + j*sbc]); T r x r triangular with strides (sr, sc).
int trsm_right(tci_handle h, const double* T, long long sr, long long sc, int r, int lower, int unit,
               double* Bm, long long sbr, long long sbc, int q) {
    if (r <= 0 || q <= 0) return TCI_OK;
    const int nblk = (r + NB - 1) / NB;
    int rc = tci_ws_reserve(h, h->ws_c, sizeof(double) * ((size_t)nblk * NB * NB + (size_t)q * NB));
    if (rc) return rc;
    double* inv = (double*)h->ws_c.ptr;
    double* tmp = inv + (size_t)nblk * NB * NB;           // q x NB scratch (column-major, ld q)
    diag_inverse_kernel<<<nblk, 64, 0, h->stream>>>(T, sr, sc, r, lower, unit, inv);
    TCI_LAUNCH_CHECK(h);
    for (int s = 0; s < nblk; ++s) {
        const int b = lower ? nblk - 1 - s : s;          // lower: last block column first; upper: first block first
        const int o = b * NB, nb = (r - o < NB) ? r - o : NB;
        // B_b -= X_done * T[done, b]
        const int d0 = lower ? o + nb : 0, dn = lower ? r - d0 : o;
        if (dn > 0) {
            GemmArgs g{q, nb, dn, -1.0, 1.0, Bm + (long long)d0 * sbc, sbr, sbc, 0,
                       T + (long long)d0 * sr + (long long)o * sc, sr, sc, 0, Bm + (long long)o * sbc, sbr, sbc, 0};
            rc = launch_gemm(h, g, 1);
            if (rc) return rc;
        }
        // X_b = B_b * inv(T_bb): GEMM into scratch (the product reads B_b), then copy back
        GemmArgs g1{q, nb, nb, 1.0, 0.0, Bm + (long long)o * sbc, sbr, sbc, 0, inv + (size_t)b * NB * NB, 1, NB, 0, tmp, 1, q, 0};
        rc = launch_gemm(h, g1, 1);
        if (rc) return rc;
        copy_strided_kernel<<<dim3((q + 255) / 256, ycap(nb)), 256, 0, h->stream>>>(tmp, 1, q, Bm + (long long)o * sbc, sbr, sbc, q, nb);
        TCI_LAUNCH_CHECK(h);
    }
    return TCI_OK;
}

} // namespace

namespace {

// out[perm[i], :] = (i < nident) ? e_i : W[i - woff, :]   (rows scattered through a 0-based permutation)
__global__ void scatter_rows_kernel(double* __restrict__ out, long long ldo, const double* __restrict__ W, long long ldw,
                                    const int* __restrict__ perm, int m, int r, int nident) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    for (int j = blockIdx.y; j < r; j += gridDim.y) {
        const double v = (i < nident) ? (i == j ? 1.0 : 0.0) : W[(long long)j * ldw + (i - nident)];
        out[(long long)j * ldo + perm[i]] = v;
    }
}
// out[:, perm[j]] = (j < nident) ? e_j : W[:, j - nident]
__global__ void scatter_cols_kernel(double* __restrict__ out, long long ldo, const double* __restrict__ W, long long ldw,
                                    const int* __restrict__ perm, int r, int n, int nident) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= r) return;
    for (int j = blockIdx.y; j < n; j += gridDim.y) {
        const double v = (j < nident) ? (i == j ? 1.0 : 0.0) : W[(long long)(j - nident) * ldw + i];
        out[(long long)perm[j] * ldo + i] = v;
    }
}
// W[:, j] = src[:, perm[j]]
__global__ void gather_cols_kernel(double* __restrict__ W, long long ldw, const double* __restrict__ src, long long lds,
                                   const int* __restrict__ perm, int q, int r) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= q) return;
    for (int j = blockIdx.y; j < r; j += gridDim.y) W[(long long)j * ldw + i] = src[(long long)perm[j] * lds + i];
}

int upload_perm(tci_handle h, const int64_t* h_perm, int len, int* d_dst, size_t) {
    std::vector<int> tmp(len);                             // pageable: staged by cudaMemcpyAsync before it returns, no sync needed
    for (int i = 0; i < len; ++i) tmp[i] = (int)(h_perm[i] - 1);
    TCI_CUDA(h, cudaMemcpyAsync(d_dst, tmp.data(), sizeof(int) * len, cudaMemcpyHostToDevice, h->stream));
    return TCI_OK;
}

} // namespace

namespace {

// ---- small pivot matrices: the whole solve X = B * inv(P) (tensorci2.jl:626) in ONE launch --------------------------------
// fillsitetensors! calls this once per site and sweep with P of size chi x chi; for the low-rank functions the reference is
// tested on chi is a few tens, and the general path below (copy, rrLU with its host round trip, two permutation uploads,
// gather, two blocked triangular solves, scatter: a dozen launches) is pure launch latency.  Here every CTA factorises P in
// shared memory by full-pivot LU (redundantly: chi <= 64, a few microseconds) and then each thread carries one row of B
// through the two substitutions.  P[rp, cp] = L U  =>  X[:, rp] = (B[:, cp] U^-1) L^-1.

__global__ void __launch_bounds__(kSolveRows)
solve_small_kernel(const double* __restrict__ B, long long ldb, const double* __restrict__ P, long long ldp, int q, int r,
                   double* __restrict__ X, long long ldx, int* __restrict__ flag) {
    extern __shared__ double sm[];
    const int lds = r | 1;                                 // odd leading dimension: conflict-free row and column walks
    double* S = sm;                                        // S[j * lds + i] = P(i, j), becomes L \ U in swapped coordinates
    double* W = S + (size_t)lds * r;                       // W[k * kSolveRows + tid]: this thread's row, column k
    __shared__ int rp[kSolveMaxR], cp[kSolveMaxR];
    __shared__ SmallLuScratch<kSolveRows> scratch;
    const int tid = threadIdx.x;
    for (int e = tid; e < r * r; e += kSolveRows) { const int i = e % r, j = e / r; S[j * lds + i] = P[(long long)j * ldp + i]; }
    __syncthreads();
    const bool singular = small_lu_fullpivot<kSolveRows>(S, lds, r, rp, cp, scratch);
    if (tid == 0) *flag = singular ? 1 : 0;                // every CTA writes the same value
    if (singular) return;
    const long long row = (long long)blockIdx.x * kSolveRows + tid;
    const bool valid = row < q;
    for (int k = 0; k < r; ++k) W[k * kSolveRows + tid] = valid ? B[(long long)cp[k] * ldb + row] : 0.0;
    small_solve_row<kSolveRows>(W, S, lds, r);
    if (valid) for (int k = 0; k < r; ++k) X[(long long)rp[k] * ldx + row] = W[k * kSolveRows + tid];
}

int solve_right_small(tci_handle h, const double* d_B, int Q, long long ldb, const double* d_P, int R, long long ldp, double* d_X,
                      long long ldx) {
    int rc = tci_ws_reserve(h, h->ws_b, 256); if (rc) return rc;
    rc = tci_pinned_reserve(h, 64); if (rc) return rc;
    int* d_flag = (int*)h->ws_b.ptr;
    const size_t smem = sizeof(double) * ((size_t)(R | 1) * R + (size_t)R * kSolveRows);
    if (smem > 48 * 1024 && !h->solve_smem_optin) {
        const size_t most = sizeof(double) * ((size_t)(kSolveMaxR | 1) * kSolveMaxR + (size_t)kSolveMaxR * kSolveRows);
        TCI_CUDA(h, cudaFuncSetAttribute(solve_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)most));
        h->solve_smem_optin = 1;
    }
    solve_small_kernel<<<(Q + kSolveRows - 1) / kSolveRows, kSolveRows, smem, h->stream>>>(d_B, ldb, d_P, ldp, Q, R, d_X, ldx, d_flag);
    TCI_LAUNCH_CHECK(h);
    int* pin = (int*)h->pinned;
    TCI_CUDA(h, cudaMemcpyAsync(pin, d_flag, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    TCI_CUDA(h, cudaStreamSynchronize(h->stream));
    if (*pin) return tci_fail(h, TCI_ERR_INVALID, "tci_solve_right_f64: pivot matrix is singular (rank < %d)", R);
    return TCI_OK;
}

} // namespace

extern "C" {

int tci_gemm_f64(tci_handle h, int64_t M, int64_t N, int64_t K, double alpha,
                 const double* d_A, int64_t sam, int64_t sak, const double* d_B, int64_t sbk, int64_t sbn,
                 double beta, double* d_C, int64_t scm, int64_t scn,
                 int64_t batch, int64_t bsa, int64_t bsb, int64_t bsc) {
    if (!h) return TCI_ERR_INVALID;
    if (M < 0 || N < 0 || K < 0 || batch < 0 || M > INT32_MAX || N > INT32_MAX || K > INT32_MAX || batch > 65535)
        return tci_fail(h, TCI_ERR_INVALID, "tci_gemm_f64: bad sizes");
    if (M == 0 || N == 0 || batch == 0) return TCI_OK;
    if (!d_C || (K > 0 && (!d_A || !d_B))) return tci_fail(h, TCI_ERR_INVALID, "tci_gemm_f64: NULL operand");
    TCI_CUDA(h, cudaSetDevice(h->device));
    GemmArgs g{(int)M, (int)N, (int)K, alpha, beta, d_A, sam, sak, bsa, d_B, sbk, sbn, bsb, d_C, scm, scn, bsc};
    return launch_gemm(h, g, (int)batch);
}

int tci_trsm_right_f64(tci_handle h, const double* d_T, int64_t st_row, int64_t st_col, int64_t r, int lower, int unit,
                       double* d_B, int64_t sb_row, int64_t sb_col, int64_t q) {
    if (!h) return TCI_ERR_INVALID;
    if (r < 0 || q < 0 || r > INT32_MAX || q > INT32_MAX) return tci_fail(h, TCI_ERR_INVALID, "tci_trsm_right_f64: bad sizes");
    TCI_CUDA(h, cudaSetDevice(h->device));
    return trsm_right(h, d_T, st_row, st_col, (int)r, lower, unit, d_B, sb_row, sb_col, (int)q);
}

int tci_luci_factors_f64(tci_handle h, int64_t m, int64_t n, int64_t r, int leftorthogonal,
                         const int64_t* h_rowperm, const int64_t* h_colperm,
                         const double* d_L, int64_t ldl, const double* d_U, int64_t ldu,
                         double* d_left, int64_t ldleft, double* d_right, int64_t ldright) {
    if (!h) return TCI_ERR_INVALID;
    if (m < 0 || n < 0 || r < 0 || r > m || r > n) return tci_fail(h, TCI_ERR_INVALID, "tci_luci_factors_f64: bad sizes");
    if (r == 0 || (!d_left && !d_right)) return TCI_OK;
    if (!h_rowperm || !h_colperm || !d_L || !d_U) return tci_fail(h, TCI_ERR_INVALID, "tci_luci_factors_f64: NULL input");
    TCI_CUDA(h, cudaSetDevice(h->device));
    const int M = (int)m, N = (int)n, R = (int)r;
    const size_t permbytes = ((sizeof(int) * (size_t)(M + N)) + 255) & ~(size_t)255;
    const size_t wl = (size_t)M * R, wr = (size_t)R * N;
    int rc = tci_ws_reserve(h, h->ws_a, permbytes + sizeof(double) * (wl > wr ? wl : wr) + 256);
    if (rc) return rc;
    rc = tci_pinned_reserve(h, permbytes + 64);
    if (rc) return rc;
    int* d_rp = (int*)h->ws_a.ptr; int* d_cp = d_rp + M;
    double* W = (double*)((char*)h->ws_a.ptr + permbytes);
    rc = upload_perm(h, h_rowperm, M, d_rp, 0); if (rc) return rc;
    rc = upload_perm(h, h_colperm, N, d_cp, sizeof(int) * (size_t)M); if (rc) return rc;
    if (d_left) {
        if (leftorthogonal) {                             // colstimespivotinv  matrixluci.jl:194-213
            const int q = M - R;
            if (q > 0) {
                copy_strided_kernel<<<dim3((q + 255) / 256, ycap(R)), 256, 0, h->stream>>>(d_L + R, 1, ldl, W, 1, q, q, R);
                TCI_LAUNCH_CHECK(h);
                rc = trsm_right(h, d_L, 1, ldl, R, 1, 0, W, 1, q, q);   // diagonal read from L (1.0 where the LU is left-orthogonal)
                if (rc) return rc;
            }
            scatter_rows_kernel<<<dim3((M + 255) / 256, ycap(R)), 256, 0, h->stream>>>(d_left, ldleft, W, q > 0 ? q : 1, d_rp, M, R, R);
            TCI_LAUNCH_CHECK(h);
        } else {                                          // colmatrix  matrixluci.jl:161-165
            GemmArgs g{M, R, R, 1.0, 0.0, d_L, 1, ldl, 0, d_U, 1, ldu, 0, W, 1, M, 0};
            rc = launch_gemm(h, g, 1); if (rc) return rc;
            scatter_rows_kernel<<<dim3((M + 255) / 256, ycap(R)), 256, 0, h->stream>>>(d_left, ldleft, W, M, d_rp, M, R, 0);
            TCI_LAUNCH_CHECK(h);
        }
    }
    if (d_right) {
        if (leftorthogonal) {                             // rowmatrix  matrixluci.jl:175-177
            GemmArgs g{R, N, R, 1.0, 0.0, d_L, 1, ldl, 0, d_U, 1, ldu, 0, W, 1, R, 0};
            rc = launch_gemm(h, g, 1); if (rc) return rc;
            scatter_cols_kernel<<<dim3((R + 255) / 256, ycap(N)), 256, 0, h->stream>>>(d_right, ldright, W, R, d_cp, R, N, 0);
            TCI_LAUNCH_CHECK(h);
        } else {                                          // pivotinvtimesrows  matrixluci.jl:227-241
            const int q = N - R;
            if (q > 0) {
                copy_strided_kernel<<<dim3((R + 255) / 256, ycap(q)), 256, 0, h->stream>>>(d_U + (long long)R * ldu, 1, ldu, W, 1, R, R, q);
                TCI_LAUNCH_CHECK(h);
                // U1 X = U2  <=>  X^T U1^T = U2^T : right solve with the lower-triangular view U1^T
                rc = trsm_right(h, d_U, ldu, 1, R, 1, 0, W, R, 1, q);
                if (rc) return rc;
            }
            scatter_cols_kernel<<<dim3((R + 255) / 256, ycap(N)), 256, 0, h->stream>>>(d_right, ldright, W, R, d_cp, R, N, R);
            TCI_LAUNCH_CHECK(h);
        }
    }
    return TCI_OK;
}

int tci_solve_right_f64(tci_handle h, const double* d_B, int64_t q, int64_t ldb, const double* d_P, int64_t r, int64_t ldp,
                        double* d_X, int64_t ldx) {
    if (!h) return TCI_ERR_INVALID;
    if (q < 0 || r < 0) return tci_fail(h, TCI_ERR_INVALID, "tci_solve_right_f64: bad sizes");
    if (q == 0 || r == 0) return TCI_OK;
    TCI_CUDA(h, cudaSetDevice(h->device));
    const int Q = (int)q, R = (int)r;
    {
        static const bool small_off = getenv("TCI_SOLVE_SMALL") && atoi(getenv("TCI_SOLVE_SMALL")) == 0;
        if (R <= kSolveMaxR && !small_off) return solve_right_small(h, d_B, Q, ldb, d_P, R, ldp, d_X, ldx);
    }
    const long long ldw = (R + 1) & ~1;
    const size_t permbytes = ((sizeof(int) * (size_t)(2 * R)) + 255) & ~(size_t)255;
    int rc = tci_ws_reserve(h, h->ws_b, permbytes + sizeof(double) * ((size_t)3 * ldw * R + (size_t)Q * R) + 256);
    if (rc) return rc;
    int* d_rp = (int*)h->ws_b.ptr; int* d_cp = d_rp + R;
    double* Pw = (double*)((char*)h->ws_b.ptr + permbytes);
    double* Lw = Pw + (size_t)ldw * R; double* Uw = Lw + (size_t)ldw * R; double* W = Uw + (size_t)ldw * R;
    copy_strided_kernel<<<dim3((R + 255) / 256, ycap(R)), 256, 0, h->stream>>>(d_P, 1, ldp, Pw, 1, ldw, R, R);
    TCI_LAUNCH_CHECK(h);
    std::vector<int64_t> rp(R), cp(R);
    int64_t npiv = 0; double err = 0.0;
    rc = tci_rrlu_inplace_f64(h, Pw, R, R, ldw, R, 0.0, 0.0, 1, rp.data(), cp.data(), &npiv, &err, Lw, ldw, Uw, ldw, nullptr);
    if (rc) return rc;
    if (npiv != R) return tci_fail(h, TCI_ERR_INVALID, "tci_solve_right_f64: pivot matrix is singular (rank %lld of %d)", (long long)npiv, R);
    rc = tci_pinned_reserve(h, permbytes + 64); if (rc) return rc;
    rc = upload_perm(h, rp.data(), R, d_rp, 0); if (rc) return rc;
    rc = upload_perm(h, cp.data(), R, d_cp, sizeof(int) * (size_t)R); if (rc) return rc;
    gather_cols_kernel<<<dim3((Q + 255) / 256, ycap(R)), 256, 0, h->stream>>>(W, Q, d_B, ldb, d_cp, Q, R);
    TCI_LAUNCH_CHECK(h);
    rc = trsm_right(h, Uw, 1, ldw, R, 0, 0, W, 1, Q, Q); if (rc) return rc;      // W <- W U^-1
    rc = trsm_right(h, Lw, 1, ldw, R, 1, 1, W, 1, Q, Q); if (rc) return rc;      // W <- W L^-1
    scatter_cols_kernel<<<dim3((Q + 255) / 256, ycap(R)), 256, 0, h->stream>>>(d_X, ldx, W, Q, d_rp, Q, R, 0);
    TCI_LAUNCH_CHECK(h);
    return TCI_OK;
}

} // extern "C"
