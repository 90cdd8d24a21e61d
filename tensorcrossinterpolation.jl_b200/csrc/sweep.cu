// sweep.cu — device-resident TensorCI2 sweeps: a whole HALF-SWEEP (every bond in order) is ONE kernel launch.
//
// For the low-rank functions the reference is quoted on (README Lorentzian: rank 12; 20-d Lorentzian: rank 18; quantics
// R = 40: rank 17) a Pi matrix has 34^2 ... 360^2 entries and a half-sweep is a chain of N - 1 dependent bond updates of a
// few microseconds each.  Driving that chain from the host — per bond: upload two index tables, K1, rrLU, read the pivots
// back, build the next bond's index sets — costs a launch latency and a host synchronisation per bond, more than the
// arithmetic.  Here the bond loop of sweep2site! (src/tensorci2.jl:1226-1250) runs on the device:
//
//   per bond b (tensorci2.jl:825-930, :full branch)          where
//   -------------------------------------------------------  ----------------------------------------------------------
//   Icombined = union(kronecker(Iset[b], d_b), extraIset)    never materialised: a row of Pi is (I row, site digit) or a
//   Jcombined = union(kronecker(d_b+1, Jset[b+1]), extraJ)   kept extra; first-seen union = "extra whose prefix is not in
//                                                            Iset[b]" (the sets hold unique rows), order kept by a scan
//   Pi = f(Icombined x Jcombined), updatemaxsample!          monoid functor states per row / per column (functors.h), one
//                                                            merge + finish per element, straight into SHARED memory
//   MatrixLUCI(Pi) -> rrlu (matrixlu.jl:346-396)             in shared memory, WITHOUT moving data: rows / columns keep
//                                                            their physical place and two position maps simulate the
//                                                            reference's swaps (the tie rule of submatrixargmax :46-87 is
//                                                            evaluated on positions, i.e. in swapped coordinates); every
//                                                            element gets the identical `a - x*y` (one multiply, one
//                                                            subtract, IEEE division for x), so pivots, errors and index
//                                                            sets are bit-identical to the step kernels of rrlu.cu
//   Iset[b+1] = Icombined[rowindices], Jset[b] = ...          rows gathered from the tables into the output arena; the
//                                                            functor states of the chosen rows are carried to the next
//                                                            bond (no re-fold of the set that moves with the sweep)
//
// Per pivot: two block barriers (winner partials | scaled pivot column), a warp-level argmax in three redux.sync + one
// ballot (64-bit key = bits(|a|^2) + 1, then the smallest (column position, row position) among equals), no swap traffic,
// no extraction inside the loop: everything the factorisation defines stays where it was computed.
//
// Cluster variant (CS = 8 CTAs, distributed shared memory) for Pi of a few hundred KB to ~1.7 MB (the 20-d Lorentzian's
// 360^2): columns are dealt round-robin to the CTAs; per pivot every CTA publishes its best candidate (32-byte record
// into every peer's shared memory) TOGETHER with that candidate's scaled column (a global scratch slot, L2-resident), so
// ONE cluster barrier per pivot decides the winner and makes its column visible.
//
// The host (csrc/tci2_driver.cpp, or a Julia caller) stays the owner of the index sets between launches: it packs them
// into one int32 arena, gets one arena back.  tci_dsweep_* are declared in include/tci_b200.h.
#include "context.h"
#include "functors.h"
#include "solve_small.h"
#include <cooperative_groups.h>
#include <chrono>
#include <cmath>
#include <cstring>
#include <type_traits>
#include <vector>

namespace cg = cooperative_groups;
using namespace tci_functors;

namespace {

// ---- arena formats (int32 words) ------------------------------------------------------------------------------------
// input:  [0] N  [1] forward  [2] s0  [3] s1 (steps [s0, s1) of the sweep order; bond b0 = forward ? s : N-2-s)
//         [4] maxbonddim (<= 0: none)  [5] useextra  [6] reserved  [7] reserved
//         localdims[N]
//         per site b: nI, offI, nJ, offJ, nXI, offXI, nXJ, offXJ      (Iset[b] rows of length b, Jset[b] rows of length N-b-1,
//                                                                      extraIset / extraJset = the history entries, same lengths)
//         tables (row-major, 1-based values)
// output: [0] status (0 ok, 1 stopped: step [1] does not fit, 2 lu.L contains NaNs, 3 lu.U contains NaNs at step [1])
//         [1] next step  [2] words used  [3..7] reserved
//         per step (s - s0), 6 words: r, offA (new Iset[b0+1], r rows of length b0+1), offB (new Jset[b0], rows of length
//                            N-b0-1), offE (pivoterrors(luci): r+1 doubles at an even word offset), m, n (size of Pi)
//         data
constexpr int kHdr = 8;
constexpr int kDesc = 12;
constexpr int kStepRec = 6;
constexpr int kRMax = 512;                   // carried functor states (rows of the set that moves with the sweep)
constexpr int kMaxCS = 8;
constexpr int kMovCap = 2048;                // resident mode: ints per buffer of the moving set's rows (two buffers)
constexpr unsigned FULL = 0xffffffffu;

// Argmax candidates.  key = bits of |a|^2 read as a SIGNED 64-bit integer (non-negative doubles order like their bit patterns);
// NaN, inactive elements and "nothing yet" carry a negative key (NaN never wins, matrixlu.jl:55), so validity is `key >= 0`.
// pos = colpos << 16 | rowpos: the element's place in the reference's swapped layout, which decides ties (smallest column,
// then smallest row) and identifies the element (rowat / colat).
constexpr long long kNoKey = (long long)0x8000000000000000ULL;

struct alignas(16) Win {
    long long key;
    double x;                                // value of the candidate (cluster exchange)
    unsigned pos;
    int owner, pad0, pad1;
};

// argmax over the warp of (key descending, pos ascending): three redux.sync, the result is uniform — no shuffles
__device__ __forceinline__ void warp_argmax(long long& key, unsigned& pos) {
    const unsigned hi = (unsigned)((unsigned long long)key >> 32) ^ 0x80000000u, lo = (unsigned)key;
    const unsigned mhi = __reduce_max_sync(FULL, hi);
    const unsigned mlo = __reduce_max_sync(FULL, hi == mhi ? lo : 0u);
    const bool eq = hi == mhi && lo == mlo;
    pos = __reduce_min_sync(FULL, eq ? pos : 0xffffffffu);
    key = (long long)(((unsigned long long)(mhi ^ 0x80000000u) << 32) | mlo);
}

__device__ __forceinline__ double jl_max(double a, double b) {                  // Julia's max propagates NaN
    if (isnan(a) || isnan(b)) return nan("");
    return a > b ? a : b;
}

template <int CS> __device__ __forceinline__ void cluster_or_block_sync() {
    if constexpr (CS > 1) cg::this_cluster().sync(); else __syncthreads();
}

__device__ __forceinline__ unsigned long long mixh(int site, int v) {           // per-(site, value) hash term of an index row
    unsigned long long z = ((unsigned long long)(unsigned)site << 32) | (unsigned)v;
    z += 0x9e3779b97f4a7c15ULL; z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL; z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
    return z ^ (z >> 31);
}

struct Carve {                               // dynamic shared memory of one bond
    double* A; int lds, mp;                  // mp = length of rowpos / xs incl. padding (a multiple of 128 unless the flat map is used)
    double* xs; double* ys;
    double* piv;
    void* Lst; void* Rst;
    int* rowpos; int* rowat; int* colpos; int* colat;
    int* xkeepI; int* xkeepJ;
    size_t bytes;
};

constexpr int kFlat = 8;                     // flat element -> thread map: at most this many elements per thread

// (only the 256-thread kernels use it: the larger variants are dispatched for blocks that are too big for it anyway, and without the
// flat path their register budget of 128 holds the update loop without spills)
template <int NT> __host__ __device__ inline bool use_flat(int m, int nloc) { return NT <= 256 && (long long)m * nloc <= (long long)kFlat * NT; }

template <class S>
__host__ __device__ inline Carve carve(unsigned char* base, int m, int n, int nloc, int rcap, int nXI, int nXJ, bool flat) {
    Carve c;
    size_t o = 0;
    auto take = [&](size_t bytes) { void* p = base ? (void*)(base + o) : nullptr; o += (bytes + 15) & ~(size_t)15; return p; };
    c.mp = flat ? m : (m + 127) / 128 * 128;             // rowpos / xs: padded to whole groups of 128 rows (position -1: never active)
    c.lds = (flat ? m : (m + 31) / 32 * 32) | 1;         // a group's loads past the end of a column read the next column: harmless
    if (!flat && nloc > 0) nloc += 1;                    // ... and past the last column this spare one
    c.A = (double*)take(sizeof(double) * (size_t)c.lds * (size_t)(nloc > 0 ? nloc : 1));
    c.xs = (double*)take(sizeof(double) * (size_t)c.mp);
    c.ys = (double*)take(sizeof(double) * (size_t)(nloc > 0 ? nloc : 1));
    c.piv = (double*)take(sizeof(double) * (size_t)(rcap + 1));
    c.Lst = take(sizeof(S) * (size_t)m);
    c.Rst = take(sizeof(S) * (size_t)n);
    c.rowpos = (int*)take(sizeof(int) * (size_t)c.mp);
    c.rowat = (int*)take(sizeof(int) * (size_t)m);
    c.colpos = (int*)take(sizeof(int) * (size_t)n);
    c.colat = (int*)take(sizeof(int) * (size_t)n);
    c.xkeepI = (int*)take(sizeof(int) * (size_t)(nXI > 0 ? nXI : 1));
    c.xkeepJ = (int*)take(sizeof(int) * (size_t)(nXJ > 0 ? nXJ : 1));
    c.bytes = o;
    return c;
}

struct SweepParams {
    const int* in;
    int* out;
    long long out_cap;                       // words
    double reltol, abstol;
    unsigned long long* acc;                 // 16 bytes: bits of max |f|, NaN flag (same accumulator as tci_batcheval_acc_f64)
    double* xs_g;                            // cluster variant: [2][CS][mcap] candidate columns
    int mcap;
    int smem_bytes;                          // dynamic shared memory available per CTA
    long long* prof;                         // diagnostics (TCI_DSWEEP_PROF=1): clock64 cycles thread 0 of CTA 0 spent per phase, else NULL
    void* st_g;                              // functor state of every input table row (prologue), Rtot entries
    unsigned long long* hs_g;                // hash of every input table row
    int in_words;                            // length of the input arena
    int resident;                            // single-CTA variant: keep the input arena, the row states and the row hashes in shared
};                                           // memory for the whole launch (every table access is then ~30 cycles instead of an L2 round trip)

// ---- own elements of the matrix --------------------------------------------------------------------------------------------
// Columns j with j % CS == crank live in this CTA (local index jl = j / CS).  Small blocks (at most kFlat elements per thread)
// use a FLAT element -> thread map held in registers for the whole factorisation: perfectly balanced, no padding.  Larger
// blocks deal (column, group of 128 rows) pairs round-robin to the warps: lane l of the warp handles rows g*128 + l + 32 u,
// u < 4, of its pair — four independent elements in flight per thread, rows padded to a multiple of 128 (rowpos = -1 there).
struct Own {
    int m, nown, crank, flat, nslots, ngroups;   // nown = own columns; nslots = used flat slots; ngroups = mp / 128
    int e[kFlat];                                // flat: i | jl << 16, or -1
};

template <int NT, int CS>
__device__ __forceinline__ Own own_elements(const Carve& c, int m, int n, int nloc, int crank) {
    Own o;
    o.m = m; o.crank = crank;
    o.nown = n > crank ? (n - crank + CS - 1) / CS : 0;
    const long long tot = (long long)m * o.nown;
    o.flat = use_flat<NT>(m, nloc);
    o.nslots = o.flat ? (int)((tot + NT - 1) / NT) : 0;
    o.ngroups = c.mp / 128;
#pragma unroll
    for (int q = 0; q < kFlat; ++q) {
        const long long idx = (long long)threadIdx.x + (long long)NT * q;
        o.e[q] = (o.flat && idx < tot) ? (int)(idx % m) | ((int)(idx / m) << 16) : -1;
    }
    return o;
}

// body(i, jl, j) for every own element (used by the fill, where the functor dominates)
template <int NT, int CS, class Body>
__device__ __forceinline__ void foreach_own(const Own& o, Body body) {
    if (o.flat) {
#pragma unroll
        for (int q = 0; q < kFlat; ++q) {
            if (q < o.nslots && o.e[q] >= 0) { const int i = o.e[q] & 0xffff, jl = o.e[q] >> 16; body(i, jl, o.crank + CS * jl); }
        }
    } else {
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        const int npairs = o.nown * o.ngroups;
        for (int pr = warp; pr < npairs; pr += NT / 32) {
            const int jl = pr / o.ngroups, g = pr - jl * o.ngroups, j = o.crank + CS * jl;
#pragma unroll
            for (int u = 0; u < 4; ++u) { const int i = g * 128 + lane + 32 * u; if (i < o.m) body(i, jl, j); }
        }
    }
}

__device__ __forceinline__ long long abs2_key(double v) {                        // bits(v * v); negative for NaN
    return __double_as_longlong(fmax(__dmul_rn(v, v), -1.0));                    // fmax(NaN, x) = x
}

// Candidate update for one element, branch-free (predicates and selects only; written in PTX because the compiler turns the C++ form
// into three branches per element, and the reference's test functions are symmetric, so ties are common).  (q = v^2, pos) replaces
// (bq, bpos) iff q > bq, or q == bq and pos < bpos — the reference's scan order (matrixlu.jl:46-87: largest abs2, then smallest
// column, then smallest row, in swapped coordinates).
// q = v^2 is a non-negative double or NaN, and non-negative doubles order like their bit patterns, so the FP64 compares decide
// exactly what compares of the integer keys (warp_argmax) decide — in 3 predicate instructions.  A NaN q (a NaN entry, or an element the caller poisoned because it is not active) compares false everywhere: never taken.
// bq = -1.0 means "no candidate yet".
__device__ __forceinline__ void take_elem_f(double& bq, unsigned& bpos, double q, unsigned pos) {
    asm("{\n\t.reg .pred eq, b;\n\t"
        "setp.eq.f64 eq, %2, %0;\n\t"
        "setp.lt.and.u32 b, %3, %1, eq;\n\t"
        "setp.gt.or.f64 b, %2, %0, b;\n\t"
        "selp.f64 %0, %2, %0, b;\n\t"
        "selp.b32 %1, %3, %1, b;\n\t}"
        : "+d"(bq), "+r"(bpos) : "d"(q), "r"(pos));
}

// keep (bk, bi, bj) = the first element in the reference's scan order among those seen: larger key, then smaller position
#define TCI_DS_TAKE(k, i, j, bk, bi, bj)                                                                                 \
    if ((k) > (bk)) { (bk) = (k); (bi) = (i); (bj) = (j); }                                                             \
    else if ((k) == (bk) && (k) >= 0) {                                                                                  \
        const unsigned p1__ = ((unsigned)c.colpos[(j)] << 16) | (unsigned)c.rowpos[(i)];                                 \
        const unsigned p0__ = ((unsigned)c.colpos[(bj)] << 16) | (unsigned)c.rowpos[(bi)];                               \
        if (p1__ < p0__) { (bi) = (i); (bj) = (j); }                                                                     \
    }

struct KP { long long key; unsigned pos; unsigned pad; };

// The loop of _optimizerrlu! (matrixlu.jl:356-369) on the matrix in shared memory.  (bkey, bi, bj) holds this thread's candidate
// of the FIRST search (fused into the fill by the caller).  Nothing moves: c.rowat / c.colat (position -> physical index) and
// c.rowpos / c.colpos simulate the reference's swaps; on return the first r entries of rowat / colat are the pivots in order,
// c.piv their values, and every L / U entry sits where it was computed:  L(p, t) = A[rowat[p], colat[t]] (/ piv[t] below the
// diagonal when leftorthogonal), U(t, q) = A[rowat[t], colat[q]] (/ piv[t] right of the diagonal otherwise).  `fwd` is
// leftorthogonal.  Per pivot: [update + search] -> warp argmax -> partials -> barrier 1 -> every warp reduces the partials
// (-> cluster: records + speculative scaled columns, ONE cluster barrier) -> stop rule -> stage pivot column / row, one thread
// swaps the position maps -> barrier 2.
template <int NT, int CS>
__device__ __forceinline__ int smem_rrlu(const Carve& c, const Own& own, const int m, const int n, const int rcap, const int fwd,
                                         const double reltol, const double abstol, long long bkey, int bi, int bj, KP (*s_w)[32],
                                         Win (*s_cl)[kMaxCS], const SweepParams& P, const int crank, double* error_out,
                                         long long* pc = nullptr) {
    constexpr int NW = NT / 32;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    double* A = c.A;
    const int lds = c.lds;
    int np = 0;
    double maxerror = 0.0, error = nan("");
    long long tc0 = clock64();
#define TCI_DS_TICK(slot) { if (pc) { const long long t1__ = clock64(); pc[slot] += t1__ - tc0; tc0 = t1__; } }
    // this thread's candidate as (key, position): the searches below track the position directly — one branch-free
    // lexicographic compare per element instead of a compare with a divergent tie branch that re-reads the position maps
    unsigned bpos = bkey >= 0 ? ((unsigned)c.colpos[bj] << 16) | (unsigned)c.rowpos[bi] : 0xffffffffu;
    double bq = bkey >= 0 ? __longlong_as_double(bkey) : -1.0;                      // the candidate's v^2 (take_elem_f)
    const double qnan = __longlong_as_double(0x7ff8000000000000LL);
    for (;;) {
        bkey = __double_as_longlong(bq);                                            // -1.0 -> negative key: no candidate
        if (bkey < 0) bpos = 0xffffffffu;
        warp_argmax(bkey, bpos);
        KP* sw = s_w[np & 1];
        if (lane == 0) { sw[warp].key = bkey; sw[warp].pos = bpos; }
        __syncthreads();                                                            // barrier 1
        long long gkey = lane < NW ? sw[lane].key : kNoKey;
        unsigned gpos = lane < NW ? sw[lane].pos : 0xffffffffu;
        warp_argmax(gkey, gpos);                                                    // every warp: the CTA's winner
        TCI_DS_TICK(0)
        int pi = -1, pj = -1, owner = crank;
        double piv = nan("");
        if (gkey >= 0) { pi = c.rowat[gpos & 0xffffu]; pj = c.colat[gpos >> 16]; piv = A[(size_t)(pj / CS) * lds + pi]; }
        if constexpr (CS > 1) {
            // publish this CTA's candidate record to every peer and its candidate's scaled column to the scratch slot
            cg::cluster_group cl = cg::this_cluster();
            const int par = np & 1;
            if (tid < CS) {
                Win w; w.key = gkey; w.x = piv; w.pos = gpos; w.owner = crank; w.pad0 = 0; w.pad1 = 0;
                *cl.map_shared_rank(&s_cl[par][crank], tid) = w;
            }
            if (gkey >= 0) {
                double* slot = P.xs_g + ((size_t)par * CS + crank) * P.mcap;
                const double* col = A + (size_t)(pj / CS) * lds;
                if (fwd) { for (int i = tid; i < m; i += NT) slot[i] = col[i] / piv; }
                else     { for (int i = tid; i < m; i += NT) slot[i] = col[i]; }
            }
            TCI_DS_TICK(12)
            cl.sync();                                                              // the one cluster barrier of the step
            TCI_DS_TICK(13)
            {   // the winner among the CS records: one record per lane, the warp-wide argmax (uniform), then the first lane that holds it
                const long long myk = lane < CS ? s_cl[par][lane].key : kNoKey;
                const unsigned myp = lane < CS ? s_cl[par][lane].pos : 0xffffffffu;
                gkey = myk; gpos = myp;
                warp_argmax(gkey, gpos);
                owner = __ffs(__ballot_sync(FULL, lane < CS && myk == gkey && myp == gpos)) - 1;
            }
            piv = s_cl[par][owner].x;
            if (gkey >= 0) { pi = c.rowat[gpos & 0xffffu]; pj = c.colat[gpos >> 16]; }
        }
        TCI_DS_TICK(1)
        if (gkey < 0) { pi = c.rowat[np]; pj = c.colat[np]; piv = nan(""); }        // all-NaN block: the reference keeps (k, k)
        const double err = fabs(piv);                                               // :360
        error = err;
        if ((err < reltol * maxerror || err < abstol) && np > 0) break;             // :363
        maxerror = jl_max(maxerror, err);                                           // :367
        const int k = np;
        // stage the pivot column x[i] = A[i, pj] (/ pivot when leftorthogonal, :303-305) and the pivot row y[jl] = A[pi, j]
        // (/ pivot otherwise, :307-309) of the own columns
        if constexpr (CS > 1) {
            if (gkey >= 0) {
                const double* slot = P.xs_g + ((size_t)(np & 1) * CS + owner) * P.mcap;
                for (int i = tid; i < m; i += NT) c.xs[i] = __ldcg(slot + i);
            } else { for (int i = tid; i < m; i += NT) c.xs[i] = nan(""); }
        } else {
            const double* col = A + (size_t)pj * lds;
            if (fwd) { for (int i = tid; i < m; i += NT) c.xs[i] = col[i] / piv; }
            else     { for (int i = tid; i < m; i += NT) c.xs[i] = col[i]; }
        }
        for (int jl = NT - 1 - tid; jl < own.nown; jl += NT) {                      // (from the last threads: the first ones do x)
            const double y = A[(size_t)jl * lds + pi];
            c.ys[jl] = fwd ? y : y / piv;
        }
        if (tid == NT - 32) {                                                       // swaprow!/swapcol! on the position maps (:254-275)
            const int rk = c.rowat[k], pp = c.rowpos[pi];
            c.rowat[k] = pi; c.rowat[pp] = rk; c.rowpos[rk] = pp; c.rowpos[pi] = k;
            const int ck = c.colat[k], qq = c.colpos[pj];
            c.colat[k] = pj; c.colat[qq] = ck; c.colpos[ck] = qq; c.colpos[pj] = k;
            c.piv[k] = piv;
        }
        __syncthreads();                                                            // barrier 2
        TCI_DS_TICK(2)
        np = k + 1;
        if (np >= rcap) break;                                                      // :356
        bq = -1.0; bpos = 0xffffffffu;
        double bq1 = -1.0; unsigned bpos1 = 0xffffffffu;                            // a second, independent candidate chain (odd elements): the
        // rule is a total order, so the two are merged by one more take at the end — half the dependent chain per element
        // :316-320 + the next search (take_elem_f): loads first, then arithmetic, then the candidate selection
        if (own.flat) {
#pragma unroll
            for (int q0 = 0; q0 < kFlat; q0 += 2) {
                if (q0 < own.nslots) {
                    int rp[2], cp[2]; double av[2], xv[2], yv[2]; double* ap[2]; bool ok[2];
#pragma unroll
                    for (int u = 0; u < 2; ++u) {
                        const int e = own.e[q0 + u];
                        ok[u] = e >= 0;
                        const int ee = ok[u] ? e : 0, jl = ee >> 16, ei = ee & 0xffff;
                        ap[u] = A + (size_t)jl * lds + ei;
                        rp[u] = c.rowpos[ei]; cp[u] = c.colpos[crank + CS * jl]; av[u] = *ap[u]; xv[u] = c.xs[ei]; yv[u] = c.ys[jl];
                    }
#pragma unroll
                    for (int u = 0; u < 2; ++u) {
                        const bool act = ok[u] && min(rp[u], cp[u]) >= np;
                        const double v = __dsub_rn(av[u], __dmul_rn(xv[u], yv[u]));
                        if (act) *ap[u] = v;
                        if (u == 0) take_elem_f(bq, bpos, act ? __dmul_rn(v, v) : qnan, ((unsigned)cp[u] << 16) | (unsigned)rp[u]);
                        else take_elem_f(bq1, bpos1, act ? __dmul_rn(v, v) : qnan, ((unsigned)cp[u] << 16) | (unsigned)rp[u]);
                    }
                }
            }
        } else {
            // warp w keeps row group g = w % ngroups for the whole pass and walks the columns w / ngroups, + NW / ngroups, ...:
            // the pivot column entries and positions of its 4 x 32 rows are loaded once, not once per element
            const int ng = own.ngroups, wpg = NW / ng;                              // warps per row group (the last NW % ng warps idle)
            const int g = warp % ng, jl0 = warp / ng;
            if (jl0 < wpg) {
                const int i0 = g * 128 + lane;
                int rp[4]; double xv[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {                                        // rows that are not active carry x = NaN: their v^2 is NaN,
                    rp[u] = c.rowpos[i0 + 32 * u];                                  // which take_elem_f never takes (padding rows have position -1)
                    xv[u] = rp[u] >= np ? c.xs[i0 + 32 * u] : qnan;
                }
                TCI_DS_TICK(14)
                if (__any_sync(FULL, rp[0] >= np || rp[1] >= np || rp[2] >= np || rp[3] >= np)) {
                    for (int jl = jl0; jl < own.nown; jl += wpg) {
                        const int cpj = c.colpos[crank + CS * jl];
                        if (cpj < np) continue;
                        const unsigned cp16 = (unsigned)cpj << 16;
                        const double y = c.ys[jl];
                        double* col = A + (size_t)jl * lds + i0;
                        double av[4];
#pragma unroll
                        for (int u = 0; u < 4; ++u) av[u] = col[32 * u];
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const double v = __dsub_rn(av[u], __dmul_rn(xv[u], y));
                            if (rp[u] >= np) col[32 * u] = v;
                            if ((u & 1) == 0) take_elem_f(bq, bpos, __dmul_rn(v, v), cp16 | (unsigned)rp[u]);
                            else take_elem_f(bq1, bpos1, __dmul_rn(v, v), cp16 | (unsigned)rp[u]);
                        }
                    }
                }
            }
        }
        take_elem_f(bq, bpos, bq1, bpos1);
        TCI_DS_TICK(3)
    }
    const int r = np;
    if (r >= (m < n ? m : n)) error = 0.0;                                          // :391-393
    __syncthreads();
    *error_out = error;
    return r;
}

// "lu.L contains NaNs" / "lu.U contains NaNs" (matrixlu.jl:376-381): bit 0 / bit 1 of *s_flags (shared, zeroed by the caller).
template <int NT, int CS>
__device__ __forceinline__ void lu_nan_flags(const Carve& c, const int m, const int n, const int r, const int fwd, const int crank, int* s_flags) {
    constexpr int NW = NT / 32;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const double* A = c.A;
    const int lds = c.lds;
    int flags = 0;
    // The scaled entries a / pivot are not formed here: an IEEE quotient is NaN exactly when a or the pivot is NaN, or both are
    // zero, or both are infinite — three compares instead of a division per entry.
    for (int t = warp; t < r; t += NW) {
        const int ct = c.colat[t], rt = c.rowat[t];
        const double pv = c.piv[t];
        const bool pvnan = pv != pv, pvzero = pv == 0.0, pvinf = isinf(pv);
        if (ct % CS == crank) {
            const double* col = A + (size_t)(ct / CS) * lds;
            for (int i = lane; i < m; i += 32) {
                const int p = c.rowpos[i];
                if (p < t) continue;
                const double a = col[i];
                bool bad = a != a;
                if (fwd && p > t) bad = bad || pvnan || (pvzero && a == 0.0) || (pvinf && isinf(a));
                if (bad) flags |= 1;
            }
        }
        for (int j = crank + CS * lane; j < n; j += CS * 32) {
            const int q = c.colpos[j];
            if (q < t) continue;
            const double a = A[(size_t)(j / CS) * lds + rt];
            bool bad = a != a;
            if (!fwd && q > t) bad = bad || pvnan || (pvzero && a == 0.0) || (pvinf && isinf(a));
            if (bad) flags |= 2;
        }
    }
    if (flags) atomicOr(s_flags, flags);
    __syncthreads();
    if constexpr (CS > 1) {
        cg::cluster_group cl = cg::this_cluster();
        if (tid == 0 && *s_flags && crank != 0) atomicOr(cl.map_shared_rank(s_flags, 0), *s_flags);
        cl.sync();
        const int fl = *cl.map_shared_rank(s_flags, 0);
        cl.sync();
        if (tid == 0) *s_flags = fl;
        __syncthreads();
    }
}

// ---- prologue: functor state and hash of every input table row ---------------------------------------------------------------
// Table (site b, w): w = 0 Iset[b] (sites 0 .. b-1), 1 Jset[b] (sites b+1 .. N-1), 2 / 3 the history entries of the same
// shapes.  desc[kDesc b + 8 + w] is the index of the table's first row in the scratch arrays.
template <class F, int NT, int CS>
__device__ __forceinline__ void fold_all_rows(const F& f, const int* in, typename F::State* st, unsigned long long* hs, int crank) {
    typedef typename F::State S;
    const int N = in[0], rtot = in[6];
    const int* desc = in + kHdr + N;
    for (int g = crank * NT + threadIdx.x; g < rtot; g += NT * CS) {
        int lo = 0, hi = 4 * N - 1;                        // last table whose base <= g
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (desc[kDesc * (mid >> 2) + 8 + (mid & 3)] <= g) lo = mid; else hi = mid - 1;
        }
        const int b = lo >> 2, w = lo & 3;
        const int r = g - desc[kDesc * b + 8 + w];
        const bool left = (w & 1) == 0;
        const int len = left ? b : N - b - 1, site0 = left ? 0 : b + 1;
        const int* row = in + desc[kDesc * b + 2 * w + 1] + (size_t)r * len;
        S s = f.zero();
        unsigned long long hh = 0;
        for (int p = 0; p < len; ++p) { const int v = row[p]; f.push(s, site0 + p, v); hh += mixh(site0 + p, v); }
        st[g] = s;
        hs[g] = hh;
    }
}

// One half-sweep.  All CTAs of the cluster run the same control flow on the same inputs (index bookkeeping is replicated,
// it is a few hundred integer operations); only the columns of Pi are divided.
template <class F, int NT, int CS>
__global__ void __launch_bounds__(NT, 1)
dsweep_halfsweep_kernel(F f, SweepParams P) {
    typedef typename F::State S;
    extern __shared__ __align__(16) unsigned char dyn[];
    __shared__ KP s_w[2][32];
    __shared__ Win s_cl[2][kMaxCS];
    __shared__ __align__(16) unsigned char s_carry_raw[kRMax * sizeof(S)];
    __shared__ unsigned long long s_carryh[kRMax];
    __shared__ int s_keepI, s_keepJ, s_flags;
    S* carry = reinterpret_cast<S*>(s_carry_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int crank = 0;
    if constexpr (CS > 1) crank = (int)cg::this_cluster().block_rank();
    const bool writer = crank == 0;

    const int* in = P.in;
    S* stw = reinterpret_cast<S*>(P.st_g);
    unsigned long long* hsw = P.hs_g;
    unsigned char* dynb = dyn;               // where the per-bond carve starts
    int smem_avail = P.smem_bytes;
    int* s_mov[2] = {nullptr, nullptr};      // resident mode: the moving set's rows also stay in shared memory (double-buffered: a bond
    int mpar = 0;                            // reads the previous bond's rows while it writes its own)
    if (CS == 1 && P.resident) {             // tables, states and hashes live in shared memory for the whole launch
        const int rtot = in[6];
        int* s_in = reinterpret_cast<int*>(dyn);
        for (int t = tid; t < P.in_words; t += NT) s_in[t] = in[t];
        size_t o = ((size_t)P.in_words * 4 + 15) & ~(size_t)15;
        stw = reinterpret_cast<S*>(dyn + o); o += ((size_t)rtot * sizeof(S) + 15) & ~(size_t)15;
        hsw = reinterpret_cast<unsigned long long*>(dyn + o); o += ((size_t)rtot * 8 + 15) & ~(size_t)15;
        s_mov[0] = reinterpret_cast<int*>(dyn + o); o += sizeof(int) * kMovCap;
        s_mov[1] = reinterpret_cast<int*>(dyn + o); o += sizeof(int) * kMovCap;
        dynb = dyn + o; smem_avail -= (int)o;
        __syncthreads();
        in = s_in;
    }
    const int N = in[0], fwd = in[1], s0 = in[2], s1 = in[3], maxbond = in[4], useextra = in[5];
    const int* ld = in + kHdr;
    const int* desc = ld + N;
    int* out = P.out;
    long long used = kHdr + (long long)kStepRec * (s1 - s0);
    if (writer && tid == 0) { out[0] = 0; out[1] = s1; out[2] = (int)used; }
    __shared__ long long pcnt[16];                       // diagnostics: thread 0 of CTA 0 only
    long long* pc = (P.prof && writer && tid == 0) ? pcnt : nullptr;
    if (pc) for (int q = 0; q < 16; ++q) pcnt[q] = 0;
    long long tk0 = clock64();
#define TCI_DS_PHASE(slot) { if (pc) { const long long t1__ = clock64(); pc[slot] += t1__ - tk0; tk0 = t1__; } }

    fold_all_rows<F, NT, CS>(f, in, stw, hsw, crank);
    __threadfence();
    cluster_or_block_sync<CS>();
    TCI_DS_PHASE(9)
    const S* stg = stw;
    const unsigned long long* hsg = hsw;

    const int* mov = nullptr;                // the set that moves with the sweep (Iset[b0] forward, Jset[b0+1] backward) ...
    int nmov = -1;                           // ... -1: take it from the input arena (first step of the launch)
    double mx = 0.0; bool isn = false;       // running max |f| of this thread

    for (int s = s0; s < s1; ++s) {
        const int b0 = fwd ? s : N - 2 - s;
        const int dL = ld[b0], dR = ld[b0 + 1];
        const int* dI = desc + kDesc * b0;            // Iset[b0]
        const int* dJ = desc + kDesc * (b0 + 1);      // Jset[b0 + 1]
        const int lenI = b0, lenJ = N - b0 - 2;
        const int* tabI = in + dI[1]; int nI = dI[0];
        const int* tabJ = in + dJ[3]; int nJ = dJ[2];
        const bool carried = nmov >= 0;
        const bool movI = carried && fwd, movJ = carried && !fwd;
        if (movI) { tabI = mov; nI = nmov; }
        if (movJ) { tabJ = mov; nJ = nmov; }
        const S* stI = stg + dI[8]; const unsigned long long* hI = hsg + dI[8];
        const S* stJ = stg + dJ[9]; const unsigned long long* hJ = hsg + dJ[9];
        // extras: extraIset = history Iset[b0 + 1] (rows of length b0 + 1), extraJset = history Jset[b0] (length N - b0 - 1)
        const int* dXI = desc + kDesc * (b0 + 1);
        const int* dXJ = desc + kDesc * b0;
        const int nXI = useextra ? dXI[4] : 0, nXJ = useextra ? dXJ[6] : 0;
        const int* tabXI = in + dXI[5];
        const int* tabXJ = in + dXJ[7];
        const S* stXI = stg + dXI[10]; const unsigned long long* hXI = hsg + dXI[10];
        const S* stXJ = stg + dXJ[11]; const unsigned long long* hXJ = hsg + dXJ[11];
        const int lenXI = b0 + 1, lenXJ = N - b0 - 1;

        // ---- which extras survive the union (tensorci2.jl:842-843): an extra is a duplicate iff its first b0 entries are a row
        //      of Iset[b0] (the kronecker part holds every digit) / its last entries a row of Jset[b0+1].  Hash match, then an
        //      exact compare; the kept ones in order (first-seen union).  Provisional lists at the start of the dynamic buffer.
        int* flg = reinterpret_cast<int*>(dynb);                                    // [nXI + nXJ] 1 = keep
        int* keepI0 = flg + nXI + nXJ;
        int* keepJ0 = keepI0 + (nXI > 0 ? nXI : 1);
        if ((size_t)(2 * (nXI + nXJ) + 4) * sizeof(int) > (size_t)smem_avail) {
            if (writer && tid == 0) { out[0] = 1; out[1] = s; }
            break;
        }
        for (int t = tid; t < nXI + nXJ; t += NT) {
            const bool left = t < nXI;
            const int e = left ? t : t - nXI;
            const int* row = left ? tabXI + (size_t)e * lenXI : tabXJ + (size_t)e * lenXJ;
            const int nb = left ? nI : nJ, blen = left ? lenI : lenJ;
            const int* btab = left ? tabI : tabJ;
            const unsigned long long hp = left ? hXI[e] - mixh(b0, row[lenXI - 1]) : hXJ[e] - mixh(b0 + 1, row[0]);
            const int* sub = left ? row : row + 1;
            const bool bm = left ? movI : movJ;
            const unsigned long long* hb = left ? hI : hJ;
            int cand = -1;                                                          // first base row with the same hash
            for (int r = nb - 1; r >= 0; --r) { const unsigned long long hr = bm ? s_carryh[r] : hb[r]; cand = hr == hp ? r : cand; }
            bool dup = false;
            while (cand >= 0) {                                                     // (a second candidate needs a 64-bit hash collision)
                const int* br = btab + (size_t)cand * blen;
                int diff = 0;
                for (int p = 0; p < blen; ++p) diff |= br[p] ^ sub[p];              // no early exit: the loads are independent
                if (diff == 0) { dup = true; break; }
                int nxt = -1;
                for (int r = nb - 1; r > cand; --r) { const unsigned long long hr = bm ? s_carryh[r] : hb[r]; nxt = hr == hp ? r : nxt; }
                cand = nxt;
            }
            flg[t] = dup ? 0 : 1;
        }
        __syncthreads();
        if (warp < 2) {                                                             // stable compaction: warp 0 left, warp 1 right
            const int cnt = warp == 0 ? nXI : nXJ;
            const int* fl = warp == 0 ? flg : flg + nXI;
            int* keep = warp == 0 ? keepI0 : keepJ0;
            int kept = 0;
            for (int e0 = 0; e0 < cnt; e0 += 32) {
                const int e = e0 + lane;
                const bool kp = e < cnt && fl[e] != 0;
                const unsigned bal = __ballot_sync(FULL, kp);
                if (kp) keep[kept + __popc(bal & ((1u << lane) - 1u))] = e;
                kept += __popc(bal);
            }
            if (lane == 0) { if (warp == 0) s_keepI = kept; else s_keepJ = kept; }
        }
        __syncthreads();
        TCI_DS_PHASE(4)
        const int nKI = s_keepI, nKJ = s_keepJ;
        const int mk = nI * dL, nk = dR * nJ;          // kronecker parts
        const int m = mk + nKI, n = nk + nKJ;
        const int mn = m < n ? m : n;
        const int rcap = (maxbond > 0 && maxbond < mn) ? maxbond : mn;
        const int nloc = (n + CS - 1) / CS;
        // stash the keep lists in registers of the first threads while the buffer is re-carved
        int myKeepI = -1, myKeepJ = -1;
        if (tid < nKI) myKeepI = keepI0[tid];
        if (tid < nKJ) myKeepJ = keepJ0[tid];
        const Carve c = carve<S>(dynb, m, n, nloc, rcap, nXI, nXJ, use_flat<NT>(m, nloc));
        const long long need_out = used + (long long)rcap * (lenXI + lenXJ) + 2 * (rcap + 1) + 2;
        const bool fits = m >= 1 && n >= 1 && m <= 1024 && n <= 65535 && c.bytes <= (size_t)smem_avail && rcap <= kRMax
                          && nKI <= NT && nKJ <= NT && need_out <= P.out_cap && (CS == 1 || m <= P.mcap);
        if (!fits) {                                   // uniform over the cluster: same inputs everywhere
            if (writer && tid == 0) { out[0] = 1; out[1] = s; }
            break;
        }
        __syncthreads();
        if (tid < nKI) c.xkeepI[tid] = myKeepI;
        if (tid < nKJ) c.xkeepJ[tid] = myKeepJ;
        S* Lst = reinterpret_cast<S*>(c.Lst);
        S* Rst = reinterpret_cast<S*>(c.Rst);
        __syncthreads();

        // ---- functor states of every row and every column of Pi
        for (int t = tid; t < m + n; t += NT) {
            if (t < m) {
                const int i = t;
                S st;
                if (i < mk) { const int ir = i % nI; st = movI ? carry[ir] : stI[ir]; f.push(st, b0, i / nI + 1); }
                else st = stXI[c.xkeepI[i - mk]];
                Lst[i] = st;
                c.rowpos[i] = i; c.rowat[i] = i;
            } else {
                const int j = t - m;
                S st;
                if (j < nk) { const int jr = j / dR; st = movJ ? carry[jr] : stJ[jr]; f.push(st, b0 + 1, j % dR + 1); }
                else st = stXJ[c.xkeepJ[j - nk]];
                Rst[j] = st;
                c.colpos[j] = j; c.colat[j] = j;
            }
        }
        for (int t = m + tid; t < c.mp; t += NT) { c.rowpos[t] = -1; c.xs[t] = 0.0; }           // padding rows: never active
        if (tid == 0) s_flags = 0;
        __syncthreads();
        TCI_DS_PHASE(5)

        // ---- Pi (own columns) + the first pivot search (positions are the identity)
        double* A = c.A;
        const int lds = c.lds;
        const Own own = own_elements<NT, CS>(c, m, n, nloc, crank);
        long long bkey = kNoKey; int bi = 0, bj = 0;
        foreach_own<NT, CS>(own, [&](int i, int jl, int j) {
            const double v = f.finish(f.merge(Lst[i], Rst[j]));
            A[(size_t)jl * lds + i] = v;
            const double a = fabs(v);
            if (a != a) isn = true; else mx = a > mx ? a : mx;
            const long long kq = abs2_key(v);
            TCI_DS_TAKE(kq, i, j, bkey, bi, bj)
        });
        TCI_DS_PHASE(6)
        double error;
        const int r = smem_rrlu<NT, CS>(c, own, m, n, rcap, fwd, P.reltol, P.abstol, bkey, bi, bj, s_w, s_cl, P, crank, &error, pc);
        if (pc) { tk0 = clock64(); pc[10] += r; pc[11] += 1; }
        lu_nan_flags<NT, CS>(c, m, n, r, fwd, crank, &s_flags);
        TCI_DS_PHASE(7)
        if (s_flags) {
            if (writer && tid == 0) { out[0] = (s_flags & 1) ? 2 : 3; out[1] = s; }
            break;
        }

        // ---- Iset[b0+1] = Icombined[rowindices], Jset[b0] = Jcombined[colindices] (:915-916), pivoterrors(luci)
        const long long offA = used, offB = offA + (long long)r * lenXI;
        long long offE = offB + (long long)r * lenXJ; offE += offE & 1;
        const long long endw = offE + 2LL * (r + 1);
        const bool movres = CS == 1 && s_mov[0] != nullptr && (long long)r * (fwd ? lenXI : lenXJ) <= kMovCap;
        int* smov = movres ? s_mov[mpar] : nullptr;
        if (writer) {
            int* oa = out + offA; int* ob = out + offB;
            const int nA = r * lenXI, nB = r * lenXJ;                                  // one loop over both sets: their table reads overlap
            for (int e0 = tid; e0 < nA + nB; e0 += NT) {
                if (e0 < nA) {
                    const int e = e0, kk = e / lenXI, p = e % lenXI, i = c.rowat[kk];
                    int v;
                    if (i < mk) v = (p < lenI) ? tabI[(size_t)(i % nI) * lenI + p] : i / nI + 1;
                    else v = tabXI[(size_t)c.xkeepI[i - mk] * lenXI + p];
                    oa[e] = v;
                    if (smov && fwd) smov[e] = v;
                } else {
                    const int e = e0 - nA, kk = e / lenXJ, p = e % lenXJ, j = c.colat[kk];
                    int v;
                    if (j < nk) v = (p == 0) ? j % dR + 1 : tabJ[(size_t)(j / dR) * lenJ + p - 1];
                    else v = tabXJ[(size_t)c.xkeepJ[j - nk] * lenXJ + p];
                    ob[e] = v;
                    if (smov && !fwd) smov[e] = v;
                }
            }
            double* oe = reinterpret_cast<double*>(out + offE);
            for (int t = tid; t <= r; t += NT) oe[t] = t < r ? fabs(c.piv[t]) : error;
            if (tid == 0) {
                int* rec = out + kHdr + kStepRec * (s - s0);
                rec[0] = r; rec[1] = (int)offA; rec[2] = (int)offB; rec[3] = (int)offE; rec[4] = m; rec[5] = n;
                out[2] = (int)endw;
            }
        }
        // carry the functor states and hashes of the chosen rows (forward) / columns (backward) to the next bond
        unsigned long long myh[ (kRMax + NT - 1) / NT ];
#pragma unroll
        for (int q = 0; q < (kRMax + NT - 1) / NT; ++q) {
            const int t = tid + q * NT;
            myh[q] = 0;
            if (t < r) {
                if (fwd) { const int i = c.rowat[t]; myh[q] = i < mk ? (movI ? s_carryh[i % nI] : hI[i % nI]) + mixh(b0, i / nI + 1) : hXI[c.xkeepI[i - mk]]; }
                else { const int j = c.colat[t]; myh[q] = j < nk ? (movJ ? s_carryh[j / dR] : hJ[j / dR]) + mixh(b0 + 1, j % dR + 1) : hXJ[c.xkeepJ[j - nk]]; }
            }
        }
        __syncthreads();                               // every read of the old carried hashes is done
#pragma unroll
        for (int q = 0; q < (kRMax + NT - 1) / NT; ++q) { const int t = tid + q * NT; if (t < r) s_carryh[t] = myh[q]; }
        for (int t = tid; t < r; t += NT) carry[t] = fwd ? Lst[c.rowat[t]] : Rst[c.colat[t]];
        used = endw;
        if (movres) { mov = smov; mpar ^= 1; }         // the next bond reads the rows from shared memory: no fence on the global copy
        else { mov = fwd ? out + offA : out + offB; __threadfence(); }
        nmov = r;
        cluster_or_block_sync<CS>();                   // the new rows and the carried states are visible to every thread / CTA
        TCI_DS_PHASE(8)
    }
    if (pc) for (int q = 0; q < 16; ++q) P.prof[q] = pcnt[q];

    // ---- updatemaxsample! (tensorci2.jl:636-638): one atomic per warp
    unsigned long long b = (unsigned long long)__double_as_longlong(mx);
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) { const unsigned long long o = __shfl_down_sync(FULL, b, d); b = o > b ? o : b; }
    const unsigned anyn = __ballot_sync(FULL, isn);
    if (lane == 0) { if (b) atomicMax(P.acc, b); if (anyn) atomicOr(reinterpret_cast<int*>(P.acc + 1), 1); }
}

// ---- fillsitetensors! (src/globalsearch.jl:202-208 -> setsitetensor! src/tensorci2.jl:599-629) for ALL sites in one launch ------
// CTA = (site b, block of 128 rows of Pi1).  Site b < N-1:  T_b = Pi1 * inv(P),  Pi1 = f((Iset[b] x s_b) x Jset[b]),
// P = f(Iset[b+1] x Jset[b]) (:617-626): the CTA evaluates P into shared memory, factorises it by full-pivot LU
// (solve_small.h, the arithmetic of tci_solve_right_f64), evaluates its rows of Pi1 on the fly and solves.  Last site:
// plain evaluation (:612-615).  Every evaluated value is merged into the max |f| accumulator.
struct FillParams {
    const int* in;                           // input arena (same format as the half-sweep's; extras unused)
    double* const* cores;                    // device array: cores[b] = (chil * d) x chir, dense column-major
    const int* blockmap;                     // per CTA: site, first row
    unsigned long long* acc;
    int* flag;                               // set to 1 + site when a pivot matrix is singular
};

template <class F>
__global__ void __launch_bounds__(kSolveRows)
dsweep_fill_kernel(F f, FillParams P) {
    typedef typename F::State S;
    extern __shared__ __align__(16) unsigned char dynf[];
    __shared__ int rp[kSolveMaxR], cp[kSolveMaxR];
    __shared__ SmallLuScratch<kSolveRows> scratch;
    __shared__ __align__(16) unsigned char s_R_raw[kSolveMaxR * sizeof(S)];
    __shared__ __align__(16) unsigned char s_L_raw[kSolveMaxR * sizeof(S)];
    S* Rst = reinterpret_cast<S*>(s_R_raw);
    S* LPst = reinterpret_cast<S*>(s_L_raw);
    const int tid = threadIdx.x, lane = tid & 31;
    const int* in = P.in;
    const int N = in[0];
    const int* ld = in + kHdr; const int* desc = ld + N;
    const int b0 = P.blockmap[2 * blockIdx.x], row0 = P.blockmap[2 * blockIdx.x + 1];
    const int d = ld[b0];
    const int nI = desc[kDesc * b0 + 0], nJ = desc[kDesc * b0 + 2];
    const int* tabI = in + desc[kDesc * b0 + 1];
    const int* tabJ = in + desc[kDesc * b0 + 3];
    const int lenI = b0, lenJ = N - b0 - 1;
    const int q = nI * d;
    double* T = P.cores[b0];
    double mx = 0.0; bool isn = false;
    const int row = row0 + tid;
    S st = f.zero();
    if (row < q) {
        const int* r_ = tabI + (size_t)(row % nI) * lenI;
        for (int p = 0; p < lenI; ++p) f.push(st, p, r_[p]);
        f.push(st, b0, row / nI + 1);
    }
    if (b0 == N - 1) {                                                              // :612-615
        for (int j = 0; j < nJ; ++j) {
            if (row < q) {
                const double v = f.finish(f.merge(st, f.zero()));
                T[(size_t)j * q + row] = v;
                const double a = fabs(v);
                if (a != a) isn = true; else mx = a > mx ? a : mx;
            }
        }
    } else {
        const int r = nJ;                                                           // == |Iset[b0+1]| (checked by the host, :623)
        const int lds = r | 1;
        double* Sm = reinterpret_cast<double*>(dynf);
        double* W = Sm + (size_t)lds * r;
        const int* tabI1 = in + desc[kDesc * (b0 + 1) + 1];
        for (int t = tid; t < 2 * r; t += kSolveRows) {
            S s2 = f.zero();
            if (t < r) { const int* r_ = tabJ + (size_t)t * lenJ; for (int p = 0; p < lenJ; ++p) f.push(s2, b0 + 1 + p, r_[p]); Rst[t] = s2; }
            else { const int* r_ = tabI1 + (size_t)(t - r) * (b0 + 1); for (int p = 0; p <= b0; ++p) f.push(s2, p, r_[p]); LPst[t - r] = s2; }
        }
        __syncthreads();
        for (int e = tid; e < r * r; e += kSolveRows) {
            const int i = e % r, j = e / r;
            const double v = f.finish(f.merge(LPst[i], Rst[j]));
            Sm[j * lds + i] = v;
            const double a = fabs(v);
            if (a != a) isn = true; else mx = a > mx ? a : mx;
        }
        __syncthreads();
        const bool singular = small_lu_fullpivot<kSolveRows>(Sm, lds, r, rp, cp, scratch);
        if (singular) { if (tid == 0) atomicMax(P.flag, 1 + b0); }
        else {
            double* W2 = W;
            for (int k = 0; k < r; ++k) {                                           // B[:, cp]: Pi1[row, cp[k]]
                double v = 0.0;
                if (row < q) {
                    v = f.finish(f.merge(st, Rst[cp[k]]));
                    const double a = fabs(v);
                    if (a != a) isn = true; else mx = a > mx ? a : mx;
                }
                W2[k * kSolveRows + tid] = v;
            }
            small_solve_row<kSolveRows>(W2, Sm, lds, r);
            if (row < q) for (int k = 0; k < r; ++k) T[(size_t)rp[k] * q + row] = W2[k * kSolveRows + tid];
        }
    }
    unsigned long long b = (unsigned long long)__double_as_longlong(mx);
#pragma unroll
    for (int dd = 16; dd > 0; dd >>= 1) { const unsigned long long o = __shfl_down_sync(FULL, b, dd); b = o > b ? o : b; }
    const unsigned anyn = __ballot_sync(FULL, isn);
    if (lane == 0) { if (b) atomicMax(P.acc, b); if (anyn) atomicOr(reinterpret_cast<int*>(P.acc + 1), 1); }
}

// ---- sweep1site!(tci, f, :forward; reltol, abstol, maxbonddim, updatetensors=true) (src/tensorci2.jl:659-725) in ONE launch ----
// Per bond b: Pi = f((Iset[b] x s_b) x Jset[b]) (:681-686), MatrixLUCI (:688-694), Iset[b+1] / Jset[b] from the pivots
// (:696-698), T_b = left(luci) = P_row^T [I_r ; L2 inv(L1)] (src/matrixluci.jl:194-213), NaN check (:705-707); the last tensor
// by evaluation (:719-723).  One CTA; the same swap-free rrLU as the half-sweep kernel.
struct Sweep1Params {
    SweepParams sp;
    double* const* cores;
    const long long* corecap;                // doubles available in cores[b]
};

template <class F, int NT>
__global__ void __launch_bounds__(NT, 1)
dsweep_sweep1site_kernel(F f, Sweep1Params Q) {
    typedef typename F::State S;
    const SweepParams& P = Q.sp;
    extern __shared__ __align__(16) unsigned char dyn[];
    __shared__ KP s_w[2][32];
    __shared__ Win s_cl[2][kMaxCS];
    __shared__ __align__(16) unsigned char s_carry_raw[kRMax * sizeof(S)];
    __shared__ int s_flags;
    S* carry = reinterpret_cast<S*>(s_carry_raw);
    const int tid = threadIdx.x, lane = tid & 31;
    const int* in = P.in;
    const int N = in[0], maxbond = in[4];
    const int* ld = in + kHdr; const int* desc = ld + N;
    int* out = P.out;
    long long used = kHdr + (long long)kStepRec * (N - 1);
    if (tid == 0) { out[0] = 0; out[1] = N - 1; out[2] = (int)used; }
    __shared__ long long pcnt1[8];                       // diagnostics (TCI_DSWEEP_PROF=1): cycles of thread 0 per phase
    long long* pc = (P.prof && tid == 0) ? pcnt1 : nullptr;
    if (pc) for (int q = 0; q < 8; ++q) pcnt1[q] = 0;
    long long tk0 = clock64();
#define TCI_S1_PHASE(slot) { if (pc) { const long long t1__ = clock64(); pc[slot] += t1__ - tk0; tk0 = t1__; } }
    fold_all_rows<F, NT, 1>(f, in, reinterpret_cast<S*>(P.st_g), P.hs_g, 0);
    __syncthreads();
    TCI_S1_PHASE(0)
    const S* stg = reinterpret_cast<const S*>(P.st_g);
    const int* mov = in + desc[1]; int nmov = desc[0];       // Iset[0]
    bool carried = false;
    double mx = 0.0; bool isn = false;
    bool stopped = false;
    for (int b0 = 0; b0 < N - 1; ++b0) {
        const int d = ld[b0];
        const int* tabI = mov; const int nI = nmov; const int lenI = b0;
        const int nJ = desc[kDesc * b0 + 2]; const int* tabJ = in + desc[kDesc * b0 + 3]; const int lenJ = N - b0 - 1;
        const S* stI = stg + desc[kDesc * b0 + 8];           // only used while the input Iset[0] is current
        const S* stJ = stg + desc[kDesc * b0 + 9];
        const int m = nI * d, n = nJ, mn = m < n ? m : n;
        const int rcap = (maxbond > 0 && maxbond < mn) ? maxbond : mn;
        Carve c = carve<S>(dyn, m, n, n, rcap, 0, 0, use_flat<NT>(m, n));
        double* L1s = reinterpret_cast<double*>(dyn + c.bytes);                     // scaled unit-lower factor of the pivot block
        const size_t bytes = c.bytes + sizeof(double) * (size_t)rcap * rcap;
        const long long need_out = used + (long long)rcap * (lenI + 1 + lenJ) + 2 * (rcap + 1) + 2;
        const bool fits = m >= 1 && n >= 1 && m <= 1024 && n <= 65535 && bytes <= (size_t)P.smem_bytes && rcap <= kRMax
                          && need_out <= P.out_cap && (long long)m * rcap <= Q.corecap[b0];
        if (!fits) { if (tid == 0) { out[0] = 1; out[1] = b0; } stopped = true; break; }
        S* Lst = reinterpret_cast<S*>(c.Lst);
        S* Rst = reinterpret_cast<S*>(c.Rst);
        __syncthreads();
        for (int t = tid; t < m + n; t += NT) {
            if (t < m) {
                const int ir = t % nI;
                S st = carried ? carry[ir] : stI[ir];
                f.push(st, b0, t / nI + 1);
                Lst[t] = st; c.rowpos[t] = t; c.rowat[t] = t;
            } else {
                const int j = t - m;
                Rst[j] = stJ[j]; c.colpos[j] = j; c.colat[j] = j;
            }
        }
        for (int t = m + tid; t < c.mp; t += NT) { c.rowpos[t] = -1; c.xs[t] = 0.0; }
        if (tid == 0) s_flags = 0;
        __syncthreads();
        TCI_S1_PHASE(1)
        double* A = c.A; const int lds = c.lds;
        const Own own = own_elements<NT, 1>(c, m, n, n, 0);
        long long bkey = kNoKey; int bi = 0, bj = 0;
        foreach_own<NT, 1>(own, [&](int i, int jl, int j) {
            const double v = f.finish(f.merge(Lst[i], Rst[j]));
            A[(size_t)jl * lds + i] = v;
            const double a = fabs(v);
            if (a != a) isn = true; else mx = a > mx ? a : mx;
            const long long kq = abs2_key(v);
            TCI_DS_TAKE(kq, i, j, bkey, bi, bj)
        });
        TCI_S1_PHASE(2)
        double error;
        const int r = smem_rrlu<NT, 1>(c, own, m, n, rcap, 1, P.reltol, P.abstol, bkey, bi, bj, s_w, s_cl, P, 0, &error);
        TCI_S1_PHASE(3)
        lu_nan_flags<NT, 1>(c, m, n, r, 1, 0, &s_flags);
        TCI_S1_PHASE(4)
        if (s_flags) { if (tid == 0) { out[0] = (s_flags & 1) ? 2 : 3; out[1] = b0; } stopped = true; break; }
        // ---- new index sets and pivot errors
        const int lenA = b0 + 1, lenB = lenJ;
        const long long offA = used, offB = offA + (long long)r * lenA;
        long long offE = offB + (long long)r * lenB; offE += offE & 1;
        const long long endw = offE + 2LL * (r + 1);
        {
            int* oa = out + offA; int* ob = out + offB;
            for (int e = tid; e < r * lenA; e += NT) {
                const int kk = e / lenA, p = e % lenA, i = c.rowat[kk];
                oa[e] = (p < lenI) ? tabI[(size_t)(i % nI) * lenI + p] : i / nI + 1;
            }
            for (int e = tid; e < r * lenB; e += NT) { const int kk = e / lenB, p = e % lenB; ob[e] = tabJ[(size_t)c.colat[kk] * lenJ + p]; }
            double* oe = reinterpret_cast<double*>(out + offE);
            for (int t = tid; t <= r; t += NT) oe[t] = t < r ? fabs(c.piv[t]) : error;
            if (tid == 0) {
                int* rec = out + kHdr + kStepRec * b0;
                rec[0] = r; rec[1] = (int)offA; rec[2] = (int)offB; rec[3] = (int)offE; rec[4] = m; rec[5] = n;
                out[2] = (int)endw;
            }
        }
        TCI_S1_PHASE(5)
        // ---- T_b = left(luci): rows at pivot positions are unit vectors, the others L2 inv(L1) by back substitution, in place
        for (int e = tid; e < r * r; e += NT) {
            const int j = e / r, t = e % r;                                         // L1(j, t), j > t
            if (j > t) L1s[j * r + t] = A[(size_t)c.colat[t] * lds + c.rowat[j]] / c.piv[t];
        }
        __syncthreads();
        double* T = Q.cores[b0];
        bool tnan = false;
        for (int i = tid; i < m; i += NT) {
            const int p = c.rowpos[i];
            if (p < r) { for (int k = 0; k < r; ++k) T[(size_t)k * m + i] = (k == p) ? 1.0 : 0.0; continue; }
            for (int t = r - 1; t >= 0; --t) {
                double x = A[(size_t)c.colat[t] * lds + i] / c.piv[t];              // L2(p, t)
                for (int j = t + 1; j < r; ++j) x = x - A[(size_t)c.colat[j] * lds + i] * L1s[j * r + t];
                A[(size_t)c.colat[t] * lds + i] = x;                                // X(p, t), final
                if (x != x) tnan = true;
                T[(size_t)t * m + i] = x;
            }
        }
        if (tnan) atomicOr(&s_flags, 4);
        for (int t = tid; t < r; t += NT) carry[t] = Lst[c.rowat[t]];
        __syncthreads();
        if (s_flags) { if (tid == 0) { out[0] = 4; out[1] = b0; } stopped = true; break; }        // "Error: NaN in tensor T[b]" :705-707
        used = endw; mov = out + offA; nmov = r; carried = true;
        __threadfence();
        __syncthreads();
        TCI_S1_PHASE(6)
    }
    if (pc) for (int q = 0; q < 8; ++q) P.prof[q] = pcnt1[q];
    if (!stopped) {                                                                // last tensor by evaluation (:719-723)
        const int b0 = N - 1, d = ld[b0];
        const int nI = nmov;
        const int nJ = desc[kDesc * b0 + 2];
        const S* stI = stg + desc[kDesc * b0 + 8];
        const long long q = (long long)nI * d;
        if (q * nJ > Q.corecap[b0]) { if (tid == 0) { out[0] = 1; out[1] = N - 1; } }
        else {
            double* T = Q.cores[b0];
            for (long long e = tid; e < q * nJ; e += NT) {
                const int row = (int)(e % q);
                const int ir = row % nI;
                S st = carried ? carry[ir] : stI[ir];
                f.push(st, b0, row / nI + 1);
                const double v = f.finish(f.merge(st, f.zero()));
                T[e] = v;
                const double a = fabs(v);
                if (a != a) isn = true; else mx = a > mx ? a : mx;
            }
        }
    }
    unsigned long long b = (unsigned long long)__double_as_longlong(mx);
#pragma unroll
    for (int dd = 16; dd > 0; dd >>= 1) { const unsigned long long o = __shfl_down_sync(FULL, b, dd); b = o > b ? o : b; }
    const unsigned anyn = __ballot_sync(FULL, isn);
    if (lane == 0) { if (b) atomicMax(P.acc, b); if (anyn) atomicOr(reinterpret_cast<int*>(P.acc + 1), 1); }
}

} // namespace

// ---- host side ---------------------------------------------------------------------------------------------------------
namespace {

struct SweepShape { int maxm = 0, maxn = 0; long long maxelems = 0; };

SweepShape predict(const int32_t* in) {
    // Pi sizes of the steps of this launch from the INPUT set sizes (the set that moves with the sweep may grow or shrink;
    // the kernel re-checks every bond and stops when one does not fit)
    SweepShape sh;
    const int N = in[0], fwd = in[1], s0 = in[2], s1 = in[3], useextra = in[5];
    const int32_t* ld = in + kHdr; const int32_t* desc = ld + N;
    for (int s = s0; s < s1; ++s) {
        const int b0 = fwd ? s : N - 2 - s;
        const int nI = desc[kDesc * b0 + 0], nJ = desc[kDesc * (b0 + 1) + 2];
        const int nXI = useextra ? desc[kDesc * (b0 + 1) + 4] : 0, nXJ = useextra ? desc[kDesc * b0 + 6] : 0;
        const int m = nI * ld[b0] + nXI, n = ld[b0 + 1] * nJ + nXJ;
        sh.maxm = std::max(sh.maxm, m); sh.maxn = std::max(sh.maxn, n);
        sh.maxelems = std::max(sh.maxelems, (long long)m * n);
    }
    return sh;
}

template <class F, int NT, int CS>
int launch_halfsweep(tci_handle h, const F& f, const SweepParams& P, int dyn_bytes) {
    auto kern = dsweep_halfsweep_kernel<F, NT, CS>;
    static thread_local int optin_done = 0;              // per instantiation
    if (!optin_done) {
        TCI_CUDA(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn_bytes));
        optin_done = 1;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(CS); cfg.blockDim = dim3(NT); cfg.dynamicSmemBytes = (size_t)dyn_bytes; cfg.stream = h->stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = CS; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = CS > 1 ? 1 : 0;
    TCI_CUDA(h, cudaLaunchKernelEx(&cfg, kern, f, P));
    h->launches++;
    return TCI_OK;
}

constexpr int kDynBytes = 208 * 1024;                    // + up to ~15 KB static (carried states and hashes, candidate records) <= 227 KB

template <class F>
int run_halfsweep(tci_handle h, const F& f, SweepParams P, const SweepShape& sh, int force_cs) {
    P.smem_bytes = kDynBytes;
    const size_t lds1 = (size_t)((sh.maxm + 31) / 32 * 32) | 1;                // padded column length, one spare column
    const size_t one = sizeof(double) * lds1 * ((size_t)sh.maxn + 1) + 48 * (lds1 + 128 + (size_t)sh.maxn) + 4096;
    int cs = one <= (size_t)kDynBytes ? 1 : kMaxCS;
    if (force_cs == 1 || force_cs == kMaxCS) cs = force_cs;
    // resident tables (single CTA): arena + state + 8 B hash per row, if 1.5 x the predicted Pi carve still fits beside them (the
    // moving set may grow; a bond that does not fit the reduced space is handed back to the host like any other that does not fit)
    const size_t nrows = (size_t)P.resident;                                     // the row count on entry
    const size_t res = (((size_t)P.in_words * 4 + 15) & ~(size_t)15) + ((nrows * sizeof(typename F::State) + 15) & ~(size_t)15)
                       + ((nrows * 8 + 15) & ~(size_t)15) + 2 * sizeof(int) * kMovCap;
    static const bool no_res = getenv("TCI_DSWEEP_NORESIDENT") != nullptr;
    P.resident = (cs == 1 && !no_res && res + one + one / 2 <= (size_t)kDynBytes) ? 1 : 0;
    // Block size of the larger variants: 512 threads, not 1024.  At 1024 the 64-register ceiling spills in the update loop (120 B of
    // local memory per thread in the single-CTA kernel, 336 B in the cluster kernel); at 512 nothing spills and the update, which is
    // issue-bound, loses nothing: C1 2.83 -> 2.52 ms, C3 13.25 -> 11.7 ms (A/B in one run).
    if (cs == 1) {
        const char* ne = getenv("TCI_DSWEEP_NT");          // experiments: force the block size of the single-CTA variant (256 | 512)
        const int nt = ne ? atoi(ne) : (sh.maxelems <= 4096 ? 256 : 512);
        if (nt == 256) return launch_halfsweep<F, 256, 1>(h, f, P, kDynBytes);
        return launch_halfsweep<F, 512, 1>(h, f, P, kDynBytes);
    }
    return launch_halfsweep<F, 512, kMaxCS>(h, f, P, kDynBytes);
}


// staging of one launch: arena words, then (8-byte aligned) extra blocks; ONE host-to-device copy
struct Staging {
    std::vector<unsigned char> bytes;
    size_t add(const void* p, size_t n, size_t align = 8) {
        size_t off = (bytes.size() + align - 1) / align * align;
        bytes.resize(off + n);
        if (n) memcpy(bytes.data() + off, p, n);
        return off;
    }
};

template <class F>
int launch_fill(tci_handle h, const F& f, const FillParams& P, int nblocks, int rmax) {
    auto kern = dsweep_fill_kernel<F>;
    const size_t smem = sizeof(double) * ((size_t)(rmax | 1) * rmax + (size_t)rmax * kSolveRows);
    static thread_local int optin_done = 0;
    if (smem > 48 * 1024 && !optin_done) {
        const size_t most = sizeof(double) * ((size_t)(kSolveMaxR | 1) * kSolveMaxR + (size_t)kSolveMaxR * kSolveRows);
        TCI_CUDA(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)most));
        optin_done = 1;
    }
    kern<<<nblocks, kSolveRows, smem, h->stream>>>(f, P);
    TCI_LAUNCH_CHECK(h);
    return TCI_OK;
}

template <class F>
int launch_sweep1(tci_handle h, const F& f, Sweep1Params Q, long long maxelems) {
    Q.sp.smem_bytes = kDynBytes;
    static const int force_nt = getenv("TCI_DSWEEP_S1NT") ? atoi(getenv("TCI_DSWEEP_S1NT")) : 0;     // experiments
    if (force_nt == 256 || (force_nt != 512 && maxelems <= 8192)) {       // measured at C3 (360 x 18 per bond): 256 threads 1.06 ms, 1024 threads 1.28 ms
        auto kern = dsweep_sweep1site_kernel<F, 256>;
        static thread_local int optin_a = 0;
        if (!optin_a) { TCI_CUDA(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kDynBytes)); optin_a = 1; }
        kern<<<1, 256, kDynBytes, h->stream>>>(f, Q);
    } else {
        auto kern = dsweep_sweep1site_kernel<F, 512>;
        static thread_local int optin_b = 0;
        if (!optin_b) { TCI_CUDA(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kDynBytes)); optin_b = 1; }
        kern<<<1, 512, kDynBytes, h->stream>>>(f, Q);
    }
    TCI_LAUNCH_CHECK(h);
    return TCI_OK;
}

int check_arena(tci_handle h, tci_functor f, const int32_t* h_in, int64_t in_words, const char* who) {
    if (!f || !h_in || in_words < kHdr) return tci_fail(h, TCI_ERR_INVALID, "%s: bad arguments", who);
    if (!tci_dsweep_supported(h, f)) return tci_fail(h, TCI_ERR_INVALID, "%s: needs a built-in device functor", who);
    const int N = h_in[0];
    if (N < 2 || N > 128 || in_words < kHdr + (int64_t)N * (1 + kDesc)) return tci_fail(h, TCI_ERR_INVALID, "%s: malformed input arena", who);
    if (f->kind == TCI_F_DENSE && f->ndims != N) return tci_fail(h, TCI_ERR_INVALID, "dense functor has %d dims, localdims has %d", f->ndims, N);
    const int32_t* desc = h_in + kHdr + N;
    int64_t rows = 0;
    for (int b = 0; b < N; ++b)
        for (int w = 0; w < 4; ++w) {
            if (desc[kDesc * b + 8 + w] != rows) return tci_fail(h, TCI_ERR_INVALID, "%s: row bases of the input arena are inconsistent", who);
            rows += desc[kDesc * b + 2 * w];
        }
    if (rows != h_in[6]) return tci_fail(h, TCI_ERR_INVALID, "%s: total row count of the input arena is inconsistent", who);
    return TCI_OK;
}

// read the append-only output arena back: a prefix first, the rest only when more was written
int read_out(tci_handle h, const int* d_out, int32_t* h_out, int64_t out_cap_words, int64_t* h_used_words, const char* who) {
    const int64_t first = std::min<int64_t>(out_cap_words, 16384);
    int rc = tci_pinned_reserve(h, (size_t)first * 4); if (rc) return rc;
    TCI_CUDA(h, cudaMemcpyAsync(h->pinned, d_out, (size_t)first * 4, cudaMemcpyDeviceToHost, h->stream));
    TCI_CUDA(h, cudaStreamSynchronize(h->stream));
    memcpy(h_out, h->pinned, (size_t)first * 4);
    const int64_t used = h_out[2];
    if (used > out_cap_words) return tci_fail(h, TCI_ERR_INVALID, "%s: output arena overrun", who);
    if (used > first) {
        TCI_CUDA(h, cudaMemcpyAsync(h_out + first, d_out + first, (size_t)(used - first) * 4, cudaMemcpyDeviceToHost, h->stream));
        TCI_CUDA(h, cudaStreamSynchronize(h->stream));
    }
    if (h_used_words) *h_used_words = used;
    return TCI_OK;
}

#define TCI_DS_DISPATCH(rc, f, CALL)                                                                                         \
    switch ((f)->kind) {                                                                                                     \
    case TCI_F_LORENTZ: { LorentzF F_{(f)->p[0]}; rc = CALL; } break;                                                        \
    case TCI_F_SUM: { SumF F_{(f)->p[0]}; rc = CALL; } break;                                                                \
    case TCI_F_QEXPSIN: { QuanticsF<TCI_F_QEXPSIN> F_{(f)->p[0], (f)->p[1], (f)->p[2]}; rc = CALL; } break;                  \
    case TCI_F_QEXPPLUSSIN: { QuanticsF<TCI_F_QEXPPLUSSIN> F_{(f)->p[0], (f)->p[1], (f)->p[2]}; rc = CALL; } break;          \
    case TCI_F_QEXP2: { QuanticsF<TCI_F_QEXP2> F_{(f)->p[0], (f)->p[1], (f)->p[2]}; rc = CALL; } break;                      \
    case TCI_F_HASH: { HashF F_{(f)->p[0], (f)->p[1]}; rc = CALL; } break;                                                   \
    case TCI_F_DENSE: { DenseF F_{(f)->d_table, (f)->d_strides}; rc = CALL; } break;                                         \
    default: rc = tci_fail(h, TCI_ERR_INVALID, "unknown functor kind %d", (f)->kind);                                        \
    }

} // namespace

extern "C" {

int tci_dsweep_supported(tci_handle h, tci_functor f) {
    (void)h;
    if (!f) return 0;
    switch (f->kind) {
    case TCI_F_LORENTZ: case TCI_F_SUM: case TCI_F_QEXPSIN: case TCI_F_QEXPPLUSSIN: case TCI_F_QEXP2: case TCI_F_HASH: case TCI_F_DENSE: return 1;
    default: return 0;
    }
}

int tci_dsweep_halfsweep_f64(tci_handle h, tci_functor f, const int32_t* h_in, int64_t in_words, double reltol, double abstol,
                             void* d_acc16, int32_t* h_out, int64_t out_cap_words, int64_t* h_used_words) {
    if (!h) return TCI_ERR_INVALID;
    { int rc0 = check_arena(h, f, h_in, in_words, "tci_dsweep_halfsweep_f64"); if (rc0) return rc0; }
    if (!h_out || !d_acc16 || out_cap_words < kHdr) return tci_fail(h, TCI_ERR_INVALID, "tci_dsweep_halfsweep_f64: bad arguments");
    const int N = h_in[0], s0 = h_in[2], s1 = h_in[3];
    if (s0 < 0 || s1 > N - 1 || s0 > s1) return tci_fail(h, TCI_ERR_INVALID, "tci_dsweep_halfsweep_f64: malformed input arena");
    if (out_cap_words < kHdr + (int64_t)kStepRec * (s1 - s0)) return tci_fail(h, TCI_ERR_INVALID, "tci_dsweep_halfsweep_f64: output arena too small");
    TCI_CUDA(h, cudaSetDevice(h->device));
    static const bool hostprof = getenv("TCI_DSWEEP_HOSTPROF") != nullptr;          // diagnostics: host wall time per stage of this call
    auto now = [] { return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double th0 = hostprof ? now() : 0.0;
    const SweepShape sh = predict(h_in);
    const int mcap = ((std::max(sh.maxm, 1) * 2 + 63) / 64) * 64;                  // room for the moving set to grow
    const size_t in_b = ((size_t)in_words * 4 + 255) & ~(size_t)255, out_b = ((size_t)out_cap_words * 4 + 255) & ~(size_t)255;
    const size_t xs_b = sizeof(double) * 2 * kMaxCS * (size_t)mcap;
    const size_t rtot = (size_t)std::max(h_in[6], 0), st_b = (rtot * 16 + 255) & ~(size_t)255, hs_b = (rtot * 8 + 255) & ~(size_t)255;
    int rc = tci_ws_reserve(h, h->ws_c, in_b + out_b + xs_b + 512 + st_b + hs_b); if (rc) return rc;
    unsigned char* base = (unsigned char*)h->ws_c.ptr;
    SweepParams P;
    P.in = (const int*)base; P.out = (int*)(base + in_b); P.out_cap = out_cap_words; P.reltol = reltol; P.abstol = abstol;
    P.acc = (unsigned long long*)d_acc16; P.xs_g = (double*)(base + in_b + out_b); P.mcap = mcap; P.smem_bytes = 0;
    const bool prof = getenv("TCI_DSWEEP_PROF") != nullptr;
    P.prof = prof ? (long long*)(base + in_b + out_b + xs_b) : nullptr;
    P.st_g = base + in_b + out_b + xs_b + 512; P.hs_g = (unsigned long long*)(base + in_b + out_b + xs_b + 512 + st_b);
    P.in_words = (int)in_words; P.resident = (int)rtot;    // run_halfsweep turns the row count into the residency decision
    const double th1 = hostprof ? now() : 0.0;
    TCI_CUDA(h, cudaMemcpyAsync(base, h_in, (size_t)in_words * 4, cudaMemcpyHostToDevice, h->stream));
    const double th2 = hostprof ? now() : 0.0;
    const char* fe = getenv("TCI_DSWEEP_CLUSTER");         // tests: 1 / 8 force the single-CTA / cluster variant
    const int force_cs = fe ? atoi(fe) : 0;
    switch (f->kind) {
    case TCI_F_LORENTZ: rc = run_halfsweep(h, LorentzF{f->p[0]}, P, sh, force_cs); break;
    case TCI_F_SUM: rc = run_halfsweep(h, SumF{f->p[0]}, P, sh, force_cs); break;
    case TCI_F_QEXPSIN: rc = run_halfsweep(h, QuanticsF<TCI_F_QEXPSIN>{f->p[0], f->p[1], f->p[2]}, P, sh, force_cs); break;
    case TCI_F_QEXPPLUSSIN: rc = run_halfsweep(h, QuanticsF<TCI_F_QEXPPLUSSIN>{f->p[0], f->p[1], f->p[2]}, P, sh, force_cs); break;
    case TCI_F_QEXP2: rc = run_halfsweep(h, QuanticsF<TCI_F_QEXP2>{f->p[0], f->p[1], f->p[2]}, P, sh, force_cs); break;
    case TCI_F_HASH: rc = run_halfsweep(h, HashF{f->p[0], f->p[1]}, P, sh, force_cs); break;
    case TCI_F_DENSE: rc = run_halfsweep(h, DenseF{f->d_table, f->d_strides}, P, sh, force_cs); break;
    default: rc = tci_fail(h, TCI_ERR_INVALID, "unknown functor kind %d", f->kind);
    }
    if (rc) return rc;
    const double th3 = hostprof ? now() : 0.0;
    if (hostprof) cudaStreamSynchronize(h->stream);
    const double th4 = hostprof ? now() : 0.0;
    rc = read_out(h, P.out, h_out, out_cap_words, h_used_words, "tci_dsweep_halfsweep_f64");
    if (hostprof)
        fprintf(stderr, "dsweep host (us): check+reserve=%.1f h2d(%lld words)=%.1f launch=%.1f kernel_wait=%.1f read_out(%lld words)=%.1f\n", th1 - th0,
                (long long)in_words, th2 - th1, th3 - th2, th4 - th3, h_used_words ? (long long)*h_used_words : -1LL, now() - th4);
    if (rc == TCI_OK && prof) {
        long long pcs[16];
        TCI_CUDA(h, cudaMemcpy(pcs, P.prof, sizeof pcs, cudaMemcpyDeviceToHost));
        static const char* names[16] = {"lu.reduce", "lu.exchange", "lu.stage", "lu.update", "extras", "states", "fill", "lu+nan", "output", "prologue",
                                        "pivots", "bonds", "x.publish", "x.clustersync", "upd.preamble", "-"};
        fprintf(stderr, "dsweep prof (cycles):");
        for (int q = 0; q < 15; ++q) fprintf(stderr, " %s=%lld", names[q], pcs[q]);
        fprintf(stderr, "\n");
    }
    return rc;
}

int tci_dsweep_fill_f64(tci_handle h, tci_functor f, const int32_t* h_in, int64_t in_words, void* d_acc16, double* const* h_d_cores) {
    if (!h) return TCI_ERR_INVALID;
    int rc = check_arena(h, f, h_in, in_words, "tci_dsweep_fill_f64"); if (rc) return rc;
    if (!d_acc16 || !h_d_cores) return tci_fail(h, TCI_ERR_INVALID, "tci_dsweep_fill_f64: NULL argument");
    const int N = h_in[0];
    const int32_t* ld = h_in + kHdr; const int32_t* desc = ld + N;
    std::vector<int> blockmap;
    int rmax = 1;
    for (int b = 0; b < N; ++b) {
        const int nI = desc[kDesc * b + 0], nJ = desc[kDesc * b + 2];
        if (b < N - 1) {
            if (desc[kDesc * (b + 1) + 0] != nJ) return tci_fail(h, TCI_ERR_NOT_SQUARE, "Pivot matrix at bond %d is not square!", b + 1);
            if (nJ > kSolveMaxR) return tci_fail(h, TCI_ERR_INVALID, "tci_dsweep_fill_f64: bond dimension %d > %d", nJ, kSolveMaxR);
            rmax = std::max(rmax, nJ);
        }
        if (!h_d_cores[b]) return tci_fail(h, TCI_ERR_INVALID, "tci_dsweep_fill_f64: NULL core buffer");
        const long long q = (long long)nI * ld[b];
        for (long long r0 = 0; r0 < q; r0 += kSolveRows) { blockmap.push_back(b); blockmap.push_back((int)r0); }
    }
    if (blockmap.empty()) return TCI_OK;
    TCI_CUDA(h, cudaSetDevice(h->device));
    Staging st;
    st.add(h_in, (size_t)in_words * 4);
    const size_t off_c = st.add(h_d_cores, sizeof(double*) * (size_t)N);
    const size_t off_b = st.add(blockmap.data(), sizeof(int) * blockmap.size());
    const int zero = 0;
    const size_t off_f = st.add(&zero, sizeof(int));
    rc = tci_ws_reserve(h, h->ws_c, st.bytes.size() + 256); if (rc) return rc;
    unsigned char* base = (unsigned char*)h->ws_c.ptr;
    TCI_CUDA(h, cudaMemcpyAsync(base, st.bytes.data(), st.bytes.size(), cudaMemcpyHostToDevice, h->stream));
    FillParams P;
    P.in = (const int*)base; P.cores = (double* const*)(base + off_c); P.blockmap = (const int*)(base + off_b);
    P.acc = (unsigned long long*)d_acc16; P.flag = (int*)(base + off_f);
    const int nblocks = (int)(blockmap.size() / 2);
    TCI_DS_DISPATCH(rc, f, launch_fill(h, F_, P, nblocks, rmax))
    if (rc) return rc;
    rc = tci_pinned_reserve(h, 64); if (rc) return rc;
    TCI_CUDA(h, cudaMemcpyAsync(h->pinned, P.flag, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    TCI_CUDA(h, cudaStreamSynchronize(h->stream));
    const int flag = *(int*)h->pinned;
    if (flag) return tci_fail(h, TCI_ERR_INVALID, "tci_solve_right_f64: pivot matrix is singular (site %d)", flag);
    return TCI_OK;
}

int tci_dsweep_sweep1site_f64(tci_handle h, tci_functor f, const int32_t* h_in, int64_t in_words, double reltol, double abstol,
                              void* d_acc16, double* const* h_d_cores, const int64_t* h_core_capacity, int32_t* h_out,
                              int64_t out_cap_words, int64_t* h_used_words) {
    if (!h) return TCI_ERR_INVALID;
    int rc = check_arena(h, f, h_in, in_words, "tci_dsweep_sweep1site_f64"); if (rc) return rc;
    const int N = h_in[0];
    if (!d_acc16 || !h_d_cores || !h_core_capacity || !h_out || out_cap_words < kHdr + (int64_t)kStepRec * (N - 1))
        return tci_fail(h, TCI_ERR_INVALID, "tci_dsweep_sweep1site_f64: bad arguments");
    const int32_t* ld = h_in + kHdr; const int32_t* desc = ld + N;
    long long maxelems = 0;
    for (int b = 0; b < N - 1; ++b) {
        const long long nIb = b == 0 ? desc[0] : desc[kDesc * (b - 1) + 2];          // |Iset[b]| after the previous bond <= |Jset[b-1]|
        maxelems = std::max(maxelems, nIb * ld[b] * (long long)desc[kDesc * b + 2]);
    }
    TCI_CUDA(h, cudaSetDevice(h->device));
    Staging st;
    st.add(h_in, (size_t)in_words * 4);
    const size_t off_c = st.add(h_d_cores, sizeof(double*) * (size_t)N);
    std::vector<long long> cap(h_core_capacity, h_core_capacity + N);
    const size_t off_k = st.add(cap.data(), sizeof(long long) * (size_t)N);
    const size_t in_b = (st.bytes.size() + 255) & ~(size_t)255, out_b = ((size_t)out_cap_words * 4 + 255) & ~(size_t)255;
    const size_t rtot = (size_t)std::max(h_in[6], 0), st_b = (rtot * 16 + 255) & ~(size_t)255, hs_b = (rtot * 8 + 255) & ~(size_t)255;
    rc = tci_ws_reserve(h, h->ws_c, in_b + out_b + st_b + hs_b + 512); if (rc) return rc;
    unsigned char* base = (unsigned char*)h->ws_c.ptr;
    TCI_CUDA(h, cudaMemcpyAsync(base, st.bytes.data(), st.bytes.size(), cudaMemcpyHostToDevice, h->stream));
    const bool prof1 = getenv("TCI_DSWEEP_PROF") != nullptr;
    Sweep1Params Q;
    Q.sp.in = (const int*)base; Q.sp.out = (int*)(base + in_b); Q.sp.out_cap = out_cap_words; Q.sp.reltol = reltol; Q.sp.abstol = abstol;
    Q.sp.acc = (unsigned long long*)d_acc16; Q.sp.xs_g = nullptr; Q.sp.mcap = 0; Q.sp.smem_bytes = kDynBytes; Q.sp.in_words = (int)in_words; Q.sp.resident = 0;
    Q.sp.prof = prof1 ? (long long*)(base + in_b + out_b + st_b + hs_b + 256) : nullptr;
    Q.sp.st_g = base + in_b + out_b; Q.sp.hs_g = (unsigned long long*)(base + in_b + out_b + st_b);
    Q.cores = (double* const*)(base + off_c); Q.corecap = (const long long*)(base + off_k);
    TCI_DS_DISPATCH(rc, f, launch_sweep1(h, F_, Q, maxelems))
    if (rc) return rc;
    rc = read_out(h, Q.sp.out, h_out, out_cap_words, h_used_words, "tci_dsweep_sweep1site_f64");
    if (rc == TCI_OK && prof1) {
        long long pcs[8];
        TCI_CUDA(h, cudaMemcpy(pcs, Q.sp.prof, sizeof pcs, cudaMemcpyDeviceToHost));
        fprintf(stderr, "dsweep1 prof (cycles): prologue=%lld states=%lld fill=%lld lu=%lld nan=%lld output=%lld left+carry=%lld\n", pcs[0], pcs[1],
                pcs[2], pcs[3], pcs[4], pcs[5], pcs[6]);
    }
    return rc;
}

} // extern "C"
