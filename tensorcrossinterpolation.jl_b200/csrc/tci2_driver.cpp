// tci2_driver.cpp — compiled host driver of the TensorCI2 sweeps, written ABOVE the C ABI.
//
// The north star keeps sweep strategy, index-set bookkeeping and convergence control in the host language and lets it
// call the kernels through the C ABI.  This file is that host side in C++: it restates src/tensorci2.jl's control flow
// (citations per function) and produces every number through the public entry points of include/tci_b200.h only
// (tci_batcheval_acc_f64 / tci_maxacc_*, tci_rrlu_inplace_f64, tci_luci_factors_f64, tci_solve_right_f64, tci_maxabs_f64,
// tci_trsm_right_f64, tci_tt_evaluate_f64, tci_malloc/free/memcpy) — no CUDA header, no internal symbol.  It exists because a half-sweep
// over small Pi matrices is a chain of hundreds of microsecond-sized device calls: an interpreted host pays ~200 us of
// bookkeeping per bond on top of them, a compiled one a few.  Being pure C-ABI client code it also compiles against any
// other implementation of those entry points, which is how its bookkeeping is tested without a GPU
// (tests/native_mock/, test infrastructure only).
//
// Index sets are row-major int32 tables (count x length, 1-based values), exactly what tci_batcheval_f64 takes.
// Matrices are column-major FP64 in device memory.  Bonds/sites in the C entry points are 1-based like the reference.
#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <limits>
#include <string>
#include <vector>

#include "../../include/tci_b200.h"

namespace {

struct IdxTable {                       // `n` multi-indices of `len` ints each (len may be 0: n copies of the empty index)
    int len = 0;
    int64_t n = 0;
    std::vector<int32_t> v;
    const int32_t* row(int64_t i) const { return v.data() + (size_t)i * len; }
    const int32_t* data() const { return v.empty() ? nullptr : v.data(); }
};

struct DevBuf {                         // grow-only device buffer of doubles
    double* p = nullptr;
    size_t cap = 0;
};

struct Core {                           // site tensor (chil, d, chir), dense column-major = (chil*d) x chir matrix
    DevBuf buf;
    int64_t chil = 0, d = 0, chir = 0;
    bool valid = false;
};

inline double jl_max(double a, double b) {          // Julia's max: NaN propagates
    if (a != a || b != b) return std::numeric_limits<double>::quiet_NaN();
    return a > b ? a : b;
}

inline double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

inline int64_t pad16(int64_t rows) { int64_t ld = (rows + 15) / 16 * 16; return ld > 0 ? ld : 1; }

inline uint64_t hash_row(const int32_t* r, int len) {
    uint64_t x = 0x9E3779B97F4A7C15ull ^ (uint64_t)len;
    for (int k = 0; k < len; ++k) {
        x ^= (uint64_t)(uint32_t)r[k];
        x *= 0xBF58476D1CE4E5B9ull;
        x ^= x >> 29;
    }
    return x;
}

// Rows of `t` with duplicates dropped, first occurrence kept, order kept: Base.union / unique on a vector of multi-indices.
IdxTable first_seen_unique(const IdxTable& t) {
    if (t.n <= 1) return t;
    IdxTable out; out.len = t.len;
    if (t.len == 0) { out.n = 1; return out; }
    size_t cap = 16; while (cap < (size_t)t.n * 2) cap <<= 1;
    std::vector<int64_t> slot(cap, -1);                 // open addressing: slot -> row index in `out`
    out.v.reserve(t.v.size());
    for (int64_t i = 0; i < t.n; ++i) {
        const int32_t* r = t.row(i);
        size_t s = (size_t)hash_row(r, t.len) & (cap - 1);
        bool dup = false;
        while (slot[s] >= 0) {
            if (std::memcmp(out.row(slot[s]), r, sizeof(int32_t) * t.len) == 0) { dup = true; break; }
            s = (s + 1) & (cap - 1);
        }
        if (dup) continue;
        slot[s] = out.n++;
        out.v.insert(out.v.end(), r, r + t.len);
    }
    return out;
}

IdxTable union_rows(IdxTable a, const IdxTable* b) {
    if (b && b->n > 0) {
        a.v.insert(a.v.end(), b->v.begin(), b->v.end());
        a.n += b->n;
    }
    return first_seen_unique(a);
}

// [is..., j] for is in Iset (fastest), j in 1:d — tensorci2.jl:512-517
IdxTable kronecker_left(const IdxTable& I, int d) {
    IdxTable out; out.len = I.len + 1; out.n = I.n * d;
    out.v.resize((size_t)out.n * out.len);
    int32_t* w = out.v.data();
    for (int j = 1; j <= d; ++j)
        for (int64_t i = 0; i < I.n; ++i) {
            if (I.len) std::memcpy(w, I.row(i), sizeof(int32_t) * I.len);
            w[I.len] = j;
            w += out.len;
        }
    return out;
}

// [i, js...] for i in 1:d (fastest), js in Jset — tensorci2.jl:524-529
IdxTable kronecker_right(int d, const IdxTable& J) {
    IdxTable out; out.len = J.len + 1; out.n = J.n * d;
    out.v.resize((size_t)out.n * out.len);
    int32_t* w = out.v.data();
    for (int64_t j = 0; j < J.n; ++j)
        for (int i = 1; i <= d; ++i) {
            w[0] = i;
            if (J.len) std::memcpy(w + 1, J.row(j), sizeof(int32_t) * J.len);
            w += out.len;
        }
    return out;
}

IdxTable select_rows(const IdxTable& t, const int64_t* idx1, int64_t count) {      // t[idx] with 1-based positions
    IdxTable out; out.len = t.len; out.n = count;
    out.v.resize((size_t)count * t.len);
    for (int64_t k = 0; k < count; ++k)
        if (t.len) std::memcpy(out.v.data() + (size_t)k * t.len, t.row(idx1[k] - 1), sizeof(int32_t) * t.len);
    return out;
}

bool contains_row(const IdxTable& t, const int32_t* r) {
    for (int64_t i = 0; i < t.n; ++i)
        if (t.len == 0 || std::memcmp(t.row(i), r, sizeof(int32_t) * t.len) == 0) return true;
    return false;
}

void pushunique(IdxTable& t, const int32_t* r) {                                   // util.jl:94-98
    if (contains_row(t, r)) return;
    t.v.insert(t.v.end(), r, r + t.len);
    t.n++;
}

// 1-based positions in `tab` of the rows of `sub` that occur in it (first occurrence), in `sub` order:
// findfirst(isequal(i), Icombined) filtered for nothing — tensorci2.jl:866-867
std::vector<int64_t> positions_of(const IdxTable& sub, const IdxTable& tab) {
    std::vector<int64_t> out;
    if (sub.n == 0 || tab.n == 0) return out;
    if (tab.len == 0) { out.assign((size_t)sub.n, 1); return out; }
    size_t cap = 16; while (cap < (size_t)tab.n * 2) cap <<= 1;
    std::vector<int64_t> slot(cap, -1);
    for (int64_t i = 0; i < tab.n; ++i) {
        const int32_t* r = tab.row(i);
        size_t s = (size_t)hash_row(r, tab.len) & (cap - 1);
        bool dup = false;
        while (slot[s] >= 0) {
            if (std::memcmp(tab.row(slot[s]), r, sizeof(int32_t) * tab.len) == 0) { dup = true; break; }
            s = (s + 1) & (cap - 1);
        }
        if (!dup) slot[s] = i;
    }
    for (int64_t k = 0; k < sub.n; ++k) {
        const int32_t* r = sub.row(k);
        size_t s = (size_t)hash_row(r, tab.len) & (cap - 1);
        while (slot[s] >= 0) {
            if (std::memcmp(tab.row(slot[s]), r, sizeof(int32_t) * tab.len) == 0) { out.push_back(slot[s] + 1); break; }
            s = (s + 1) & (cap - 1);
        }
    }
    return out;
}

inline uint64_t splitmix64(uint64_t& s) {
    s += 0x9E3779B97F4A7C15ull;
    uint64_t z = s;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// pushrandomsubset!(subset, 1:setsize, n) — util.jl:214-219 with randomsubset :173-191 (draw, then deleteat!).  The reference
// draws from Julia's global RNG, which has no portable restatement; the stream here is the counter-based splitmix64 the
// Python host (matrixlu.Rng) and the CPU checker use, so all three pick the same rows/columns for a given seed.
void pushrandomsubset(std::vector<int64_t>& subset, int64_t setsize, int64_t n, uint64_t& rng) {
    std::vector<char> in((size_t)setsize + 1, 0);
    for (int64_t v : subset) in[(size_t)v] = 1;
    std::vector<int64_t> c;
    for (int64_t v = 1; v <= setsize; ++v) if (!in[(size_t)v]) c.push_back(v);
    n = std::min<int64_t>(n, (int64_t)c.size());
    for (int64_t i = 0; i < n; ++i) {
        const size_t idx = (size_t)(splitmix64(rng) % (uint64_t)c.size());
        subset.push_back(c[idx]);
        c.erase(c.begin() + (std::ptrdiff_t)idx);
    }
}

} // namespace

struct tci_tci2_s {
    tci_handle h = nullptr;
    int N = 0;
    std::vector<int32_t> localdims;
    std::vector<IdxTable> I, J;                         // Iset[b] has length b, Jset[b] has length N-b-1 (0-based site b)
    std::vector<Core> T;
    std::vector<double> pivoterrors, bonderrors;
    double maxsamplevalue = 0.0;
    bool have_hist = false;                             // Iset_history[end] / Jset_history[end] (only the last entry is ever read)
    std::vector<IdxTable> histI, histJ;
    int64_t nhist = 0;
    DevBuf pi, L, U, pi1, pmat, pts;                    // scratch: Pi, rrLU factors, setsitetensor! operands, point values
    void* acc = nullptr;                                // device-side max |f| accumulator: updatemaxsample! without a host round
                                                        // trip per batch; read back where a sweep ends (sync_maxsample)
    std::vector<int64_t> rowperm, colperm;
    std::vector<double> pe;
    std::string err;
    uint64_t rng = 0;                                   // random starts of the global pivot finder
    int engine = 0;                                     // 1: device-resident half-sweeps (tci_dsweep_*) where they apply
    std::vector<int32_t> arena_in, arena_out;           // packed index sets for / from the device-resident sweeps
    int64_t ndevsweeps = 0, ndevbonds = 0;
    int pivotsearch = 0;                                // 0 :full, 1 :rook (tensorci2.jl:845-914)
    uint64_t rookrng = 0;                               // stream behind pushrandomsubset! of the rook search
    // statistics (host wall clock around the calls; the batch-eval and rrLU entry points synchronise)
    int64_t nevals = 0, nbondupdates = 0;
    double t_eval = 0, t_lu = 0, t_factors = 0, t_host = 0;
};

namespace {

int fail(tci_tci2 t, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    t->err = buf;
    return code;
}

int libfail(tci_tci2 t, int code) {                     // a C-ABI call failed: keep its message
    const char* m = tci_last_error(t->h);
    t->err = m ? m : "tci_b200 call failed";
    return code;
}

#define TCI2_CALL(t, expr) do { int rc__ = (expr); if (rc__ != TCI_OK) return libfail((t), rc__); } while (0)

int reserve(tci_tci2 t, DevBuf& b, size_t doubles) {
    if (b.cap >= doubles && b.p) return TCI_OK;
    if (b.p) { TCI2_CALL(t, tci_free_pooled(t->h, b.p)); b.p = nullptr; b.cap = 0; }
    // stream-ordered pool (cudaMallocAsync): a regrow neither synchronises nor costs a cudaMalloc; start small and double
    size_t want = std::max<size_t>(2 * doubles, 2048);
    void* p = nullptr;
    TCI2_CALL(t, tci_malloc_pooled(t->h, &p, want * sizeof(double)));
    b.p = (double*)p; b.cap = want;
    return TCI_OK;
}

void invalidate(tci_tci2 t) { for (auto& c : t->T) c.valid = false; }       // tensorci2.jl:206-211

void updateerrors(tci_tci2 t, int b0, const std::vector<double>& errs) {     // tensorci2.jl:247-260
    t->bonderrors[b0] = errs.back();
    const size_t k = std::max(t->pivoterrors.size(), errs.size());
    std::vector<double> out(k);
    for (size_t i = 0; i < k; ++i)
        out[i] = jl_max(i < t->pivoterrors.size() ? t->pivoterrors[i] : 0.0, i < errs.size() ? errs[i] : 0.0);
    t->pivoterrors.swap(out);
}

double pivoterror(const tci_tci2_s* t) {                                     // tensorci2.jl:277-279
    double m = -std::numeric_limits<double>::infinity();
    for (double x : t->bonderrors) m = jl_max(m, x);
    return m;
}

// filltensor (tensorci2.jl:475-497) + updatemaxsample! (:636-638): element (i, c, j) -> out[i + c*sc + j*sj]
int fill(tci_tci2 t, tci_functor f, const IdxTable& I, const IdxTable& J, double* d_out, int64_t sc, int64_t sj, int64_t ncenter_elems) {
    if (I.n == 0 || J.n == 0) return TCI_OK;
    const double t0 = now_s();
    TCI2_CALL(t, tci_batcheval_acc_f64(t->h, f, t->localdims.data(), t->N, I.data(), I.n, I.len, J.data(), J.n, J.len, d_out, sc, sj,
                                       t->acc));
    t->nevals += I.n * J.n * ncenter_elems;
    t->t_eval += now_s() - t0;
    return TCI_OK;
}

// maxsamplevalue <- the device accumulator (every fill since the last read has been merged into it on the device)
int sync_maxsample(tci_tci2 t) {
    double v = 0.0;
    TCI2_CALL(t, tci_maxacc_get(t->h, t->acc, &v));
    t->maxsamplevalue = v;
    return TCI_OK;
}

int set_maxsample(tci_tci2 t, double v) {
    t->maxsamplevalue = v;
    TCI2_CALL(t, tci_maxacc_set(t->h, t->acc, v));
    return TCI_OK;
}

struct Luci {                       // what updatepivots!/sweep1site! read off a MatrixLUCI
    int64_t m = 0, n = 0, r = 0, ldl = 1, ldu = 1;
    bool leftorth = true;
    std::vector<double> errs;       // pivoterrors(luci) = [|diag|..., error]
};

// MatrixLUCI(Pi; reltol, abstol, maxrank, leftorthogonal) (matrixluci.jl:55) on t->pi (m x n, ld): rrLU in place,
// L/U extracted to t->L / t->U, permutations to t->rowperm / t->colperm.
int luci_factorize(tci_tci2 t, int64_t m, int64_t n, int64_t ld, int64_t maxbonddim, double reltol, double abstol, bool leftorth, Luci* out) {
    const double t0 = now_s();
    const int64_t mn = std::min(m, n);
    const int64_t mr = (maxbonddim <= 0 || maxbonddim > mn) ? mn : maxbonddim;
    out->m = m; out->n = n; out->leftorth = leftorth;
    out->ldl = std::max<int64_t>(m, 1); out->ldu = std::max<int64_t>(mr, 1);
    int rc = reserve(t, t->L, (size_t)out->ldl * std::max<int64_t>(mr, 1)); if (rc) return rc;
    rc = reserve(t, t->U, (size_t)out->ldu * std::max<int64_t>(n, 1)); if (rc) return rc;
    t->rowperm.resize((size_t)std::max<int64_t>(m, 1));
    t->colperm.resize((size_t)std::max<int64_t>(n, 1));
    t->pe.assign((size_t)mr + 1, 0.0);
    int64_t npivot = 0; double error = 0.0;
    TCI2_CALL(t, tci_rrlu_inplace_f64(t->h, t->pi.p, m, n, ld, mr, reltol, abstol, leftorth ? 1 : 0, t->rowperm.data(), t->colperm.data(),
                                      &npivot, &error, t->L.p, out->ldl, t->U.p, out->ldu, t->pe.data()));
    out->r = npivot;
    out->errs.assign(t->pe.begin(), t->pe.begin() + npivot + 1);
    t->t_lu += now_s() - t0;
    return TCI_OK;
}

// left(luci) (m x r) or right(luci) (r x n) (matrixluci.jl:256-283) straight into the storage of site tensor b0
int core_from_factor(tci_tci2 t, const Luci& lu, int b0, bool left) {
    const double t0 = now_s();
    Core& c = t->T[b0];
    const int64_t d = t->localdims[b0];
    const int64_t chil = t->I[b0].n, chir = t->J[b0].n;
    const int64_t rows = left ? lu.m : lu.r, cols = left ? lu.r : lu.n;
    if (rows * cols != chil * d * chir) return fail(t, TCI_ERR_INVALID, "setsitetensor!: size mismatch");
    int rc = reserve(t, c.buf, (size_t)rows * cols); if (rc) return rc;
    if (rows * cols > 0)
        TCI2_CALL(t, tci_luci_factors_f64(t->h, lu.m, lu.n, lu.r, lu.leftorth ? 1 : 0, t->rowperm.data(), t->colperm.data(), t->L.p, lu.ldl,
                                          t->U.p, lu.ldu, left ? c.buf.p : nullptr, std::max<int64_t>(rows, 1), left ? nullptr : c.buf.p,
                                          std::max<int64_t>(rows, 1)));
    c.chil = chil; c.d = d; c.chir = chir; c.valid = true;
    t->t_factors += now_s() - t0;
    return TCI_OK;
}

// arrlu(Float64, f, (m, n), I0, J0; maxrank, reltol, abstol, leftorthogonal, numrookiter) — matrixlu.jl:500-569 — on the Pi matrix
// defined by the combined index sets, which is never formed: alternating row and column STRIPS of it (SubMatrix,
// tensorci2.jl:764-804) are evaluated (K1) and factorised (rrLU) until the pivots stop moving, then the factors are completed
// by two triangular solves (cols2Lmatrix!/rows2Umatrix!, matrixlu.jl:627-674).  Every strip's rrLU writes L and U straight
// into the full-size factor buffers (leading dimensions m and maxrank), so the completion happens in place.
int arrlu(tci_tci2 t, tci_functor f, const IdxTable& Ic, const IdxTable& Jc, std::vector<int64_t> I0, std::vector<int64_t> J0,
          int64_t maxbonddim, double reltol, double abstol, bool leftorth, Luci* out) {
    const int numrookiter = 5;
    const int64_t m = Ic.n, n = Jc.n, mn = std::min(m, n);
    const int64_t maxrank = (maxbonddim <= 0 || maxbonddim > mn) ? mn : maxbonddim;
    out->m = m; out->n = n; out->leftorth = leftorth; out->r = 0;
    out->ldl = std::max<int64_t>(m, 1); out->ldu = std::max<int64_t>(maxrank, 1);
    if (mn == 0) { out->errs.assign(1, 0.0); return TCI_OK; }
    int rc = reserve(t, t->L, (size_t)out->ldl * (size_t)std::max<int64_t>(maxrank, 1)); if (rc) return rc;
    rc = reserve(t, t->U, (size_t)out->ldu * (size_t)n); if (rc) return rc;
    std::vector<int64_t> rowperm((size_t)m), colperm((size_t)n), srow, scol, tmp;
    for (int64_t i = 0; i < m; ++i) rowperm[(size_t)i] = i + 1;
    for (int64_t j = 0; j < n; ++j) colperm[(size_t)j] = j + 1;
    bool islowrank = false;
    int64_t r = 0, lastrows = 0, lastcols = 0;
    std::vector<double> pe;
    for (;;) {
        if (leftorth) pushrandomsubset(J0, n, std::max<int64_t>(1, (int64_t)J0.size()), t->rookrng);     // :521-525
        else          pushrandomsubset(I0, m, std::max<int64_t>(1, (int64_t)I0.size()), t->rookrng);
        for (int rookiter = 1; rookiter <= numrookiter; ++rookiter) {
            const bool colmove = ((rookiter % 2 == 0) == leftorth);                                       // :529
            const std::vector<int64_t>& rows = colmove ? I0 : rowperm;                                    // :532-536
            const std::vector<int64_t>& cols = colmove ? colperm : J0;
            const int64_t kr = (int64_t)rows.size(), kc = (int64_t)cols.size(), ld = pad16(kr);
            IdxTable left = select_rows(Ic, rows.data(), kr), right = select_rows(Jc, cols.data(), kc);
            rc = reserve(t, t->pi, (size_t)ld * (size_t)std::max<int64_t>(kc, 1)); if (rc) return rc;
            rc = fill(t, f, left, right, t->pi.p, kr, ld, 1); if (rc) return rc;
            const double t0 = now_s();
            const int64_t kmn = std::min(kr, kc), mr = std::min(maxrank, kmn);
            srow.assign((size_t)std::max<int64_t>(kr, 1), 0); scol.assign((size_t)std::max<int64_t>(kc, 1), 0);
            pe.assign((size_t)mr + 1, 0.0);
            int64_t npivot = 0; double error = 0.0;
            TCI2_CALL(t, tci_rrlu_inplace_f64(t->h, t->pi.p, kr, kc, ld, mr, reltol, abstol, leftorth ? 1 : 0, srow.data(), scol.data(), &npivot,
                                              &error, t->L.p, out->ldl, t->U.p, out->ldu, pe.data()));
            t->t_lu += now_s() - t0;
            // _optimizerrlu! swaps lu.rowpermutation / lu.colpermutation at the POSITIONS of the strip (:254-275, :539-540)
            tmp.assign(rowperm.begin(), rowperm.begin() + kr);
            for (int64_t i = 0; i < kr; ++i) rowperm[(size_t)i] = tmp[(size_t)(srow[(size_t)i] - 1)];
            tmp.assign(colperm.begin(), colperm.begin() + kc);
            for (int64_t j = 0; j < kc; ++j) colperm[(size_t)j] = tmp[(size_t)(scol[(size_t)j] - 1)];
            r = npivot; lastrows = kr; lastcols = kc;
            pe.resize((size_t)r + 1);
            islowrank = islowrank || r < kmn;                                                             // :541
            std::vector<int64_t> ri(rowperm.begin(), rowperm.begin() + r), ci(colperm.begin(), colperm.begin() + r);
            if (ri == I0 && ci == J0) break;                                                              // :544-546
            I0.swap(ri); J0.swap(ci);                                                                     // :548-549
        }
        if (islowrank || (int64_t)I0.size() >= maxrank) break;                                            // :552-554
    }
    out->r = r; out->errs = pe;
    if (r > 0 && lastrows < m) {                                                                          // :558-564
        std::vector<char> in((size_t)m + 1, 0); for (int64_t v : I0) in[(size_t)v] = 1;
        std::vector<int64_t> I2; for (int64_t v = 1; v <= m; ++v) if (!in[(size_t)v]) I2.push_back(v);
        rowperm = I0; rowperm.insert(rowperm.end(), I2.begin(), I2.end());
        const int64_t q = (int64_t)I2.size();
        IdxTable left = select_rows(Ic, I2.data(), q), right = select_rows(Jc, J0.data(), r);
        rc = fill(t, f, left, right, t->L.p + r, q, out->ldl, 1); if (rc) return rc;                      // rows r.. of L, in place
        const double t0 = now_s();
        TCI2_CALL(t, tci_trsm_right_f64(t->h, t->U.p, 1, out->ldu, r, 0, 0, t->L.p + r, 1, out->ldl, q)); // cols2Lmatrix! :627-648
        t->t_factors += now_s() - t0;
    }
    if (r > 0 && lastcols < n) {                                                                          // :567-573
        std::vector<char> in((size_t)n + 1, 0); for (int64_t v : J0) in[(size_t)v] = 1;
        std::vector<int64_t> J2; for (int64_t v = 1; v <= n; ++v) if (!in[(size_t)v]) J2.push_back(v);
        colperm = J0; colperm.insert(colperm.end(), J2.begin(), J2.end());
        const int64_t q = (int64_t)J2.size();
        IdxTable left = select_rows(Ic, I0.data(), r), right = select_rows(Jc, J2.data(), q);
        double* U2 = t->U.p + (size_t)r * (size_t)out->ldu;                                               // columns r.. of U, in place
        rc = fill(t, f, left, right, U2, r, out->ldu, 1); if (rc) return rc;
        const double t0 = now_s();
        // rows2Umatrix! :655-674: L1 X = U2  <=>  X^T L1^T = U2^T — the right solve on transposed strides
        TCI2_CALL(t, tci_trsm_right_f64(t->h, t->L.p, out->ldl, 1, r, 0, 0, U2, out->ldu, 1, q));
        t->t_factors += now_s() - t0;
    }
    t->rowperm = rowperm; t->colperm = colperm;
    if (t->rowperm.empty()) t->rowperm.assign(1, 1);
    if (t->colperm.empty()) t->colperm.assign(1, 1);
    return TCI_OK;
}

// updatepivots! (tensorci2.jl:825-930), pivotsearch = :full or :rook.  b0: 0-based bond (between sites b0 and b0+1).
// want_cores = false: the caller is about to invalidate or recompute the site tensors anyway (every later updatepivots!
// invalidates them, fillsitetensors! recomputes all of them), so left(luci)/right(luci) need not be formed.
int updatepivots(tci_tci2 t, tci_functor f, int b0, bool leftorth, double reltol, double abstol, int64_t maxbonddim,
                 const IdxTable* extraI, const IdxTable* extraJ, bool want_cores = true) {
    invalidate(t);
    const double th = now_s();
    const bool noextra = (!extraI || extraI->n == 0) && (!extraJ || extraJ->n == 0);
    IdxTable Ic = union_rows(kronecker_left(t->I[b0], t->localdims[b0]), extraI);            // :842
    IdxTable Jc = union_rows(kronecker_right(t->localdims[b0 + 1], t->J[b0 + 1]), extraJ);   // :843
    t->t_host += now_s() - th;
    const int64_t m = Ic.n, n = Jc.n, ld = pad16(m);
    Luci lu;
    int rc = TCI_OK;
    bool full = t->pivotsearch == 0;
    if (t->pivotsearch == 1) {                                                               // :863-905
        rc = arrlu(t, f, Ic, Jc, positions_of(t->I[b0 + 1], Ic), positions_of(t->J[b0], Jc), maxbonddim, reltol, abstol, leftorth, &lu);
        if (rc) return rc;
        full = lu.r == 0;                                                                    // :886-899 fall back to the full search
    }
    if (full) {
        rc = reserve(t, t->pi, (size_t)ld * std::max<int64_t>(n, 1)); if (rc) return rc;
        rc = fill(t, f, Ic, Jc, t->pi.p, m, ld, 1); if (rc) return rc;                       // :851-854
        rc = luci_factorize(t, m, n, ld, maxbonddim, reltol, abstol, leftorth, &lu); if (rc) return rc;   // :856-862
    }
    t->I[b0 + 1] = select_rows(Ic, t->rowperm.data(), lu.r);                                 // :915
    t->J[b0] = select_rows(Jc, t->colperm.data(), lu.r);                                     // :916
    if (noextra && want_cores) {                                                             // :917-920
        rc = core_from_factor(t, lu, b0, true); if (rc) return rc;
        rc = core_from_factor(t, lu, b0 + 1, false); if (rc) return rc;
    }
    updateerrors(t, b0, lu.errs);                                                            // :921-924
    t->nbondupdates++;
    return TCI_OK;
}

// setsitetensor!(tci, f, b) (tensorci2.jl:599-629)
int setsitetensor(tci_tci2 t, tci_functor f, int b0) {
    const int64_t nI = t->I[b0].n, nJ = t->J[b0].n, d = t->localdims[b0];
    Core& c = t->T[b0];
    const int64_t q = nI * d;
    if (b0 == t->N - 1) {                                                                    // :612-615: plain evaluation
        int rc = reserve(t, c.buf, (size_t)q * std::max<int64_t>(nJ, 1)); if (rc) return rc;
        rc = fill(t, f, t->I[b0], t->J[b0], c.buf.p, nI, q, d); if (rc) return rc;
        c.chil = nI; c.d = d; c.chir = nJ; c.valid = true;
        return TCI_OK;
    }
    const int64_t ld1 = pad16(q);
    int rc = reserve(t, t->pi1, (size_t)ld1 * std::max<int64_t>(nJ, 1)); if (rc) return rc;
    rc = fill(t, f, t->I[b0], t->J[b0], t->pi1.p, nI, ld1, d); if (rc) return rc;             // Pi1, :617-620
    const int64_t nI1 = t->I[b0 + 1].n, ldp = pad16(nI1);
    rc = reserve(t, t->pmat, (size_t)ldp * std::max<int64_t>(nJ, 1)); if (rc) return rc;
    rc = fill(t, f, t->I[b0 + 1], t->J[b0], t->pmat.p, nI1, ldp, 1); if (rc) return rc;      // P, :622
    if (nI1 != nJ) return fail(t, TCI_ERR_NOT_SQUARE, "Pivot matrix at bond %d is not square!", b0 + 1);   // :623
    const double t0 = now_s();
    rc = reserve(t, c.buf, (size_t)q * std::max<int64_t>(nJ, 1)); if (rc) return rc;
    TCI2_CALL(t, tci_solve_right_f64(t->h, t->pi1.p, q, ld1, t->pmat.p, nJ, ldp, c.buf.p, std::max<int64_t>(q, 1)));   // :626
    c.chil = nI; c.d = d; c.chir = nI1; c.valid = true;
    t->t_factors += now_s() - t0;
    return TCI_OK;
}

int device_fillsitetensors(tci_tci2 t, tci_functor f, bool* done);

int fillsitetensors(tci_tci2 t, tci_functor f) {                                             // globalsearch.jl:202-208
    if (t->engine == 1 && tci_dsweep_supported(t->h, f)) {
        bool done = false;
        int rc = device_fillsitetensors(t, f, &done); if (rc) return rc;
        if (done) return TCI_OK;
    }
    for (int b0 = 0; b0 < t->N; ++b0) { int rc = setsitetensor(t, f, b0); if (rc) return rc; }
    return TCI_OK;
}

// sweep1site! (tensorci2.jl:659-725)
int device_sweep1site(tci_tci2 t, tci_functor f, double reltol, double abstol, int64_t maxbonddim, int* next_step);

int sweep1site(tci_tci2 t, tci_functor f, bool fwd, double reltol, double abstol, int64_t maxbonddim, bool updatetensors) {
    t->pivoterrors.clear();                                                                  // flushpivoterror!
    invalidate(t);
    const int N = t->N;
    int first = 0;
    if (t->engine == 1 && fwd && updatetensors && tci_dsweep_supported(t->h, f)) {           // the whole pass as ONE launch (csrc/sweep.cu)
        int rc = device_sweep1site(t, f, reltol, abstol, maxbonddim, &first); if (rc) return rc;
        if (first > N - 1) return sync_maxsample(t);                                          // every bond and the last tensor done
    }
    for (int step = first; step < N - 1; ++step) {
        const int b0 = fwd ? step : N - 1 - step;
        const int d = t->localdims[b0];
        IdxTable Is = fwd ? kronecker_left(t->I[b0], d) : t->I[b0];                          // :679-680
        IdxTable Js = fwd ? t->J[b0] : kronecker_right(d, t->J[b0]);
        const int64_t m = Is.n, n = Js.n, ld = pad16(m), nI = t->I[b0].n;
        int rc = reserve(t, t->pi, (size_t)ld * std::max<int64_t>(n, 1)); if (rc) return rc;
        // (|I|, d, |J|) reshaped to rows x cols: forward rows = (I, s); backward cols = (s, J)       :681-686
        rc = fwd ? fill(t, f, t->I[b0], t->J[b0], t->pi.p, nI, ld, d) : fill(t, f, t->I[b0], t->J[b0], t->pi.p, ld, ld * d, d);
        if (rc) return rc;
        Luci lu;
        rc = luci_factorize(t, m, n, ld, maxbonddim, reltol, abstol, fwd, &lu); if (rc) return rc;   // :688-694
        if (fwd) {                                                                            // :696-702
            t->I[b0 + 1] = select_rows(Is, t->rowperm.data(), lu.r);
            t->J[b0] = select_rows(Js, t->colperm.data(), lu.r);
        } else {
            t->I[b0] = select_rows(Is, t->rowperm.data(), lu.r);
            t->J[b0 - 1] = select_rows(Js, t->colperm.data(), lu.r);
        }
        if (updatetensors) {                                                                  // :703-711
            rc = core_from_factor(t, lu, b0, fwd); if (rc) return rc;
            const Core& c = t->T[b0];
            const int64_t rows = c.chil * c.d;
            double mx = 0.0;
            if (rows * c.chir > 0) TCI2_CALL(t, tci_maxabs_f64(t->h, c.buf.p, rows, c.chir, rows, &mx));
            if (mx != mx) return fail(t, TCI_ERR_NAN_TENSOR, "Error: NaN in tensor T[%d]", b0 + 1);
        }
        updateerrors(t, fwd ? b0 : b0 - 1, lu.errs);                                         // :712-716
    }
    if (updatetensors) {                                                                      // :719-723: last tensor by evaluation
        const int last = fwd ? N - 1 : 0;
        Core& c = t->T[last];
        const int64_t nI = t->I[last].n, nJ = t->J[last].n, d = t->localdims[last];
        int rc = reserve(t, c.buf, (size_t)(nI * d) * std::max<int64_t>(nJ, 1)); if (rc) return rc;
        rc = fill(t, f, t->I[last], t->J[last], c.buf.p, nI, nI * d, d); if (rc) return rc;
        c.chil = nI; c.d = d; c.chir = nJ; c.valid = true;
    }
    return sync_maxsample(t);
}

int makecanonical(tci_tci2 t, tci_functor f, double reltol, double abstol, int64_t maxbonddim) {     // tensorci2.jl:738-749
    int rc = sweep1site(t, f, true, 0.0, 0.0, 0, false); if (rc) return rc;
    rc = sweep1site(t, f, false, reltol, abstol, maxbonddim, false); if (rc) return rc;
    return sweep1site(t, f, true, reltol, abstol, maxbonddim, true);
}

inline bool forwardsweep(int strategy, int iter) {      // sweepstrategies.jl:41-50; 0 = :forward, 1 = :backandforth, else :backward
    return strategy == 0 || (strategy == 1 && iter % 2 == 1);
}

// ---- device-resident half-sweep (csrc/sweep.cu) ---------------------------------------------------------------------------
// The bond loop of one half-sweep (tensorci2.jl:1226-1250) as ONE launch per stretch of bonds whose Pi fits shared memory:
// the index sets (and the history entries used as extras) go down as one int32 arena, the new sets and pivot errors come
// back as one arena.  A bond that does not fit is done by the per-bond path above and the sweep continues behind it.
constexpr int kArenaHdr = 8, kArenaDesc = 12, kArenaStep = 6;

bool device_bond_fits(int64_t m, int64_t n) {             // mirror of the kernel's own check (8-CTA cluster, 208 KB each)
    if (m < 1 || n < 1 || m > 1024 || n > 65535) return false;
    const int64_t nloc = (n + 7) / 8 + 1, lds = ((m + 31) / 32 * 32) | 1;        // padded column length, one spare column
    return 8 * lds * nloc + 48 * (m + 128 + n) + 8192 <= 208 * 1024;
}

void append_table(std::vector<int32_t>& a, const IdxTable* t, int32_t* cnt_off) {
    cnt_off[0] = t ? (int32_t)t->n : 0;
    cnt_off[1] = (int32_t)a.size();
    if (t && !t->v.empty()) a.insert(a.end(), t->v.begin(), t->v.end());
}

void pack_sets(tci_tci2 t, std::vector<int32_t>& a, bool fwd, int s0, int s1, int64_t maxbonddim, const std::vector<IdxTable>* extraI,
               const std::vector<IdxTable>* extraJ) {
    const int N = t->N;
    const bool useextra = extraI != nullptr;
    a.assign((size_t)kArenaHdr + (size_t)N * (1 + kArenaDesc), 0);
    a[0] = N; a[1] = fwd ? 1 : 0; a[2] = s0; a[3] = s1;
    a[4] = (maxbonddim <= 0 || maxbonddim > 0x7fffffff) ? 0 : (int32_t)maxbonddim; a[5] = useextra ? 1 : 0;
    for (int b = 0; b < N; ++b) a[(size_t)kArenaHdr + b] = t->localdims[b];
    int32_t rows = 0;
    for (int b = 0; b < N; ++b) {
        int32_t d[12];
        const IdxTable* tabs[4] = {&t->I[b], &t->J[b], useextra ? &(*extraI)[b] : nullptr, useextra ? &(*extraJ)[b] : nullptr};
        for (int w = 0; w < 4; ++w) {
            append_table(a, tabs[w], d + 2 * w);
            d[8 + w] = rows;                              // index of the table's first row among all rows of the arena
            rows += d[2 * w];
        }
        std::memcpy(a.data() + kArenaHdr + N + (size_t)kArenaDesc * b, d, sizeof d);
    }
    a[6] = rows;
}

int device_halfsweep(tci_tci2 t, tci_functor f, bool fwd, double abstol, int64_t maxbonddim, const std::vector<IdxTable>* extraI,
                     const std::vector<IdxTable>* extraJ) {
    const int N = t->N;
    const bool useextra = extraI != nullptr;
    invalidate(t);
    int s = 0;
    while (s < N - 1) {
        const int b0 = fwd ? s : N - 2 - s;
        {   // does the bond at the head of the stretch fit?  (the kernel re-checks every later bond with the sizes it produces)
            const int64_t m = t->I[b0].n * t->localdims[b0] + (useextra ? (*extraI)[b0 + 1].n : 0);
            const int64_t n = t->localdims[b0 + 1] * t->J[b0 + 1].n + (useextra ? (*extraJ)[b0].n : 0);
            if (!device_bond_fits(m, n)) {
                int rc = updatepivots(t, f, b0, fwd, 1e-14, abstol, maxbonddim, useextra ? &(*extraI)[b0 + 1] : nullptr,
                                      useextra ? &(*extraJ)[b0] : nullptr, false);
                if (rc) return rc;
                ++s;
                continue;
            }
        }
        const double t0 = now_s();
        std::vector<int32_t>& a = t->arena_in;
        pack_sets(t, a, fwd, s, N - 1, maxbonddim, extraI, extraJ);
        static const bool hostprof = getenv("TCI_DSWEEP_HOSTPROF") != nullptr;
        const double tp1 = now_s();
        // output capacity: every bond may return up to rcap rows of N ints twice; rcap is bounded by what fits the kernel (512)
        int64_t rcap = 512;
        if (maxbonddim > 0 && maxbonddim < rcap) rcap = maxbonddim;
        const int64_t cap = kArenaHdr + (int64_t)kArenaStep * (N - 1) + (int64_t)(N - 1 - s) * (rcap * (N + 1) + 2 * (rcap + 2));
        t->arena_out.resize((size_t)cap);
        int64_t used = 0;
        TCI2_CALL(t, tci_dsweep_halfsweep_f64(t->h, f, a.data(), (int64_t)a.size(), 1e-14, abstol, t->acc, t->arena_out.data(), cap, &used));
        const double tp2 = now_s();
        const int32_t* o = t->arena_out.data();
        const int status = o[0], next = o[1];
        t->ndevsweeps++;
        for (int st = s; st < next; ++st) {
            const int bb = fwd ? st : N - 2 - st;
            const int32_t* rec = o + kArenaHdr + (size_t)kArenaStep * (st - s);
            const int r = rec[0];
            IdxTable& nI = t->I[bb + 1]; IdxTable& nJ = t->J[bb];
            nI.n = r; nI.v.assign(o + rec[1], o + rec[1] + (size_t)r * nI.len);
            nJ.n = r; nJ.v.assign(o + rec[2], o + rec[2] + (size_t)r * nJ.len);
            std::vector<double> errs((size_t)r + 1);
            std::memcpy(errs.data(), o + rec[3], sizeof(double) * ((size_t)r + 1));
            updateerrors(t, bb, errs);
            t->nevals += (int64_t)rec[4] * rec[5];
            t->nbondupdates++; t->ndevbonds++;
        }
        t->t_lu += now_s() - t0;
        if (hostprof) fprintf(stderr, "driver halfsweep (us): pack=%.1f call=%.1f unpack=%.1f\n", 1e6 * (tp1 - t0), 1e6 * (tp2 - tp1), 1e6 * (now_s() - tp2));
        if (status == 2) return fail(t, TCI_ERR_NAN_L, "lu.L contains NaNs");
        if (status == 3) return fail(t, TCI_ERR_NAN_U, "lu.U contains NaNs");
        s = next;
        if (status == 1 && s < N - 1) {                     // this bond outgrew shared memory: per-bond path, then go on
            const int bb = fwd ? s : N - 2 - s;
            int rc = updatepivots(t, f, bb, fwd, 1e-14, abstol, maxbonddim, useextra ? &(*extraI)[bb + 1] : nullptr,
                                  useextra ? &(*extraJ)[bb] : nullptr, false);
            if (rc) return rc;
            ++s;
        }
    }
    return TCI_OK;
}

// fillsitetensors! (globalsearch.jl:202-208) as ONE launch over all sites (tci_dsweep_fill_f64) when every bond is <= 64
int device_fillsitetensors(tci_tci2 t, tci_functor f, bool* done) {
    const int N = t->N;
    *done = false;
    for (int b = 0; b + 1 < N; ++b) {
        if (t->J[b].n > 64 || t->J[b].n < 1 || t->I[b].n < 1) return TCI_OK;
        if (t->I[b + 1].n != t->J[b].n) return fail(t, TCI_ERR_NOT_SQUARE, "Pivot matrix at bond %d is not square!", b + 1);   // :623
    }
    if (t->I[N - 1].n < 1 || t->J[N - 1].n < 1) return TCI_OK;
    const double t0 = now_s();
    std::vector<double*> ptrs((size_t)N);
    for (int b = 0; b < N; ++b) {
        Core& c = t->T[b];
        const int64_t q = t->I[b].n * t->localdims[b], r = t->J[b].n;
        int rc = reserve(t, c.buf, (size_t)(q * r)); if (rc) return rc;
        ptrs[(size_t)b] = c.buf.p;
    }
    pack_sets(t, t->arena_in, true, 0, 0, 0, nullptr, nullptr);
    TCI2_CALL(t, tci_dsweep_fill_f64(t->h, f, t->arena_in.data(), (int64_t)t->arena_in.size(), t->acc, ptrs.data()));
    for (int b = 0; b < N; ++b) {
        Core& c = t->T[b];
        c.chil = t->I[b].n; c.d = t->localdims[b]; c.chir = t->J[b].n; c.valid = true;
        t->nevals += c.chil * c.d * c.chir + (b + 1 < N ? c.chir * c.chir : 0);
    }
    t->t_factors += now_s() - t0;
    *done = true;
    return TCI_OK;
}

// sweep1site!(tci, f, :forward; updatetensors=true) (tensorci2.jl:659-725) as ONE launch (tci_dsweep_sweep1site_f64).  *next_step =
// the first step the kernel did NOT do (N: everything including the last tensor; the caller continues from there per bond).
int device_sweep1site(tci_tci2 t, tci_functor f, double reltol, double abstol, int64_t maxbonddim, int* next_step) {
    const int N = t->N;
    *next_step = 0;
    for (int b = 0; b < N; ++b) if (t->I[b].n < 1 || t->J[b].n < 1) return TCI_OK;
    const double t0 = now_s();
    std::vector<double*> ptrs((size_t)N);
    std::vector<int64_t> caps((size_t)N);
    for (int b = 0; b < N; ++b) {                          // |Iset[b]| after bond b-1 <= |Jset[b-1]|, the new |Jset[b]| <= the old one
        const int64_t nIb = b == 0 ? t->I[0].n : t->J[b - 1].n;
        caps[(size_t)b] = nIb * t->localdims[b] * std::max<int64_t>(t->J[b].n, 1);
        int rc = reserve(t, t->T[b].buf, (size_t)caps[(size_t)b]); if (rc) return rc;
        ptrs[(size_t)b] = t->T[b].buf.p;
        caps[(size_t)b] = (int64_t)t->T[b].buf.cap;
    }
    pack_sets(t, t->arena_in, true, 0, N - 1, maxbonddim, nullptr, nullptr);
    int64_t rcap = 512;
    if (maxbonddim > 0 && maxbonddim < rcap) rcap = maxbonddim;
    const int64_t cap = kArenaHdr + (int64_t)kArenaStep * (N - 1) + (int64_t)(N - 1) * (rcap * (N + 1) + 2 * (rcap + 2));
    t->arena_out.resize((size_t)cap);
    int64_t used = 0;
    TCI2_CALL(t, tci_dsweep_sweep1site_f64(t->h, f, t->arena_in.data(), (int64_t)t->arena_in.size(), reltol, abstol, t->acc, ptrs.data(),
                                           caps.data(), t->arena_out.data(), cap, &used));
    const int32_t* o = t->arena_out.data();
    const int status = o[0], next = o[1];
    t->ndevsweeps++;
    for (int b0 = 0; b0 < next && b0 < N - 1; ++b0) {
        const int32_t* rec = o + kArenaHdr + (size_t)kArenaStep * b0;
        const int r = rec[0];
        const int64_t nIb = t->I[b0].n;                    // already the new Iset[b0] (set by the previous step)
        IdxTable& nI = t->I[b0 + 1]; IdxTable& nJ = t->J[b0];
        nI.n = r; nI.v.assign(o + rec[1], o + rec[1] + (size_t)r * nI.len);
        nJ.n = r; nJ.v.assign(o + rec[2], o + rec[2] + (size_t)r * nJ.len);
        std::vector<double> errs((size_t)r + 1);
        std::memcpy(errs.data(), o + rec[3], sizeof(double) * ((size_t)r + 1));
        updateerrors(t, b0, errs);
        Core& c = t->T[b0];
        c.chil = nIb; c.d = t->localdims[b0]; c.chir = r; c.valid = true;
        t->nevals += (int64_t)rec[4] * rec[5];
        t->nbondupdates++; t->ndevbonds++;
    }
    t->t_lu += now_s() - t0;
    if (status == 2) return fail(t, TCI_ERR_NAN_L, "lu.L contains NaNs");
    if (status == 3) return fail(t, TCI_ERR_NAN_U, "lu.U contains NaNs");
    if (status == 4) return fail(t, TCI_ERR_NAN_TENSOR, "Error: NaN in tensor T[%d]", next + 1);
    if (status == 0) {                                     // the last tensor was evaluated too
        Core& c = t->T[N - 1];
        c.chil = t->I[N - 1].n; c.d = t->localdims[N - 1]; c.chir = t->J[N - 1].n; c.valid = true;
        t->nevals += c.chil * c.d * c.chir;
        *next_step = N;
    } else {
        *next_step = next;                                 // status 1: bond `next` (or the last tensor, next == N - 1) does not fit
    }
    return TCI_OK;
}

// sweep2site! (tensorci2.jl:1195-1258)
int sweep2site(tci_tci2 t, tci_functor f, int niter, int iter1, double abstol, int64_t maxbonddim, int strategy, bool strictlynested,
               bool dofill) {
    invalidate(t);
    const int N = t->N;
    for (int it = iter1; it < iter1 + niter; ++it) {
        std::vector<IdxTable> extraI, extraJ;                                                // :1212-1218
        const bool useextra = !strictlynested && t->have_hist;
        if (useextra) { extraI.swap(t->histI); extraJ.swap(t->histJ); }
        t->histI = t->I; t->histJ = t->J; t->have_hist = true; t->nhist++;                   // push!(Iset_history, deepcopy(Iset))
        t->pivoterrors.clear();                                                              // flushpivoterror!
        const bool fwd = forwardsweep(strategy, it);
        // only the last bond's tensors can survive the call, and only without the final fill: that case keeps the per-bond path
        const bool lastkeeps = !dofill && it == iter1 + niter - 1;
        if (t->engine == 1 && t->pivotsearch == 0 && !lastkeeps && tci_dsweep_supported(t->h, f)) {
            int rc = device_halfsweep(t, f, fwd, abstol, maxbonddim, useextra ? &extraI : nullptr, useextra ? &extraJ : nullptr);
            if (rc) return rc;
            continue;
        }
        for (int step = 0; step < N - 1; ++step) {                                           // :1226-1250
            const int b0 = fwd ? step : N - 2 - step;
            // only the tensors of the very last bond update can survive this call, and only without the final fill
            const bool keep = !dofill && it == iter1 + niter - 1 && step == N - 2;
            int rc = updatepivots(t, f, b0, fwd, 1e-14, abstol, maxbonddim, useextra ? &extraI[b0 + 1] : nullptr,
                                  useextra ? &extraJ[b0] : nullptr, keep);
            if (rc) return rc;
        }
    }
    if (dofill) { int rc = fillsitetensors(t, f); if (rc) return rc; }                       // :1253-1255
    return sync_maxsample(t);
}

bool convergencecriterion(const std::vector<int64_t>& ranks, const std::vector<double>& errors, const std::vector<int64_t>& nglobal,
                          double tol, int64_t maxbonddim, int ncheck, bool checkglobal) {    // tensorci2.jl:947-966
    if ((int)errors.size() < ncheck) return false;
    const size_t s = errors.size() - ncheck;
    bool allerr = true, allglob = true, allmax = true;
    int64_t minrank = std::numeric_limits<int64_t>::max();
    for (size_t i = s; i < errors.size(); ++i) {
        allerr = allerr && (errors[i] < tol);
        allglob = allglob && (nglobal[i] == 0);
        minrank = std::min(minrank, ranks[i]);
        allmax = allmax && (maxbonddim > 0 && ranks[i] >= maxbonddim);
    }
    return (allerr && (checkglobal ? allglob : true) && minrank == ranks.back()) || allmax;
}

int64_t rank_of(const tci_tci2_s* t) {
    int64_t r = 0;
    for (int b = 1; b < t->N; ++b) r = std::max(r, t->I[b].n);
    return r;
}

// DefaultGlobalPivotFinder (globalpivotfinder.jl:205-264): nsearch random starts, one line search per site on |f - tt|;
// all sum_p d_p candidates of all starts go through ONE batch evaluation of f and ONE of the tensor train.
int find_global_pivots(tci_tci2 t, tci_functor f, double abstol, int nsearch, int maxnglobal, double tolmargin, IdxTable* found) {
    const int N = t->N;
    found->len = N; found->n = 0; found->v.clear();
    if (nsearch <= 0) return TCI_OK;
    int64_t per = 0; for (int p = 0; p < N; ++p) per += t->localdims[p];
    IdxTable pts; pts.len = N; pts.n = per * nsearch; pts.v.resize((size_t)pts.n * N);
    std::vector<int32_t> start(N);
    int32_t* w = pts.v.data();
    for (int s = 0; s < nsearch; ++s) {
        for (int p = 0; p < N; ++p) start[p] = 1 + (int32_t)(splitmix64(t->rng) % (uint64_t)t->localdims[p]);
        for (int p = 0; p < N; ++p)
            for (int v = 1; v <= t->localdims[p]; ++v) { std::memcpy(w, start.data(), sizeof(int32_t) * N); w[p] = v; w += N; }
    }
    int rc = reserve(t, t->pts, (size_t)pts.n); if (rc) return rc;
    // the points are the left set, the right set is the single empty multi-index
    TCI2_CALL(t, tci_batcheval_f64(t->h, f, t->localdims.data(), N, pts.data(), pts.n, N, nullptr, 1, 0, t->pts.p, 0, 0, nullptr));
    std::vector<double> fv((size_t)pts.n), tv((size_t)pts.n);
    TCI2_CALL(t, tci_memcpy_d2h(t->h, fv.data(), t->pts.p, sizeof(double) * (size_t)pts.n));
    std::vector<const double*> ptrs(N); std::vector<int64_t> shapes(3 * (size_t)N), lds(N);
    for (int b = 0; b < N; ++b) {
        const Core& c = t->T[b];
        if (!c.valid) return fail(t, TCI_ERR_INVALID, "site tensors are not available (call fillsitetensors! or a sweep first)");
        ptrs[b] = c.buf.p; shapes[3 * b] = c.chil; shapes[3 * b + 1] = c.d; shapes[3 * b + 2] = c.chir; lds[b] = std::max<int64_t>(c.chil * c.d, 1);
    }
    TCI2_CALL(t, tci_tt_evaluate_f64(t->h, ptrs.data(), shapes.data(), lds.data(), N, pts.data(), pts.n, tv.data()));
    for (int s = 0; s < nsearch; ++s) {
        double best = 0.0; int64_t bestk = -1;
        for (int64_t k = s * per; k < (s + 1) * per; ++k) {
            const double e = std::fabs(fv[k] - tv[k]);
            if (e > best) { best = e; bestk = k; }                          // strict '>' keeps the first maximum like the reference loop
        }
        if (best > abstol * tolmargin && bestk >= 0) {
            found->v.insert(found->v.end(), pts.row(bestk), pts.row(bestk) + N);
            found->n++;
        }
    }
    if (found->n > maxnglobal) { found->n = maxnglobal; found->v.resize((size_t)maxnglobal * N); }
    return TCI_OK;
}

// device pointers / shapes / leading dimensions of the site tensors, as tci_tt_sum_f64 / tci_tt_evaluate_f64 take them
int core_args(tci_tci2 t, std::vector<const double*>& ptrs, std::vector<int64_t>& shapes, std::vector<int64_t>& lds) {
    const int N = t->N;
    ptrs.resize((size_t)N); shapes.resize(3 * (size_t)N); lds.resize((size_t)N);
    for (int b = 0; b < N; ++b) {
        const Core& c = t->T[b];
        if (!c.valid || c.chil * c.chir == 0)
            return fail(t, TCI_ERR_INVALID, "site tensors are not available (call fillsitetensors! or a sweep first)");
        ptrs[(size_t)b] = c.buf.p; shapes[3 * (size_t)b] = c.chil; shapes[3 * (size_t)b + 1] = c.d; shapes[3 * (size_t)b + 2] = c.chir;
        lds[(size_t)b] = std::max<int64_t>(c.chil * c.d, 1);
    }
    return TCI_OK;
}

int addglobalpivots(tci_tci2 t, const int32_t* pivots, int64_t np) {                          // tensorci2.jl:335-357
    const int N = t->N;
    for (int64_t k = 0; k < np; ++k) {
        const int32_t* p = pivots + (size_t)k * N;
        for (int b = 0; b < N; ++b) {
            pushunique(t->I[b], p);                 // first b entries
            pushunique(t->J[b], p + b + 1);         // last N-b-1 entries
        }
    }
    if (np > 0) invalidate(t);
    return TCI_OK;
}

int sanitycheck(tci_tci2 t) {                                                                 // globalsearch.jl:226-233
    for (int b = 0; b + 1 < t->N; ++b)
        if (t->I[b + 1].n != t->J[b].n) return fail(t, TCI_ERR_NOT_SQUARE, "Pivot matrix at bond %d is not square!", b + 1);
    return TCI_OK;
}

} // namespace

extern "C" {

// optfirstpivot (util.jl:260-291): greedy one-site-at-a-time search for a large |f|.  The d_i candidates of a site differ
// only in coordinate i, so they go through ONE batch evaluation; keeping the first value that beats the current maximum is
// exactly the reference's sequential strict-`>` scan.
int tci_optfirstpivot_f64(tci_handle h, tci_functor f, const int32_t* h_localdims, int nsites, int32_t* h_pivot, int maxsweep) {
    if (!h || !f || !h_localdims || !h_pivot || nsites < 1) return TCI_ERR_INVALID;
    int dmax = 1;
    for (int b = 0; b < nsites; ++b) {
        if (h_localdims[b] < 1 || h_pivot[b] < 1 || h_pivot[b] > h_localdims[b]) return TCI_ERR_INVALID;
        dmax = std::max(dmax, (int)h_localdims[b]);
    }
    void* d_vals = nullptr;
    int rc = tci_malloc_pooled(h, &d_vals, sizeof(double) * (size_t)dmax);
    if (rc) return rc;
    std::vector<double> vals((size_t)dmax);
    std::vector<int32_t> pts((size_t)dmax * nsites);
    auto eval = [&](int count) -> int {                 // the points are the left set, the right set is the empty index
        int r = tci_batcheval_f64(h, f, h_localdims, nsites, pts.data(), count, nsites, nullptr, 1, 0, (double*)d_vals, 0, 0, nullptr);
        if (r) return r;
        return tci_memcpy_d2h(h, vals.data(), d_vals, sizeof(double) * (size_t)count);
    };
    std::memcpy(pts.data(), h_pivot, sizeof(int32_t) * nsites);
    rc = eval(1);
    double valf = std::fabs(vals[0]);
    for (int sweep = 0; sweep < maxsweep && rc == TCI_OK; ++sweep) {
        const double prev = valf;
        for (int i = 0; i < nsites && rc == TCI_OK; ++i) {
            const int d = h_localdims[i];
            for (int v = 0; v < d; ++v) {
                std::memcpy(pts.data() + (size_t)v * nsites, h_pivot, sizeof(int32_t) * nsites);
                pts[(size_t)v * nsites + i] = v + 1;
            }
            rc = eval(d);
            if (rc) break;
            for (int v = 0; v < d; ++v)
                if (std::fabs(vals[v]) > valf) { valf = std::fabs(vals[v]); h_pivot[i] = v + 1; }
        }
        if (prev == valf) break;
    }
    tci_free_pooled(h, d_vals);
    return rc;
}

int tci_tci2_create(tci_handle h, const int32_t* h_localdims, int nsites, tci_tci2* out) {
    if (!h || !h_localdims || !out || nsites < 2) return TCI_ERR_INVALID;                     // "localdims should have at least 2 elements!"
    for (int b = 0; b < nsites; ++b) if (h_localdims[b] < 1) return TCI_ERR_INVALID;
    tci_tci2 t = new tci_tci2_s;
    t->h = h; t->N = nsites;
    t->localdims.assign(h_localdims, h_localdims + nsites);
    t->I.resize(nsites); t->J.resize(nsites); t->T.resize(nsites);
    for (int b = 0; b < nsites; ++b) { t->I[b].len = b; t->J[b].len = nsites - b - 1; }
    t->bonderrors.assign((size_t)nsites - 1, 0.0);
    if (tci_malloc_pooled(h, &t->acc, 64) != TCI_OK || tci_maxacc_set(h, t->acc, 0.0) != TCI_OK) {
        if (t->acc) tci_free_pooled(h, t->acc);
        delete t;
        return TCI_ERR_NOMEM;
    }
    *out = t;
    return TCI_OK;
}

int tci_tci2_destroy(tci_tci2 t) {
    if (!t) return TCI_OK;
    DevBuf* bufs[] = {&t->pi, &t->L, &t->U, &t->pi1, &t->pmat, &t->pts};
    for (DevBuf* b : bufs) if (b->p) tci_free_pooled(t->h, b->p);
    for (auto& c : t->T) if (c.buf.p) tci_free_pooled(t->h, c.buf.p);
    if (t->acc) tci_free_pooled(t->h, t->acc);
    delete t;
    return TCI_OK;
}

const char* tci_tci2_last_error(tci_tci2 t) { return t ? t->err.c_str() : "NULL tci_tci2"; }

int tci_tci2_seed(tci_tci2 t, uint64_t seed) { if (!t) return TCI_ERR_INVALID; t->rng = seed; return TCI_OK; }

int tci_tci2_set_engine(tci_tci2 t, int engine) {
    if (!t || (engine != 0 && engine != 1)) return TCI_ERR_INVALID;
    t->engine = engine;
    return TCI_OK;
}

int tci_tci2_set_pivotsearch(tci_tci2 t, int rook, uint64_t seed) {
    if (!t || (rook != 0 && rook != 1)) return TCI_ERR_INVALID;
    t->pivotsearch = rook; t->rookrng = seed;
    return TCI_OK;
}

int tci_tci2_addglobalpivots(tci_tci2 t, const int32_t* h_pivots, int64_t npivots) {
    if (!t || npivots < 0 || (npivots > 0 && !h_pivots)) return TCI_ERR_INVALID;
    for (int64_t k = 0; k < npivots; ++k)
        for (int b = 0; b < t->N; ++b) {
            const int32_t v = h_pivots[(size_t)k * t->N + b];
            if (v < 1 || v > t->localdims[b]) return fail(t, TCI_ERR_INVALID, "pivot %lld: index %d at site %d is outside 1:%d", (long long)k + 1, v, b + 1, t->localdims[b]);
        }
    return addglobalpivots(t, h_pivots, npivots);
}

int tci_tci2_init_maxsample(tci_tci2 t, tci_functor f, const int32_t* h_pivots, int64_t npivots) {
    if (!t || !f || npivots < 0 || (npivots > 0 && !h_pivots)) return TCI_ERR_INVALID;
    double m = -std::numeric_limits<double>::infinity();
    if (npivots > 0) {                                                                        // maximum(abs(f(p)) for p in pivots), :112
        int rc = reserve(t, t->pts, (size_t)npivots); if (rc) return rc;
        double mx = 0.0;
        TCI2_CALL(t, tci_batcheval_f64(t->h, f, t->localdims.data(), t->N, h_pivots, npivots, t->N, nullptr, 1, 0, t->pts.p, 0, 0, &mx));
        m = jl_max(m, mx);
    }
    int rc = set_maxsample(t, m); if (rc) return rc;
    if (!(std::fabs(m) > 0.0)) return fail(t, TCI_ERR_INVALID, "maxsamplevalue is zero!");     // :113
    invalidate(t);
    return TCI_OK;
}

int tci_tci2_set_indexset(tci_tci2 t, int which, int site, const int32_t* h_rows, int64_t count) {
    if (!t || site < 1 || site > t->N || count < 0 || (which != 0 && which != 1)) return TCI_ERR_INVALID;
    IdxTable& tab = which == 0 ? t->I[site - 1] : t->J[site - 1];
    if (count > 0 && tab.len > 0 && !h_rows) return TCI_ERR_INVALID;
    tab.n = count;
    tab.v.assign(h_rows, h_rows + (tab.len > 0 ? (size_t)count * tab.len : 0));
    invalidate(t);
    return TCI_OK;
}

int tci_tci2_indexset(tci_tci2 t, int which, int site, int32_t* h_rows, int64_t* h_count) {
    if (!t || site < 1 || site > t->N || (which != 0 && which != 1)) return TCI_ERR_INVALID;
    const IdxTable& tab = which == 0 ? t->I[site - 1] : t->J[site - 1];
    if (h_count) *h_count = tab.n;
    if (h_rows && !tab.v.empty()) std::memcpy(h_rows, tab.v.data(), sizeof(int32_t) * tab.v.size());
    return TCI_OK;
}

int tci_tci2_set_maxsamplevalue(tci_tci2 t, double v) { if (!t) return TCI_ERR_INVALID; return set_maxsample(t, v); }

int tci_tci2_updatepivots(tci_tci2 t, tci_functor f, int bond, int leftorthogonal, double reltol, double abstol, int64_t maxbonddim,
                          const int32_t* h_extraI, int64_t nextraI, const int32_t* h_extraJ, int64_t nextraJ) {
    if (!t || !f || bond < 1 || bond > t->N - 1 || nextraI < 0 || nextraJ < 0) return TCI_ERR_INVALID;
    IdxTable eI, eJ;
    eI.len = bond; eI.n = nextraI; if (nextraI > 0) eI.v.assign(h_extraI, h_extraI + (size_t)nextraI * eI.len);
    eJ.len = t->N - bond; eJ.n = nextraJ; if (nextraJ > 0) eJ.v.assign(h_extraJ, h_extraJ + (size_t)nextraJ * eJ.len);
    int rc = updatepivots(t, f, bond - 1, leftorthogonal != 0, reltol, abstol, maxbonddim, nextraI ? &eI : nullptr, nextraJ ? &eJ : nullptr);
    if (rc) return rc;
    return sync_maxsample(t);
}

int tci_tci2_sweep2site(tci_tci2 t, tci_functor f, int niter, int iter1, double abstol, int64_t maxbonddim, int sweepstrategy,
                        int strictlynested, int fillsitetensors_) {
    if (!t || !f || niter < 0) return TCI_ERR_INVALID;
    return sweep2site(t, f, niter, iter1, abstol, maxbonddim, sweepstrategy, strictlynested != 0, fillsitetensors_ != 0);
}

int tci_tci2_sweep1site(tci_tci2 t, tci_functor f, int forward, double reltol, double abstol, int64_t maxbonddim, int updatetensors) {
    if (!t || !f) return TCI_ERR_INVALID;
    return sweep1site(t, f, forward != 0, reltol, abstol, maxbonddim, updatetensors != 0);
}

int tci_tci2_makecanonical(tci_tci2 t, tci_functor f, double reltol, double abstol, int64_t maxbonddim) {
    if (!t || !f) return TCI_ERR_INVALID;
    return makecanonical(t, f, reltol, abstol, maxbonddim);
}

int tci_tci2_fillsitetensors(tci_tci2 t, tci_functor f) {
    if (!t || !f) return TCI_ERR_INVALID;
    int rc = fillsitetensors(t, f); if (rc) return rc;
    return sync_maxsample(t);
}

int tci_tci2_optimize(tci_tci2 t, tci_functor f, double tolerance, int64_t maxbonddim, int maxiter, int sweepstrategy, int normalizeerror,
                      int ncheckhistory, int nsearchglobalpivot, int maxnglobalpivot, double tolmarginglobalsearch, int strictlynested,
                      int checkconvglobalpivot, int64_t* h_ranks, double* h_errors, int* h_niter) {
    if (!t || !f || maxiter < 0) return TCI_ERR_INVALID;
    if (nsearchglobalpivot > 0 && nsearchglobalpivot < maxnglobalpivot) return fail(t, TCI_ERR_INVALID, "nsearchglobalpivot < maxnglobalpivot!");
    if (maxbonddim <= 0 && !(tolerance > 0.0))
        return fail(t, TCI_ERR_INVALID, "Specify either tolerance > 0 or some maxbonddim; otherwise, the convergence criterion is not reachable!");
    std::vector<int64_t> ranks, nglobal;
    std::vector<double> errors;
    for (int it = 1; it <= maxiter; ++it) {                                                   // tensorci2.jl:1084-1154
        const double norm = normalizeerror ? t->maxsamplevalue : 1.0;
        const double abstol = tolerance * norm;
        int rc = sweep2site(t, f, 2, 1, abstol, maxbonddim, sweepstrategy, strictlynested != 0, true); if (rc) return rc;
        errors.push_back(pivoterror(t));
        IdxTable found;
        rc = find_global_pivots(t, f, abstol, nsearchglobalpivot, maxnglobalpivot, tolmarginglobalsearch, &found); if (rc) return rc;
        rc = addglobalpivots(t, found.data(), found.n); if (rc) return rc;
        nglobal.push_back(found.n);
        ranks.push_back(rank_of(t));
        if (convergencecriterion(ranks, errors, nglobal, abstol, maxbonddim, ncheckhistory, checkconvglobalpivot != 0)) break;
    }
    const double norm = normalizeerror ? t->maxsamplevalue : 1.0;                             // :1157-1171
    int rc = sweep1site(t, f, true, 1e-14, tolerance * norm, maxbonddim, true); if (rc) return rc;
    rc = sanitycheck(t); if (rc) return rc;
    for (size_t i = 0; i < ranks.size(); ++i) {
        if (h_ranks) h_ranks[i] = ranks[i];
        if (h_errors) h_errors[i] = errors[i] / norm;
    }
    if (h_niter) *h_niter = (int)ranks.size();
    return TCI_OK;
}

int tci_tci2_core(tci_tci2 t, int site, const double** d_core, int64_t* h_shape3) {
    if (!t || site < 1 || site > t->N) return TCI_ERR_INVALID;
    const Core& c = t->T[site - 1];
    if (!c.valid) return fail(t, TCI_ERR_INVALID, "site tensors are not available (call fillsitetensors! or a sweep first)");
    if (d_core) *d_core = c.buf.p;
    if (h_shape3) { h_shape3[0] = c.chil; h_shape3[1] = c.d; h_shape3[2] = c.chir; }
    return TCI_OK;
}

/* evaluate(tci, indexset) (src/abstracttensortrain.jl:328-342) for a batch of points, and sum(tci) (:428-441), on the object's
 * own site tensors: with these a C or Julia caller gets from a device functor to an interpolated value or an integral without
 * any other host library. */
int tci_tci2_evaluate(tci_tci2 t, const int32_t* h_points, int64_t npoints, double* h_out) {
    if (!t || npoints < 0 || (npoints > 0 && (!h_points || !h_out))) return TCI_ERR_INVALID;
    std::vector<const double*> ptrs; std::vector<int64_t> shapes, lds;
    int rc = core_args(t, ptrs, shapes, lds); if (rc) return rc;
    if (npoints == 0) return TCI_OK;
    TCI2_CALL(t, tci_tt_evaluate_f64(t->h, ptrs.data(), shapes.data(), lds.data(), t->N, h_points, npoints, h_out));
    return TCI_OK;
}

int tci_tci2_sum(tci_tci2 t, double* h_out) {
    if (!t || !h_out) return TCI_ERR_INVALID;
    std::vector<const double*> ptrs; std::vector<int64_t> shapes, lds;
    int rc = core_args(t, ptrs, shapes, lds); if (rc) return rc;
    TCI2_CALL(t, tci_tt_sum_f64(t->h, ptrs.data(), shapes.data(), lds.data(), t->N, h_out));
    return TCI_OK;
}

/* Everything a host wants after optimize!, in two calls instead of 4 N: all site tensors copied (device to device, enqueued, no
 * synchronisation) into ONE caller buffer, dense and back to back — h_shapes[3b..] = (chi_l, d, chi_r), h_offsets[b] = first
 * element of tensor b in d_dst (query the total with d_dst = NULL first) — and all index sets packed into one int32 buffer. */
int tci_tci2_export_cores(tci_tci2 t, double* d_dst, int64_t capacity, int64_t* h_shapes, int64_t* h_offsets, int64_t* h_total) {
    if (!t || !h_shapes || !h_offsets || !h_total) return TCI_ERR_INVALID;
    int64_t off = 0;
    for (int b = 0; b < t->N; ++b) {
        const Core& c = t->T[b];
        if (!c.valid) return fail(t, TCI_ERR_INVALID, "site tensors are not available (call fillsitetensors! or a sweep first)");
        h_shapes[3 * b] = c.chil; h_shapes[3 * b + 1] = c.d; h_shapes[3 * b + 2] = c.chir;
        h_offsets[b] = off;
        off += c.chil * c.d * c.chir;
    }
    *h_total = off;
    if (!d_dst) return TCI_OK;
    if (capacity < off) return fail(t, TCI_ERR_INVALID, "tci_tci2_export_cores: buffer too small");
    for (int b = 0; b < t->N; ++b) {
        const Core& c = t->T[b];
        const int64_t sz = c.chil * c.d * c.chir;
        if (sz > 0) TCI2_CALL(t, tci_memcpy_d2d(t->h, d_dst + h_offsets[b], c.buf.p, sizeof(double) * (size_t)sz));
    }
    return TCI_OK;
}

int tci_tci2_export_indexsets(tci_tci2 t, int32_t* h_dst, int64_t capacity, int64_t* h_counts2, int64_t* h_total) {
    if (!t || !h_counts2 || !h_total) return TCI_ERR_INVALID;            /* h_counts2[2b] = |Iset[b]|, [2b+1] = |Jset[b]|; I_0, J_0, I_1, ... */
    int64_t off = 0;
    for (int b = 0; b < t->N; ++b) {
        h_counts2[2 * b] = t->I[b].n; h_counts2[2 * b + 1] = t->J[b].n;
        off += (int64_t)t->I[b].v.size() + (int64_t)t->J[b].v.size();
    }
    *h_total = off;
    if (!h_dst) return TCI_OK;
    if (capacity < off) return fail(t, TCI_ERR_INVALID, "tci_tci2_export_indexsets: buffer too small");
    int32_t* w = h_dst;
    for (int b = 0; b < t->N; ++b) {
        if (!t->I[b].v.empty()) { std::memcpy(w, t->I[b].v.data(), sizeof(int32_t) * t->I[b].v.size()); w += t->I[b].v.size(); }
        if (!t->J[b].v.empty()) { std::memcpy(w, t->J[b].v.data(), sizeof(int32_t) * t->J[b].v.size()); w += t->J[b].v.size(); }
    }
    return TCI_OK;
}

int tci_tci2_sitetensors_available(tci_tci2 t) {
    if (!t) return 0;
    for (const auto& c : t->T) if (!c.valid || c.chil * c.chir == 0) return 0;
    return 1;
}

int tci_tci2_errors(tci_tci2 t, double* h_bonderrors, double* h_pivoterrors, int64_t* h_npivoterrors, double* h_maxsamplevalue) {
    if (!t) return TCI_ERR_INVALID;
    if (h_bonderrors) std::copy(t->bonderrors.begin(), t->bonderrors.end(), h_bonderrors);
    if (h_pivoterrors) std::copy(t->pivoterrors.begin(), t->pivoterrors.end(), h_pivoterrors);
    if (h_npivoterrors) *h_npivoterrors = (int64_t)t->pivoterrors.size();
    if (h_maxsamplevalue) *h_maxsamplevalue = t->maxsamplevalue;
    return TCI_OK;
}

int tci_tci2_stats(tci_tci2 t, int64_t* h_counts2, double* h_seconds4) {
    if (!t) return TCI_ERR_INVALID;
    if (h_counts2) { h_counts2[0] = t->nevals; h_counts2[1] = t->nbondupdates; }
    if (h_seconds4) { h_seconds4[0] = t->t_eval; h_seconds4[1] = t->t_lu; h_seconds4[2] = t->t_factors; h_seconds4[3] = t->t_host; }
    return TCI_OK;
}

} // extern "C"
