// mock_cabi.cpp — TEST INFRASTRUCTURE ONLY.  A host-memory stand-in for the handful of C-ABI entry points that
// csrc/tci2_driver.cpp calls, backed by the CPU oracle (oracle/_build/libtci_oracle.so).  Linking the UNCHANGED driver
// source against this file lets the driver's bookkeeping (index sets, sweep order, histories, error merging, convergence)
// be checked bit for bit against the oracle's own TensorCI2 on a machine without a GPU (tests/test_native_driver_mock.py).
// "Device" pointers are plain host pointers here; a tci_functor is an oracle Func*.  Nothing in the product links this.
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/tci_b200.h"

extern "C" {
const char* orc_last_error();
int orc_rrlu(const double* A, int64_t m, int64_t n, int64_t lda, int64_t maxrank, double reltol, double abstol, int leftorthogonal,
             int64_t* rowperm, int64_t* colperm, int64_t* npivot, double* error, double* L, int64_t ldl, double* U, int64_t ldu,
             double* pivoterrors);
int orc_luci_factors(int64_t m, int64_t n, int64_t r, int leftorthogonal, const int64_t* rowperm, const int64_t* colperm,
                     const double* L, int64_t ldl, const double* U, int64_t ldu, double* left, double* right);
int orc_solve_right(const double* P, int64_t r, const double* B, int64_t q, double* X);
int orc_batcheval(void* f, const int32_t* localdims, int nsites, const int32_t* left, int64_t nleft, int nl, const int32_t* right,
                  int64_t nright, int nr, int M, double* out);
int orc_tt_evaluate(const double* cores, const int64_t* shapes, int nsites, const int32_t* idx, int64_t npoints, double* out);
int orc_tt_sum(const double* cores, const int64_t* shapes, int nsites, double* out);
}

struct tci_context { std::string err; int64_t ncalls = 0; };

static int mfail(tci_handle h, int code, const char* what) { if (h) h->err = what; return code; }

extern "C" {

int mock_create(tci_handle* out) { *out = new tci_context; return TCI_OK; }
int mock_destroy(tci_handle h) { delete h; return TCI_OK; }
int64_t mock_ncalls(tci_handle h) { return h->ncalls; }

const char* tci_last_error(tci_handle h) { return h ? h->err.c_str() : ""; }
int tci_malloc(tci_handle, void** p, size_t bytes) { *p = std::malloc(bytes ? bytes : 1); return *p ? TCI_OK : TCI_ERR_NOMEM; }
int tci_free(tci_handle, void* p) { std::free(p); return TCI_OK; }
int tci_malloc_pooled(tci_handle h, void** p, size_t bytes) { return tci_malloc(h, p, bytes); }
int tci_free_pooled(tci_handle h, void* p) { return tci_free(h, p); }
int tci_memcpy_d2h(tci_handle, void* dst, const void* src, size_t bytes) { std::memcpy(dst, src, bytes); return TCI_OK; }
int tci_memcpy_d2d(tci_handle, void* dst, const void* src, size_t bytes) { std::memcpy(dst, src, bytes); return TCI_OK; }

static void merge_maxabs(double* h_maxabs, const double* x, size_t n) {
    if (!h_maxabs) return;
    double m = std::fabs(*h_maxabs);
    bool nan = m != m;
    for (size_t i = 0; i < n; ++i) { const double a = std::fabs(x[i]); if (a != a) nan = true; else if (a > m) m = a; }
    *h_maxabs = nan ? NAN : m;
}

int tci_batcheval_f64(tci_handle h, tci_functor f, const int32_t* h_localdims, int nsites, const int32_t* h_left, int64_t nleft, int nl,
                      const int32_t* h_right, int64_t nright, int nr, double* d_out, int64_t stride_c, int64_t stride_j, double* h_maxabs) {
    h->ncalls++;
    const int M = nsites - nl - nr;
    if (M < 0) return mfail(h, TCI_ERR_INVALID, "Invalid number of central indices");
    if (nleft == 0 || nright == 0) return TCI_OK;
    int64_t ncent = 1;
    for (int x = 0; x < M; ++x) ncent *= h_localdims[nl + x];
    std::vector<double> dense((size_t)(nleft * ncent * nright));
    static const int32_t dummy = 0;
    if (orc_batcheval((void*)f, h_localdims, nsites, h_left ? h_left : &dummy, nleft, nl, h_right ? h_right : &dummy, nright, nr, M, dense.data()))
        return mfail(h, TCI_ERR_INVALID, orc_last_error());
    const int64_t sc = stride_c > 0 ? stride_c : nleft, sj = stride_j > 0 ? stride_j : sc * ncent;
    for (int64_t j = 0; j < nright; ++j)
        for (int64_t c = 0; c < ncent; ++c)
            for (int64_t i = 0; i < nleft; ++i) d_out[i + c * sc + j * sj] = dense[(size_t)(i + nleft * (c + ncent * j))];
    merge_maxabs(h_maxabs, dense.data(), dense.size());
    return TCI_OK;
}

struct MockAcc { double mx; int nan; int pad; };         // same 16 bytes as the device accumulator: bits of the maximum, NaN flag

int tci_maxacc_set(tci_handle, void* acc, double v) {
    MockAcc* a = (MockAcc*)acc; a->mx = 0.0; a->nan = 0; a->pad = 0;
    if (v != v) a->nan = 1; else a->mx = std::fabs(v);
    return TCI_OK;
}
int tci_maxacc_get(tci_handle, const void* acc, double* out) {
    const MockAcc* a = (const MockAcc*)acc;
    *out = a->nan ? NAN : a->mx;
    return TCI_OK;
}
int tci_batcheval_acc_f64(tci_handle h, tci_functor f, const int32_t* h_localdims, int nsites, const int32_t* h_left, int64_t nleft, int nl,
                          const int32_t* h_right, int64_t nright, int nr, double* d_out, int64_t stride_c, int64_t stride_j, void* acc) {
    MockAcc* a = (MockAcc*)acc;
    double m = a->nan ? NAN : a->mx;
    const int rc = tci_batcheval_f64(h, f, h_localdims, nsites, h_left, nleft, nl, h_right, nright, nr, d_out, stride_c, stride_j, &m);
    if (rc) return rc;
    if (m != m) a->nan = 1; else a->mx = m;
    return TCI_OK;
}

int tci_rrlu_inplace_f64(tci_handle h, double* d_A, int64_t m, int64_t n, int64_t lda, int64_t maxrank, double reltol, double abstol,
                         int leftorthogonal, int64_t* h_rowperm, int64_t* h_colperm, int64_t* h_npivot, double* h_error, double* d_L,
                         int64_t ldl, double* d_U, int64_t ldu, double* h_pivoterrors) {
    h->ncalls++;
    const int64_t mn = m < n ? m : n;
    const int64_t mr = (maxrank <= 0 || maxrank > mn) ? mn : maxrank;
    if (mn == 0) { *h_npivot = 0; *h_error = 0.0; if (h_pivoterrors) h_pivoterrors[0] = 0.0; return TCI_OK; }
    const int st = orc_rrlu(d_A, m, n, lda, mr, reltol, abstol, leftorthogonal, h_rowperm, h_colperm, h_npivot, h_error, d_L, ldl, d_U, ldu,
                            h_pivoterrors);
    if (st == -1) return mfail(h, TCI_ERR_NAN_L, "lu.L contains NaNs");
    if (st == -2) return mfail(h, TCI_ERR_NAN_U, "lu.U contains NaNs");
    if (st) return mfail(h, TCI_ERR_INVALID, orc_last_error());
    return TCI_OK;
}

int tci_luci_factors_f64(tci_handle h, int64_t m, int64_t n, int64_t r, int leftorthogonal, const int64_t* h_rowperm, const int64_t* h_colperm,
                         const double* d_L, int64_t ldl, const double* d_U, int64_t ldu, double* d_left, int64_t ldleft, double* d_right,
                         int64_t ldright) {
    h->ncalls++;
    if (r == 0 || (!d_left && !d_right)) return TCI_OK;
    std::vector<double> left((size_t)(m * r)), right((size_t)(r * n));
    if (orc_luci_factors(m, n, r, leftorthogonal, h_rowperm, h_colperm, d_L, ldl, d_U, ldu, left.data(), right.data()))
        return mfail(h, TCI_ERR_INVALID, orc_last_error());
    if (d_left) for (int64_t k = 0; k < r; ++k) std::memcpy(d_left + k * ldleft, &left[(size_t)(k * m)], sizeof(double) * m);
    if (d_right) for (int64_t j = 0; j < n; ++j) std::memcpy(d_right + j * ldright, &right[(size_t)(j * r)], sizeof(double) * r);
    return TCI_OK;
}

int tci_solve_right_f64(tci_handle h, const double* d_B, int64_t q, int64_t ldb, const double* d_P, int64_t r, int64_t ldp, double* d_X,
                        int64_t ldx) {
    h->ncalls++;
    if (q == 0 || r == 0) return TCI_OK;
    std::vector<double> P((size_t)(r * r)), B((size_t)(q * r)), X((size_t)(q * r));
    for (int64_t j = 0; j < r; ++j) {
        std::memcpy(&P[(size_t)(j * r)], d_P + j * ldp, sizeof(double) * r);
        std::memcpy(&B[(size_t)(j * q)], d_B + j * ldb, sizeof(double) * q);
    }
    if (orc_solve_right(P.data(), r, B.data(), q, X.data())) return mfail(h, TCI_ERR_INVALID, "tci_solve_right_f64: pivot matrix is singular");
    for (int64_t j = 0; j < r; ++j) std::memcpy(d_X + j * ldx, &X[(size_t)(j * q)], sizeof(double) * q);
    return TCI_OK;
}

// X * T = B in place on B (q x r), T triangular; same operation order as the checker's cols2Lmatrix / rows2Umatrix
// (column k scaled by the diagonal, then eliminated from the columns that follow), so the stand-in is bit-identical to it.
int tci_trsm_right_f64(tci_handle h, const double* T, int64_t st_row, int64_t st_col, int64_t r, int lower, int unit, double* B,
                       int64_t sb_row, int64_t sb_col, int64_t q) {
    h->ncalls++;
    auto t = [&](int64_t i, int64_t j) { return T[i * st_row + j * st_col]; };
    auto b = [&](int64_t i, int64_t j) -> double& { return B[i * sb_row + j * sb_col]; };
    for (int64_t kk = 0; kk < r; ++kk) {
        const int64_t k = lower ? r - 1 - kk : kk;
        if (!unit) for (int64_t i = 0; i < q; ++i) b(i, k) = b(i, k) / t(k, k);
        for (int64_t j = lower ? 0 : k + 1; j < (lower ? k : r); ++j) {
            const double y = t(k, j);
            for (int64_t i = 0; i < q; ++i) b(i, j) = b(i, j) - b(i, k) * y;
        }
    }
    return TCI_OK;
}

int tci_maxabs_f64(tci_handle h, const double* d_x, int64_t rows, int64_t cols, int64_t ld, double* h_maxabs) {
    h->ncalls++;
    for (int64_t j = 0; j < cols; ++j) merge_maxabs(h_maxabs, d_x + j * ld, (size_t)rows);
    return TCI_OK;
}

int tci_tt_evaluate_f64(tci_handle h, const double* const* h_core_ptrs, const int64_t* h_shapes, const int64_t* h_lds, int nsites,
                        const int32_t* h_indices, int64_t npoints, double* h_out) {
    h->ncalls++;
    std::vector<double> flat;
    for (int b = 0; b < nsites; ++b) {
        const int64_t rows = h_shapes[3 * b] * h_shapes[3 * b + 1], cols = h_shapes[3 * b + 2], ld = h_lds ? h_lds[b] : rows;
        for (int64_t j = 0; j < cols; ++j) flat.insert(flat.end(), h_core_ptrs[b] + j * ld, h_core_ptrs[b] + j * ld + rows);
    }
    if (orc_tt_evaluate(flat.data(), h_shapes, nsites, h_indices, npoints, h_out)) return mfail(h, TCI_ERR_INVALID, orc_last_error());
    return TCI_OK;
}

int tci_tt_sum_f64(tci_handle h, const double* const* h_core_ptrs, const int64_t* h_shapes, const int64_t* h_lds, int nsites, double* h_out) {
    h->ncalls++;
    std::vector<double> flat;
    for (int b = 0; b < nsites; ++b) {
        const int64_t rows = h_shapes[3 * b] * h_shapes[3 * b + 1], cols = h_shapes[3 * b + 2], ld = h_lds ? h_lds[b] : rows;
        for (int64_t j = 0; j < cols; ++j) flat.insert(flat.end(), h_core_ptrs[b] + j * ld, h_core_ptrs[b] + j * ld + rows);
    }
    if (orc_tt_sum(flat.data(), h_shapes, nsites, h_out)) return mfail(h, TCI_ERR_INVALID, orc_last_error());
    return TCI_OK;
}

// ---- tci_dsweep_*: the arena protocol of the device-resident half-sweep (csrc/sweep.cu) restated on the host with the oracle's
// arithmetic, so the driver's packing / unpacking / fall-back logic is checked without a GPU.  `mock_set_dsweep_limit(h, e)`
// makes bonds with more than e Pi entries "not fit" (status 1), which exercises the driver's per-bond fall-back.
static int64_t g_dsweep_limit = 1 << 30;
void mock_set_dsweep_limit(tci_handle, int64_t elems) { g_dsweep_limit = elems; }
int tci_dsweep_supported(tci_handle, tci_functor f) { return f != nullptr; }

int tci_dsweep_halfsweep_f64(tci_handle h, tci_functor f, const int32_t* in, int64_t, double reltol, double abstol, void* acc,
                             int32_t* out, int64_t out_cap, int64_t* used_words) {
    h->ncalls++;
    const int N = in[0], fwd = in[1], s0 = in[2], s1 = in[3], maxbond = in[4], useextra = in[5];
    const int32_t* ld = in + 8; const int32_t* desc = ld + N;
    int64_t used = 8 + 6LL * (s1 - s0);
    out[0] = 0; out[1] = s1; out[2] = (int)used;
    typedef std::vector<std::vector<int32_t>> Rows;
    auto rows_of = [&](int cnt, int off, int len) { Rows r((size_t)cnt); for (int i = 0; i < cnt; ++i) r[(size_t)i].assign(in + off + (size_t)i * len, in + off + (size_t)(i + 1) * len); return r; };
    Rows mov; bool have_mov = false;
    MockAcc* a = (MockAcc*)acc;
    for (int s = s0; s < s1; ++s) {
        const int b0 = fwd ? s : N - 2 - s;
        const int32_t* dI = desc + 12 * b0; const int32_t* dJ = desc + 12 * (b0 + 1);
        Rows I = rows_of(dI[0], dI[1], b0), J = rows_of(dJ[2], dJ[3], N - b0 - 2);
        if (have_mov) { if (fwd) I = mov; else J = mov; }
        Rows Ic, Jc;
        for (int d = 1; d <= ld[b0]; ++d) for (auto r : I) { r.push_back(d); Ic.push_back(r); }
        for (const auto& r : J) for (int d = 1; d <= ld[b0 + 1]; ++d) { std::vector<int32_t> x; x.push_back(d); x.insert(x.end(), r.begin(), r.end()); Jc.push_back(x); }
        if (useextra) {
            const int32_t* dXI = desc + 12 * (b0 + 1); const int32_t* dXJ = desc + 12 * b0;
            for (const auto& r : rows_of(dXI[4], dXI[5], b0 + 1)) { bool dup = false; for (const auto& y : Ic) if (y == r) { dup = true; break; } if (!dup) Ic.push_back(r); }
            for (const auto& r : rows_of(dXJ[6], dXJ[7], N - b0 - 1)) { bool dup = false; for (const auto& y : Jc) if (y == r) { dup = true; break; } if (!dup) Jc.push_back(r); }
        }
        const int64_t m = (int64_t)Ic.size(), n = (int64_t)Jc.size(), mn = m < n ? m : n;
        const int64_t rcap = (maxbond > 0 && maxbond < mn) ? maxbond : mn;
        const int lenA = b0 + 1, lenB = N - b0 - 1;
        if (m * n > g_dsweep_limit || used + rcap * (lenA + lenB) + 2 * (rcap + 1) + 2 > out_cap) { out[0] = 1; out[1] = s; break; }
        std::vector<int32_t> lt, rt;
        for (const auto& r : Ic) lt.insert(lt.end(), r.begin(), r.end());
        for (const auto& r : Jc) rt.insert(rt.end(), r.begin(), r.end());
        std::vector<double> Pi((size_t)(m * n));
        if (orc_batcheval((void*)f, ld, N, lt.data(), m, lenA, rt.data(), n, lenB, 0, Pi.data())) return mfail(h, TCI_ERR_INVALID, orc_last_error());
        double mx = a->nan ? NAN : a->mx;
        merge_maxabs(&mx, Pi.data(), Pi.size());
        if (mx != mx) a->nan = 1; else a->mx = mx;
        std::vector<int64_t> rp((size_t)m), cp((size_t)n);
        std::vector<double> pe((size_t)rcap + 1);
        int64_t npiv = 0; double err = 0.0;
        const int st = orc_rrlu(Pi.data(), m, n, m, rcap, reltol, abstol, fwd, rp.data(), cp.data(), &npiv, &err, nullptr, m, nullptr, rcap, pe.data());
        if (st == -1) { out[0] = 2; out[1] = s; break; }
        if (st == -2) { out[0] = 3; out[1] = s; break; }
        if (st) return mfail(h, TCI_ERR_INVALID, orc_last_error());
        const int64_t offA = used, offB = offA + npiv * lenA; int64_t offE = offB + npiv * lenB; offE += offE & 1;
        Rows newI, newJ;
        for (int64_t k = 0; k < npiv; ++k) { newI.push_back(Ic[(size_t)(rp[(size_t)k] - 1)]); newJ.push_back(Jc[(size_t)(cp[(size_t)k] - 1)]); }
        for (int64_t k = 0; k < npiv; ++k) { std::memcpy(out + offA + k * lenA, newI[(size_t)k].data(), sizeof(int32_t) * lenA); std::memcpy(out + offB + k * lenB, newJ[(size_t)k].data(), sizeof(int32_t) * lenB); }
        std::memcpy(out + offE, pe.data(), sizeof(double) * (size_t)(npiv + 1));
        int32_t* rec = out + 8 + 6 * (s - s0);
        rec[0] = (int)npiv; rec[1] = (int)offA; rec[2] = (int)offB; rec[3] = (int)offE; rec[4] = (int)m; rec[5] = (int)n;
        used = offE + 2 * (npiv + 1);
        out[2] = (int)used;
        mov = fwd ? newI : newJ; have_mov = true;
    }
    if (used_words) *used_words = used;
    return TCI_OK;
}

int tci_dsweep_fill_f64(tci_handle h, tci_functor f, const int32_t* in, int64_t, void* acc, double* const* cores) {
    h->ncalls++;
    const int N = in[0];
    const int32_t* ld = in + 8; const int32_t* desc = ld + N;
    MockAcc* a = (MockAcc*)acc;
    for (int b = 0; b < N; ++b) {
        const int nI = desc[12 * b], nJ = desc[12 * b + 2], d = ld[b];
        const int32_t* tI = in + desc[12 * b + 1]; const int32_t* tJ = in + desc[12 * b + 3];
        const int64_t q = (int64_t)nI * d;
        static const int32_t dummy = 0;
        std::vector<double> pi1((size_t)(q * nJ));
        if (orc_batcheval((void*)f, ld, N, b > 0 ? tI : &dummy, nI, b, N - b - 1 > 0 ? tJ : &dummy, nJ, N - b - 1, 1, pi1.data()))
            return mfail(h, TCI_ERR_INVALID, orc_last_error());
        double mx = a->nan ? NAN : a->mx;
        merge_maxabs(&mx, pi1.data(), pi1.size());
        if (b == N - 1) { std::memcpy(cores[b], pi1.data(), sizeof(double) * pi1.size()); }
        else {
            const int nI1 = desc[12 * (b + 1)];
            if (nI1 != nJ) return mfail(h, TCI_ERR_NOT_SQUARE, "Pivot matrix at bond is not square!");
            std::vector<double> Pm((size_t)nJ * nJ), X((size_t)(q * nJ));
            if (orc_batcheval((void*)f, ld, N, in + desc[12 * (b + 1) + 1], nI1, b + 1, N - b - 1 > 0 ? tJ : &dummy, nJ, N - b - 1, 0, Pm.data()))
                return mfail(h, TCI_ERR_INVALID, orc_last_error());
            merge_maxabs(&mx, Pm.data(), Pm.size());
            if (orc_solve_right(Pm.data(), nJ, pi1.data(), q, X.data())) return mfail(h, TCI_ERR_INVALID, "tci_solve_right_f64: pivot matrix is singular");
            std::memcpy(cores[b], X.data(), sizeof(double) * X.size());
        }
        if (mx != mx) a->nan = 1; else a->mx = mx;
    }
    return TCI_OK;
}

int tci_dsweep_sweep1site_f64(tci_handle h, tci_functor f, const int32_t* in, int64_t, double reltol, double abstol, void* acc,
                              double* const* cores, const int64_t* corecap, int32_t* out, int64_t out_cap, int64_t* used_words) {
    h->ncalls++;
    const int N = in[0], maxbond = in[4];
    const int32_t* ld = in + 8; const int32_t* desc = ld + N;
    int64_t used = 8 + 6LL * (N - 1);
    out[0] = 0; out[1] = N - 1; out[2] = (int)used;
    MockAcc* a = (MockAcc*)acc;
    std::vector<int32_t> mov(in + desc[1], in + desc[1] + 0);       // Iset[0]: rows of length 0
    int nmov = desc[0];
    static const int32_t dummy = 0;
    bool stopped = false;
    for (int b0 = 0; b0 < N - 1; ++b0) {
        const int d = ld[b0], nI = nmov, nJ = desc[12 * b0 + 2];
        const int32_t* tJ = in + desc[12 * b0 + 3];
        const int64_t m = (int64_t)nI * d, n = nJ, mn = m < n ? m : n;
        const int64_t rcap = (maxbond > 0 && maxbond < mn) ? maxbond : mn;
        if (m * n > g_dsweep_limit || m * rcap > corecap[b0] || used + rcap * (N + 1) + 2 * (rcap + 2) > out_cap) { out[0] = 1; out[1] = b0; stopped = true; break; }
        std::vector<double> Pi((size_t)(m * n));
        if (orc_batcheval((void*)f, ld, N, b0 > 0 ? mov.data() : &dummy, nI, b0, N - b0 - 1 > 0 ? tJ : &dummy, nJ, N - b0 - 1, 1, Pi.data()))
            return mfail(h, TCI_ERR_INVALID, orc_last_error());
        double mx = a->nan ? NAN : a->mx;
        merge_maxabs(&mx, Pi.data(), Pi.size());
        if (mx != mx) a->nan = 1; else a->mx = mx;
        std::vector<int64_t> rp((size_t)m), cp((size_t)n);
        std::vector<double> pe((size_t)rcap + 1), Lm((size_t)(m * rcap)), Um((size_t)(rcap * n));
        int64_t r = 0; double err = 0.0;
        const int st = orc_rrlu(Pi.data(), m, n, m, rcap, reltol, abstol, 1, rp.data(), cp.data(), &r, &err, Lm.data(), m, Um.data(), rcap, pe.data());
        if (st == -1) { out[0] = 2; out[1] = b0; stopped = true; break; }
        if (st == -2) { out[0] = 3; out[1] = b0; stopped = true; break; }
        if (st) return mfail(h, TCI_ERR_INVALID, orc_last_error());
        std::vector<double> left((size_t)(m * r)), right((size_t)(r * n));
        if (r > 0 && orc_luci_factors(m, n, r, 1, rp.data(), cp.data(), Lm.data(), m, Um.data(), rcap, left.data(), right.data()))
            return mfail(h, TCI_ERR_INVALID, orc_last_error());
        bool tnan = false;
        for (double v : left) if (v != v) tnan = true;
        std::memcpy(cores[b0], left.data(), sizeof(double) * left.size());
        if (tnan) { out[0] = 4; out[1] = b0; stopped = true; break; }
        const int lenA = b0 + 1, lenB = N - b0 - 1;
        const int64_t offA = used, offB = offA + r * lenA; int64_t offE = offB + r * lenB; offE += offE & 1;
        std::vector<int32_t> newI((size_t)(r * lenA));
        for (int64_t k = 0; k < r; ++k) {
            const int64_t i = rp[(size_t)k] - 1;
            for (int p = 0; p < b0; ++p) newI[(size_t)(k * lenA + p)] = mov[(size_t)((i % nI) * b0 + p)];
            newI[(size_t)(k * lenA + b0)] = (int32_t)(i / nI + 1);
            std::memcpy(out + offB + k * lenB, tJ + (cp[(size_t)k] - 1) * lenB, sizeof(int32_t) * lenB);
        }
        std::memcpy(out + offA, newI.data(), sizeof(int32_t) * newI.size());
        std::memcpy(out + offE, pe.data(), sizeof(double) * (size_t)(r + 1));
        int32_t* rec = out + 8 + 6 * b0;
        rec[0] = (int)r; rec[1] = (int)offA; rec[2] = (int)offB; rec[3] = (int)offE; rec[4] = (int)m; rec[5] = (int)n;
        used = offE + 2 * (r + 1);
        out[2] = (int)used;
        mov = newI; nmov = (int)r;
    }
    if (!stopped) {
        const int b0 = N - 1, d = ld[b0], nJ = desc[12 * b0 + 2];
        const int64_t q = (int64_t)nmov * d;
        if (q * nJ > corecap[b0]) { out[0] = 1; out[1] = N - 1; }
        else {
            std::vector<double> T((size_t)(q * nJ));
            if (orc_batcheval((void*)f, ld, N, mov.data(), nmov, b0, &dummy, nJ, 0, 1, T.data())) return mfail(h, TCI_ERR_INVALID, orc_last_error());
            double mx = a->nan ? NAN : a->mx;
            merge_maxabs(&mx, T.data(), T.size());
            if (mx != mx) a->nan = 1; else a->mx = mx;
            std::memcpy(cores[b0], T.data(), sizeof(double) * T.size());
        }
    }
    if (used_words) *used_words = used;
    return TCI_OK;
}

} // extern "C"
