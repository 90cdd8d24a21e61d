// oracle/tci_oracle.cpp
//
// TEST INFRASTRUCTURE ONLY.  A plain CPU restatement (C++17, no GPU, no BLAS) of the
// TensorCI2 two-site-sweep hot path of TensorCrossInterpolation.jl.  It is the parity
// checker for tests/, __graft_entry__.smoke() and the `cpu_baseline` / `--impl reference`
// legs of bench.py.  Nothing in the product path may import, link or call it.
//
// Parity status: pinned against the literal known-answer tests of the reference
// (test/test_matrixlu.jl, test/test_matrixluci.jl, test/test_batcheval.jl,
// test/test_tensorci2.jl, test/test_tensortrain.jl, test/test_contraction.jl; fixtures
// under tests/golden/).  The reference itself (Julia) cannot run in this image, so
// configs C1..C5 have no reference-generated outputs: for those the oracle IS the golden
// result ("parity unpinned against a live reference run").
//
// Build: oracle/Makefile  (g++ -O2 -ffp-contract=off: Julia never contracts a*b-c to FMA).
// Every function cites the reference file:line it restates (paths relative to the
// reference checkout, e.g. src/matrixlu.jl).
//
// Conventions: matrices are column-major like Julia; indices are 1-based in every
// quantity that is an index *value* (multi-indices, permutations), 0-based only for
// C++ container positions.

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <functional>
#include <limits>
#include <map>
#include <omp.h>
#include <stdexcept>
#include <string>
#include <vector>

namespace orc {

typedef std::vector<int32_t> MultiIndex;          // src/abstracttensortrain.jl:33
typedef std::vector<MultiIndex> IndexList;

struct Mat {                                       // column-major dense matrix
    int64_t m = 0, n = 0;
    std::vector<double> a;
    Mat() {}
    Mat(int64_t m_, int64_t n_, double v = 0.0) : m(m_), n(n_), a((size_t)m_ * n_, v) {}
    double& operator()(int64_t i, int64_t j) { return a[(size_t)j * m + i]; }
    double operator()(int64_t i, int64_t j) const { return a[(size_t)j * m + i]; }
};

static Mat matmul(const Mat& A, const Mat& B) {    // plain triple loop, column-major friendly
    if (A.n != B.m) throw std::runtime_error("matmul: shape mismatch");
    Mat C(A.m, B.n, 0.0);
    for (int64_t j = 0; j < B.n; ++j)
        for (int64_t k = 0; k < A.n; ++k) {
            const double b = B(k, j);
            const double* ac = &A.a[(size_t)k * A.m];
            double* cc = &C.a[(size_t)j * C.m];
            for (int64_t i = 0; i < A.m; ++i) cc[i] += ac[i] * b;
        }
    return C;
}

// ---------------------------------------------------------------------------------------
// submatrixargmax — src/matrixlu.jl:46-87 (general form) and :133-135 (start-index form).
// Columns outer, rows inner, replace only on strict '>', start from typemin => NaN never
// wins and ties resolve to the smallest column, then the smallest row.
// `mode`: 0 = identity (signed compare), 1 = abs2 (x*x).
// rows/cols are 0-based positions here; result is 0-based.
// ---------------------------------------------------------------------------------------
static void submatrixargmax(const double* A, int64_t lda, int mode,
                            const int64_t* rows, int64_t nrows,
                            const int64_t* cols, int64_t ncols,
                            int64_t* outr, int64_t* outc) {
    if (nrows <= 0) throw std::invalid_argument("rows must not be empty");
    if (ncols <= 0) throw std::invalid_argument("cols must not be empty");
    double best = -std::numeric_limits<double>::infinity();
    int64_t mr = rows[0], mc = cols[0];
    for (int64_t cj = 0; cj < ncols; ++cj) {
        const int64_t c = cols[cj];
        for (int64_t ri = 0; ri < nrows; ++ri) {
            const int64_t r = rows[ri];
            const double x = A[(size_t)c * lda + r];
            const double v = mode ? x * x : x;
            if (v > best) { best = v; mr = r; mc = c; }
        }
    }
    *outr = mr; *outc = mc;
}

// Fast contiguous form used by the LU loop: argmax of abs2 over A[k:m, k:n] (0-based k).
static void argmax_abs2_from(const Mat& A, int64_t k, int64_t* outr, int64_t* outc) {
    if (k >= A.m) throw std::invalid_argument("rows must not be empty");
    if (k >= A.n) throw std::invalid_argument("cols must not be empty");
    double best = -std::numeric_limits<double>::infinity();
    int64_t mr = k, mc = k;
    for (int64_t c = k; c < A.n; ++c) {
        const double* col = &A.a[(size_t)c * A.m];
        for (int64_t r = k; r < A.m; ++r) {
            const double v = col[r] * col[r];
            if (v > best) { best = v; mr = r; mc = c; }
        }
    }
    *outr = mr; *outc = mc;
}

// ---------------------------------------------------------------------------------------
// rrLU — src/matrixlu.jl:200-231 (type), :254-275 (swaps), :295-322 (addpivot!),
// :346-396 (_optimizerrlu!), :420-463 (rrlu!/rrlu).
// ---------------------------------------------------------------------------------------
struct RRLU {
    std::vector<int64_t> rowperm, colperm;         // 1-based, full length, pivots first
    Mat L, U;                                      // m x r, r x n
    bool leftorthogonal = true;
    int64_t npivot = 0;
    double error = std::numeric_limits<double>::quiet_NaN();
};

static void swaprow(RRLU& lu, Mat& A, int64_t a, int64_t b) {      // :254-262
    std::swap(lu.rowperm[a], lu.rowperm[b]);
    if (a == b) return;
    for (int64_t j = 0; j < A.n; ++j) std::swap(A(a, j), A(b, j));
}
static void swapcol(RRLU& lu, Mat& A, int64_t a, int64_t b) {      // :269-275
    std::swap(lu.colperm[a], lu.colperm[b]);
    if (a == b) return;
    for (int64_t i = 0; i < A.m; ++i) std::swap(A(i, a), A(i, b));
}

static void addpivot(RRLU& lu, Mat& A, int64_t pr, int64_t pc) {   // :295-322
    const int64_t k = lu.npivot;                   // 0-based position of the new pivot
    lu.npivot += 1;
    swaprow(lu, A, k, pr);
    swapcol(lu, A, k, pc);
    const double piv = A(k, k);
    if (lu.leftorthogonal) {
        for (int64_t i = k + 1; i < A.m; ++i) A(i, k) = A(i, k) / piv;     // true division
    } else {
        for (int64_t j = k + 1; j < A.n; ++j) A(k, j) = A(k, j) / piv;
    }
    const double* x = &A.a[(size_t)k * A.m];       // column k
    for (int64_t j = k + 1; j < A.n; ++j) {
        const double y = A(k, j);
        double* col = &A.a[(size_t)j * A.m];
        for (int64_t i = k + 1; i < A.m; ++i) col[i] = col[i] - x[i] * y;  // mul then sub, no FMA
    }
}

static inline double jl_max(double a, double b) {                  // Julia's max propagates NaN
    if (std::isnan(a) || std::isnan(b)) return std::numeric_limits<double>::quiet_NaN();
    return a > b ? a : b;
}

// returns 0 ok, -1 "lu.L contains NaNs", -2 "lu.U contains NaNs"  (:376-381)
static int optimizerrlu(RRLU& lu, Mat& A, int64_t maxrank, double reltol, double abstol) {
    maxrank = std::min(maxrank, std::min(A.m, A.n));               // :353
    double maxerror = 0.0;
    while (lu.npivot < maxrank) {                                  // :356
        const int64_t k = lu.npivot;
        int64_t pr, pc;
        argmax_abs2_from(A, k, &pr, &pc);                          // :359
        lu.error = std::fabs(A(pr, pc));                           // :360
        if ((std::fabs(lu.error) < reltol * maxerror || std::fabs(lu.error) < abstol) && lu.npivot > 0)
            break;                                                 // :363
        maxerror = jl_max(maxerror, lu.error);                     // :367
        addpivot(lu, A, pr, pc);
    }
    const int64_t r = lu.npivot;
    lu.L = Mat(A.m, r, 0.0);                                       // tril(A[:, 1:r])   :372
    for (int64_t j = 0; j < r; ++j)
        for (int64_t i = j; i < A.m; ++i) lu.L(i, j) = A(i, j);
    lu.U = Mat(r, A.n, 0.0);                                       // triu(A[1:r, :])   :373
    for (int64_t j = 0; j < A.n; ++j)
        for (int64_t i = 0; i <= std::min(j, r - 1); ++i) lu.U(i, j) = A(i, j);
    for (double v : lu.L.a) if (std::isnan(v)) return -1;
    for (double v : lu.U.a) if (std::isnan(v)) return -2;
    if (lu.leftorthogonal) { for (int64_t i = 0; i < r; ++i) lu.L(i, i) = 1.0; }   // :384-388
    else                   { for (int64_t i = 0; i < r; ++i) lu.U(i, i) = 1.0; }
    if (lu.npivot >= std::min(A.m, A.n)) lu.error = 0.0;           // :391-393
    return 0;
}

// Thread-parallel execution of the SAME loop (matrixlu.jl:346-369) for big matrices (the 8192 x 8192 C2 workload takes
// ~70 s on one thread).  The reference's loops are scalar Julia; nothing here changes a value or a decision: every element
// of the trailing block gets the identical `a - x*y` (one multiply, one subtract), the search keeps per-thread winners over
// contiguous column ranges with the same strict '>' scan and merges them in column order with strict '>', which is the
// sequential first-maximum rule.  tests/test_oracle_golden.py checks it against the single-thread loop bit for bit.
static int optimizerrlu_mt(RRLU& lu, Mat& A, int64_t maxrank, double reltol, double abstol, int nthreads) {
    maxrank = std::min(maxrank, std::min(A.m, A.n));
    double maxerror = 0.0;
    const int64_t m = A.m, n = A.n;
    struct alignas(64) Best { double v; int64_t r, c; };
    std::vector<Best> best((size_t)nthreads);
    while (lu.npivot < maxrank) {
        const int64_t k = lu.npivot;
#pragma omp parallel num_threads(nthreads)
        {
            const int t = omp_get_thread_num(), T = omp_get_num_threads();
            const int64_t c0 = k + (n - k) * t / T, c1 = k + (n - k) * (t + 1) / T;
            Best b{-std::numeric_limits<double>::infinity(), -1, -1};
            for (int64_t c = c0; c < c1; ++c) {
                const double* col = &A.a[(size_t)c * m];
                for (int64_t r = k; r < m; ++r) { const double v = col[r] * col[r]; if (v > b.v) { b.v = v; b.r = r; b.c = c; } }
            }
            best[(size_t)t] = b;
            if (t == 0) for (int u = T; u < nthreads; ++u) best[(size_t)u] = Best{-std::numeric_limits<double>::infinity(), -1, -1};
        }
        double bv = -std::numeric_limits<double>::infinity(); int64_t pr = k, pc = k;
        for (int t = 0; t < nthreads; ++t) if (best[(size_t)t].r >= 0 && best[(size_t)t].v > bv) { bv = best[(size_t)t].v; pr = best[(size_t)t].r; pc = best[(size_t)t].c; }
        lu.error = std::fabs(A(pr, pc));
        if ((std::fabs(lu.error) < reltol * maxerror || std::fabs(lu.error) < abstol) && lu.npivot > 0) break;
        maxerror = jl_max(maxerror, lu.error);
        lu.npivot += 1;
        swaprow(lu, A, k, pr);
        swapcol(lu, A, k, pc);
        const double piv = A(k, k);
        if (lu.leftorthogonal) { for (int64_t i = k + 1; i < m; ++i) A(i, k) = A(i, k) / piv; }
        else                   { for (int64_t j = k + 1; j < n; ++j) A(k, j) = A(k, j) / piv; }
        const double* x = &A.a[(size_t)k * m];
#pragma omp parallel for schedule(static) num_threads(nthreads)
        for (int64_t j = k + 1; j < n; ++j) {
            const double y = A(k, j);
            double* col = &A.a[(size_t)j * m];
            for (int64_t i = k + 1; i < m; ++i) col[i] = col[i] - x[i] * y;
        }
    }
    // the extraction below is the tail of optimizerrlu (maxrank already reached or the stop rule fired): run it with no further pivots
    return optimizerrlu(lu, A, lu.npivot, reltol, abstol) ;
}

static int rrlu_inplace(Mat& A, RRLU& lu, int64_t maxrank, double reltol, double abstol, bool leftorth) {
    lu = RRLU();                                                   // :228-230
    lu.leftorthogonal = leftorth;
    lu.rowperm.resize(A.m); lu.colperm.resize(A.n);
    for (int64_t i = 0; i < A.m; ++i) lu.rowperm[i] = i + 1;
    for (int64_t j = 0; j < A.n; ++j) lu.colperm[j] = j + 1;
    return optimizerrlu(lu, A, maxrank, reltol, abstol);
}

// ---------------------------------------------------------------------------------------
// arrlu (rook pivot search) — src/matrixlu.jl:500-569, helpers :627-674, util.jl:173-219.
// The reference draws from Julia's global RNG (`rand(1:length(c))`, util.jl:186), whose stream
// cannot be restated; oracle and product share the counter-based generator below instead, so the
// random columns differ from a Julia run while every other step follows the reference line by line.
// ---------------------------------------------------------------------------------------
struct Rng {
    uint64_t s = 0;
    uint64_t next() {                              // splitmix64 stream
        s += 0x9e3779b97f4a7c15ULL;
        uint64_t z = s; z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL; z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
        return z ^ (z >> 31);
    }
    int64_t randint(int64_t n) { return (int64_t)(next() % (uint64_t)n); }      // 0-based stand-in for rand(1:n)
};
typedef std::function<Mat(const std::vector<int64_t>&, const std::vector<int64_t>&)> BatchF;   // 1-based rows, cols -> matrix

static void pushrandomsubset(std::vector<int64_t>& subset, int64_t setsize, int64_t n, Rng& rng) {   // util.jl:173-219
    std::vector<char> in(setsize + 1, 0);
    for (int64_t v : subset) in[v] = 1;
    std::vector<int64_t> c;                                        // setdiff(1:setsize, subset), ascending
    for (int64_t v = 1; v <= setsize; ++v) if (!in[v]) c.push_back(v);
    n = std::min<int64_t>(n, (int64_t)c.size());
    for (int64_t i = 0; i < n; ++i) {
        const int64_t idx = rng.randint((int64_t)c.size());
        subset.push_back(c[idx]);
        c.erase(c.begin() + idx);                                  // deleteat!
    }
}
static void cols2Lmatrix(Mat& C, const Mat& P) {                   // :627-648  (C <- C * inv(triu P), column by column)
    if (C.n != P.n) throw std::runtime_error("C and P matrices must have same number of columns in `cols2Lmatrix!`.");
    if (P.m != P.n) throw std::runtime_error("P matrix must be square in `cols2Lmatrix!`.");
    for (int64_t k = 0; k < P.m; ++k) {
        for (int64_t i = 0; i < C.m; ++i) C(i, k) = C(i, k) / P(k, k);
        for (int64_t j = k + 1; j < C.n; ++j) {
            const double y = P(k, j);
            for (int64_t i = 0; i < C.m; ++i) C(i, j) = C(i, j) - C(i, k) * y;
        }
    }
}
static void rows2Umatrix(Mat& R, const Mat& P) {                   // :655-674  (R <- inv(tril P) * R, row by row)
    if (R.m != P.m) throw std::runtime_error("R and P matrices must have same number of rows in `rows2Umatrix!`.");
    if (P.m != P.n) throw std::runtime_error("P matrix must be square in `rows2Umatrix!`.");
    for (int64_t k = 0; k < P.m; ++k) {
        for (int64_t j = 0; j < R.n; ++j) R(k, j) = R(k, j) / P(k, k);
        for (int64_t j = 0; j < R.n; ++j) {
            const double y = R(k, j);
            for (int64_t i = k + 1; i < R.m; ++i) R(i, j) = R(i, j) - P(i, k) * y;
        }
    }
}
static Mat topleft(const Mat& A, int64_t r) { Mat B(r, r); for (int64_t j = 0; j < r; ++j) for (int64_t i = 0; i < r; ++i) B(i, j) = A(i, j); return B; }

static int arrlu(const BatchF& f, int64_t m, int64_t n, std::vector<int64_t> I0, std::vector<int64_t> J0, RRLU& lu,
                 int64_t maxrank, double reltol, double abstol, bool leftorth, int numrookiter, Rng& rng) {
    lu = RRLU(); lu.leftorthogonal = leftorth;                     // :513
    lu.rowperm.resize(m); lu.colperm.resize(n);
    for (int64_t i = 0; i < m; ++i) lu.rowperm[i] = i + 1;
    for (int64_t j = 0; j < n; ++j) lu.colperm[j] = j + 1;
    lu.L = Mat(0, 0); lu.U = Mat(0, 0);
    bool islowrank = false;
    maxrank = std::min(maxrank, std::min(m, n));                   // :515
    while (true) {
        if (leftorth) pushrandomsubset(J0, n, std::max<int64_t>(1, (int64_t)J0.size()), rng);     // :521-525
        else          pushrandomsubset(I0, m, std::max<int64_t>(1, (int64_t)I0.size()), rng);
        for (int rookiter = 1; rookiter <= numrookiter; ++rookiter) {
            const bool colmove = ((rookiter % 2 == 0) == leftorth);                                // :529
            Mat sub = colmove ? f(I0, lu.colperm) : f(lu.rowperm, J0);                             // :532-536
            lu.npivot = 0;                                                                         // :539
            // _optimizerrlu! swaps lu.rowpermutation / lu.colpermutation at the POSITIONS of the submatrix (:254-275)
            int st = optimizerrlu(lu, sub, maxrank, reltol, abstol);
            if (st) return st;
            islowrank = islowrank || (lu.npivot < std::min(sub.m, sub.n));                         // :541
            std::vector<int64_t> ri(lu.rowperm.begin(), lu.rowperm.begin() + lu.npivot);
            std::vector<int64_t> ci(lu.colperm.begin(), lu.colperm.begin() + lu.npivot);
            if (ri == I0 && ci == J0) break;                                                       // :544-546
            J0 = ci; I0 = ri;                                                                      // :548-549
        }
        if (islowrank || (int64_t)I0.size() >= maxrank) break;                                     // :552-554
    }
    const int64_t r = lu.npivot;
    if (lu.L.m < m) {                                                                              // :558-564
        std::vector<char> in(m + 1, 0); for (int64_t v : I0) in[v] = 1;
        std::vector<int64_t> I2; for (int64_t v = 1; v <= m; ++v) if (!in[v]) I2.push_back(v);
        lu.rowperm = I0; lu.rowperm.insert(lu.rowperm.end(), I2.begin(), I2.end());
        Mat L2 = f(I2, J0);
        cols2Lmatrix(L2, topleft(lu.U, r));
        Mat L(m, r);
        for (int64_t j = 0; j < r; ++j) { for (int64_t i = 0; i < r; ++i) L(i, j) = lu.L(i, j);
                                          for (int64_t i = 0; i < L2.m; ++i) L(r + i, j) = L2(i, j); }
        lu.L = L;
    }
    if (lu.U.n < n) {                                                                              // :567-573
        std::vector<char> in(n + 1, 0); for (int64_t v : J0) in[v] = 1;
        std::vector<int64_t> J2; for (int64_t v = 1; v <= n; ++v) if (!in[v]) J2.push_back(v);
        lu.colperm = J0; lu.colperm.insert(lu.colperm.end(), J2.begin(), J2.end());
        Mat U2 = f(I0, J2);
        rows2Umatrix(U2, topleft(lu.L, r));
        Mat U(r, n);
        for (int64_t j = 0; j < r; ++j) for (int64_t i = 0; i < r; ++i) U(i, j) = lu.U(i, j);
        for (int64_t j = 0; j < U2.n; ++j) for (int64_t i = 0; i < r; ++i) U(i, r + j) = U2(i, j);
        lu.U = U;
    }
    return 0;
}

static Mat lu_left(const RRLU& lu, bool permute) {                 // :716-724
    if (!permute) return lu.L;
    Mat l(lu.L.m, lu.L.n);
    for (int64_t j = 0; j < lu.L.n; ++j)
        for (int64_t i = 0; i < lu.L.m; ++i) l(lu.rowperm[i] - 1, j) = lu.L(i, j);
    return l;
}
static Mat lu_right(const RRLU& lu, bool permute) {                // :738-746
    if (!permute) return lu.U;
    Mat u(lu.U.m, lu.U.n);
    for (int64_t j = 0; j < lu.U.n; ++j)
        for (int64_t i = 0; i < lu.U.m; ++i) u(i, lu.colperm[j] - 1) = lu.U(i, j);
    return u;
}
static std::vector<double> lu_pivoterrors(const RRLU& lu) {        // :756-762, :799-801
    std::vector<double> e;
    for (int64_t i = 0; i < lu.npivot; ++i)
        e.push_back(std::fabs(lu.leftorthogonal ? lu.U(i, i) : lu.L(i, i)));
    e.push_back(lu.error);
    return e;
}

// ---------------------------------------------------------------------------------------
// MatrixLUCI factors — src/matrixluci.jl:161-283.
// ---------------------------------------------------------------------------------------
static Mat luci_colmatrix(const RRLU& lu) {                        // :161-165  left(lu) * U[:, 1:r]
    const int64_t r = lu.npivot;
    Mat U1(r, r);
    for (int64_t j = 0; j < r; ++j) for (int64_t i = 0; i < r; ++i) U1(i, j) = lu.U(i, j);
    return matmul(lu_left(lu, true), U1);
}
static Mat luci_rowmatrix(const RRLU& lu) {                        // :175-177  L[1:r, :] * right(lu)
    const int64_t r = lu.npivot;
    Mat L1(r, r);
    for (int64_t j = 0; j < r; ++j) for (int64_t i = 0; i < r; ++i) L1(i, j) = lu.L(i, j);
    return matmul(L1, lu_right(lu, true));
}
static Mat luci_colstimespivotinv(const RRLU& lu) {                // :194-213
    const int64_t r = lu.npivot, m = lu.L.m;
    Mat res(m, r, 0.0);
    for (int64_t i = 0; i < std::min(m, r); ++i) res(i, i) = 1.0;
    if (r < m) {
        // X * L1 = L2  (right division by a lower-triangular matrix), back substitution over columns
        for (int64_t i = r; i < m; ++i) {
            for (int64_t j = r - 1; j >= 0; --j) {
                double s = lu.L(i, j);
                for (int64_t k = j + 1; k < r; ++k) s -= res(i, k) * lu.L(k, j);
                res(i, j) = s / lu.L(j, j);
            }
        }
    }
    Mat out(m, r);
    for (int64_t j = 0; j < r; ++j)
        for (int64_t i = 0; i < m; ++i) out(lu.rowperm[i] - 1, j) = res(i, j);       // :211
    return out;
}
static Mat luci_pivotinvtimesrows(const RRLU& lu) {                // :227-241
    const int64_t r = lu.npivot, n = lu.U.n;
    Mat res(r, n, 0.0);
    for (int64_t i = 0; i < std::min(n, r); ++i) res(i, i) = 1.0;
    if (r < n) {
        for (int64_t j = r; j < n; ++j) {                          // U1 \ U2, back substitution
            for (int64_t i = r - 1; i >= 0; --i) {
                double s = lu.U(i, j);
                for (int64_t k = i + 1; k < r; ++k) s -= lu.U(i, k) * res(k, j);
                res(i, j) = s / lu.U(i, i);
            }
        }
    }
    Mat out(r, n);
    for (int64_t j = 0; j < n; ++j)
        for (int64_t i = 0; i < r; ++i) out(i, lu.colperm[j] - 1) = res(i, j);       // :239
    return out;
}
static Mat luci_left(const RRLU& lu) {                             // :256-262
    return lu.leftorthogonal ? luci_colstimespivotinv(lu) : luci_colmatrix(lu);
}
static Mat luci_right(const RRLU& lu) {                            // :277-283
    return lu.leftorthogonal ? luci_rowmatrix(lu) : luci_pivotinvtimesrows(lu);
}

// ---------------------------------------------------------------------------------------
// Target functions.  Built-in kinds restate the functions the reference's README, tests and
// BASELINE configs use; `cb` is an arbitrary host callback (tests only).
// ---------------------------------------------------------------------------------------
enum FuncKind {
    F_CALLBACK = 0,
    F_LORENTZ = 1,        // p0 / (1 + sum v_i^2)                 README.md:21-29, test_tensorci2.jl:249
    F_QEXPSIN = 2,        // exp(-p0 x) sin(2 pi p1 x^p2)         test_tensorci2.jl:437
    F_QEXPPLUSSIN = 3,    // exp(-p0 x) + p1 sin(p2 x)            test_globalsearch.jl:14
    F_SUM = 4,            // p0 * sum v_i                         test_batcheval.jl:18
    F_QEXP2 = 5,          // exp(-p0 x) + p1 exp(-p2 x)           test_tensorci2.jl:63,:146
    F_DENSE = 6,          // column-major table lookup T[v...]    test_tensorci2.jl:29,:464
    F_HASH = 7,           // p0 * u(v) + p1 / (1 + sum v_i^2), u = integer hash of v in [-1, 1): a full-rank synthetic
                          // function (not in the reference) that drives every bond to maxbonddim (bench only)
    F_BATCHCB = 8         // a BatchEvaluator (cachedtensortrain.jl:16-31, batcheval.jl:196-214): the host supplies whole
                          // (left, right, M) batches.  Tests use it to feed the oracle the function VALUES the device
                          // functor produced (hard part 1: libm and libdevice differ in the last ulp), so that
                          // index sets can be compared bit for bit and the values separately.
};
static inline uint64_t splitmix64(uint64_t z) {
    z += 0x9e3779b97f4a7c15ULL; z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL; z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
    return z ^ (z >> 31);
}
typedef double (*orc_callback)(const int32_t* idx, int32_t n, void* user);
// out[i + c*nleft + j*nleft*ncent] = f([left_i..., c..., right_j...]); tables are row-major (count x len), 1-based values
typedef void (*orc_batch_callback)(const int32_t* left, int64_t nleft, int32_t nl, const int32_t* right, int64_t nright, int32_t nr,
                                   int32_t M, double* out, void* user);

struct Func {
    int kind = F_LORENTZ;
    std::vector<double> p;
    std::vector<int32_t> dims;                     // F_DENSE: table dims
    std::vector<double> table;
    orc_callback cb = nullptr; void* user = nullptr;
    orc_batch_callback bcb = nullptr;

    // quantics coordinate x = sum_i (b_i - 1) 2^{-i}, i = 1..R  (bit 1 most significant;
    // the in-tree statement is test/test_tensorci2.jl:66, x = (i-1)/2^R)
    static double qx(const int32_t* v, int n) {
        double x = 0.0, w = 0.5;
        for (int i = 0; i < n; ++i) { x += (double)(v[i] - 1) * w; w *= 0.5; }
        return x;
    }
    double operator()(const int32_t* v, int n) const {
        switch (kind) {
        case F_CALLBACK: return cb(v, n, user);
        case F_BATCHCB: { double out = 0.0; bcb(v, 1, n, nullptr, 1, 0, 0, &out, user); return out; }   // point form of a BatchEvaluator
        case F_LORENTZ: { int64_t s = 0; for (int i = 0; i < n; ++i) s += (int64_t)v[i] * v[i];
                          return p[0] / (double)(1 + s); }
        case F_QEXPSIN: { double x = qx(v, n);
                          return std::exp(-p[0] * x) * std::sin(2.0 * M_PI * p[1] * std::pow(x, p[2])); }
        case F_QEXPPLUSSIN: { double x = qx(v, n); return std::exp(-p[0] * x) + p[1] * std::sin(p[2] * x); }
        case F_SUM: { int64_t s = 0; for (int i = 0; i < n; ++i) s += v[i]; return p[0] * (double)s; }
        case F_QEXP2: { double x = qx(v, n); return std::exp(-p[0] * x) + p[1] * std::exp(-p[2] * x); }
        case F_HASH: { uint64_t hsum = 0; int64_t s2 = 0;
                       for (int i = 0; i < n; ++i) { hsum += (uint64_t)v[i] * (splitmix64((uint64_t)i + 1) | 1ULL); s2 += (int64_t)v[i] * v[i]; }
                       const double u = (double)(splitmix64(hsum) >> 11) * 0x1.0p-52 - 1.0;
                       return p[0] * u + p[1] / (double)(1 + s2); }
        case F_DENSE: { int64_t off = 0, st = 1;
                        for (int i = 0; i < n; ++i) { off += (int64_t)(v[i] - 1) * st; st *= dims[i]; }
                        return table[(size_t)off]; }
        }
        throw std::runtime_error("unknown function kind");
    }
    double operator()(const MultiIndex& v) const { return (*this)(v.data(), (int)v.size()); }
};

// ---------------------------------------------------------------------------------------
// _batchevaluate_dispatch — src/batcheval.jl:131-175.  result[i, c, j], i fastest in memory
// (Julia column-major), centre multi-index enumerated first-index-fastest (:160).
// ---------------------------------------------------------------------------------------
static std::vector<double> batchevaluate(const Func& f, const std::vector<int32_t>& localdims,
                                         const IndexList& left, const IndexList& right, int M,
                                         std::vector<int64_t>* shape) {
    shape->clear();
    if (left.empty() || right.empty()) { shape->assign(M + 2, 0); return {}; }       // :140-142
    const int nl = (int)left[0].size(), nr = (int)right[0].size();
    const int L = M + nl + nr;
    int64_t ncent = 1;
    for (int x = 0; x < M; ++x) ncent *= localdims[nl + x];
    const int64_t nI = (int64_t)left.size(), nJ = (int64_t)right.size();
    std::vector<double> res((size_t)(nI * ncent * nJ));
    shape->push_back(nI);
    for (int x = 0; x < M; ++x) shape->push_back(localdims[nl + x]);
    shape->push_back(nJ);
    if (f.kind == F_BATCHCB) {                                                       // batcheval.jl:196-214: the evaluator gets the batch
        std::vector<int32_t> lt((size_t)nI * nl), rt((size_t)nJ * nr);
        for (int64_t i = 0; i < nI; ++i) std::copy(left[i].begin(), left[i].end(), lt.begin() + (size_t)i * nl);
        for (int64_t j = 0; j < nJ; ++j) std::copy(right[j].begin(), right[j].end(), rt.begin() + (size_t)j * nr);
        f.bcb(lt.data(), nI, nl, rt.data(), nJ, nr, M, res.data(), f.user);
        return res;
    }
#pragma omp parallel
    {
        std::vector<int32_t> idx(L);
#pragma omp for collapse(2) schedule(static)
        for (int64_t j = 0; j < nJ; ++j)
            for (int64_t c = 0; c < ncent; ++c) {
                int64_t rem = c;
                for (int x = 0; x < M; ++x) { idx[nl + x] = (int32_t)(rem % localdims[nl + x]) + 1; rem /= localdims[nl + x]; }
                for (int x = 0; x < nr; ++x) idx[nl + M + x] = right[j][x];
                for (int64_t i = 0; i < nI; ++i) {
                    for (int x = 0; x < nl; ++x) idx[x] = left[i][x];
                    res[(size_t)((j * ncent + c) * nI + i)] = f(idx.data(), L);
                }
            }
    }
    return res;
}

// ---------------------------------------------------------------------------------------
// TensorCI2 — src/tensorci2.jl.
// ---------------------------------------------------------------------------------------
struct Tensor3 { int64_t d0 = 0, d1 = 0, d2 = 0; std::vector<double> a; };   // column-major (d0,d1,d2)

static bool contains(const IndexList& c, const MultiIndex& x) {
    for (const auto& y : c) if (y == x) return true;
    return false;
}
static void pushunique(IndexList& c, const MultiIndex& x) { if (!contains(c, x)) c.push_back(x); }   // util.jl:94-98

static IndexList kronecker_left(const IndexList& I, int d) {       // :512-517  [is..., j], is fastest
    IndexList out; out.reserve(I.size() * d);
    for (int j = 1; j <= d; ++j)
        for (const auto& is : I) { MultiIndex x(is); x.push_back(j); out.push_back(x); }
    return out;
}
static IndexList kronecker_right(int d, const IndexList& J) {      // :524-529  [i, js...], i fastest
    IndexList out; out.reserve(J.size() * d);
    for (const auto& js : J)
        for (int i = 1; i <= d; ++i) { MultiIndex x; x.reserve(js.size() + 1); x.push_back(i); x.insert(x.end(), js.begin(), js.end()); out.push_back(x); }
    return out;
}
static IndexList set_union(const IndexList& a, const IndexList& b) {   // Base.union on vectors: first-seen order
    std::map<MultiIndex, int> seen;
    IndexList out;
    for (const auto& x : a) if (seen.emplace(x, 1).second) out.push_back(x);
    for (const auto& x : b) if (seen.emplace(x, 1).second) out.push_back(x);
    return out;
}

struct TCI2 {
    std::vector<IndexList> Iset, Jset;             // :51-52
    std::vector<int32_t> localdims;
    std::vector<Tensor3> sitetensors;
    std::vector<double> pivoterrors, bonderrors;
    double maxsamplevalue = 0.0;
    std::vector<std::vector<IndexList>> Iset_history, Jset_history;
    int bondorder = 0;                             // 0: sequential bonds (the reference, :1226-1250); 1: odd bonds then even bonds
                                                   // (the parallel-bond schedule of the product; NOT in the reference)
    int pivotsearch = 0;                           // 0: :full, 1: :rook (:863-905)
    Rng rng;                                       // stream behind pushrandomsubset! in the rook search
    Rng grng;                                      // stream behind the random starts of the global pivot finder
    int64_t nevals = 0;                            // bookkeeping for the bench only
    double t_eval = 0, t_lu = 0;                   // seconds spent in Pi fill and in MatrixLUCI

    int n() const { return (int)localdims.size(); }

    explicit TCI2(const std::vector<int32_t>& ld) : localdims(ld) {            // :76-92
        if (ld.size() <= 1) throw std::runtime_error("localdims should have at least 2 elements!");
        Iset.resize(ld.size()); Jset.resize(ld.size()); sitetensors.resize(ld.size());
        bonderrors.assign(ld.size() - 1, 0.0);
    }
    void invalidatesitetensors() { for (auto& t : sitetensors) t = Tensor3(); }   // :206-211
    bool sitetensorsavailable() const { for (auto& t : sitetensors) if (t.a.empty()) return false; return true; }

    void addglobalpivots(const IndexList& pivots) {                // :335-357
        for (const auto& p : pivots)
            if ((int)p.size() != n()) throw std::runtime_error("Please specify a pivot as one index per leg of the MPS.");
        for (const auto& p : pivots)
            for (int b = 0; b < n(); ++b) {
                pushunique(Iset[b], MultiIndex(p.begin(), p.begin() + b));
                pushunique(Jset[b], MultiIndex(p.begin() + b + 1, p.end()));
            }
        if (!pivots.empty()) invalidatesitetensors();
    }
    void init(const Func& f, const IndexList& pivots) {            // :105-116
        addglobalpivots(pivots);
        double mx = -std::numeric_limits<double>::infinity();
        for (const auto& p : pivots) mx = jl_max(mx, std::fabs(f(p)));
        maxsamplevalue = mx;
        if (!(std::fabs(maxsamplevalue) > 0.0)) throw std::runtime_error("maxsamplevalue is zero!");
        invalidatesitetensors();
    }
    std::vector<int64_t> linkdims() const {                        // :189-191
        std::vector<int64_t> l; for (int b = 0; b + 1 < n(); ++b) l.push_back((int64_t)Iset[b + 1].size()); return l;
    }
    int64_t rank() const { int64_t r = 0; for (auto x : linkdims()) r = std::max(r, x); return r; }

    void updateerrors(int b, const std::vector<double>& e) {       // :252-260, :281-289
        bonderrors[b] = e.back();
        std::vector<double> out(std::max(pivoterrors.size(), e.size()), 0.0);
        for (size_t i = 0; i < out.size(); ++i) {
            const double x = i < pivoterrors.size() ? pivoterrors[i] : 0.0;
            const double y = i < e.size() ? e[i] : 0.0;
            out[i] = jl_max(x, y);
        }
        pivoterrors = out;
    }
    double pivoterror() const {                                    // :243-245, :277-279
        double m = -std::numeric_limits<double>::infinity();
        for (double x : bonderrors) m = jl_max(m, x);
        return m;
    }
    void updatemaxsample(const std::vector<double>& s) {           // :636-638, util.jl:34-43
        double m = maxsamplevalue;
        for (double x : s) m = jl_max(std::fabs(m), std::fabs(x));
        maxsamplevalue = m;
    }
    Mat fillmatrix(const Func& f, const IndexList& I, const IndexList& J, int M, int64_t rows, int64_t cols);
    void setsitetensor(int b, const Mat& T) {                      // :536-545
        Tensor3 t; t.d0 = (int64_t)Iset[b].size(); t.d1 = localdims[b]; t.d2 = (int64_t)Jset[b].size();
        if (t.d0 * t.d1 * t.d2 != T.m * T.n) throw std::runtime_error("setsitetensor!: size mismatch");
        t.a = T.a; sitetensors[b] = t;
    }
    int updatepivots(int b, const Func& f, bool leftorth, double reltol, double abstol, int64_t maxbonddim,
                     const IndexList& extraI, const IndexList& extraJ);
    void setsitetensor_f(const Func& f, int b);
    void fillsitetensors(const Func& f) { for (int b = 0; b < n(); ++b) setsitetensor_f(f, b); }   // globalsearch.jl:202-208
    int sweep1site(const Func& f, bool forward, double reltol, double abstol, int64_t maxbonddim, bool updatetensors);
    int makecanonical(const Func& f, double reltol, double abstol, int64_t maxbonddim);
    int sweep0site(const Func& f, int b, double reltol, double abstol);
    int sweep2site(const Func& f, int niter, int iter1, double abstol, int64_t maxbonddim, int sweepstrategy,
                   bool strictlynested, bool fillsite);
    double evaluate(const MultiIndex& idx) const;
};

static double now_s();

Mat TCI2::fillmatrix(const Func& f, const IndexList& I, const IndexList& J, int M, int64_t rows, int64_t cols) {
    // filltensor :475-497 then reshape to (rows, cols) (column-major reshape is a no-op on the data)
    const double t0 = now_s();
    std::vector<int64_t> shape;
    Mat P(rows, cols);
    P.a = batchevaluate(f, localdims, I, J, M, &shape);
    if ((int64_t)P.a.size() != rows * cols) throw std::runtime_error("filltensor: unexpected size");
    nevals += rows * cols;
    t_eval += now_s() - t0;
    return P;
}

// Error codes shared by the sweep functions: 0 ok, -1/-2 NaNs in L/U (matrixlu.jl:376-381),
// -3 NaN in a site tensor (tensorci2.jl:705-707), -4 pivot matrix not square (:623).

int TCI2::updatepivots(int b, const Func& f, bool leftorth, double reltol, double abstol, int64_t maxbonddim,
                       const IndexList& extraI, const IndexList& extraJ) {       // :825-930, :full branch
    invalidatesitetensors();
    IndexList Ic = set_union(kronecker_left(Iset[b], localdims[b]), extraI);      // :842
    IndexList Jc = set_union(kronecker_right(localdims[b + 1], Jset[b + 1]), extraJ);   // :843
    RRLU lu;
    double t0 = now_s();
    bool full = pivotsearch == 0;
    if (pivotsearch == 1) {                                                        // :863-905
        std::vector<int64_t> I0, J0;                                               // :866-867 positions of the current pivots
        for (const auto& x : Iset[b + 1]) { auto it = std::find(Ic.begin(), Ic.end(), x); if (it != Ic.end()) I0.push_back(it - Ic.begin() + 1); }
        for (const auto& x : Jset[b])     { auto it = std::find(Jc.begin(), Jc.end(), x); if (it != Jc.end()) J0.push_back(it - Jc.begin() + 1); }
        double submax = 0.0;                                                       // SubMatrix.maxsamplevalue :769-790
        BatchF pif = [&](const std::vector<int64_t>& ir, const std::vector<int64_t>& ic) {
            IndexList rows, cols;
            for (int64_t i : ir) rows.push_back(Ic[i - 1]);
            for (int64_t j : ic) cols.push_back(Jc[j - 1]);
            Mat S = fillmatrix(f, rows, cols, 0, (int64_t)rows.size(), (int64_t)cols.size());
            for (double v : S.a) submax = jl_max(submax, std::fabs(v));
            return S;
        };
        t0 = now_s();
        int st = arrlu(pif, (int64_t)Ic.size(), (int64_t)Jc.size(), I0, J0, lu, maxbonddim, reltol, abstol, leftorth, 5, rng);
        if (st) return st;
        updatemaxsample({submax});                                                 // :881
        full = lu.npivot == 0;                                                     // :886 fall back to the full search
    }
    if (full) {
        Mat Pi = fillmatrix(f, Ic, Jc, 0, (int64_t)Ic.size(), (int64_t)Jc.size());    // :848-852
        updatemaxsample(Pi.a);                                                         // :855
        t0 = now_s();
        int st = rrlu_inplace(Pi, lu, maxbonddim, reltol, abstol, leftorth);           // :856-862 (Pi is a fresh array: no copy needed)
        if (st) return st;
    }
    IndexList newI, newJ;
    for (int64_t k = 0; k < lu.npivot; ++k) newI.push_back(Ic[lu.rowperm[k] - 1]); // :919
    for (int64_t k = 0; k < lu.npivot; ++k) newJ.push_back(Jc[lu.colperm[k] - 1]); // :920
    Iset[b + 1] = newI; Jset[b] = newJ;
    if (extraI.empty() && extraJ.empty()) {                                        // :923-926
        setsitetensor(b, luci_left(lu));
        setsitetensor(b + 1, luci_right(lu));
    }
    t_lu += now_s() - t0;
    updateerrors(b, lu_pivoterrors(lu));                                           // :928
    return 0;
}

// Dense solve X * P = B (i.e. transpose(transpose(P) \ transpose(B)), tensorci2.jl:626) by LU with
// partial pivoting of P^T — the algorithm LAPACK getrf/getrs applies for Julia's `\` on a square matrix.
static int solve_right(const Mat& P, const Mat& B, Mat& X) {
    const int64_t r = P.m;
    if (P.n != r || B.n != r) return -4;
    Mat A(r, r);                                   // A = P^T
    for (int64_t j = 0; j < r; ++j) for (int64_t i = 0; i < r; ++i) A(i, j) = P(j, i);
    std::vector<int64_t> piv(r);
    for (int64_t k = 0; k < r; ++k) {
        int64_t p = k; double best = std::fabs(A(k, k));
        for (int64_t i = k + 1; i < r; ++i) if (std::fabs(A(i, k)) > best) { best = std::fabs(A(i, k)); p = i; }
        piv[k] = p;
        if (p != k) for (int64_t j = 0; j < r; ++j) std::swap(A(k, j), A(p, j));
        const double d = A(k, k);
        for (int64_t i = k + 1; i < r; ++i) A(i, k) /= d;
        for (int64_t j = k + 1; j < r; ++j) {
            const double y = A(k, j);
            for (int64_t i = k + 1; i < r; ++i) A(i, j) -= A(i, k) * y;
        }
    }
    // solve A * Y = B^T column by column; Y^T = X (rows of B are the right-hand sides)
    X = Mat(B.m, r);
    std::vector<double> y(r);
    for (int64_t q = 0; q < B.m; ++q) {
        for (int64_t i = 0; i < r; ++i) y[i] = B(q, i);
        for (int64_t k = 0; k < r; ++k) std::swap(y[k], y[piv[k]]);
        for (int64_t k = 0; k < r; ++k) for (int64_t i = k + 1; i < r; ++i) y[i] -= A(i, k) * y[k];
        for (int64_t k = r - 1; k >= 0; --k) { y[k] /= A(k, k); for (int64_t i = 0; i < k; ++i) y[i] -= A(i, k) * y[k]; }
        for (int64_t i = 0; i < r; ++i) X(q, i) = y[i];
    }
    return 0;
}

void TCI2::setsitetensor_f(const Func& f, int b) {                  // :599-629 (leftorthogonal=true only)
    const int64_t rows = (int64_t)Iset[b].size() * localdims[b], cols = (int64_t)Jset[b].size();
    Mat Pi1 = fillmatrix(f, Iset[b], Jset[b], 1, rows, cols);       // :607-609
    updatemaxsample(Pi1.a);
    if (b == n() - 1) { setsitetensor(b, Pi1); return; }            // :613-617
    Mat P = fillmatrix(f, Iset[b + 1], Jset[b], 0, (int64_t)Iset[b + 1].size(), (int64_t)Jset[b].size());   // :620-622
    if (Iset[b + 1].size() != Jset[b].size())
        throw std::runtime_error("Pivot matrix at bond " + std::to_string(b + 1) + " is not square!");
    Mat T;
    const double t0 = now_s();
    if (solve_right(P, Pi1, T)) throw std::runtime_error("setsitetensor!: bad shapes");
    t_lu += now_s() - t0;
    Tensor3 t; t.d0 = (int64_t)Iset[b].size(); t.d1 = localdims[b]; t.d2 = (int64_t)Iset[b + 1].size();   // :627
    t.a = T.a; sitetensors[b] = t;
}

int TCI2::sweep1site(const Func& f, bool forward, double reltol, double abstol, int64_t maxbonddim, bool updatetensors) {
    // :659-725
    pivoterrors.clear();
    invalidatesitetensors();
    const int N = n();
    for (int step = 0; step < N - 1; ++step) {
        const int b = forward ? step : N - 1 - step;                // 0-based site: 1:N-1 or N:-1:2
        IndexList Is = forward ? kronecker_left(Iset[b], localdims[b]) : Iset[b];
        IndexList Js = forward ? Jset[b] : kronecker_right(localdims[b], Jset[b]);
        Mat Pi = fillmatrix(f, Iset[b], Jset[b], 1, (int64_t)Is.size(), (int64_t)Js.size());   // :683-685
        updatemaxsample(Pi.a);
        const double t0 = now_s();
        RRLU lu;
        int st = rrlu_inplace(Pi, lu, maxbonddim, reltol, abstol, forward);        // :688-694
        if (st) return st;
        IndexList newI, newJ;
        for (int64_t k = 0; k < lu.npivot; ++k) newI.push_back(Is[lu.rowperm[k] - 1]);
        for (int64_t k = 0; k < lu.npivot; ++k) newJ.push_back(Js[lu.colperm[k] - 1]);
        Iset[forward ? b + 1 : b] = newI;                                          // :697
        Jset[forward ? b : b - 1] = newJ;                                          // :698
        if (updatetensors) setsitetensor(b, forward ? luci_left(lu) : luci_right(lu));   // :701-703
        for (double v : sitetensors[b].a) if (std::isnan(v)) return -3;            // :705-707
        t_lu += now_s() - t0;
        updateerrors(forward ? b : b - 1, lu_pivoterrors(lu));                     // :709
    }
    if (updatetensors) {                                                           // :713-723
        const int last = forward ? N - 1 : 0;
        const int64_t rows = forward ? (int64_t)Iset[last].size() : localdims[0];
        const int64_t cols = forward ? localdims[last] : (int64_t)Jset[0].size();
        setsitetensor(last, fillmatrix(f, Iset[last], Jset[last], 1, rows, cols));
    }
    return 0;
}

int TCI2::sweep0site(const Func& f, int b, double reltol, double abstol) {       // sweep0site! / rmbadpivots! :559-589 (b 0-based)
    invalidatesitetensors();
    const int64_t r0 = (int64_t)Iset[b + 1].size(), c0 = (int64_t)Jset[b].size();
    Mat P = fillmatrix(f, Iset[b + 1], Jset[b], 0, r0, c0);                     // :566-568
    updatemaxsample(P.a);                                                       // :569
    RRLU lu;
    int st = rrlu_inplace(P, lu, std::numeric_limits<int64_t>::max(), reltol, abstol, true);   // MatrixLUCI(P; reltol, abstol, leftorthogonal=true) :572-577
    if (st) return st;
    int64_t ndiag = 0;                                                          // :580 (a count over the whole diagonal, not a prefix)
    for (int64_t i = 0; i < lu.npivot; ++i) {
        const double u = lu.U(i, i), u0 = lu.U(0, 0);
        if (std::fabs(u) > abstol && std::fabs(u / u0) > reltol) ++ndiag;
    }
    IndexList newI, newJ;
    for (int64_t k = 0; k < ndiag; ++k) newI.push_back(Iset[b + 1][lu.rowperm[k] - 1]);      // :583
    for (int64_t k = 0; k < ndiag; ++k) newJ.push_back(Jset[b][lu.colperm[k] - 1]);          // :584
    Iset[b + 1] = newI; Jset[b] = newJ;
    return 0;
}

int TCI2::makecanonical(const Func& f, double reltol, double abstol, int64_t maxbonddim) {     // :738-749
    const int64_t inf = std::numeric_limits<int64_t>::max();
    int st = sweep1site(f, true, 0.0, 0.0, inf, false); if (st) return st;
    st = sweep1site(f, false, reltol, abstol, maxbonddim, false); if (st) return st;
    return sweep1site(f, true, reltol, abstol, maxbonddim, true);
}

// sweepstrategy: 0 = :forward, 1 = :backandforth (sweepstrategies.jl:41-50)
int TCI2::sweep2site(const Func& f, int niter, int iter1, double abstol, int64_t maxbonddim, int sweepstrategy,
                     bool strictlynested, bool fillsite) {                         // :1195-1258
    invalidatesitetensors();
    const int N = n();
    for (int iter = iter1; iter < iter1 + niter; ++iter) {
        std::vector<IndexList> extraI(N), extraJ(N);
        if (!strictlynested && !Iset_history.empty()) { extraI = Iset_history.back(); extraJ = Jset_history.back(); }
        Iset_history.push_back(Iset); Jset_history.push_back(Jset);               // :1220-1221
        pivoterrors.clear();                                                       // :1223
        const bool fwd = (sweepstrategy == 0) || (sweepstrategy == 1 && (iter % 2 == 1));
        for (int pass = 0; pass < (bondorder == 1 ? 2 : 1); ++pass)
            for (int s = 0; s < N - 1; ++s) {
                const int b = fwd ? s : N - 2 - s;
                if (bondorder == 1 && (b % 2) != pass) continue;          // pass 0: bonds 1, 3, 5, ... (1-based); pass 1: 2, 4, ...
                int st = updatepivots(b, f, fwd, 1e-14, abstol, maxbonddim, extraI[b + 1], extraJ[b]);   // :1227-1249
                if (st) return st;
            }
    }
    if (fillsite) fillsitetensors(f);                                              // :1254-1256
    return 0;
}

double TCI2::evaluate(const MultiIndex& idx) const {               // abstracttensortrain.jl:328-342
    if ((int)idx.size() != n()) throw std::invalid_argument("evaluate: wrong number of indices");
    std::vector<double> v(1, 1.0);
    for (int b = 0; b < n(); ++b) {
        const Tensor3& T = sitetensors[b];
        if ((int64_t)v.size() != T.d0) throw std::runtime_error("evaluate: site tensors not available");
        std::vector<double> w((size_t)T.d2, 0.0);
        for (int64_t c = 0; c < T.d2; ++c) {
            double s = 0.0;
            for (int64_t a = 0; a < T.d0; ++a) s += v[a] * T.a[(size_t)((c * T.d1 + (idx[b] - 1)) * T.d0 + a)];
            w[c] = s;
        }
        v.swap(w);
    }
    if (v.size() != 1) throw std::runtime_error("evaluate: result is not a scalar");
    return v[0];
}

static bool convergencecriterion(const std::vector<int64_t>& ranks, const std::vector<double>& errors,
                                 const std::vector<int64_t>& nglobalpivots, double tolerance, int64_t maxbonddim,
                                 int ncheckhistory, bool checkconvglobalpivot) {   // :947-966
    if ((int)errors.size() < ncheckhistory) return false;
    const size_t nr = std::min(ranks.size(), (size_t)ncheckhistory);
    const size_t ng = std::min(nglobalpivots.size(), (size_t)ncheckhistory);
    bool allerr = true;
    for (size_t i = errors.size() - ncheckhistory; i < errors.size(); ++i) allerr = allerr && (errors[i] < tolerance);
    bool allg = true;
    if (checkconvglobalpivot) for (size_t i = nglobalpivots.size() - ng; i < nglobalpivots.size(); ++i) allg = allg && (nglobalpivots[i] == 0);
    int64_t minrank = std::numeric_limits<int64_t>::max(); bool allmax = true;
    for (size_t i = ranks.size() - nr; i < ranks.size(); ++i) { minrank = std::min(minrank, ranks[i]); allmax = allmax && (ranks[i] >= maxbonddim); }
    return (allerr && allg && minrank == ranks.back()) || allmax;
}

struct OptimizeResult { std::vector<int64_t> ranks; std::vector<double> errors; int status = 0; };

// DefaultGlobalPivotFinder — globalpivotfinder.jl:205-264: nsearch random starts; from each, one line search per site over
// |f - tt| (strict '>' keeps the first maximum, the start's own coordinate is restored after each site); a start contributes its
// best point when that error exceeds abstol * tolmarginglobalsearch; the list is cut to maxnglobalpivot.  The reference draws
// the starts from Julia's default RNG (`rand(rng, 1:d)`, :219), which cannot be restated; oracle and product share the seeded
// splitmix64 stream `grng` instead (all starts are drawn before any search, start-major, site-minor, as in :219).  f is read
// through ONE batch (left set = the candidate points, right set = the empty index): the values are those of the point calls.
static IndexList find_global_pivots(TCI2& tci, const Func& f, double abstol, int nsearch, int maxnglobalpivot, double tolmargin) {
    const int L = tci.n();
    IndexList found;
    if (nsearch <= 0) return found;
    IndexList initial((size_t)nsearch, MultiIndex((size_t)L));
    for (int s = 0; s < nsearch; ++s)
        for (int p = 0; p < L; ++p) initial[(size_t)s][(size_t)p] = 1 + (int32_t)tci.grng.randint(tci.localdims[(size_t)p]);
    IndexList pts;
    for (const auto& point : initial)
        for (int p = 0; p < L; ++p)
            for (int v = 1; v <= tci.localdims[(size_t)p]; ++v) { MultiIndex c(point); c[(size_t)p] = v; pts.push_back(c); }
    std::vector<int64_t> shape;
    const std::vector<double> fv = batchevaluate(f, tci.localdims, pts, IndexList(1, MultiIndex()), 0, &shape);
    size_t k = 0;
    for (const auto& point : initial) {
        double best_error = 0.0;
        MultiIndex best_point(point);
        for (int p = 0; p < L; ++p)
            for (int v = 1; v <= tci.localdims[(size_t)p]; ++v, ++k) {
                const double error = std::fabs(fv[k] - tci.evaluate(pts[k]));              // :236
                if (error > best_error) { best_error = error; best_point = pts[k]; }       // :239-242
            }
        if (best_error > abstol * tolmargin) found.push_back(best_point);                  // :249-251
    }
    if ((int64_t)found.size() > maxnglobalpivot) found.resize((size_t)maxnglobalpivot);    // :255-257
    return found;
}

// optimize! — :1018-1172 with the DefaultGlobalPivotFinder (nsearchglobalpivot = 0 turns the search off).
static OptimizeResult optimize(TCI2& tci, const Func& f, double tol, int64_t maxbonddim, int maxiter,
                               int sweepstrategy, bool normalizeerror, int ncheckhistory, bool strictlynested,
                               bool checkconvglobalpivot, int nsearchglobalpivot = 0, int maxnglobalpivot = 5,
                               double tolmarginglobalsearch = 10.0) {
    OptimizeResult R;
    std::vector<int64_t> nglobalpivots;
    if (nsearchglobalpivot > 0 && nsearchglobalpivot < maxnglobalpivot) throw std::invalid_argument("nsearchglobalpivot < maxnglobalpivot!");   // :1056-1058
    if (maxbonddim >= std::numeric_limits<int64_t>::max() && tol <= 0)
        throw std::invalid_argument("Specify either tolerance > 0 or some maxbonddim; otherwise, the convergence criterion is not reachable!");
    for (int iter = 1; iter <= maxiter; ++iter) {
        const double errnorm = normalizeerror ? tci.maxsamplevalue : 1.0;          // :1089-1090
        const double abstol = tol * errnorm;
        R.status = tci.sweep2site(f, 2, 1, abstol, maxbonddim, sweepstrategy, strictlynested, true);   // :1098-1108
        if (R.status) return R;
        R.errors.push_back(tci.pivoterror());                                      // :1118
        const IndexList found = find_global_pivots(tci, f, abstol, nsearchglobalpivot, maxnglobalpivot, tolmarginglobalsearch);   // :1126-1131
        tci.addglobalpivots(found);                                                // :1132
        nglobalpivots.push_back((int64_t)found.size());                            // :1133
        R.ranks.push_back(tci.rank());                                             // :1140
        if (convergencecriterion(R.ranks, R.errors, nglobalpivots, abstol, maxbonddim, ncheckhistory, checkconvglobalpivot))
            break;                                                                 // :1147-1154
    }
    const double errnorm = normalizeerror ? tci.maxsamplevalue : 1.0;              // :1160-1161
    const double abstol = tol * errnorm;
    R.status = tci.sweep1site(f, true, 1e-14, abstol, maxbonddim, true);           // :1162-1167
    if (R.status) return R;
    for (int b = 0; b + 1 < tci.n(); ++b)                                          // _sanitycheck globalsearch.jl:226-233
        if (tci.Iset[b + 1].size() != tci.Jset[b].size()) throw std::runtime_error("Pivot matrix at bond " + std::to_string(b + 1) + " is not square!");
    for (auto& e : R.errors) e = e / errnorm;                                      // :1171
    return R;
}

// ---------------------------------------------------------------------------------------
// TensorTrain sum / evaluate / fulltensor — src/abstracttensortrain.jl:328-342, :428-441,
// src/tensortrain.jl:580-600.  Cores are (chi_l, d, chi_r) column-major, d = product of site legs.
// ---------------------------------------------------------------------------------------
static double tt_sum(const std::vector<Tensor3>& tt) {
    std::vector<double> v(1, 1.0);                 // sum(tt[1], dims=(1,2)) folds the leading bond (size 1) too
    bool first = true;
    for (const Tensor3& T : tt) {
        std::vector<double> w((size_t)T.d2, 0.0);
        if (first) {
            for (int64_t c = 0; c < T.d2; ++c) { double s = 0.0; for (int64_t x = 0; x < T.d0 * T.d1; ++x) s += T.a[(size_t)(c * T.d0 * T.d1 + x)]; w[c] = s; }
            first = false;
        } else {
            if ((int64_t)v.size() != T.d0) throw std::runtime_error("sum: inconsistent bond dimensions");
            Mat S(T.d0, T.d2, 0.0);                // sum(T, dims=2)[:, 1, :]
            for (int64_t c = 0; c < T.d2; ++c) for (int64_t s = 0; s < T.d1; ++s) for (int64_t a = 0; a < T.d0; ++a)
                S(a, c) += T.a[(size_t)((c * T.d1 + s) * T.d0 + a)];
            for (int64_t c = 0; c < T.d2; ++c) { double s = 0.0; for (int64_t a = 0; a < T.d0; ++a) s += v[a] * S(a, c); w[c] = s; }
        }
        v.swap(w);
    }
    if (v.size() != 1) throw std::runtime_error("sum: result is not a scalar");
    return v[0];
}
static double tt_evaluate(const std::vector<Tensor3>& tt, const int32_t* idx) {
    std::vector<double> v(1, 1.0);
    for (size_t b = 0; b < tt.size(); ++b) {
        const Tensor3& T = tt[b];
        if ((int64_t)v.size() != T.d0) throw std::runtime_error("evaluate: inconsistent bond dimensions");
        std::vector<double> w((size_t)T.d2, 0.0);
        for (int64_t c = 0; c < T.d2; ++c) { double s = 0.0; for (int64_t a = 0; a < T.d0; ++a) s += v[a] * T.a[(size_t)((c * T.d1 + (idx[b] - 1)) * T.d0 + a)]; w[c] = s; }
        v.swap(w);
    }
    return v[0];
}

} // namespace orc

#include <chrono>
namespace orc {
static double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}
}

// =======================================================================================
// C API (ctypes) — used by tests/, smoke() and bench.py's CPU baseline only.
// All matrices column-major; index values 1-based; status 0 = ok.
// =======================================================================================
using namespace orc;

static thread_local std::string g_err;
#define ORC_TRY try {
#define ORC_CATCH(code) } catch (const std::exception& e) { g_err = e.what(); return code; }

static Func make_func(int kind, const double* params, int nparams, const int32_t* dims, int ndims,
                      const double* table, orc_callback cb, void* user) {
    Func f; f.kind = kind; f.p.assign(params, params + nparams);
    if (kind == F_DENSE) {
        f.dims.assign(dims, dims + ndims);
        size_t tot = 1; for (int i = 0; i < ndims; ++i) tot *= dims[i];
        f.table.assign(table, table + tot);
    }
    f.cb = cb; f.user = user;
    if (kind == F_BATCHCB) { f.bcb = (orc_batch_callback)(void*)cb; f.cb = nullptr; }
    return f;
}
static IndexList to_list(const int32_t* flat, int64_t count, int len) {   // row-major table: count rows of `len` ints
    IndexList out((size_t)count);
    for (int64_t i = 0; i < count; ++i) out[(size_t)i].assign(flat + i * len, flat + (i + 1) * len);
    return out;
}

extern "C" {

const char* orc_last_error() { return g_err.c_str(); }

// submatrixargmax with explicit 0-based row/col lists; mode 0 identity, 1 abs2
int orc_submatrixargmax(const double* A, int64_t lda, int mode, const int64_t* rows, int64_t nrows,
                        const int64_t* cols, int64_t ncols, int64_t* outr, int64_t* outc) {
    ORC_TRY submatrixargmax(A, lda, mode, rows, nrows, cols, ncols, outr, outc); return 0; ORC_CATCH(-10)
}

// Threads the next orc_rrlu calls of this thread use (1 = the plain scalar loop, the reference's execution model).
static thread_local int g_rrlu_threads = 1;
void orc_rrlu_set_threads(int n) { g_rrlu_threads = n < 1 ? 1 : n; }
int orc_max_threads() { return omp_get_max_threads(); }

// rrlu (copies A).  L is m x ldl_cols (first npivot columns valid), U is maxr x n with leading dim ldu.
// pivoterrors must hold min(maxrank,m,n)+1 doubles; returns npivot+1 entries.
int orc_rrlu(const double* A, int64_t m, int64_t n, int64_t lda, int64_t maxrank, double reltol, double abstol,
             int leftorthogonal, int64_t* rowperm, int64_t* colperm, int64_t* npivot, double* error,
             double* L, int64_t ldl, double* U, int64_t ldu, double* pivoterrors) {
    ORC_TRY
    Mat W(m, n);
    for (int64_t j = 0; j < n; ++j) std::memcpy(&W.a[(size_t)j * m], A + (size_t)j * lda, sizeof(double) * m);
    RRLU lu;
    int st;
    if (g_rrlu_threads > 1) {
        lu.leftorthogonal = leftorthogonal != 0;
        lu.rowperm.resize((size_t)m); lu.colperm.resize((size_t)n);
        for (int64_t i = 0; i < m; ++i) lu.rowperm[(size_t)i] = i + 1;
        for (int64_t j = 0; j < n; ++j) lu.colperm[(size_t)j] = j + 1;
        st = optimizerrlu_mt(lu, W, maxrank, reltol, abstol, g_rrlu_threads);
    } else {
        st = rrlu_inplace(W, lu, maxrank, reltol, abstol, leftorthogonal != 0);
    }
    *npivot = lu.npivot; *error = lu.error;
    std::copy(lu.rowperm.begin(), lu.rowperm.end(), rowperm);
    std::copy(lu.colperm.begin(), lu.colperm.end(), colperm);
    if (st) return st;
    if (L) for (int64_t j = 0; j < lu.npivot; ++j) std::memcpy(L + (size_t)j * ldl, &lu.L.a[(size_t)j * m], sizeof(double) * m);
    if (U) for (int64_t j = 0; j < n; ++j) for (int64_t i = 0; i < lu.npivot; ++i) U[(size_t)j * ldu + i] = lu.U(i, j);
    auto pe = lu_pivoterrors(lu);
    std::copy(pe.begin(), pe.end(), pivoterrors);
    return 0;
    ORC_CATCH(-10)
}

// arrlu on a dense matrix used as f(i, j) = A[i, j] (the form test/test_matrixlu.jl:79 calls).  I0/J0 1-based.
// L is m x min(maxrank,m,n) (ld ldl), U has ld ldu; *nevals = matrix elements sampled.
int orc_arrlu(const double* A, int64_t m, int64_t n, int64_t lda, const int64_t* I0, int64_t nI0, const int64_t* J0, int64_t nJ0,
              int64_t maxrank, double reltol, double abstol, int leftorthogonal, int numrookiter, uint64_t seed,
              int64_t* rowperm, int64_t* colperm, int64_t* npivot, double* error,
              double* L, int64_t ldl, double* U, int64_t ldu, double* pivoterrors, int64_t* nevals) {
    ORC_TRY
    int64_t count = 0;
    BatchF f = [&](const std::vector<int64_t>& ir, const std::vector<int64_t>& ic) {
        Mat S((int64_t)ir.size(), (int64_t)ic.size());
        for (size_t j = 0; j < ic.size(); ++j) for (size_t i = 0; i < ir.size(); ++i) S((int64_t)i, (int64_t)j) = A[(size_t)(ic[j] - 1) * lda + (ir[i] - 1)];
        count += S.m * S.n;
        return S;
    };
    RRLU lu; Rng rng; rng.s = seed;
    int st = arrlu(f, m, n, std::vector<int64_t>(I0, I0 + nI0), std::vector<int64_t>(J0, J0 + nJ0), lu, maxrank, reltol, abstol,
                   leftorthogonal != 0, numrookiter, rng);
    *npivot = lu.npivot; *error = lu.error; *nevals = count;
    std::copy(lu.rowperm.begin(), lu.rowperm.end(), rowperm);
    std::copy(lu.colperm.begin(), lu.colperm.end(), colperm);
    if (st) return st;
    for (int64_t j = 0; j < lu.npivot; ++j) for (int64_t i = 0; i < m; ++i) L[(size_t)j * ldl + i] = lu.L(i, j);
    for (int64_t j = 0; j < n; ++j) for (int64_t i = 0; i < lu.npivot; ++i) U[(size_t)j * ldu + i] = lu.U(i, j);
    auto pe = lu_pivoterrors(lu);
    std::copy(pe.begin(), pe.end(), pivoterrors);
    return 0;
    ORC_CATCH(-10)
}

// MatrixLUCI left (m x r) and right (r x n) from rrlu outputs (L m x r ld ldl, U r x n ld ldu)
int orc_luci_factors(int64_t m, int64_t n, int64_t r, int leftorthogonal, const int64_t* rowperm, const int64_t* colperm,
                     const double* L, int64_t ldl, const double* U, int64_t ldu, double* left, double* right) {
    ORC_TRY
    RRLU lu; lu.leftorthogonal = leftorthogonal != 0; lu.npivot = r;
    lu.rowperm.assign(rowperm, rowperm + m); lu.colperm.assign(colperm, colperm + n);
    lu.L = Mat(m, r); lu.U = Mat(r, n);
    for (int64_t j = 0; j < r; ++j) for (int64_t i = 0; i < m; ++i) lu.L(i, j) = L[(size_t)j * ldl + i];
    for (int64_t j = 0; j < n; ++j) for (int64_t i = 0; i < r; ++i) lu.U(i, j) = U[(size_t)j * ldu + i];
    Mat l = luci_left(lu), rr = luci_right(lu);
    std::copy(l.a.begin(), l.a.end(), left);
    std::copy(rr.a.begin(), rr.a.end(), right);
    return 0;
    ORC_CATCH(-10)
}

// X = B * inv(P)  (tensorci2.jl:626); B is q x r, P is r x r
int orc_solve_right(const double* P, int64_t r, const double* B, int64_t q, double* X) {
    ORC_TRY
    Mat Pm(r, r), Bm(q, r), Xm;
    std::copy(P, P + r * r, Pm.a.begin()); std::copy(B, B + q * r, Bm.a.begin());
    int st = solve_right(Pm, Bm, Xm); if (st) return st;
    std::copy(Xm.a.begin(), Xm.a.end(), X);
    return 0;
    ORC_CATCH(-10)
}

void* orc_func_create(int kind, const double* params, int nparams, const int32_t* dims, int ndims,
                      const double* table, orc_callback cb, void* user) {
    try { return new Func(make_func(kind, params, nparams, dims, ndims, table, cb, user)); }
    catch (const std::exception& e) { g_err = e.what(); return nullptr; }
}
void orc_func_destroy(void* f) { delete (Func*)f; }
double orc_func_eval(void* f, const int32_t* idx, int n) { return (*(Func*)f)(idx, n); }

// filltensor / _batchevaluate_dispatch: out has nleft * prod(centre dims) * nright doubles
int orc_batcheval(void* f, const int32_t* localdims, int nsites, const int32_t* left, int64_t nleft, int nl,
                  const int32_t* right, int64_t nright, int nr, int M, double* out) {
    ORC_TRY
    std::vector<int32_t> ld(localdims, localdims + nsites);
    if (nl + nr + M != nsites) throw std::runtime_error("Invalid number of central indices");   // tensorci2.jl:492
    std::vector<int64_t> shape;
    auto res = batchevaluate(*(Func*)f, ld, to_list(left, nleft, nl), to_list(right, nright, nr), M, &shape);
    std::copy(res.begin(), res.end(), out);
    return 0;
    ORC_CATCH(-10)
}

// ---- TensorCI2 handle -------------------------------------------------------------------
struct OrcTCI { TCI2 tci; Func* f; OrcTCI(const std::vector<int32_t>& ld, Func* f_) : tci(ld), f(f_) {} };

void* orc_tci_create(void* f, const int32_t* localdims, int nsites, const int32_t* pivots, int64_t npivots) {
    try {
        std::vector<int32_t> ld(localdims, localdims + nsites);
        OrcTCI* h = new OrcTCI(ld, (Func*)f);
        if (npivots > 0) h->tci.init(*h->f, to_list(pivots, npivots, nsites));
        return h;
    } catch (const std::exception& e) { g_err = e.what(); return nullptr; }
}
void orc_tci_destroy(void* h) { delete (OrcTCI*)h; }
void orc_tci_set_bondorder(void* h, int v) { ((OrcTCI*)h)->tci.bondorder = v; }
void orc_tci_set_pivotsearch(void* h, int rook, uint64_t seed) { ((OrcTCI*)h)->tci.pivotsearch = rook; ((OrcTCI*)h)->tci.rng.s = seed; }

// optimize!: ranks/errors must hold maxiter entries; *niter receives the number written.
int orc_tci_optimize_gp(void* h_, double tolerance, int64_t maxbonddim, int maxiter, int sweepstrategy, int normalizeerror,
                        int ncheckhistory, int strictlynested, int checkconvglobalpivot, int nsearchglobalpivot, int maxnglobalpivot,
                        double tolmarginglobalsearch, uint64_t seed, int64_t* ranks, double* errors, int* niter);
int orc_tci_optimize(void* h_, double tolerance, int64_t maxbonddim, int maxiter, int sweepstrategy, int normalizeerror,
                     int ncheckhistory, int strictlynested, int checkconvglobalpivot,
                     int64_t* ranks, double* errors, int* niter) {
    return orc_tci_optimize_gp(h_, tolerance, maxbonddim, maxiter, sweepstrategy, normalizeerror, ncheckhistory, strictlynested,
                               checkconvglobalpivot, 0, 5, 10.0, 0, ranks, errors, niter);
}
// the same with the DefaultGlobalPivotFinder: nsearchglobalpivot starts per iteration drawn from the stream seeded with `seed`
int orc_tci_optimize_gp(void* h_, double tolerance, int64_t maxbonddim, int maxiter, int sweepstrategy, int normalizeerror,
                        int ncheckhistory, int strictlynested, int checkconvglobalpivot, int nsearchglobalpivot, int maxnglobalpivot,
                        double tolmarginglobalsearch, uint64_t seed, int64_t* ranks, double* errors, int* niter) {
    ORC_TRY
    OrcTCI* h = (OrcTCI*)h_;
    h->tci.grng.s = seed;
    OptimizeResult R = optimize(h->tci, *h->f, tolerance, maxbonddim, maxiter, sweepstrategy, normalizeerror != 0,
                                ncheckhistory, strictlynested != 0, checkconvglobalpivot != 0, nsearchglobalpivot, maxnglobalpivot,
                                tolmarginglobalsearch);
    *niter = (int)R.ranks.size();
    std::copy(R.ranks.begin(), R.ranks.end(), ranks);
    std::copy(R.errors.begin(), R.errors.end(), errors);
    return R.status;
    ORC_CATCH(-10)
}
// updatepivots! on 1-based bond b with no extra index sets (the form test_tensorci2.jl:259 calls)
int orc_tci_updatepivots(void* h_, int b, int leftorthogonal, double reltol, double abstol, int64_t maxbonddim) {
    ORC_TRY OrcTCI* h = (OrcTCI*)h_;
    return h->tci.updatepivots(b - 1, *h->f, leftorthogonal != 0, reltol, abstol, maxbonddim, IndexList(), IndexList());
    ORC_CATCH(-10)
}
int orc_tci_sweep2site(void* h_, int niter, int iter1, double abstol, int64_t maxbonddim, int sweepstrategy,
                       int strictlynested, int fillsitetensors) {
    ORC_TRY OrcTCI* h = (OrcTCI*)h_;
    return h->tci.sweep2site(*h->f, niter, iter1, abstol, maxbonddim, sweepstrategy, strictlynested != 0, fillsitetensors != 0);
    ORC_CATCH(-10)
}
int orc_tci_sweep1site(void* h_, int forward, double reltol, double abstol, int64_t maxbonddim, int updatetensors) {
    ORC_TRY OrcTCI* h = (OrcTCI*)h_;
    return h->tci.sweep1site(*h->f, forward != 0, reltol, abstol, maxbonddim, updatetensors != 0);
    ORC_CATCH(-10)
}
int orc_tci_sweep0site(void* h_, int b, double reltol, double abstol) {
    ORC_TRY OrcTCI* h = (OrcTCI*)h_;
    if (b < 1 || b >= h->tci.n()) return -11;
    return h->tci.sweep0site(*h->f, b - 1, reltol, abstol);
    ORC_CATCH(-10)
}
int orc_tci_makecanonical(void* h_, double reltol, double abstol, int64_t maxbonddim) {
    ORC_TRY OrcTCI* h = (OrcTCI*)h_;
    return h->tci.makecanonical(*h->f, reltol, abstol, maxbonddim);
    ORC_CATCH(-10)
}
int orc_tci_addglobalpivots(void* h_, const int32_t* pivots, int64_t npivots) {
    ORC_TRY OrcTCI* h = (OrcTCI*)h_;
    h->tci.addglobalpivots(to_list(pivots, npivots, h->tci.n())); return 0;
    ORC_CATCH(-10)
}
int orc_tci_fillsitetensors(void* h_) {
    ORC_TRY OrcTCI* h = (OrcTCI*)h_; h->tci.fillsitetensors(*h->f); return 0; ORC_CATCH(-10)
}
// which: 0 = Iset, 1 = Jset; b 1-based.  Returns the number of multi-indices, *len their length.
int64_t orc_tci_setsize(void* h_, int which, int b, int* len) {
    OrcTCI* h = (OrcTCI*)h_;
    const IndexList& s = which ? h->tci.Jset[b - 1] : h->tci.Iset[b - 1];
    *len = which ? h->tci.n() - b : b - 1;
    return (int64_t)s.size();
}
void orc_tci_getset(void* h_, int which, int b, int32_t* out) {
    OrcTCI* h = (OrcTCI*)h_;
    const IndexList& s = which ? h->tci.Jset[b - 1] : h->tci.Iset[b - 1];
    for (const auto& x : s) { std::copy(x.begin(), x.end(), out); out += x.size(); }
}
void orc_tci_sitetensor_shape(void* h_, int b, int64_t* shape) {
    const Tensor3& t = ((OrcTCI*)h_)->tci.sitetensors[b - 1]; shape[0] = t.d0; shape[1] = t.d1; shape[2] = t.d2;
}
void orc_tci_sitetensor(void* h_, int b, double* out) {
    const Tensor3& t = ((OrcTCI*)h_)->tci.sitetensors[b - 1]; std::copy(t.a.begin(), t.a.end(), out);
}
double orc_tci_maxsamplevalue(void* h_) { return ((OrcTCI*)h_)->tci.maxsamplevalue; }
int64_t orc_tci_npivoterrors(void* h_) { return (int64_t)((OrcTCI*)h_)->tci.pivoterrors.size(); }
void orc_tci_pivoterrors(void* h_, double* out) { auto& v = ((OrcTCI*)h_)->tci.pivoterrors; std::copy(v.begin(), v.end(), out); }
void orc_tci_bonderrors(void* h_, double* out) { auto& v = ((OrcTCI*)h_)->tci.bonderrors; std::copy(v.begin(), v.end(), out); }
int orc_tci_evaluate(void* h_, const int32_t* idx, double* out) {
    ORC_TRY OrcTCI* h = (OrcTCI*)h_;
    *out = h->tci.evaluate(MultiIndex(idx, idx + h->tci.n())); return 0;
    ORC_CATCH(-10)
}
// counters for the CPU baseline: evaluations and seconds in Pi fill / factorisation
void orc_tci_stats(void* h_, int64_t* nevals, double* t_eval, double* t_lu) {
    OrcTCI* h = (OrcTCI*)h_; *nevals = h->tci.nevals; *t_eval = h->tci.t_eval; *t_lu = h->tci.t_lu;
}

// ---- TensorTrain sum / evaluate on flat cores ----------------------------------------------
// cores: concatenated column-major (chi_l, d, chi_r) blocks; shapes: 3 * nsites int64
static std::vector<Tensor3> to_tt(const double* cores, const int64_t* shapes, int nsites) {
    std::vector<Tensor3> tt((size_t)nsites);
    for (int b = 0; b < nsites; ++b) {
        Tensor3& t = tt[(size_t)b]; t.d0 = shapes[3 * b]; t.d1 = shapes[3 * b + 1]; t.d2 = shapes[3 * b + 2];
        const size_t sz = (size_t)(t.d0 * t.d1 * t.d2);
        t.a.assign(cores, cores + sz); cores += sz;
    }
    return tt;
}
int orc_tt_sum(const double* cores, const int64_t* shapes, int nsites, double* out) {
    ORC_TRY *out = tt_sum(to_tt(cores, shapes, nsites)); return 0; ORC_CATCH(-10)
}
int orc_tt_evaluate(const double* cores, const int64_t* shapes, int nsites, const int32_t* idx, int64_t npoints, double* out) {
    ORC_TRY
    auto tt = to_tt(cores, shapes, nsites);
    for (int64_t p = 0; p < npoints; ++p) out[p] = tt_evaluate(tt, idx + p * nsites);
    return 0;
    ORC_CATCH(-10)
}

} // extern "C"
