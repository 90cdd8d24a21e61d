// functors.h — the built-in device functors (monoid form) shared by the batch-evaluation kernels (batcheval.cu) and the
// device-resident sweep kernels (sweep.cu).  A functor folds a small State over the sites of a multi-index (`push`), states
// of disjoint site ranges combine (`merge`) and `finish` maps the state of the full index to the function value.
#pragma once

enum { TCI_F_LORENTZ = 1, TCI_F_QEXPSIN = 2, TCI_F_QEXPPLUSSIN = 3, TCI_F_SUM = 4, TCI_F_QEXP2 = 5, TCI_F_DENSE = 6, TCI_F_HASH = 7, TCI_F_USER = 100 };

struct tci_functor_s {
    int kind = 0;
    double p[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    int nparams = 0;
    double* d_table = nullptr;           // F_DENSE
    long long* d_strides = nullptr;
    int ndims = 0;
    double* d_params = nullptr;          // F_USER
    cudaLibrary_t mod = nullptr;         // NVRTC cubin loaded through the runtime API (no libcuda link dependency)
    cudaKernel_t fn = nullptr;
};

namespace tci_functors {

// ---- built-in monoid functors ---------------------------------------------------------------------
struct LorentzF {                        // p0 / (1 + sum v^2): README.md:21-29
    double coeff;
    typedef long long State;
    __device__ State zero() const { return 0; }
    __device__ void push(State& s, int, int v) const { s += (long long)v * v; }
    __device__ State merge(State a, State b) const { return a + b; }
    __device__ double finish(State s) const {
        const long long d = 1 + s;
        // exact int64 -> double for 0 <= d < 2^52 with one add (the I2F.F64.S64 conversion runs on the slow XU path): the
        // bits 0x433 << 52 | d are the double 2^52 + d
        const double x = (d >= 0 && d < (1LL << 52)) ? __longlong_as_double(0x4330000000000000LL | d) - 4503599627370496.0 : (double)d;
        // coeff == 1 (the README function): the correctly rounded reciprocal — the same bits as the IEEE division 1.0 / x
        if (coeff == 1.0) return __drcp_rn(x);
        return coeff / x;
    }
};
struct LorentzUnitF {                    // coeff == 1: 1 / (1 + sum v^2) through the correctly rounded reciprocal (the same bits as
    typedef long long State;             // the IEEE division 1.0 / x of LorentzF, without the multiply-and-correct tail of a division)
    __device__ State zero() const { return 0; }
    __device__ void push(State& s, int, int v) const { s += (long long)v * v; }
    __device__ State merge(State a, State b) const { return a + b; }
    __device__ double finish(State s) const {
        const long long d = 1 + s;
        const double x = (d >= 0 && d < (1LL << 52)) ? __longlong_as_double(0x4330000000000000LL | d) - 4503599627370496.0 : (double)d;
        return __drcp_rn(x);
    }
};
struct SumF {                            // p0 * sum v: test/test_batcheval.jl:18
    double scale;
    typedef long long State;
    __device__ State zero() const { return 0; }
    __device__ void push(State& s, int, int v) const { s += v; }
    __device__ State merge(State a, State b) const { return a + b; }
    __device__ double finish(State s) const { return scale * (double)s; }
};
// quantics coordinate x = sum_i (b_i - 1) 2^-(i)  (1-based site i; exact in FP64 for <= 52 bits, so the
// order of the partial sums does not matter)
template <int KIND>
struct QuanticsF {
    double p0, p1, p2;
    typedef double State;
    __device__ State zero() const { return 0.0; }
    // (v - 1) * 2^-(site+1): the power of two is built from its exponent field (exact, the same value ldexp returns; site < 1022)
    __device__ void push(State& s, int site, int v) const {
        s += (double)(v - 1) * __longlong_as_double((long long)(1022 - site) << 52);
    }
    __device__ State merge(State a, State b) const { return a + b; }
    __device__ double finish(State x) const {
        if (KIND == TCI_F_QEXPSIN) return exp(-p0 * x) * sin(2.0 * 3.14159265358979323846 * p1 * pow(x, p2));   // test_tensorci2.jl:437
        if (KIND == TCI_F_QEXPPLUSSIN) return exp(-p0 * x) + p1 * sin(p2 * x);                                  // test_globalsearch.jl:14
        return exp(-p0 * x) + p1 * exp(-p2 * x);                                                               // test_tensorci2.jl:63,:146
    }
};
// p0 * u(v) + p1 / (1 + sum v^2), u = integer hash of the multi-index in [-1, 1): a synthetic FULL-RANK function (not in the
// reference) whose TT rank saturates any maxbonddim; used to benchmark sweeps in the HBM-bound regime.
struct HashF {
    double amp, lor;
    struct State { unsigned long long h; long long s2; };
    __device__ static unsigned long long mix(unsigned long long z) {
        z += 0x9e3779b97f4a7c15ULL; z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL; z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
        return z ^ (z >> 31);
    }
    __device__ State zero() const { return State{0ULL, 0LL}; }
    __device__ void push(State& s, int site, int v) const { s.h += (unsigned long long)v * (mix((unsigned long long)site + 1) | 1ULL); s.s2 += (long long)v * v; }
    __device__ State merge(State a, State b) const { return State{a.h + b.h, a.s2 + b.s2}; }
    __device__ double finish(State s) const {
        const double u = (double)(mix(s.h) >> 11) * 0x1.0p-52 - 1.0;
        return amp * u + lor / (double)(1 + s.s2);
    }
};
struct DenseF {                          // column-major table lookup
    const double* table; const long long* strides;
    typedef long long State;
    __device__ State zero() const { return 0; }
    __device__ void push(State& s, int site, int v) const { s += (long long)(v - 1) * strides[site]; }
    __device__ State merge(State a, State b) const { return a + b; }
    __device__ double finish(State s) const { return table[s]; }
};

} // namespace tci_functors
