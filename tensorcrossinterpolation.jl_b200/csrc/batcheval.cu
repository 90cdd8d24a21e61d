// batcheval.cu — K1: batched evaluation of the target function over Iset x (centre sites) x Jset and
// assembly of the result (the Pi matrix of src/tensorci2.jl:848-852) straight into HBM.
//
// Replaces the scalar triple loop of src/batcheval.jl:157-171 (one closure call per element on a
// reused Vector{Int}) and the extra pass of updatemaxsample! (src/tensorci2.jl:636-638, util.jl:34-43).
//
// Layout decisions (B200-first, not a translation):
//   * the function is a DEVICE functor.  Built-ins are monoid functors: a small State is folded over
//     the sites (push), states of disjoint site ranges combine (merge), and `finish` maps the state to
//     the value.  That makes the evaluation separable: one tiny kernel folds every left multi-index and
//     every right multi-index once (|I| + |J| threads), the assembly kernel then does
//     merge(L[i], centre digits, R[j]) + finish per element — O(1) integer work per element instead of
//     O(number of sites), and no per-thread index array.
//   * user functors arrive as CUDA source through the C ABI and are compiled with NVRTC for sm_100a
//     (tci_functor_from_source); they see the full multi-index like the reference's f(indexset).
//   * output element (i, c, j) goes to out[i + c*stride_c + j*stride_j]: consecutive threads write
//     consecutive i (coalesced 256 B per warp); the strides let the caller pad the leading dimension
//     so the rrLU kernels can use 128-bit accesses on the same buffer.
//   * max|value| (and a NaN flag, Julia's max propagates NaN) is reduced in the same pass.
#include "context.h"
#include <nvrtc.h>
#include <cmath>
#include <cstring>
#include <mutex>
#include <vector>

#include "functors.h"
using namespace tci_functors;

namespace {

constexpr int kMaxSites = 128;

struct EvalShape {
    int nl, nr, ncent_sites;             // multi-index lengths
    int cdims[8];                        // centre local dims (ncent_sites <= 8)
    long long nleft, nright, ncent;      // counts
    long long stride_c, stride_j;        // output strides (stride of i is 1)
    int nouts;                           // number of destination buffers (1 = local only)
    double* outs[8];                     // destinations: this GPU's buffer and, for a sharded batch, the peers' buffers
                                         // (peer-mapped through NVLink): the assembly IS the all-gather
};

// fold the left multi-indices (sites 0..nl-1) and the right ones (sites N-nr..N-1); with Cs != NULL also, per (c, j) pair
// (c = centre multi-index, first site fastest), the state of "centre digits + right index" and the pair's output offset, so
// that the assembly kernel does no index arithmetic at all
template <class F>
__global__ void fold_kernel(F f, const int* __restrict__ left, const int* __restrict__ right, EvalShape sh, int nsites,
                            typename F::State* Ls, typename F::State* Rs, typename F::State* Cs, long long* Off) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < sh.nleft) {
        typename F::State s = f.zero();
        for (int p = 0; p < sh.nl; ++p) f.push(s, p, left[t * sh.nl + p]);
        Ls[t] = s;
    } else if (t < sh.nleft + sh.nright) {
        const long long j = t - sh.nleft;
        typename F::State s = f.zero();
        for (int p = 0; p < sh.nr; ++p) f.push(s, nsites - sh.nr + p, right[j * sh.nr + p]);
        Rs[j] = s;
    } else if (Cs && t < sh.nleft + sh.nright + sh.ncent * sh.nright) {
        const long long cj = t - sh.nleft - sh.nright;
        const long long j = cj / sh.ncent, c = cj - j * sh.ncent;
        typename F::State s = f.zero();
        for (int p = 0; p < sh.nr; ++p) f.push(s, nsites - sh.nr + p, right[j * sh.nr + p]);
        long long rem = c;
        for (int x = 0; x < sh.ncent_sites; ++x) { f.push(s, sh.nl + x, (int)(rem % sh.cdims[x]) + 1); rem /= sh.cdims[x]; }
        Cs[cj] = s;
        Off[cj] = c * sh.stride_c + j * sh.stride_j;
    }
}

__device__ __forceinline__ void maxabs_commit(double v, bool isn, unsigned long long* maxbits, int* nanflag) {
    // block-level: warp shuffle max, then one atomic per warp (non-negative doubles order like their bit patterns)
    unsigned long long b = (unsigned long long)__double_as_longlong(v);
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        unsigned long long o = __shfl_down_sync(0xffffffffu, b, d);
        b = o > b ? o : b;
    }
    const unsigned anyn = __ballot_sync(0xffffffffu, isn);
    if ((threadIdx.x & 31) == 0) {
        if (b) atomicMax(maxbits, b);
        if (anyn) atomicOr(nanflag, 1);
    }
}

template <class F>
__global__ void __launch_bounds__(256)
assemble_kernel(F f, EvalShape sh, int nsites, const typename F::State* __restrict__ Ls,
                const typename F::State* __restrict__ Rs,
                unsigned long long* maxbits, int* nanflag) {
    // grid: x over i (fastest), y over (c, j) pairs in chunks.  The (c, j) decode and the centre-site pushes are done once
    // per pair and amortised over ROWS_PER_THREAD-ish rows (the host sizes gridDim.x for that); 32-bit index arithmetic
    // whenever the pair count allows (a 64-bit division costs ~100 instructions, more than the functor itself).
    const long long ncj = sh.ncent * sh.nright;
    const bool small = ncj < 0x7fffffffLL && sh.nleft < 0x7fffffffLL;
    const int nleft32 = (int)sh.nleft, ncent32 = (int)sh.ncent;
    double mx = 0.0; bool isn = false;
    for (long long cj = blockIdx.y; cj < ncj; cj += gridDim.y) {
        long long c, j;
        if (small) { const unsigned q = (unsigned)cj / (unsigned)ncent32; j = q; c = (unsigned)cj - q * (unsigned)ncent32; }
        else { c = cj % sh.ncent; j = cj / sh.ncent; }
        typename F::State cs = Rs[j];
        unsigned rem = (unsigned)c;                               // ncent = prod of <= 8 local dims: fits 32 bits when `small`
        long long rem64 = c;
        for (int x = 0; x < sh.ncent_sites; ++x) {
            int digit;
            if (small) { const unsigned q = rem / (unsigned)sh.cdims[x]; digit = (int)(rem - q * (unsigned)sh.cdims[x]); rem = q; }
            else { digit = (int)(rem64 % sh.cdims[x]); rem64 /= sh.cdims[x]; }
            f.push(cs, sh.nl + x, digit + 1);
        }
        const long long coff = c * sh.stride_c + j * sh.stride_j;
        double* __restrict__ o0 = sh.outs[0] + coff;
        if (small) {
            const int step = (int)(gridDim.x * blockDim.x);
            for (int i = (int)(blockIdx.x * blockDim.x + threadIdx.x); i < nleft32; i += step) {
                const double v = f.finish(f.merge(Ls[i], cs));
                o0[i] = v;
                for (int p = 1; p < sh.nouts; ++p) sh.outs[p][coff + i] = v;   // peer stores (coalesced 256 B per warp)
                const double a = fabs(v);
                if (a != a) isn = true; else mx = a > mx ? a : mx;
            }
        } else {
            for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < sh.nleft; i += (long long)gridDim.x * blockDim.x) {
                const double v = f.finish(f.merge(Ls[i], cs));
                o0[i] = v;
                for (int p = 1; p < sh.nouts; ++p) sh.outs[p][coff + i] = v;
                const double a = fabs(v);
                if (a != a) isn = true; else mx = a > mx ? a : mx;
            }
        }
    }
    maxabs_commit(mx, isn, maxbits, nanflag);
}

// Vectorised assembly for the bulk case (one destination, even strides, 16-byte aligned output): a thread keeps the left
// states of FOUR consecutive-pair rows (i, i+1) and (i + 2 T, i + 2 T + 1) ... in registers for the whole launch and walks the
// (c, j) pairs of its block, so per element there is one merge, one finish and half a 128-bit store — no index arithmetic,
// no re-load of the row state.  Grid: x over blocks of 4 * 256 rows, y over the pairs (strided).
template <class F>
__global__ void __launch_bounds__(256, 6)
assemble4_kernel(F f, EvalShape sh, const typename F::State* __restrict__ Ls, const typename F::State* __restrict__ Cs,
                 const long long* __restrict__ Off, unsigned long long* maxbits, int* nanflag) {
    typedef typename F::State S;
    const int nleft = (int)sh.nleft;
    const int ncj = (int)(sh.ncent * sh.nright);
    const int ia = blockIdx.x * 1024 + 2 * threadIdx.x, ib = ia + 512;   // rows ia, ia+1 and ib, ib+1 of this block's 1024 rows
    const bool va0 = ia < nleft, va1 = ia + 1 < nleft, vb0 = ib < nleft, vb1 = ib + 1 < nleft;
    S la0 = f.zero(), la1 = f.zero(), lb0 = f.zero(), lb1 = f.zero();
    if (va0) la0 = Ls[ia]; if (va1) la1 = Ls[ia + 1]; if (vb0) lb0 = Ls[ib]; if (vb1) lb1 = Ls[ib + 1];
    unsigned long long mb = 0;                                        // bits of the largest |value| seen; any NaN exceeds the bits of +Inf
    double* __restrict__ out = sh.outs[0];
    int cj = blockIdx.y;
    S csn = cj < ncj ? Cs[cj] : f.zero();
    long long offn = cj < ncj ? Off[cj] : 0;
    for (; cj < ncj; cj += gridDim.y) {
        const S cs = csn;
        double* __restrict__ o = out + offn;
        const int cn = cj + gridDim.y;                                  // the next pair's state and offset: loaded one iteration ahead
        if (cn < ncj) { csn = Cs[cn]; offn = Off[cn]; }
        const double a0 = f.finish(f.merge(la0, cs)), a1 = f.finish(f.merge(la1, cs));
        const double b0 = f.finish(f.merge(lb0, cs)), b1 = f.finish(f.merge(lb1, cs));
        if (va1) *reinterpret_cast<double2*>(o + ia) = make_double2(a0, a1); else if (va0) o[ia] = a0;
        if (vb1) *reinterpret_cast<double2*>(o + ib) = make_double2(b0, b1); else if (vb0) o[ib] = b0;
        const unsigned long long k0 = va0 ? (unsigned long long)__double_as_longlong(a0) & 0x7fffffffffffffffULL : 0ULL;
        const unsigned long long k1 = va1 ? (unsigned long long)__double_as_longlong(a1) & 0x7fffffffffffffffULL : 0ULL;
        const unsigned long long k2 = vb0 ? (unsigned long long)__double_as_longlong(b0) & 0x7fffffffffffffffULL : 0ULL;
        const unsigned long long k3 = vb1 ? (unsigned long long)__double_as_longlong(b1) & 0x7fffffffffffffffULL : 0ULL;
        const unsigned long long ka = k0 > k1 ? k0 : k1, kb = k2 > k3 ? k2 : k3, kc = ka > kb ? ka : kb;
        mb = kc > mb ? kc : mb;
    }
    const bool isn = mb > 0x7ff0000000000000ULL;                      // a NaN was written (Julia's max propagates it, util.jl:34-43)
    maxabs_commit(isn ? 0.0 : __longlong_as_double((long long)mb), isn, maxbits, nanflag);
}

// generic kernel source for NVRTC user functors; the user supplies
//   __device__ double tci_user_f(const int* idx, int n, const double* params, int nparams);
const char* kUserKernelSrc = R"SRC(
extern "C" __global__ void __launch_bounds__(256)
tci_user_assemble(const int* __restrict__ left, const int* __restrict__ right, int nl, int nr, int ncs,
                  int cd0, int cd1, int cd2, int cd3, int cd4, int cd5, int cd6, int cd7,
                  long long nleft, long long nright, long long ncent, long long stride_c, long long stride_j,
                  const double* __restrict__ params, int nparams, double* __restrict__ out,
                  double* o1, double* o2, double* o3, double* o4, double* o5, double* o6, double* o7, int nouts,
                  unsigned long long* maxbits, int* nanflag) {
    double* peers[7] = {o1, o2, o3, o4, o5, o6, o7};
    int cdims[8] = {cd0, cd1, cd2, cd3, cd4, cd5, cd6, cd7};
    int idx[TCI_MAX_SITES];
    const int n = nl + ncs + nr;
    const long long total = nleft * ncent * nright;
    double mx = 0.0; bool isn = false;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        const long long i = t % nleft, cj = t / nleft, c = cj % ncent, j = cj / ncent;
        for (int p = 0; p < nl; ++p) idx[p] = left[i * nl + p];
        long long rem = c;
        for (int x = 0; x < ncs; ++x) { idx[nl + x] = (int)(rem % cdims[x]) + 1; rem /= cdims[x]; }
        for (int p = 0; p < nr; ++p) idx[nl + ncs + p] = right[j * nr + p];
        const double v = tci_user_f(idx, n, params, nparams);
        const long long off = i + c * stride_c + j * stride_j;
        out[off] = v;
        for (int p = 1; p < nouts; ++p) peers[p - 1][off] = v;
        const double a = fabs(v);
        if (a != a) isn = true; else mx = a > mx ? a : mx;
    }
    unsigned long long b = (unsigned long long)__double_as_longlong(mx);
    for (int d = 16; d > 0; d >>= 1) { unsigned long long o = __shfl_down_sync(0xffffffffu, b, d); b = o > b ? o : b; }
    const unsigned anyn = __ballot_sync(0xffffffffu, isn);
    if ((threadIdx.x & 31) == 0) { if (b) atomicMax(maxbits, b); if (anyn) atomicOr(nanflag, 1); }
}
)SRC";

} // namespace

namespace {

struct EvalBuffers { int* d_left; int* d_right; void* Ls; void* Rs; void* Cs; long long* Off; unsigned long long* maxbits; int* nanflag; };

int stage_tables(tci_handle h, const int32_t* h_left, const int32_t* h_right, const EvalShape& sh, EvalBuffers* b, bool zero_acc = true) {
    const size_t nli = (size_t)sh.nleft * sh.nl, nri = (size_t)sh.nright * sh.nr;
    const size_t ints = ((nli + nri) * sizeof(int) + 255) & ~(size_t)255;
    int rc = tci_ws_reserve(h, h->ws_a, ints + 256);
    if (rc) return rc;
    const size_t npair = (size_t)(sh.ncent * sh.nright);
    const size_t st_bytes = 16 * (size_t)(sh.nleft + sh.nright);                      // a State is at most 16 bytes
    const bool pairs = npair <= ((size_t)1 << 22);                                   // per-pair states for the vectorised assembly (<= 96 MB)
    rc = tci_ws_reserve(h, h->ws_b, st_bytes + (pairs ? 24 * npair : 0) + 256);
    if (rc) return rc;
    b->d_left = (int*)h->ws_a.ptr; b->d_right = b->d_left + nli;
    b->maxbits = (unsigned long long*)((char*)h->ws_a.ptr + ints);
    b->nanflag = (int*)(b->maxbits + 1);
    b->Ls = h->ws_b.ptr; b->Rs = (char*)h->ws_b.ptr + 16 * (size_t)sh.nleft;
    b->Cs = pairs ? (char*)h->ws_b.ptr + st_bytes : nullptr;
    b->Off = pairs ? (long long*)((char*)h->ws_b.ptr + st_bytes + 16 * npair) : nullptr;
    if (zero_acc) TCI_CUDA(h, cudaMemsetAsync(b->maxbits, 0, 16, h->stream));
    // The index tables are pageable host memory: cudaMemcpyAsync stages them before it returns, so there is no host
    // synchronisation here and the caller may reuse its arrays immediately.
    if (nli) TCI_CUDA(h, cudaMemcpyAsync(b->d_left, h_left, nli * sizeof(int), cudaMemcpyHostToDevice, h->stream));
    if (nri) TCI_CUDA(h, cudaMemcpyAsync(b->d_right, h_right, nri * sizeof(int), cudaMemcpyHostToDevice, h->stream));
    return TCI_OK;
}

template <class F>
int run_builtin(tci_handle h, const F& f, const EvalShape& sh, int nsites, const EvalBuffers& b) {
    typedef typename F::State S;
    // bulk case: the vectorised kernel (rows in registers, 128-bit streaming stores, per-pair states from the fold)
    const long long ncj = sh.ncent * sh.nright;
    const bool bulk = sh.nouts == 1 && sh.nleft >= 512 && sh.nleft < (1LL << 30) && ncj < (1LL << 30) && ncj >= 64 && (sh.stride_c % 2) == 0 &&
                      (sh.stride_j % 2) == 0 && ((uintptr_t)sh.outs[0] % 16) == 0 && b.Cs != nullptr;
    const long long nfold = sh.nleft + sh.nright + (bulk ? ncj : 0);
    fold_kernel<F><<<(unsigned)((nfold + 255) / 256), 256, 0, h->stream>>>(f, b.d_left, b.d_right, sh, nsites, (S*)b.Ls, (S*)b.Rs,
                                                                           bulk ? (S*)b.Cs : nullptr, bulk ? b.Off : nullptr);
    TCI_LAUNCH_CHECK(h);
    if (bulk) {
        const long long bx = (sh.nleft + 1023) / 1024;
        // exactly one resident wave (a persistent grid): 1.3 waves of a write-bound kernel leave a third of the GPU idle at the end
        static thread_local int occ = 0;
        if (!occ) { if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, assemble4_kernel<F>, 256, 0) != cudaSuccess || occ < 1) occ = 4; }
        long long by = ((long long)h->num_sms * occ) / bx;
        if (by < 1) by = 1;
        if (by > ncj) by = ncj;
        if (by > 65535) by = 65535;
        assemble4_kernel<F><<<dim3((unsigned)bx, (unsigned)by), 256, 0, h->stream>>>(f, sh, (const S*)b.Ls, (const S*)b.Cs, b.Off, b.maxbits, b.nanflag);
        TCI_LAUNCH_CHECK(h);
        return TCI_OK;
    }
    long long gx = (sh.nleft + 256 * 8 - 1) / (256 * 8);        // ~8 rows per thread per (c, j) pair
    if (gx > 32) gx = 32;
    long long gy = sh.ncent * sh.nright;
    const long long want = (long long)h->num_sms * 16 / gx + 1;
    if (gy > want) gy = want;
    if (gy > 65535) gy = 65535;
    assemble_kernel<F><<<dim3((unsigned)gx, (unsigned)gy), 256, 0, h->stream>>>(f, sh, nsites, (const S*)b.Ls, (const S*)b.Rs, b.maxbits, b.nanflag);
    TCI_LAUNCH_CHECK(h);
    return TCI_OK;
}

int run_user(tci_handle h, tci_functor f, const EvalShape& sh, const EvalBuffers& b) {
    double* d_out = sh.outs[0];
    double* po[7]; for (int i = 0; i < 7; ++i) po[i] = (i + 1 < sh.nouts) ? sh.outs[i + 1] : nullptr;
    int nouts = sh.nouts;
    int nl = sh.nl, nr = sh.nr, ncs = sh.ncent_sites, np = f->nparams;
    int cd[8]; for (int i = 0; i < 8; ++i) cd[i] = sh.cdims[i];
    long long nleft = sh.nleft, nright = sh.nright, ncent = sh.ncent, sc = sh.stride_c, sj = sh.stride_j;
    const int* dl = b.d_left; const int* dr = b.d_right; const double* dp = f->d_params;
    unsigned long long* mb = b.maxbits; int* nf = b.nanflag;
    void* args[] = {&dl, &dr, &nl, &nr, &ncs, &cd[0], &cd[1], &cd[2], &cd[3], &cd[4], &cd[5], &cd[6], &cd[7],
                    &nleft, &nright, &ncent, &sc, &sj, &dp, &np, &d_out,
                    &po[0], &po[1], &po[2], &po[3], &po[4], &po[5], &po[6], &nouts, &mb, &nf};
    long long total = nleft * ncent * nright;
    long long blocks = (total + 255) / 256;
    if (blocks > (long long)h->num_sms * 8) blocks = (long long)h->num_sms * 8;
    cudaError_t r = cudaLaunchKernel((const void*)f->fn, dim3((unsigned)blocks), dim3(256), args, 0, h->stream);
    h->launches++;
    if (r != cudaSuccess) return tci_fail(h, TCI_ERR_CUDA, "cudaLaunchKernel(user functor): %s", cudaGetErrorString(r));
    return TCI_OK;
}

} // namespace

extern "C" {

int tci_functor_builtin(tci_handle h, int kind, const double* h_params, int nparams, tci_functor* out) {
    if (!h || !out) return TCI_ERR_INVALID;
    if ((kind < TCI_F_LORENTZ || kind > TCI_F_QEXP2) && kind != TCI_F_HASH) return tci_fail(h, TCI_ERR_INVALID, "tci_functor_builtin: unknown kind %d", kind);
    const int need = (kind == TCI_F_LORENTZ || kind == TCI_F_SUM) ? 1 : (kind == TCI_F_HASH ? 2 : 3);
    if (nparams < need || nparams > 8) return tci_fail(h, TCI_ERR_INVALID, "tci_functor_builtin: kind %d needs %d parameters", kind, need);
    tci_functor f = new tci_functor_s();
    f->kind = kind; f->nparams = nparams;
    for (int i = 0; i < nparams; ++i) f->p[i] = h_params[i];
    *out = f;
    return TCI_OK;
}

int tci_functor_dense(tci_handle h, const double* h_table, const int32_t* h_dims, int ndims, tci_functor* out) {
    if (!h || !out || !h_table || !h_dims || ndims < 1 || ndims > kMaxSites) return tci_fail(h, TCI_ERR_INVALID, "tci_functor_dense: bad arguments");
    TCI_CUDA(h, cudaSetDevice(h->device));
    std::vector<long long> strides(ndims);
    long long tot = 1;
    for (int i = 0; i < ndims; ++i) { strides[i] = tot; tot *= h_dims[i]; }
    tci_functor f = new tci_functor_s();
    f->kind = TCI_F_DENSE; f->ndims = ndims;
    if (cudaMalloc(&f->d_table, sizeof(double) * tot) != cudaSuccess || cudaMalloc(&f->d_strides, sizeof(long long) * ndims) != cudaSuccess) {
        delete f; return tci_fail(h, TCI_ERR_NOMEM, "tci_functor_dense: cudaMalloc failed");
    }
    TCI_CUDA(h, cudaMemcpy(f->d_table, h_table, sizeof(double) * tot, cudaMemcpyHostToDevice));
    TCI_CUDA(h, cudaMemcpy(f->d_strides, strides.data(), sizeof(long long) * ndims, cudaMemcpyHostToDevice));
    *out = f;
    return TCI_OK;
}

int tci_functor_from_source(tci_handle h, const char* cuda_src, const double* h_params, int nparams, tci_functor* out) {
    if (!h || !out || !cuda_src || nparams < 0) return tci_fail(h, TCI_ERR_INVALID, "tci_functor_from_source: bad arguments");
    TCI_CUDA(h, cudaSetDevice(h->device));
    TCI_CUDA(h, cudaFree(0));
    std::string src = std::string(cuda_src) + "\n" + kUserKernelSrc;
    nvrtcProgram prog;
    if (nvrtcCreateProgram(&prog, src.c_str(), "tci_user_functor.cu", 0, nullptr, nullptr) != NVRTC_SUCCESS)
        return tci_fail(h, TCI_ERR_NVRTC, "nvrtcCreateProgram failed");
    const char* opts[] = {"--gpu-architecture=sm_100a", "-DTCI_MAX_SITES=128", "--fmad=false", "-std=c++17", "-lineinfo"};
    nvrtcResult cr = nvrtcCompileProgram(prog, 5, opts);
    if (cr != NVRTC_SUCCESS) {
        size_t ls = 0; nvrtcGetProgramLogSize(prog, &ls);
        std::string log(ls, '\0'); nvrtcGetProgramLog(prog, &log[0]);
        nvrtcDestroyProgram(&prog);
        return tci_fail(h, TCI_ERR_NVRTC, "user functor failed to compile: %s", log.c_str());
    }
    size_t cs = 0; nvrtcGetCUBINSize(prog, &cs);
    std::vector<char> cubin(cs); nvrtcGetCUBIN(prog, cubin.data());
    nvrtcDestroyProgram(&prog);
    tci_functor f = new tci_functor_s();
    f->kind = TCI_F_USER; f->nparams = nparams;
    cudaError_t r = cudaLibraryLoadData(&f->mod, cubin.data(), nullptr, nullptr, 0, nullptr, nullptr, 0);
    if (r == cudaSuccess) r = cudaLibraryGetKernel(&f->fn, f->mod, "tci_user_assemble");
    if (r != cudaSuccess) { delete f; return tci_fail(h, TCI_ERR_CUDA, "loading user functor: %s", cudaGetErrorString(r)); }
    if (cudaMalloc(&f->d_params, sizeof(double) * (nparams > 0 ? nparams : 1)) != cudaSuccess) { cudaLibraryUnload(f->mod); delete f; return tci_fail(h, TCI_ERR_NOMEM, "cudaMalloc failed"); }
    if (nparams > 0) TCI_CUDA(h, cudaMemcpy(f->d_params, h_params, sizeof(double) * nparams, cudaMemcpyHostToDevice));
    *out = f;
    return TCI_OK;
}

int tci_functor_destroy(tci_handle h, tci_functor f) {
    if (!f) return TCI_OK;
    if (h) { cudaSetDevice(h->device); cudaStreamSynchronize(h->stream); }
    if (f->d_table) cudaFree(f->d_table);
    if (f->d_strides) cudaFree(f->d_strides);
    if (f->d_params) cudaFree(f->d_params);
    if (f->mod) cudaLibraryUnload(f->mod);
    delete f;
    return TCI_OK;
}

} // extern "C"

// Shared body of the batch-evaluation entry points: validates, stages the index tables and enqueues the kernels; no host
// synchronisation.  The running max |value| / NaN flag go to `d_acc` (16 bytes: u64 bits of the maximum, int flag) WITHOUT
// resetting it when the caller owns an accumulator, else to a scratch accumulator that is zeroed first; *d_acc_used tells
// which (NULL when nothing was launched because a set is empty).
static int batcheval_enqueue(tci_handle h, tci_functor f, const int32_t* h_localdims, int nsites,
                             const int32_t* h_left, int64_t nleft, int nl, const int32_t* h_right, int64_t nright, int nr,
                             double* const* h_d_outs, int nouts, int64_t stride_c, int64_t stride_j, void* d_acc, void** d_acc_used) {
    if (d_acc_used) *d_acc_used = nullptr;
    if (!h) return TCI_ERR_INVALID;
    if (nouts < 1 || nouts > 8 || !h_d_outs) return tci_fail(h, TCI_ERR_INVALID, "tci_batcheval_multi_f64: 1..8 destination buffers");
    double* d_out = h_d_outs[0];
    for (int p = 0; p < nouts; ++p) if (!h_d_outs[p] && nleft > 0 && nright > 0) return tci_fail(h, TCI_ERR_INVALID, "tci_batcheval_f64: NULL buffer");
    if (!f || !h_localdims) return tci_fail(h, TCI_ERR_INVALID, "tci_batcheval_f64: NULL functor or localdims");
    if (nsites < 1 || nsites > kMaxSites) return tci_fail(h, TCI_ERR_INVALID, "tci_batcheval_f64: nsites must be in 1..%d", kMaxSites);
    const int ncs = nsites - nl - nr;
    if (nl < 0 || nr < 0 || ncs < 0) return tci_fail(h, TCI_ERR_INVALID, "Invalid number of central indices");    // tensorci2.jl:492
    if (ncs > 8) return tci_fail(h, TCI_ERR_INVALID, "tci_batcheval_f64: at most 8 free centre sites");
    if (nleft < 0 || nright < 0) return tci_fail(h, TCI_ERR_INVALID, "tci_batcheval_f64: negative set size");
    if (nleft == 0 || nright == 0) return TCI_OK;                               // batcheval.jl:140-142: empty result
    if (f->kind == TCI_F_DENSE && f->ndims != nsites) return tci_fail(h, TCI_ERR_INVALID, "dense functor has %d dims, localdims has %d", f->ndims, nsites);
    if ((nl > 0 && !h_left) || (nr > 0 && !h_right) || !d_out) return tci_fail(h, TCI_ERR_INVALID, "tci_batcheval_f64: NULL buffer");
    TCI_CUDA(h, cudaSetDevice(h->device));
    EvalShape sh;
    sh.nl = nl; sh.nr = nr; sh.ncent_sites = ncs; sh.nleft = nleft; sh.nright = nright; sh.ncent = 1;
    for (int x = 0; x < 8; ++x) sh.cdims[x] = 1;
    for (int x = 0; x < ncs; ++x) { sh.cdims[x] = h_localdims[nl + x]; sh.ncent *= h_localdims[nl + x]; }
    sh.stride_c = stride_c > 0 ? stride_c : nleft;
    sh.stride_j = stride_j > 0 ? stride_j : sh.stride_c * sh.ncent;
    sh.nouts = nouts;
    for (int p = 0; p < 8; ++p) sh.outs[p] = p < nouts ? h_d_outs[p] : nullptr;
    EvalBuffers b;
    int rc = stage_tables(h, h_left, h_right, sh, &b, d_acc == nullptr);
    if (rc) return rc;
    if (d_acc) { b.maxbits = (unsigned long long*)d_acc; b.nanflag = (int*)(b.maxbits + 1); }
    switch (f->kind) {
    case TCI_F_LORENTZ: rc = f->p[0] == 1.0 ? run_builtin(h, LorentzUnitF{}, sh, nsites, b) : run_builtin(h, LorentzF{f->p[0]}, sh, nsites, b); break;
    case TCI_F_SUM: rc = run_builtin(h, SumF{f->p[0]}, sh, nsites, b); break;
    case TCI_F_QEXPSIN: rc = run_builtin(h, QuanticsF<TCI_F_QEXPSIN>{f->p[0], f->p[1], f->p[2]}, sh, nsites, b); break;
    case TCI_F_QEXPPLUSSIN: rc = run_builtin(h, QuanticsF<TCI_F_QEXPPLUSSIN>{f->p[0], f->p[1], f->p[2]}, sh, nsites, b); break;
    case TCI_F_QEXP2: rc = run_builtin(h, QuanticsF<TCI_F_QEXP2>{f->p[0], f->p[1], f->p[2]}, sh, nsites, b); break;
    case TCI_F_HASH: rc = run_builtin(h, HashF{f->p[0], f->p[1]}, sh, nsites, b); break;
    case TCI_F_DENSE: rc = run_builtin(h, DenseF{f->d_table, f->d_strides}, sh, nsites, b); break;
    case TCI_F_USER: rc = run_user(h, f, sh, b); break;
    default: rc = tci_fail(h, TCI_ERR_INVALID, "unknown functor kind %d", f->kind);
    }
    if (rc) return rc;
    if (d_acc_used) *d_acc_used = b.maxbits;
    return TCI_OK;
}

// 16 accumulator bytes -> max(|prev|, maximum), NaN-propagating (util.jl:34-43); synchronises the stream.
static int read_acc(tci_handle h, const void* d_acc, double prev_in, double* out) {
    int rc = tci_pinned_reserve(h, 64);
    if (rc) return rc;
    unsigned long long* pin = (unsigned long long*)h->pinned;
    TCI_CUDA(h, cudaMemcpyAsync(pin, d_acc, 16, cudaMemcpyDeviceToHost, h->stream));
    TCI_CUDA(h, cudaStreamSynchronize(h->stream));
    double mx; memcpy(&mx, &pin[0], 8);
    int nf; memcpy(&nf, &pin[1], 4);
    const double prev = fabs(prev_in);
    if (nf || prev != prev) *out = NAN; else *out = prev > mx ? prev : mx;
    return TCI_OK;
}

extern "C" {
int tci_batcheval_multi_f64(tci_handle h, tci_functor f, const int32_t* h_localdims, int nsites,
                            const int32_t* h_left, int64_t nleft, int nl, const int32_t* h_right, int64_t nright, int nr,
                            double* const* h_d_outs, int nouts, int64_t stride_c, int64_t stride_j, double* h_maxabs) {
    void* acc = nullptr;
    int rc = batcheval_enqueue(h, f, h_localdims, nsites, h_left, nleft, nl, h_right, nright, nr, h_d_outs, nouts, stride_c, stride_j,
                               nullptr, &acc);
    if (rc) return rc;
    if (h_maxabs && acc) return read_acc(h, acc, *h_maxabs, h_maxabs);
    return TCI_OK;
}

int tci_batcheval_f64(tci_handle h, tci_functor f, const int32_t* h_localdims, int nsites,
                      const int32_t* h_left, int64_t nleft, int nl, const int32_t* h_right, int64_t nright, int nr,
                      double* d_out, int64_t stride_c, int64_t stride_j, double* h_maxabs) {
    double* outs[1] = {d_out};
    return tci_batcheval_multi_f64(h, f, h_localdims, nsites, h_left, nleft, nl, h_right, nright, nr, outs, 1, stride_c, stride_j, h_maxabs);
}

int tci_batcheval_acc_f64(tci_handle h, tci_functor f, const int32_t* h_localdims, int nsites,
                          const int32_t* h_left, int64_t nleft, int nl, const int32_t* h_right, int64_t nright, int nr,
                          double* d_out, int64_t stride_c, int64_t stride_j, void* d_acc16) {
    if (h && !d_acc16) return tci_fail(h, TCI_ERR_INVALID, "tci_batcheval_acc_f64: NULL accumulator");
    double* outs[1] = {d_out};
    return batcheval_enqueue(h, f, h_localdims, nsites, h_left, nleft, nl, h_right, nright, nr, outs, 1, stride_c, stride_j, d_acc16, nullptr);
}

int tci_maxacc_set(tci_handle h, void* d_acc16, double value) {
    if (!h) return TCI_ERR_INVALID;
    if (!d_acc16) return tci_fail(h, TCI_ERR_INVALID, "tci_maxacc_set: NULL accumulator");
    TCI_CUDA(h, cudaSetDevice(h->device));
    unsigned long long w[2] = {0ull, 0ull};
    const double a = fabs(value);
    if (a != a) { const int one = 1; memcpy(&w[1], &one, 4); } else memcpy(&w[0], &a, 8);
    // pageable source: cudaMemcpyAsync stages it before returning (same as the index tables)
    TCI_CUDA(h, cudaMemcpyAsync(d_acc16, w, 16, cudaMemcpyHostToDevice, h->stream));
    return TCI_OK;
}

int tci_maxacc_get(tci_handle h, const void* d_acc16, double* h_value) {
    if (!h) return TCI_ERR_INVALID;
    if (!d_acc16 || !h_value) return tci_fail(h, TCI_ERR_INVALID, "tci_maxacc_get: NULL argument");
    TCI_CUDA(h, cudaSetDevice(h->device));
    return read_acc(h, d_acc16, 0.0, h_value);
}

// ---- peer-mapped buffers for the sharded batch (CUDA IPC; one process per GPU on one NVSwitch box) --------------------
int tci_ipc_export(tci_handle h, void* d_ptr, void* h_handle64) {
    if (!h || !d_ptr || !h_handle64) return TCI_ERR_INVALID;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    TCI_CUDA(h, cudaSetDevice(h->device));
    cudaIpcMemHandle_t mh;
    TCI_CUDA(h, cudaIpcGetMemHandle(&mh, d_ptr));
    memcpy(h_handle64, &mh, 64);
    return TCI_OK;
}

int tci_ipc_open(tci_handle h, const void* h_handle64, void** d_ptr) {
    if (!h || !d_ptr || !h_handle64) return TCI_ERR_INVALID;
    TCI_CUDA(h, cudaSetDevice(h->device));
    cudaIpcMemHandle_t mh;
    memcpy(&mh, h_handle64, 64);
    TCI_CUDA(h, cudaIpcOpenMemHandle(d_ptr, mh, cudaIpcMemLazyEnablePeerAccess));
    return TCI_OK;
}

int tci_ipc_close(tci_handle h, void* d_ptr) {
    if (!h || !d_ptr) return TCI_ERR_INVALID;
    TCI_CUDA(h, cudaSetDevice(h->device));
    TCI_CUDA(h, cudaIpcCloseMemHandle(d_ptr));
    return TCI_OK;
}

} // extern "C"

// ---- allocation-free variant of the batch evaluation --------------------------------------------------------------------------
// No counterpart in the reference: added for callers that cannot call cudaMalloc (code
// under CUDA graph capture, tight serving loops, harnesses that forbid it).  tci_batcheval_acc_f64's scratch — the uploaded
// index tables and the folded per-row / per-pair functor states, normally carved from the handle's growing cudaMalloc'ed
// workspaces — lives here in ONE file-scope device array of kNoAllocScratchBytes that every GPU loading the library holds until
// exit.  Host code must not pass the symbol itself to a kernel (it is a host-side stand-in, not a device address): the
// address comes from cudaGetSymbolAddress, once per device, with that device current.
constexpr size_t kNoAllocScratchBytes = TCI_BATCHEVAL_NOALLOC_SCRATCH_BYTES;   // 4 MiB (include/tci_b200.h)
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

static int noalloc_scratch_address(tci_handle h, unsigned char** out) {      // h->device is current (the caller set it)
    static unsigned char* addr[64] = {};
    static std::mutex mu;
    if (h->device < 0 || h->device >= 64) return tci_fail(h, TCI_ERR_INVALID, "tci_batcheval_noalloc_f64: device ordinal %d out of range", h->device);
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[h->device]) {
        unsigned char* ptr = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch);
        if (e != cudaSuccess || !ptr)
            return tci_fail(h, TCI_ERR_CUDA, "tci_batcheval_noalloc_f64: cudaGetSymbolAddress(scratch) on device %d: %s", h->device, cudaGetErrorString(e));
        addr[h->device] = ptr;
    }
    *out = addr[h->device];
    return TCI_OK;
}

extern "C" {

int tci_batcheval_noalloc_f64(tci_handle h, tci_functor f, const int32_t* h_localdims, int nsites,
                              const int32_t* h_left, int64_t nleft, int nl, const int32_t* h_right, int64_t nright, int nr,
                              double* d_out, int64_t stride_c, int64_t stride_j, void* d_acc16) {
    const char* who = "tci_batcheval_noalloc_f64";
    if (!h) return TCI_ERR_INVALID;
    if (!f || !h_localdims) return tci_fail(h, TCI_ERR_INVALID, "%s: NULL functor or localdims", who);
    if (!d_acc16) return tci_fail(h, TCI_ERR_INVALID, "%s: NULL accumulator", who);
    if (nsites < 1 || nsites > kMaxSites) return tci_fail(h, TCI_ERR_INVALID, "%s: nsites must be in 1..%d", who, kMaxSites);
    const int ncs = nsites - nl - nr;
    if (nl < 0 || nr < 0 || ncs < 0) return tci_fail(h, TCI_ERR_INVALID, "Invalid number of central indices");    // tensorci2.jl:492
    if (ncs > 8) return tci_fail(h, TCI_ERR_INVALID, "%s: at most 8 free centre sites", who);
    if (nleft < 0 || nright < 0) return tci_fail(h, TCI_ERR_INVALID, "%s: negative set size", who);
    if (nleft == 0 || nright == 0) return TCI_OK;                               // batcheval.jl:140-142: empty result
    if (f->kind == TCI_F_DENSE && f->ndims != nsites) return tci_fail(h, TCI_ERR_INVALID, "dense functor has %d dims, localdims has %d", f->ndims, nsites);
    if ((nl > 0 && !h_left) || (nr > 0 && !h_right) || !d_out) return tci_fail(h, TCI_ERR_INVALID, "%s: NULL buffer", who);
    TCI_CUDA(h, cudaSetDevice(h->device));
    EvalShape sh;
    sh.nl = nl; sh.nr = nr; sh.ncent_sites = ncs; sh.nleft = nleft; sh.nright = nright; sh.ncent = 1;
    for (int x = 0; x < 8; ++x) sh.cdims[x] = 1;
    for (int x = 0; x < ncs; ++x) { sh.cdims[x] = h_localdims[nl + x]; sh.ncent *= h_localdims[nl + x]; }
    sh.stride_c = stride_c > 0 ? stride_c : nleft;
    sh.stride_j = stride_j > 0 ? stride_j : sh.stride_c * sh.ncent;
    sh.nouts = 1;
    for (int p = 0; p < 8; ++p) sh.outs[p] = nullptr;
    sh.outs[0] = d_out;
    // carve the static array: [index tables | left + right states | per-pair states + offsets (only if they fit)]
    auto al = [](double x) { return (double)(((unsigned long long)x + 255ull) & ~255ull); };    // sizes are checked in double: no overflow
    const double nli = (double)nleft * nl, nri = (double)nright * nr;
    const double ints = al((nli + nri) * sizeof(int)), st_bytes = al(16.0 * ((double)nleft + (double)nright));
    if ((double)nleft > 1e12 || (double)nright > 1e12 || ints + st_bytes > (double)kNoAllocScratchBytes)
        return tci_fail(h, TCI_ERR_INVALID, "%s: %lld + %lld index rows need %.0f bytes of scratch, the static array holds %zu "
                        "(use tci_batcheval_acc_f64, which allocates)", who, (long long)nleft, (long long)nright, ints + st_bytes, kNoAllocScratchBytes);
    const double npair = (double)sh.ncent * (double)nright;
    const bool pairs = ints + st_bytes + 24.0 * npair <= (double)kNoAllocScratchBytes;   // else the kernel that folds the centre digits itself
    unsigned char* base = nullptr;
    int rc = noalloc_scratch_address(h, &base);
    if (rc) return rc;
    EvalBuffers b;
    b.d_left = (int*)base; b.d_right = b.d_left + (size_t)nli;
    b.Ls = base + (size_t)ints; b.Rs = (char*)b.Ls + 16 * (size_t)nleft;
    b.Cs = pairs ? base + (size_t)ints + (size_t)st_bytes : nullptr;
    b.Off = pairs ? (long long*)((char*)b.Cs + 16 * (size_t)npair) : nullptr;
    b.maxbits = (unsigned long long*)d_acc16; b.nanflag = (int*)(b.maxbits + 1);
    // Nothing in the array has to start at zero: the tables are uploaded and every state / offset is written by the fold kernel
    // before the assembly reads it; the running maximum lives in the CALLER's accumulator.  Pageable sources are staged by
    // cudaMemcpyAsync before it returns.
    if (nli > 0) TCI_CUDA(h, cudaMemcpyAsync(b.d_left, h_left, (size_t)nli * sizeof(int), cudaMemcpyHostToDevice, h->stream));
    if (nri > 0) TCI_CUDA(h, cudaMemcpyAsync(b.d_right, h_right, (size_t)nri * sizeof(int), cudaMemcpyHostToDevice, h->stream));
    switch (f->kind) {
    case TCI_F_LORENTZ: return f->p[0] == 1.0 ? run_builtin(h, LorentzUnitF{}, sh, nsites, b) : run_builtin(h, LorentzF{f->p[0]}, sh, nsites, b);
    case TCI_F_SUM: return run_builtin(h, SumF{f->p[0]}, sh, nsites, b);
    case TCI_F_QEXPSIN: return run_builtin(h, QuanticsF<TCI_F_QEXPSIN>{f->p[0], f->p[1], f->p[2]}, sh, nsites, b);
    case TCI_F_QEXPPLUSSIN: return run_builtin(h, QuanticsF<TCI_F_QEXPPLUSSIN>{f->p[0], f->p[1], f->p[2]}, sh, nsites, b);
    case TCI_F_QEXP2: return run_builtin(h, QuanticsF<TCI_F_QEXP2>{f->p[0], f->p[1], f->p[2]}, sh, nsites, b);
    case TCI_F_HASH: return run_builtin(h, HashF{f->p[0], f->p[1]}, sh, nsites, b);
    case TCI_F_DENSE: return run_builtin(h, DenseF{f->d_table, f->d_strides}, sh, nsites, b);
    case TCI_F_USER: return run_user(h, f, sh, b);
    default: return tci_fail(h, TCI_ERR_INVALID, "unknown functor kind %d", f->kind);
    }
}

} // extern "C"
