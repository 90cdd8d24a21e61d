// tt.cu — K4b: TensorTrain `sum` and `evaluate` on device-resident cores.
//
// Replaces src/abstracttensortrain.jl:428-441 (sum: reduce each core over its site legs, then a row-vector x
// matrix chain) and :328-342 (evaluate: product of the slices T[:, i, :]).  A core is the column-major array
// (chi_l, d, chi_r) stored as a (chi_l*d) x chi_r matrix with leading dimension ld >= chi_l*d.
//
//   sum:      w[c] = sum_{a,s} v[a] T[a,s,c] — the site reduction and the vector-matrix product are one pass that
//             reads the core exactly once (HBM-bound: bytes = size of the cores), one warp per output column c,
//             lanes along the contiguous (a,s) axis.
//   evaluate: one CTA per point; the running row vector lives in shared memory; per site one warp per output
//             column, lanes along a (coalesced reads of the slice T[:, i, c]).
#include "context.h"
#include <cstring>
#include <vector>

namespace {

struct CoreDesc { const double* p; int chil, d, chir; long long ld; };

__global__ void tt_sum_site_kernel(CoreDesc c, const double* __restrict__ v, double* __restrict__ w, int first) {
    // first != 0: v is implicit ones over the left bond (sum(tt[1], dims=(1,2)) folds the left bond too)
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= c.chir) return;
    const double* col = c.p + (long long)warp * c.ld;
    const int len = c.chil * c.d;
    double s = 0.0;
    for (int e = lane; e < len; e += 32) {
        const double x = col[e];
        s += first ? x : v[e % c.chil] * x;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
    if (lane == 0) w[warp] = s;
}

constexpr int kMaxChi = 4096;       // shared-memory row vector: 2 x kMaxChi doubles = 64 KB
__global__ void __launch_bounds__(256)
tt_evaluate_kernel(const CoreDesc* __restrict__ cores, int nsites, const int* __restrict__ idx, long long npoints,
                   double* __restrict__ out, int maxchi) {
    extern __shared__ double sm[];
    double* v = sm; double* w = sm + maxchi;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
    for (long long p = blockIdx.x; p < npoints; p += gridDim.x) {
        if (threadIdx.x == 0) v[0] = 1.0;
        __syncthreads();
        for (int b = 0; b < nsites; ++b) {
            const CoreDesc c = cores[b];
            const int i = idx[p * nsites + b] - 1;
            for (int col = warp; col < c.chir; col += nwarps) {
                const double* src = c.p + (long long)col * c.ld + (long long)i * c.chil;
                double s = 0.0;
                for (int a = lane; a < c.chil; a += 32) s += v[a] * src[a];
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
                if (lane == 0) w[col] = s;
            }
            __syncthreads();
            double* t = v; v = w; w = t;
            __syncthreads();
        }
        if (threadIdx.x == 0) out[p] = v[0];
        __syncthreads();
    }
}

int stage_cores(tci_handle h, const double* const* h_core_ptrs, const int64_t* h_shapes, const int64_t* h_lds, int nsites,
                std::vector<CoreDesc>* cores, int* maxchi) {
    cores->resize(nsites);
    *maxchi = 1;
    for (int b = 0; b < nsites; ++b) {
        CoreDesc& c = (*cores)[b];
        c.p = h_core_ptrs[b]; c.chil = (int)h_shapes[3 * b]; c.d = (int)h_shapes[3 * b + 1]; c.chir = (int)h_shapes[3 * b + 2];
        c.ld = h_lds ? h_lds[b] : (long long)c.chil * c.d;
        if (c.chil < 1 || c.d < 1 || c.chir < 1 || c.ld < (long long)c.chil * c.d || !c.p)
            return tci_fail(h, TCI_ERR_INVALID, "tensor train core %d has a bad shape", b + 1);
        if (b > 0 && (*cores)[b - 1].chir != c.chil)
            return tci_fail(h, TCI_ERR_INVALID, "The tensors at %d and %d must have consistent dimensions for a tensor train.", b, b + 1);  // tensortrain.jl:70-74
        if (c.chil > *maxchi) *maxchi = c.chil;
        if (c.chir > *maxchi) *maxchi = c.chir;
    }
    return TCI_OK;
}

} // namespace

extern "C" {

int tci_tt_sum_f64(tci_handle h, const double* const* h_core_ptrs, const int64_t* h_shapes, const int64_t* h_lds,
                   int nsites, double* h_out) {
    if (!h) return TCI_ERR_INVALID;
    if (nsites < 1 || !h_core_ptrs || !h_shapes || !h_out) return tci_fail(h, TCI_ERR_INVALID, "tci_tt_sum_f64: bad arguments");
    TCI_CUDA(h, cudaSetDevice(h->device));
    std::vector<CoreDesc> cores; int maxchi;
    int rc = stage_cores(h, h_core_ptrs, h_shapes, h_lds, nsites, &cores, &maxchi);
    if (rc) return rc;
    rc = tci_ws_reserve(h, h->ws_c, sizeof(double) * 2 * (size_t)maxchi + 64);
    if (rc) return rc;
    double* v = (double*)h->ws_c.ptr; double* w = v + maxchi;
    for (int b = 0; b < nsites; ++b) {
        const int threads = 256, warps_per_block = threads / 32;
        const int blocks = (cores[b].chir + warps_per_block - 1) / warps_per_block;
        tt_sum_site_kernel<<<blocks, threads, 0, h->stream>>>(cores[b], v, w, b == 0);
        TCI_LAUNCH_CHECK(h);
        double* t = v; v = w; w = t;
    }
    if (cores[nsites - 1].chir != 1) return tci_fail(h, TCI_ERR_INVALID, "sum: the last bond dimension must be 1");
    rc = tci_pinned_reserve(h, 64);
    if (rc) return rc;
    TCI_CUDA(h, cudaMemcpyAsync(h->pinned, v, sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    TCI_CUDA(h, cudaStreamSynchronize(h->stream));
    memcpy(h_out, h->pinned, sizeof(double));
    return TCI_OK;
}

int tci_tt_evaluate_f64(tci_handle h, const double* const* h_core_ptrs, const int64_t* h_shapes, const int64_t* h_lds,
                        int nsites, const int32_t* h_indices, int64_t npoints, double* h_out) {
    if (!h) return TCI_ERR_INVALID;
    if (nsites < 1 || !h_core_ptrs || !h_shapes || npoints < 0) return tci_fail(h, TCI_ERR_INVALID, "tci_tt_evaluate_f64: bad arguments");
    if (npoints == 0) return TCI_OK;
    if (!h_indices || !h_out) return tci_fail(h, TCI_ERR_INVALID, "tci_tt_evaluate_f64: NULL buffer");
    TCI_CUDA(h, cudaSetDevice(h->device));
    std::vector<CoreDesc> cores; int maxchi;
    int rc = stage_cores(h, h_core_ptrs, h_shapes, h_lds, nsites, &cores, &maxchi);
    if (rc) return rc;
    if (cores[0].chil != 1 || cores[nsites - 1].chir != 1) return tci_fail(h, TCI_ERR_INVALID, "evaluate: boundary bond dimensions must be 1");
    if (maxchi > kMaxChi) return tci_fail(h, TCI_ERR_INVALID, "evaluate: bond dimension %d exceeds %d", maxchi, kMaxChi);
    for (int64_t p = 0; p < npoints; ++p)
        for (int b = 0; b < nsites; ++b) {
            const int v = h_indices[p * nsites + b];
            if (v < 1 || v > cores[b].d) return tci_fail(h, TCI_ERR_INVALID, "evaluate: index %d at site %d is out of range 1..%d", v, b + 1, cores[b].d);
        }
    const size_t cb = (sizeof(CoreDesc) * nsites + 255) & ~(size_t)255, ib = (sizeof(int) * (size_t)npoints * nsites + 255) & ~(size_t)255;
    rc = tci_ws_reserve(h, h->ws_c, cb + ib + sizeof(double) * (size_t)npoints + 256);
    if (rc) return rc;
    rc = tci_pinned_reserve(h, cb + ib + sizeof(double) * (size_t)npoints + 256);
    if (rc) return rc;
    char* pin = (char*)h->pinned; char* dev = (char*)h->ws_c.ptr;
    memcpy(pin, cores.data(), sizeof(CoreDesc) * nsites);
    memcpy(pin + cb, h_indices, sizeof(int) * (size_t)npoints * nsites);
    TCI_CUDA(h, cudaMemcpyAsync(dev, pin, cb + ib, cudaMemcpyHostToDevice, h->stream));
    double* d_out = (double*)(dev + cb + ib);
    const size_t smem = sizeof(double) * 2 * (size_t)maxchi;
    if (smem > 48 * 1024) TCI_CUDA(h, cudaFuncSetAttribute(tt_evaluate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    long long blocks = npoints < (long long)h->num_sms * 4 ? npoints : (long long)h->num_sms * 4;
    tt_evaluate_kernel<<<(unsigned)blocks, 256, smem, h->stream>>>((const CoreDesc*)dev, nsites, (const int*)(dev + cb), npoints, d_out, maxchi);
    TCI_LAUNCH_CHECK(h);
    TCI_CUDA(h, cudaMemcpyAsync(pin + cb + ib, d_out, sizeof(double) * (size_t)npoints, cudaMemcpyDeviceToHost, h->stream));
    TCI_CUDA(h, cudaStreamSynchronize(h->stream));
    memcpy(h_out, pin + cb + ib, sizeof(double) * (size_t)npoints);
    return TCI_OK;
}

} // extern "C"
