// tensor.cu — K4c support: strided tensor copies (permutedims / slicing), element-wise maps, permutation scatters.
//
// Replaces the `permutedims` + slicing that surround every GEMM of the reference's contraction code
// (src/contraction.jl:211-212 in _contract, :256-258 slices in _extend_cache, :541-556 in batchevaluate, :595 in
// _contractsitetensors, :770 in contract_zipup) and the row/column un-permutation of left(lu)/right(lu)
// (src/matrixlu.jl:716-746).  All of it is HBM-bound data movement: 8 B read + 8 B written per element; writes are
// fully coalesced (the output is dense column-major), reads follow the input strides.
#include "context.h"
#include <cmath>
#include <cstring>

namespace {

constexpr int kMaxDims = 8;

struct CopyDesc {
    int ndim;
    long long shape[kMaxDims];       // output shape, first index fastest
    long long istride[kMaxDims];     // input element stride of each OUTPUT dimension
    long long total;
};

__global__ void __launch_bounds__(256)
strided_copy_kernel(const double* __restrict__ in, double* __restrict__ out, CopyDesc d) {
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < d.total; t += (long long)gridDim.x * blockDim.x) {
        long long rem = t, off = 0;
#pragma unroll
        for (int k = 0; k < kMaxDims; ++k) {
            if (k < d.ndim) { const long long s = d.shape[k]; const long long q = rem / s; off += (rem - q * s) * d.istride[k]; rem = q; }
        }
        out[t] = in[off];
    }
}

// 2-D tiled transpose through shared memory for the case that dominates in practice: the fastest OUTPUT dimension is
// strided in the input while some other dimension is contiguous in the input.  Tile: 32 x 32 doubles, padded.
// View: out[i + I*(j + J*rest)] = in[i*si + j*1 + off_rest], i.e. input contiguous along j.
struct TransDesc {
    long long I, J, si;              // sizes of the two swapped dimensions, input stride of i (input stride of j is 1)
    long long oj;                    // output stride of j (elements)
    int nrest;
    long long rshape[kMaxDims], ristride[kMaxDims], rostride[kMaxDims];
    long long nbatch;
};

__global__ void __launch_bounds__(256)
tiled_transpose_kernel(const double* __restrict__ in, double* __restrict__ out, TransDesc d) {
    __shared__ double tile[32][33];
    const long long ti = (d.I + 31) / 32, tj = (d.J + 31) / 32;
    const long long ntiles = ti * tj * d.nbatch;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;      // 32 x 8
    for (long long t = blockIdx.x; t < ntiles; t += gridDim.x) {
        const long long bi = t % ti, bj = (t / ti) % tj;
        long long rem = t / (ti * tj), ioff = 0, ooff = 0;
        for (int k = 0; k < d.nrest; ++k) { const long long q = rem / d.rshape[k]; const long long r = rem - q * d.rshape[k]; ioff += r * d.ristride[k]; ooff += r * d.rostride[k]; rem = q; }
        // read: threads along j (contiguous in the input)
#pragma unroll
        for (int r = 0; r < 32; r += 8) {
            const long long i = bi * 32 + ty + r, j = bj * 32 + tx;
            if (i < d.I && j < d.J) tile[ty + r][tx] = in[ioff + i * d.si + j];
        }
        __syncthreads();
        // write: threads along i (contiguous in the output)
#pragma unroll
        for (int r = 0; r < 32; r += 8) {
            const long long i = bi * 32 + tx, j = bj * 32 + ty + r;
            if (i < d.I && j < d.J) out[ooff + i + j * d.oj] = tile[tx][ty + r];
        }
        __syncthreads();
    }
}

__global__ void elementwise_kernel(double* __restrict__ x, long long n, int kind, double p0, double p1) {
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) {
        const double v = x[t];
        double r;
        switch (kind) {
        case 1: r = p0 * v + p1; break;          // affine
        case 2: r = v * v; break;                // abs2
        case 3: r = fabs(v); break;              // abs
        case 4: r = exp(p0 * v); break;
        case 5: r = v / p0; break;               // true division (divide!, tensortrain.jl:412-415)
        default: r = v;
        }
        x[t] = r;
    }
}

// max |x| over a (rows x cols, ld) matrix; non-negative doubles order like their bit patterns; NaN sets the flag
__global__ void maxabs_kernel(const double* __restrict__ x, long long rows, long long cols, long long ld,
                              unsigned long long* maxbits, int* nanflag) {
    const long long total = rows * cols;
    double mx = 0.0; bool isn = false;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        const double a = fabs(x[(t / rows) * ld + (t % rows)]);
        if (a != a) isn = true; else mx = a > mx ? a : mx;
    }
    unsigned long long b = (unsigned long long)__double_as_longlong(mx);
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) { unsigned long long o = __shfl_down_sync(0xffffffffu, b, d); b = o > b ? o : b; }
    const unsigned anyn = __ballot_sync(0xffffffffu, isn);
    if ((threadIdx.x & 31) == 0) { if (b) atomicMax(maxbits, b); if (anyn) atomicOr(nanflag, 1); }
}

// out[perm[i], j] = in[i, j]  (rows) / out[i, perm[j]] = in[i, j]  (cols); perm 0-based on the device
__global__ void unpermute_kernel(const double* __restrict__ in, long long ldi, double* __restrict__ out, long long ldo,
                                 const int* __restrict__ perm, int rows, int cols, int by_rows) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows) return;
    for (int j = blockIdx.y; j < cols; j += gridDim.y) {       // grid.y is capped at 65535: wide matrices loop
        const double v = in[(long long)j * ldi + i];
        if (by_rows) out[(long long)j * ldo + perm[i]] = v; else out[(long long)perm[j] * ldo + i] = v;
    }
}

} // namespace

extern "C" {

int tci_strided_copy_f64(tci_handle h, const double* d_in, int64_t in_offset, const int64_t* h_shape, const int64_t* h_in_strides,
                         int ndim, double* d_out) {
    if (!h) return TCI_ERR_INVALID;
    if (ndim < 0 || ndim > kMaxDims || (ndim > 0 && (!h_shape || !h_in_strides))) return tci_fail(h, TCI_ERR_INVALID, "tci_strided_copy_f64: ndim must be in 0..%d", kMaxDims);
    long long total = 1;
    for (int k = 0; k < ndim; ++k) { if (h_shape[k] < 0) return tci_fail(h, TCI_ERR_INVALID, "tci_strided_copy_f64: negative extent"); total *= h_shape[k]; }
    if (total == 0) return TCI_OK;
    if (!d_in || !d_out) return tci_fail(h, TCI_ERR_INVALID, "tci_strided_copy_f64: NULL buffer");
    TCI_CUDA(h, cudaSetDevice(h->device));
    // drop extent-1 dimensions and merge dimensions that are contiguous in the input view
    long long shp[kMaxDims], str[kMaxDims]; int nd = 0;
    for (int k = 0; k < ndim; ++k) {
        if (h_shape[k] == 1) continue;
        if (nd > 0 && str[nd - 1] * shp[nd - 1] == h_in_strides[k]) shp[nd - 1] *= h_shape[k];
        else { shp[nd] = h_shape[k]; str[nd] = h_in_strides[k]; ++nd; }
    }
    const double* src = d_in + in_offset;
    if (nd == 0) { TCI_CUDA(h, cudaMemcpyAsync(d_out, src, sizeof(double), cudaMemcpyDeviceToDevice, h->stream)); return TCI_OK; }
    if (nd == 1 && str[0] == 1) { TCI_CUDA(h, cudaMemcpyAsync(d_out, src, sizeof(double) * total, cudaMemcpyDeviceToDevice, h->stream)); return TCI_OK; }
    // tiled transpose when dim 0 is strided and another dimension is unit-stride in the input
    int jdim = -1;
    if (str[0] != 1) for (int k = 1; k < nd; ++k) if (str[k] == 1 && shp[k] >= 8) { jdim = k; break; }
    if (jdim > 0 && shp[0] >= 8) {
        TransDesc d; d.I = shp[0]; d.si = str[0]; d.J = shp[jdim]; d.nrest = 0; d.nbatch = 1;
        long long ostr = 1;
        for (int k = 0; k < nd; ++k) {
            if (k == jdim) d.oj = ostr;
            else if (k != 0) { d.rshape[d.nrest] = shp[k]; d.ristride[d.nrest] = str[k]; d.rostride[d.nrest] = ostr; d.nbatch *= shp[k]; ++d.nrest; }
            ostr *= shp[k];
        }
        long long tiles = ((d.I + 31) / 32) * ((d.J + 31) / 32) * d.nbatch;
        long long blocks = tiles < (long long)h->num_sms * 16 ? tiles : (long long)h->num_sms * 16;
        tiled_transpose_kernel<<<(unsigned)blocks, 256, 0, h->stream>>>(src, d_out, d);
        TCI_LAUNCH_CHECK(h);
        return TCI_OK;
    }
    CopyDesc d; d.ndim = nd; d.total = total;
    for (int k = 0; k < kMaxDims; ++k) { d.shape[k] = k < nd ? shp[k] : 1; d.istride[k] = k < nd ? str[k] : 0; }
    long long blocks = (total + 255) / 256;
    if (blocks > (long long)h->num_sms * 16) blocks = (long long)h->num_sms * 16;
    strided_copy_kernel<<<(unsigned)blocks, 256, 0, h->stream>>>(src, d_out, d);
    TCI_LAUNCH_CHECK(h);
    return TCI_OK;
}

int tci_elementwise_f64(tci_handle h, int kind, double p0, double p1, double* d_x, int64_t n) {
    if (!h) return TCI_ERR_INVALID;
    if (kind < 1 || kind > 5) return tci_fail(h, TCI_ERR_INVALID, "tci_elementwise_f64: unknown kind %d", kind);
    if (n <= 0) return TCI_OK;
    if (!d_x) return tci_fail(h, TCI_ERR_INVALID, "tci_elementwise_f64: NULL buffer");
    TCI_CUDA(h, cudaSetDevice(h->device));
    long long blocks = (n + 255) / 256;
    if (blocks > (long long)h->num_sms * 16) blocks = (long long)h->num_sms * 16;
    elementwise_kernel<<<(unsigned)blocks, 256, 0, h->stream>>>(d_x, n, kind, p0, p1);
    TCI_LAUNCH_CHECK(h);
    return TCI_OK;
}

int tci_maxabs_f64(tci_handle h, const double* d_x, int64_t rows, int64_t cols, int64_t ld, double* h_maxabs) {
    if (!h) return TCI_ERR_INVALID;
    if (!h_maxabs || rows < 0 || cols < 0 || ld < rows) return tci_fail(h, TCI_ERR_INVALID, "tci_maxabs_f64: bad arguments");
    if (rows == 0 || cols == 0) return TCI_OK;
    if (!d_x) return tci_fail(h, TCI_ERR_INVALID, "tci_maxabs_f64: NULL buffer");
    TCI_CUDA(h, cudaSetDevice(h->device));
    int rc = tci_ws_reserve(h, h->ws_c, 256); if (rc) return rc;
    rc = tci_pinned_reserve(h, 64); if (rc) return rc;
    unsigned long long* mb = (unsigned long long*)h->ws_c.ptr; int* nf = (int*)(mb + 1);
    TCI_CUDA(h, cudaMemsetAsync(mb, 0, 16, h->stream));
    long long blocks = (rows * cols + 255) / 256;
    if (blocks > (long long)h->num_sms * 8) blocks = (long long)h->num_sms * 8;
    maxabs_kernel<<<(unsigned)blocks, 256, 0, h->stream>>>(d_x, rows, cols, ld, mb, nf);
    TCI_LAUNCH_CHECK(h);
    unsigned long long* pin = (unsigned long long*)h->pinned;
    TCI_CUDA(h, cudaMemcpyAsync(pin, mb, 16, cudaMemcpyDeviceToHost, h->stream));
    TCI_CUDA(h, cudaStreamSynchronize(h->stream));
    double mx; memcpy(&mx, &pin[0], 8);
    int f; memcpy(&f, &pin[1], 4);
    const double prev = fabs(*h_maxabs);
    if (f || prev != prev) *h_maxabs = NAN; else *h_maxabs = prev > mx ? prev : mx;
    return TCI_OK;
}

int tci_unpermute_f64(tci_handle h, const double* d_in, int64_t ldi, double* d_out, int64_t ldo, const int64_t* h_perm,
                      int64_t rows, int64_t cols, int by_rows) {
    if (!h) return TCI_ERR_INVALID;
    if (rows < 0 || cols < 0 || rows > INT32_MAX || cols > INT32_MAX) return tci_fail(h, TCI_ERR_INVALID, "tci_unpermute_f64: bad sizes");
    if (rows == 0 || cols == 0) return TCI_OK;
    if (!d_in || !d_out || !h_perm) return tci_fail(h, TCI_ERR_INVALID, "tci_unpermute_f64: NULL buffer");
    TCI_CUDA(h, cudaSetDevice(h->device));
    const int64_t len = by_rows ? rows : cols;
    int rc = tci_ws_reserve(h, h->ws_c, sizeof(int) * (size_t)len + 256); if (rc) return rc;
    rc = tci_pinned_reserve(h, sizeof(int) * (size_t)len + 64); if (rc) return rc;
    int* pin = (int*)h->pinned;
    for (int64_t i = 0; i < len; ++i) pin[i] = (int)(h_perm[i] - 1);
    TCI_CUDA(h, cudaMemcpyAsync(h->ws_c.ptr, pin, sizeof(int) * (size_t)len, cudaMemcpyHostToDevice, h->stream));
    TCI_CUDA(h, cudaStreamSynchronize(h->stream));
    unpermute_kernel<<<dim3((unsigned)((rows + 255) / 256), (unsigned)(cols > 65535 ? 65535 : cols)), 256, 0, h->stream>>>(d_in, ldi, d_out, ldo, (const int*)h->ws_c.ptr, (int)rows, (int)cols, by_rows);
    TCI_LAUNCH_CHECK(h);
    return TCI_OK;
}

} // extern "C"
