// context.cu — handle lifetime, error strings, raw memory helpers of the C ABI (include/tci_b200.h).
#include "context.h"

static thread_local std::string g_create_err;

int tci_fail(tci_handle h, int code, const char* fmt, ...) {
    char buf[2048];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (h) h->err = buf; else g_create_err = buf;
    return code;
}

int tci_ws_reserve(tci_handle h, tci_workspace& ws, size_t bytes) {
    if (ws.bytes >= bytes) return TCI_OK;
    if (ws.ptr) { TCI_CUDA(h, cudaStreamSynchronize(h->stream)); TCI_CUDA(h, cudaFree(ws.ptr)); ws.ptr = nullptr; ws.bytes = 0; }
    size_t want = bytes + bytes / 8 + 256;
    cudaError_t e = cudaMalloc(&ws.ptr, want);
    if (e != cudaSuccess) { ws.ptr = nullptr; return tci_fail(h, TCI_ERR_NOMEM, "cudaMalloc(%zu) failed: %s", want, cudaGetErrorString(e)); }
    ws.bytes = want;
    return TCI_OK;
}

int tci_pinned_reserve(tci_handle h, size_t bytes) {
    if (h->pinned_bytes >= bytes) return TCI_OK;
    if (h->pinned) { TCI_CUDA(h, cudaStreamSynchronize(h->stream)); TCI_CUDA(h, cudaFreeHost(h->pinned)); h->pinned = nullptr; h->pinned_bytes = 0; }
    size_t want = bytes + bytes / 4 + 4096;
    TCI_CUDA(h, cudaMallocHost(&h->pinned, want));
    h->pinned_bytes = want;
    return TCI_OK;
}

extern "C" {

const char* tci_version(void) { return "tci_b200 0.1 (sm_100a)"; }

int tci_create(tci_handle* out, int device) {
    if (!out) return tci_fail(nullptr, TCI_ERR_INVALID, "tci_create: out is NULL");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return tci_fail(nullptr, TCI_ERR_CUDA, "tci_create: no CUDA device (%s); this library has no CPU fallback",
                        e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    if (device < 0 || device >= count) return tci_fail(nullptr, TCI_ERR_INVALID, "tci_create: device %d out of range [0,%d)", device, count);
    tci_context* h = new tci_context();
    h->device = device;
    if ((e = cudaSetDevice(device)) != cudaSuccess) { int c = tci_fail(nullptr, TCI_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e)); delete h; return c; }
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, device);
    h->num_sms = prop.multiProcessorCount;
    if ((e = cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking)) != cudaSuccess) {
        int c = tci_fail(nullptr, TCI_ERR_CUDA, "cudaStreamCreate: %s", cudaGetErrorString(e)); delete h; return c;
    }
    h->stream = h->own_stream;
    for (int i = 0; i < 4; ++i) cudaEventCreate(&h->ev[i]);
    *out = h;
    return TCI_OK;
}

int tci_destroy(tci_handle h) {
    if (!h) return TCI_OK;
    cudaSetDevice(h->device);
    cudaStreamSynchronize(h->stream);
    tci_workspace* wss[] = {&h->ws_matrix, &h->ws_lu, &h->ws_small, &h->ws_a, &h->ws_b, &h->ws_c, &h->dist_parked};
    for (auto* w : wss) if (w->ptr) cudaFree(w->ptr);
    if (h->pinned) cudaFreeHost(h->pinned);
    for (int i = 0; i < 4; ++i) if (h->ev[i]) cudaEventDestroy(h->ev[i]);
    for (auto e : h->ev_pool) cudaEventDestroy(e);
    if (h->own_stream) cudaStreamDestroy(h->own_stream);
    delete h;
    return TCI_OK;
}

const char* tci_last_error(tci_handle h) { return h ? h->err.c_str() : g_create_err.c_str(); }

int tci_set_stream(tci_handle h, void* s) {
    if (!h) return TCI_ERR_INVALID;
    h->stream = (cudaStream_t)s;                          // NULL is CUDA's legacy default stream
    return TCI_OK;
}
int tci_reset_stream(tci_handle h) {
    if (!h) return TCI_ERR_INVALID;
    h->stream = h->own_stream;
    return TCI_OK;
}
int tci_synchronize(tci_handle h) {
    if (!h) return TCI_ERR_INVALID;
    TCI_CUDA(h, cudaStreamSynchronize(h->stream));
    return TCI_OK;
}
int64_t tci_launch_count(tci_handle h) { return h ? h->launches : 0; }

int tci_malloc(tci_handle h, void** d_ptr, size_t bytes) {
    if (!h || !d_ptr) return TCI_ERR_INVALID;
    TCI_CUDA(h, cudaSetDevice(h->device));
    cudaError_t e = cudaMalloc(d_ptr, bytes ? bytes : 1);
    if (e != cudaSuccess) return tci_fail(h, TCI_ERR_NOMEM, "cudaMalloc(%zu): %s", bytes, cudaGetErrorString(e));
    return TCI_OK;
}
int tci_free(tci_handle h, void* d_ptr) {
    if (!h) return TCI_ERR_INVALID;
    TCI_CUDA(h, cudaStreamSynchronize(h->stream));
    TCI_CUDA(h, cudaFree(d_ptr));
    return TCI_OK;
}
/* Stream-ordered allocation from the device's default memory pool (cudaMallocAsync): the native driver's scratch grows over the
 * first sweeps, and cudaMalloc / cudaFree each cost hundreds of microseconds plus a device-wide sync.  The pool keeps what is
 * freed (release threshold = unbounded), so a regrow is a few microseconds and never synchronises. */
int tci_malloc_pooled(tci_handle h, void** d_ptr, size_t bytes) {
    if (!h || !d_ptr) return TCI_ERR_INVALID;
    TCI_CUDA(h, cudaSetDevice(h->device));
    static bool configured[64] = {};
    if (h->device >= 0 && h->device < 64 && !configured[h->device]) {
        cudaMemPool_t pool;
        TCI_CUDA(h, cudaDeviceGetDefaultMemPool(&pool, h->device));
        unsigned long long keep = ~0ull;
        TCI_CUDA(h, cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
        configured[h->device] = true;
    }
    cudaError_t e = cudaMallocAsync(d_ptr, bytes ? bytes : 1, h->stream);
    if (e != cudaSuccess) return tci_fail(h, TCI_ERR_NOMEM, "cudaMallocAsync(%zu): %s", bytes, cudaGetErrorString(e));
    return TCI_OK;
}
int tci_free_pooled(tci_handle h, void* d_ptr) {          /* ordered after the work already enqueued on the handle's stream */
    if (!h) return TCI_ERR_INVALID;
    if (d_ptr) TCI_CUDA(h, cudaFreeAsync(d_ptr, h->stream));
    return TCI_OK;
}
int tci_memcpy_h2d(tci_handle h, void* d_dst, const void* h_src, size_t bytes) {
    if (!h) return TCI_ERR_INVALID;
    TCI_CUDA(h, cudaMemcpyAsync(d_dst, h_src, bytes, cudaMemcpyHostToDevice, h->stream));
    TCI_CUDA(h, cudaStreamSynchronize(h->stream));
    return TCI_OK;
}
int tci_memcpy_d2d(tci_handle h, void* d_dst, const void* d_src, size_t bytes) {          /* enqueued on the handle's stream, no sync */
    if (!h) return TCI_ERR_INVALID;
    TCI_CUDA(h, cudaMemcpyAsync(d_dst, d_src, bytes, cudaMemcpyDeviceToDevice, h->stream));
    return TCI_OK;
}
int tci_memcpy_d2h(tci_handle h, void* h_dst, const void* d_src, size_t bytes) {
    if (!h) return TCI_ERR_INVALID;
    TCI_CUDA(h, cudaMemcpyAsync(h_dst, d_src, bytes, cudaMemcpyDeviceToHost, h->stream));
    TCI_CUDA(h, cudaStreamSynchronize(h->stream));
    return TCI_OK;
}

} // extern "C"
