/* tci_b200.h — C ABI of the B200-native TensorCI2 hot path.
 *
 * Drop-in boundary for TensorCrossInterpolation.jl's two-site sweep.  The reference has no FFI of its
 * own (it is pure Julia); each entry point below names the Julia method it replaces, i.e. what a
 * `ccall` shim on the reference side binds (see INTEGRATION.md for that shim).
 *
 * Conventions
 *   - every function returns an int status: TCI_OK (0) or a negative tci_status code; the message of
 *     the last failure on a handle is `tci_last_error(h)`;
 *   - matrices/tensors are column-major FP64 (Julia layout); `d_` = device pointer, `h_` = host pointer;
 *   - index VALUES crossing the boundary (permutations, multi-indices) are 1-based like Julia;
 *   - one tci_handle owns one CUDA device, one stream and its workspaces; it is not thread-safe
 *     (the reference never calls these paths concurrently: src/tensorci2.jl:1225-1251 is sequential);
 *   - there is no CPU fallback: without a CUDA device tci_create fails with TCI_ERR_CUDA.
 */
#ifndef TCI_B200_H
#define TCI_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct tci_context* tci_handle;
typedef struct tci_functor_s* tci_functor;

typedef enum {
    TCI_OK = 0,
    TCI_ERR_INVALID = -1,       /* bad argument (ArgumentError / DimensionMismatch in the reference) */
    TCI_ERR_CUDA = -2,          /* CUDA runtime/driver failure, or no device */
    TCI_ERR_NAN_L = -3,         /* "lu.L contains NaNs"            src/matrixlu.jl:376-378 */
    TCI_ERR_NAN_U = -4,         /* "lu.U contains NaNs"            src/matrixlu.jl:379-381 */
    TCI_ERR_NOT_SQUARE = -5,    /* "Pivot matrix at bond b is not square!"  src/tensorci2.jl:623 */
    TCI_ERR_NAN_TENSOR = -6,    /* "Error: NaN in tensor T[b]"     src/tensorci2.jl:705-707 */
    TCI_ERR_NVRTC = -7,         /* user functor failed to compile (log in tci_last_error) */
    TCI_ERR_NOMEM = -8
} tci_status;

/* ---- context ------------------------------------------------------------------------------- */
int tci_create(tci_handle* out, int device);
int tci_destroy(tci_handle h);
const char* tci_last_error(tci_handle h);               /* h may be NULL: last tci_create failure */
const char* tci_version(void);
/* Run on a caller-owned cudaStream_t (e.g. torch's current stream); NULL = the handle's own stream. */
int tci_set_stream(tci_handle h, void* cuda_stream);
int tci_synchronize(tci_handle h);
/* Number of kernels this library has launched on the handle since creation (bench.py's gpu_launches). */
int64_t tci_launch_count(tci_handle h);
/* Raw device memory for hosts that have no allocator of their own (the Julia shim). */
int tci_malloc(tci_handle h, void** d_ptr, size_t bytes);
int tci_free(tci_handle h, void* d_ptr);
int tci_memcpy_h2d(tci_handle h, void* d_dst, const void* h_src, size_t bytes);
int tci_memcpy_d2h(tci_handle h, void* h_dst, const void* d_src, size_t bytes);

/* ---- K2 + K3: full-pivot rank-revealing LU --------------------------------------------------
 * Replaces  rrlu!(A; maxrank, reltol, abstol, leftorthogonal)        src/matrixlu.jl:420-430
 *           (loop _optimizerrlu! :346-396, submatrixargmax :46-87, addpivot! :295-322).
 * d_A (m x n, leading dimension lda) is overwritten exactly as the reference overwrites A: after the
 * call it holds the swapped, partially eliminated matrix whose first npivot columns/rows are L and U.
 * Outputs:
 *   h_rowperm[m], h_colperm[n]   full 1-based permutations, pivots first          (:201-202)
 *   h_npivot, h_error            lu.npivot, lu.error (0 when npivot == min(m,n))  (:206-207, :391-393)
 *   d_L (m x npivot, ld ldl), d_U (npivot x n, ld ldu): tril/triu with the unit diagonal on L
 *                                (leftorthogonal) or U (:372-388); ldl >= m, ldu >= min(maxrank,m,n);
 *                                either may be NULL to skip the extraction
 *   h_pivoterrors[npivot+1]      pivoterrors(lu) = [|diag|..., error]             (:799-801); may be NULL
 * maxrank <= 0 or > min(m,n) means min(m,n) (typemax(Int) in the reference).
 * Returns TCI_ERR_NAN_L / TCI_ERR_NAN_U where the reference throws.
 */
int tci_rrlu_inplace_f64(tci_handle h, double* d_A, int64_t m, int64_t n, int64_t lda,
                         int64_t maxrank, double reltol, double abstol, int leftorthogonal,
                         int64_t* h_rowperm, int64_t* h_colperm, int64_t* h_npivot, double* h_error,
                         double* d_L, int64_t ldl, double* d_U, int64_t ldu, double* h_pivoterrors);

/* Replaces  rrlu(A; ...)  src/matrixlu.jl:455-463 with A, L, U in HOST memory (the call a Julia/Python
 * user makes on an ordinary Matrix{Float64}): copies A to the device (A is not modified), factorises,
 * copies L (m x npivot, ld ldl) and U (npivot x n, ld ldu) back.  h_L/h_U may be NULL. */
int tci_rrlu_host_f64(tci_handle h, const double* h_A, int64_t m, int64_t n, int64_t lda,
                      int64_t maxrank, double reltol, double abstol, int leftorthogonal,
                      int64_t* h_rowperm, int64_t* h_colperm, int64_t* h_npivot, double* h_error,
                      double* h_L, int64_t ldl, double* h_U, int64_t ldu, double* h_pivoterrors);

/* Replaces  submatrixargmax(abs2, A, k)  src/matrixlu.jl:133-135 (k is 1-based like the reference):
 * (row, col), 1-based, of the largest |A[i,j]|^2 over i >= k, j >= k; ties -> smallest column, then
 * smallest row; NaN never wins. */
int tci_submatrixargmax_abs2_f64(tci_handle h, const double* d_A, int64_t m, int64_t n, int64_t lda,
                                 int64_t k, int64_t* h_row, int64_t* h_col);

/* Device time (ms, CUDA events on the handle's stream) the last rrlu call spent in
 * [0] the whole factorisation loop, [1] the Schur-update kernel launches alone (only when profiling is on,
 * else 0), [2] L/U extraction; and the number of pivot steps run.
 * tci_rrlu_set_profiling(h, 1) brackets every update-kernel launch with an event pair (adds a little
 * launch overhead, so bench.py uses it on a separate pass). */
int tci_rrlu_set_profiling(tci_handle h, int on);
int tci_rrlu_last_timing(tci_handle h, double* h_ms3, int64_t* h_steps);

#ifdef __cplusplus
}
#endif
#endif /* TCI_B200_H */
