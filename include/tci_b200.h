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
/* Run on a caller-owned cudaStream_t (e.g. torch's current stream; NULL = CUDA's legacy default stream), so the
 * caller's own copies/kernels on that stream are ordered with this library's.  tci_reset_stream returns to the
 * handle's private non-blocking stream (the default after tci_create). */
int tci_set_stream(tci_handle h, void* cuda_stream);
int tci_reset_stream(tci_handle h);
int tci_synchronize(tci_handle h);
/* Number of kernels this library has launched on the handle since creation (bench.py's gpu_launches). */
int64_t tci_launch_count(tci_handle h);
/* Raw device memory for hosts that have no allocator of their own (the Julia shim). */
int tci_malloc(tci_handle h, void** d_ptr, size_t bytes);
int tci_free(tci_handle h, void* d_ptr);
/* Stream-ordered variants (cudaMallocAsync on the handle's stream, default pool with an unbounded release threshold): no
 * device-wide synchronisation, a regrow costs microseconds.  Used by the native driver for its growing scratch (the
 * reference's Julia arrays come from the GC's pools the same way, tensorci2.jl:560-575 allocates Pi per bond). */
int tci_malloc_pooled(tci_handle h, void** d_ptr, size_t bytes);
int tci_free_pooled(tci_handle h, void* d_ptr);
int tci_memcpy_h2d(tci_handle h, void* d_dst, const void* h_src, size_t bytes);
int tci_memcpy_d2h(tci_handle h, void* h_dst, const void* d_src, size_t bytes);
int tci_memcpy_d2d(tci_handle h, void* d_dst, const void* d_src, size_t bytes);     /* enqueued on the handle's stream, not synchronised */

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
/* Per-launch update-kernel times (ms) of the last profiled rrlu; returns how many there are, copies at most cap. */
int64_t tci_rrlu_last_launch_ms(tci_handle h, float* h_ms, int64_t cap);

/* ---- Column-sharded rrLU over several GPUs (SURVEY 8e, row "rrLU"; an experiment next to the single-GPU default) ----------
 * Same factorisation as tci_rrlu_inplace_f64 (src/matrixlu.jl:346-396), with the columns of A dealt to `nranks` GPUs:
 * this rank holds the nloc columns [col0, col0 + nloc) as d_A (m x nloc, ld lda, overwritten).  One "record"
 * (tci_rrlu_dist_record_doubles doubles: best local candidate + that candidate's column) per rank and pivot step is the only
 * data exchanged; the caller gathers the records of all ranks (rank order) between the calls — an all-gather, or peer stores:
 *     tci_rrlu_dist_search(..., d_record);                       first search
 *     repeat maxrank times:  d_gathered <- all-gather(d_record); tci_rrlu_dist_step(..., d_gathered, d_record);
 *     tci_rrlu_dist_finish(...)
 * Every rank takes identical decisions (pivot choice with the reference's tie rule on the swapped column order, stop rule), so
 * permutations, pivots and factors are bit-identical to the single-GPU call.  finish returns the full L (m x npivot) on every
 * rank, this rank's columns of U (npivot x nloc, in LOCAL column order) and h_locpos[j] = 0-based position of local column j
 * in the swapped order (U[:, locpos[j]] = Uloc[:, j]). */
typedef struct tci_rrlu_dist_s* tci_rrlu_dist;
int tci_rrlu_dist_create(tci_handle h, int64_t m, int64_t n, int64_t nloc, int64_t col0, int rank, int nranks, int64_t maxrank,
                         double reltol, double abstol, int leftorthogonal, tci_rrlu_dist* out);
int64_t tci_rrlu_dist_record_doubles(tci_rrlu_dist d);
int tci_rrlu_dist_search(tci_handle h, tci_rrlu_dist d, double* d_A, int64_t lda, double* d_record);
int tci_rrlu_dist_step(tci_handle h, tci_rrlu_dist d, double* d_A, int64_t lda, const double* d_gathered, double* d_record);
int tci_rrlu_dist_status(tci_handle h, tci_rrlu_dist d, int64_t* h_npivot, int* h_stop);
/* Debug aid: with TCI_DIST_TRACE=1 in the environment at create time the step kernels record 8 device timestamps
 * (%globaltimer, ns) per pivot step — 0 swap kernel running, 1 headers arrived, 2 pivot column stored / arrived, 3 update
 * kernel running, 4 block 0 done, 5 all tiles done, 6 header released; h_stamps receives 8 * nsteps values. */
int tci_rrlu_dist_trace(tci_handle h, tci_rrlu_dist d, uint64_t* h_stamps, int64_t nsteps);
/* Fused exchange: instead of the caller's all-gather the step kernels talk through MAILBOXES (peer-mapped device memory:
 * allocate tci_rrlu_dist_mailbox_bytes with tci_malloc, share it with tci_ipc_export/open), two rounds per pivot step:
 * the update kernel's last block stores a 48-byte candidate header into every rank's mailbox, every block of the next
 * (swap) kernel picks the winner from the G headers, and the winner's owner stores the pivot column into the peers'
 * mailboxes from that same kernel.  Packets are self-validating (16 bytes = 2 x {32 data bits | 32-bit step number}), so
 * there are no flags and no fences; a receiver gives up after 4 s.  Two launches per step.  h_boxes[g] = rank g's mailbox
 * as mapped in THIS process (h_boxes[rank] = own).  All ranks must have called set_mailboxes (which zeroes the own
 * mailbox) before any rank calls run; run enqueues the whole factorisation (search + maxrank steps) without a host
 * synchronisation. */
int64_t tci_rrlu_dist_mailbox_bytes(tci_rrlu_dist d);
int tci_rrlu_dist_set_mailboxes(tci_handle h, tci_rrlu_dist d, void* const* h_boxes);
int tci_rrlu_dist_run(tci_handle h, tci_rrlu_dist d, double* d_A, int64_t lda);
int tci_rrlu_dist_finish(tci_handle h, tci_rrlu_dist d, const double* d_A, int64_t lda, int64_t* h_rowperm, int64_t* h_colperm,
                         int64_t* h_npivot, double* h_error, double* d_L, int64_t ldl, double* d_Uloc, int64_t ldu,
                         int64_t* h_locpos, double* h_pivoterrors);
int tci_rrlu_dist_destroy(tci_handle h, tci_rrlu_dist d);

/* ---- K1: batched function evaluation + Pi assembly ------------------------------------------------
 * The target function is a DEVICE functor (the reference's `f(indexset)::Float64`, evaluated on the GPU).
 * Built-in kinds (parameters in h_params):
 *   1 LORENTZ      p0 / (1 + sum_i v_i^2)                         README.md:21-29, test/test_tensorci2.jl:249
 *   2 QEXPSIN      exp(-p0 x) sin(2 pi p1 x^p2), x = sum_i (v_i - 1) 2^-i     test/test_tensorci2.jl:437
 *   3 QEXPPLUSSIN  exp(-p0 x) + p1 sin(p2 x)                      test/test_globalsearch.jl:14
 *   4 SUM          p0 * sum_i v_i                                 test/test_batcheval.jl:18
 *   5 QEXP2        exp(-p0 x) + p1 exp(-p2 x)                     test/test_tensorci2.jl:63,:146
 *   7 HASH         p0 u(v) + p1 / (1 + sum_i v_i^2), u = integer hash of v in [-1, 1): synthetic full-rank function (not in
 *                  the reference) that saturates any maxbonddim; benchmarks only
 * tci_functor_dense: lookup in a column-major table T[v_1, ..., v_N] (copied to the device).
 * tci_functor_from_source: the plugin path — CUDA C++ source that defines
 *     __device__ double tci_user_f(const int* idx, int n, const double* params, int nparams);
 *   (idx = the full 1-based multi-index, as the reference passes to f) is compiled with NVRTC for sm_100a.
 */
enum { TCI_F_LORENTZ_ID = 1, TCI_F_QEXPSIN_ID = 2, TCI_F_QEXPPLUSSIN_ID = 3, TCI_F_SUM_ID = 4, TCI_F_QEXP2_ID = 5, TCI_F_HASH_ID = 7 };
int tci_functor_builtin(tci_handle h, int kind, const double* h_params, int nparams, tci_functor* out);
int tci_functor_dense(tci_handle h, const double* h_table, const int32_t* h_dims, int ndims, tci_functor* out);
int tci_functor_from_source(tci_handle h, const char* cuda_src, const double* h_params, int nparams, tci_functor* out);
int tci_functor_destroy(tci_handle h, tci_functor f);

/* Replaces  _batchevaluate_dispatch / (f::BatchEvaluator)(leftindexset, rightindexset, Val(M))
 *           src/batcheval.jl:131-175, :196-214 as called by filltensor src/tensorci2.jl:475-497,
 *           and  updatemaxsample!  src/tensorci2.jl:636-638.
 * h_left: nleft x nl int32 (row = one 1-based multi-index), h_right: nright x nr; M = nsites - nl - nr free
 * centre sites (<= 8), enumerated first-index-fastest (batcheval.jl:160).  Element (i, c, j) is written to
 * d_out[i + c*stride_c + j*stride_j]; stride_c <= 0 means nleft and stride_j <= 0 means stride_c*prod(centre
 * dims), i.e. the dense column-major Array{Float64,M+2} of the reference.  Empty index sets write nothing
 * (batcheval.jl:140-142).  h_maxabs (may be NULL) is in/out: max(|*h_maxabs|, max |value|), NaN-propagating.
 */
int tci_batcheval_f64(tci_handle h, tci_functor f, const int32_t* h_localdims, int nsites,
                      const int32_t* h_left, int64_t nleft, int nl, const int32_t* h_right, int64_t nright, int nr,
                      double* d_out, int64_t stride_c, int64_t stride_j, double* h_maxabs);

/* The same evaluation without the host round trip for the maximum: max |value| (and a NaN flag) are merged into a 16-byte
 * DEVICE accumulator the caller owns (tci_malloc), across as many batches as it likes; nothing is synchronised.  The
 * accumulator is initialised with tci_maxacc_set(value) (= the running maxsamplevalue) and read with tci_maxacc_get, which
 * synchronises and returns max(|initial|, every value since), NaN-propagating — `updatemaxsample!` (src/tensorci2.jl:636-638)
 * deferred to where the host next needs the number (the tolerance of the next sweep, src/tensorci2.jl:1089-1090). */
int tci_batcheval_acc_f64(tci_handle h, tci_functor f, const int32_t* h_localdims, int nsites,
                          const int32_t* h_left, int64_t nleft, int nl, const int32_t* h_right, int64_t nright, int nr,
                          double* d_out, int64_t stride_c, int64_t stride_j, void* d_acc16);
/* NO COUNTERPART IN THE REFERENCE — added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arguments, same kernels and
 * bit-identical results as tci_batcheval_acc_f64, but the call allocates NOTHING (no cudaMalloc / cudaFree / allocating
 * library routine): the uploaded index tables and the folded functor states live in one file-scope `static __device__` array
 * of TCI_BATCHEVAL_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until exit.
 *   limit:  256-aligned(4 (nleft nl + nright nr)) + 256-aligned(16 (nleft + nright)) <= TCI_BATCHEVAL_NOALLOC_SCRATCH_BYTES, else
 *           TCI_ERR_INVALID with a message naming both numbers (tci_last_error); the per-pair states of the vectorised
 *           assembly (24 bytes per (centre, right) pair) are used only when they fit too, else the plain assembly kernel runs.
 *   d_acc16 must not be NULL (the running maximum is the caller's, see above); nothing in the array has to start at zero.
 *   EVERY call on a device shares the array, which keeps the previous call's contents: two calls that overlap on the GPU
 *   (different streams / handles on the same device) RACE.  Keep them on one stream or order them with events.
 *   The first call on each device resolves the array's address (cudaGetSymbolAddress, with that device current); make that
 *   call outside stream capture. */
#define TCI_BATCHEVAL_NOALLOC_SCRATCH_BYTES (4u << 20)
int tci_batcheval_noalloc_f64(tci_handle h, tci_functor f, const int32_t* h_localdims, int nsites,
                              const int32_t* h_left, int64_t nleft, int nl, const int32_t* h_right, int64_t nright, int nr,
                              double* d_out, int64_t stride_c, int64_t stride_j, void* d_acc16);
int tci_maxacc_set(tci_handle h, void* d_acc16, double value);
int tci_maxacc_get(tci_handle h, const void* d_acc16, double* h_value);

/* Sharded batch: the same evaluation, but every value is stored to `nouts` (1..8) destination buffers —
 * h_d_outs[0] is this GPU's buffer, the others are the peers' buffers mapped into this process (tci_ipc_open): the
 * assembly kernel IS the all-gather (peer stores over NVLink), there is no separate exchange pass.  The caller passes its
 * own block of the right index set and destination pointers already offset to that block's first column. */
int tci_batcheval_multi_f64(tci_handle h, tci_functor f, const int32_t* h_localdims, int nsites,
                            const int32_t* h_left, int64_t nleft, int nl, const int32_t* h_right, int64_t nright, int nr,
                            double* const* h_d_outs, int nouts, int64_t stride_c, int64_t stride_j, double* h_maxabs);
/* CUDA IPC plumbing for those buffers (memory from tci_malloc): a 64-byte handle to send to the peer processes, and the
 * peer side's open/close. */
int tci_ipc_export(tci_handle h, void* d_ptr, void* h_handle64);
int tci_ipc_open(tci_handle h, const void* h_handle64, void** d_ptr);
int tci_ipc_close(tci_handle h, void* d_ptr);

/* ---- K4a: FP64 dense algebra for site tensors (FP64 tensor pipe, DMMA) ------------------------------
 * tci_gemm_f64: C = alpha * A * B + beta * C, batched.  Element strides make transposes/reshaped views free:
 *   A(m,k) = d_A[m*sam + k*sak + z*bsa], B(k,n) = d_B[k*sbk + n*sbn + z*bsb], C(m,n) = d_C[m*scm + n*scn + z*bsc].
 * Replaces the `amat * bmat` of _contract (src/contraction.jl:214) and every `*` on the path
 * (src/matrixluci.jl:164,:176, src/abstracttensortrain.jl:341,:436).
 */
int tci_gemm_f64(tci_handle h, int64_t M, int64_t N, int64_t K, double alpha,
                 const double* d_A, int64_t sam, int64_t sak, const double* d_B, int64_t sbk, int64_t sbn,
                 double beta, double* d_C, int64_t scm, int64_t scn,
                 int64_t batch, int64_t bsa, int64_t bsb, int64_t bsc);

/* X * T = B in place on B (q x r); T (r x r) lower (lower != 0) or upper triangular, optionally with an implicit
 * unit diagonal; element strides as above.  Replaces `L2 / LowerTriangular(L1)` and `UpperTriangular(U1) \ U2`
 * (src/matrixluci.jl:207, :235; the left solve is the right solve on transposed strides). */
int tci_trsm_right_f64(tci_handle h, const double* d_T, int64_t st_row, int64_t st_col, int64_t r, int lower, int unit,
                       double* d_B, int64_t sb_row, int64_t sb_col, int64_t q);

/* Replaces  left(luci::MatrixLUCI) / right(luci)  src/matrixluci.jl:256-283 (colstimespivotinv :194-213,
 * rowmatrix :175-177, colmatrix :161-165, pivotinvtimesrows :227-241) from the outputs of tci_rrlu_*:
 * d_left (m x r, ld ldleft) and d_right (r x n, ld ldright); either may be NULL. */
int tci_luci_factors_f64(tci_handle h, int64_t m, int64_t n, int64_t r, int leftorthogonal,
                         const int64_t* h_rowperm, const int64_t* h_colperm,
                         const double* d_L, int64_t ldl, const double* d_U, int64_t ldu,
                         double* d_left, int64_t ldleft, double* d_right, int64_t ldright);

/* Replaces  Tmat = transpose(transpose(P) \ transpose(Pi1))  src/tensorci2.jl:626:  X = B * inv(P) with
 * B (q x r, ld ldb), P (r x r, ld ldp), X (q x r, ld ldx).  P is factorised by the full-pivot rrLU above
 * (at least as stable as the LAPACK partial-pivot LU the reference calls), then two triangular solves. */
int tci_solve_right_f64(tci_handle h, const double* d_B, int64_t q, int64_t ldb, const double* d_P, int64_t r, int64_t ldp,
                        double* d_X, int64_t ldx);

/* ---- Thin SVD ---------------------------------------------------------------------------------------
 * Replaces  LinearAlgebra.svd(A)  in _factorize(A, :SVD)  src/tensortrain.jl:244-262 (used by compress!(tt, :SVD) :302-348,
 * add/subtract src/abstracttensortrain.jl:524-545, contract_naive/contract_zipup src/contraction.jl:616-637, :751-788).
 * A (m x n, ld lda, OVERWRITTEN) = U diag(S) Vt with k = min(m, n): d_U (m x k, ld ldu), h_S[k] in descending order,
 * d_Vt (k x n, ld ldvt); d_U / d_Vt may be NULL.  One-sided Jacobi rotations (csrc/svd.cu); columns of U that belong to an
 * exactly zero singular value are zero.  h_sweeps (may be NULL) receives the number of Jacobi sweeps. */
int tci_svd_f64(tci_handle h, double* d_A, int64_t m, int64_t n, int64_t lda, double* d_U, int64_t ldu, double* h_S,
                double* d_Vt, int64_t ldvt, int* h_sweeps);
/* M[i, j] *= s[i] (by_rows != 0) or s[j]: the `Diagonal(S) * Vt` / `U * Diagonal(S)` of src/tensortrain.jl:252-259. */
int tci_diag_scale_f64(tci_handle h, double* d_M, int64_t rows, int64_t cols, int64_t ld, const double* h_s, int by_rows);

/* ---- K4c support: strided tensor copies, element-wise maps, permutation scatters ---------------------------
 * tci_strided_copy_f64: out (dense column-major, extents h_shape[0..ndim), first index fastest) <- the view of d_in
 *   that starts at element in_offset and has element stride h_in_strides[k] along output dimension k (ndim <= 8).
 *   Replaces `permutedims` and slicing around the GEMMs of the contraction code (src/contraction.jl:211-212, :256-258,
 *   :541-556, :595, :770).
 * tci_elementwise_f64: x <- g(x) in place; kind 1: p0*x + p1, 2: x^2 (abs2), 3: |x|, 4: exp(p0*x), 5: x / p0.  Replaces the
 *   `res .= obj.f.(res)` of Contraction.batchevaluate (src/contraction.jl:569-571) for the maps the tests use.
 * tci_unpermute_f64: out[perm[i], :] = in[i, :] (by_rows != 0) or out[:, perm[j]] = in[:, j]; h_perm 1-based.
 *   Replaces left(lu; permute=true) / right(lu; permute=true) (src/matrixlu.jl:716-746). */
int tci_strided_copy_f64(tci_handle h, const double* d_in, int64_t in_offset, const int64_t* h_shape, const int64_t* h_in_strides,
                         int ndim, double* d_out);
int tci_elementwise_f64(tci_handle h, int kind, double p0, double p1, double* d_x, int64_t n);
int tci_unpermute_f64(tci_handle h, const double* d_in, int64_t ldi, double* d_out, int64_t ldo, const int64_t* h_perm,
                      int64_t rows, int64_t cols, int by_rows);
/* updatemaxsample! (src/tensorci2.jl:636-638, util.jl:34-43) for a Pi matrix that was not produced by tci_batcheval_f64:
 * *h_maxabs = max(|*h_maxabs|, max |x|), NaN-propagating; x is rows x cols with leading dimension ld. */
int tci_maxabs_f64(tci_handle h, const double* d_x, int64_t rows, int64_t cols, int64_t ld, double* h_maxabs);

/* ---- K4b: TensorTrain sum / evaluate on device-resident cores ---------------------------------------
 * A core is the column-major (chi_l, d, chi_r) array of the reference (d = product of the site legs), i.e. a
 * (chi_l*d) x chi_r matrix; h_core_ptrs[b] is its DEVICE pointer, h_shapes[3b..3b+2] = (chi_l, d, chi_r),
 * h_lds[b] its leading dimension (NULL: dense).
 * tci_tt_sum_f64       replaces sum(tt)                 src/abstracttensortrain.jl:428-441
 * tci_tt_evaluate_f64  replaces evaluate(tt, indexset)  src/abstracttensortrain.jl:328-342 for a batch of
 *                      npoints multi-indices (h_indices: npoints x nsites, 1-based); values to h_out. */
int tci_tt_sum_f64(tci_handle h, const double* const* h_core_ptrs, const int64_t* h_shapes, const int64_t* h_lds,
                   int nsites, double* h_out);
int tci_tt_evaluate_f64(tci_handle h, const double* const* h_core_ptrs, const int64_t* h_shapes, const int64_t* h_lds,
                        int nsites, const int32_t* h_indices, int64_t npoints, double* h_out);

/* ---- Device-resident sweeps (csrc/sweep.cu): the bond loop of a half-sweep in ONE launch ----------------------------------
 * Replaces the loop `for bondindex in 1:n-1 updatepivots!(...)` of sweep2site! (src/tensorci2.jl:1226-1250; each iteration =
 * updatepivots! :825-930, :full branch: kronecker/union :842-843, filltensor :848-852, updatemaxsample! :855, MatrixLUCI ->
 * rrlu src/matrixlu.jl:346-396, new Iset[b+1]/Jset[b] :915-916, pivoterrors) for Pi matrices that fit shared memory (one CTA:
 * ~210 KB; a cluster of 8 CTAs with distributed shared memory: ~1.7 MB).  The host keeps the index sets between launches and
 * exchanges them as int32 arenas:
 *   h_in  (in_words):  [0] N  [1] forward (1/0)  [2] s0  [3] s1  — steps [s0, s1) of the sweep order, bond b0 = forward ? s : N-2-s
 *                      (0-based) —  [4] maxbonddim (<= 0: none)  [5] useextra  [6] total number of table rows  [7] 0;  localdims[N];
 *                      per site b (12 words): nI, offI, nJ, offJ, nXI, offXI, nXJ, offXJ — counts and word offsets of Iset[b]
 *                      (rows of length b), Jset[b] (rows of length N-b-1) and of the history entries used as extraIset /
 *                      extraJset — then baseI, baseJ, baseXI, baseXJ: the running row count before each table in the order
 *                      I_0, J_0, XI_0, XJ_0, I_1, ...; then the tables (row-major, 1-based values).  The sets hold unique rows.
 *   h_out (capacity out_cap_words, *h_used_words written): [0] status — 0 all steps done, 1 stopped before step [1] because
 *                      its Pi does not fit (run that bond through tci_tci2_updatepivots-style calls and continue), 2 / 3
 *                      "lu.L / lu.U contains NaNs" at step [1];  [2] words used;  per step s-s0 (6 words at 8 + 6 (s-s0)):
 *                      r, offA (new Iset[b0+1]: r rows of length b0+1), offB (new Jset[b0]: r rows of length N-b0-1), offE
 *                      (pivoterrors(luci): r+1 doubles, 8-byte aligned), m, n (the size of that bond's Pi).
 * reltol / abstol: the rrlu tolerances of every bond (sweep2site! passes reltol = 1e-14, src/tensorci2.jl:1236);
 * leftorthogonal = forward.  max |f| of every evaluated element is merged into d_acc16 (see tci_batcheval_acc_f64).
 * Built-in functor kinds only (tci_dsweep_supported); results are bit-identical to the per-bond entry points. */
int tci_dsweep_supported(tci_handle h, tci_functor f);
/* fillsitetensors!(tci, f) (src/globalsearch.jl:202-208; per site setsitetensor! src/tensorci2.jl:599-629: T_b = Pi1 inv(P), the last
 * tensor by evaluation) for ALL sites in one launch.  h_in: the arena above (steps / extras unused); h_d_cores[b] = device buffer
 * of |Iset[b]| * d_b * |Jset[b]| doubles for site tensor b (dense column-major).  Every bond dimension must be <= 64.  Returns
 * TCI_ERR_NOT_SQUARE where the reference throws "Pivot matrix at bond b is not square!" (:623). */
int tci_dsweep_fill_f64(tci_handle h, tci_functor f, const int32_t* h_in, int64_t in_words, void* d_acc16, double* const* h_d_cores);
/* sweep1site!(tci, f, :forward; reltol, abstol, maxbonddim, updatetensors=true) (src/tensorci2.jl:659-725) in one launch: per bond
 * Pi = f((Iset[b] x s_b) x Jset[b]), MatrixLUCI, the new Iset[b+1] / Jset[b], T_b = left(luci) (src/matrixluci.jl:194-213, :256-262),
 * the last tensor by evaluation.  h_in: the arena above with [4] = maxbonddim; h_d_cores[b] / h_core_capacity[b]: buffer and its
 * size in doubles (|Jset[b-1]| * d_b * |Jset[b]| suffices).  h_out as above (one record per bond); status 4 = "Error: NaN in
 * tensor T[b]" (:705-707) at bond [1]; status 1 with [1] = N-1 means only the last tensor is left to the caller. */
int tci_dsweep_sweep1site_f64(tci_handle h, tci_functor f, const int32_t* h_in, int64_t in_words, double reltol, double abstol,
                              void* d_acc16, double* const* h_d_cores, const int64_t* h_core_capacity, int32_t* h_out,
                              int64_t out_cap_words, int64_t* h_used_words);
int tci_dsweep_halfsweep_f64(tci_handle h, tci_functor f, const int32_t* h_in, int64_t in_words, double reltol, double abstol,
                             void* d_acc16, int32_t* h_out, int64_t out_cap_words, int64_t* h_used_words);

/* ---- Compiled host driver of the TensorCI2 sweeps (csrc/tci2_driver.cpp) -------------------------------------------------
 * The reference keeps sweep strategy, index-set bookkeeping and convergence control in its host language and so does this
 * library's boundary: a Julia host calls the entry points above once per bond.  For hosts whose per-call overhead is large
 * next to a microsecond-sized kernel (Python), the same host logic is available compiled: a tci_tci2 object is the
 * `TensorCI2{Float64}` state (src/tensorci2.jl:50-93) — Iset/Jset as int32 tables on the host, site tensors in device memory —
 * and the calls below are the reference's functions of the same names, written purely on top of the entry points above
 * (every number still comes from those kernels).  Sites and bonds are 1-based, index values
 * 1-based; maxbonddim <= 0 means unbounded (typemax(Int)).  sweepstrategy: 0 :forward, 1 :backandforth, 2 :backward
 * (src/sweepstrategies.jl:41-50).  The message of the last failure on an object is tci_tci2_last_error(t). */
/* optfirstpivot(f, localdims, firstpivot; maxsweep) src/util.jl:260-291: h_pivot (nsites, 1-based) is improved in place by a
 * greedy one-site-at-a-time search for a large |f| (one batch evaluation per site). */
int tci_optfirstpivot_f64(tci_handle h, tci_functor f, const int32_t* h_localdims, int nsites, int32_t* h_pivot, int maxsweep);
typedef struct tci_tci2_s* tci_tci2;
int tci_tci2_create(tci_handle h, const int32_t* h_localdims, int nsites, tci_tci2* out);       /* TensorCI2{V}(localdims) :95-103 */
int tci_tci2_destroy(tci_tci2 t);
const char* tci_tci2_last_error(tci_tci2 t);
/* Seed of the global pivot finder's random starts (the reference draws from Julia's default RNG, which has no portable
 * restatement: global pivots are not comparable with a Julia run). */
int tci_tci2_seed(tci_tci2 t, uint64_t seed);
/* engine 0 (default): one C-ABI call per bond phase (batch evaluation, rrLU, factors).  engine 1: half-sweeps whose Pi matrices
 * fit shared memory run as ONE launch each (tci_dsweep_halfsweep_f64); bonds that do not fit, the rook search and plugin
 * functors take the per-bond calls.  Index sets, ranks and errors are identical either way. */
int tci_tci2_set_engine(tci_tci2 t, int engine);
/* Pivot search strategy of updatepivots! (and therefore of sweep2site! / optimize!): rook = 0 is :full (the default), 1 is :rook
 * (src/tensorci2.jl:863-905 -> arrlu src/matrixlu.jl:500-569 over SubMatrix strips :764-804: Pi is never formed).  `seed` starts
 * the stream behind pushrandomsubset! (src/util.jl:173-219; Julia's global RNG in the reference). */
int tci_tci2_set_pivotsearch(tci_tci2 t, int rook, uint64_t seed);
/* addglobalpivots!(tci, pivots) :335-357 — h_pivots: npivots x nsites. */
int tci_tci2_addglobalpivots(tci_tci2 t, const int32_t* h_pivots, int64_t npivots);
/* The constructor's maxsamplevalue = maximum(abs(f(p)) for p in pivots) and its zero check :105-116. */
int tci_tci2_init_maxsample(tci_tci2 t, tci_functor f, const int32_t* h_pivots, int64_t npivots);
/* State in / out: which = 0 Iset[site] (rows of length site-1), 1 Jset[site] (rows of length nsites-site).
 * tci_tci2_indexset: h_rows may be NULL to query the count only. */
int tci_tci2_set_indexset(tci_tci2 t, int which, int site, const int32_t* h_rows, int64_t count);
int tci_tci2_indexset(tci_tci2 t, int which, int site, int32_t* h_rows, int64_t* h_count);
int tci_tci2_set_maxsamplevalue(tci_tci2 t, double v);
/* updatepivots!(tci, b, f, leftorthogonal; reltol, abstol, maxbonddim, extraIset, extraJset) :825-930 (:full branch).
 * h_extraI: nextraI rows of length bond, h_extraJ: nextraJ rows of length nsites-bond; either may be NULL/0. */
int tci_tci2_updatepivots(tci_tci2 t, tci_functor f, int bond, int leftorthogonal, double reltol, double abstol, int64_t maxbonddim,
                          const int32_t* h_extraI, int64_t nextraI, const int32_t* h_extraJ, int64_t nextraJ);
/* sweep2site!(tci, f, niter; iter1, abstol, maxbonddim, sweepstrategy, strictlynested, fillsitetensors) :1195-1258 */
int tci_tci2_sweep2site(tci_tci2 t, tci_functor f, int niter, int iter1, double abstol, int64_t maxbonddim, int sweepstrategy,
                        int strictlynested, int fillsitetensors);
/* sweep1site!(tci, f, sweepdirection; reltol, abstol, maxbonddim, updatetensors) :659-725; forward != 0 is :forward */
int tci_tci2_sweep1site(tci_tci2 t, tci_functor f, int forward, double reltol, double abstol, int64_t maxbonddim, int updatetensors);
int tci_tci2_makecanonical(tci_tci2 t, tci_functor f, double reltol, double abstol, int64_t maxbonddim);      /* :738-749 */
int tci_tci2_fillsitetensors(tci_tci2 t, tci_functor f);                      /* src/globalsearch.jl:202-208, setsitetensor! :599-629 */
/* optimize!(tci, f; tolerance, maxbonddim, maxiter, sweepstrategy, normalizeerror, ncheckhistory, nsearchglobalpivot,
 * maxnglobalpivot, tolmarginglobalsearch, strictlynested, checkconvglobalpivot) :1018-1172 with the DefaultGlobalPivotFinder
 * (src/globalpivotfinder.jl:205-264; all candidates of an iteration in one batch evaluation).  h_ranks / h_errors receive
 * *h_niter <= maxiter entries (errors already divided by the normalisation, as the reference returns them). */
int tci_tci2_optimize(tci_tci2 t, tci_functor f, double tolerance, int64_t maxbonddim, int maxiter, int sweepstrategy, int normalizeerror,
                      int ncheckhistory, int nsearchglobalpivot, int maxnglobalpivot, double tolmarginglobalsearch, int strictlynested,
                      int checkconvglobalpivot, int64_t* h_ranks, double* h_errors, int* h_niter);
/* Site tensor `site` as a dense column-major (chi_l, d, chi_r) array in device memory; the pointer stays valid until the next
 * call that changes the object.  Fails when the site tensors are invalidated (:206-220). */
int tci_tci2_core(tci_tci2 t, int site, const double** d_core, int64_t* h_shape3);
int tci_tci2_sitetensors_available(tci_tci2 t);
/* The whole result in two calls: all site tensors copied device-to-device (enqueued, not synchronised) into ONE caller buffer,
 * dense and back to back — h_shapes3[3b..] = (chi_l, d, chi_r), h_offsets[b] = first element of tensor b, *h_total = doubles
 * needed (call with d_dst = NULL to size the buffer) — and all index sets packed as I_0, J_0, I_1, J_1, ... (row-major int32
 * tables; h_counts2[2b] = |Iset[b]|, [2b+1] = |Jset[b]|; h_dst = NULL queries *h_total). */
int tci_tci2_export_cores(tci_tci2 t, double* d_dst, int64_t capacity, int64_t* h_shapes3, int64_t* h_offsets, int64_t* h_total);
int tci_tci2_export_indexsets(tci_tci2 t, int32_t* h_dst, int64_t capacity, int64_t* h_counts2, int64_t* h_total);
/* evaluate(tci, indexset) src/abstracttensortrain.jl:328-342 at npoints multi-indices (h_points: npoints x nsites, 1-based; values
 * to h_out) and sum(tci) :428-441, on the object's own site tensors (tci_tt_evaluate_f64 / tci_tt_sum_f64 underneath). */
int tci_tci2_evaluate(tci_tci2 t, const int32_t* h_points, int64_t npoints, double* h_out);
int tci_tci2_sum(tci_tci2 t, double* h_out);
/* bonderrors (nsites-1), pivoterrors (query the count with h_pivoterrors = NULL first), maxsamplevalue; any may be NULL. */
int tci_tci2_errors(tci_tci2 t, double* h_bonderrors, double* h_pivoterrors, int64_t* h_npivoterrors, double* h_maxsamplevalue);
/* h_counts2 = {function evaluations, bond updates}; h_seconds4 = host wall time in {evaluation, rrLU, factors, index sets}. */
int tci_tci2_stats(tci_tci2 t, int64_t* h_counts2, double* h_seconds4);

#ifdef __cplusplus
}
#endif
#endif /* TCI_B200_H */
