/* ssb200.h -- C ABI of the B200-native l1 hot path (ss::homotopy / ss::irls solve loops).
 *
 * This is the drop-in boundary.  In the reference the four `Policy::run` overloads
 *   homotopy_policy::run(A, y, tol, max_iter, x)   src/lib.cpp:30-46
 *   irls_policy::run(state, y, tol, max_iter, x)   src/lib.cpp:61-73
 * dispatch through kernelpp::run<solve_homotopy / solve_irls> (kernel_invoke.h:236-249) to the
 * CPU solver loops (src/solvers/homotopy-cpu.cpp:185-297, src/solvers/irls-cpu.cpp:62-146), which
 * do their arithmetic through the OpenBLAS wrapper (src/linalg/blas_wrapper.h:99-400).  The entry
 * points below replace that dispatch + BLAS layer for this path: a maintainer re-implements the four
 * overloads as calls into this header (INTEGRATION.md shows the stub) and nothing above
 * the include/ss headers changes.
 *
 * Conventions
 *   - plain pointers and sizes only; no C++ / torch types cross the boundary.
 *   - every function returns an ssb200_status; nothing throws, nothing aborts (the reference
 *     abort()s when BLAS cannot be loaded, src/linalg/blas_wrapper.cpp:39-49; here that is
 *     SSB200_COMPUTE_MODE_UNAVAILABLE).  ssb200_status_string() gives the kernelpp wording
 *     (third_party/kernelpp/include/kernelpp/kernel.h:122-133) so `maybe<report>` error strings
 *     stay what callers saw before.
 *   - strides are in ELEMENTS, as ss::ndspan carries them (include/ss/ndspan.h:28-31); any
 *     row-major / column-major / padded / sub-view dictionary is accepted
 *     (src/linalg/blas_wrapper.h:73-93; bindings/python/tests/test_binding.py:31-68).
 *   - there is NO CPU fallback: without an sm_100 device the create call fails.
 *   - a handle may be used from one host thread at a time (the reference's solve is
 *     re-entrant; here the handle owns its stream and workspace, so callers that want
 *     concurrency create one handle per thread).
 */
#ifndef SSB200_H
#define SSB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum ssb200_status {
    SSB200_OK = 0,
    SSB200_KERNEL_FAILED = 1,            /* "Kernel Failed"            kernel.h:122-133 */
    SSB200_COMPUTE_MODE_DISABLED = 2,    /* "Compute Mode Disabled"   */
    SSB200_COMPUTE_MODE_UNAVAILABLE = 3, /* "Compute Mode Unavailable": no sm_100 device */
    SSB200_KERNEL_NOT_DEFINED = 4,       /* "Kernel Not Defined"      */
    SSB200_CANCELLED = 5,                /* "Cancelled"               */
    SSB200_INVALID_ARGUMENT = 6          /* shape / pointer checks the reference only assert()s
                                            (homotopy-cpu.cpp:193-199) */
} ssb200_status;

const char* ssb200_status_string(int status);
/* detail of the last failure on the calling thread (CUDA error text, shape message) */
const char* ssb200_last_error(void);
/* {major, minor, patch} of the reference API this library stands in for (binding.cpp:42-46) */
void ssb200_version(int v[3]);
/* number of visible sm_100 devices (0 when there is none / no driver) */
int ssb200_device_count(void);

/* ------------------------------------------------------------------------------------------
 * Homotopy.  Replaces ss::solver<T, homotopy_policy> state + solve (include/ss/ss.h:98-115).
 *
 * create: uploads the M x N dictionary once and repacks it atom-major ("A^T row-major"), builds
 *   the TF32 hi/lo split used by the tensor-core contraction and the Gram matrix A^T A.  The
 *   reference state is a NON-owning view of A (include/ss/policies.h:42); here the device copy is
 *   owned by the handle, so a caller that edits A in place between solves (the reference bench
 *   does, src/solvers/homotopy_bench.cpp:38-48) calls ssb200_homotopy_refresh_*.
 * solve: one signal, host pointers (homotopy-cpu.cpp:185-275).  x (length N) is overwritten.
 *   *solution_error is |c|_inf at exit, *iter the number of loop passes (policies.h:25-32).
 * solve_batch: `batch` independent signals, row b of Y at Y + b*ldy (host memory), x_b written
 *   to X + b*ldx (host memory).  New functionality (the reference has no batching, ss.h:50);
 *   each signal's result equals what solve() returns for it.
 *   A batch row and a solve() of the same signal follow the same algorithm but not the same kernels (one cooperative
 *   kernel per single solve, SIMT or tcgen05 contraction by batch size), so they agree to rounding, not bit for bit.
 * solve_batch_dev: same with DEVICE pointers for Y, X, iters, errors (no host copies).  `stream`
 *   (a cudaStream_t, may be NULL) is the stream the inputs were produced on: it is synchronised
 *   first, the work then runs on the handle's own streams, and the call returns when the results
 *   are complete (visible to every stream).
 * create_dev: the device is synchronised first (dA may come from any stream, e.g. an NCCL broadcast).
 * max_iterations: any value; the active set holds min(max_iterations + 1, N) atoms and that bound must
 *   not exceed 4096 (SSB200_INVALID_ARGUMENT otherwise, before any work is queued).
 */
int ssb200_homotopy_create_f32(const float* A, size_t m, size_t n, ptrdiff_t row_stride, ptrdiff_t col_stride,
                               int device, void** handle);
int ssb200_homotopy_create_f64(const double* A, size_t m, size_t n, ptrdiff_t row_stride, ptrdiff_t col_stride,
                               int device, void** handle);
/* device-resident dictionary (dA on `device`, strides in elements) -- used after an NCCL broadcast of A */
int ssb200_homotopy_create_dev_f32(const float* dA, size_t m, size_t n, ptrdiff_t row_stride, ptrdiff_t col_stride,
                                   int device, void** handle);
int ssb200_homotopy_create_dev_f64(const double* dA, size_t m, size_t n, ptrdiff_t row_stride, ptrdiff_t col_stride,
                                   int device, void** handle);
int ssb200_homotopy_refresh_f32(void* handle, const float* A, ptrdiff_t row_stride, ptrdiff_t col_stride);
int ssb200_homotopy_refresh_f64(void* handle, const double* A, ptrdiff_t row_stride, ptrdiff_t col_stride);

/* Residual-correlation mode.  The reference re-forms c = A^T (y - A x) from scratch every iteration
 * (src/solvers/homotopy-cpu.cpp:87-98, called at :215 and :255).
 *   EXACT        does the same: r = y - A_G x_G is rebuilt from the active atoms every iteration and c = A^T r is contracted
 *                together with q = A^T (A d) (two contraction rows per signal-iteration).
 *   INCREMENTAL  carries c -= gamma q from iteration to iteration (one contraction row per signal-iteration).  Algebraically
 *                the same; on coherent (nearly parallel) dictionaries the rounding drift of c leaves the reference's path
 *                (more solves run to max_iterations), so it is only taken where it was measured to stay inside the
 *                reference's own build-to-build noise.
 *   AUTO         (default) EXACT when the dictionary's mutual coherence max_{i != j} |a_i . a_j| / (|a_i| |a_j|) exceeds 0.7
 *                (e.g. the reference's bench workload, homotopy_bench.cpp:14-48), INCREMENTAL otherwise.
 * get_residual_info reports the mode in effect for the handle's dictionary and its mutual coherence. */
enum { SSB200_RESIDUAL_AUTO = 0, SSB200_RESIDUAL_EXACT = 1, SSB200_RESIDUAL_INCREMENTAL = 2 };
int ssb200_homotopy_set_residual_mode(void* handle, int mode);
int ssb200_homotopy_get_residual_info(void* handle, int* mode_in_effect, double* coherence);

int ssb200_homotopy_solve_f32(void* handle, const float* y, ptrdiff_t y_stride, float tolerance, uint32_t max_iterations,
                              float* x, ptrdiff_t x_stride, uint32_t* iter, double* solution_error);
int ssb200_homotopy_solve_f64(void* handle, const double* y, ptrdiff_t y_stride, double tolerance, uint32_t max_iterations,
                              double* x, ptrdiff_t x_stride, uint32_t* iter, double* solution_error);

int ssb200_homotopy_solve_batch_f32(void* handle, const float* Y, size_t batch, size_t ldy, float tolerance,
                                    uint32_t max_iterations, float* X, size_t ldx, uint32_t* iters, double* errors);
int ssb200_homotopy_solve_batch_f64(void* handle, const double* Y, size_t batch, size_t ldy, double tolerance,
                                    uint32_t max_iterations, double* X, size_t ldx, uint32_t* iters, double* errors);

/* solve_batch with the SPARSE form of the result in host memory (new): row b's non-zero coefficients in index order as
 * (idx[b*cap + k], val[b*cap + k]) for k < count[b]; entries at k >= count[b] are unspecified.  cap >= min(max_iterations + 1, N),
 * the most non-zeros a solution can have.  No dense row is formed on the host: 8 (cap) + 16 bytes per signal cross PCIe and are
 * written once, where the dense form writes 4 N bytes per signal (what bounds the dense entry when several GPUs of one box
 * return results to the same host memory). */
int ssb200_homotopy_solve_batch_sparse_f32(void* handle, const float* Y, size_t batch, size_t ldy, float tolerance,
                                           uint32_t max_iterations, int32_t* idx, float* val, int32_t* count, size_t cap,
                                           uint32_t* iters, double* errors);
int ssb200_homotopy_solve_batch_sparse_f64(void* handle, const double* Y, size_t batch, size_t ldy, double tolerance,
                                           uint32_t max_iterations, int32_t* idx, double* val, int32_t* count, size_t cap,
                                           uint32_t* iters, double* errors);

int ssb200_homotopy_solve_batch_dev_f32(void* handle, const float* dY, size_t batch, size_t ldy, float tolerance,
                                        uint32_t max_iterations, float* dX, size_t ldx, uint32_t* d_iters,
                                        double* d_errors, void* stream);
int ssb200_homotopy_solve_batch_dev_f64(void* handle, const double* dY, size_t batch, size_t ldy, double tolerance,
                                        uint32_t max_iterations, double* dX, size_t ldx, uint32_t* d_iters,
                                        double* d_errors, void* stream);

/* ------------------------------------------------------------------------------------------
 * Multi-GPU group (new: the reference is one signal per call on one CPU, include/ss/ss.h:50).
 * One process, one communicator over the listed GPUs of the box (ncclCommInitAll), one host thread per GPU.
 *   create: uploads A to devices[0], replicates it on the others with ONE ncclBroadcast over NVLink, builds a homotopy
 *     handle per GPU.  NCCL is bound at run time (libnccl.so.2); a one-device group never touches it.
 *   solve_batch: host pointers, same contract as ssb200_homotopy_solve_batch_*: the batch is cut into ndev contiguous
 *     shards (the first batch % ndev shards one signal longer), every GPU's thread moves its shard in and out itself.
 *   solve_batch_dev: dY, dX, d_iters, d_errors live on devices[0] (dense rows: ldy == M, ldx == N when ndev > 1); shards
 *     are scattered with ncclSend/ncclRecv, solved, and the coefficient rows and reports gathered back the same way.  No
 *     other collective exists on the path: signals are independent.
 */
int ssb200_group_create_f32(const float* A, size_t m, size_t n, ptrdiff_t row_stride, ptrdiff_t col_stride,
                            const int* devices, int ndev, void** group);
int ssb200_group_create_f64(const double* A, size_t m, size_t n, ptrdiff_t row_stride, ptrdiff_t col_stride,
                            const int* devices, int ndev, void** group);
int ssb200_group_solve_batch_f32(void* group, const float* Y, size_t batch, size_t ldy, float tolerance, uint32_t max_iterations,
                                 float* X, size_t ldx, uint32_t* iters, double* errors);
int ssb200_group_solve_batch_f64(void* group, const double* Y, size_t batch, size_t ldy, double tolerance, uint32_t max_iterations,
                                 double* X, size_t ldx, uint32_t* iters, double* errors);
int ssb200_group_solve_batch_dev_f32(void* group, const float* dY, size_t batch, size_t ldy, float tolerance, uint32_t max_iterations,
                                     float* dX, size_t ldx, uint32_t* d_iters, double* d_errors, void* stream);
int ssb200_group_solve_batch_dev_f64(void* group, const double* dY, size_t batch, size_t ldy, double tolerance, uint32_t max_iterations,
                                     double* dX, size_t ldx, uint32_t* d_iters, double* d_errors, void* stream);
int ssb200_group_set_residual_mode(void* group, int mode);
int ssb200_group_size(void* group);
void ssb200_group_destroy(void* group);

/* ------------------------------------------------------------------------------------------
 * IRLS.  Replaces ss::solver<T, irls_policy>: irls_state(A) = Householder QR of A
 * (src/lib.cpp:51-57, src/linalg/qr_decomposition.h:93-190; requires M >= N, :101) and the
 * solve loop (src/solvers/irls-cpu.cpp:62-124).  spd_failure is a successful result
 * (include/ss/policies.h:58-71); solution_error is the final epsilon (irls-cpu.cpp:123).
 */
int ssb200_irls_create_f32(const float* A, size_t m, size_t n, ptrdiff_t row_stride, ptrdiff_t col_stride,
                           int device, void** handle);
int ssb200_irls_create_f64(const double* A, size_t m, size_t n, ptrdiff_t row_stride, ptrdiff_t col_stride,
                           int device, void** handle);
int ssb200_irls_solve_f32(void* handle, const float* y, ptrdiff_t y_stride, float tolerance, uint32_t max_iterations,
                          float* x, ptrdiff_t x_stride, uint32_t* iter, double* solution_error, uint8_t* spd_failure);
int ssb200_irls_solve_f64(void* handle, const double* y, ptrdiff_t y_stride, double tolerance, uint32_t max_iterations,
                          double* x, ptrdiff_t x_stride, uint32_t* iter, double* solution_error, uint8_t* spd_failure);
int ssb200_irls_solve_batch_f32(void* handle, const float* Y, size_t batch, size_t ldy, float tolerance,
                                uint32_t max_iterations, float* X, size_t ldx, uint32_t* iters, double* errors,
                                uint8_t* spd_failures);
int ssb200_irls_solve_batch_f64(void* handle, const double* Y, size_t batch, size_t ldy, double tolerance,
                                uint32_t max_iterations, double* X, size_t ldx, uint32_t* iters, double* errors,
                                uint8_t* spd_failures);
int ssb200_irls_solve_batch_dev_f32(void* handle, const float* dY, size_t batch, size_t ldy, float tolerance,
                                    uint32_t max_iterations, float* dX, size_t ldx, uint32_t* d_iters, double* d_errors,
                                    uint8_t* d_spd_failures, void* stream);
int ssb200_irls_solve_batch_dev_f64(void* handle, const double* dY, size_t batch, size_t ldy, double tolerance,
                                    uint32_t max_iterations, double* dX, size_t ldx, uint32_t* d_iters, double* d_errors,
                                    uint8_t* d_spd_failures, void* stream);

/* frees a homotopy or IRLS handle (solver destructor, include/ss/ss.h:52-55) */
void ssb200_destroy(void* handle);

/* ------------------------------------------------------------------------------------------
 * Utilities next to the path (include/ss/ss.h:79-93, src/lib.cpp:78-112), batched on the device.
 *   reconstruct: Y[b] = A * X[b]          (ss::reconstruct_signal)
 *   norm_l1:     A[:, j] /= sum_i |A[i, j]|  in place on the host array (ss::norm_l1)
 */
int ssb200_reconstruct_batch_f32(void* homotopy_handle, const float* X, size_t batch, size_t ldx, float* Y, size_t ldy);
int ssb200_reconstruct_batch_f64(void* homotopy_handle, const double* X, size_t batch, size_t ldx, double* Y, size_t ldy);
/* the same with DEVICE pointers (dX: batch x N with leading dimension ldx, dY: batch x M), enqueued on `stream` and synchronised */
int ssb200_reconstruct_batch_dev_f32(void* homotopy_handle, const float* dX, size_t batch, size_t ldx, float* dY, size_t ldy, void* stream);
int ssb200_reconstruct_batch_dev_f64(void* homotopy_handle, const double* dX, size_t batch, size_t ldx, double* dY, size_t ldy, void* stream);
/* sparse form of a dense result (new): for every row of dX its non-zero entries in index order as
 * (d_idx[b*cap + k], d_val[b*cap + k]), k < min(d_count[b], cap); unused slots hold (-1, 0).  d_count[b] is the number of
 * non-zeros found and may exceed cap (the row was cut).  A homotopy solution has at most min(max_iterations + 1, N)
 * non-zeros, so cap = that bound never cuts.  All pointers are DEVICE pointers on `device`. */
int ssb200_compress_rows_dev_f32(const float* dX, size_t batch, size_t n, size_t ldx, size_t cap, int32_t* d_idx, float* d_val,
                                 int32_t* d_count, int device, void* stream);
int ssb200_compress_rows_dev_f64(const double* dX, size_t batch, size_t n, size_t ldx, size_t cap, int32_t* d_idx, double* d_val,
                                 int32_t* d_count, int device, void* stream);
int ssb200_norm_l1_f32(float* A, size_t m, size_t n, ptrdiff_t row_stride, ptrdiff_t col_stride, int device);
int ssb200_norm_l1_f64(double* A, size_t m, size_t n, ptrdiff_t row_stride, ptrdiff_t col_stride, int device);
/* ss::norm_l1 on a DEVICE matrix (dA: m x n row-major with leading dimension lda, normalised in place), allocating NOTHING.
 * No counterpart in the reference: added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it).  The only scratch -- the n column sums -- is a
 * static __device__ array of SSB200_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until the process exits, so
 * n * sizeof(element) may not exceed it (n <= 262144 float32 / 131072 float64 columns): larger matrices are refused with
 * SSB200_INVALID_ARGUMENT and the message names the limit.  The call makes no cudaMalloc / cudaFree / allocating library
 * call; it enqueues two kernels on `stream` (a cudaStream_t of `device`, NULL = the default stream) and returns WITHOUT
 * synchronising.  All calls on one GPU share the scratch array: calls that overlap on the GPU race on it -- keep them on one
 * stream or order them with events.  Nothing in the array has to start at zero (every slot read was written by the same
 * call).  The column sums are formed in the row order of ssb200_norm_l1_*, so both entries give the same values. */
#define SSB200_NOALLOC_SCRATCH_BYTES (1 << 20)
int ssb200_norm_l1_dev_noalloc_f32(float* dA, size_t m, size_t n, size_t lda, int device, void* stream);
int ssb200_norm_l1_dev_noalloc_f64(double* dA, size_t m, size_t n, size_t lda, int device, void* stream);

/* ------------------------------------------------------------------------------------------
 * Instrumentation (bench.py / tests; not part of the reference API).
 *   counters[0] = kernel launches since the last reset
 *   counters[1] = launches of the contraction kernel (Q = A^T P)
 *   counters[2] = signal-rows pushed through the contraction kernel
 *   counters[3] = lockstep iterations executed
 *   timing (when enabled the contraction launches are bracketed by CUDA events on the handle's
 *   stream):  ms[0] = summed contraction time, ms[1] = summed step-kernel time
 *   path: 0 = choose by batch size (product behaviour), 1 = no tcgen05 (float32: register-tiled FFMA; float64: the DMMA kernel),
 *         2 = force the tcgen05 contraction (float32 only; fp16 operand containers, the default flavour),
 *         3 = force the tcgen05 contraction with tf32 operand containers (the literal 3xTF32 scheme),
 *         4 = plain FFMA / DFMA tiles for both dtypes (the checker of the tensor-core kernels)
 *   float64 batches of more than 4 signals run on the FP64 tensor pipe (mma.sync.m8n8k4.f64) by default
 */
int ssb200_reset_counters(void* handle);
int ssb200_get_counters(void* handle, uint64_t counters[4]);
int ssb200_set_timing(void* handle, int enabled);
int ssb200_get_timing(void* handle, double ms[2]);
/* per-kernel-class profile of the homotopy lockstep loop (timing enabled): summed device time and the signal-rows processed
 *   class 0 contraction, 1 all step kernels, 2 scan kernel with Q streamed from HBM, 3 scan kernel with q formed from the
 *   Gram matrix, 4 algebra kernel, 5 monolithic step / init kernels */
int ssb200_get_kernel_profile(void* handle, double ms[6], uint64_t rows[6]);
/* algorithmic HBM bytes of the launches of classes 2-5 above (c read + written, q read when streamed, the K x K inverse read +
 * written with K bounded by the iteration number, the next right-hand side written), summed like the profile's rows */
int ssb200_get_kernel_bytes(void* handle, uint64_t bytes[8]);   /* [6] = L2 bytes of the Gram rows read by class 3, [7] unused */
/* bytes[0] = host->device, bytes[1] = device->host bytes the host-pointer batch entries copied since the last reset */
int ssb200_get_transfer_bytes(void* handle, uint64_t bytes[2]);
int ssb200_set_contraction_path(void* handle, int path);
/* measured FP64 peak of `device` in TFLOP/s (bench.py's roofline denominator for the float64 paths): kind 0 = independent DFMA
 * chains on every SM, kind 1 = independent mma.sync.m8n8k4.f64 (DMMA) tiles on every SM */
int ssb200_fp64_peak_probe(int device, int kind, double* tflops);
/* test / bench hook for the contraction kernel alone: Q[r][n] = sum_m A[m][n] * P[r][m] for `rows` host rows
 * (P: rows x M, Q: rows x N, both dense row-major), computed by path 1 (SIMT) or 2 (tcgen05, float32);
 * *ms_per_launch = mean device time of `reps` back-to-back launches (CUDA events on the handle's stream). */
int ssb200_debug_contract_f32(void* handle, const float* P, size_t rows, float* Q, int path, int reps, double* ms_per_launch);
int ssb200_debug_contract_f64(void* handle, const double* P, size_t rows, double* Q, int path, int reps, double* ms_per_launch);

#ifdef __cplusplus
}
#endif
#endif /* SSB200_H */
