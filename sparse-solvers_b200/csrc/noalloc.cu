// ssb200_norm_l1_dev_noalloc_{f32,f64} (include/ssb200.h): ss::norm_l1 on a DEVICE matrix without a single allocation.
// No counterpart in the reference: added for callers that cannot call cudaMalloc
// (CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// The existing entry (capi.cu: norm_l1) stages the host matrix in a cudaMalloc'ed buffer.  Here the caller owns the matrix on
// the device and the only scratch -- the N column sums -- lives in a file-scope __device__ array of
// SSB200_NOALLOC_SCRATCH_BYTES, which every GPU that loads the library holds until the process exits.
//   kernel 1  one thread per column adds |A[i][j]| over the rows IN ROW ORDER (the order of capi.cu's norm_l1_kernel, so the
//             sums -- and with them the quotients -- are the same bits) and writes the sum to scratch[j]
//   kernel 2  every element is divided by its column's sum (grid over row blocks x column blocks)
// Every scratch slot kernel 2 reads was written by kernel 1 of the same call, so nothing has to start at zero and the last
// call's contents are harmless.  All calls on one GPU share the array: two calls that overlap on the GPU (different streams,
// different host threads) race on it -- serialise them.
//
// In host code `g_noalloc_scratch` is only a stand-in for the symbol; the device address comes from cudaGetSymbolAddress,
// asked once per device with that device current (scratch_on), and is what the kernels are given.
// The two kernels are enqueued on the caller's stream and NOT synchronised: no allocation, no copy, no host wait.
#include <cmath>
#include <mutex>
#include <string>

#include "solver_base.h"

namespace ssb {

static __device__ __align__(256) unsigned char g_noalloc_scratch[SSB200_NOALLOC_SCRATCH_BYTES];

constexpr int NOALLOC_MAX_DEVICES = 64;

// device address of the scratch array on `device`, which must be the current device
static void* scratch_on(int device)
{
    static std::mutex mu;
    static void* addr[NOALLOC_MAX_DEVICES] = {};
    SSB_REQUIRE(device >= 0 && device < NOALLOC_MAX_DEVICES, "device index out of range");
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[device]) {
        void* p = nullptr;
        SSB_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        SSB_REQUIRE(p != nullptr, "cudaGetSymbolAddress gave a null scratch address");
        addr[device] = p;
    }
    return addr[device];
}

template <typename T>
__global__ void __launch_bounds__(128) noalloc_column_sums_kernel(const T* __restrict__ A, size_t lda, int M, int N, T* __restrict__ sums)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= N) return;
    T s = T(0);
    for (int i = 0; i < M; ++i) s += fabs(A[(size_t)i * lda + j]);
    sums[j] = s;
}

constexpr int NOALLOC_ROWS = 16;   // rows per CTA of the scaling kernel

template <typename T>
__global__ void __launch_bounds__(128) noalloc_scale_columns_kernel(T* __restrict__ A, size_t lda, int M, int N, const T* __restrict__ sums)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= N) return;
    const T s = sums[j];
    const int i0 = blockIdx.y * NOALLOC_ROWS, i1 = min(M, i0 + NOALLOC_ROWS);
    for (int i = i0; i < i1; ++i) A[(size_t)i * lda + j] /= s;
}

template <typename T>
static int norm_l1_dev_noalloc(T* dA, size_t m, size_t n, size_t lda, int device, void* stream)
{
    try {
        SSB_REQUIRE(dA && m > 0 && n > 0, "empty matrix");
        SSB_REQUIRE(lda >= n, "lda is smaller than the row length");
        SSB_REQUIRE(m <= (size_t)INT32_MAX && n <= (size_t)INT32_MAX, "matrix dimensions exceed int32");
        if (n * sizeof(T) > (size_t)SSB200_NOALLOC_SCRATCH_BYTES)
            throw invalid_argument{ "ssb200_norm_l1_dev_noalloc: " + std::to_string(n) + " columns need " + std::to_string(n * sizeof(T)) +
                                    " bytes of scratch, the static array holds SSB200_NOALLOC_SCRATCH_BYTES = " +
                                    std::to_string((size_t)SSB200_NOALLOC_SCRATCH_BYTES) + " (use ssb200_norm_l1_* for larger matrices)" };
        const size_t row_blocks = (m + NOALLOC_ROWS - 1) / NOALLOC_ROWS;
        SSB_REQUIRE(row_blocks <= 65535, "more than 65535 * 16 rows");
        device_guard g(device);
        T* sums = static_cast<T*>(scratch_on(device));
        cudaStream_t s = (cudaStream_t)stream;
        const unsigned col_blocks = (unsigned)((n + 127) / 128);
        noalloc_column_sums_kernel<T><<<col_blocks, 128, 0, s>>>(dA, lda, (int)m, (int)n, sums);
        SSB_CUDA(cudaGetLastError());
        noalloc_scale_columns_kernel<T><<<dim3(col_blocks, (unsigned)row_blocks), 128, 0, s>>>(dA, lda, (int)m, (int)n, sums);
        SSB_CUDA(cudaGetLastError());
        return SSB200_OK;
    } catch (const cuda_failure& e) {
        set_last_error(std::string(cudaGetErrorString(e.err)) + " in `" + e.what + "` at " + e.file + ":" + std::to_string(e.line));
        cudaGetLastError();
        return (e.err == cudaErrorNoDevice || e.err == cudaErrorInsufficientDriver || e.err == cudaErrorInvalidDevice)
                   ? SSB200_COMPUTE_MODE_UNAVAILABLE : SSB200_KERNEL_FAILED;
    } catch (const invalid_argument& e) {
        set_last_error(e.what);
        return SSB200_INVALID_ARGUMENT;
    } catch (...) {
        set_last_error("unknown failure");
        return SSB200_KERNEL_FAILED;
    }
}

}  // namespace ssb

extern "C" {

int ssb200_norm_l1_dev_noalloc_f32(float* dA, size_t m, size_t n, size_t lda, int device, void* stream)
{ return ssb::norm_l1_dev_noalloc<float>(dA, m, n, lda, device, stream); }

int ssb200_norm_l1_dev_noalloc_f64(double* dA, size_t m, size_t n, size_t lda, int device, void* stream)
{ return ssb::norm_l1_dev_noalloc<double>(dA, m, n, lda, device, stream); }

}  // extern "C"
