// extern "C" surface of libssb200 (include/ssb200.h).  Nothing throws across this boundary.
#include <cmath>
#include <cstring>
#include <new>
#include <string>

#include "homotopy.cuh"
#include "irls.cuh"
#include "group.cuh"

namespace ssb {

static thread_local std::string g_last_error;
void set_last_error(const std::string& msg) { g_last_error = msg; }

template <typename F>
static int guarded(F&& f)
{
    try {
        f();
        return SSB200_OK;
    } catch (const cuda_failure& e) {
        set_last_error(std::string(cudaGetErrorString(e.err)) + " in `" + e.what + "` at " + e.file + ":" + std::to_string(e.line));
        cudaGetLastError();   // clear the sticky-less error state
        if (e.err == cudaErrorNoDevice || e.err == cudaErrorInsufficientDriver || e.err == cudaErrorInvalidDevice)
            return SSB200_COMPUTE_MODE_UNAVAILABLE;
        return SSB200_KERNEL_FAILED;
    } catch (const invalid_argument& e) {
        set_last_error(e.what);
        return SSB200_INVALID_ARGUMENT;
    } catch (const std::bad_alloc&) {
        set_last_error("host allocation failed");
        return SSB200_KERNEL_FAILED;
    } catch (...) {
        set_last_error("unknown failure");
        return SSB200_KERNEL_FAILED;
    }
}

// the product has no CPU path and no pre-Blackwell path: anything but sm_100 is "unavailable"
static void require_sm100(int device)
{
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) throw cuda_failure{ e == cudaSuccess ? cudaErrorNoDevice : e, "cudaGetDeviceCount", __FILE__, __LINE__ };
    if (device < 0 || device >= count) throw cuda_failure{ cudaErrorInvalidDevice, "device index", __FILE__, __LINE__ };
    int major = 0;
    SSB_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
    if (major != 10) throw cuda_failure{ cudaErrorInvalidDevice, "device is not sm_100 (B200)", __FILE__, __LINE__ };
}

template <typename T>
static int homotopy_create(const T* A, size_t m, size_t n, ptrdiff_t rs, ptrdiff_t cs, int device, void** handle, bool on_device)
{
    return guarded([&] {
        SSB_REQUIRE(A && handle, "null pointer");
        *handle = nullptr;
        require_sm100(device);
        device_guard g(device);
        auto* h = new Homotopy<T>(device);
        try {
            if (on_device) h->set_dictionary_dev(A, m, n, rs, cs);
            else h->set_dictionary_host(A, m, n, rs, cs);
        } catch (...) { delete h; throw; }
        *handle = static_cast<SolverBase*>(h);
    });
}

template <typename T>
static Homotopy<T>* as_homotopy(void* handle)
{
    SSB_REQUIRE(handle, "null handle");
    auto* b = static_cast<SolverBase*>(handle);
    SSB_REQUIRE(b->kind == (sizeof(T) == 4 ? SolverKind::HomotopyF32 : SolverKind::HomotopyF64), "handle is not a homotopy solver of this dtype");
    return static_cast<Homotopy<T>*>(b);
}

template <typename T>
static Irls<T>* as_irls(void* handle)
{
    SSB_REQUIRE(handle, "null handle");
    auto* b = static_cast<SolverBase*>(handle);
    SSB_REQUIRE(b->kind == (sizeof(T) == 4 ? SolverKind::IrlsF32 : SolverKind::IrlsF64), "handle is not an IRLS solver of this dtype");
    return static_cast<Irls<T>*>(b);
}

// strided single-signal entry: gather y, solve a batch of one, scatter x
template <typename T>
static int homotopy_solve_one(void* handle, const T* y, ptrdiff_t ys, T tol, uint32_t max_iter, T* x, ptrdiff_t xs,
                              uint32_t* iter, double* err)
{
    return guarded([&] {
        auto* h = as_homotopy<T>(handle);
        SSB_REQUIRE(y && x && iter && err, "null pointer");
        device_guard g(h->device);
        std::vector<T> yy(h->M), xx(h->N);
        for (int i = 0; i < h->M; ++i) yy[i] = y[(ptrdiff_t)i * ys];
        h->solve_batch_host(yy.data(), 1, h->M, tol, max_iter, xx.data(), h->N, iter, err);
        for (int j = 0; j < h->N; ++j) x[(ptrdiff_t)j * xs] = xx[j];
    });
}

template <typename T>
static Group<T>* as_group(void* group)
{
    SSB_REQUIRE(group, "null group");
    auto* b = static_cast<GroupBase*>(group);
    SSB_REQUIRE(b->is_f64 == (sizeof(T) == 8), "group is not of this dtype");
    return static_cast<Group<T>*>(b);
}

template <typename T>
static int group_create(const T* A, size_t m, size_t n, ptrdiff_t rs, ptrdiff_t cs, const int* devices, int ndev, void** group)
{
    return guarded([&] {
        SSB_REQUIRE(A && devices && group, "null pointer");
        *group = nullptr;
        SSB_REQUIRE(ndev >= 1, "a group needs at least one device");
        for (int i = 0; i < ndev; ++i) require_sm100(devices[i]);
        int prev = -1;
        cudaGetDevice(&prev);
        auto* g = new Group<T>();
        try { g->create(A, m, n, rs, cs, devices, ndev); } catch (...) { delete g; if (prev >= 0) cudaSetDevice(prev); throw; }
        if (prev >= 0) cudaSetDevice(prev);
        *group = static_cast<GroupBase*>(g);
    });
}

template <typename T>
static int irls_create(const T* A, size_t m, size_t n, ptrdiff_t rs, ptrdiff_t cs, int device, void** handle)
{
    return guarded([&] {
        SSB_REQUIRE(A && handle, "null pointer");
        *handle = nullptr;
        require_sm100(device);
        device_guard g(device);
        auto* h = new Irls<T>(device);
        try { h->set_dictionary_host(A, m, n, rs, cs); } catch (...) { delete h; throw; }
        *handle = static_cast<SolverBase*>(h);
    });
}

template <typename T>
static int irls_solve_one(void* handle, const T* y, ptrdiff_t ys, T tol, uint32_t max_iter, T* x, ptrdiff_t xs,
                          uint32_t* iter, double* err, uint8_t* spd)
{
    return guarded([&] {
        auto* h = as_irls<T>(handle);
        SSB_REQUIRE(y && x && iter && err && spd, "null pointer");
        device_guard g(h->device);
        std::vector<T> yy(h->M), xx(h->N);
        for (int i = 0; i < h->M; ++i) yy[i] = y[(ptrdiff_t)i * ys];
        h->solve_batch_host(yy.data(), 1, h->M, tol, max_iter, xx.data(), h->N, iter, err, spd);
        for (int j = 0; j < h->N; ++j) x[(ptrdiff_t)j * xs] = xx[j];
    });
}

// A[:, j] /= sum_i |A[i, j]|   (ss::norm_l1, src/lib.cpp:106-112, src/linalg/norms.h:22-33)
template <typename T>
__global__ void norm_l1_kernel(T* A, int M, int N)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= N) return;
    T s = T(0);
    for (int i = 0; i < M; ++i) s += fabs(A[(size_t)i * N + j]);
    for (int i = 0; i < M; ++i) A[(size_t)i * N + j] /= s;
}

template <typename T>
static int norm_l1(T* A, size_t m, size_t n, ptrdiff_t rs, ptrdiff_t cs, int device)
{
    return guarded([&] {
        SSB_REQUIRE(A && m > 0 && n > 0, "empty matrix");
        require_sm100(device);
        device_guard g(device);
        std::vector<T> dense(m * n);
        for (size_t i = 0; i < m; ++i) for (size_t j = 0; j < n; ++j) dense[i * n + j] = A[(ptrdiff_t)i * rs + (ptrdiff_t)j * cs];
        dev_buffer d;
        d.reserve(dense.size() * sizeof(T));
        SSB_CUDA(cudaMemcpy(d.p, dense.data(), dense.size() * sizeof(T), cudaMemcpyHostToDevice));
        norm_l1_kernel<T><<<(unsigned)((n + 127) / 128), 128>>>(d.as<T>(), (int)m, (int)n);
        SSB_CUDA(cudaGetLastError());
        SSB_CUDA(cudaMemcpy(dense.data(), d.p, dense.size() * sizeof(T), cudaMemcpyDeviceToHost));
        for (size_t i = 0; i < m; ++i) for (size_t j = 0; j < n; ++j) A[(ptrdiff_t)i * rs + (ptrdiff_t)j * cs] = dense[i * n + j];
    });
}

}  // namespace ssb

using namespace ssb;

extern "C" {

const char* ssb200_status_string(int status)
{
    switch (status) {   // third_party/kernelpp/include/kernelpp/kernel.h:122-133
        case SSB200_OK: return "Success";
        case SSB200_KERNEL_FAILED: return "Kernel Failed";
        case SSB200_COMPUTE_MODE_DISABLED: return "Compute Mode Disabled";
        case SSB200_COMPUTE_MODE_UNAVAILABLE: return "Compute Mode Unavailable";
        case SSB200_KERNEL_NOT_DEFINED: return "Kernel Not Defined";
        case SSB200_CANCELLED: return "Cancelled";
        case SSB200_INVALID_ARGUMENT: return "Invalid Argument";
        default: return "Unknown";
    }
}
const char* ssb200_last_error(void) { return g_last_error.c_str(); }
void ssb200_version(int v[3]) { v[0] = 0; v[1] = 8; v[2] = 8; }   // reference CMakeLists.txt:4
int ssb200_device_count(void)
{
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess) { cudaGetLastError(); return 0; }
    int ok = 0;
    for (int d = 0; d < count; ++d) {
        int major = 0;
        if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, d) == cudaSuccess && major == 10) ok++;
    }
    return ok;
}

#define SSB_HOMOTOPY_API(SFX, T)                                                                                        \
int ssb200_homotopy_create_##SFX(const T* A, size_t m, size_t n, ptrdiff_t rs, ptrdiff_t cs, int device, void** h)      \
{ return homotopy_create<T>(A, m, n, rs, cs, device, h, false); }                                                       \
int ssb200_homotopy_create_dev_##SFX(const T* A, size_t m, size_t n, ptrdiff_t rs, ptrdiff_t cs, int device, void** h)  \
{ return homotopy_create<T>(A, m, n, rs, cs, device, h, true); }                                                        \
int ssb200_homotopy_refresh_##SFX(void* h, const T* A, ptrdiff_t rs, ptrdiff_t cs)                                      \
{ return guarded([&] { auto* s = as_homotopy<T>(h); SSB_REQUIRE(A, "null pointer"); device_guard g(s->device);          \
                       s->set_dictionary_host(A, s->M, s->N, rs, cs); }); }                                             \
int ssb200_homotopy_solve_##SFX(void* h, const T* y, ptrdiff_t ys, T tol, uint32_t mi, T* x, ptrdiff_t xs,              \
                                uint32_t* iter, double* err)                                                            \
{ return homotopy_solve_one<T>(h, y, ys, tol, mi, x, xs, iter, err); }                                                  \
int ssb200_homotopy_solve_batch_##SFX(void* h, const T* Y, size_t batch, size_t ldy, T tol, uint32_t mi, T* X,          \
                                      size_t ldx, uint32_t* iters, double* errs)                                        \
{ return guarded([&] { auto* s = as_homotopy<T>(h); SSB_REQUIRE(batch == 0 || (Y && X && iters && errs), "null pointer"); \
                       device_guard g(s->device); s->solve_batch_host(Y, batch, ldy, tol, mi, X, ldx, iters, errs); }); } \
int ssb200_homotopy_solve_batch_sparse_##SFX(void* h, const T* Y, size_t batch, size_t ldy, T tol, uint32_t mi,        \
                                             int32_t* idx, T* val, int32_t* count, size_t cap, uint32_t* iters, double* errs) \
{ return guarded([&] { auto* s = as_homotopy<T>(h); SSB_REQUIRE(batch == 0 || (Y && idx && val && count && iters && errs), "null pointer"); \
                       device_guard g(s->device); typename Homotopy<T>::SparseOut sp{ idx, val, count, cap };               \
                       s->solve_batch_host(Y, batch, ldy, tol, mi, nullptr, 0, iters, errs, &sp); }); }                     \
int ssb200_homotopy_solve_batch_dev_##SFX(void* h, const T* Y, size_t batch, size_t ldy, T tol, uint32_t mi, T* X,      \
                                          size_t ldx, uint32_t* iters, double* errs, void* stream)                      \
{ return guarded([&] { auto* s = as_homotopy<T>(h); SSB_REQUIRE(batch == 0 || (Y && X && iters && errs), "null pointer"); \
                       device_guard g(s->device);                                                                       \
                       s->solve_batch_dev(Y, batch, ldy, tol, mi, X, ldx, iters, errs, (cudaStream_t)stream); }); }     \
int ssb200_reconstruct_batch_##SFX(void* h, const T* X, size_t batch, size_t ldx, T* Y, size_t ldy)                     \
{ return guarded([&] { auto* s = as_homotopy<T>(h); SSB_REQUIRE(batch == 0 || (X && Y), "null pointer");                \
                       device_guard g(s->device); s->reconstruct_host(X, batch, ldx, Y, ldy); }); }                     \
int ssb200_reconstruct_batch_dev_##SFX(void* h, const T* dX, size_t batch, size_t ldx, T* dY, size_t ldy, void* stream)     \
{ return guarded([&] { auto* s = as_homotopy<T>(h); SSB_REQUIRE(batch == 0 || (dX && dY), "null pointer");               \
                       device_guard g(s->device); s->reconstruct_dev(dX, batch, ldx, dY, ldy, (cudaStream_t)stream); }); } \
int ssb200_compress_rows_dev_##SFX(const T* dX, size_t batch, size_t n, size_t ldx, size_t cap, int32_t* d_idx, T* d_val, \
                                   int32_t* d_count, int device, void* stream)                                          \
{ return guarded([&] { SSB_REQUIRE(batch == 0 || (dX && d_idx && d_val && d_count), "null pointer");                    \
                       SSB_REQUIRE(n > 0 && cap > 0 && ldx >= n && n < ((size_t)1 << 31) && cap < ((size_t)1 << 31), "bad shape"); \
                       require_sm100(device); device_guard g(device);                                                   \
                       if (batch == 0) return;                                                                          \
                       compress_rows_kernel<T><<<(unsigned)((batch + 3) / 4), 128, 0, (cudaStream_t)stream>>>(          \
                           dX, ldx, (int)n, batch, (int)cap, d_idx, d_val, d_count);                                    \
                       SSB_CUDA(cudaGetLastError()); SSB_CUDA(cudaStreamSynchronize((cudaStream_t)stream)); }); }        \
int ssb200_debug_contract_##SFX(void* h, const T* P, size_t rows, T* Q, int path, int reps, double* ms_per_launch)      \
{ return guarded([&] { auto* s = as_homotopy<T>(h); SSB_REQUIRE(P && Q, "null pointer"); device_guard g(s->device);     \
                       double ms = s->debug_contract(P, rows, Q, path, reps); if (ms_per_launch) *ms_per_launch = ms; }); } \
int ssb200_norm_l1_##SFX(T* A, size_t m, size_t n, ptrdiff_t rs, ptrdiff_t cs, int device)                              \
{ return norm_l1<T>(A, m, n, rs, cs, device); }

SSB_HOMOTOPY_API(f32, float)
SSB_HOMOTOPY_API(f64, double)

#define SSB_GROUP_API(SFX, T)                                                                                           \
int ssb200_group_create_##SFX(const T* A, size_t m, size_t n, ptrdiff_t rs, ptrdiff_t cs, const int* devices, int ndev, void** g) \
{ return group_create<T>(A, m, n, rs, cs, devices, ndev, g); }                                                          \
int ssb200_group_solve_batch_##SFX(void* g, const T* Y, size_t batch, size_t ldy, T tol, uint32_t mi, T* X, size_t ldx,  \
                                   uint32_t* iters, double* errs)                                                       \
{ return guarded([&] { auto* s = as_group<T>(g); SSB_REQUIRE(batch == 0 || (Y && X && iters && errs), "null pointer");   \
                       SSB_REQUIRE(mi > 0, "max_iterations must be positive"); if (batch == 0) return;                  \
                       int prev = -1; cudaGetDevice(&prev);                                                             \
                       try { s->solve_batch_host(Y, batch, ldy, tol, mi, X, ldx, iters, errs); }                        \
                       catch (...) { if (prev >= 0) cudaSetDevice(prev); throw; }                                       \
                       if (prev >= 0) cudaSetDevice(prev); }); }                                                        \
int ssb200_group_solve_batch_dev_##SFX(void* g, const T* dY, size_t batch, size_t ldy, T tol, uint32_t mi, T* dX,        \
                                       size_t ldx, uint32_t* d_iters, double* d_errs, void* stream)                     \
{ return guarded([&] { auto* s = as_group<T>(g); SSB_REQUIRE(batch == 0 || (dY && dX && d_iters && d_errs), "null pointer"); \
                       SSB_REQUIRE(mi > 0, "max_iterations must be positive"); if (batch == 0) return;                  \
                       int prev = -1; cudaGetDevice(&prev);                                                             \
                       try { s->solve_batch_dev(dY, batch, ldy, tol, mi, dX, ldx, d_iters, d_errs, (cudaStream_t)stream); } \
                       catch (...) { if (prev >= 0) cudaSetDevice(prev); throw; }                                       \
                       if (prev >= 0) cudaSetDevice(prev); }); }

SSB_GROUP_API(f32, float)
SSB_GROUP_API(f64, double)

int ssb200_group_size(void* g)
{ if (!g) return 0; auto* b = static_cast<GroupBase*>(g);
  return b->is_f64 ? (int)static_cast<Group<double>*>(b)->devices.size() : (int)static_cast<Group<float>*>(b)->devices.size(); }
int ssb200_group_set_residual_mode(void* g, int mode)
{ return guarded([&] { SSB_REQUIRE(g && mode >= 0 && mode <= 2, "bad residual mode"); auto* b = static_cast<GroupBase*>(g);
                       if (b->is_f64) static_cast<Group<double>*>(b)->set_residual_mode(mode);
                       else static_cast<Group<float>*>(b)->set_residual_mode(mode); }); }
void ssb200_group_destroy(void* g)
{
    if (!g) return;
    int prev = -1;
    cudaGetDevice(&prev);
    delete static_cast<GroupBase*>(g);
    if (prev >= 0) cudaSetDevice(prev);
}

#define SSB_IRLS_API(SFX, T)                                                                                            \
int ssb200_irls_create_##SFX(const T* A, size_t m, size_t n, ptrdiff_t rs, ptrdiff_t cs, int device, void** h)          \
{ return irls_create<T>(A, m, n, rs, cs, device, h); }                                                                  \
int ssb200_irls_solve_##SFX(void* h, const T* y, ptrdiff_t ys, T tol, uint32_t mi, T* x, ptrdiff_t xs, uint32_t* iter,  \
                            double* err, uint8_t* spd)                                                                  \
{ return irls_solve_one<T>(h, y, ys, tol, mi, x, xs, iter, err, spd); }                                                 \
int ssb200_irls_solve_batch_##SFX(void* h, const T* Y, size_t batch, size_t ldy, T tol, uint32_t mi, T* X, size_t ldx,  \
                                  uint32_t* iters, double* errs, uint8_t* spd)                                          \
{ return guarded([&] { auto* s = as_irls<T>(h); SSB_REQUIRE(batch == 0 || (Y && X && iters && errs && spd), "null pointer"); \
                       device_guard g(s->device); s->solve_batch_host(Y, batch, ldy, tol, mi, X, ldx, iters, errs, spd); }); } \
int ssb200_irls_solve_batch_dev_##SFX(void* h, const T* Y, size_t batch, size_t ldy, T tol, uint32_t mi, T* X,          \
                                      size_t ldx, uint32_t* iters, double* errs, uint8_t* spd, void* stream)            \
{ return guarded([&] { auto* s = as_irls<T>(h); SSB_REQUIRE(batch == 0 || (Y && X && iters && errs && spd), "null pointer"); \
                       device_guard g(s->device);                                                                       \
                       s->solve_batch_dev(Y, batch, ldy, tol, mi, X, ldx, iters, errs, spd, (cudaStream_t)stream); }); }

SSB_IRLS_API(f32, float)
SSB_IRLS_API(f64, double)

void ssb200_destroy(void* handle)
{
    if (!handle) return;
    auto* b = static_cast<SolverBase*>(handle);
    int prev = -1;
    cudaGetDevice(&prev);
    cudaSetDevice(b->device);
    delete b;
    if (prev >= 0) cudaSetDevice(prev);
}

int ssb200_reset_counters(void* h)
{ return guarded([&] { SSB_REQUIRE(h, "null handle"); auto* b = static_cast<SolverBase*>(h);
                       for (auto& c : b->counters) c = 0; for (auto& m : b->ms) m = 0.0; for (auto& r : b->kernel_rows) r = 0; for (auto& r : b->kernel_bytes) r = 0; b->h2d_bytes = 0; b->d2h_bytes = 0; }); }
int ssb200_get_counters(void* h, uint64_t counters[4])
{ return guarded([&] { SSB_REQUIRE(h && counters, "null pointer"); auto* b = static_cast<SolverBase*>(h);
                       for (int i = 0; i < 4; ++i) counters[i] = b->counters[i]; }); }
int ssb200_set_timing(void* h, int enabled)
{ return guarded([&] { SSB_REQUIRE(h, "null handle"); static_cast<SolverBase*>(h)->timing = enabled != 0; }); }
int ssb200_get_timing(void* h, double ms[2])
{ return guarded([&] { SSB_REQUIRE(h && ms, "null pointer"); auto* b = static_cast<SolverBase*>(h); ms[0] = b->ms[0]; ms[1] = b->ms[1]; }); }
int ssb200_get_kernel_profile(void* h, double ms[6], uint64_t rows[6])
{ return guarded([&] { SSB_REQUIRE(h && ms && rows, "null pointer"); auto* b = static_cast<SolverBase*>(h);
                       for (int i = 0; i < SolverBase::NCLS; ++i) { ms[i] = b->ms[i]; rows[i] = b->kernel_rows[i]; } }); }
int ssb200_get_kernel_bytes(void* h, uint64_t bytes[8])
{ return guarded([&] { SSB_REQUIRE(h && bytes, "null pointer"); auto* b = static_cast<SolverBase*>(h);
                       for (int i = 0; i < 8; ++i) bytes[i] = b->kernel_bytes[i]; }); }
int ssb200_get_transfer_bytes(void* h, uint64_t bytes[2])
{ return guarded([&] { SSB_REQUIRE(h && bytes, "null pointer"); auto* b = static_cast<SolverBase*>(h);
                       bytes[0] = b->h2d_bytes.load(); bytes[1] = b->d2h_bytes.load(); }); }
int ssb200_homotopy_set_residual_mode(void* h, int mode)
{ return guarded([&] { SSB_REQUIRE(h && mode >= 0 && mode <= 2, "bad residual mode");
                       auto* b = static_cast<SolverBase*>(h);
                       SSB_REQUIRE(b->kind == SolverKind::HomotopyF32 || b->kind == SolverKind::HomotopyF64, "handle is not a homotopy solver");
                       if (b->kind == SolverKind::HomotopyF32) static_cast<Homotopy<float>*>(b)->residual_mode = mode;
                       else static_cast<Homotopy<double>*>(b)->residual_mode = mode; }); }
int ssb200_homotopy_get_residual_info(void* h, int* mode_in_effect, double* coherence)
{ return guarded([&] { SSB_REQUIRE(h, "null handle");
                       auto* b = static_cast<SolverBase*>(h);
                       SSB_REQUIRE(b->kind == SolverKind::HomotopyF32 || b->kind == SolverKind::HomotopyF64, "handle is not a homotopy solver");
                       bool ex; double mu;
                       if (b->kind == SolverKind::HomotopyF32) { auto* s = static_cast<Homotopy<float>*>(b); ex = s->exact_residual(); mu = s->coherence(); }
                       else { auto* s = static_cast<Homotopy<double>*>(b); ex = s->exact_residual(); mu = s->coherence(); }
                       if (mode_in_effect) *mode_in_effect = ex ? SSB200_RESIDUAL_EXACT : SSB200_RESIDUAL_INCREMENTAL;
                       if (coherence) *coherence = mu; }); }
#if SSB_SINGLE_PROF
// variant builds only (tools/single_prof.py): the phase timer of the single-signal kernel, read and cleared
int ssb200_debug_single_profile(uint64_t out[32])
{
    return guarded([&] {
        unsigned long long h[32], z[32] = {};
        SSB_CUDA(cudaMemcpyFromSymbol(h, g_single_prof, sizeof(h)));
        SSB_CUDA(cudaMemcpyToSymbol(g_single_prof, z, sizeof(z)));
        for (int i = 0; i < 32; ++i) out[i] = h[i];
    });
}
#endif

int ssb200_fp64_peak_probe(int device, int kind, double* tflops)
{ return guarded([&] { SSB_REQUIRE(tflops && (kind == 0 || kind == 1), "bad arguments");
                       require_sm100(device); device_guard g(device);
                       int sms = 0; SSB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
                       const int blocks = sms * 8, iters = 1 << 14;
                       dev_buffer out; out.reserve((size_t)blocks * 256 * sizeof(double));
                       cudaEvent_t e0, e1; SSB_CUDA(cudaEventCreate(&e0)); SSB_CUDA(cudaEventCreate(&e1));
                       float best = 1e30f;
                       for (int rep = 0; rep < 4; ++rep) {        // first round warms up; best of the rest
                           SSB_CUDA(cudaEventRecord(e0));
                           if (kind == 0) fp64_dfma_probe_kernel<<<blocks, 256>>>(out.as<double>(), iters, 1e-3);
                           else fp64_dmma_probe_kernel<<<blocks, 256>>>(out.as<double>(), iters, 1e-3);
                           SSB_CUDA(cudaEventRecord(e1)); SSB_CUDA(cudaEventSynchronize(e1));
                           float ms = 0.f; SSB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
                           if (rep > 0) best = std::min(best, ms);
                       }
                       cudaEventDestroy(e0); cudaEventDestroy(e1);
                       // DFMA: 16 chains x 2 flop per thread-iteration; DMMA: 8 tiles x 512 flop per warp-iteration
                       const double flop = kind == 0 ? (double)blocks * 256 * iters * 32.0 : (double)blocks * 8 * iters * 8 * 512.0;
                       *tflops = flop / (best * 1e-3) / 1e12; }); }
int ssb200_set_contraction_path(void* h, int path)
{ return guarded([&] { SSB_REQUIRE(h && path >= 0 && path <= 4, "bad path"); static_cast<SolverBase*>(h)->contraction_path = path; }); }

}  // extern "C"
