// Shared host/device helpers of libssb200 (sm_100a only).
#pragma once

#include <cuda_runtime.h>

#include <atomic>
#include <cfloat>
#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <string>

#include <nvtx3/nvToolsExt.h>   // header-only; a no-op unless a profiler injects itself

#include "../../include/ssb200.h"

namespace ssb {

// ---- error plumbing: CUDA failures become an ssb200_status + a thread-local message ----------
void set_last_error(const std::string& msg);

struct cuda_failure { cudaError_t err; const char* what; const char* file; int line; };

#define SSB_CUDA(expr)                                                                  \
    do {                                                                                \
        cudaError_t _e = (expr);                                                        \
        if (_e != cudaSuccess) throw ::ssb::cuda_failure{ _e, #expr, __FILE__, __LINE__ }; \
    } while (0)

struct invalid_argument { std::string what; };
#define SSB_REQUIRE(cond, msg)                                   \
    do { if (!(cond)) throw ::ssb::invalid_argument{ msg }; } while (0)

// ---- small RAII device buffer --------------------------------------------------------------
struct dev_buffer {
    void* p = nullptr;
    size_t bytes = 0;
    dev_buffer() = default;
    dev_buffer(const dev_buffer&) = delete;
    dev_buffer& operator=(const dev_buffer&) = delete;
    ~dev_buffer() { release(); }
    void release() { if (p) cudaFree(p); p = nullptr; bytes = 0; }
    // grows only; contents are not preserved
    void reserve(size_t nbytes) {
        if (nbytes <= bytes) return;
        release();
        SSB_CUDA(cudaMalloc(&p, nbytes));
        bytes = nbytes;
    }
    template <typename U> U* as() const { return static_cast<U*>(p); }
};

struct pinned_buffer {
    void* p = nullptr;
    size_t bytes = 0;
    ~pinned_buffer() { if (p) cudaFreeHost(p); }
    void reserve(size_t nbytes) {
        if (nbytes <= bytes) return;
        if (p) cudaFreeHost(p);
        p = nullptr; bytes = 0;
        SSB_CUDA(cudaMallocHost(&p, nbytes));
        bytes = nbytes;
    }
    template <typename U> U* as() const { return static_cast<U*>(p); }
};

inline size_t round_up(size_t v, size_t to) { return (v + to - 1) / to * to; }

// NVTX range around the launches of one kernel class (contraction / scan / algebra / step): ncu and nsys captures name themselves
struct nvtx_range {
    explicit nvtx_range(const char* name) { nvtxRangePushA(name); }
    ~nvtx_range() { nvtxRangePop(); }
    nvtx_range(const nvtx_range&) = delete;
    nvtx_range& operator=(const nvtx_range&) = delete;
};

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is a per-DEVICE attribute: one bit per device and kernel instantiation
inline void optin_shared_memory(const void* fn, std::atomic<uint64_t>& done, int bytes = 200 * 1024)
{
    int dev = 0;
    SSB_CUDA(cudaGetDevice(&dev));
    const uint64_t bit = 1ull << (dev & 63);
    if (dev < 64 && (done.load(std::memory_order_acquire) & bit)) return;
    SSB_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    done.fetch_or(bit, std::memory_order_release);
}

template <typename T> struct num_traits;
template <> struct num_traits<float>  { static __host__ __device__ float  max() { return FLT_MAX; } static __host__ __device__ float  eps() { return FLT_EPSILON; } };
template <> struct num_traits<double> { static __host__ __device__ double max() { return DBL_MAX; } static __host__ __device__ double eps() { return DBL_EPSILON; } };

// ---- device reductions ----------------------------------------------------------------------
#ifdef __CUDACC__

// separately rounded multiply / add: keeps nvcc from contracting them into an FMA where the reference's build (x86-64
// baseline, no FMA instruction) rounds twice and an EXACT outcome depends on it
__device__ __forceinline__ float  mul_rn(float a, float b)   { return __fmul_rn(a, b); }
__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ float  add_rn(float a, float b)   { return __fadd_rn(a, b); }
__device__ __forceinline__ double add_rn(double a, double b) { return __dadd_rn(a, b); }

// (value, index) candidates ordered lexicographically: smaller value first, then lower index --
// the left-most strict decrease of the reference's scan (homotopy-cpu.cpp:156-160).
template <typename T>
__device__ __forceinline__ bool cand_less(T va, uint32_t ia, T vb, uint32_t ib) {
    return va < vb || (va == vb && ia < ib);
}

template <typename T>
__device__ __forceinline__ void warp_argmin(T& v, uint32_t& i) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        T ov = __shfl_xor_sync(0xffffffffu, v, o);
        uint32_t oi = __shfl_xor_sync(0xffffffffu, i, o);
        if (cand_less(ov, oi, v, i)) { v = ov; i = oi; }
    }
}

// first index of the largest value (cblas i?amax, homotopy-cpu.cpp:32-37)
template <typename T>
__device__ __forceinline__ void warp_argmax_first(T& v, uint32_t& i) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        T ov = __shfl_xor_sync(0xffffffffu, v, o);
        uint32_t oi = __shfl_xor_sync(0xffffffffu, i, o);
        if (ov > v || (ov == v && oi < i)) { v = ov; i = oi; }
    }
}

template <typename T>
__device__ __forceinline__ T warp_max(T v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// The same reductions for float32 from warp-wide integer reductions (REDUX) instead of five rounds of shuffles: a step candidate is
// strictly positive (or the FLT_MAX sentinel) and a |c| partial is >= 0, so the bits order like the values; results are the same bits.
// C2: the scan kernels 14.7 -> 14.3 ms per step.  float64 keeps the shuffle forms.
#ifndef SSB_SHUFFLE_REDUCE
#define SSB_SHUFFLE_REDUCE 0      // 1: the shuffle forms everywhere (A/B builds, tools/build_variant.sh)
#endif
__device__ __forceinline__ void warp_argmin_fast(float& v, uint32_t& i) {
    if (SSB_SHUFFLE_REDUCE) { warp_argmin(v, i); return; }
    const uint32_t kv = __float_as_uint(v);
    const uint32_t mv = __reduce_min_sync(0xffffffffu, kv);
    i = __reduce_min_sync(0xffffffffu, kv == mv ? i : 0xffffffffu);
    v = __uint_as_float(mv);
}
__device__ __forceinline__ void warp_argmin_fast(double& v, uint32_t& i) { warp_argmin(v, i); }
__device__ __forceinline__ float warp_max_nonneg(float v) { return SSB_SHUFFLE_REDUCE ? warp_max(v) : __uint_as_float(__reduce_max_sync(0xffffffffu, __float_as_uint(v))); }
__device__ __forceinline__ double warp_max_nonneg(double v) { return warp_max(v); }

template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

#endif  // __CUDACC__

}  // namespace ssb
