// Host driver of the batched homotopy solve: owns the device copy of the dictionary and the
// per-chunk workspace, runs the lockstep loop  contraction -> step kernel -> compaction.
// Stands in for homotopy_policy::run + solve_homotopy::op<CPU,T> (src/lib.cpp:30-46,
// src/solvers/homotopy-cpu.cpp:277-297).
#pragma once

#include <algorithm>
#include <vector>

#include "contract_simt.cuh"
#include "homotopy_kernels.cuh"
#include "solver_base.h"

namespace ssb {

// dA (device, element strides rs/cs) -> atom-major At[Npad][Mp], zero padded
template <typename T>
__global__ void pack_dictionary_kernel(const T* __restrict__ A, long long rs, long long cs, int M, int N,
                                       T* __restrict__ At, int Mp, int Npad)
{
    const size_t total = (size_t)Npad * Mp;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const int n = (int)(e / Mp), m = (int)(e % Mp);
        At[e] = (n < N && m < M) ? A[(long long)m * rs + (long long)n * cs] : T(0);
    }
}

__global__ void tf32_split_kernel(const float* __restrict__ src, float* __restrict__ hi, float* __restrict__ lo, size_t n)
{
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (size_t)gridDim.x * blockDim.x) {
        const float v = src[e];
        const float h = tf32_rna(v);
        hi[e] = h;
        lo[e] = tf32_rna(v - h);
    }
}

// Y[b][m] = sum_n A[m][n] X[b][n]   (ss::reconstruct_signal, src/lib.cpp:78-104), one CTA per signal
template <typename T>
__global__ void __launch_bounds__(128)
reconstruct_kernel(const T* __restrict__ At, int Mp, int M, int N, const T* __restrict__ X, size_t ldx,
                   T* __restrict__ Y, size_t ldy)
{
    const size_t b = blockIdx.x;
    const T* x = X + b * ldx;
    for (int m = threadIdx.x; m < M; m += blockDim.x) {
        T acc = T(0);
        for (int n = 0; n < N; ++n) {
            const T xv = x[n];
            if (xv != T(0)) acc += At[(size_t)n * Mp + m] * xv;
        }
        Y[b * ldy + m] = acc;
    }
}

template <typename T>
class Homotopy : public SolverBase {
public:
    static constexpr int STEP_THREADS = 128;

    Homotopy(int dev)
        : SolverBase(sizeof(T) == 4 ? SolverKind::HomotopyF32 : SolverKind::HomotopyF64, dev) {}
    ~Homotopy() override { if (plan_[0]) tc_plan_destroy(plan_[0]); if (plan_[1]) tc_plan_destroy(plan_[1]); if (plan_y_) tc_plan_destroy(plan_y_); }

    int M = 0, N = 0, Mp = 0, Npad = 0;

    // A on the host, element strides (any layout the reference's ndspan can describe)
    void set_dictionary_host(const T* A, size_t m, size_t n, ptrdiff_t rs, ptrdiff_t cs) {
        check_shape(m, n);
        // repack on the host into atom-major, then one contiguous upload
        std::vector<T> packed((size_t)Npad * Mp, T(0));
        for (size_t j = 0; j < n; ++j)
            for (size_t i = 0; i < m; ++i)
                packed[j * Mp + i] = A[(ptrdiff_t)i * rs + (ptrdiff_t)j * cs];
        At_.reserve(packed.size() * sizeof(T));
        SSB_CUDA(cudaMemcpyAsync(At_.p, packed.data(), packed.size() * sizeof(T), cudaMemcpyHostToDevice, stream));
        SSB_CUDA(cudaStreamSynchronize(stream));
        finish_dictionary();
    }
    // A already on this device
    void set_dictionary_dev(const T* dA, size_t m, size_t n, ptrdiff_t rs, ptrdiff_t cs) {
        check_shape(m, n);
        At_.reserve((size_t)Npad * Mp * sizeof(T));
        pack_dictionary_kernel<T><<<1024, 256, 0, stream>>>(dA, (long long)rs, (long long)cs, M, N, At_.as<T>(), Mp, Npad);
        SSB_CUDA(cudaGetLastError());
        counters[0]++;
        finish_dictionary();
    }

    void solve_batch_dev(const T* dY, size_t batch, size_t ldy, T tol, uint32_t max_iter, T* dX, size_t ldx,
                         uint32_t* d_iters, double* d_errs, cudaStream_t st);
    void solve_batch_host(const T* Y, size_t batch, size_t ldy, T tol, uint32_t max_iter, T* X, size_t ldx,
                          uint32_t* iters, double* errs);
    void reconstruct_host(const T* X, size_t batch, size_t ldx, T* Y, size_t ldy);
    // test / bench hook: Q[r][n] = sum_m A[m][n] P[r][m] through the chosen contraction kernel, `reps` launches;
    // returns the mean device time per launch in ms
    double debug_contract(const T* P, size_t rows, T* Q, int path, int reps);

private:
    dev_buffer At_, At_hi_, At_lo_, G_;
    // per-chunk workspace
    dev_buffer C_, Q_, P_[2], Plo_[2], gam_, dG_, Ginv_, lam_, K_, act_[2], counts_;
    dev_buffer stY_, stX_, stIt_, stErr_;   // staging for the host-pointer entry points
    pinned_buffer h_count_;
    TcPlan* plan_[2] = { nullptr, nullptr };
    TcPlan* plan_y_ = nullptr;
    size_t ws_rows_ = 0;   // rows the workspace was sized for
    int ws_kmax_ = 0;

    void check_shape(size_t m, size_t n) {
        SSB_REQUIRE(m > 0 && n > 0, "dictionary must be non-empty");
        SSB_REQUIRE(m < (1u << 24) && n < (1u << 26), "dictionary too large");
        SSB_REQUIRE(M == 0 || ((size_t)M == m && (size_t)N == n), "refresh must keep the dictionary shape");
        M = (int)m; N = (int)n;
        Mp = (int)round_up(m, 32);
        Npad = (int)round_up(n, 128);
    }

    void finish_dictionary() {
        // Gram matrix G = A^T A, [N][Npad]: the contraction with P := At
        G_.reserve((size_t)Npad * Npad * sizeof(T));
        counters[0] += launch_contract_simt<T>(At_.as<T>(), Mp, Npad, Mp, At_.as<T>(), Mp, Npad, G_.as<T>(), Npad, stream);
        if constexpr (sizeof(T) == 4) {
            At_hi_.reserve((size_t)Npad * Mp * sizeof(T));
            At_lo_.reserve((size_t)Npad * Mp * sizeof(T));
            tf32_split_kernel<<<1024, 256, 0, stream>>>(At_.as<float>(), At_hi_.as<float>(), At_lo_.as<float>(), (size_t)Npad * Mp);
            SSB_CUDA(cudaGetLastError());
            counters[0]++;
        }
        SSB_CUDA(cudaStreamSynchronize(stream));
        drop_plans();
    }
    void drop_plans() {
        for (auto& p : plan_) { if (p) tc_plan_destroy(p); p = nullptr; }
        if (plan_y_) { tc_plan_destroy(plan_y_); plan_y_ = nullptr; }
    }

    size_t bytes_per_signal(int kmax) const {
        return (size_t)(2 * Npad + 4 * Mp + (size_t)kmax * kmax + 2 * kmax + 4) * sizeof(T) + 6 * sizeof(int);
    }
    size_t choose_chunk(size_t batch, int kmax) const {
        size_t free_b = 0, total_b = 0;
        SSB_CUDA(cudaMemGetInfo(&free_b, &total_b));
        const size_t budget = std::min<size_t>((size_t)(0.6 * (double)free_b) + ws_bytes(), (size_t)24 << 30);
        size_t chunk = std::max<size_t>(1, budget / bytes_per_signal(kmax));
        chunk = std::min<size_t>(chunk, 65536);
        if (chunk >= 256) chunk = chunk / 256 * 256;
        return std::min(chunk, batch);
    }
    size_t ws_bytes() const {
        return C_.bytes + Q_.bytes + P_[0].bytes + P_[1].bytes + Plo_[0].bytes + Plo_[1].bytes + Ginv_.bytes;
    }
    void reserve_workspace(size_t rows, int kmax);
    bool use_tc(size_t rows) const {
        if (sizeof(T) != 4) return false;
        if (contraction_path == 1) return false;
        if (!tc_supported(Npad, Mp)) return false;
        if (contraction_path == 2) return true;
        return rows >= 64;   // below that the tiles are mostly padding and the SIMT kernels win
    }
    void contract(bool tc, int which, size_t rows, cudaStream_t st);
    void solve_chunk(const T* dY, size_t rows, size_t ldy, T tol, uint32_t max_iter, int kmax, T* dX, size_t ldx,
                     uint32_t* d_iters, double* d_errs, cudaStream_t st);
};

}  // namespace ssb

#include "homotopy_impl.cuh"
