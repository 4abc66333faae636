// Host driver of the batched homotopy solve: owns the device copy of the dictionary and the
// per-chunk workspace, runs the lockstep loop  contraction -> step kernel -> compaction.
// Stands in for homotopy_policy::run + solve_homotopy::op<CPU,T> (src/lib.cpp:30-46,
// src/solvers/homotopy-cpu.cpp:277-297).
#pragma once

#include <sched.h>

#include <algorithm>
#include <atomic>
#include <exception>
#include <future>
#include <cstring>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

#include "contract_simt.cuh"
#include "homotopy_kernels.cuh"
#include "homotopy_split.cuh"
#include "homotopy_single.cuh"
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

// fp16 containers of the dictionary: every atom row times its own power of two, hi = fp16(a s), lo = fp16((a s - hi) 2^11);
// a_inv[n] = 1 / s_n undoes the scale in the contraction epilogue; amax_bits collects max |A| (positive floats order as
// unsigned integers).  One warp per atom.
__global__ void f16_split_kernel(const float* __restrict__ src, __half* __restrict__ hi, __half* __restrict__ lo,
                                 float* __restrict__ a_inv, unsigned int* __restrict__ amax_bits, int n_rows, int mp)
{
    const int lane = threadIdx.x & 31;
    for (int n = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); n < n_rows; n += gridDim.x * (blockDim.x >> 5)) {
        const float* row = src + (size_t)n * mp;
        float mx = 0.f;
        for (int m = lane; m < mp; m += 32) mx = fmaxf(mx, fabsf(row[m]));
        mx = warp_max(mx);
        const int e = f16_scale_exp(mx);
        const float sc = ldexpf(1.f, e);
        if (lane == 0) { a_inv[n] = ldexpf(1.f, -e); if (mx < FLT_MAX) atomicMax(amax_bits, __float_as_uint(mx)); }
        for (int m = lane; m < mp; m += 32) {
            const float v = row[m] * sc;
            const __half h = __float2half_rn(v);
            hi[(size_t)n * mp + m] = h;
            lo[(size_t)n * mp + m] = __float2half_rn((v - __half2float(h)) * 2048.0f);
        }
    }
}

// mutual coherence of the dictionary from its Gram matrix: max_{i != j} |G_ij| / sqrt(G_ii G_jj) (positive floats order as
// unsigned integers).  Decides the residual mode and whether q may be formed from rows of G (Homotopy::exact_residual).
template <typename T>
__global__ void coherence_kernel(const T* __restrict__ G, int ldg, int N, unsigned int* __restrict__ mu_bits)
{
    float mx = 0.f;
    const size_t total = (size_t)N * N;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const int i = (int)(e / N), j = (int)(e % N);
        if (i == j) continue;
        const double gii = (double)G[(size_t)i * ldg + i], gjj = (double)G[(size_t)j * ldg + j];
        if (gii > 0.0 && gjj > 0.0) {
            const float v = (float)(fabs((double)G[(size_t)i * ldg + j]) / sqrt(gii * gjj));
            if (v < FLT_MAX) mx = fmaxf(mx, v);
        }
    }
    mx = warp_max(mx);
    if ((threadIdx.x & 31) == 0 && mx > 0.f) atomicMax(mu_bits, __float_as_uint(mx));
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

// Row compaction for the sparse result form (X dense is N values per signal, its support at most
// max_iterations + 1): one warp per signal walks its row, keeps the non-zero entries in index order as (index, value)
// pairs, at most `cap` of them; count[b] is the number found (may exceed cap: the caller then knows the row was cut).
template <typename T>
__global__ void __launch_bounds__(128)
compress_rows_kernel(const T* __restrict__ X, size_t ldx, int N, size_t batch, int cap,
                     int* __restrict__ idx, T* __restrict__ val, int* __restrict__ count)
{
    const size_t b = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (b >= batch) return;
    const int lane = threadIdx.x & 31;
    const T* x = X + b * ldx;
    int found = 0;
    for (int n0 = 0; n0 < N; n0 += 32) {
        const int n = n0 + lane;
        const T v = n < N ? x[n] : T(0);
        const unsigned m = __ballot_sync(0xffffffffu, v != T(0));
        const int slot = found + __popc(m & ((1u << lane) - 1u));
        if (v != T(0) && slot < cap) { idx[b * cap + slot] = n; val[b * cap + slot] = v; }
        found += __popc(m);
    }
    for (int k = found + lane; k < cap; k += 32) { idx[b * cap + k] = -1; val[b * cap + k] = T(0); }
    if (lane == 0) count[b] = found;
}

template <typename T>
class Homotopy : public SolverBase {
public:
    static constexpr int STEP_THREADS = 128;
    static constexpr int MAX_LANES = 2;
    // the lockstep loop runs at most MAX_RUN_AHEAD iterations ahead of the newest signal count the host has seen; the
    // per-iteration counts live in a ring of that many slots + 2 (any max_iterations, no per-call allocation)
    static constexpr uint32_t MAX_RUN_AHEAD = 64, COUNT_RING = MAX_RUN_AHEAD + 2;

    Homotopy(int dev)
        : SolverBase(sizeof(T) == 4 ? SolverKind::HomotopyF32 : SolverKind::HomotopyF64, dev) {}
    ~Homotopy() override { lanes_.clear(); }

    int M = 0, N = 0, Mp = 0, Npad = 0;

    // A on the host, element strides (any layout the reference's ndspan can describe)
    void set_dictionary_host(const T* A, size_t m, size_t n, ptrdiff_t rs, ptrdiff_t cs) {
        check_shape(m, n);
        // repack on the host into atom-major, then one contiguous upload
        std::vector<T> packed((size_t)Npad * Mp, T(0));
        for (size_t j = 0; j < n; ++j)
            for (size_t i = 0; i < m; ++i)
                packed[j * Mp + i] = A[(ptrdiff_t)i * rs + (ptrdiff_t)j * cs];
        alloc_dictionary();
        SSB_CUDA(cudaMemcpyAsync(At_, packed.data(), packed.size() * sizeof(T), cudaMemcpyHostToDevice, stream));
        SSB_CUDA(cudaStreamSynchronize(stream));
        finish_dictionary();
    }
    // A already on this device
    void set_dictionary_dev(const T* dA, size_t m, size_t n, ptrdiff_t rs, ptrdiff_t cs) {
        check_shape(m, n);
        // dA may have been produced on any stream of this device (an NCCL broadcast on torch's stream, a pageable upload):
        // the create call has no stream argument, so everything queued on the device so far is waited for
        SSB_CUDA(cudaDeviceSynchronize());
        alloc_dictionary();
        pack_dictionary_kernel<T><<<1024, 256, 0, stream>>>(dA, (long long)rs, (long long)cs, M, N, At_, Mp, Npad);
        SSB_CUDA(cudaGetLastError());
        counters[0]++;
        finish_dictionary();
    }

    void solve_batch_dev(const T* dY, size_t batch, size_t ldy, T tol, uint32_t max_iter, T* dX, size_t ldx,
                         uint32_t* d_iters, double* d_errs, cudaStream_t user_stream);
    // sparse form of the result on the host: row b's non-zeros in index order as (idx[b * cap + k], val[b * cap + k]), k < count[b]
    struct SparseOut { int32_t* idx; T* val; int32_t* count; size_t cap; };
    // X may be null when `sp` is given: the coefficients then come back in the sparse form only (no dense rows are written)
    void solve_batch_host(const T* Y, size_t batch, size_t ldy, T tol, uint32_t max_iter, T* X, size_t ldx,
                          uint32_t* iters, double* errs, const SparseOut* sp = nullptr);
    void reconstruct_host(const T* X, size_t batch, size_t ldx, T* Y, size_t ldy);
    void reconstruct_dev(const T* dX, size_t batch, size_t ldx, T* dY, size_t ldy, cudaStream_t st);
    // test / bench hook: Q[r][n] = sum_m A[m][n] P[r][m] through the chosen contraction kernel, `reps` launches;
    // returns the mean device time per launch in ms
    double debug_contract(const T* P, size_t rows, T* Q, int path, int reps);

    // How the residual correlation c = A^T (y - A x) (homotopy-cpu.cpp:87-98, re-formed at :255 every iteration) is kept:
    //   exact        r = y - A_G x_G re-formed every iteration and contracted next to the direction (kernel header of
    //                homotopy_kernels.cuh): the reference's own numerics, two contraction rows per signal-iteration
    //   incremental  c -= gamma q: one contraction row per signal-iteration; drifts on coherent dictionaries
    // mode 0 (auto) takes the exact form whenever the dictionary's mutual coherence exceeds COHERENCE_LIMIT -- the regime
    // where the incremental form was measured to leave the reference's path (tests/studies, profiles/r2_parity.json);
    // SSB200_RESIDUAL=exact|incremental overrides an `auto` handle (A/B runs).
    static constexpr float COHERENCE_LIMIT = 0.7f;
    int residual_mode = 0;               // 0 auto, 1 exact, 2 incremental
    float coherence() const { return coherence_; }
    bool coherent() const { return !(coherence_ <= COHERENCE_LIMIT); }
    bool exact_residual() const {
        int mode = residual_mode;
        if (mode == 0) if (const char* e = getenv("SSB200_RESIDUAL")) mode = e[0] == 'e' ? 1 : e[0] == 'i' ? 2 : 0;
        return mode == 1 || (mode == 0 && coherent());
    }

private:
    // One lane = one sub-batch in flight: its own stream and workspace.  Two lanes run concurrently so that one
    // lane's tensor-bound contraction overlaps the other's HBM/latency-bound step kernel (and, on the host-pointer
    // path, its H2D / D2H copies).
    struct Lane {
        cudaStream_t stream = nullptr;
        dev_buffer C, C0, Q, P[2], Plo[2], Pinv[2], gam, dG, Ginv, lam, K, pick, mask, act[2], counts, ymax;
        // host-pointer entry: double-buffered staging and the stream its copy-outs run on
        dev_buffer stY[2], stX[2], stIt[2], stErr[2];
        // ... sparse return: the coefficients travel as (index, value) x Kmax + count per signal and are scattered into the
        // caller's zero-filled X by a host task (40 MB instead of 537 MB over PCIe at C2)
        dev_buffer stIdx[2], stVal[2], stCnt[2];
        pinned_buffer hIdx[2], hVal[2], hCnt[2];
        std::future<void> fin[2];
        cudaStream_t copy_out = nullptr;
        cudaEvent_t ev_out[2] = { nullptr, nullptr };
        int toggle = 0;
        pinned_buffer h_count;
        TcPlan* plan[2] = { nullptr, nullptr };
        bool plan_f16 = false;            // container flavour the plans were built for
        size_t ws_rows = 0;
        int ws_kmax = 0;
        int ws_nb = 1;                    // row blocks of P and Q: 2 in exact-residual mode (directions | residuals)
        // instrumentation, merged into the handle after the solve
        uint64_t counters[4] = { 0, 0, 0, 0 };
        uint64_t rows[6] = { 0, 0, 0, 0, 0, 0 };   // signal-rows per launch class (SolverBase::kernel_rows)
        uint64_t bytes[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };  // algorithmic bytes per launch class (SolverBase::kernel_bytes)
        struct span { cudaEvent_t a, b; int cls; };
        std::vector<span> spans;
        std::vector<cudaEvent_t> pool;
        std::vector<cudaEvent_t> count_events, iter_events;   // one pair per iteration in flight (count read-back)
        cudaStream_t side = nullptr;                          // the read-backs' own stream
        dev_buffer fin_ws;                                    // the finisher's scratch
        Lane() { SSB_CUDA(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking)); }
        ~Lane() {
            drop_plans();
            for (auto e : pool) cudaEventDestroy(e);
            for (auto e : count_events) cudaEventDestroy(e);
            for (auto e : iter_events) cudaEventDestroy(e);
            if (side) cudaStreamDestroy(side);
            for (auto& s : spans) { cudaEventDestroy(s.a); cudaEventDestroy(s.b); }
            if (copy_out) { cudaStreamDestroy(copy_out); cudaEventDestroy(ev_out[0]); cudaEventDestroy(ev_out[1]); }
            if (stream) cudaStreamDestroy(stream);
        }
        void drop_plans() { for (auto& p : plan) { if (p) tc_plan_destroy(p); p = nullptr; } }
        cudaEvent_t event() {
            if (!pool.empty()) { cudaEvent_t e = pool.back(); pool.pop_back(); return e; }
            cudaEvent_t e; SSB_CUDA(cudaEventCreate(&e)); return e;
        }
        int begin(bool on, int cls) {
            if (!on) return -1;
            span s{ event(), event(), cls };
            SSB_CUDA(cudaEventRecord(s.a, stream));
            spans.push_back(s);
            return (int)spans.size() - 1;
        }
        void end(int id) { if (id >= 0) SSB_CUDA(cudaEventRecord(spans[id].b, stream)); }
        size_t ws_bytes() const { return C.bytes + C0.bytes + Q.bytes + P[0].bytes + P[1].bytes + Plo[0].bytes + Plo[1].bytes + Ginv.bytes; }
    };

    dev_buffer dict_;                 // At | At_hi | At_lo | G in one slab (one L2 persistence window covers it)
    T* At_ = nullptr; T* At_hi_ = nullptr; T* At_lo_ = nullptr; T* G_ = nullptr;
    __half* At16_hi_ = nullptr; __half* At16_lo_ = nullptr; float* a_inv_ = nullptr; unsigned int* amax_bits_ = nullptr;
    float amax_ = 0.f;
    float coherence_ = 0.f;
    std::vector<std::unique_ptr<Lane>> lanes_;
    std::mutex merge_mu_;
    size_t l2_set_aside_ = 0;

    void check_shape(size_t m, size_t n) {
        SSB_REQUIRE(m > 0 && n > 0, "dictionary must be non-empty");
        SSB_REQUIRE(m < (1u << 24) && n < (1u << 26), "dictionary too large");
        SSB_REQUIRE(M == 0 || ((size_t)M == m && (size_t)N == n), "refresh must keep the dictionary shape");
        M = (int)m; N = (int)n;
        Mp = (int)round_up(m, 32);
        Npad = (int)round_up(n, 128);
    }
    void alloc_dictionary() {
        const size_t a = (size_t)Npad * Mp, g = (size_t)Npad * Npad;
        // float32: At | tf32 hi | tf32 lo | fp16 (hi, lo) | G | a_inv | amax ; float64: At | G
        const size_t parts = sizeof(T) == 4 ? 4 : 1;
        const size_t extra = sizeof(T) == 4 ? (size_t)Npad + 4 : 4;   // + a_inv (float32) + amax / coherence bits
        dict_.reserve((parts * a + g + extra) * sizeof(T));
        At_ = dict_.as<T>();
        At_hi_ = sizeof(T) == 4 ? At_ + a : nullptr;
        At_lo_ = sizeof(T) == 4 ? At_ + 2 * a : nullptr;
        G_ = At_ + parts * a;
        if (sizeof(T) == 4) {
            At16_hi_ = reinterpret_cast<__half*>(At_ + 3 * a);
            At16_lo_ = At16_hi_ + a;
            a_inv_ = reinterpret_cast<float*>(G_ + g);
            amax_bits_ = reinterpret_cast<unsigned int*>(a_inv_ + Npad);
        } else {
            amax_bits_ = reinterpret_cast<unsigned int*>(G_ + g);
        }
    }
    void finish_dictionary();
    Lane& lane(int i) {
        while ((int)lanes_.size() <= i) lanes_.emplace_back(new Lane());
        return *lanes_[i];
    }
    void apply_l2_policy(cudaStream_t st);

    struct plan_key {
        size_t batch; int kmax; bool host_path; int env_lanes, env_chunk; bool exact, timing;
        bool operator==(const plan_key& o) const {
            return batch == o.batch && kmax == o.kmax && host_path == o.host_path && env_lanes == o.env_lanes && env_chunk == o.env_chunk &&
                   exact == o.exact && timing == o.timing;
        }
    };
    plan_key plan_key_{};
    size_t plan_chunk_ = 0;
    int plan_lanes_ = 0;
    bool plan_valid_ = false;

    size_t bytes_per_signal(int kmax) const {   // exact mode: Q and the right-hand sides twice, c0 kept
        const size_t nb = exact_residual() ? 2 : 1;
        return (size_t)((1 + nb + (nb - 1)) * Npad + 4 * nb * Mp + (size_t)kmax * kmax + 2 * kmax + 4) * sizeof(T) + 6 * sizeof(int);
    }
    // signals per chunk and number of lanes for a batch
    void plan_batch(size_t batch, int kmax, bool host_path, size_t& chunk, int& nlanes);
    void reserve_workspace(Lane& L, size_t rows, int kmax);
    static int active_set_capacity(uint32_t max_iter, int n) { return (int)std::min<uint64_t>((uint64_t)max_iter + 1, (uint64_t)n); }
    void check_capacity(int kmax) const;
    bool use_tc(size_t rows) const {
        if (sizeof(T) != 4) return false;
        if (contraction_path == 1 || contraction_path == 4) return false;
        if (!tc_supported(Npad, Mp)) return false;
        if (contraction_path >= 2) return true;
        return rows >= 64;   // below that the tiles are mostly padding and the SIMT kernels win
    }
    // tensor-path containers: fp16 with power-of-two row scales (default), tf32 on request (path 3 / SSB200_TF32=1)
    bool use_f16() const {
        if (contraction_path == 3) return false;
        if (const char* e = getenv("SSB200_TF32")) if (atoi(e) != 0) return false;   // tuning knob
        return true;
    }
    void contract(Lane& L, bool tc, int which, size_t rows, const int* rows_dev = nullptr, bool count_rows = false, int nblk = 1);
    void solve_chunk(Lane& L, const T* dY, size_t rows, size_t ldy, T tol, uint32_t max_iter, T* dX, size_t ldx,
                     uint32_t* d_iters, double* d_errs);
    void merge(Lane& L);
    // the finisher: the last <= FIN_MAX signals of a chunk taken from their lockstep state to the end by the cooperative
    // single-signal kernel in resume mode (homotopy_single.cuh); false = shape not supported, the lockstep loop goes on
    static constexpr int FIN_MAX = 4, FIN_AFTER = 6;
    bool launch_finisher(Lane& L, const HomArgs<T>& a, uint32_t it0, const int* d_act, const int* d_count);
    std::mutex coop_mu_;                 // one cooperative grid at a time per handle (two lanes must not wait on each other's CTAs)
    // one signal: the whole solve in one cooperative kernel (homotopy_single.cuh); false = shape not supported
    bool solve_single(const T* y, bool y_on_host, T tol, uint32_t max_iter, T* x, bool x_on_host, uint32_t* iters, double* errs,
                      bool reports_on_host);
    // the same kernel as thread-block clusters (one cluster per signal, `nclusters` signals at a time): fills a.cnt and launches on
    // `st`; false = does not apply (dictionary too large to stream through one cluster, device refuses the cluster, SSB200_SINGLE_CLUSTER=0)
    bool launch_single_cluster(SingleArgs<T>& a, int nclusters, cudaStream_t st);
    dev_buffer single_ws_;
    pinned_buffer single_rep_, single_pin_;
    template <typename F> void run_lanes(int nlanes, size_t nchunks, F&& body);
};

}  // namespace ssb

#include "homotopy_impl.cuh"
