// Host-side handle shared by the homotopy and IRLS drivers: device, stream, counters, event timing.
#pragma once

#include <atomic>
#include <vector>

#include "common.cuh"

namespace ssb {

enum class SolverKind : int { HomotopyF32 = 1, HomotopyF64 = 2, IrlsF32 = 3, IrlsF64 = 4 };

struct device_guard {
    int prev = -1;
    explicit device_guard(int dev) { cudaGetDevice(&prev); if (prev != dev) SSB_CUDA(cudaSetDevice(dev)); else prev = -1; }
    ~device_guard() { if (prev >= 0) cudaSetDevice(prev); }
};

struct SolverBase {
    SolverKind kind;
    int device = 0;
    cudaStream_t stream = nullptr;       // the handle's own stream (host-pointer entry points)
    uint64_t counters[4] = { 0, 0, 0, 0 };
    bool timing = false;
    // event-timed launch classes: 0 contraction, 1 all step kernels, 2 scan (Q streamed), 3 scan (q from the Gram matrix),
    // 4 algebra, 5 monolithic step / init; kernel_rows[cls] = signal-rows those launches processed
    static constexpr int NCLS = 6;
    double ms[NCLS] = { 0.0, 0.0, 0.0, 0.0, 0.0, 0.0 };
    uint64_t kernel_rows[NCLS] = { 0, 0, 0, 0, 0, 0 };
    uint64_t kernel_bytes[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };   // algorithmic HBM bytes of those launches (classes 2-5); [6] = L2 bytes of the Gram rows
    std::atomic<uint64_t> h2d_bytes{ 0 }, d2h_bytes{ 0 };   // bytes the host-pointer batch entries moved over PCIe since the last reset
    int contraction_path = 0;            // 0 auto, 1 no tcgen05 (float32: FFMA tiles; float64: DMMA), 2 tcgen05 (fp16 containers),
                                         // 3 tcgen05 with tf32 containers, 4 plain FFMA / DFMA tiles (the checker of the tensor kernels)

    // pairs of events bracketing launches of class `cls` on the current solve's stream
    struct span { cudaEvent_t a, b; int cls; };
    std::vector<span> spans;
    std::vector<cudaEvent_t> event_pool;

    SolverBase(SolverKind k, int dev) : kind(k), device(dev) {
        SSB_CUDA(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    }
    virtual ~SolverBase() {
        for (auto e : event_pool) cudaEventDestroy(e);
        for (auto& s : spans) { cudaEventDestroy(s.a); cudaEventDestroy(s.b); }
        if (stream) cudaStreamDestroy(stream);
    }

    cudaEvent_t get_event() {
        if (!event_pool.empty()) { cudaEvent_t e = event_pool.back(); event_pool.pop_back(); return e; }
        cudaEvent_t e;
        SSB_CUDA(cudaEventCreate(&e));
        return e;
    }
    // usage: auto t = begin_span(cls, st); launch...; end_span(t, st);
    int begin_span(int cls, cudaStream_t st) {
        if (!timing) return -1;
        span s{ get_event(), get_event(), cls };
        SSB_CUDA(cudaEventRecord(s.a, st));
        spans.push_back(s);
        return (int)spans.size() - 1;
    }
    void end_span(int id, cudaStream_t st) {
        if (id < 0) return;
        SSB_CUDA(cudaEventRecord(spans[id].b, st));
    }
    // call after the stream has been synchronised
    void collect_spans() {
        for (auto& s : spans) {
            float t = 0.f;
            if (cudaEventElapsedTime(&t, s.a, s.b) == cudaSuccess) ms[s.cls] += t;
            event_pool.push_back(s.a);
            event_pool.push_back(s.b);
        }
        spans.clear();
    }
};

// a multi-GPU group of homotopy handles (group.cuh)
struct GroupBase { bool is_f64 = false; virtual ~GroupBase() {} };

// contract_tc.cu: Q[r][n] = sum_m At[n][m] P[r][m] on the tcgen05 pipe, 3xTF32 (hi*hi + hi*lo + lo*hi).
// At_hi/At_lo: [n_pad][mp], P_hi/P_lo: [rows_alloc][mp] with rows_alloc % 256 == 0 (rows beyond
// `rows` are computed but not stored), Q: [rows][ldq].  n_pad % 128 == 0, mp % 32 == 0.
struct TcPlan;   // holds the TMA descriptors of one (At, P, Q) buffer set
// f16: operands are fp16 containers (hi, lo * 2^11) of rows scaled by powers of two, a_inv / p_inv undo the scales in the
// epilogue; otherwise tf32 containers (float arrays) and no scales
TcPlan* tc_plan_create(bool f16, const void* At_hi, const void* At_lo, int n_pad, int mp, const void* P_hi, const void* P_lo,
                       int rows_alloc, const float* a_inv, const float* p_inv);
void tc_plan_destroy(TcPlan*);
// rows is an upper bound when rows_dev (device int) is given: the kernel uses min(rows, *rows_dev).
// nblk > 1: the same `rows` rows of nblk row blocks that start blk_off rows apart in P and in Q (blk_off % 256 == 0) --
// the exact-residual mode contracts the directions' block and the residuals' block in one launch
void tc_contract(TcPlan* plan, int rows, float* Q, int ldq, cudaStream_t stream, const int* rows_dev = nullptr,
                 int nblk = 1, int blk_off = 0);
bool tc_supported(int n_pad, int mp);

}  // namespace ssb
