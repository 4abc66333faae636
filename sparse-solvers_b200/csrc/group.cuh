// Multi-GPU group of homotopy handles in ONE process (SURVEY.md 8(b) last row, 8(e)): one communicator over the GPUs of a
// box (ncclCommInitAll), the dictionary replicated once with ncclBroadcast, the batch cut into contiguous shards, one host
// thread per GPU, results gathered -- and nothing else exchanged (signals are independent; the reference has no counterpart,
// its solve is one signal per call, include/ss/ss.h:50).
//   host-pointer entry    every GPU's thread copies its shard in and out itself: the "gather" is the caller's X array
//   device-pointer entry  Y, X and the reports live on the group's first GPU: shards travel over NVLink with
//                         ncclSend / ncclRecv, the dense coefficient rows come back the same way
// NCCL is bound at run time (dlopen of libnccl.so.2: the copy the process already loaded -- e.g. torch's --, else
// $SSB200_NCCL_LIB, else the system one), so libssb200.so has no link-time dependency on it and single-GPU users never touch
// it.  A Python process that will import torch later should load torch's copy first (sparsesolvers.HomotopyGroup does): two
// NCCL versions share one soname, and whichever is loaded first is the one every later DT_NEEDED resolves to.
#pragma once

#include <dlfcn.h>
#include <nccl.h>

#include <memory>
#include <thread>
#include <vector>

#include "homotopy.cuh"

namespace ssb {

struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;

    static NcclApi& get() {
        static NcclApi api;
        static std::once_flag once;
        std::call_once(once, [] {
            // the copy the process already has (torch's bundled one, when torch came first) > $SSB200_NCCL_LIB > the system one.
            // Never RTLD_GLOBAL: a second NCCL of another version must not shadow the symbols of the first.
            api.lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
            if (!api.lib) if (const char* e = getenv("SSB200_NCCL_LIB")) api.lib = dlopen(e, RTLD_NOW | RTLD_LOCAL);
            for (const char* name : { "libnccl.so.2", "libnccl.so" }) {
                if (api.lib) break;
                api.lib = dlopen(name, RTLD_NOW | RTLD_LOCAL);
            }
            if (!api.lib) return;
            auto sym = [&](const char* n) { return dlsym(api.lib, n); };
            api.CommInitAll = reinterpret_cast<decltype(api.CommInitAll)>(sym("ncclCommInitAll"));
            api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
            api.Broadcast = reinterpret_cast<decltype(api.Broadcast)>(sym("ncclBroadcast"));
            api.Send = reinterpret_cast<decltype(api.Send)>(sym("ncclSend"));
            api.Recv = reinterpret_cast<decltype(api.Recv)>(sym("ncclRecv"));
            api.GroupStart = reinterpret_cast<decltype(api.GroupStart)>(sym("ncclGroupStart"));
            api.GroupEnd = reinterpret_cast<decltype(api.GroupEnd)>(sym("ncclGroupEnd"));
            api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
        });
        return api;
    }
    bool ok() const { return lib && CommInitAll && CommDestroy && Broadcast && Send && Recv && GroupStart && GroupEnd; }
};

#define SSB_NCCL(expr)                                                                                      \
    do {                                                                                                    \
        ncclResult_t _r = (expr);                                                                           \
        if (_r != ncclSuccess) {                                                                            \
            auto& _api = ::ssb::NcclApi::get();                                                              \
            throw ::ssb::invalid_argument{ std::string("NCCL: ") + (_api.GetErrorString ? _api.GetErrorString(_r) : "error") + " in " #expr }; \
        }                                                                                                   \
    } while (0)

// contiguous, balanced split: the first (batch % n) shards get one more signal (sparsesolvers/distributed.py: shard_bounds)
inline void shard_bounds(size_t batch, int n, int i, size_t& lo, size_t& hi) {
    const size_t base = batch / n, rem = batch % n;
    lo = (size_t)i * base + std::min<size_t>((size_t)i, rem);
    hi = lo + base + ((size_t)i < rem ? 1 : 0);
}

template <typename T>
struct Group : GroupBase {
    std::vector<int> devices;
    std::vector<std::unique_ptr<Homotopy<T>>> solvers;
    std::vector<ncclComm_t> comms;
    std::vector<cudaStream_t> streams;
    // device-pointer entry: every non-root GPU's shard buffers
    struct Shard { dev_buffer Y, X, it, err; };
    std::vector<std::unique_ptr<Shard>> shards;
    int M = 0, N = 0;

    Group() { is_f64 = sizeof(T) == 8; }
    ~Group() override {
        auto& api = NcclApi::get();
        for (size_t i = 0; i < devices.size(); ++i) {
            cudaSetDevice(devices[i]);
            if (i < solvers.size()) solvers[i].reset();
            if (i < shards.size()) shards[i].reset();
            if (i < streams.size() && streams[i]) cudaStreamDestroy(streams[i]);
            if (i < comms.size() && comms[i] && api.CommDestroy) api.CommDestroy(comms[i]);
        }
    }

    static constexpr ncclDataType_t nccl_type() { return sizeof(T) == 4 ? ncclFloat32 : ncclFloat64; }

    void create(const T* A, size_t m, size_t n, ptrdiff_t rs, ptrdiff_t cs, const int* devs, int ndev) {
        SSB_REQUIRE(ndev >= 1 && ndev <= 64, "bad device count");
        devices.assign(devs, devs + ndev);
        for (int i = 0; i < ndev; ++i)
            for (int j = 0; j < i; ++j) SSB_REQUIRE(devices[i] != devices[j], "a device may appear once in a group");
        M = (int)m; N = (int)n;
        auto& api = NcclApi::get();
        comms.assign(ndev, nullptr);
        streams.assign(ndev, nullptr);
        for (int i = 0; i < ndev; ++i) shards.emplace_back(new Shard());
        if (ndev > 1) {
            SSB_REQUIRE(api.ok(), "libnccl.so.2 could not be loaded: a multi-GPU group needs NCCL");
            SSB_NCCL(api.CommInitAll(comms.data(), ndev, devices.data()));
        }
        // dense row-major copy of the caller's view on the first GPU, then one broadcast over NVLink
        std::vector<T> dense(m * n);
        for (size_t i = 0; i < m; ++i) for (size_t j = 0; j < n; ++j) dense[i * n + j] = A[(ptrdiff_t)i * rs + (ptrdiff_t)j * cs];
        std::vector<dev_buffer> dA(ndev);
        for (int i = 0; i < ndev; ++i) {
            SSB_CUDA(cudaSetDevice(devices[i]));
            SSB_CUDA(cudaStreamCreateWithFlags(&streams[i], cudaStreamNonBlocking));
            dA[i].reserve(m * n * sizeof(T));
        }
        SSB_CUDA(cudaSetDevice(devices[0]));
        SSB_CUDA(cudaMemcpyAsync(dA[0].p, dense.data(), m * n * sizeof(T), cudaMemcpyHostToDevice, streams[0]));
        if (ndev > 1) {
            SSB_NCCL(api.GroupStart());
            for (int i = 0; i < ndev; ++i)
                SSB_NCCL(api.Broadcast(dA[0].p, dA[i].p, m * n, nccl_type(), 0, comms[i], streams[i]));
            SSB_NCCL(api.GroupEnd());
        }
        for (int i = 0; i < ndev; ++i) {
            SSB_CUDA(cudaSetDevice(devices[i]));
            SSB_CUDA(cudaStreamSynchronize(streams[i]));
            solvers.emplace_back(new Homotopy<T>(devices[i]));
            solvers.back()->set_dictionary_dev(dA[i].template as<T>(), m, n, (ptrdiff_t)n, 1);
            dA[i].release();
        }
    }

    void set_residual_mode(int mode) { for (auto& s : solvers) s->residual_mode = mode; }

    // runs body(i) on one host thread per GPU, rethrows the first failure
    template <typename F>
    void per_device(F&& body) {
        const int ndev = (int)devices.size();
        std::vector<std::exception_ptr> err(ndev);
        std::vector<std::thread> th;
        auto run = [&](int i) {
            try { SSB_CUDA(cudaSetDevice(devices[i])); body(i); } catch (...) { err[i] = std::current_exception(); }
        };
        for (int i = 1; i < ndev; ++i) th.emplace_back(run, i);
        run(0);
        for (auto& t : th) t.join();
        for (auto& e : err) if (e) std::rethrow_exception(e);
    }

    void solve_batch_host(const T* Y, size_t batch, size_t ldy, T tol, uint32_t max_iter, T* X, size_t ldx, uint32_t* iters, double* errs) {
        const int ndev = (int)devices.size();
        per_device([&](int i) {
            size_t lo, hi;
            shard_bounds(batch, ndev, i, lo, hi);
            if (hi > lo) solvers[i]->solve_batch_host(Y + lo * ldy, hi - lo, ldy, tol, max_iter, X + lo * ldx, ldx, iters + lo, errs + lo);
        });
    }

    // dY / dX / reports on devices[0]; `stream` is the stream they were produced on
    void solve_batch_dev(const T* dY, size_t batch, size_t ldy, T tol, uint32_t max_iter, T* dX, size_t ldx, uint32_t* d_iters,
                         double* d_errs, cudaStream_t stream) {
        const int ndev = (int)devices.size();
        auto& api = NcclApi::get();
        SSB_REQUIRE(ldy >= (size_t)M && ldx >= (size_t)N, "leading dimensions too small");
        SSB_CUDA(cudaSetDevice(devices[0]));
        SSB_CUDA(cudaStreamSynchronize(stream));
        if (ndev == 1) { solvers[0]->solve_batch_dev(dY, batch, ldy, tol, max_iter, dX, ldx, d_iters, d_errs, nullptr); return; }
        SSB_REQUIRE(ldy == (size_t)M && ldx == (size_t)N, "the multi-GPU device entry takes dense rows (ldy == M, ldx == N)");
        std::vector<size_t> lo(ndev), hi(ndev);
        for (int i = 0; i < ndev; ++i) shard_bounds(batch, ndev, i, lo[i], hi[i]);
        for (int i = 1; i < ndev; ++i) {
            SSB_CUDA(cudaSetDevice(devices[i]));
            const size_t r = hi[i] - lo[i];
            shards[i]->Y.reserve(r * M * sizeof(T)); shards[i]->X.reserve(r * N * sizeof(T));
            shards[i]->it.reserve(r * sizeof(uint32_t)); shards[i]->err.reserve(r * sizeof(double));
        }
        // scatter the signal shards: root sends, every other GPU receives (one NCCL group = one fused launch per GPU)
        SSB_NCCL(api.GroupStart());
        for (int i = 1; i < ndev; ++i) {
            if (hi[i] == lo[i]) continue;
            SSB_NCCL(api.Send(dY + lo[i] * M, (hi[i] - lo[i]) * M, nccl_type(), i, comms[0], streams[0]));
            SSB_NCCL(api.Recv(shards[i]->Y.p, (hi[i] - lo[i]) * M, nccl_type(), 0, comms[i], streams[i]));
        }
        SSB_NCCL(api.GroupEnd());
        per_device([&](int i) {
            SSB_CUDA(cudaStreamSynchronize(streams[i]));
            const size_t r = hi[i] - lo[i];
            if (r == 0) return;
            if (i == 0) solvers[0]->solve_batch_dev(dY, r, M, tol, max_iter, dX, N, d_iters, d_errs, nullptr);
            else solvers[i]->solve_batch_dev(shards[i]->Y.template as<T>(), r, M, tol, max_iter, shards[i]->X.template as<T>(), N,
                                             shards[i]->it.template as<uint32_t>(), shards[i]->err.template as<double>(), nullptr);
        });
        // gather: coefficient rows and reports back to the root's arrays
        SSB_NCCL(api.GroupStart());
        for (int i = 1; i < ndev; ++i) {
            const size_t r = hi[i] - lo[i];
            if (r == 0) continue;
            SSB_NCCL(api.Send(shards[i]->X.p, r * N, nccl_type(), 0, comms[i], streams[i]));
            SSB_NCCL(api.Recv(dX + lo[i] * N, r * N, nccl_type(), i, comms[0], streams[0]));
            SSB_NCCL(api.Send(shards[i]->it.p, r * sizeof(uint32_t), ncclUint8, 0, comms[i], streams[i]));
            SSB_NCCL(api.Recv(d_iters + lo[i], r * sizeof(uint32_t), ncclUint8, i, comms[0], streams[0]));
            SSB_NCCL(api.Send(shards[i]->err.p, r * sizeof(double), ncclUint8, 0, comms[i], streams[i]));
            SSB_NCCL(api.Recv(d_errs + lo[i], r * sizeof(double), ncclUint8, i, comms[0], streams[0]));
        }
        SSB_NCCL(api.GroupEnd());
        for (int i = 0; i < ndev; ++i) { SSB_CUDA(cudaSetDevice(devices[i])); SSB_CUDA(cudaStreamSynchronize(streams[i])); }
        SSB_CUDA(cudaSetDevice(devices[0]));
    }
};

}  // namespace ssb
