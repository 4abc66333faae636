// Host driver of the IRLS solve.  Stands in for irls_state (src/lib.cpp:51-57), irls_policy::run
// (src/lib.cpp:61-73) and solve_irls::op<CPU,T> (src/solvers/irls-cpu.cpp:126-146).
#pragma once

#include <algorithm>
#include <cmath>
#include <limits>
#include <vector>

#include "contract_simt.cuh"
#include "homotopy_kernels.cuh"
#include "irls_kernels.cuh"
#include "solver_base.h"

namespace ssb {

template <typename T>
__global__ void fill_kernel(T* p, size_t n, T v) {
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (size_t)gridDim.x * blockDim.x) p[e] = v;
}

// Householder QR of a dense row-major M x N matrix (M >= N): thin Q (M x N) and upper R (N x N).
// Same algorithm and sign convention as the reference's JAMA-style class
// (src/linalg/qr_decomposition.h:93-190: hypot-accumulated column norms, R_kk = -norm * sign).
// One-time, per dictionary; runs on the host in T precision like the reference's constructor.
template <typename T>
void householder_qr(std::vector<T>& a, size_t M, size_t N, std::vector<T>& Q, std::vector<T>& R)
{
    std::vector<T> rdiag(N, T(0));
    for (size_t k = 0; k < N; ++k) {
        T nrm = T(0);
        for (size_t i = k; i < M; ++i) nrm = (T)std::hypot((double)nrm, (double)a[i * N + k]);
        if (nrm != T(0)) {
            if (a[k * N + k] < T(0)) nrm = -nrm;
            for (size_t i = k; i < M; ++i) a[i * N + k] /= nrm;
            a[k * N + k] += T(1);
            const T akk = a[k * N + k];
            std::vector<T> s(N, T(0));
            for (size_t i = k; i < M; ++i) {
                const T v = a[i * N + k];
                for (size_t j = k + 1; j < N; ++j) s[j] += v * a[i * N + j];
            }
            for (size_t j = k + 1; j < N; ++j) s[j] /= -akk;
            for (size_t i = k; i < M; ++i) {
                const T v = a[i * N + k];
                for (size_t j = k + 1; j < N; ++j) a[i * N + j] += v * s[j];
            }
        }
        rdiag[k] = -nrm;
    }
    R.assign(N * N, T(0));
    for (size_t i = 0; i < N; ++i) {
        R[i * N + i] = rdiag[i];
        for (size_t j = i + 1; j < N; ++j) R[i * N + j] = a[i * N + j];
    }
    Q.assign(M * N, T(0));
    std::vector<T> s(N);
    for (size_t kk = N; kk-- > 0;) {
        const size_t k = kk;
        for (size_t i = 0; i < M; ++i) Q[i * N + k] = T(0);
        Q[k * N + k] = T(1);
        const T akk = a[k * N + k];
        if (akk == T(0)) continue;
        std::fill(s.begin(), s.end(), T(0));
        for (size_t i = k; i < M; ++i) {
            const T v = a[i * N + k];
            for (size_t j = k; j < N; ++j) s[j] += v * Q[i * N + j];
        }
        for (size_t j = k; j < N; ++j) s[j] = -s[j] / akk;
        for (size_t i = k; i < M; ++i) {
            const T v = a[i * N + k];
            for (size_t j = k; j < N; ++j) Q[i * N + j] += s[j] * v;
        }
    }
}

template <typename T>
class Irls : public SolverBase {
public:
    static constexpr int THREADS = 256;
    Irls(int dev) : SolverBase(sizeof(T) == 4 ? SolverKind::IrlsF32 : SolverKind::IrlsF64, dev) {}

    int M = 0, N = 0, Mp = 0, Npad = 0;
    bool wide = false;   // M < N: outside the reference's domain (qr_decomposition.h:101)

    void set_dictionary_host(const T* A, size_t m, size_t n, ptrdiff_t rs, ptrdiff_t cs) {
        SSB_REQUIRE(m > 0 && n > 0, "dictionary must be non-empty");
        M = (int)m; N = (int)n;
        Mp = (int)round_up(m, 32);
        Npad = (int)round_up(n, 128);
        wide = m < n;
        if (wide) return;
        std::vector<T> a(m * n);
        for (size_t i = 0; i < m; ++i)
            for (size_t j = 0; j < n; ++j) a[i * n + j] = A[(ptrdiff_t)i * rs + (ptrdiff_t)j * cs];
        std::vector<T> Q, R;
        householder_qr<T>(a, m, n, Q, R);
        // Q atom-major (one column of Q contiguous), zero padded; R column-major
        std::vector<T> qt((size_t)Npad * Mp, T(0)), rc(n * n, T(0));
        for (size_t j = 0; j < n; ++j)
            for (size_t i = 0; i < m; ++i) qt[j * Mp + i] = Q[i * n + j];
        for (size_t k = 0; k < n; ++k)
            for (size_t i = 0; i < n; ++i) rc[i * n + k] = R[k * n + i];
        Qt_.reserve(qt.size() * sizeof(T));
        Rc_.reserve(rc.size() * sizeof(T));
        SSB_CUDA(cudaMemcpyAsync(Qt_.p, qt.data(), qt.size() * sizeof(T), cudaMemcpyHostToDevice, stream));
        SSB_CUDA(cudaMemcpyAsync(Rc_.p, rc.data(), rc.size() * sizeof(T), cudaMemcpyHostToDevice, stream));
        SSB_CUDA(cudaStreamSynchronize(stream));
    }

    void solve_batch_dev(const T* dY, size_t batch, size_t ldy, T tol, uint32_t max_iter, T* dX, size_t ldx,
                         uint32_t* d_iters, double* d_errs, uint8_t* d_spd, cudaStream_t st)
    {
        SSB_REQUIRE(M > 0, "no dictionary");
        SSB_REQUIRE(max_iter > 0, "max_iterations must be positive");
        if (batch == 0) return;
        SSB_REQUIRE(ldy >= (size_t)M && ldx >= (size_t)N, "leading dimensions too small");
        if (!st) st = stream;
        if (wide) {
            // what the reference build returns for M < N: no iteration, SPD failure, x = 0/0
            for (size_t b = 0; b < batch; ++b)
                fill_kernel<T><<<8, 256, 0, st>>>(dX + b * ldx, (size_t)N, std::numeric_limits<T>::quiet_NaN());
            fill_kernel<uint32_t><<<64, 256, 0, st>>>(d_iters, batch, 0u);
            fill_kernel<double><<<64, 256, 0, st>>>(d_errs, batch, 1.0);
            fill_kernel<uint8_t><<<64, 256, 0, st>>>(d_spd, batch, (uint8_t)1);
            SSB_CUDA(cudaGetLastError());
            counters[0] += batch + 3;
            SSB_CUDA(cudaStreamSynchronize(st));
            return;
        }
        // (the chunk of the previous call with the same batch is reused: cudaMemGetInfo takes 10-80 ms every few calls on a busy box)
        size_t chunk = planned_chunk_;
        if (planned_batch_ != batch || chunk == 0) {
            size_t free_b = 0, total_b = 0;
            SSB_CUDA(cudaMemGetInfo(&free_b, &total_b));
            const size_t per = ((size_t)N * N + 4 * (size_t)N + Npad + Mp) * sizeof(T);
            const size_t budget = std::min<size_t>((size_t)(0.6 * (double)free_b) + L_.bytes, (size_t)32 << 30);
            chunk = std::max<size_t>(1, std::min<size_t>(budget / per, 16384));
            chunk = std::min(chunk, batch);
            planned_batch_ = batch; planned_chunk_ = chunk;
        }
        L_.reserve(chunk * (size_t)N * N * sizeof(T));
        vec_.reserve(chunk * 4 * (size_t)N * sizeof(T));
        S_.reserve(round_up(chunk, 64) * Npad * sizeof(T));
        P_.reserve(round_up(chunk, 64) * Mp * sizeof(T));
        for (size_t b0 = 0; b0 < batch; b0 += chunk) {
            const size_t rows = std::min(chunk, batch - b0);
            const int blocks = (int)std::min<size_t>(8192, (rows + 7) / 8);
            stage_signals_kernel<T><<<blocks, 256, 0, st>>>(dY + b0 * ldy, ldy, M, (int)rows, P_.as<T>(), nullptr, nullptr, Mp, 0, nullptr);
            SSB_CUDA(cudaGetLastError());
            counters[0]++;
            // S = Q^T Y  (irls-cpu.cpp:53, the same for every iteration of a signal)
            int span = begin_span(0, st);
            counters[0] += launch_contract_simt<T>(Qt_.as<T>(), Mp, Npad, Mp, P_.as<T>(), Mp, (int)rows, S_.as<T>(), Npad, st);
            end_span(span, st);
            counters[1]++; counters[2] += rows;
            IrlsArgs<T> a{};
            a.N = N; a.lde = Npad; a.Tq = P_.as<T>(); a.Qt = Qt_.as<T>(); a.M = M; a.ldq = Mp; a.Rc = Rc_.as<T>(); a.S = S_.as<T>();
            a.X = dX + b0 * ldx; a.ldx = ldx; a.L = L_.as<T>(); a.vec = vec_.as<T>();
            a.iters = d_iters + b0; a.errs = d_errs + b0; a.spd = d_spd + b0;
            a.tol = tol; a.max_iter = max_iter;
            span = begin_span(1, st);
            irls_solve_kernel<T, THREADS><<<(unsigned)rows, THREADS, 0, st>>>(a);
            SSB_CUDA(cudaGetLastError());
            end_span(span, st);
            counters[0]++;
            SSB_CUDA(cudaStreamSynchronize(st));
        }
        collect_spans();
    }

    void solve_batch_host(const T* Y, size_t batch, size_t ldy, T tol, uint32_t max_iter, T* X, size_t ldx,
                          uint32_t* iters, double* errs, uint8_t* spd)
    {
        if (batch == 0) return;
        SSB_REQUIRE(ldy >= (size_t)M && ldx >= (size_t)N, "leading dimensions too small");
        dev_buffer dY, dX, dIt, dErr, dSpd;
        dY.reserve(batch * M * sizeof(T));
        dX.reserve(batch * N * sizeof(T));
        dIt.reserve(batch * sizeof(uint32_t));
        dErr.reserve(batch * sizeof(double));
        dSpd.reserve(batch);
        SSB_CUDA(cudaMemcpy2DAsync(dY.p, M * sizeof(T), Y, ldy * sizeof(T), M * sizeof(T), batch, cudaMemcpyHostToDevice, stream));
        solve_batch_dev(dY.as<T>(), batch, M, tol, max_iter, dX.as<T>(), N, dIt.as<uint32_t>(), dErr.as<double>(),
                        dSpd.as<uint8_t>(), stream);
        SSB_CUDA(cudaMemcpy2DAsync(X, ldx * sizeof(T), dX.p, N * sizeof(T), N * sizeof(T), batch, cudaMemcpyDeviceToHost, stream));
        SSB_CUDA(cudaMemcpyAsync(iters, dIt.p, batch * sizeof(uint32_t), cudaMemcpyDeviceToHost, stream));
        SSB_CUDA(cudaMemcpyAsync(errs, dErr.p, batch * sizeof(double), cudaMemcpyDeviceToHost, stream));
        SSB_CUDA(cudaMemcpyAsync(spd, dSpd.p, batch, cudaMemcpyDeviceToHost, stream));
        SSB_CUDA(cudaStreamSynchronize(stream));
    }

private:
    dev_buffer Qt_, Rc_, L_, vec_, S_, P_;
    size_t planned_batch_ = 0, planned_chunk_ = 0;
};

}  // namespace ssb
