// SIMT forms of the residual-correlation contraction  Q[r][n] = sum_m At[n][m] * P[r][m]
//   (reference: the two transposed xgemv calls per iteration, src/solvers/homotopy-cpu.cpp:97,:120).
//
//   contract_gemv  -- 1..4 signals: the single-signal latency path.  One warp per atom, the atom's
//                     M contiguous values are read once with 16-byte loads (HBM/L2-bound: the whole
//                     dictionary streams through every launch).
//   contract_simt  -- register-tiled FFMA/DFMA GEMM, used for float64 batches and for float32 batches
//                     too small to fill tcgen05 tiles (and as the checker of the tensor-core kernel).
//
// Layouts: At is atom-major [Npad][lda] (lda = Mpad, zero padded), P is signal-major [rows][ldp],
// Q is signal-major [rows][ldq].  Npad % 128 == 0, Mpad % 32 == 0, so no bounds checks on atoms / m.
#pragma once

#include <cstdlib>

#include "common.cuh"
#include "contract_dmma.cuh"

namespace ssb {

template <typename T, int R>
__global__ void __launch_bounds__(256)
contract_gemv_kernel(const T* __restrict__ At, int lda, int n_atoms, int mdim,
                     const T* __restrict__ P, int ldp, T* __restrict__ Q, int ldq)
{
    constexpr int V = 16 / sizeof(T);  // elements per 16-byte load
    const int lane = threadIdx.x & 31;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int nwarps = (gridDim.x * blockDim.x) >> 5;
    for (int n = warp; n < n_atoms; n += nwarps) {
        const T* a = At + (size_t)n * lda;
        T acc[R];
#pragma unroll
        for (int r = 0; r < R; ++r) acc[r] = T(0);
        for (int m = lane * V; m < mdim; m += 32 * V) {
            T av[V];
            *reinterpret_cast<int4*>(av) = __ldg(reinterpret_cast<const int4*>(a + m));
#pragma unroll
            for (int r = 0; r < R; ++r) {
                T pv[V];
                *reinterpret_cast<int4*>(pv) = __ldg(reinterpret_cast<const int4*>(P + (size_t)r * ldp + m));
#pragma unroll
                for (int v = 0; v < V; ++v) acc[r] += av[v] * pv[v];
            }
        }
#pragma unroll
        for (int r = 0; r < R; ++r) {
            T s = warp_sum(acc[r]);
            if (lane == 0) Q[(size_t)r * ldq + n] = s;
        }
    }
}

// 64 atoms x 64 signals per CTA, K-step 16, 256 threads, 4x4 outputs per thread.
template <typename T>
__global__ void __launch_bounds__(256)
contract_simt_kernel(const T* __restrict__ At, int lda, const T* __restrict__ P, int ldp, int rows,
                     T* __restrict__ Q, int ldq, int mdim)
{
    constexpr int BM = 64, BR = 64, BK = 16, PAD = 4;
    __shared__ T As[BK][BM + PAD];
    __shared__ T Ps[BK][BR + PAD];

    const int t = threadIdx.x;
    const int tx = t & 15, ty = t >> 4;
    const int n0 = blockIdx.x * BM;
    const int r0 = blockIdx.y * BR;

    const int la = t >> 2;          // tile row loaded by this thread (atom or signal)
    const int lk = (t & 3) * 4;     // first of its 4 consecutive k
    const T* a_src = At + (size_t)(n0 + la) * lda + lk;
    const bool p_ok = (r0 + la) < rows;
    const T* p_src = P + (size_t)(p_ok ? (r0 + la) : 0) * ldp + lk;

    T acc[4][4];
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int i = 0; i < 4; ++i) acc[j][i] = T(0);

    for (int k0 = 0; k0 < mdim; k0 += BK) {
        T av[4], pv[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) { av[e] = a_src[k0 + e]; pv[e] = p_ok ? p_src[k0 + e] : T(0); }
        __syncthreads();
#pragma unroll
        for (int e = 0; e < 4; ++e) { As[lk + e][la] = av[e]; Ps[lk + e][la] = pv[e]; }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < BK; ++k) {
            T a[4], p[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) { a[i] = As[k][tx * 4 + i]; p[i] = Ps[k][ty * 4 + i]; }
#pragma unroll
            for (int j = 0; j < 4; ++j)
#pragma unroll
                for (int i = 0; i < 4; ++i) acc[j][i] += a[i] * p[j];
        }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int r = r0 + ty * 4 + j;
        if (r < rows) {
            T* q = Q + (size_t)r * ldq + n0 + tx * 4;
#pragma unroll
            for (int i = 0; i < 4; ++i) q[i] = acc[j][i];
        }
    }
}

// float64 batches go to the FP64 tensor pipe (contract_dmma.cuh) unless the shape rules it out; SSB200_DMMA=0 keeps the
// register-tiled DFMA kernel (A/B runs, tests/test_gpu_contract.py)
inline bool dmma_enabled() {
    static const bool on = [] { const char* e = getenv("SSB200_DMMA"); return !(e && atoi(e) == 0); }();
    return on;
}

// returns the number of kernel launches (always 1)
template <typename T>
inline int launch_contract_simt(const T* At, int lda, int n_pad, int mdim, const T* P, int ldp, int rows,
                                T* Q, int ldq, cudaStream_t stream, bool allow_dmma = true)
{
    if (rows <= 0) return 0;
    if constexpr (sizeof(T) == 8) {
        if (rows > 4 && allow_dmma && dmma_enabled() && n_pad % DM_ATOMS == 0 && mdim % DM_BK == 0 && lda % 2 == 0 && ldp % 2 == 0 &&
            (reinterpret_cast<uintptr_t>(At) & 15) == 0 && (reinterpret_cast<uintptr_t>(P) & 15) == 0) {
            static std::atomic<uint64_t> configured{ 0 };
            optin_shared_memory((const void*)contract_dmma_kernel, configured, DM_SMEM_BYTES);
            dim3 grid(n_pad / DM_ATOMS, (rows + DM_SIGS - 1) / DM_SIGS);
            contract_dmma_kernel<<<grid, 256, DM_SMEM_BYTES, stream>>>(At, lda, P, ldp, rows, Q, ldq, mdim);
            SSB_CUDA(cudaGetLastError());
            return 1;
        }
    }
    if (rows <= 4) {
        const int blocks = (n_pad + 7) / 8;  // 8 warps per CTA, one atom per warp
        if (rows == 1)      contract_gemv_kernel<T, 1><<<blocks, 256, 0, stream>>>(At, lda, n_pad, mdim, P, ldp, Q, ldq);
        else if (rows == 2) contract_gemv_kernel<T, 2><<<blocks, 256, 0, stream>>>(At, lda, n_pad, mdim, P, ldp, Q, ldq);
        else if (rows == 3) contract_gemv_kernel<T, 3><<<blocks, 256, 0, stream>>>(At, lda, n_pad, mdim, P, ldp, Q, ldq);
        else                contract_gemv_kernel<T, 4><<<blocks, 256, 0, stream>>>(At, lda, n_pad, mdim, P, ldp, Q, ldq);
    } else {
        dim3 grid(n_pad / 64, (rows + 63) / 64);
        contract_simt_kernel<T><<<grid, 256, 0, stream>>>(At, lda, P, ldp, rows, Q, ldq, mdim);
    }
    SSB_CUDA(cudaGetLastError());
    return 1;
}

}  // namespace ssb
