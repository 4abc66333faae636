// float64 form of the batched residual correlation  Q[r][n] = sum_m At[n][m] * P[r][m]  on the FP64 tensor pipe
// (the reference's transposed xgemv calls in double, src/solvers/homotopy-cpu.cpp:97,:116-120; IRLS S = Q^T y,
// src/solvers/irls-cpu.cpp:53).  tcgen05 has no f64 kind, so this is warp-level mma.sync.m8n8k4.f64 (DMMA):
//   D (8 atoms x 8 signals) += A (8 atoms x 4 m, row-major) * B (4 m x 8 signals, column-major)
// Both operands are m-contiguous in memory (At is atom-major, P signal-major), which is exactly the (row, col) operand
// pair: lane (g = lane / 4, t = lane % 4) holds A[g][t] = At[n0 + g][m0 + t] and B[t][g] = P[r0 + g][m0 + t].
//
// CTA: 128 atoms x 64 signals, 256 threads = 8 warps as 4 (atoms) x 2 (signals), a warp owns 32 x 32 = 4 x 4 DMMA tiles
// (32 accumulator doubles per thread).  Operand chunks of 16 m-values are staged in shared memory with cp.async, two stages;
// rows are padded to 20 doubles so that the 16 lanes of a half-warp (g = 0..3, t = 0..3 -> bank g * 4 + t) hit 16
// different 8-byte banks.  Rows of P past `rows` are zero-filled by cp.async (src-size 0).
#pragma once

#include "common.cuh"

namespace ssb {

constexpr int DM_ATOMS = 128, DM_SIGS = 64, DM_BK = 16, DM_LD = 20, DM_STAGES = 2;
constexpr int DM_STAGE_DOUBLES = (DM_ATOMS + DM_SIGS) * DM_LD;
constexpr int DM_SMEM_BYTES = DM_STAGES * DM_STAGE_DOUBLES * 8;   // 61,440 bytes: opt-in

__device__ __forceinline__ void dmma_m8n8k4(double (&c)[2], double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
                 : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem, bool valid) {
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem);
    const int bytes = valid ? 16 : 0;   // src-size 0: the 16 bytes are zero-filled
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(s), "l"(gmem), "r"(bytes) : "memory");
}

__global__ void __launch_bounds__(256)
contract_dmma_kernel(const double* __restrict__ At, int lda, const double* __restrict__ P, int ldp, int rows,
                     double* __restrict__ Q, int ldq, int mdim)
{
    extern __shared__ __align__(16) double dm_smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int wm = warp & 3, wr = warp >> 2;
    const int n0 = blockIdx.x * DM_ATOMS, r0 = blockIdx.y * DM_SIGS;

    // stage loader: (128 + 64) rows x 16 doubles = 192 rows x 8 chunks of 16 bytes = 1536 chunks, 6 per thread
    auto load_stage = [&](int stage, int k0) {
        double* sA = dm_smem + stage * DM_STAGE_DOUBLES;
        double* sP = sA + DM_ATOMS * DM_LD;
#pragma unroll
        for (int i = 0; i < 6; ++i) {
            const int ch = tid + i * 256, row = ch >> 3, c2 = (ch & 7) * 2;
            const bool inside = k0 + c2 < mdim;                      // mdim is a multiple of 2 (Mp % 32 == 0)
            if (row < DM_ATOMS) cp_async16(sA + row * DM_LD + c2, At + (size_t)(n0 + row) * lda + k0 + c2, inside);
            else {
                const int pr = row - DM_ATOMS;
                const bool ok = inside && (r0 + pr) < rows;
                cp_async16(sP + pr * DM_LD + c2, P + (size_t)(ok ? r0 + pr : 0) * ldp + (inside ? k0 + c2 : 0), ok);
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };

    double acc[4][4][2];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) { acc[i][j][0] = 0.0; acc[i][j][1] = 0.0; }

    const int nk = (mdim + DM_BK - 1) / DM_BK;
    load_stage(0, 0);
    for (int kc = 0; kc < nk; ++kc) {
        if (kc + 1 < nk) { load_stage((kc + 1) & 1, (kc + 1) * DM_BK); asm volatile("cp.async.wait_group 1;" ::: "memory"); }
        else asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncthreads();
        const double* sA = dm_smem + (kc & 1) * DM_STAGE_DOUBLES + (wm * 32 + g) * DM_LD + t;
        const double* sP = dm_smem + (kc & 1) * DM_STAGE_DOUBLES + DM_ATOMS * DM_LD + (wr * 32 + g) * DM_LD + t;
#pragma unroll
        for (int kk = 0; kk < DM_BK; kk += 4) {
            double a[4], b[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) { a[i] = sA[i * 8 * DM_LD + kk]; b[i] = sP[i * 8 * DM_LD + kk]; }
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) dmma_m8n8k4(acc[i][j], a[i], b[j]);
        }
        __syncthreads();   // the stage is refilled two chunks later
    }
    // D fragment: atom = i * 8 + g, signals = j * 8 + 2 t + {0, 1}
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            const int r = r0 + wr * 32 + j * 8 + 2 * t + e;
            if (r < rows) {
                double* q = Q + (size_t)r * ldq + n0 + wm * 32 + g;
#pragma unroll
                for (int i = 0; i < 4; ++i) q[i * 8] = acc[i][j][e];
            }
        }
}

// FP64 peak probes (bench.py --config C4: the roofline denominator is measured, not taken from a datasheet): every thread runs
// `iters` rounds of 16 independent DFMA chains / 8 independent DMMA tiles; out keeps the compiler from dropping them
__global__ void __launch_bounds__(256) fp64_dfma_probe_kernel(double* out, int iters, double x)
{
    double a[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = x + threadIdx.x + i;
    for (int it = 0; it < iters; ++it)
#pragma unroll
        for (int i = 0; i < 16; ++i) a[i] = fma(a[i], 0.999999, x);
    double s = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += a[i];
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void __launch_bounds__(256) fp64_dmma_probe_kernel(double* out, int iters, double x)
{
    double c[8][2];
#pragma unroll
    for (int i = 0; i < 8; ++i) { c[i][0] = x + i; c[i][1] = x - i; }
    const double a = 0.5 + 1e-9 * threadIdx.x, b = 1.0 - 1e-9 * threadIdx.x;
    for (int it = 0; it < iters; ++it)
#pragma unroll
        for (int i = 0; i < 8; ++i) dmma_m8n8k4(c[i], a, b);
    double s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += c[i][0] + c[i][1];
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

}  // namespace ssb
