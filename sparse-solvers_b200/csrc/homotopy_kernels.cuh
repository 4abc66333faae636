// Per-signal kernels of the batched homotopy solve (sm_100a).
//
// One CTA owns one signal for one lockstep iteration.  Everything the reference does between two
// full-dictionary products is in homotopy_step_kernel:
//   min-gamma search over the join and leave candidate sets      homotopy-cpu.cpp:122-163
//   active-set toggle + online update of (A_G^T A_G)^-1           homotopy-cpu.cpp:166-183,
//                                                                 src/linalg/online_inverse.h:183-293
//   x += gamma * d (old direction), correlation update, |c|_inf   homotopy-cpu.cpp:252-255,:270
//   new direction d_G = inv * sign(c_G, tol)                      homotopy-cpu.cpp:259-266
//   p = A_G d_G for the next contraction  q = A^T p               homotopy-cpu.cpp:112-120
//
// The residual correlation (residual_vector, homotopy-cpu.cpp:87-98, called at :255) comes in two forms:
//   exact        r = y - A_G x_G is re-formed from the active atoms every iteration, c' = A^T r travels through the same
//                contraction as q = A^T p (two stacked row blocks), and the step search and lambda = |c'|_inf read that
//                fresh c'; the direction signs, the stop test and the report use c' - gamma q, ONE step from the fresh base
//                (never accumulated).  Default wherever the dictionary is coherent (Homotopy::exact_residual).
//   incremental  c -= gamma*q carried from iteration to iteration (q = A^T A d is already there): one contraction per
//                iteration.  Inside the reference's own two-BLAS-builds noise floor on incoherent dictionaries only.
// What differs from the reference by construction (all algebraically identical):
//   - A_G^T v for a joining atom v is read from the precomputed Gram matrix G = A^T A instead of
//     K dot products (online_inverse.h:217-218);
//   - the active set is kept in slot (insertion) order with a bitmask for membership, not as a
//     sorted rank index (src/linalg/rank_index.h): no result depends on the order, only the
//     summation order inside inv*u does.  A leaving atom's slot is refilled by the last slot.
#pragma once

#include <cuda_fp16.h>

#include "common.cuh"

namespace ssb {

template <typename T>
struct HomArgs {
    int N;            // atoms
    int ldc;          // leading dimension of C and Q (= Npad)
    int Mp;           // padded signal length (= Mpad), also ld of At
    int ldp;          // leading dimension of the P buffers
    int Kmax;         // capacity of the per-signal active set
    const T* At;      // [Npad][Mp]  atom-major dictionary, zero padded
    const T* G;       // [N][ldg]    Gram matrix A^T A
    int ldg;
    T* C;             // [Bc][ldc]   correlations, by signal
    const T* Q;       // [rows][ldc] contraction result, by position in `act`
    T* X;             // signal-major coefficients (caller's output)
    size_t ldx;
    uint32_t* gam;    // [Bc][Kmax]  active atoms in slot order
    T* dG;            // [Bc][Kmax]  direction on the active atoms
    T* Ginv;          // [Bc][Kmax*Kmax], leading dimension Kmax
    T* lam;           // [Bc]        |c|_inf
    int* Kcur;        // [Bc]
    uint32_t* iters;  // [Bc]        report: loop passes
    double* errs;     // [Bc]        report: solution_error
    const int* act;   // [rows]      position -> signal (nullptr: identity)
    int* act_next;    // compacted list for the next iteration
    int* next_count;
    const int* count_in;   // signals entering this iteration, on the device (nullptr: the grid is exact)
    int* count_zero;       // ring slot of the iteration after next's count: cleared here, its last reader is long done
    uint32_t* mask;   // [Bc][ldc/32] membership bits of the active set (kept by init / algebra / step, read by the scan)
    uint32_t* pick;   // [Bc]        atom chosen by the min-gamma search (split kernels: scan -> algebra)
    T* Pn;            // [rows][ldp] next right-hand sides (full precision, or the TF32 hi part)
    T* Pn_lo;         //             lo part (only when split != 0)
    int split;        // how the right-hand sides are written: 0 full precision, 1 tf32 (hi, lo) floats, 2 fp16 (hi, lo * 2^11)
                      // of the row times a power of two (Pn / Pn_lo then point at __half arrays, row stride ldp halfs)
    float* Pn_inv;    // [rows]      split 2: 1 / scale of every right-hand side
    int skip_rhs;     //             split kernels: do not form p (the next scan takes q from G)
    float amax;       //             split 2: max |A|, bounds |p| <= amax * sum |d|
    T tol;
    uint32_t max_iter;
    uint32_t it;      // lockstep iteration number, 1-based
    int dbg;          // timing experiments only (SSB200_DBG): 1 = stop after the c update, 2 = stop before p
    // ---- exact-residual mode (exact != 0)
    int exact;
    const T* Qc;      // [rows][ldc] fresh correlations c' = A^T r, by position (the contraction's second row block)
    T* C0;            // [Bc][ldc]   A^T y, kept for the Gram iterations (c' = c0 - G_G x_G); nullptr when they are off
    const T* Y;       // [Bc][ldy]   the signals, caller layout (r = y - A_G x_G)
    size_t ldy;
    int M;            // signal length (rows of A); Mp - M padding values of r are zero
    const float* ymax; // [Bc]       max |y| per signal (bounds r for the fp16 containers)
    T* Rn;            // [rows][ldp] next residual rows, same container format as Pn
    T* Rn_lo;
    float* Rn_inv;
};

__device__ __forceinline__ float tf32_rna(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return __uint_as_float(r);
}

// Power-of-two scale that brings a row bounded by `bound` into the fp16 range: bound * 2^e <= 2^14.  A loose bound
// only costs headroom: fp16 keeps 11 significant bits over 29 binades, so a bound 2^8 above the true maximum still leaves
// every element within 2^-21 of it a normal number.
__device__ __forceinline__ int f16_scale_exp(float bound) {
    if (!(bound > 0.f) || !(bound < FLT_MAX)) return 0;
    int ex;
    frexpf(bound, &ex);                 // bound = m * 2^ex, m in [0.5, 1)
    return max(-100, min(100, 14 - ex));
}

// writes V consecutive values of p: full precision, a TF32 (hi, lo) pair, or an fp16 (hi, lo * 2^11) pair of p * scale.
// tf32 and fp16 both carry 11 significant bits, so the two splits have the same accuracy (hi + lo covers 22 bits);
// the fp16 form feeds tcgen05.mma.kind::f16 at twice the tf32 rate and half the operand bytes.
template <typename T, int V>
__device__ __forceinline__ void store_rhs(int split, void* hi_row, void* lo_row, int m0, const T (&v)[V], float scale) {
    if constexpr (sizeof(T) == 4) {
        if (split == 1) {
            float h[V], l[V];
#pragma unroll
            for (int e = 0; e < V; ++e) { h[e] = tf32_rna(v[e]); l[e] = tf32_rna(v[e] - h[e]); }
            *reinterpret_cast<int4*>(static_cast<float*>(hi_row) + m0) = *reinterpret_cast<const int4*>(h);
            *reinterpret_cast<int4*>(static_cast<float*>(lo_row) + m0) = *reinterpret_cast<const int4*>(l);
            return;
        }
        if (split == 2) {
            __half h[V], l[V];
#pragma unroll
            for (int e = 0; e < V; ++e) {
                const float ps = v[e] * scale;
                h[e] = __float2half_rn(ps);
                l[e] = __float2half_rn((ps - __half2float(h[e])) * 2048.0f);
            }
            *reinterpret_cast<uint2*>(static_cast<__half*>(hi_row) + m0) = *reinterpret_cast<const uint2*>(h);
            *reinterpret_cast<uint2*>(static_cast<__half*>(lo_row) + m0) = *reinterpret_cast<const uint2*>(l);
            return;
        }
    }
    *reinterpret_cast<int4*>(static_cast<T*>(hi_row) + m0) = *reinterpret_cast<const int4*>(v);
}

// row `row` of the next right-hand sides, as bytes (the element size depends on the split)
template <typename T>
__device__ __forceinline__ void rhs_rows(const HomArgs<T>& a, int row, void*& hi, void*& lo) {
    const size_t esz = a.split == 2 ? 2 : sizeof(T);
    hi = reinterpret_cast<unsigned char*>(a.Pn) + (size_t)row * a.ldp * esz;
    lo = a.split ? reinterpret_cast<unsigned char*>(a.Pn_lo) + (size_t)row * a.ldp * esz : nullptr;
}

// Y (caller layout) -> first right-hand sides, zero padded to ldp; one warp per signal (split 2 needs the row maximum)
template <typename T>
__global__ void stage_signals_kernel(const T* __restrict__ Y, size_t ldy, int M, int rows,
                                     void* __restrict__ P, void* __restrict__ Plo, float* __restrict__ Pinv, int ldp, int split,
                                     float* __restrict__ ymax)
{
    constexpr int V = 16 / sizeof(T);
    const int lane = threadIdx.x & 31;
    const size_t esz = split == 2 ? 2 : sizeof(T);
    for (size_t r = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); r < (size_t)rows; r += (size_t)gridDim.x * (blockDim.x >> 5)) {
        float scale = 1.f;
        if (split == 2 || ymax) {
            float mx = 0.f;
            for (int m = lane; m < M; m += 32) mx = fmaxf(mx, fabsf((float)Y[r * ldy + m]));
            mx = warp_max(mx);
            if (ymax && lane == 0) ymax[r] = mx;
        if (split == 2) {
            const int e = f16_scale_exp(mx);
            scale = ldexpf(1.f, e);
            if (lane == 0) Pinv[r] = ldexpf(1.f, -e);
        }
        }
        void* hi = static_cast<unsigned char*>(P) + r * ldp * esz;
        void* lo = Plo ? static_cast<unsigned char*>(Plo) + r * ldp * esz : nullptr;
        for (int m0 = lane * V; m0 < ldp; m0 += 32 * V) {
            T v[V];
#pragma unroll
            for (int k = 0; k < V; ++k) v[k] = (m0 + k < M) ? Y[r * ldy + m0 + k] : T(0);
            store_rhs<T, V>(split, hi, lo, m0, v, scale);
        }
    }
}

// join-candidate step (lambda -+ c_i) / (1 -+ q_i) (homotopy-cpu.cpp:143-153).  2N of these per signal-iteration make
// the IEEE division sequence the largest single consumer of issue slots in the step kernel, so float32 ORDERS the candidates
// with the hardware reciprocal (<= 2 ulp; exact cases such as x/1 stay exact) and the one step that wins is then recomputed
// with IEEE division (winning_step): the gamma applied to x and c is the reference's quotient.  -DSSB_IEEE_ORDER=1 orders with
// IEEE division as well (A/B of the parity counts, tests/studies/noise_floor.py).
#ifndef SSB_IEEE_ORDER
#define SSB_IEEE_ORDER 0
#endif
__device__ __forceinline__ float step_ratio(float num, float den) { return SSB_IEEE_ORDER ? __fdiv_rn(num, den) : __fdividef(num, den); }
__device__ __forceinline__ double step_ratio(double num, double den) { return num / den; }

// the step of join candidate i in IEEE arithmetic: the smaller strictly positive one of its two ratios (:143-153);
// `fallback` (the ordering value) when neither qualifies, which only a 2-ulp disagreement at the sign boundary can cause
template <typename T>
__device__ __forceinline__ T winning_step(T lam, T c, T q, T fallback) {
    const T TMAX = num_traits<T>::max();
    T best = TMAX;
    const T dl = T(1) - q, dr = T(1) + q;
    if (dl != T(0)) { const T t = (lam - c) / dl; if (t > T(0) && t < best) best = t; }
    if (dr != T(0)) { const T t = (lam + c) / dr; if (t > T(0) && t < best) best = t; }
    return best < TMAX ? best : fallback;
}

template <typename T>
__device__ __forceinline__ T sign_tol(T v, T tol) { return v > tol ? T(1) : (v < -tol ? T(-1) : T(0)); }

// row `row` of the next residual block (exact mode), as bytes
template <typename T>
__device__ __forceinline__ void res_rows(const HomArgs<T>& a, int row, void*& hi, void*& lo) {
    const size_t esz = a.split == 2 ? 2 : sizeof(T);
    hi = reinterpret_cast<unsigned char*>(a.Rn) + (size_t)row * a.ldp * esz;
    lo = a.split ? reinterpret_cast<unsigned char*>(a.Rn_lo) + (size_t)row * a.ldp * esz : nullptr;
}

// p = sum_k d[k] * At[gam[k]][:]  for the CTA's signal, written to row `row` of the next P; in exact mode also the residual
// r = y - sum_k x[k] * At[gam[k]][:] (homotopy-cpu.cpp:94-96 over the active atoms: every other x is zero or the
// rounding residue of a removed atom) to row `row` of the next R, from the same atom rows
template <typename T, int THREADS>
__device__ __forceinline__ void write_rhs(const HomArgs<T>& a, int row, int sig, const uint32_t* s_gam, const T* s_d, const T* s_x, int K)
{
    constexpr int V = 16 / sizeof(T);
    void *pn, *pl, *rn = nullptr, *rl = nullptr;
    rhs_rows(a, row, pn, pl);
    const bool res = a.exact != 0;
    if (res) res_rows(a, row, rn, rl);
    float scale = 1.f, rscale = 1.f;
    if (a.split == 2) {   // |p_m| <= amax * sum |d_k| ; |r_m| <= max |y| + amax * sum |x_k|
        float sd = 0.f, sx = 0.f;
        for (int k = 0; k < K; ++k) { sd += fabsf((float)s_d[k]); if (res) sx += fabsf((float)s_x[k]); }
        const int e = f16_scale_exp(sd * a.amax);
        scale = ldexpf(1.f, e);
        if (threadIdx.x == 0) a.Pn_inv[row] = ldexpf(1.f, -e);
        if (res) {
            const int er = f16_scale_exp(a.ymax[sig] + sx * a.amax);
            rscale = ldexpf(1.f, er);
            if (threadIdx.x == 0) a.Rn_inv[row] = ldexpf(1.f, -er);
        }
    }
    const T* y = res ? a.Y + (size_t)sig * a.ldy : nullptr;
    for (int m0 = threadIdx.x * V; m0 < a.Mp; m0 += THREADS * V) {
        T acc[V], accr[V];
#pragma unroll
        for (int e = 0; e < V; ++e) { acc[e] = T(0); accr[e] = T(0); }
        // eight atoms per round: their rows are fetched together (eight independent 16-byte loads in flight per thread -- with
        // one CTA per signal and K of them in a row this loop is L2 latency, not bandwidth), then added in k order
        int k = 0;
        for (; k + 8 <= K; k += 8) {
            T av[8][V];
#pragma unroll
            for (int u = 0; u < 8; ++u)
                *reinterpret_cast<int4*>(av[u]) = __ldg(reinterpret_cast<const int4*>(a.At + (size_t)s_gam[k + u] * a.Mp + m0));
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const T dk = s_d[k + u];
#pragma unroll
                for (int e = 0; e < V; ++e) acc[e] += av[u][e] * dk;
                if (res) {
                    const T xk = s_x[k + u];
#pragma unroll
                    for (int e = 0; e < V; ++e) accr[e] += av[u][e] * xk;
                }
            }
        }
        for (; k < K; ++k) {
            T av[V];
            *reinterpret_cast<int4*>(av) = __ldg(reinterpret_cast<const int4*>(a.At + (size_t)s_gam[k] * a.Mp + m0));
            const T dk = s_d[k];
#pragma unroll
            for (int e = 0; e < V; ++e) acc[e] += av[e] * dk;
            if (res) {
                const T xk = s_x[k];
#pragma unroll
                for (int e = 0; e < V; ++e) accr[e] += av[e] * xk;
            }
        }
        store_rhs<T, V>(a.split, pn, pl, m0, acc, scale);
        if (res) {
#pragma unroll
            for (int e = 0; e < V; ++e) accr[e] = (m0 + e < a.M ? y[m0 + e] : T(0)) - accr[e];
            store_rhs<T, V>(a.split, rn, rl, m0, accr, rscale);
        }
    }
}

// Initial state (homotopy-cpu.cpp:201-229): x = 0, c = A^T y (in Q), lambda = |c|_inf with the FIRST
// maximal index, G = {idx}, inv = 1/|a_idx|^2, d = sign(lambda, tol) * inv -- the sign of |c|, not of c
// (the reference quirk at :223-224, reproduced).
template <typename T, int THREADS>
__global__ void __launch_bounds__(THREADS) homotopy_init_kernel(HomArgs<T> a)
{
    constexpr int V = 16 / sizeof(T);
    constexpr int NW = THREADS / 32;
    __shared__ T red_v[NW];
    __shared__ uint32_t red_i[NW];
    __shared__ uint32_t s_gam[1];
    __shared__ T s_d[1], s_x[1];

    const int s = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const T* q = a.Q + (size_t)s * a.ldc;
    T* c = a.C + (size_t)s * a.ldc;
    T* x = a.X + (size_t)s * a.ldx;

    T best = T(-1);
    uint32_t bidx = 0xffffffffu;
    for (int i0 = tid * V; i0 < a.ldc; i0 += THREADS * V) {
        T qv[V];
        *reinterpret_cast<int4*>(qv) = *reinterpret_cast<const int4*>(q + i0);
        *reinterpret_cast<int4*>(c + i0) = *reinterpret_cast<const int4*>(qv);
        if (a.C0) *reinterpret_cast<int4*>(a.C0 + (size_t)s * a.ldc + i0) = *reinterpret_cast<const int4*>(qv);
#pragma unroll
        for (int e = 0; e < V; ++e) {
            const int i = i0 + e;
            if (i < a.N) {
                x[i] = T(0);
                const T av = fabs(qv[e]);
                if (av > best) { best = av; bidx = (uint32_t)i; }
            }
        }
    }
    warp_argmax_first(best, bidx);
    if (lane == 0) { red_v[warp] = best; red_i[warp] = bidx; }
    __syncthreads();
    if (warp == 0) {
        best = lane < NW ? red_v[lane] : T(-1);
        bidx = lane < NW ? red_i[lane] : 0xffffffffu;
        warp_argmax_first(best, bidx);
        if (lane == 0) {
            if (bidx == 0xffffffffu) { bidx = 0; best = fabs(q[0]); }  // all-NaN correlations
            const T nrm = sqrt(a.G[(size_t)bidx * a.ldg + bidx]);       // online_inverse.h:193-201
            const T inv = T(1) / (nrm * nrm);
            const T d0 = sign_tol(best, a.tol) * inv;
            a.gam[(size_t)s * a.Kmax] = bidx;
            a.dG[(size_t)s * a.Kmax] = d0;
            a.Ginv[(size_t)s * a.Kmax * a.Kmax] = inv;
            a.lam[s] = best;
            a.Kcur[s] = 1;
            s_gam[0] = bidx;
            s_d[0] = d0;
            s_x[0] = T(0);
        }
    }
    __syncthreads();
    if (a.mask) {
        const uint32_t g = s_gam[0];
        for (int w = tid; w < (a.ldc >> 5); w += THREADS) a.mask[(size_t)s * (a.ldc >> 5) + w] = (w == (int)(g >> 5)) ? 1u << (g & 31) : 0u;
    }
    write_rhs<T, THREADS>(a, s, s, s_gam, s_d, s_x, 1);   // exact mode: r = y - 0
}

// out[r] = sum_k G[r][k] v[k] for the K x K inverse (the reference's K x K xgemv, homotopy-cpu.cpp:263 and
// online_inverse.h:225).  On chip: one thread per row, k in order (odd row stride: conflict-free); in HBM: one
// warp per row so that the loads coalesce.
template <typename T, int THREADS, bool ON_CHIP>
__device__ __forceinline__ void matvec(const T* G, int ld, int K, const T* v, T* out, T* out2)
{
    if constexpr (ON_CHIP) {
        for (int r = threadIdx.x; r < K; r += THREADS) {
            T acc = T(0);
            const T* row = G + (size_t)r * ld;
            for (int k = 0; k < K; ++k) acc += row[k] * v[k];
            out[r] = acc;
            if (out2) out2[r] = acc;
        }
    } else {
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        for (int r = warp; r < K; r += THREADS / 32) {
            T acc = T(0);
            for (int k = lane; k < K; k += 32) acc += G[(size_t)r * ld + k] * v[k];
            acc = warp_sum(acc);
            if (lane == 0) { out[r] = acc; if (out2) out2[r] = acc; }
        }
    }
}

// One homotopy iteration for one signal (homotopy-cpu.cpp:236-272).
//
// Latency is what bounds this kernel (profiles/r1_ncu_step_and_contract_v1.txt: long-scoreboard + barrier stalls, DRAM
// at 26 %), so everything that does not depend on the step size is fetched up front in ONE round trip -- c and q into
// registers (NCH 16-byte chunks per thread), the active set, its direction and the K x K inverse into shared
// memory -- and c / q are read from HBM exactly once per iteration.
//   NCH  chunks of c,q a thread keeps in registers; 0 = dictionary too wide, c and q are re-read for the update
//   GS   inverse held in shared memory (row stride ks) for the iteration; false = worked on in place in HBM
//   KR   active-set slots a thread keeps in registers: K <= KR * THREADS (3 by default; 16 for active sets up to 4096)
template <typename T, int THREADS, int NCH, bool GS, int KR = 3>
__global__ void __launch_bounds__(THREADS) homotopy_step_kernel(HomArgs<T> a, int kb, int ks)
{
    constexpr int V = 16 / sizeof(T);
    constexpr int NW = THREADS / 32;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ T red_v[NW];
    __shared__ uint32_t red_i[NW];
    __shared__ T sh_gamma, sh_lam;
    __shared__ uint32_t sh_idx;
    __shared__ int sh_slot, sh_row, sh_join;

    const int KC = (a.Kmax + 2) & ~1;   // even, keeps the arrays behind it 8-byte aligned
    T* s_d = reinterpret_cast<T*>(smem_raw);
    T* s_u1 = s_d + KC;
    T* s_u2 = s_u1 + KC;
    T* s_cg = s_u2 + KC;
    T* s_x = s_cg + KC;
    uint32_t* s_gam = reinterpret_cast<uint32_t*>(s_x + KC);
    uint32_t* mask = s_gam + KC;
    const int mask_words = (a.ldc + 31) >> 5;
    T* s_G = reinterpret_cast<T*>(mask + ((mask_words + 3) & ~3));

    const int pos = blockIdx.x;
    if (a.count_zero && pos == 0 && threadIdx.x == 0) *a.count_zero = 0;   // before the early exit: an empty iteration clears it too
    if (a.count_in && pos >= *a.count_in) return;   // grid sized from a stale (larger) count
    const int s = a.act ? a.act[pos] : pos;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    uint32_t* gam = a.gam + (size_t)s * a.Kmax;
    T* dG = a.dG + (size_t)s * a.Kmax;
    T* Ginv = a.Ginv + (size_t)s * a.Kmax * a.Kmax;
    T* c = a.C + (size_t)s * a.ldc;
    const T* q = a.Q + (size_t)pos * a.ldc;
    // exact mode: the search reads the fresh c' = A^T r of this iteration (second row block of the contraction)
    const T* csrc = a.exact ? a.Qc + (size_t)pos * a.ldc : c;
    T* x = a.X + (size_t)s * a.ldx;
    const T TMAX = num_traits<T>::max();
    T* Gm = GS ? s_G : Ginv;              // where the inverse is worked on
    const int ldm = GS ? ks : a.Kmax;

    // ---- one round trip: everything the iteration needs that does not depend on gamma
    T cv[NCH > 0 ? NCH : 1][V], qv[NCH > 0 ? NCH : 1][V];
    if constexpr (NCH > 0) {
#pragma unroll
        for (int ch = 0; ch < NCH; ++ch) {
            const int i0 = (ch * THREADS + tid) * V;
            if (i0 < a.ldc) {
                *reinterpret_cast<int4*>(cv[ch]) = *reinterpret_cast<const int4*>(csrc + i0);
                *reinterpret_cast<int4*>(qv[ch]) = *reinterpret_cast<const int4*>(q + i0);
            }
        }
    }
    if constexpr (NCH > 0) {
        if (a.dbg == 3) {   // timing experiment: pure c/q streaming
#pragma unroll
            for (int ch = 0; ch < NCH; ++ch) {
                const int i0 = (ch * THREADS + tid) * V;
                if (i0 < a.ldc) {
#pragma unroll
                    for (int e = 0; e < V; ++e) cv[ch][e] -= T(0.5) * qv[ch][e];
                    *reinterpret_cast<int4*>(c + i0) = *reinterpret_cast<const int4*>(cv[ch]);
                }
            }
            if (tid == 0) { const int row = atomicAdd(a.next_count, 1); a.act_next[row] = s; }
            return;
        }
    }
    const int K = a.Kcur[s];
    T lam = a.lam[s];
    for (int k = tid; k < kb; k += THREADS) { s_gam[k] = gam[k]; s_d[k] = dG[k]; }   // K <= kb = min(it, Kmax)
    if (a.exact) {   // lambda = |c'|_inf over the same vector the search reads (:270 on the re-formed c, :255)
        T mx0 = T(0);
        if constexpr (NCH > 0) {
#pragma unroll
            for (int ch = 0; ch < NCH; ++ch) {
                const int i0 = (ch * THREADS + tid) * V;
#pragma unroll
                for (int e = 0; e < V; ++e) if (i0 + e < a.N) mx0 = max(mx0, fabs(cv[ch][e]));
            }
        } else {
            for (int i = tid; i < a.N; i += THREADS) mx0 = max(mx0, fabs(csrc[i]));
        }
        mx0 = warp_max_nonneg(mx0);
        if (lane == 0) red_v[warp] = mx0;   // read after the next barrier, rewritten only after the one that follows it
    }
    if constexpr (GS) {
        // flat copy, 8 independent loads in flight per thread (row = e / kb by multiply-high: exact for e*kb < 2^32)
        const uint32_t total = (uint32_t)(kb * kb), magic = (uint32_t)((0x100000000ull + kb - 1) / kb);
        for (uint32_t e0 = tid; e0 < total; e0 += 8 * THREADS) {
            T v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const uint32_t e = e0 + u * THREADS;
                const uint32_t r = __umulhi(e, magic);
                if (e < total) v[u] = Ginv[(size_t)r * a.Kmax + (e - r * kb)];
            }
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const uint32_t e = e0 + u * THREADS;
                const uint32_t r = __umulhi(e, magic);
                if (e < total) s_G[r * ks + (e - r * kb)] = v[u];
            }
        }
    }
    for (int w = tid; w < mask_words; w += THREADS) mask[w] = 0u;
    __syncthreads();
    if (a.exact) {
        lam = red_v[0];
#pragma unroll
        for (int w = 1; w < NW; ++w) lam = max(lam, red_v[w]);
    }
    // per-slot values stay in registers (slot k lives in thread k % THREADS) so that no load is waited for
    // before its value is needed: x at the active atoms, the Gram row of a joining atom, c at the active atoms
    T xk[KR], u1r[KR], cgr[KR];   // K <= KR * THREADS (checked by the host)
#pragma unroll
    for (int u = 0; u < KR; ++u) {
        const int k = tid + u * THREADS;
        if (k < K) {
            const uint32_t g = s_gam[k];
            atomicOr(&mask[g >> 5], 1u << (g & 31));
            xk[u] = x[g];
        }
    }
    __syncthreads();

    // ---- min-gamma search (find_max_gamma part 2, :122-163): smallest strictly positive step, ties -> lowest index
    T best = TMAX;
    uint32_t bidx = 0xffffffffu;
    // float32: candidates are compared as integer keys bits(t) - 1, unsigned: the order of positive floats, while
    // zero, negatives, inf and NaN all land above bits(FLT_MAX) - 1 and lose -- one min per candidate instead of
    // two compares and two selects, and the (rare) new minimum is located afterwards.
    uint32_t bestkey = __float_as_uint(FLT_MAX) - 1u;
    auto scan = [&](const T (&cc)[V], const T (&qq)[V], int i0) {
        const uint32_t mw = mask[i0 >> 5] >> (i0 & 31);
        if constexpr (sizeof(T) == 4) {
            uint32_t key[V];
#pragma unroll
            for (int e = 0; e < V; ++e) {
                const float t1 = step_ratio(lam - cc[e], 1.0f - qq[e]), t2 = step_ratio(lam + cc[e], 1.0f + qq[e]);
                const uint32_t k12 = min(__float_as_uint(t1) - 1u, __float_as_uint(t2) - 1u);
                key[e] = ((mw >> e) & 1u) || (i0 + e >= a.N) ? 0xffffffffu : k12;
            }
            uint32_t m = key[0];
#pragma unroll
            for (int e = 1; e < V; ++e) m = min(m, key[e]);
            if (m < bestkey) {
                bestkey = m;
#pragma unroll
                for (int e = V - 1; e >= 0; --e) if (key[e] == m) bidx = (uint32_t)(i0 + e);
            }
        } else {
#pragma unroll
            for (int e = 0; e < V; ++e) {
                const int i = i0 + e;
                if (i < a.N && !((mw >> e) & 1u)) {
                    const T dl = T(1) - qq[e], dr = T(1) + qq[e];
                    if (dl != T(0)) { const T t = step_ratio(lam - cc[e], dl); if (t > T(0) && t < best) { best = t; bidx = (uint32_t)i; } }
                    if (dr != T(0)) { const T t = step_ratio(lam + cc[e], dr); if (t > T(0) && t < best) { best = t; bidx = (uint32_t)i; } }
                }
            }
        }
    };
    if constexpr (NCH > 0) {
#pragma unroll
        for (int ch = 0; ch < NCH; ++ch) {
            const int i0 = (ch * THREADS + tid) * V;
            if (i0 < a.N) scan(cv[ch], qv[ch], i0);
        }
    } else {
        for (int i0 = tid * V; i0 < a.N; i0 += THREADS * V) {
            T cc[V], qq[V];
            *reinterpret_cast<int4*>(cc) = *reinterpret_cast<const int4*>(csrc + i0);
            *reinterpret_cast<int4*>(qq) = *reinterpret_cast<const int4*>(q + i0);
            scan(cc, qq, i0);
        }
    }
    if constexpr (sizeof(T) == 4) best = bidx == 0xffffffffu ? TMAX : __uint_as_float(bestkey + 1u);
#pragma unroll
    for (int u = 0; u < KR; ++u) {   // leave candidates, -x_i / d_i (:129-137)
        const int k = tid + u * THREADS;
        if (k < K) {
            const uint32_t i = s_gam[k];
            const T t = -xk[u] / s_d[k];
            if (t > T(0) && t < TMAX && cand_less(t, i, best, bidx)) { best = t; bidx = i; }
        }
    }
    warp_argmin_fast(best, bidx);
    if (lane == 0) { red_v[warp] = best; red_i[warp] = bidx; }
    __syncthreads();
    if (warp == 0) {
        best = lane < NW ? red_v[lane] : TMAX;
        bidx = lane < NW ? red_i[lane] : 0xffffffffu;
        warp_argmin_fast(best, bidx);
        if (lane == 0) {
            sh_join = bidx != 0xffffffffu;
            if (bidx == 0xffffffffu) bidx = 0;   // no candidate: (max, 0) -- the caller toggles column 0 (:122-123)
            sh_gamma = best;
            sh_idx = bidx;
            sh_slot = -1;
        }
    }
    __syncthreads();
    const uint32_t idx = sh_idx;
    const bool present = (mask[idx >> 5] >> (idx & 31)) & 1u;
    if constexpr (sizeof(T) == 4) {
        // the winning join step again, in IEEE arithmetic (the ordering used the hardware reciprocal)
        if (sh_join && !present && !SSB_IEEE_ORDER) {
            if constexpr (NCH > 0) {
                if ((uint32_t)tid == (idx / V) % THREADS) {   // the thread that holds atom idx: selects, then ONE pair of divisions
                    T cw = T(0), qw = T(0);
#pragma unroll
                    for (int ch = 0; ch < NCH; ++ch) {
                        const int i0 = (ch * THREADS + tid) * V;
#pragma unroll
                        for (int e = 0; e < V; ++e)
                            if ((uint32_t)(i0 + e) == idx) { cw = cv[ch][e]; qw = qv[ch][e]; }
                    }
                    sh_gamma = winning_step(lam, cw, qw, sh_gamma);
                }
            } else if (tid == 0) sh_gamma = winning_step(lam, csrc[idx], q[idx], sh_gamma);
            __syncthreads();   // sh_join and present are block-uniform
        }
    }
    const T gamma = sh_gamma;
    if (present)
        for (int k = tid; k < K; k += THREADS) if (s_gam[k] == idx) sh_slot = k;
    const int Knew = present ? K - 1 : K + 1;

    if (Knew == 0) {   // :248-249: the last atom left; x, c and lambda stay as they were
        if (tid == 0) { a.iters[s] = a.it; a.errs[s] = (double)lam; a.Kcur[s] = 0; }
        return;
    }

    // ---- loads that depend on idx, issued before the streaming update so their latency is hidden
    T dot = T(0);
    if (!present) {
#pragma unroll
        for (int u = 0; u < KR; ++u) {                                                         // u1 = A_G^T v (online_inverse.h:218)
            const int k = tid + u * THREADS;
            if (k < K) u1r[u] = a.G[(size_t)idx * a.ldg + s_gam[k]];
        }
        dot = a.G[(size_t)idx * a.ldg + idx];                                                 // v.v
        if (a.exact && tid == 0) s_x[K] = x[idx];   // 0, or the rounding residue of an atom that left earlier
    }

    // ---- x += gamma * d with the OLD direction (:252): a leaving atom lands on ~0, a joining one stays 0.  Multiply and add are
    // rounded separately, as in the reference's build: the leaving atom's x + fl(gamma d) with gamma = fl(-x / d) is then
    // EXACTLY zero most of the time, where a fused multiply-add leaves the division's rounding error behind almost always --
    // and a non-zero residue of the "wrong" sign makes the atom leave again the moment it rejoins (:129-137), after which it
    // sits outside the active set with the largest |c| and the solve runs to max_iterations (measured on the reference's own
    // bench workload: 12 of 32 solves with the fused form against 3 of 32 for the reference, tests/studies/)
#pragma unroll
    for (int u = 0; u < KR; ++u) {
        const int k = tid + u * THREADS;
        if (k < K) { const T xn = add_rn(xk[u], mul_rn(gamma, s_d[k])); x[s_gam[k]] = xn; if (a.exact) s_x[k] = xn; }
    }

    // ---- c = A^T (y - A x)  ==  c - gamma * q ;  lambda = |c|_inf (:255, :270).  Exact mode: c is this iteration's fresh
    // c' = A^T r, so the result is one step from a re-formed base; incremental mode: c carries over from iteration to iteration
    T mx = T(0);
    auto update = [&](T (&cc)[V], const T (&qq)[V], int i0) {
#pragma unroll
        for (int e = 0; e < V; ++e) {
            cc[e] = cc[e] - gamma * qq[e];
            if (i0 + e < a.N) mx = max(mx, fabs(cc[e]));
        }
        *reinterpret_cast<int4*>(c + i0) = *reinterpret_cast<const int4*>(cc);
    };
    if constexpr (NCH > 0) {
#pragma unroll
        for (int ch = 0; ch < NCH; ++ch) {
            const int i0 = (ch * THREADS + tid) * V;
            if (i0 < a.N) update(cv[ch], qv[ch], i0);
        }
    } else {
        for (int i0 = tid * V; i0 < a.N; i0 += THREADS * V) {
            T cc[V], qq[V];
            *reinterpret_cast<int4*>(cc) = *reinterpret_cast<const int4*>(csrc + i0);
            *reinterpret_cast<int4*>(qq) = *reinterpret_cast<const int4*>(q + i0);
            update(cc, qq, i0);
        }
    }
    mx = warp_max_nonneg(mx);
    if (lane == 0) red_v[warp] = mx;
    if (!present) {
#pragma unroll
        for (int u = 0; u < KR; ++u) { const int k = tid + u * THREADS; if (k < K) s_u1[k] = u1r[u]; }
    }
    __syncthreads();   // also orders sh_slot, s_u1 and the c stores
    if (tid == 0) {
        T m = red_v[0];
        for (int w = 1; w < NW; ++w) m = max(m, red_v[w]);
        sh_lam = m;
    }
    // c at the (old-slot) active atoms and at a joining atom, for the signs; read back through L2 while the
    // inverse is being updated
    const int ncg = K + (present ? 0 : 1);
#pragma unroll
    for (int u = 0; u < KR; ++u) {
        const int k = tid + u * THREADS;
        if (k < ncg) cgr[u] = __ldcg(c + (k < K ? s_gam[k] : idx));
    }

    if (a.dbg == 1) { if (tid == 0) { const int row = atomicAdd(a.next_count, 1); a.act_next[row] = s; } return; }
    // ---- toggle idx and update inv = (A_G^T A_G)^-1 (inverse_add_or_remove, :166-183)
    if (!present) {
        // online_inverse.h:203-249 (bordering); the new atom takes slot K
        matvec<T, THREADS, GS>(Gm, ldm, K, s_u1, s_u2, nullptr);                               // u2 = inv u1
        __syncthreads();
        T uu = T(0);
        for (int k = lane; k < K; k += 32) uu += s_u1[k] * s_u2[k];
        uu = warp_sum(uu);
        const T dd = T(1) / (dot - uu);                                                       // :228
        if (K >= 24) {                                                                        // inv += d u2 u2^T
            // a warp per row, lanes along it: per element a load, a multiply-add and a store (the flat loop below pays an index
            // split and two more loads per element -- ncu r2, C3: this kernel is issue-bound on exactly that pass); same
            // expression per element, so the same bits
            for (int r = warp; r < K; r += NW) {
                const T t = dd * s_u2[r];
                T* g = Gm + (size_t)r * ldm;
                for (int k = lane; k < K; k += 32) g[k] += t * s_u2[k];
            }
        } else {
            // row = e / K by multiply-high, exact while e * K < 2^32, i.e. K < 1625; a plain division beyond
            const uint32_t total = (uint32_t)(K * K), magic = (uint32_t)((0x100000000ull + K - 1) / K);
            for (uint32_t e = tid; e < total; e += THREADS) {
                const uint32_t r = __umulhi(e, magic), k = e - r * K;
                Gm[(size_t)r * ldm + k] += (dd * s_u2[r]) * s_u2[k];
            }
        }
        for (int r = tid; r < K; r += THREADS) {                                              // border -d u2, corner d
            const T u3 = -dd * s_u2[r];
            Gm[(size_t)r * ldm + K] = u3;
            Gm[(size_t)K * ldm + r] = u3;
        }
        if (tid == 0) { Gm[(size_t)K * ldm + K] = dd; s_gam[K] = idx; gam[K] = idx; }
#pragma unroll
        for (int u = 0; u < KR; ++u) { const int k = tid + u * THREADS; if (k < ncg) s_cg[k] = cgr[u]; }
    } else {
        // online_inverse.h:253-293; the leaving atom sits in slot j
        const int j = sh_slot;
        const T dd = Gm[(size_t)j * ldm + j];
        const T sc = -(T(1) / dd);
        for (int r = tid; r < K; r += THREADS) s_u1[r] = Gm[(size_t)r * ldm + j] * sc;         // u = -col/d
        __syncthreads();
        for (int r = warp; r < K; r += NW) {                                                  // inv -= d u u^T
            if (r == j) continue;
            const T ur = -dd * s_u1[r];
            for (int k = lane; k < K; k += 32) if (k != j) Gm[(size_t)r * ldm + k] += ur * s_u1[k];
        }
#pragma unroll
        for (int u = 0; u < KR; ++u) { const int k = tid + u * THREADS; if (k < ncg) s_cg[k] = cgr[u]; }
        __syncthreads();
        const int last = K - 1;
        if (j != last) {   // refill slot j with the last slot (reads touch row/col `last` only, writes never do)
            for (int k = tid; k < last; k += THREADS) {
                const int kk = (k == j) ? last : k;
                const T vrow = Gm[(size_t)last * ldm + kk];
                const T vcol = Gm[(size_t)kk * ldm + last];
                Gm[(size_t)j * ldm + k] = vrow;
                Gm[(size_t)k * ldm + j] = vcol;
            }
            if (tid == 0) { const uint32_t g = s_gam[last]; s_gam[j] = g; gam[j] = g; s_cg[j] = s_cg[last]; if (a.exact) s_x[j] = s_x[last]; }
        }
    }
    __syncthreads();

    // ---- d_G = inv * sign(c_G, tol) (:259-263); the tolerance is also the sign dead-band
    for (int k = tid; k < Knew; k += THREADS) s_cg[k] = sign_tol(s_cg[k], a.tol);
    __syncthreads();
    matvec<T, THREADS, GS>(Gm, ldm, Knew, s_cg, s_d, dG);
    if constexpr (GS) {   // the updated inverse goes back to HBM for the next iteration
        const uint32_t total = (uint32_t)(Knew * Knew), magic = (uint32_t)((0x100000000ull + Knew - 1) / Knew);
        for (uint32_t e = tid; e < total; e += THREADS) {
            const uint32_t r = __umulhi(e, magic), k = e - r * Knew;
            Ginv[(size_t)r * a.Kmax + k] = s_G[r * ks + k];
        }
    }
    const T lam_new = sh_lam;
    const bool cont = (a.it < a.max_iter) && (lam_new > a.tol);   // :272
    if (tid == 0) {
        a.lam[s] = lam_new;
        a.Kcur[s] = Knew;
        if (a.mask) a.mask[(size_t)s * (a.ldc >> 5) + (idx >> 5)] ^= 1u << (idx & 31);
        if (cont) {
            const int row = atomicAdd(a.next_count, 1);
            a.act_next[row] = s;
            sh_row = row;
        } else {
            a.iters[s] = a.it;
            a.errs[s] = (double)lam_new;   // :274
        }
    }
    __syncthreads();
    if (cont && a.dbg != 2) write_rhs<T, THREADS>(a, sh_row, s, s_gam, s_d, s_x, Knew);
}

// dynamic shared memory of the step kernel: per-slot vectors, membership mask, and (when held on chip) the inverse
template <typename T>
inline size_t homotopy_step_smem(int Kmax, int ldc, int g_rows, int ks) {
    const size_t mask_words = (size_t)((ldc + 31) / 32 + 3) & ~(size_t)3;
    return (size_t)((Kmax + 2) & ~1) * (5 * sizeof(T) + sizeof(uint32_t)) + mask_words * sizeof(uint32_t) + (size_t)g_rows * ks * sizeof(T) + 16;
}

}  // namespace ssb
