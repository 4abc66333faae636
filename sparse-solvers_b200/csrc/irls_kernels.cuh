// IRLS solve loop on the device (reference: src/solvers/irls-cpu.cpp:39-124 with
// src/linalg/cholesky_decomposition.h:57-103).  One CTA owns one signal for its WHOLE solve: unlike
// homotopy there is no full-dictionary product inside the loop once Q^T y is known, so nothing has
// to run in lockstep across signals.
//
// With A = QR (thin, M >= N) the reference forms, every iteration,
//     G = Q^T (Q o w),  L = chol(tril(G)),  s = L^-T L^-1 (Q^T y),  x = R^-1 Q^T (Q s).
// Q^T y comes from one batched contraction; t = Q s and x = Q^T t are formed as the reference forms them (two products with
// Q, irls-cpu.cpp:55-57).  Replacing them by E s with a precomputed E = Q^T Q is the same algebra, but once the iterates blow
// up (at the C4 shape they pass 1e23 by the fifth iteration and the solve ends when a weight drops under eps^2) what is
// amplified is the rounding structure of exactly these products, and the shortcut ended one solve in four an iteration late.
// G is formed the reference's way -- column j of Q scaled by w_j first, then the dot products over
// the M rows in order, multiply and add rounded separately -- because the way most solves END depends
// on it: the last pivots are w_j minus rounding-level terms, and whether sqrt(pivot) comes out
// <= eps ("not SPD", cholesky_decomposition.h:80-83) or NaN (which passes that test) is decided by
// those roundings.  A cheaper G_ij = E_ij * w_j flips that outcome on ~10 % of signals (measured).
#pragma once

#include "common.cuh"

namespace ssb {

template <typename T>
struct IrlsArgs {
    int N;              // columns of A
    int lde;            // leading dimension of E and S (= Npad)
    T* Tq;              // [rows][ldq] scratch: t = Q s (the staged signal row, free once Q^T y is formed)
    const T* Qt;        // [Npad][ldq] columns of Q, contiguous
    int M, ldq;
    const T* Rc;        // [N][N]     R stored column-major: Rc[i*N + k] = R[k][i]
    const T* S;         // [rows][lde] Q^T y
    T* X; size_t ldx;   // output
    T* L;               // [rows][N*N] Cholesky factor, column-major: L[k*N + i] = l_ik
    T* vec;             // [rows][4*N]  w | xnext | s | (spare)
    uint32_t* iters; double* errs; uint8_t* spd;
    T tol; uint32_t max_iter;
};

// (mul_rn / add_rn, common.cuh: separately rounded multiply / add -- an FMA is something the reference's build does not use)

template <typename T, int THREADS>
__device__ __forceinline__ T block_sum(T v, T* red) {
    constexpr int NW = THREADS / 32;
    v = warp_sum(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    T t = T(0);
#pragma unroll
    for (int w = 0; w < NW; ++w) t += red[w];
    return t;
}

// acc[x][y] += sum over k in [k0, k1), IN ORDER, of u(k, i) * v(k, j) for the 4 x 4 elements this thread owns in the
// 64 x 64 tile (it, jt): i = 64 it + ti + 16 x, j = 64 jt + tj + 16 y.  Multiply and add are rounded separately, one
// term at a time, exactly like the scalar loops they replace; the 16 chains of a thread advance together (that is what
// fills the FP64 pipe) and each operand row is read once per tile instead of once per element.
//   GRAM   u(k, i) = Q[k][i],  v(k, j) = Q[k][j] * w_j   (G = Q^T (Q o w), irls-cpu.cpp:47-48; rows of Qt are columns of Q)
//   !GRAM  u(k, i) = L[i][k],  v(k, j) = L[j][k]         (left-looking Cholesky update, cholesky_decomposition.h:77)
// Rows / columns past N and rows past M are zero padding: they add +0.
template <typename T, int THREADS, bool GRAM>
__device__ __forceinline__ void tile_mac(const T* __restrict__ src, int ld, int N, const T* __restrict__ w, int it, int jt,
                                         int k0, int k1, T (&acc)[4][4], T (*sm_i)[65], T (*sm_j)[65])
{
    static_assert(THREADS == 256, "16 x 16 threads, 4 x 4 elements each");
    constexpr int TB = 64, MC = 16;
    const int tid = threadIdx.x, ti = tid & 15, tj = tid >> 4;
    for (int kc = k0; kc < k1; kc += MC) {
        for (int e = tid; e < TB * MC; e += THREADS) {
            if constexpr (GRAM) {
                const int r = e / MC, kk = e % MC;
                const int cj = jt * TB + r;
                sm_i[kk][r] = src[(size_t)(it * TB + r) * ld + kc + kk];
                sm_j[kk][r] = mul_rn(src[(size_t)cj * ld + kc + kk], cj < N ? w[cj] : T(0));
            } else {
                const int kk = e / TB, r = e % TB;
                const int ri = it * TB + r, rj = jt * TB + r;
                sm_i[kk][r] = ri < N ? src[(size_t)(kc + kk) * ld + ri] : T(0);
                sm_j[kk][r] = rj < N ? src[(size_t)(kc + kk) * ld + rj] : T(0);
            }
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < MC; ++kk) {
            T u[4], v[4];
#pragma unroll
            for (int x = 0; x < 4; ++x) { u[x] = sm_i[kk][ti + 16 * x]; v[x] = sm_j[kk][tj + 16 * x]; }
#pragma unroll
            for (int x = 0; x < 4; ++x)
#pragma unroll
                for (int y = 0; y < 4; ++y) acc[x][y] = add_rn(acc[x][y], mul_rn(u[x], v[y]));
        }
        __syncthreads();
    }
}

template <typename T, int THREADS>
__global__ void __launch_bounds__(THREADS) irls_solve_kernel(IrlsArgs<T> a)
{
    constexpr int NW = THREADS / 32;
    __shared__ T red[NW];
    __shared__ T red2[NW];
    __shared__ T sh_piv;
    __shared__ int sh_bad;
    __shared__ T sm_i[16][65], sm_j[16][65];   // operand chunks of the Gram / Cholesky tiles
    __shared__ T diag_acc[64];

    const int N = a.N;
    const size_t b = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    T* x = a.X + b * a.ldx;
    T* L = a.L + b * (size_t)N * N;
    T* w = a.vec + b * 4 * (size_t)N;
    T* xn = w + N;
    T* sv = xn + N;
    const T* qty = a.S + b * (size_t)a.lde;
    const T EPS = num_traits<T>::eps();
    const double ex = ((double)(T)0.9) / 2.0 - 1.0;   // p = 0.9 (irls-cpu.cpp:70)

    for (int j = tid; j < N; j += THREADS) { x[j] = T(0); w[j] = T(1); xn[j] = T(1); }   // :82-85
    uint32_t iter = 0;
    int spd_error = 0;
    T abstol = T(1), eps = T(1), second = T(1);
    __syncthreads();

    do {
        // ---- irls_newton (:39-60).  G = Q^T (Q o w) on the lower triangle (:47-48), left-looking Cholesky.
        int bad = 0;
        constexpr int TB = 64;
        const int NT = (N + TB - 1) / TB;
        const int ti = tid & 15, tj = tid >> 4;
        // G = Q^T (Q o w), lower triangle, into L (:47-48)
        for (int jt = 0; jt < NT; ++jt)
            for (int it = jt; it < NT; ++it) {
                T acc[4][4] = {};
                tile_mac<T, THREADS, true>(a.Qt, a.ldq, N, w, it, jt, 0, a.ldq, acc, sm_i, sm_j);
#pragma unroll
                for (int x = 0; x < 4; ++x)
#pragma unroll
                    for (int y = 0; y < 4; ++y) {
                        const int i = it * TB + ti + 16 * x, j = jt * TB + tj + 16 * y;
                        if (i < N && j < N && i >= j) L[(size_t)j * N + i] = acc[x][y];
                    }
            }
        __syncthreads();
        // Left-looking Cholesky (cholesky_decomposition.h:57-86), blocked by 64 columns: the part of every element's sum
        // that runs over finished blocks (k < 64 jt) is formed tile by tile and parked in the unused upper triangle
        // (the diagonal in shared memory); the columns of the block then finish their sums one after the other.
        for (int jt = 0; jt < NT; ++jt) {
            const int kend = jt * TB;
            if (kend > 0) {
                for (int it = jt; it < NT; ++it) {
                    T acc[4][4] = {};
                    tile_mac<T, THREADS, false>(L, N, N, nullptr, it, jt, 0, kend, acc, sm_i, sm_j);
#pragma unroll
                    for (int x = 0; x < 4; ++x)
#pragma unroll
                        for (int y = 0; y < 4; ++y) {
                            const int i = it * TB + ti + 16 * x, j = jt * TB + tj + 16 * y;
                            if (i < N && j < N && i >= j) { if (i == j) diag_acc[j - kend] = acc[x][y]; else L[(size_t)i * N + j] = acc[x][y]; }
                        }
                }
                __syncthreads();
            }
            const int jend = min(N, kend + TB);
            for (int j = kend; j < jend; ++j) {
            for (int i = j + tid; i < N; i += THREADS) {
                T acc = kend > 0 ? (i == j ? diag_acc[j - kend] : L[(size_t)i * N + j]) : T(0);
                for (int k = kend; k < j; ++k) acc = add_rn(acc, mul_rn(L[(size_t)k * N + i], L[(size_t)k * N + j]));
                L[(size_t)j * N + i] = L[(size_t)j * N + i] - acc;
            }
            __syncthreads();
            const T ajj = sqrt(L[(size_t)j * N + j]);
            if (ajj <= EPS) bad = 1;           // NaN pivots pass, as in the reference
            const T inv = T(1) / ajj;
            __syncthreads();
            for (int i = j + tid; i < N; i += THREADS) L[(size_t)j * N + i] *= inv;
            __syncthreads();
            }
        }
        if (bad) { spd_error = 1; break; }      // :94-97, x keeps the previous iterate

        // s = L^-T L^-1 (Q^T y)  (cholesky_decomposition.h:93-103)
        for (int j = tid; j < N; j += THREADS) sv[j] = qty[j];
        __syncthreads();
        for (int i = 0; i < N; ++i) {           // forward, column sweeps
            const T zi = sv[i] / L[(size_t)i * N + i];
            __syncthreads();
            if (tid == 0) sv[i] = zi;
            for (int k = i + 1 + tid; k < N; k += THREADS) sv[k] -= L[(size_t)i * N + k] * zi;
            __syncthreads();
        }
        for (int i = N - 1; i >= 0; --i) {      // backward, dot with column i of L below the diagonal
            T acc = T(0);
            for (int k = i + 1 + tid; k < N; k += THREADS) acc += L[(size_t)i * N + k] * sv[k];
            acc = block_sum<T, THREADS>(acc, red);
            if (tid == 0) sv[i] = (sv[i] - acc) / L[(size_t)i * N + i];
            __syncthreads();
        }
        // t = Q s (:55), xnext = Q^T t (:57)
        T* tq = a.Tq + b * (size_t)a.ldq;
        for (int m = tid; m < a.ldq; m += THREADS) {       // column j of Q is row j of Qt: coalesced over m, j in order
            T acc = T(0);
            for (int j = 0; j < N; ++j) acc += a.Qt[(size_t)j * a.ldq + m] * sv[j];
            tq[m] = acc;
        }
        __syncthreads();
        for (int r = warp; r < N; r += NW) {
            T acc = T(0);
            for (int m = lane; m < a.M; m += 32) acc += a.Qt[(size_t)r * a.ldq + m] * tq[m];
            acc = warp_sum(acc);
            if (lane == 0) xn[r] = acc;
        }
        __syncthreads();
        // R xnext = xnext, R upper triangular (:58), column sweeps from the last unknown
        for (int i = N - 1; i >= 0; --i) {
            const T xi = xn[i] / a.Rc[(size_t)i * N + i];
            __syncthreads();
            if (tid == 0) xn[i] = xi;
            for (int k = tid; k < i; k += THREADS) xn[k] -= a.Rc[(size_t)i * N + k] * xi;
            __syncthreads();
        }

        // ---- thresholding and reweighting (:100-114)
        T mx = -num_traits<T>::max();
        for (int j = tid; j < N; j += THREADS) mx = max(mx, xn[j]);   // signed max
        mx = warp_max(mx);
        if (lane == 0) red[warp] = mx;
        __syncthreads();
        mx = red[0];
        for (int q = 1; q < NW; ++q) mx = max(mx, red[q]);
        // a NaN pivot (sqrt of a negative rounding-level pivot passes the SPD test) turns every entry into NaN;
        // the reference's running max / nth_element then carry the NaN of entry 0 through (irls-cpu.cpp:100,:107)
        const T xn0 = xn[0];
        const bool all_nan = xn0 != xn0;
        if (all_nan) mx = xn0;
        abstol = mx * a.tol;
        __syncthreads();
        T top = -num_traits<T>::max(), sec = -num_traits<T>::max();
        for (int j = tid; j < N; j += THREADS) {
            T v = xn[j];
            if (v < abstol) v = T(0);           // also zeroes every negative entry
            xn[j] = v;
            x[j] = v;
            if (v > top) { sec = top; top = v; } else if (v > sec) sec = v;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {      // second-largest value (nth_element at :107)
            const T ot = __shfl_xor_sync(0xffffffffu, top, o), os = __shfl_xor_sync(0xffffffffu, sec, o);
            const T nt = max(top, ot);
            sec = max(min(top, ot), max(sec, os));
            top = nt;
        }
        if (lane == 0) { red[warp] = top; red2[warp] = sec; }
        __syncthreads();
        top = red[0]; sec = red2[0];
        for (int q = 1; q < NW; ++q) {
            const T ot = red[q], os = red2[q];
            const T nt = max(top, ot);
            sec = max(min(top, ot), max(sec, os));
            top = nt;
        }
        second = (N > 1) ? sec : top;
        if (all_nan) second = xn0;
        if (second / (T)N < eps) eps = second / (T)N;                 // :110
        __syncthreads();
        T sw = T(0);
        for (int j = tid; j < N; j += THREADS) {
            const T base = x[j] * x[j] + eps;
            const T wv = (T)pow((double)base, ex);                    // :113
            w[j] = wv;
            sw += wv;
        }
        sw = block_sum<T, THREADS>(sw, red);
        for (int j = tid; j < N; j += THREADS) w[j] /= sw;            // :114
        __syncthreads();
        iter++;
    } while (iter < a.max_iter && second > abstol);                   // :118

    __syncthreads();
    T sx = T(0);
    for (int j = tid; j < N; j += THREADS) sx += x[j];
    sx = block_sum<T, THREADS>(sx, red);
    for (int j = tid; j < N; j += THREADS) x[j] /= sx;                // :121
    if (tid == 0) { a.iters[b] = iter; a.errs[b] = (double)eps; a.spd[b] = (uint8_t)spd_error; }
    (void)sh_piv; (void)sh_bad;
}

}  // namespace ssb
