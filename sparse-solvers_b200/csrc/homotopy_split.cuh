// Split form of the homotopy iteration (sm_100a): the same arithmetic as homotopy_step_kernel, bit for bit, cut
// where its two bounds meet.
//   homotopy_scan_kernel     one CTA per active signal, HBM-bound streaming: min-gamma search over join and leave
//                            candidates (homotopy-cpu.cpp:122-163), x += gamma d (:252), c -= gamma q, |c|_inf (:255,
//                            :270), stop test and compaction (:272).  No inverse, no gather: few registers, high occupancy.
//   homotopy_algebra_kernel  one WARP per continuing signal, latency-bound small algebra with no block barrier:
//                            toggle + online inverse update (online_inverse.h:183-293), d = inv sign(c_G) (:259-263),
//                            p = A_G d for the next contraction (:112-116).
// The monolithic kernel stays as the fallback for active sets too large for a warp's shared-memory slice.
#pragma once

#include "homotopy_kernels.cuh"

#ifndef SSB_SCAN_MINB
#define SSB_SCAN_MINB 8   // resident scan CTAs of 128 threads per SM the register allocation aims for
#endif

namespace ssb {

template <typename T, int THREADS, int NCH>
__global__ void __launch_bounds__(THREADS, SSB_SCAN_MINB * 128 / THREADS) homotopy_scan_kernel(HomArgs<T> a)
{
    constexpr int V = 16 / sizeof(T);
    constexpr int NW = THREADS / 32;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ T red_v[NW];
    __shared__ uint32_t red_i[NW];
    __shared__ T sh_gamma;
    __shared__ uint32_t sh_idx;

    const int KC = (a.Kmax + 2) & ~1;
    T* s_d = reinterpret_cast<T*>(smem_raw);
    uint32_t* s_gam = reinterpret_cast<uint32_t*>(s_d + KC);
    uint32_t* mask = s_gam + KC;
    const int mask_words = (a.ldc + 31) >> 5;

    const int pos = blockIdx.x;
    if (a.count_in && pos >= *a.count_in) return;   // grid sized from a stale (larger) count
    const int s = a.act ? a.act[pos] : pos;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    T* c = a.C + (size_t)s * a.ldc;
    const T* q = a.Q + (size_t)pos * a.ldc;
    T* x = a.X + (size_t)s * a.ldx;
    const T TMAX = num_traits<T>::max();

    // one round trip: c and q into registers, the active set and its direction into shared memory
    T cv[NCH > 0 ? NCH : 1][V], qv[NCH > 0 ? NCH : 1][V];
    if constexpr (NCH > 0) {
#pragma unroll
        for (int ch = 0; ch < NCH; ++ch) {
            const int i0 = (ch * THREADS + tid) * V;
            if (i0 < a.ldc) {
                *reinterpret_cast<int4*>(cv[ch]) = *reinterpret_cast<const int4*>(c + i0);
                *reinterpret_cast<int4*>(qv[ch]) = *reinterpret_cast<const int4*>(q + i0);
            }
        }
    }
    const int K = a.Kcur[s];
    const T lam = a.lam[s];
    for (int k = tid; k < K; k += THREADS) { s_gam[k] = a.gam[(size_t)s * a.Kmax + k]; s_d[k] = a.dG[(size_t)s * a.Kmax + k]; }
    for (int w = tid; w < mask_words; w += THREADS) mask[w] = 0u;
    __syncthreads();
    constexpr int KR = 3;   // K <= KR * THREADS (checked by the host)
    T xk[KR];
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

    // ---- min-gamma search (find_max_gamma part 2, :122-163): smallest strictly positive step, ties -> lowest index.
    // float32 candidates are compared as integer keys bits(t) - 1 (see homotopy_step_kernel).
    T best = TMAX;
    uint32_t bidx = 0xffffffffu;
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
            *reinterpret_cast<int4*>(cc) = *reinterpret_cast<const int4*>(c + i0);
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
    warp_argmin(best, bidx);
    if (lane == 0) { red_v[warp] = best; red_i[warp] = bidx; }
    __syncthreads();
    if (warp == 0) {
        best = lane < NW ? red_v[lane] : TMAX;
        bidx = lane < NW ? red_i[lane] : 0xffffffffu;
        warp_argmin(best, bidx);
        if (lane == 0) {
            if (bidx == 0xffffffffu) bidx = 0;   // no candidate: (max, 0) -- the caller toggles column 0 (:122-123)
            sh_gamma = best;
            sh_idx = bidx;
        }
    }
    __syncthreads();
    const T gamma = sh_gamma;
    const uint32_t idx = sh_idx;
    const bool present = (mask[idx >> 5] >> (idx & 31)) & 1u;
    const int Knew = present ? K - 1 : K + 1;
    if (Knew == 0) {   // :248-249: the last atom left; x, c and lambda stay as they were
        if (tid == 0) { a.iters[s] = a.it; a.errs[s] = (double)lam; a.Kcur[s] = 0; }
        return;
    }

    // ---- x += gamma * d with the OLD direction (:252): a leaving atom lands on ~0, a joining one stays 0
#pragma unroll
    for (int u = 0; u < KR; ++u) {
        const int k = tid + u * THREADS;
        if (k < K) x[s_gam[k]] = xk[u] + gamma * s_d[k];
    }

    // ---- c = A^T (y - A x)  ==  c - gamma * q ;  lambda = |c|_inf (:255, :270)
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
            *reinterpret_cast<int4*>(cc) = *reinterpret_cast<const int4*>(c + i0);
            *reinterpret_cast<int4*>(qq) = *reinterpret_cast<const int4*>(q + i0);
            update(cc, qq, i0);
        }
    }
    mx = warp_max(mx);
    if (lane == 0) red_v[warp] = mx;
    __syncthreads();
    if (tid == 0) {
        T m = red_v[0];
        for (int w = 1; w < NW; ++w) m = max(m, red_v[w]);
        a.lam[s] = m;
        a.pick[s] = idx;
        if ((a.it < a.max_iter) && (m > a.tol)) {   // :272 -- the toggle, inverse and direction follow in the algebra kernel
            const int row = atomicAdd(a.next_count, 1);
            a.act_next[row] = s;
        } else {
            a.iters[s] = a.it;
            a.errs[s] = (double)m;   // :274
            a.Kcur[s] = Knew;
        }
    }
}

template <typename T>
inline size_t homotopy_scan_smem(int Kmax, int ldc) {
    const size_t mask_words = (size_t)((ldc + 31) / 32 + 3) & ~(size_t)3;
    return (size_t)((Kmax + 2) & ~1) * (sizeof(T) + sizeof(uint32_t)) + mask_words * sizeof(uint32_t) + 16;
}

// One warp per continuing signal (row of act_next).  K <= kb before the toggle; the inverse is staged in the warp's
// shared-memory slice with the odd row stride ks (lane = row: conflict-free), worked on, and written back.
template <typename T, int W>
__global__ void __launch_bounds__(W * 32) homotopy_algebra_kernel(HomArgs<T> a, int kb, int ks, int slice_bytes)
{
    constexpr int V = 16 / sizeof(T);
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int row = blockIdx.x * W + warp;
    if (row >= *a.next_count) return;
    const int s = a.act_next[row];

    const int KC = (kb + 3) & ~1;
    T* s_G = reinterpret_cast<T*>(smem_raw + (size_t)warp * slice_bytes);
    T* s_u1 = s_G + (size_t)(kb + 1) * ks + ((kb + 1) * ks & 1);
    T* s_u2 = s_u1 + KC;
    T* s_cg = s_u2 + KC;
    T* s_d = s_cg + KC;
    uint32_t* s_gam = reinterpret_cast<uint32_t*>(s_d + KC);

    uint32_t* gam = a.gam + (size_t)s * a.Kmax;
    T* Ginv = a.Ginv + (size_t)s * a.Kmax * a.Kmax;
    const T* c = a.C + (size_t)s * a.ldc;
    const int K = a.Kcur[s];
    const uint32_t idx = a.pick[s];

    // active set, membership of idx, and the inverse: all issued before anything is waited for
    int slot = -1;
    for (int k0 = 0; k0 < K; k0 += 32) {
        const int k = k0 + lane;
        uint32_t g = 0xffffffffu;
        if (k < K) { g = gam[k]; s_gam[k] = g; }
        const unsigned hit = __ballot_sync(0xffffffffu, g == idx);
        if (hit) slot = k0 + __ffs(hit) - 1;
    }
    const bool present = slot >= 0;
    {
        const uint32_t total = (uint32_t)(K * K), magic = (uint32_t)((0x100000000ull + K - 1) / K);
        for (uint32_t e0 = lane; e0 < total; e0 += 8 * 32) {
            T v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const uint32_t e = e0 + u * 32;
                const uint32_t r = __umulhi(e, magic);
                if (e < total) v[u] = Ginv[(size_t)r * a.Kmax + (e - r * K)];
            }
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const uint32_t e = e0 + u * 32;
                const uint32_t r = __umulhi(e, magic);
                if (e < total) s_G[r * ks + (e - r * K)] = v[u];
            }
        }
    }
    const int ncg = K + (present ? 0 : 1);
    T dot = T(0);
    __syncwarp();
    for (int k = lane; k < ncg; k += 32) s_cg[k] = c[k < K ? s_gam[k] : idx];
    if (!present) {
        for (int k = lane; k < K; k += 32) s_u1[k] = a.G[(size_t)idx * a.ldg + s_gam[k]];   // u1 = A_G^T v (online_inverse.h:218)
        dot = a.G[(size_t)idx * a.ldg + idx];                                              // v.v
    }
    __syncwarp();
    const int Knew = present ? K - 1 : K + 1;

    // ---- toggle idx and update inv = (A_G^T A_G)^-1 (inverse_add_or_remove, :166-183)
    if (!present) {
        // online_inverse.h:203-249 (bordering); the new atom takes slot K
        for (int r = lane; r < K; r += 32) {                                                // u2 = inv u1
            T acc = T(0);
            const T* g = s_G + (size_t)r * ks;
            for (int k = 0; k < K; ++k) acc += g[k] * s_u1[k];
            s_u2[r] = acc;
        }
        __syncwarp();
        T uu = T(0);
        for (int k = lane; k < K; k += 32) uu += s_u1[k] * s_u2[k];
        uu = warp_sum(uu);
        const T dd = T(1) / (dot - uu);                                                     // :228
        {                                                                                   // inv += d u2 u2^T
            const uint32_t total = (uint32_t)(K * K), magic = (uint32_t)((0x100000000ull + K - 1) / K);
            for (uint32_t e = lane; e < total; e += 32) {
                const uint32_t r = __umulhi(e, magic), k = e - r * K;
                s_G[r * ks + k] += (dd * s_u2[r]) * s_u2[k];
            }
        }
        for (int r = lane; r < K; r += 32) {                                                // border -d u2, corner d
            const T u3 = -dd * s_u2[r];
            s_G[r * ks + K] = u3;
            s_G[K * ks + r] = u3;
        }
        if (lane == 0) { s_G[K * ks + K] = dd; s_gam[K] = idx; gam[K] = idx; }
    } else {
        // online_inverse.h:253-293; the leaving atom sits in slot j
        const int j = slot;
        const T dd = s_G[j * ks + j];
        const T sc = -(T(1) / dd);
        for (int r = lane; r < K; r += 32) s_u1[r] = s_G[r * ks + j] * sc;                   // u = -col/d
        __syncwarp();
        {                                                                                   // inv -= d u u^T
            const uint32_t total = (uint32_t)(K * K), magic = (uint32_t)((0x100000000ull + K - 1) / K);
            for (uint32_t e = lane; e < total; e += 32) {
                const uint32_t r = __umulhi(e, magic), k = e - r * K;
                if ((int)r != j && (int)k != j) s_G[r * ks + k] += (-dd * s_u1[r]) * s_u1[k];
            }
        }
        __syncwarp();
        const int last = K - 1;
        if (j != last) {   // refill slot j with the last slot (reads touch row/col `last` only, writes never do)
            for (int k = lane; k < last; k += 32) {
                const int kk = (k == j) ? last : k;
                const T vrow = s_G[last * ks + kk];
                const T vcol = s_G[kk * ks + last];
                s_G[j * ks + k] = vrow;
                s_G[k * ks + j] = vcol;
            }
            if (lane == 0) { const uint32_t g = s_gam[last]; s_gam[j] = g; gam[j] = g; s_cg[j] = s_cg[last]; }
        }
    }
    __syncwarp();

    // ---- d_G = inv * sign(c_G, tol) (:259-263); the tolerance is also the sign dead-band
    for (int k = lane; k < Knew; k += 32) s_cg[k] = sign_tol(s_cg[k], a.tol);
    __syncwarp();
    for (int r = lane; r < Knew; r += 32) {
        T acc = T(0);
        const T* g = s_G + (size_t)r * ks;
        for (int k = 0; k < Knew; ++k) acc += g[k] * s_cg[k];
        s_d[r] = acc;
        a.dG[(size_t)s * a.Kmax + r] = acc;
    }
    {   // the updated inverse goes back for the next iteration
        const uint32_t total = (uint32_t)(Knew * Knew), magic = (uint32_t)((0x100000000ull + Knew - 1) / Knew);
        for (uint32_t e = lane; e < total; e += 32) {
            const uint32_t r = __umulhi(e, magic), k = e - r * Knew;
            Ginv[(size_t)r * a.Kmax + k] = s_G[r * ks + k];
        }
    }
    if (lane == 0) a.Kcur[s] = Knew;
    __syncwarp();

    // ---- p = A_G d for the next contraction (:112-116), k ascending per element as in write_rhs
    void *pn, *pl;
    rhs_rows(a, row, pn, pl);
    float scale = 1.f;
    if (a.split == 2) {   // |p_m| <= amax * sum |d_k|
        float sd = 0.f;
        for (int k = lane; k < Knew; k += 32) sd += fabsf((float)s_d[k]);
        sd = warp_sum(sd);
        const int e = f16_scale_exp(sd * a.amax);
        scale = ldexpf(1.f, e);
        if (lane == 0) a.Pn_inv[row] = ldexpf(1.f, -e);
    }
    for (int m0 = lane * V; m0 < a.Mp; m0 += 32 * V) {
        T acc[V];
#pragma unroll
        for (int e = 0; e < V; ++e) acc[e] = T(0);
        for (int k0 = 0; k0 < Knew; k0 += 8) {
            T av[8][V];
#pragma unroll
            for (int u = 0; u < 8; ++u)
                if (k0 + u < Knew)
                    *reinterpret_cast<int4*>(av[u]) = __ldg(reinterpret_cast<const int4*>(a.At + (size_t)s_gam[k0 + u] * a.Mp + m0));
#pragma unroll
            for (int u = 0; u < 8; ++u)
                if (k0 + u < Knew) {
                    const T dk = s_d[k0 + u];
#pragma unroll
                    for (int e = 0; e < V; ++e) acc[e] += av[u][e] * dk;
                }
        }
        store_rhs<T, V>(a.split, pn, pl, m0, acc, scale);
    }
}

// shared memory of one warp of the algebra kernel (bytes, multiple of 16)
template <typename T>
inline size_t homotopy_algebra_slice(int kb, int ks) {
    const size_t KC = (size_t)((kb + 3) & ~1);
    const size_t g = (size_t)(kb + 1) * ks + 1;
    return (g * sizeof(T) + KC * (4 * sizeof(T) + sizeof(uint32_t)) + 15) & ~(size_t)15;
}

}  // namespace ssb
