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

// (value, index) min across the warp; bit 31 of the index marks a LEAVE candidate and does not take part in the order
// (an atom is either active or not, so two candidates never share the other 31 bits)
template <typename T>
__device__ __forceinline__ void warp_argmin_flagged(T& v, uint32_t& i) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        T ov = __shfl_xor_sync(0xffffffffu, v, o);
        uint32_t oi = __shfl_xor_sync(0xffffffffu, i, o);
        if (cand_less(ov, oi & 0x7fffffffu, v, i & 0x7fffffffu)) { v = ov; i = oi; }
    }
}

// the same result from three warp-wide integer reductions (REDUX) instead of five rounds of two shuffles and a compare: a float
// candidate is strictly positive (or the FLT_MAX sentinel), so its bits order like its value; ties go to the lowest index, and the
// one lane that holds the winner hands out its flag.  float64 keeps the shuffle form.
__device__ __forceinline__ void warp_argmin_flagged_fast(float& v, uint32_t& i) {
    if (SSB_SHUFFLE_REDUCE) { warp_argmin_flagged(v, i); return; }
    const uint32_t kv = __float_as_uint(v);
    const uint32_t mv = __reduce_min_sync(0xffffffffu, kv);
    const uint32_t mi = __reduce_min_sync(0xffffffffu, kv == mv ? (i & 0x7fffffffu) : 0xffffffffu);
    const bool mine = kv == mv && (i & 0x7fffffffu) == mi;
    i = __reduce_max_sync(0xffffffffu, mine ? i : 0u);
    v = __uint_as_float(mv);
}
__device__ __forceinline__ void warp_argmin_flagged_fast(double& v, uint32_t& i) { warp_argmin_flagged(v, i); }

// The kernel is bound by the dependent round trips of one CTA times the CTAs an SM holds, so everything is fetched in
// two: position -> signal, then c, q, the membership bits of this thread's atoms (a per-signal bitmask kept in HBM by
// the algebra kernel), K, lambda and -- speculatively, for every slot below the iteration's bound kb -- the active set
// and its direction; x at the active atoms follows while c and q are still arriving.  No shared-memory staging: slot k
// lives in thread k % THREADS.  Three block barriers: two for the argmin, one for |c|_inf.
//   GRAMQ  q is not read from the contraction's output but formed here as q = G[:, active] d (G = A^T A, L2-resident): while
//          the active set is small that is K rows of G per signal instead of a full 2 M N contraction, a gather of K atoms
//          and 16 KB of Q traffic -- the host uses it for the first iterations (K <= it < gram_iters)
//   KR     active-set slots per thread: 1 while the iteration's bound kb fits the CTA (every BASELINE config but the long runs),
//          else 3.  With 3 the slots past the first are dead weight that the 64-register cap turns into local-memory spills
//          right behind the dependent gam -> x[gam] loads (ncu r2: STL at 24 % of the stall samples).
template <typename T, int THREADS, int NCH, bool GRAMQ, bool EXACT, int KR = 3>
__global__ void __launch_bounds__(THREADS, SSB_SCAN_MINB * 128 / THREADS) homotopy_scan_kernel(HomArgs<T> a, int kb)
{
    constexpr int GRAM_KMAX = 32;
    __shared__ uint32_t s_gq[GRAMQ ? GRAM_KMAX : 1];
    __shared__ T s_dq[GRAMQ ? GRAM_KMAX : 1], s_xq[GRAMQ ? GRAM_KMAX : 1];
    constexpr int V = 16 / sizeof(T);
    constexpr int NW = THREADS / 32;
    __shared__ T red_v[NW], red_m[NW];
    __shared__ uint32_t red_i[NW];
    __shared__ T sh_gamma;
    __shared__ uint32_t sh_idx;
    __shared__ int sh_join;

    const int pos = blockIdx.x;
    if (a.count_zero && pos == 0 && threadIdx.x == 0) *a.count_zero = 0;   // before the early exit: an empty iteration clears it too
    if (a.count_in && pos >= *a.count_in) return;   // grid sized from a stale (larger) count
    const int s = a.act ? a.act[pos] : pos;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    T* c = a.C + (size_t)s * a.ldc;
    const T* q = a.Q + (size_t)pos * a.ldc;
    // exact mode: the search reads this iteration's fresh c' -- the contraction's second row block, or, on a Gram iteration,
    // c0 - G_G x_G formed below from the same rows of G that give q
    const T* csrc = EXACT ? (GRAMQ ? a.C0 + (size_t)s * a.ldc : a.Qc + (size_t)pos * a.ldc) : c;
    T* x = a.X + (size_t)s * a.ldx;
    const uint32_t* maskg = a.mask + (size_t)s * (a.ldc >> 5);
    const T TMAX = num_traits<T>::max();

    T cv[NCH > 0 ? NCH : 1][V], qv[NCH > 0 ? NCH : 1][V];
    uint32_t mwr[NCH > 0 ? NCH : 1];
    if constexpr (NCH > 0) {
#pragma unroll
        for (int ch = 0; ch < NCH; ++ch) {
            const int i0 = (ch * THREADS + tid) * V;
            if (i0 < a.ldc) {
                // (evict-first loads were measured and lose 8 %: a third of q is still in L2 from the contraction)
                // (walking the positions last to first, to meet the rows the contraction wrote last: no change)
                *reinterpret_cast<int4*>(cv[ch]) = *reinterpret_cast<const int4*>(csrc + i0);
                if constexpr (!GRAMQ) *reinterpret_cast<int4*>(qv[ch]) = *reinterpret_cast<const int4*>(q + i0);
                mwr[ch] = maskg[i0 >> 5];
            }
        }
    }
    const int K = a.Kcur[s];
    T lam = a.lam[s];
    // K <= KR * THREADS (checked by the host)
    uint32_t gk[KR];
    T dk[KR], xk[KR];
#pragma unroll
    for (int u = 0; u < KR; ++u) {
        const int k = tid + u * THREADS;
        if (k < kb) { gk[u] = a.gam[(size_t)s * a.Kmax + k]; dk[u] = a.dG[(size_t)s * a.Kmax + k]; }
    }
#pragma unroll
    for (int u = 0; u < KR; ++u) {
        const int k = tid + u * THREADS;
        if (k < K) xk[u] = x[gk[u]];
    }
    if constexpr (GRAMQ && NCH > 0) {
        if (tid < kb) { s_gq[tid] = gk[0]; s_dq[tid] = dk[0]; if (tid < K) s_xq[tid] = xk[0]; }   // kb <= GRAM_KMAX <= THREADS (checked by the host)
        __syncthreads();
#pragma unroll
        for (int ch = 0; ch < NCH; ++ch) {
            const int i0 = (ch * THREADS + tid) * V;
#pragma unroll
            for (int e = 0; e < V; ++e) qv[ch][e] = T(0);
            // (measured, C2, 8.4 ms of Gram scans per step: fetching 2 / 4 rows x NCH chunks before adding their products
            // 9.0 / 27.8 ms (spills), 6 or 10 resident CTAs per SM instead of 8: 9.6 / 8.9 ms -- this form is the optimum)
            if (i0 < a.ldc)
                for (int k = 0; k < K; ++k) {
                    T gv[V];
                    *reinterpret_cast<int4*>(gv) = __ldg(reinterpret_cast<const int4*>(a.G + (size_t)s_gq[k] * a.ldg + i0));
                    const T d = s_dq[k];
#pragma unroll
                    for (int e = 0; e < V; ++e) qv[ch][e] += gv[e] * d;
                    if constexpr (EXACT) {   // c' = A^T y - G_G x_G, re-formed every iteration
                        const T xg = s_xq[k];
#pragma unroll
                        for (int e = 0; e < V; ++e) cv[ch][e] -= gv[e] * xg;
                    }
                }
        }
    }
    if constexpr (EXACT) {   // lambda = |c'|_inf over the vector the search reads (:270 on the re-formed c)
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
        if (lane == 0) red_m[warp] = mx0;
        __syncthreads();
        lam = red_m[0];
#pragma unroll
        for (int w = 1; w < NW; ++w) lam = max(lam, red_m[w]);
    }

    // ---- min-gamma search (find_max_gamma part 2, :122-163): smallest strictly positive step, ties -> lowest index.
    // float32 candidates are compared as integer keys bits(t) - 1 (see homotopy_step_kernel).
    T best = TMAX;
    uint32_t bidx = 0xffffffffu;
    uint32_t bestkey = __float_as_uint(FLT_MAX) - 1u;
    auto scan = [&](const T (&cc)[V], const T (&qq)[V], int i0, uint32_t mword) {
        const uint32_t mw = mword >> (i0 & 31);
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
            if (i0 < a.N) scan(cv[ch], qv[ch], i0, mwr[ch]);
        }
    } else {
        for (int i0 = tid * V; i0 < a.N; i0 += THREADS * V) {
            T cc[V], qq[V];
            *reinterpret_cast<int4*>(cc) = *reinterpret_cast<const int4*>(csrc + i0);
            *reinterpret_cast<int4*>(qq) = *reinterpret_cast<const int4*>(q + i0);
            scan(cc, qq, i0, maskg[i0 >> 5]);
        }
    }
    if constexpr (sizeof(T) == 4) best = bidx == 0xffffffffu ? TMAX : __uint_as_float(bestkey + 1u);
#pragma unroll
    for (int u = 0; u < KR; ++u) {   // leave candidates, -x_i / d_i (:129-137)
        const int k = tid + u * THREADS;
        if (k < K) {
            const T t = -xk[u] / dk[u];
            if (t > T(0) && t < TMAX && cand_less(t, gk[u], best, bidx & 0x7fffffffu)) { best = t; bidx = gk[u] | 0x80000000u; }
        }
    }
    warp_argmin_flagged_fast(best, bidx);
    if (lane == 0) { red_v[warp] = best; red_i[warp] = bidx; }
    __syncthreads();
    if (warp == 0) {
        best = lane < NW ? red_v[lane] : TMAX;
        bidx = lane < NW ? red_i[lane] : 0xffffffffu;
        warp_argmin_flagged_fast(best, bidx);
        if (lane == 0) {
            // no candidate: (max, 0) -- the caller toggles column 0 (:122-123), whichever side it is on
            sh_join = bidx != 0xffffffffu && !(bidx >> 31);
            if (bidx == 0xffffffffu) bidx = (maskg[0] & 1u) ? 0x80000000u : 0u;
            sh_gamma = best;
            sh_idx = bidx;
        }
    }
    __syncthreads();
    const bool present = (sh_idx >> 31) != 0;
    const uint32_t idx = sh_idx & 0x7fffffffu;
    if constexpr (sizeof(T) == 4) {
        // the winning join step again, in IEEE arithmetic (the ordering used the hardware reciprocal)
        if (sh_join && !SSB_IEEE_ORDER) {
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
            __syncthreads();   // sh_join is block-uniform
        }
    }
    const T gamma = sh_gamma;
    const int Knew = present ? K - 1 : K + 1;
    if (Knew == 0) {   // :248-249: the last atom left; x, c and lambda stay as they were
        if (tid == 0) { a.iters[s] = a.it; a.errs[s] = (double)lam; a.Kcur[s] = 0; }
        return;
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
        if (k < K) x[gk[u]] = add_rn(xk[u], mul_rn(gamma, dk[u]));
    }

    // ---- c = A^T (y - A x)  ==  c - gamma * q ;  lambda = |c|_inf (:255, :270); exact mode: one step from the fresh c'
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

// One warp per continuing signal (row of act_next).  K <= kb before the toggle; the inverse is staged in the warp's
// shared-memory slice with the odd row stride ks (lane = row: conflict-free), worked on, and written back.
template <typename T, int W, bool EXACT>
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
        if (K >= 24) {                                                                      // inv += d u2 u2^T
            // row by row, lanes along the row, this lane's u2 entries in registers: a load, a multiply-add and a store per element
            // (the flat loop below pays an index split and two more loads, but fills the lanes better while K is small)
            T u2k[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) u2k[j] = lane + 32 * j < K ? s_u2[lane + 32 * j] : T(0);
            for (int r = 0; r < K; ++r) {
                const T t = dd * s_u2[r];
                T* g = s_G + (size_t)r * ks;
#pragma unroll
                for (int j = 0; j < 4; ++j) if (lane + 32 * j < K) g[lane + 32 * j] += t * u2k[j];
                for (int k = lane + 128; k < K; k += 32) g[k] += t * s_u2[k];
            }
        } else {
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
    if (lane == 0) {
        a.Kcur[s] = Knew;
        a.mask[(size_t)s * (a.ldc >> 5) + (idx >> 5)] ^= 1u << (idx & 31);   // membership bits read by the next scan
    }
    __syncwarp();

    // ---- p = A_G d for the next contraction (:112-116), k ascending per element as in write_rhs; in exact mode also the
    // residual r = y - A_G x_G (:94-96 over the active atoms) from the same atom rows
    if (a.skip_rhs) return;   // the next iteration forms q (and c') from the Gram matrix: no right-hand side needed
    constexpr bool res = EXACT;
    T* s_x = s_u2;            // free since the inverse update
    if (res) {
        const T* x = a.X + (size_t)s * a.ldx;
        for (int k = lane; k < Knew; k += 32) s_x[k] = x[s_gam[k]];
        __syncwarp();
    }
    void *pn, *pl, *rn = nullptr, *rl = nullptr;
    rhs_rows(a, row, pn, pl);
    if (res) res_rows(a, row, rn, rl);
    float scale = 1.f, rscale = 1.f;
    if (a.split == 2) {   // |p_m| <= amax * sum |d_k| ; |r_m| <= max |y| + amax * sum |x_k|
        float sd = 0.f, sx = 0.f;
        for (int k = lane; k < Knew; k += 32) { sd += fabsf((float)s_d[k]); if (res) sx += fabsf((float)s_x[k]); }
        sd = warp_sum(sd);
        const int e = f16_scale_exp(sd * a.amax);
        scale = ldexpf(1.f, e);
        if (lane == 0) a.Pn_inv[row] = ldexpf(1.f, -e);
        if (res) {
            sx = warp_sum(sx);
            const int er = f16_scale_exp(a.ymax[s] + sx * a.amax);
            rscale = ldexpf(1.f, er);
            if (lane == 0) a.Rn_inv[row] = ldexpf(1.f, -er);
        }
    }
    const T* y = res ? a.Y + (size_t)s * a.ldy : nullptr;
    // two 16-byte chunks of p per pass, four atoms per batch: 8 independent loads in flight per lane, and the active
    // set, the direction and the row address are fetched once per atom for both chunks
    for (int mb = lane * V; mb < a.Mp; mb += 2 * 32 * V) {
        const bool two = mb + 32 * V < a.Mp;   // warp-uniform
        T acc0[V], acc1[V], rac0[V], rac1[V];
#pragma unroll
        for (int e = 0; e < V; ++e) { acc0[e] = T(0); acc1[e] = T(0); rac0[e] = T(0); rac1[e] = T(0); }
        int k = 0;
        for (; k + 4 <= Knew; k += 4) {
            T a0[4][V], a1[4][V];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const T* rp = a.At + (size_t)s_gam[k + u] * a.Mp + mb;
                *reinterpret_cast<int4*>(a0[u]) = __ldg(reinterpret_cast<const int4*>(rp));
                if (two) *reinterpret_cast<int4*>(a1[u]) = __ldg(reinterpret_cast<const int4*>(rp + 32 * V));
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const T dk = s_d[k + u];
#pragma unroll
                for (int e = 0; e < V; ++e) acc0[e] += a0[u][e] * dk;
                if (two) {
#pragma unroll
                    for (int e = 0; e < V; ++e) acc1[e] += a1[u][e] * dk;
                }
                if (res) {
                    const T xk = s_x[k + u];
#pragma unroll
                    for (int e = 0; e < V; ++e) rac0[e] += a0[u][e] * xk;
                    if (two) {
#pragma unroll
                        for (int e = 0; e < V; ++e) rac1[e] += a1[u][e] * xk;
                    }
                }
            }
        }
        for (; k < Knew; ++k) {
            T a0[V], a1[V];
            const T* rp = a.At + (size_t)s_gam[k] * a.Mp + mb;
            *reinterpret_cast<int4*>(a0) = __ldg(reinterpret_cast<const int4*>(rp));
            if (two) *reinterpret_cast<int4*>(a1) = __ldg(reinterpret_cast<const int4*>(rp + 32 * V));
            const T dk = s_d[k];
#pragma unroll
            for (int e = 0; e < V; ++e) acc0[e] += a0[e] * dk;
            if (two) {
#pragma unroll
                for (int e = 0; e < V; ++e) acc1[e] += a1[e] * dk;
            }
            if (res) {
                const T xk = s_x[k];
#pragma unroll
                for (int e = 0; e < V; ++e) rac0[e] += a0[e] * xk;
                if (two) {
#pragma unroll
                    for (int e = 0; e < V; ++e) rac1[e] += a1[e] * xk;
                }
            }
        }
        store_rhs<T, V>(a.split, pn, pl, mb, acc0, scale);
        if (two) store_rhs<T, V>(a.split, pn, pl, mb + 32 * V, acc1, scale);
        if (res) {
#pragma unroll
            for (int e = 0; e < V; ++e) rac0[e] = (mb + e < a.M ? y[mb + e] : T(0)) - rac0[e];
            store_rhs<T, V>(a.split, rn, rl, mb, rac0, rscale);
            if (two) {
#pragma unroll
                for (int e = 0; e < V; ++e) rac1[e] = (mb + 32 * V + e < a.M ? y[mb + 32 * V + e] : T(0)) - rac1[e];
                store_rhs<T, V>(a.split, rn, rl, mb + 32 * V, rac1, rscale);
            }
        }
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
