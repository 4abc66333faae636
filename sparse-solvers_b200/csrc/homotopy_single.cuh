// Single-signal latency path (BASELINE config 1): the WHOLE homotopy solve of one signal in one cooperative kernel.
// The lockstep path spends ~35 us per iteration of one signal on two launches and a host round trip; here an iteration
// is three grid barriers.  Every CTA owns a slice of atoms for the whole solve and keeps its slice of c on chip:
//   phase 1  all CTAs   q = A^T p over the CTA's atoms (warp per atom, the arithmetic of contract_gemv_kernel), join
//                       candidates of the slice; CTA 0 adds the leave candidates (homotopy-cpu.cpp:112-163)
//   phase 2  all CTAs   (gamma, idx) = min over the CTA partials (every CTA redundantly); c -= gamma q on the slice,
//                       partial |c|_inf, slice of c published for the signs (:252-255, :270)
//   phase 3  CTA 0      x += gamma d, toggle, online inverse update (in shared memory for the whole solve), new d,
//                       p = A_G d (:166-183, :259-266, online_inverse.h:183-293); all CTAs: stop test (:272)
// Same formulas, summation orders and quirks as homotopy_step_kernel; the dictionary (1-4 MB) streams from L2.
// Exact-residual mode (SingleArgs::exact, see homotopy_kernels.cuh): CTA 0 also re-forms r = y - A_G x_G in phase 3, phase 1
// takes c' = A^T r and q = A^T p from ONE pass over the CTA's atoms, and a fourth grid barrier publishes lambda = |c'|_inf
// before the search.  All divisions here are IEEE (a CTA looks at ~N/148 candidates per iteration).
#pragma once

#include <cooperative_groups.h>

#include "homotopy_split.cuh"

namespace ssb {

// phase timer of the loop as CTA 0 / thread 0 sees it (variant build -DSSB_SINGLE_PROF=1, tools/single_prof.py): clock64 deltas
// summed per phase into g_single_prof, slot 15 counts iterations
#ifndef SSB_SINGLE_PROF
#define SSB_SINGLE_PROF 0
#endif
#if SSB_SINGLE_PROF
__device__ unsigned long long g_single_prof[32];
// (sums are kept in shared memory and flushed when the solve ends: a read-modify-write of global memory per probe stalls warp 0 for
// an L2 round trip, which is what the first version of this timer mostly measured)
#define SSB_PROF(i) do { if (threadIdx.x == 0 && cta == 0) { const long long t_ = clock64(); s_prof[i] += (unsigned long long)(t_ - prof_t); prof_t = t_; } } while (0)
#else
#define SSB_PROF(i) do {} while (0)
#endif

template <typename T>
struct SingleArgs {
    int N, Npad, Mp, Kmax, ks;
    const T* At; const T* G; int ldg;
    const T* y;           // [Mp] zero padded signal; may be PINNED HOST memory (read once, at the start)
    T* ywork;             // [Mp] device copy of y for the exact mode's r = y - A_G x_G when y is host memory (else nullptr)
    T* x;                 // [N] output (zeroed by the kernel)
    T* x_out;             // [N] pinned host copy of the finished x, written by the kernel (else nullptr): a host-pointer call is then
                          //     one launch and one stream synchronise -- no staged copies
    T* P;                 // [Mp] scratch: current right-hand side
    T* R;                 // [Mp] scratch: current residual (exact mode)
    int exact;
    T* Cg;                // [Npad] scratch: c, published every iteration
    T* part_t; uint32_t* part_i; T* part_max;   // [grid] partials
    uint32_t* iters; double* errs;
    T tol; uint32_t max_iter;
    int cnt;              // atoms per CTA
    int dict_smem;        // 1: every CTA keeps its slice of the dictionary (cnt x Mp) in shared memory for the whole launch
    int stage;            // 1: SINGLE_STAGE_BYTES of shared memory are there for the gather of p = A_G d
    // Gram iterations, as in the lockstep loop (homotopy_impl.cuh: solve_chunk): on iterations it < gram_iters (K <= it) every CTA forms
    // q = G[:, active] d -- and, in exact mode, c' = A^T y - G[:, active] x -- over its slice from K rows of G = A^T A; no p, no r, no
    // pass over the dictionary.  0 for coherent dictionaries; not used in resume mode.
    uint32_t gram_iters;
    uint32_t* pub_gam; T* pub_d; T* pub_x;   // grid form: CTA 0 publishes the active set, its direction and x on it here ([Kmax + 2] each)
    // ---- resume mode: the finisher of the lockstep loop.  The last few still-running signals of a chunk (stragglers that go on
    // to max_iterations long after the rest has converged) are taken from their lockstep state to the end here, one after the
    // other: an iteration then costs grid barriers instead of two thin launches and a host round trip.
    int resume;            // 0: one signal from scratch (y, x); 1: signals act[0 .. min(*count, max_sigs)) from their state
    const int* act;        // position -> signal (nullptr: identity)
    const int* count;
    int max_sigs;
    uint32_t it0;          // iterations the lockstep loop has completed
    const T* Y; size_t ldy; int M;       // the signals, caller layout (exact mode: r = y - A_G x_G)
    T* X; size_t ldx;      // coefficient rows
    const T* C; int ldc;   // correlations by signal (incremental mode)
    const uint32_t* gamv; const T* dGv; const T* Ginvv; const T* lamv; const int* Kcurv; int Kstate;   // lockstep state, Kstate slots per signal
    uint32_t* itersv; double* errsv;     // reports, by signal
};

constexpr int SINGLE_STAGE_BYTES = 16 * 1024;   // cluster form: the atom rows of p = A_G d pass through shared memory, 4 x 16 bytes per thread at a time
template <typename T>
inline size_t homotopy_single_smem(int Mp, int cnt, int Kmax, int ks, bool dict_slice = false, bool stage = false) {
    const size_t KC = (size_t)((Kmax + 2) & ~1);
    return (2 * (size_t)Mp + 4 * (size_t)cnt + (size_t)(Kmax + 1) * ks + 6 * KC + 8) * sizeof(T) + ((size_t)cnt + KC) * sizeof(uint32_t) + 64 +
           (stage ? SINGLE_STAGE_BYTES + 16 : 0) + (dict_slice ? (size_t)cnt * Mp * sizeof(T) + 16 : 0);
}

constexpr int SINGLE_CLUSTER_MAX = 16;   // CTAs of a cluster (16 needs the non-portable opt-in, 8 is portable)

// CLUSTER = false: a cooperative grid, one CTA per SM, barriers are grid.sync() and the exchange goes through global scratch.
// CLUSTER = true:  one thread-block cluster per signal (grid = clusters x cluster size, an ordinary launch); barriers are
//                  barrier.cluster (~0.2 us against ~2 us), the partials are PUSHED into every CTA's shared memory, the right-hand
//                  sides are read from CTA 0's shared memory and c at the active atoms from its owner's (distributed shared memory):
//                  no global scratch, no fence.  Same arithmetic, same order: results are the same bits.
template <typename T, int THREADS, bool CLUSTER>
__global__ void __launch_bounds__(THREADS, 1) homotopy_single_kernel(SingleArgs<T> a)
{
    namespace cg = cooperative_groups;
    constexpr int V = 16 / sizeof(T);
    constexpr int NW = THREADS / 32;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ T red_v[NW];
    __shared__ uint32_t red_i[NW];
    __shared__ T xp_t[SINGLE_CLUSTER_MAX], xp_max[SINGLE_CLUSTER_MAX], xp_lam[SINGLE_CLUSTER_MAX];   // cluster form: the partials of every CTA, pushed by their owners
    __shared__ uint32_t xp_i[SINGLE_CLUSTER_MAX + 1];

    const int KC = (a.Kmax + 2) & ~1;
    T* s_p = reinterpret_cast<T*>(smem_raw);          // [Mp]
    T* s_r = s_p + a.Mp;                              // [Mp]   residual (exact mode)
    T* s_c = s_r + a.Mp;                              // [cnt]
    T* s_q = s_c + a.cnt;                             // [cnt]
    T* s_c2 = s_q + a.cnt;                            // [cnt]  cluster form: the updated c goes here, then the two swap
    T* s_c0 = s_c2 + a.cnt;                           // [cnt]  A^T y over the slice (exact mode, Gram iterations)
    T* s_G = s_c0 + a.cnt;                            // [(Kmax+1) * ks]   CTA 0 only from here on (the others use s_gam / s_d / s_x as copies)
    T* s_d = s_G + (size_t)(a.Kmax + 1) * a.ks;
    T* s_u1 = s_d + KC;
    T* s_u2 = s_u1 + KC;
    T* s_cg = s_u2 + KC;
    T* s_x = s_cg + KC;                               // x at the active atoms
    uint32_t* s_act = reinterpret_cast<uint32_t*>(s_x + 2 * KC);   // [cnt] membership of the slice's atoms
    uint32_t* s_gam = s_act + a.cnt;                  // [KC]

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int cta = blockIdx.x, G = gridDim.x, cid = 0, ncl = 1;
    if constexpr (CLUSTER) {
        cg::cluster_group cl = cg::this_cluster();
        cta = (int)cl.block_rank(); G = (int)cl.num_blocks(); cid = blockIdx.x / G; ncl = gridDim.x / G;
    }
    // barrier over the CTAs that share a signal; what was written before it is visible after it
    auto sync_all = [&]() {
        if constexpr (CLUSTER) cg::this_cluster().sync();
        else { __threadfence(); cg::this_grid().sync(); }
    };
    // partials: (value, index) and |c|_inf per CTA; slot G of the index array carries the membership of atom 0
    auto put_pair = [&](int slot, T v, uint32_t i) {          // one thread
        if constexpr (CLUSTER) {
            cg::cluster_group cl = cg::this_cluster();
            for (int g = 0; g < G; ++g) { cl.map_shared_rank(xp_t, g)[slot] = v; cl.map_shared_rank(xp_i, g)[slot] = i; }
        } else { a.part_t[slot] = v; a.part_i[slot] = i; }
    };
    auto put_index = [&](int slot, uint32_t i) {
        if constexpr (CLUSTER) { cg::cluster_group cl = cg::this_cluster(); for (int g = 0; g < G; ++g) cl.map_shared_rank(xp_i, g)[slot] = i; }
        else a.part_i[slot] = i;
    };
    auto put_max = [&](int slot, T v) {
        if constexpr (CLUSTER) { cg::cluster_group cl = cg::this_cluster(); for (int g = 0; g < G; ++g) cl.map_shared_rank(xp_max, g)[slot] = v; }
        else a.part_max[slot] = v;
    };
    auto get_t = [&](int g) -> T { if constexpr (CLUSTER) return xp_t[g]; else return __ldcg(a.part_t + g); };
    auto get_i = [&](int g) -> uint32_t { if constexpr (CLUSTER) return xp_i[g]; else return __ldcg(a.part_i + g); };
    auto get_max = [&](int g) -> T { if constexpr (CLUSTER) return xp_max[g]; else return __ldcg(a.part_max + g); };
    // the same inside the loop, where every thread holds the value: cluster form, thread g pushes to CTA g (G stores in flight at once)
    auto push_pair = [&](T v, uint32_t i) {
        if constexpr (CLUSTER) {
            if (tid < G) { cg::cluster_group cl = cg::this_cluster(); cl.map_shared_rank(xp_t, tid)[cta] = v; cl.map_shared_rank(xp_i, tid)[cta] = i; }
        } else if (tid == 0) { a.part_t[cta] = v; a.part_i[cta] = i; }
    };
    auto push_index = [&](int slot, uint32_t i) {
        if constexpr (CLUSTER) { if (tid < G) cg::this_cluster().map_shared_rank(xp_i, tid)[slot] = i; }
        else if (tid == 0) a.part_i[slot] = i;
    };
    auto push_max = [&](T v, bool lam_slot) {     // lam_slot: the exact mode's |c'|_inf exchange has its own array (one barrier less between uses)
        if constexpr (CLUSTER) { if (tid < G) cg::this_cluster().map_shared_rank(lam_slot ? xp_lam : xp_max, tid)[cta] = v; }
        else if (tid == 0) a.part_max[cta] = v;
    };
    // reductions over the CTA partials after a barrier, result in every thread.  Cluster form: <= 16 values, every warp reduces them
    // itself -- no block barrier
    auto reduce_max = [&](bool lam_slot) -> T {
        if constexpr (CLUSTER) {
            const T* src = lam_slot ? xp_lam : xp_max;
            return warp_max(lane < G ? src[lane] : T(0));
        } else {
            T l = T(0);
            for (int g = tid; g < G; g += THREADS) l = max(l, __ldcg(a.part_max + g));
            l = warp_max(l);
            __syncthreads();
            if (lane == 0) red_v[warp] = l;
            __syncthreads();
            l = red_v[0];
            for (int w = 1; w < NW; ++w) l = max(l, red_v[w]);
            __syncthreads();
            return l;
        }
    };
    // where CTA 0 leaves the right-hand sides of the next iteration: its own s_p / s_r (read by the others through DSMEM) or global
    T* const outP = CLUSTER ? s_p : a.P;
    T* const outR = CLUSTER ? s_r : a.R;
    const int a0 = cta * a.cnt, a1 = min(a.Npad, a0 + a.cnt);
    const T TMAX = num_traits<T>::max();

    // the CTA's slice of the dictionary: shared memory when it fits (cluster form: 1/16 of the atoms), else streamed from L2
    const T* dict = a.At + (size_t)a0 * a.Mp;                      // row of atom n: dict + (n - a0) * Mp
    T* const s_stage = a.stage ? reinterpret_cast<T*>((reinterpret_cast<uintptr_t>(s_gam + KC) + 15) & ~(uintptr_t)15) : nullptr;
    if (a.dict_smem) {
        T* s_A = reinterpret_cast<T*>((reinterpret_cast<uintptr_t>(s_gam + KC) + 15 + (a.stage ? SINGLE_STAGE_BYTES + 16 : 0)) & ~(uintptr_t)15);
        const int n16 = max(0, a1 - a0) * (a.Mp / V);
        const int4* src = reinterpret_cast<const int4*>(dict);
        for (int i = tid; i < n16; i += THREADS) reinterpret_cast<int4*>(s_A)[i] = __ldg(src + i);
        __syncthreads();
        dict = s_A;
    }
    // q over the slice: one warp per atom, lanes stride the row in 16-byte pieces, warp_sum (contract_gemv_kernel: the same sum
    // per atom in the same order); a warp keeps the rows of CB atoms in flight (a slice is 8 atoms per CTA as a grid, 64 as a cluster)
    constexpr int CB = 4;
    auto correlate = [&](const T* rhs) {
        for (int nb = a0 + warp; nb < a1; nb += NW * CB) {
            T acc[CB];
#pragma unroll
            for (int u = 0; u < CB; ++u) acc[u] = T(0);
            for (int m = lane * V; m < a.Mp; m += 32 * V) {
                T av[CB][V];
#pragma unroll
                for (int u = 0; u < CB; ++u) {
                    const int n = min(nb + u * NW, a1 - 1);
                    *reinterpret_cast<int4*>(av[u]) = *reinterpret_cast<const int4*>(dict + (size_t)(n - a0) * a.Mp + m);
                }
#pragma unroll
                for (int u = 0; u < CB; ++u)
#pragma unroll
                    for (int v = 0; v < V; ++v) acc[u] += av[u][v] * rhs[m + v];
            }
#pragma unroll
            for (int u = 0; u < CB; ++u) {
                const T t = warp_sum(acc[u]);
                if (lane == 0 && nb + u * NW < a1) s_q[nb + u * NW - a0] = t;
            }
        }
    };
    // exact mode: c' = A^T r and q = A^T p over the slice from one pass over its atoms
    auto correlate2 = [&]() {
        for (int nb = a0 + warp; nb < a1; nb += NW * CB) {
            T accq[CB], accc[CB];
#pragma unroll
            for (int u = 0; u < CB; ++u) { accq[u] = T(0); accc[u] = T(0); }
            for (int m = lane * V; m < a.Mp; m += 32 * V) {
                T av[CB][V];
#pragma unroll
                for (int u = 0; u < CB; ++u) {
                    const int n = min(nb + u * NW, a1 - 1);
                    *reinterpret_cast<int4*>(av[u]) = *reinterpret_cast<const int4*>(dict + (size_t)(n - a0) * a.Mp + m);
                }
#pragma unroll
                for (int u = 0; u < CB; ++u)
#pragma unroll
                    for (int v = 0; v < V; ++v) { accq[u] += av[u][v] * s_p[m + v]; accc[u] += av[u][v] * s_r[m + v]; }
            }
#pragma unroll
            for (int u = 0; u < CB; ++u) {
                const T tq = warp_sum(accq[u]), tc = warp_sum(accc[u]);
                if (lane == 0 && nb + u * NW < a1) { s_q[nb + u * NW - a0] = tq; s_c[nb + u * NW - a0] = tc; }
            }
        }
    };
    // block-wide lexicographic min of (value, index); bit 31 of the index may carry the leave flag
    auto block_argmin = [&](T& v, uint32_t& i) {
        warp_argmin_flagged(v, i);
        if (lane == 0) { red_v[warp] = v; red_i[warp] = i; }
        __syncthreads();
        v = lane < NW ? red_v[lane] : TMAX;
        i = lane < NW ? red_i[lane] : 0xffffffffu;
        warp_argmin_flagged(v, i);
        __syncthreads();
    };
    // (gamma, idx) over the CTA partials after the barrier, in every thread
    auto reduce_pairs = [&](T& v, uint32_t& i) {
        if constexpr (CLUSTER) {
            v = lane < G ? xp_t[lane] : TMAX;
            i = lane < G ? xp_i[lane] : 0xffffffffu;
            warp_argmin_flagged_fast(v, i);
        } else {
            v = TMAX; i = 0xffffffffu;
            for (int g = tid; g < G; g += THREADS) {
                const T pv = __ldcg(a.part_t + g); const uint32_t pi = __ldcg(a.part_i + g);
                if (cand_less(pv, pi & 0x7fffffffu, v, i & 0x7fffffffu)) { v = pv; i = pi; }
            }
            block_argmin(v, i);
        }
    };

    const int nsig = a.resume ? min(*a.count, a.max_sigs) : 1;
    for (int si = cid; si < nsig; si += ncl) {
    const int s = a.resume ? (a.act ? a.act[si] : si) : 0;
    T* xrow = a.resume ? a.X + (size_t)s * a.ldx : a.x;
    const T* ysrc = a.resume ? a.Y + (size_t)s * a.ldy : a.y;
    const T* yrow = (!a.resume && a.ywork) ? a.ywork : ysrc;
    const int ylen = a.resume ? a.M : a.Mp;          // a.y is zero padded to Mp, a caller's row is not
    uint32_t* it_out = a.resume ? a.itersv + s : a.iters;
    double* err_out = a.resume ? a.errsv + s : a.errs;
    uint32_t it = 0;
    T lam = T(0);
    int K = 1;
    // p = A_G d (:112-116), k ascending per element as write_rhs; exact mode: r = y - A_G x_G (:94-96); CTA 0 only
    auto publish_rhs = [&](int Knew) {
                const int tpr = a.Mp / V;                      // threads that cover one atom row in 16-byte pieces
                if (s_stage != nullptr && tpr <= THREADS && THREADS % tpr == 0) {
                    // every thread fetches up to four 16-byte pieces at once into the stage (all rows of a pass in flight: one L2
                    // round trip), then the first tpr threads add them up in slot order -- the sums of the loop below, bit for bit
                    const int rpp = THREADS / tpr, lr = tid / tpr, mq = (tid - lr * tpr) * V, SR = 4 * rpp;
                    T acc[V], accr[V];
#pragma unroll
                    for (int e = 0; e < V; ++e) { acc[e] = T(0); accr[e] = T(0); }
                    for (int k0 = 0; k0 < Knew; k0 += SR) {
                        const int nr = min(SR, Knew - k0);
                        int4 piece[4];
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const int r = lr + u * rpp;
                            if (r < nr) piece[u] = __ldg(reinterpret_cast<const int4*>(a.At + (size_t)s_gam[k0 + r] * a.Mp + mq));
                        }
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const int r = lr + u * rpp;
                            if (r < nr) *reinterpret_cast<int4*>(s_stage + (size_t)r * a.Mp + mq) = piece[u];
                        }
                        __syncthreads();
                        if (tid < tpr)
                            for (int r = 0; r < nr; ++r) {
                                T av[V];
                                *reinterpret_cast<int4*>(av) = *reinterpret_cast<const int4*>(s_stage + (size_t)r * a.Mp + tid * V);
                                const T dk = s_d[k0 + r], xk = s_x[k0 + r];
#pragma unroll
                                for (int e = 0; e < V; ++e) { acc[e] += av[e] * dk; accr[e] += av[e] * xk; }
                            }
                        __syncthreads();
                    }
                    if (tid < tpr) {
                        const int m0 = tid * V;
                        *reinterpret_cast<int4*>(outP + m0) = *reinterpret_cast<const int4*>(acc);
                        if (a.exact) {
#pragma unroll
                            for (int e = 0; e < V; ++e) accr[e] = (m0 + e < ylen ? yrow[m0 + e] : T(0)) - accr[e];
                            *reinterpret_cast<int4*>(outR + m0) = *reinterpret_cast<const int4*>(accr);
                        }
                    }
                    return;
                }
                for (int m0 = tid * V; m0 < a.Mp; m0 += THREADS * V) {
                    T acc[V], accr[V];
#pragma unroll
                    for (int e = 0; e < V; ++e) { acc[e] = T(0); accr[e] = T(0); }
                    int k = 0;
                    for (; k + 8 <= Knew; k += 8) {   // eight rows in flight per thread: this loop is L2 latency on ONE CTA
                        T av8[8][V];
#pragma unroll
                        for (int u = 0; u < 8; ++u)
                            *reinterpret_cast<int4*>(av8[u]) = __ldg(reinterpret_cast<const int4*>(a.At + (size_t)s_gam[k + u] * a.Mp + m0));
#pragma unroll
                        for (int u = 0; u < 8; ++u) {
                            const T dk = s_d[k + u], xk = s_x[k + u];
#pragma unroll
                            for (int e = 0; e < V; ++e) { acc[e] += av8[u][e] * dk; accr[e] += av8[u][e] * xk; }
                        }
                    }
                    for (; k < Knew; ++k) {
                        T av[V];
                        *reinterpret_cast<int4*>(av) = __ldg(reinterpret_cast<const int4*>(a.At + (size_t)s_gam[k] * a.Mp + m0));
                        const T dk = s_d[k], xk = s_x[k];
#pragma unroll
                        for (int e = 0; e < V; ++e) { acc[e] += av[e] * dk; accr[e] += av[e] * xk; }
                    }
                    *reinterpret_cast<int4*>(outP + m0) = *reinterpret_cast<const int4*>(acc);
                    if (a.exact) {
#pragma unroll
                        for (int e = 0; e < V; ++e) accr[e] = (m0 + e < ylen ? yrow[m0 + e] : T(0)) - accr[e];
                        *reinterpret_cast<int4*>(outR + m0) = *reinterpret_cast<const int4*>(accr);
                    }
                }
    };
    if (!a.resume) {
        // ---- initial state (homotopy-cpu.cpp:201-229)
        for (int m = tid; m < a.Mp; m += THREADS) { const T v = ysrc[m]; s_p[m] = v; if (cta == 0 && a.ywork) a.ywork[m] = v; }
        for (int n = a0 + tid; n < a1; n += THREADS) { s_act[n - a0] = 0u; if (n < a.N) xrow[n] = T(0); }
        __syncthreads();
        correlate(s_p);
        __syncthreads();
        {
            T best = T(-1);
            uint32_t bidx = 0xffffffffu;
            for (int n = a0 + tid; n < a1; n += THREADS) {
                s_c[n - a0] = s_q[n - a0];
                s_c0[n - a0] = s_q[n - a0];
                if (n < a.N) { const T av = fabs(s_q[n - a0]); if (av > best) { best = av; bidx = (uint32_t)n; } }
            }
            warp_argmax_first(best, bidx);
            if (lane == 0) { red_v[warp] = best; red_i[warp] = bidx; }
            __syncthreads();
            if (tid == 0) {
                for (int w = 1; w < NW; ++w) if (red_v[w] > best || (red_v[w] == best && red_i[w] < bidx)) { best = red_v[w]; bidx = red_i[w]; }
                put_pair(cta, best, bidx);
                if (cta == 0) put_max(0, fabs(s_q[0]));
            }
        }
        sync_all();
        K = 1;
        {
            T best = T(-1);
            uint32_t bidx = 0xffffffffu;
            for (int g = tid; g < G; g += THREADS) {
                const T v = get_t(g); const uint32_t i = get_i(g);
                if (v > best || (v == best && i < bidx)) { best = v; bidx = i; }
            }
            warp_argmax_first(best, bidx);
            if (lane == 0) { red_v[warp] = best; red_i[warp] = bidx; }
            __syncthreads();
            best = red_v[0]; bidx = red_i[0];
            for (int w = 1; w < NW; ++w) if (red_v[w] > best || (red_v[w] == best && red_i[w] < bidx)) { best = red_v[w]; bidx = red_i[w]; }
            if (bidx == 0xffffffffu) { bidx = 0; best = get_max(0); }   // all-NaN correlations
            lam = best;
            if ((int)bidx >= a0 && (int)bidx < a1 && tid == 0) s_act[bidx - a0] = 1u;
            if (cta == 0) {
                if (tid == 0) {
                    const T nrm = sqrt(a.G[(size_t)bidx * a.ldg + bidx]);       // online_inverse.h:193-201
                    const T inv = T(1) / (nrm * nrm);
                    s_gam[0] = bidx; s_G[0] = inv; s_d[0] = sign_tol(best, a.tol) * inv; s_x[0] = T(0);   // sign of |c|: :223-224
                    if constexpr (!CLUSTER) { if (a.gram_iters > 1) { a.pub_gam[0] = bidx; a.pub_d[0] = s_d[0]; a.pub_x[0] = T(0); } }
                }
                __syncthreads();
                const T d0 = s_d[0];
                for (int m = tid; m < a.Mp; m += THREADS) {
                    outP[m] = T(0) + a.At[(size_t)s_gam[0] * a.Mp + m] * d0;
                    if (a.exact) outR[m] = yrow[m];                    // r = y - A 0
                }
            }
        }
    } else {
        // ---- take the signal over from the lockstep loop: K, lambda, its slice of c and of the membership; CTA 0 also the active
        // set, its direction, x on it and the inverse; then the right-hand side(s) of the next iteration
        K = a.Kcurv[s];
        lam = a.lamv[s];
        const uint32_t* gst = a.gamv + (size_t)s * a.Kstate;
        for (int n = a0 + tid; n < a1; n += THREADS) { s_act[n - a0] = 0u; s_c[n - a0] = a.C[(size_t)s * a.ldc + n]; }
        __syncthreads();
        for (int k = tid; k < K; k += THREADS) { const int gi = (int)gst[k]; if (gi >= a0 && gi < a1) s_act[gi - a0] = 1u; }
        if (cta == 0) {
            for (int k = tid; k < K; k += THREADS) { const uint32_t gi = gst[k]; s_gam[k] = gi; s_d[k] = a.dGv[(size_t)s * a.Kstate + k]; s_x[k] = xrow[gi]; }
            const T* gsrc = a.Ginvv + (size_t)s * a.Kstate * a.Kstate;
            for (int e = tid; e < K * K; e += THREADS) { const int r = e / K, kk = e - r * K; s_G[(size_t)r * a.ks + kk] = gsrc[(size_t)r * a.Kstate + kk]; }
            __syncthreads();
            publish_rhs(K);
        }
        it = a.it0;
    }
    sync_all();

    // ---- the loop (homotopy-cpu.cpp:236-272)
#if SSB_SINGLE_PROF
    __shared__ unsigned long long s_prof[32];
    if (threadIdx.x < 32) s_prof[threadIdx.x] = 0;
    __syncthreads();
#endif
    [[maybe_unused]] long long prof_t = clock64();
    for (;;) {
        ++it;
        SSB_PROF(14);
        const bool gram = !a.resume && it < a.gram_iters;             // K <= it < gram_iters <= 32
        const bool gram_next = !a.resume && it + 1 < a.gram_iters;    // the next iteration needs no right-hand side
        if (gram) {
            // the active set, its direction and (exact mode) x on it: from CTA 0
            if (cta != 0) {
                if constexpr (CLUSTER) {
                    cg::cluster_group cl = cg::this_cluster();
                    const uint32_t* rg = cl.map_shared_rank(s_gam, 0);
                    const T* rd = cl.map_shared_rank(s_d, 0);
                    const T* rx = cl.map_shared_rank(s_x, 0);
                    if (tid < K) { s_gam[tid] = rg[tid]; s_d[tid] = rd[tid]; s_x[tid] = rx[tid]; }
                } else {
                    if (tid < K) { s_gam[tid] = __ldcg(a.pub_gam + tid); s_d[tid] = __ldcg(a.pub_d + tid); s_x[tid] = __ldcg(a.pub_x + tid); }
                }
            }
            __syncthreads();
            SSB_PROF(0);
            // one thread per atom of the slice, slots in order (the sums of homotopy_scan_kernel's Gram form)
            for (int n = a0 + tid; n < a1; n += THREADS) {
                T qn = T(0), cn = s_c0[n - a0];
                for (int k = 0; k < K; k += 16) {     // K <= 15 on the default 16 Gram iterations: every row in flight at once
                    T gv[16];
#pragma unroll
                    for (int u = 0; u < 16; ++u) if (k + u < K) gv[u] = __ldg(a.G + (size_t)s_gam[k + u] * a.ldg + n);
                    if (a.exact) {
#pragma unroll
                        for (int u = 0; u < 16; ++u) if (k + u < K) { qn += gv[u] * s_d[k + u]; cn -= gv[u] * s_x[k + u]; }
                    } else {
#pragma unroll
                        for (int u = 0; u < 16; ++u) if (k + u < K) qn += gv[u] * s_d[k + u];
                    }
                }
                s_q[n - a0] = qn;
                if (a.exact) s_c[n - a0] = cn;
            }
            SSB_PROF(23);
            __syncthreads();
        } else {
        if constexpr (CLUSTER) {
            if (cta != 0) {
                cg::cluster_group cl = cg::this_cluster();
                const T* rp = cl.map_shared_rank(s_p, 0);
                const T* rr = cl.map_shared_rank(s_r, 0);
                for (int m = tid; m < a.Mp; m += THREADS) { s_p[m] = rp[m]; if (a.exact) s_r[m] = rr[m]; }
            }
        } else {
            for (int m = tid; m < a.Mp; m += THREADS) { s_p[m] = __ldcg(a.P + m); if (a.exact) s_r[m] = __ldcg(a.R + m); }
        }
        __syncthreads();
        SSB_PROF(0);
        if (a.exact) correlate2(); else correlate(s_p);
        __syncthreads();
        }
        SSB_PROF(1);
        if (a.exact) {   // lambda = |c'|_inf over the re-formed correlations (:270 on :255's c)
            T mx0 = T(0);
            for (int n = a0 + tid; n < a1 && n < a.N; n += THREADS) mx0 = max(mx0, fabs(s_c[n - a0]));
            mx0 = warp_max(mx0);
            if (lane == 0) red_v[warp] = mx0;
            __syncthreads();
            T m = red_v[0];
            for (int w = 1; w < NW; ++w) m = max(m, red_v[w]);
            push_max(m, true);
            sync_all();
            lam = reduce_max(true);
            __syncthreads();
        }
        // min-gamma search (:122-163): join candidates of the slice, leave candidates in CTA 0
        T best = TMAX;
        uint32_t bidx = 0xffffffffu;
        // a slice of <= 64 atoms (8 per CTA as a grid, 64 as a 16-cluster at N = 1024) is searched by warp 0 alone: one warp-level
        // argmin and no block barrier; (value, index) is a total order, so the winner does not depend on who compares what
        const bool small = a.cnt <= 64;
        auto join_candidates = [&](int first, int stride) {
            for (int n = a0 + first; n < a1 && n < a.N; n += stride) {
                if (s_act[n - a0]) continue;
                const T cc = s_c[n - a0], qq = s_q[n - a0];
                const T dl = T(1) - qq, dr = T(1) + qq;
                if (dl != T(0)) { const T t = (lam - cc) / dl; if (t > T(0) && t < TMAX && cand_less(t, (uint32_t)n, best, bidx & 0x7fffffffu)) { best = t; bidx = (uint32_t)n; } }
                if (dr != T(0)) { const T t = (lam + cc) / dr; if (t > T(0) && t < TMAX && cand_less(t, (uint32_t)n, best, bidx & 0x7fffffffu)) { best = t; bidx = (uint32_t)n; } }
            }
            if (cta == 0)
                for (int k = first; k < K; k += stride) {
                    const T xk_ = s_x[k]; const T t = xk_ == T(0) ? T(0) : -xk_ / s_d[k];   // x = 0 (the atom that just joined) is no candidate either way: skip the division's slow path
                    if (t > T(0) && t < TMAX && cand_less(t, s_gam[k], best, bidx & 0x7fffffffu)) { best = t; bidx = s_gam[k] | 0x80000000u; }
                }
        };
        if (small) {
            if (warp == 0) {
                // <= 2 atoms and <= 1 active slot per lane (more slots: the loop): every quotient is formed first, side by side, then
                // they are compared in the order of join_candidates
                T tl[2], tr[2], tx = T(0);
                bool live[2];
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    const int n = a0 + lane + 32 * u;
                    live[u] = n < a1 && n < a.N && !s_act[min(n, a1 - 1) - a0];
                    const T cc = live[u] ? s_c[n - a0] : T(0), qq = live[u] ? s_q[n - a0] : T(0);
                    tl[u] = (lam - cc) / (T(1) - qq);
                    tr[u] = (lam + cc) / (T(1) + qq);
                }
                const bool leave = cta == 0 && lane < K;
                if (leave && s_x[lane] != T(0)) tx = -s_x[lane] / s_d[lane];   // x = 0 (the atom that just joined): no candidate, and a zero numerator takes the division's slow path
#pragma unroll
                for (int u = 0; u < 2; ++u)
                    if (live[u]) {
                        const uint32_t n = (uint32_t)(a0 + lane + 32 * u);
                        const T qq = s_q[n - a0];
                        if (T(1) - qq != T(0) && tl[u] > T(0) && tl[u] < TMAX && cand_less(tl[u], n, best, bidx & 0x7fffffffu)) { best = tl[u]; bidx = n; }
                        if (T(1) + qq != T(0) && tr[u] > T(0) && tr[u] < TMAX && cand_less(tr[u], n, best, bidx & 0x7fffffffu)) { best = tr[u]; bidx = n; }
                    }
                if (leave && tx > T(0) && tx < TMAX && cand_less(tx, s_gam[lane], best, bidx & 0x7fffffffu)) { best = tx; bidx = s_gam[lane] | 0x80000000u; }
                if (cta == 0)
                    for (int k = lane + 32; k < K; k += 32) {
                        const T xk_ = s_x[k]; const T t = xk_ == T(0) ? T(0) : -xk_ / s_d[k];   // x = 0 (the atom that just joined) is no candidate either way: skip the division's slow path
                        if (t > T(0) && t < TMAX && cand_less(t, s_gam[k], best, bidx & 0x7fffffffu)) { best = t; bidx = s_gam[k] | 0x80000000u; }
                    }
                SSB_PROF(16);
                warp_argmin_flagged_fast(best, bidx);
                SSB_PROF(17);
            }
        } else {
            join_candidates(tid, THREADS);
            block_argmin(best, bidx);
        }
        push_pair(best, bidx);
        SSB_PROF(18);
        if (cta == 0) push_index(G, s_act[0]);      // membership of atom 0, for the no-candidate case
        SSB_PROF(2);
        sync_all();
        SSB_PROF(3);

        reduce_pairs(best, bidx);
        if (bidx == 0xffffffffu) bidx = get_i(G) ? 0x80000000u : 0u;   // (max, 0): the caller toggles column 0 (:122-123)
        SSB_PROF(4);
        const T gamma = best;
        const bool present = (bidx >> 31) != 0;
        const uint32_t idx = bidx & 0x7fffffffu;
        const int Knew = present ? K - 1 : K + 1;
        if (Knew == 0) {   // :248-249
            if (cta == 0 && tid == 0) { *it_out = it; *err_out = (double)lam; }
            break;
        }
        if (Knew > a.Kmax) {   // the host gave this launch fewer slots than min(max_iterations + 1, N) (a full-rank active set has at most
                               // M atoms): nothing has been written for this iteration; the caller solves again on the lockstep path
            if (cta == 0 && tid == 0) { *it_out = 0xffffffffu; *err_out = (double)lam; }
            break;
        }
        // c -= gamma q on the slice, |c|_inf partial (:255, :270).  Cluster form: into the slice's second buffer, so that CTA 0 can form
        // c at the active atoms itself from the owners' old c and q while they are still at it -- no barrier before the algebra
        {
            T* const c_out = CLUSTER ? s_c2 : s_c;
            T mx = T(0);
            auto step_c = [&](int first, int stride) {
                for (int n = a0 + first; n < a1; n += stride) {
                    const T cn = s_c[n - a0] - gamma * s_q[n - a0];
                    c_out[n - a0] = cn;
                    if constexpr (!CLUSTER) a.Cg[n] = cn;
                    if (n < a.N) mx = max(mx, fabs(cn));
                }
            };
            if ((int)idx >= a0 && (int)idx < a1 && tid == 0) s_act[idx - a0] ^= 1u;
            if (small) {
                if (warp == 0) { step_c(lane, 32); push_max(warp_max(mx), false); }
            } else {
                step_c(tid, THREADS);
                mx = warp_max(mx);
                if (lane == 0) red_v[warp] = mx;
                __syncthreads();
                T m = red_v[0];
                for (int w = 1; w < NW; ++w) m = max(m, red_v[w]);
                push_max(m, false);
            }
        }
        SSB_PROF(5);

        // x update, toggle, online inverse, new direction, next right-hand side(s): CTA 0
        auto algebra = [&](bool publish) {
            const int ldm = a.ks;
            // what the bordering needs from global memory is asked for first and used last
            T pre_u1 = T(0), pre_dot = T(0), pre_x = T(0);
            if (!present) {
                if (tid < K) pre_u1 = a.G[(size_t)idx * a.ldg + s_gam[tid]];
                pre_dot = a.G[(size_t)idx * a.ldg + idx];
                if (tid == 0) pre_x = xrow[idx];
            }
            // x += gamma * d with the OLD direction (:252), multiply and add rounded separately (see homotopy_scan_kernel)
            for (int k = tid; k < K; k += THREADS) { const T xv = add_rn(s_x[k], mul_rn(gamma, s_d[k])); s_x[k] = xv; xrow[s_gam[k]] = xv; }
            int slot = -1;
            if (present) { for (int k = 0; k < K; ++k) if (s_gam[k] == idx) slot = k; }
            const int ncg = K + (present ? 0 : 1);
            for (int k = tid; k < ncg; k += THREADS) {
                const int n = (int)(k < K ? s_gam[k] : idx);
                if constexpr (CLUSTER) {   // the owner's c and q as they stood before the step: the owner's own expression, so its bits
                    cg::cluster_group cl = cg::this_cluster();
                    const int owner = n / a.cnt, off = n - owner * a.cnt;
                    const T co = cl.map_shared_rank(s_c, owner)[off], qo = cl.map_shared_rank(s_q, owner)[off];
                    s_cg[k] = co - gamma * qo;
                } else s_cg[k] = __ldcg(a.Cg + n);
            }
            __syncthreads();
            SSB_PROF(8);
            if (!present) {
                // online_inverse.h:203-249 (bordering); the new atom takes slot K
                if (tid < K) s_u1[tid] = pre_u1;
                for (int k = tid + THREADS; k < K; k += THREADS) s_u1[k] = a.G[(size_t)idx * a.ldg + s_gam[k]];
                const T dot = pre_dot;
                __syncthreads();
                SSB_PROF(19);
                matvec<T, THREADS, true>(s_G, ldm, K, s_u1, s_u2, nullptr);
                __syncthreads();
                SSB_PROF(20);
                T uu = T(0);
                for (int k = lane; k < K; k += 32) uu += s_u1[k] * s_u2[k];
                uu = warp_sum(uu);
                const T dd = T(1) / (dot - uu);
                SSB_PROF(21);
                for (int r = warp; r < K; r += NW) {          // a warp per row, lanes along it (no index split; the same expression)
                    const T t = dd * s_u2[r];
                    T* g = s_G + (size_t)r * ldm;
                    for (int k = lane; k < K; k += 32) g[k] += t * s_u2[k];
                }
                for (int r = tid; r < K; r += THREADS) {
                    const T u3 = -dd * s_u2[r];
                    s_G[(size_t)r * ldm + K] = u3;
                    s_G[(size_t)K * ldm + r] = u3;
                }
                if (tid == 0) { s_G[(size_t)K * ldm + K] = dd; s_gam[K] = idx; s_x[K] = pre_x; }
                SSB_PROF(22);
            } else {
                // online_inverse.h:253-293; the leaving atom sits in slot j
                const int j = slot;
                const T dd = s_G[(size_t)j * ldm + j];
                const T sc = -(T(1) / dd);
                __syncthreads();
                for (int r = tid; r < K; r += THREADS) s_u1[r] = s_G[(size_t)r * ldm + j] * sc;
                __syncthreads();
                for (int r = warp; r < K; r += NW) {
                    if (r == j) continue;
                    const T t = -dd * s_u1[r];
                    T* g = s_G + (size_t)r * ldm;
                    for (int k = lane; k < K; k += 32) if (k != j) g[k] += t * s_u1[k];
                }
                __syncthreads();
                const int last = K - 1;
                if (j != last) {
                    for (int k = tid; k < last; k += THREADS) {
                        const int kk = (k == j) ? last : k;
                        const T vrow = s_G[(size_t)last * ldm + kk];
                        const T vcol = s_G[(size_t)kk * ldm + last];
                        s_G[(size_t)j * ldm + k] = vrow;
                        s_G[(size_t)k * ldm + j] = vcol;
                    }
                    if (tid == 0) { s_gam[j] = s_gam[last]; s_cg[j] = s_cg[last]; s_x[j] = s_x[last]; }
                }
            }
            __syncthreads();
            SSB_PROF(9);
            // d_G = inv * sign(c_G, tol) (:259-263)
            for (int k = tid; k < Knew; k += THREADS) s_cg[k] = sign_tol(s_cg[k], a.tol);
            __syncthreads();
            matvec<T, THREADS, true>(s_G, ldm, Knew, s_cg, s_d, nullptr);
            __syncthreads();
            SSB_PROF(10);
            if (publish && !gram_next) publish_rhs(Knew);
            if constexpr (!CLUSTER) {
                if (gram_next) for (int k = tid; k < Knew; k += THREADS) { a.pub_gam[k] = s_gam[k]; a.pub_d[k] = s_d[k]; a.pub_x[k] = s_x[k]; }
            }
        };

        T lam_new;
        bool cont;
        if constexpr (CLUSTER) {
            // two barriers per iteration: the algebra runs beside the other CTAs' c update, its right-hand side is published -- and
            // wasted -- even when the stop test, which needs every CTA's |c|_inf, then ends the solve
            if (cta == 0) algebra(true);
            { T* const t_ = s_c; s_c = s_c2; s_c2 = t_; }
            SSB_PROF(11);
            sync_all();
            SSB_PROF(12);
            lam_new = reduce_max(false);
            cont = (it < a.max_iter) && (lam_new > a.tol);   // :272
            if (!cont && cta == 0 && tid == 0) { *it_out = it; *err_out = (double)lam_new; }   // :274
            SSB_PROF(7);
        } else {
            sync_all();
            SSB_PROF(6);
            lam_new = reduce_max(false);
            cont = (it < a.max_iter) && (lam_new > a.tol);   // :272
            SSB_PROF(7);
            if (cta == 0) {
                algebra(cont);
                if (!cont && tid == 0) { *it_out = it; *err_out = (double)lam_new; }   // :274
            }
            SSB_PROF(11);
        }
        K = Knew;
        lam = lam_new;
        if (!cont) break;
        if constexpr (!CLUSTER) { sync_all(); SSB_PROF(12); }
#if SSB_SINGLE_PROF
        if (tid == 0 && cta == 0) s_prof[15] += 1;
#endif
    }
#if SSB_SINGLE_PROF
    __syncthreads();
    if (cta == 0 && threadIdx.x < 32) atomicAdd(&g_single_prof[threadIdx.x], s_prof[threadIdx.x]);
#endif
    // scratch and shared memory are reused by the next signal; in cluster mode nobody may leave (or move on) while CTA 0 still reads
    // the others' shared memory
    if (CLUSTER || si + 1 < nsig || a.x_out) sync_all();
    if (!a.resume && a.x_out)
        for (int n = a0 + tid; n < a1 && n < a.N; n += THREADS) a.x_out[n] = __ldcg(xrow + n);
    }
}

}  // namespace ssb
