// Member definitions of ssb::Homotopy<T> (included by homotopy.cuh).
#pragma once

#include <chrono>
#include <cstdlib>
#include <functional>

namespace ssb {

template <typename T>
void Homotopy<T>::finish_dictionary()
{
    // Gram matrix G = A^T A, [Npad][Npad]: the contraction with P := At
    counters[0] += launch_contract_simt<T>(At_, Mp, Npad, Mp, At_, Mp, Npad, G_, Npad, stream);
    if constexpr (sizeof(T) == 4) {
        tf32_split_kernel<<<1024, 256, 0, stream>>>(At_, At_hi_, At_lo_, (size_t)Npad * Mp);
        SSB_CUDA(cudaGetLastError());
        SSB_CUDA(cudaMemsetAsync(amax_bits_, 0, sizeof(unsigned int), stream));
        f16_split_kernel<<<(Npad + 7) / 8, 256, 0, stream>>>(reinterpret_cast<const float*>(At_), At16_hi_, At16_lo_, a_inv_, amax_bits_, Npad, Mp);
        SSB_CUDA(cudaGetLastError());
        SSB_CUDA(cudaMemcpyAsync(&amax_, amax_bits_, sizeof(float), cudaMemcpyDeviceToHost, stream));
        counters[0] += 2;
    }
    // mutual coherence (from G): which residual form and whether the Gram shortcut may be used (exact_residual())
    SSB_CUDA(cudaMemsetAsync(amax_bits_ + 1, 0, sizeof(unsigned int), stream));
    coherence_kernel<T><<<512, 256, 0, stream>>>(G_, Npad, N, amax_bits_ + 1);
    SSB_CUDA(cudaGetLastError());
    SSB_CUDA(cudaMemcpyAsync(&coherence_, amax_bits_ + 1, sizeof(float), cudaMemcpyDeviceToHost, stream));
    counters[0]++;
    SSB_CUDA(cudaStreamSynchronize(stream));
    for (auto& l : lanes_) l->drop_plans();
    plan_valid_ = false;
}

template <typename T>
void Homotopy<T>::check_capacity(int kmax) const
{
    // the step kernel keeps up to 16 active-set slots per thread of 256 (launch_step: SSB_STEP_BIG)
    SSB_REQUIRE(kmax <= 16 * 256, "max_iterations too large: the active set is limited to 4096 atoms (min(max_iterations + 1, N))");
}

// marks the dictionary slab as persisting in L2 for work on `st` (c, q, P, Ginv traffic is streaming)
template <typename T>
void Homotopy<T>::apply_l2_policy(cudaStream_t st)
{
    int max_window = 0, max_persist = 0;
    cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, device);
    cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, device);
    if (max_window <= 0 || max_persist <= 0 || !dict_.p) { cudaGetLastError(); return; }
    if (const char* e = getenv("SSB200_L2_PERSIST")) if (atoi(e) == 0) return;                        // tuning knob
    // set aside only what the dictionary slab needs: the rest of L2 keeps serving the streaming state
    const size_t want = std::min<size_t>(round_up(dict_.bytes, (size_t)1 << 20), (size_t)max_persist);
    if (l2_set_aside_ != want) { cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want); l2_set_aside_ = want; }
    cudaStreamAttrValue v{};
    const size_t win = std::min<size_t>(dict_.bytes, (size_t)max_window);
    v.accessPolicyWindow.base_ptr = dict_.p;
    v.accessPolicyWindow.num_bytes = win;
    v.accessPolicyWindow.hitRatio = (float)std::min(1.0, (double)want / (double)win);
    v.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
    v.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
    cudaStreamSetAttribute(st, cudaStreamAttributeAccessPolicyWindow, &v);
    cudaGetLastError();
}

template <typename T>
void Homotopy<T>::plan_batch(size_t batch, int kmax, bool host_path, size_t& chunk, int& nlanes)
{
    // the plan of the previous call with the same shape is reused: cudaMemGetInfo goes through the resource manager and was
    // measured to take 10-80 ms every few calls on a busy box (tools/variance.py: the only slow part of a slow step)
    const char* env_l = getenv("SSB200_LANES"); const char* env_c = getenv("SSB200_CHUNK");
    const plan_key key{ batch, kmax, host_path, env_l ? atoi(env_l) : -1, env_c ? atoi(env_c) : -1, exact_residual(), timing };
    if (plan_valid_ && key == plan_key_) { chunk = plan_chunk_; nlanes = plan_lanes_; return; }
    size_t free_b = 0, total_b = 0;
    SSB_CUDA(cudaMemGetInfo(&free_b, &total_b));
    size_t held = 0;
    for (auto& l : lanes_) held += l->ws_bytes();
    const size_t budget = std::min<size_t>((size_t)(0.6 * (double)free_b) + held, (size_t)32 << 30);
    // host-pointer entry: two lanes, so that one lane's copies, thin tail iterations and host round trips hide under
    // the other's kernels (C2: 84.4 -> 52.9 ms/step).  Device-pointer entry: one lane with the largest chunk (two lanes
    // gain 2 % there and blur the per-kernel event timing the roofline is computed from).
    // Below ~32 K signals a lane's kernels are short enough (35-100 us) for the gaps between them to show: two lanes side by side fill
    // them (8,192 signals: 4.69 -> 4.34 ms per step, 16,384: 8.17 -> 7.82) -- unless the per-kernel events are on.
    nlanes = (host_path && batch >= 8192) ? 2 : (!host_path && !timing && batch >= 4096 && batch <= 32768) ? 2 : 1;
    if (const char* e = getenv("SSB200_LANES")) nlanes = std::max(1, std::min(MAX_LANES, atoi(e)));   // tuning knob
    size_t cap = std::max<size_t>(1, budget / nlanes / bytes_per_signal(kmax));
    size_t cap_max = 65536;
    if (const char* e = getenv("SSB200_CHUNK")) cap_max = std::max(256, atoi(e));                     // tuning knob
    cap = std::min<size_t>(cap, cap_max);
    if (cap >= 256) cap = cap / 256 * 256;
    // even split over the lanes, in multiples of 256 signals
    size_t per = (batch + nlanes - 1) / nlanes;
    if (per > 256) per = round_up(per, 256);
    const size_t pieces = (per + cap - 1) / cap;
    chunk = (per + pieces - 1) / pieces;
    if (chunk > 256) chunk = round_up(chunk, 256);
    chunk = std::min(std::min(chunk, cap), batch);
    plan_key_ = key; plan_chunk_ = chunk; plan_lanes_ = nlanes; plan_valid_ = true;
}

template <typename T>
void Homotopy<T>::reserve_workspace(Lane& L, size_t rows, int kmax)
{
    const size_t rows_alloc = round_up(rows, 256);
    const int nb = exact_residual() ? 2 : 1;       // exact mode: a block of residual rows behind the direction rows
    if (rows_alloc <= L.ws_rows && kmax <= L.ws_kmax && nb == L.ws_nb) return;
    const size_t r = std::max(rows_alloc, L.ws_rows);
    const int k = std::max(kmax, L.ws_kmax);
    L.drop_plans();
    L.C.reserve(r * Npad * sizeof(T));
    if (nb == 2) L.C0.reserve(r * Npad * sizeof(T));
    L.Q.reserve(nb * r * Npad * sizeof(T));
    for (int i = 0; i < 2; ++i) {
        const size_t want = nb * r * Mp * sizeof(T);
        const bool grew = L.P[i].bytes < want;
        L.P[i].reserve(want);
        if (grew) SSB_CUDA(cudaMemsetAsync(L.P[i].p, 0, L.P[i].bytes, L.stream));   // rows past `count` feed the tensor tiles
        if (sizeof(T) == 4) {
            const bool grew_lo = L.Plo[i].bytes < want;
            L.Plo[i].reserve(want);
            if (grew_lo) SSB_CUDA(cudaMemsetAsync(L.Plo[i].p, 0, L.Plo[i].bytes, L.stream));
        }
        L.act[i].reserve(r * sizeof(int));
        L.Pinv[i].reserve(nb * r * sizeof(float));
    }
    L.gam.reserve(r * k * sizeof(uint32_t));
    L.dG.reserve(r * k * sizeof(T));
    L.Ginv.reserve(r * (size_t)k * k * sizeof(T));
    L.lam.reserve(r * sizeof(T));
    L.K.reserve(r * sizeof(int));
    L.pick.reserve(r * sizeof(uint32_t));
    L.mask.reserve(r * (size_t)(Npad / 32) * sizeof(uint32_t));
    L.ymax.reserve(r * sizeof(float));
    L.counts.reserve(COUNT_RING * sizeof(int));
    L.h_count.reserve(COUNT_RING * sizeof(int));
    SSB_CUDA(cudaStreamSynchronize(L.stream));
    L.ws_rows = r;
    L.ws_kmax = k;
    L.ws_nb = nb;
    apply_l2_policy(L.stream);
}

// Q[b * ws_rows + r][:] = A^T P[which][b * ws_rows + r][:] for r < rows, b < nblk (block 0: directions, block 1: residuals)
template <typename T>
void Homotopy<T>::contract(Lane& L, bool tc, int which, size_t rows, const int* rows_dev, bool count_rows, int nblk)
{
    nvtx_range nv(nblk == 2 ? "ssb200 contraction [Q; C'] = A^T [P; R]" : "ssb200 contraction Q = A^T P");
    const int span = L.begin(timing, 0);
    const size_t prow = (size_t)L.ws_nb * L.ws_rows;          // rows of the P buffers
    if constexpr (sizeof(T) == 4) {
        if (tc) {
            const bool f16 = use_f16();
            if (L.plan_f16 != f16) { L.drop_plans(); L.plan_f16 = f16; }
            if (!L.plan[which]) {
                if (f16) {
                    __half* hi = L.Plo[which].template as<__half>();
                    L.plan[which] = tc_plan_create(true, At16_hi_, At16_lo_, Npad, Mp, hi, hi + prow * Mp, (int)prow,
                                                   a_inv_, L.Pinv[which].template as<float>());
                } else {
                    L.plan[which] = tc_plan_create(false, At_hi_, At_lo_, Npad, Mp, L.P[which].p, L.Plo[which].p, (int)prow,
                                                   nullptr, nullptr);
                }
            }
            tc_contract(L.plan[which], (int)rows, L.Q.template as<float>(), Npad, L.stream, rows_dev, nblk, (int)L.ws_rows);
            L.counters[0]++;
        }
    }
    if (!tc)
        for (int b = 0; b < nblk; ++b)
            L.counters[0] += launch_contract_simt<T>(At_, Mp, Npad, Mp, L.P[which].template as<T>() + (size_t)b * L.ws_rows * Mp, Mp, (int)rows,
                                                     L.Q.template as<T>() + (size_t)b * L.ws_rows * Npad, Npad, L.stream, contraction_path != 4);
    L.end(span);
    L.counters[1]++;
    if (count_rows) L.counters[2] += (uint64_t)nblk * rows;
}


constexpr int STEP_KR = 3, STEP_KR_BIG = 16;   // active-set slots per thread of the step kernel: K <= KR * threads

template <typename T, int THREADS, int NCH, bool GS, int KR = STEP_KR>
static void launch_step_variant(unsigned grid, size_t smem, cudaStream_t st, const HomArgs<T>& a, int kb, int ks)
{
    static std::atomic<uint64_t> configured{ 0 };
    if (smem > 48 * 1024) optin_shared_memory((const void*)homotopy_step_kernel<T, THREADS, NCH, GS, KR>, configured);
    homotopy_step_kernel<T, THREADS, NCH, GS, KR><<<grid, THREADS, smem, st>>>(a, kb, ks);
    SSB_CUDA(cudaGetLastError());
}

template <typename T>
static void launch_step(int threads, int nch, bool gs, unsigned grid, size_t smem, cudaStream_t st, const HomArgs<T>& a, int kb, int ks)
{
    // active sets beyond 3 slots per thread (max_iterations > 767): 256 threads, 16 slots each, inverse worked on in HBM
#define SSB_STEP_BIG(NC)                                                                          \
    if (a.Kmax > STEP_KR * threads && threads == 256 && nch == NC) {                              \
        launch_step_variant<T, 256, NC, false, STEP_KR_BIG>(grid, smem, st, a, kb, ks);           \
        return;                                                                                   \
    }
    SSB_STEP_BIG(1) SSB_STEP_BIG(2) SSB_STEP_BIG(4) SSB_STEP_BIG(0)
#undef SSB_STEP_BIG
#define SSB_STEP(TH, NC)                                                                          \
    if (threads == TH && nch == NC) {                                                             \
        if (gs) launch_step_variant<T, TH, NC, true>(grid, smem, st, a, kb, ks);                  \
        else    launch_step_variant<T, TH, NC, false>(grid, smem, st, a, kb, ks);                 \
        return;                                                                                   \
    }
    SSB_STEP(128, 1) SSB_STEP(128, 2) SSB_STEP(128, 4) SSB_STEP(256, 1) SSB_STEP(256, 2) SSB_STEP(256, 4) SSB_STEP(256, 0)
#undef SSB_STEP
    throw invalid_argument{ "no step-kernel variant for this shape" };
}


template <typename T>
static void launch_scan(int threads, int nch, bool gram, unsigned grid, cudaStream_t st, const HomArgs<T>& a, int kb)
{
    static const bool one_slot_off = [] { const char* e = getenv("SSB200_SCAN_KR3"); return e && atoi(e) != 0; }();   // A/B switch
#define SSB_SCAN_KR(TH, NC, KR)                                                                   \
        if (a.exact) { if (gram) homotopy_scan_kernel<T, TH, NC, true, true, KR><<<grid, TH, 0, st>>>(a, kb);     \
                       else      homotopy_scan_kernel<T, TH, NC, false, true, KR><<<grid, TH, 0, st>>>(a, kb); } \
        else         { if (gram) homotopy_scan_kernel<T, TH, NC, true, false, KR><<<grid, TH, 0, st>>>(a, kb);    \
                       else      homotopy_scan_kernel<T, TH, NC, false, false, KR><<<grid, TH, 0, st>>>(a, kb); }
#define SSB_SCAN(TH, NC)                                                                          \
    if (threads == TH && nch == NC) {                                                             \
        if (kb <= TH && !one_slot_off) { SSB_SCAN_KR(TH, NC, 1) } else { SSB_SCAN_KR(TH, NC, 3) }  \
        SSB_CUDA(cudaGetLastError()); return; }
    SSB_SCAN(128, 1) SSB_SCAN(128, 2) SSB_SCAN(128, 4) SSB_SCAN(256, 1) SSB_SCAN(256, 2) SSB_SCAN(256, 4) SSB_SCAN(256, 0)
#undef SSB_SCAN
#undef SSB_SCAN_KR
    throw invalid_argument{ "no scan-kernel variant for this shape" };
}

template <typename T, int W, bool EXACT>
static void launch_algebra_variant(unsigned rows, size_t slice, cudaStream_t st, const HomArgs<T>& a, int kb, int ks)
{
    static std::atomic<uint64_t> configured{ 0 };
    const size_t smem = slice * W;
    if (smem > 48 * 1024) optin_shared_memory((const void*)homotopy_algebra_kernel<T, W, EXACT>, configured);
    homotopy_algebra_kernel<T, W, EXACT><<<(rows + W - 1) / W, W * 32, smem, st>>>(a, kb, ks, (int)slice);
    SSB_CUDA(cudaGetLastError());
}

// warps per CTA of the algebra kernel for a slice size; 0 = the active set is too large for the split kernels
// The warp-per-signal algebra kernel pays off while a warp's K x K work is short: measured on C3 (K grows to 100) a slice
// limit of 8-16 KB (K <= ~45-60) runs the step in 167-176 ms per 65,536 signals, 24 KB in 191 ms, 48 KB in 332 ms; past
// the limit the CTA-per-signal step kernel takes over.
static inline int algebra_warps(size_t slice) {
    int lim = 10 * 1024;
    if (const char* e = getenv("SSB200_ALG_SLICE")) lim = atoi(e);                                     // tuning knob
    return slice <= (size_t)lim ? 4 : 0;
}

template <typename T>
static void launch_algebra(int W, unsigned rows, size_t slice, cudaStream_t st, const HomArgs<T>& a, int kb, int ks)
{
    if (a.exact) {
        if (W == 4) launch_algebra_variant<T, 4, true>(rows, slice, st, a, kb, ks);
        else if (W == 2) launch_algebra_variant<T, 2, true>(rows, slice, st, a, kb, ks);
        else launch_algebra_variant<T, 1, true>(rows, slice, st, a, kb, ks);
        return;
    }
    if (W == 4) launch_algebra_variant<T, 4, false>(rows, slice, st, a, kb, ks);
    else if (W == 2) launch_algebra_variant<T, 2, false>(rows, slice, st, a, kb, ks);
    else launch_algebra_variant<T, 1, false>(rows, slice, st, a, kb, ks);
}


template <typename T>
void Homotopy<T>::solve_chunk(Lane& L, const T* dY, size_t rows, size_t ldy, T tol, uint32_t max_iter, T* dX,
                              size_t ldx, uint32_t* d_iters, double* d_errs)
{
    cudaStream_t st = L.stream;
    const bool tc = use_tc(rows);
    const bool exact = L.ws_nb == 2;                   // reserve_workspace sized the row blocks from exact_residual()
    const size_t boff = L.ws_rows;                     // rows between the direction block and the residual block
    const size_t prow = (size_t)L.ws_nb * L.ws_rows;
    HomArgs<T> a{};
    a.N = N; a.ldc = Npad; a.Mp = Mp; a.ldp = Mp; a.Kmax = L.ws_kmax;
    a.At = At_; a.G = G_; a.ldg = Npad;
    a.C = L.C.template as<T>(); a.Q = L.Q.template as<T>(); a.X = dX; a.ldx = ldx;
    a.gam = L.gam.template as<uint32_t>(); a.dG = L.dG.template as<T>(); a.Ginv = L.Ginv.template as<T>();
    a.lam = L.lam.template as<T>(); a.Kcur = L.K.template as<int>(); a.pick = L.pick.template as<uint32_t>(); a.mask = L.mask.template as<uint32_t>();
    a.iters = d_iters; a.errs = d_errs;
    a.exact = exact ? 1 : 0;
    a.Qc = exact ? L.Q.template as<T>() + boff * Npad : nullptr;
    a.Y = dY; a.ldy = ldy; a.M = M; a.ymax = L.ymax.template as<float>();
    const int tc_split = tc ? (use_f16() ? 2 : 1) : 0;
    a.split = tc_split;
    a.amax = amax_;
    // where right-hand sides number i go for the current split (hi / full, lo, inverse scales); exact mode: the residual
    // rows of the same buffers, one block (ws_rows rows) further
    auto set_rhs = [&](int i, int split) {
        if (split == 2) {
            __half* hi = L.Plo[i].template as<__half>();
            __half* lo = hi + prow * Mp;
            a.Pn = reinterpret_cast<T*>(hi);
            a.Pn_lo = reinterpret_cast<T*>(lo);
            a.Pn_inv = L.Pinv[i].template as<float>();
            a.Rn = reinterpret_cast<T*>(hi + boff * Mp);
            a.Rn_lo = reinterpret_cast<T*>(lo + boff * Mp);
            a.Rn_inv = a.Pn_inv + boff;
        } else {
            a.Pn = L.P[i].template as<T>();
            a.Pn_lo = split ? L.Plo[i].template as<T>() : nullptr;
            a.Pn_inv = nullptr;
            a.Rn = a.Pn + boff * Mp;
            a.Rn_lo = split ? a.Pn_lo + boff * Mp : nullptr;
            a.Rn_inv = nullptr;
        }
    };
    a.tol = tol; a.max_iter = max_iter;
    if (const char* e = getenv("SSB200_DBG")) a.dbg = atoi(e);

    // step-kernel variant: c,q register-resident when the dictionary is narrow enough, inverse on chip when it fits
    constexpr int V = 16 / sizeof(T);
    const int chunks128 = (Npad + 128 * V - 1) / (128 * V), chunks256 = (Npad + 256 * V - 1) / (256 * V);
    int threads = chunks128 <= 4 ? 128 : 256;
    int nch = chunks128 <= 1 ? 1 : chunks128 <= 2 ? 2 : chunks128 <= 4 ? 4 : chunks256 <= 4 ? 4 : 0;
    bool wide = L.ws_kmax > STEP_KR * 128;            // more active-set slots than 128 threads hold: 256 threads
    if (const char* e = getenv("SSB200_STEP_THREADS")) if (atoi(e) == 256) wide = true;               // tuning knob
    if (wide) { threads = 256; nch = chunks256 <= 1 ? 1 : chunks256 <= 2 ? 2 : chunks256 <= 4 ? 4 : 0; }
    const bool big = L.ws_kmax > STEP_KR * threads;   // ... than 256 threads hold at 3 each: the 16-slot variant
    const size_t g_budget = 96 * 1024;
    bool split_ok = true;                                                                            // scan + algebra kernels
    if (const char* e = getenv("SSB200_SPLIT")) split_ok = atoi(e) != 0;                              // tuning knob
    uint32_t RUN_AHEAD = 8;
    if (const char* e = getenv("SSB200_RUN_AHEAD")) RUN_AHEAD = (uint32_t)std::max(1, std::min((int)MAX_RUN_AHEAD, atoi(e)));   // tuning knob
    // While the active set is small (K <= it < gram_iters) q = G[:, active] d -- and, in exact mode, c' = A^T y - G[:, active] x
    // from the same rows -- costs K rows of G per signal in the scan kernel: less than a contraction, a K-atom gather and the Q
    // traffic.  Sums over rows of G cancel badly when the atoms are nearly parallel (measured: tests/studies/residual_modes.py,
    // "Gram form"), so coherent dictionaries do not take this route.
    uint32_t gram_iters = 16;                                    // iterations 1 .. gram_iters - 1 take q from the Gram matrix
    if (const char* e = getenv("SSB200_GRAM_ITERS")) gram_iters = (uint32_t)std::max(0, std::min(32, atoi(e)));   // tuning knob
    if (!split_ok || nch == 0 || coherent() || algebra_warps(homotopy_algebra_slice<T>((int)gram_iters, ((int)gram_iters + 1) | 1)) == 0) gram_iters = 0;
    gram_iters = std::min<uint32_t>(gram_iters, (uint32_t)L.ws_kmax);
    a.C0 = (exact && gram_iters > 1) ? L.C0.template as<T>() : nullptr;

    // c0 = A^T y for the whole chunk (homotopy-cpu.cpp:215)
    {
        set_rhs(0, a.split);
        const int blocks = (int)std::min<size_t>(8192, (rows + 7) / 8);
        stage_signals_kernel<T><<<blocks, 256, 0, st>>>(dY, ldy, M, (int)rows, a.Pn, a.Pn_lo, a.Pn_inv, Mp, a.split,
                                                        L.ymax.template as<float>());
        SSB_CUDA(cudaGetLastError());
        L.counters[0]++;
    }
    contract(L, tc, 0, rows, nullptr, true);
    {
        a.act = nullptr; a.act_next = nullptr; a.next_count = nullptr; a.count_zero = nullptr;
        set_rhs(1, a.split);
        a.it = 0;
        const int span = L.begin(timing, 1);
        homotopy_init_kernel<T, STEP_THREADS><<<(unsigned)rows, STEP_THREADS, 0, st>>>(a);
        SSB_CUDA(cudaGetLastError());
        L.end(span);
        L.counters[0]++;
    }

    // Lockstep loop.  The number of signals still running lives on the device: the count of iteration `it` is appended to by
    // iteration it and read by iteration it + 1 (every kernel bounds itself by it), so the host never has to wait for an
    // iteration before queueing the next.  The counts sit in a ring of COUNT_RING slots (any max_iterations): iteration it
    // clears the slot iteration it + 1 will append to -- its previous occupant, COUNT_RING iterations older, has been read by
    // its kernels and copied to the host long before, because the host stays at most RUN_AHEAD < COUNT_RING - 1 iterations
    // ahead of the newest count it has seen.  The read-backs travel on a side stream: behind another lane's bulk copy-out a
    // 4-byte copy can wait milliseconds, and on the main stream that wait would stall the kernels behind it.  Grids are
    // sized from the newest count the host has (counts only fall, so a stale one is a valid upper bound).
    // The SIMT contraction takes its row count on the host, so that path waits for every iteration.
    int sync_iters = 0;
    if (const char* e = getenv("SSB200_SYNC_ITERS")) sync_iters = atoi(e);                            // debugging aid
    int cur = 1;                 // P[cur] holds the right-hand sides of this iteration
    int* d_counts = L.counts.template as<int>();        // [COUNT_RING]
    int* h_ring = L.h_count.template as<int>();         // pinned mirror
    auto slot = [](uint32_t i) { return i % COUNT_RING; };
    SSB_CUDA(cudaMemsetAsync(d_counts, 0, COUNT_RING * sizeof(int), st));
    h_ring[0] = (int)rows;
    SSB_CUDA(cudaMemcpyAsync(d_counts, h_ring, sizeof(int), cudaMemcpyHostToDevice, st));
    while (L.count_events.size() < RUN_AHEAD + 1) {
        cudaEvent_t e, f;
        SSB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); L.count_events.push_back(e);
        SSB_CUDA(cudaEventCreateWithFlags(&f, cudaEventDisableTiming)); L.iter_events.push_back(f);
    }
    if (!L.side) SSB_CUDA(cudaStreamCreateWithFlags(&L.side, cudaStreamNonBlocking));
    const bool async = tc;
    size_t known = rows;         // newest count the host has seen (upper bound of the current one)
    uint32_t seen = 0;           // ... which is the count of iteration `seen`
    uint32_t last = 0;
    uint32_t thin_run = 0;       // consecutive iterations entered with at most a thin tile's worth of signals
    bool finished_by_kernel = false;
    std::vector<int> hist{ (int)rows };                    // hist[j] = signals that continue after iteration j
    std::vector<uint8_t> route{ 0 };                        // launch class of every iteration's scan (kernel profile)
    // debugging aid (SSB200_TRACE_STALL=<ms>): host timeline of a chunk that took longer than <ms> -- when each iteration was
    // queued, how long the host waited for a count, which count it had then
    double trace_ms = 0.0;
    if (const char* e = getenv("SSB200_TRACE_STALL")) trace_ms = atof(e);
    struct trace_row { double t_begin, t_wait, t_queued; uint32_t seen; size_t known; };
    std::vector<trace_row> trace;
    const auto t_chunk0 = std::chrono::steady_clock::now();
    auto now_ms = [&] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_chunk0).count(); };
    for (uint32_t it = 1; it <= max_iter; ++it) {
        const double t_begin = trace_ms > 0 ? now_ms() : 0.0;
        const uint32_t must = async ? (it > RUN_AHEAD ? it - RUN_AHEAD : 0) : it - 1;   // this iteration's count has to be in
        for (uint32_t j = seen + 1; j < it; ++j) {
            cudaEvent_t ev = L.count_events[j % (RUN_AHEAD + 1)];
            if (j <= must) SSB_CUDA(cudaEventSynchronize(ev));
            else if (cudaEventQuery(ev) != cudaSuccess) { cudaGetLastError(); break; }
            seen = j;
            known = (size_t)h_ring[slot(j)];
            hist.push_back((int)known);
        }
        const double t_wait = trace_ms > 0 ? now_ms() : 0.0;
        if (known == 0) break;
        // Stragglers: once the chunk has been down to a thin tile's worth of signals for FIN_AFTER iterations (the bulk converges
        // within four or five iterations of each other, so these go on to max_iterations) and no more than FIN_MAX are left,
        // they are handed to the finisher.  Only such signals ever see it, so a converging signal's arithmetic does not depend
        // on the batch it travels in.
        thin_run = known <= 32 ? thin_run + 1 : 0;
        if (known <= (size_t)FIN_MAX && thin_run > (uint32_t)FIN_AFTER && it > 1 && rows > 1 &&
            launch_finisher(L, a, it - 1, L.act[cur].template as<int>(), d_counts + slot(it - 1))) {
            finished_by_kernel = true;
            break;
        }
        last = it;
        const bool gram = it < gram_iters;
        a.skip_rhs = (it + 1 < gram_iters) ? 1 : 0;
        if (!gram) contract(L, tc, cur, known, async ? d_counts + slot(it - 1) : nullptr, false, L.ws_nb);
        a.split = tc ? tc_split : 0;
        a.act = (it == 1) ? nullptr : L.act[cur].template as<int>();
        a.act_next = L.act[cur ^ 1].template as<int>();
        a.count_in = d_counts + slot(it - 1);
        a.next_count = d_counts + slot(it);
        a.count_zero = d_counts + slot(it + 1);
        set_rhs(cur ^ 1, a.split);
        a.it = it;
        const int kb = (int)std::min<uint32_t>(it, (uint32_t)L.ws_kmax);     // K <= kb before iteration `it`
        const int ks = (kb + 1) | 1;                                          // odd row stride: conflict-free column reads
        const bool gs = !big && (size_t)(kb + 1) * ks * sizeof(T) <= g_budget;
        const size_t smem = homotopy_step_smem<T>(L.ws_kmax, Npad, gs ? kb + 1 : 0, ks);
        SSB_REQUIRE(smem <= 200 * 1024, "max_iterations too large for the step kernel's shared memory");
        const size_t slice = homotopy_algebra_slice<T>(kb, ks);
        const int aw = (gram || a.skip_rhs) ? 4 : (split_ok && known >= 1024) ? algebra_warps(slice) : 0;   // few signals: one CTA per signal has the shorter critical path
        const int span = L.begin(timing, 1);
        if (aw > 0) {
            const int sp_scan = L.begin(timing, gram ? 3 : 2);
            { nvtx_range nv(gram ? "ssb200 scan (q from Gram rows)" : "ssb200 scan (Q streamed)"); launch_scan(threads, nch, gram, (unsigned)known, st, a, kb); }
            L.end(sp_scan);
            const int sp_alg = L.begin(timing, 4);
            { nvtx_range nv("ssb200 algebra"); launch_algebra(aw, (unsigned)known, slice, st, a, kb, ks); }
            L.end(sp_alg);
            L.counters[0] += 2;
            route.push_back(gram ? 3 : 2);
        } else {
            const int sp_step = L.begin(timing, 5);
            { nvtx_range nv("ssb200 step (scan + algebra, one CTA per signal)"); launch_step(threads, nch, gs, (unsigned)known, smem, st, a, kb, ks); }
            L.end(sp_step);
            L.counters[0]++;
            route.push_back(5);
        }
        L.end(span);
        if (sync_iters) {   // debugging aid: find the iteration a device fault belongs to
            int hk = -1;
            T hl = T(0);
            uint32_t hp = 0;
            const cudaError_t e = cudaStreamSynchronize(st);
            if (e == cudaSuccess) { cudaMemcpy(&hk, a.Kcur, sizeof(int), cudaMemcpyDeviceToHost); cudaMemcpy(&hl, a.lam, sizeof(T), cudaMemcpyDeviceToHost);
                                    cudaMemcpy(&hp, a.gam + (hk > 0 ? hk - 1 : 0), sizeof(uint32_t), cudaMemcpyDeviceToHost); }
            if (e != cudaSuccess || sync_iters > 1) fprintf(stderr, "[ssb200] it %u known %zu kb %d K[0] %d lam[0] %.9g lastslot %u big %d gs %d aw %d exact %d: %s\n", it, known, kb, hk, (double)hl, hp, (int)big, (int)gs, aw, (int)exact, cudaGetErrorString(e));
            SSB_CUDA(e);
        }
        L.counters[3]++;
        SSB_CUDA(cudaEventRecord(L.iter_events[it % (RUN_AHEAD + 1)], st));
        SSB_CUDA(cudaStreamWaitEvent(L.side, L.iter_events[it % (RUN_AHEAD + 1)], 0));
        SSB_CUDA(cudaMemcpyAsync(h_ring + slot(it), d_counts + slot(it), sizeof(int), cudaMemcpyDeviceToHost, L.side));
        SSB_CUDA(cudaEventRecord(L.count_events[it % (RUN_AHEAD + 1)], L.side));
        cur ^= 1;
        if (trace_ms > 0) trace.push_back({ t_begin, t_wait, now_ms(), seen, known });
    }
    const double t_loop = trace_ms > 0 ? now_ms() : 0.0;
    SSB_CUDA(cudaStreamSynchronize(st));
    SSB_CUDA(cudaStreamSynchronize(L.side));
    for (uint32_t j = seen + 1; j <= last; ++j) hist.push_back(h_ring[slot(j)]);   // at most RUN_AHEAD + 1 unread slots, none reused yet
    (void)finished_by_kernel;
    if (trace_ms > 0 && now_ms() > trace_ms) {
        fprintf(stderr, "[ssb200] slow chunk: %.2f ms (loop queued by %.2f ms, %zu rows)\n", now_ms(), t_loop, rows);
        for (size_t i = 0; i < trace.size(); ++i)
            fprintf(stderr, "[ssb200]   it %3zu begin %8.3f waited %7.3f queued +%6.3f seen %u known %zu\n", i + 1, trace[i].t_begin,
                    trace[i].t_wait - trace[i].t_begin, trace[i].t_queued - trace[i].t_wait, trace[i].seen, trace[i].known);
    }
    for (uint32_t it = std::max<uint32_t>(1, gram_iters); it <= last; ++it) L.counters[2] += (uint64_t)L.ws_nb * (uint64_t)hist[it - 1];   // signal-rows contracted in the loop
    for (uint32_t it = 1; it <= last; ++it) {
        const uint64_t n = (uint64_t)hist[it - 1];           // signals that entered iteration `it`
        const uint64_t nc = (uint64_t)hist[it];              // ... and continue after it
        const uint64_t kb = std::min<uint64_t>(it, (uint64_t)L.ws_kmax);
        // algorithmic HBM bytes of the launch: c read (exact mode: the fresh c') + written, q read when it was
        // streamed; the K x K inverse read + written (K <= kb = it); the next right-hand side(s) written in container format
        const uint64_t cq = (uint64_t)Npad * sizeof(T), inv = 2 * kb * kb * sizeof(T), rhs = (uint64_t)L.ws_nb * Mp * (tc ? 4 : sizeof(T));
        L.rows[route[it]] += n;
        L.rows[1] += n;
        if (route[it] == 2) L.bytes[2] += n * 3 * cq;
        if (route[it] == 3) { L.bytes[3] += n * 2 * cq; L.bytes[6] += n * kb * cq; }   // [6]: the K <= kb rows of G per signal, from L2
        if (route[it] != 5) { L.rows[4] += nc; L.bytes[4] += nc * (inv + ((it + 1 < gram_iters) ? 0 : rhs)); }
        else L.bytes[5] += n * 3 * cq + nc * (inv + rhs);
    }
}

// folds a lane's counters and event-timed spans into the handle (stream already synchronised)
template <typename T>
void Homotopy<T>::merge(Lane& L)
{
    std::lock_guard<std::mutex> g(merge_mu_);
    for (int i = 0; i < 4; ++i) { counters[i] += L.counters[i]; L.counters[i] = 0; }
    for (int i = 0; i < NCLS; ++i) { kernel_rows[i] += L.rows[i]; L.rows[i] = 0; }
    for (int i = 0; i < 8; ++i) { kernel_bytes[i] += L.bytes[i]; L.bytes[i] = 0; }
    for (auto& s : L.spans) {
        float t = 0.f;
        if (cudaEventElapsedTime(&t, s.a, s.b) == cudaSuccess) ms[s.cls] += t;
        L.pool.push_back(s.a);
        L.pool.push_back(s.b);
    }
    L.spans.clear();
}

// runs body(lane, chunk_index) over nchunks chunks on nlanes lanes; lane 0 is the calling thread
template <typename T>
template <typename F>
void Homotopy<T>::run_lanes(int nlanes, size_t nchunks, F&& body)
{
    std::atomic<size_t> next{ 0 };
    std::exception_ptr err[MAX_LANES] = {};
    auto worker = [&](int li) {
        try {
            SSB_CUDA(cudaSetDevice(device));
            Lane& L = lane(li);
            for (;;) {
                const size_t c = next.fetch_add(1);
                if (c >= nchunks) break;
                body(L, c);
            }
            merge(L);
        } catch (...) {
            err[li] = std::current_exception();
            next.store(nchunks);     // stop the other lane from taking new chunks
        }
    };
    for (int li = 0; li < nlanes; ++li) lane(li);          // create lanes on the calling thread
    std::vector<std::thread> helpers;
    for (int li = 1; li < nlanes; ++li) helpers.emplace_back(worker, li);
    worker(0);
    for (auto& t : helpers) t.join();
    for (int li = 0; li < nlanes; ++li) if (err[li]) std::rethrow_exception(err[li]);
}


// Cluster form of the single-signal kernel.  A cluster of 16 (8 where the device refuses the non-portable size) owns a signal: an
// iteration's three barriers are barrier.cluster instead of grid.sync over 148 CTAs, which is most of what an iteration cost.
// The dictionary streams from L2 through the cluster's SMs only, so it applies while that is cheap (<= 8 MB: C1 1 MB, C2 4 MB).
template <typename T>
bool Homotopy<T>::launch_single_cluster(SingleArgs<T>& a, int nclusters, cudaStream_t st)
{
    if (const char* e = getenv("SSB200_SINGLE_CLUSTER")) if (atoi(e) == 0) return false;                      // tuning knob / A-B switch
    if ((size_t)Npad * Mp * sizeof(T) > ((size_t)8 << 20)) return false;
    constexpr int THREADS = 256;
    auto fn = homotopy_single_kernel<T, THREADS, true>;
    static std::atomic<uint64_t> smem_done{ 0 }, nonportable_done{ 0 };
    int want = SINGLE_CLUSTER_MAX;
    if (const char* e = getenv("SSB200_CLUSTER_SIZE")) want = std::max(2, std::min(SINGLE_CLUSTER_MAX, atoi(e)));
    for (int cs = want; cs >= 2; cs >>= 1) {
        const int cnt = (Npad + cs - 1) / cs;
        // the CTA's slice of the dictionary rides in shared memory when it fits (C1: 64 KB per CTA of a 16-cluster)
        int dict = 1;
        if (const char* e = getenv("SSB200_CLUSTER_DICT")) dict = atoi(e) != 0;
        int stage = 1;
        if (const char* e = getenv("SSB200_CLUSTER_STAGE")) stage = atoi(e) != 0;
        size_t smem = homotopy_single_smem<T>(Mp, cnt, a.Kmax, a.ks, dict != 0, stage != 0);
        if (smem > 200 * 1024) { dict = 0; smem = homotopy_single_smem<T>(Mp, cnt, a.Kmax, a.ks, false, stage != 0); }
        if (smem > 200 * 1024) { stage = 0; smem = homotopy_single_smem<T>(Mp, cnt, a.Kmax, a.ks, false, false); }
        if (smem > 200 * 1024) continue;
        if (smem > 48 * 1024) optin_shared_memory((const void*)fn, smem_done);
        if (cs > 8) {
            const uint64_t bit = 1ull << (device & 63);
            if (!(nonportable_done.load(std::memory_order_acquire) & bit)) {
                if (cudaFuncSetAttribute((const void*)fn, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) { cudaGetLastError(); continue; }
                nonportable_done.fetch_or(bit, std::memory_order_release);
            }
        }
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)(cs * nclusters));
        cfg.blockDim = dim3(THREADS);
        cfg.dynamicSmemBytes = smem;
        cfg.stream = st;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = (unsigned)cs; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        int fit = 0;
        if (cudaOccupancyMaxActiveClusters(&fit, (const void*)fn, &cfg) != cudaSuccess || fit < 1) { cudaGetLastError(); continue; }
        a.cnt = cnt;
        a.dict_smem = dict;
        a.stage = stage;
        SSB_CUDA(cudaLaunchKernelEx(&cfg, fn, a));
        return true;
    }
    return false;
}

template <typename T>
bool Homotopy<T>::launch_finisher(Lane& L, const HomArgs<T>& h, uint32_t it0, const int* d_act, const int* d_count)
{
    if (const char* e = getenv("SSB200_FINISHER")) if (atoi(e) == 0) return false;                     // tuning knob
    constexpr int THREADS = 256;
    int sms = 0, coop = 0;
    SSB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    SSB_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, device));
    if (!coop) return false;
    const int kmax = L.ws_kmax;
    const int ks = (kmax + 1) | 1;
    const int grid = std::min(sms, (Npad + 7) / 8);
    const int cnt = (Npad + grid - 1) / grid;
    const size_t smem = homotopy_single_smem<T>(Mp, cnt, kmax, ks);
    if (smem > 200 * 1024) return false;
    static std::atomic<uint64_t> configured{ 0 };
    if (smem > 48 * 1024) optin_shared_memory((const void*)homotopy_single_kernel<T, THREADS, false>, configured);
    // scratch: P | R | Cg | part_t | part_max | part_i
    const size_t nT = (size_t)2 * Mp + Npad + 2 * (size_t)grid + 8;
    L.fin_ws.reserve(nT * sizeof(T) + ((size_t)grid + 8) * sizeof(uint32_t) + 64);
    T* w = L.fin_ws.template as<T>();
    SingleArgs<T> a{};
    a.N = N; a.Npad = Npad; a.Mp = Mp; a.Kmax = kmax; a.ks = ks;
    a.At = At_; a.G = G_; a.ldg = Npad;
    a.P = w; a.R = w + Mp; a.Cg = a.R + Mp; a.part_t = a.Cg + Npad; a.part_max = a.part_t + grid;
    a.part_i = reinterpret_cast<uint32_t*>(reinterpret_cast<unsigned char*>(w) + round_up((size_t)(a.part_max + grid - w) * sizeof(T), 16));
    a.exact = h.exact; a.tol = h.tol; a.max_iter = h.max_iter; a.cnt = cnt;
    a.resume = 1; a.act = d_act; a.count = d_count; a.max_sigs = FIN_MAX; a.it0 = it0;
    a.Y = h.Y; a.ldy = h.ldy; a.M = h.M; a.X = h.X; a.ldx = h.ldx; a.C = h.C; a.ldc = h.ldc;
    a.gamv = h.gam; a.dGv = h.dG; a.Ginvv = h.Ginv; a.lamv = h.lam; a.Kcurv = h.Kcur; a.Kstate = h.Kmax;
    a.itersv = h.iters; a.errsv = h.errs;
    void* params[] = { &a };
    nvtx_range nv("ssb200 finisher (stragglers to the end)");
    // one cluster per straggler, side by side (an ordinary launch); else the cooperative grid, which needs every SM: two lanes
    // launching one each could each hold half the machine and wait forever, hence one at a time per handle
    const int span = L.begin(timing, 5);
    if (!launch_single_cluster(a, FIN_MAX, L.stream)) {
        std::lock_guard<std::mutex> g(coop_mu_);
        SSB_CUDA(cudaLaunchCooperativeKernel((const void*)homotopy_single_kernel<T, THREADS, false>, dim3((unsigned)grid), dim3(THREADS), params, smem, L.stream));
    }
    L.end(span);
    L.counters[0]++;
    SSB_CUDA(cudaStreamSynchronize(L.stream));
    return true;
}

template <typename T>
bool Homotopy<T>::solve_single(const T* y, bool y_on_host, T tol, uint32_t max_iter, T* x, bool x_on_host, uint32_t* iters,
                               double* errs, bool reports_on_host)
{
    if (contraction_path >= 2) return false;                      // a forced tensor path is honoured by the lockstep code
    if (const char* e = getenv("SSB200_SINGLE")) if (atoi(e) == 0) return false;                       // tuning knob
    constexpr int THREADS = 256;
    int sms = 0, coop = 0;
    SSB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    SSB_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, device));
    if (!coop) return false;
    // Slots for the active set.  min(max_iterations + 1, N) can never overflow; a full-rank active set has at most M atoms, so when the
    // host reads the report anyway the launch gets min(.., M + 1) slots -- the reference's bench passes max_iterations = N, which would
    // otherwise keep every N >= 256 cell off this kernel -- and a solve that does ask for more (rank-deficient numerics, tiny
    // tolerances) reports so and is solved again by the lockstep path.
    int kmax = active_set_capacity(max_iter, N);
    const int grid = std::min(sms, (Npad + 7) / 8);
    const int cnt = (Npad + grid - 1) / grid;
    if (reports_on_host && x_on_host) {
        kmax = std::min(kmax, M + 1);
        // ... and no more than the K x K inverse finds room for in shared memory (float32: ~210 slots): most solves of the reference's
        // 256 x N bench cells stay below that, the others cost one abandoned launch
        const int cnt_cluster = (Npad + SINGLE_CLUSTER_MAX - 1) / SINGLE_CLUSTER_MAX;
        while (kmax > 48 && homotopy_single_smem<T>(Mp, std::max(cnt, cnt_cluster), kmax, (kmax + 1) | 1, false, true) > 200 * 1024) kmax -= 8;
    }
    const int ks = (kmax + 1) | 1;
    const size_t smem = homotopy_single_smem<T>(Mp, cnt, kmax, ks);
    if (smem > 200 * 1024) return false;
    static std::atomic<uint64_t> configured{ 0 };
    if (smem > 48 * 1024) optin_shared_memory((const void*)homotopy_single_kernel<T, THREADS, false>, configured);
    Lane& L = lane(0);
    cudaStream_t st = L.stream;
    // workspace: y (padded) | P | R | Cg | x | part_t | part_max | part_i | iters | errs
    const size_t nT = (size_t)3 * Mp + Npad + N + 2 * (size_t)grid + 8;
    const size_t pub_off = round_up(nT * sizeof(T) + ((size_t)grid + 8) * sizeof(uint32_t) + 64, 16);   // grid form, Gram iterations: d | x | gam
    single_ws_.reserve(pub_off + (size_t)(kmax + 2) * (2 * sizeof(T) + sizeof(uint32_t)) + 64);
    single_rep_.reserve(64);
    T* w = single_ws_.template as<T>();
    T* d_y = w; T* d_P = d_y + Mp; T* d_R = d_P + Mp; T* d_Cg = d_R + Mp; T* d_x = d_Cg + Npad; T* d_pt = d_x + N; T* d_pm = d_pt + grid;
    double* d_err = reinterpret_cast<double*>(reinterpret_cast<unsigned char*>(w) + round_up((size_t)(d_pm + grid - w) * sizeof(T), 16));
    uint32_t* d_it = reinterpret_cast<uint32_t*>(d_err + 1);
    uint32_t* d_pi = d_it + 2;
    // host pointers: the kernel reads y from, and leaves x and the report in, one pinned buffer -- the call is a launch and a stream
    // synchronise (staged copies of 1-4 KB to and from pageable memory were 2/3 of the 40 us a call cost beside its kernel)
    bool zero_copy = y_on_host && x_on_host && reports_on_host;
    if (const char* e = getenv("SSB200_ZERO_COPY")) if (atoi(e) == 0) zero_copy = false;               // A/B switch
    unsigned char* hp = nullptr;
    if (zero_copy) {
        single_pin_.reserve(64 + ((size_t)Mp + N) * sizeof(T));
        hp = single_pin_.template as<unsigned char>();
        T* h_y = reinterpret_cast<T*>(hp + 64);
        std::memcpy(h_y, y, (size_t)M * sizeof(T));
        for (int m = M; m < Mp; ++m) h_y[m] = T(0);
    } else {
        if (Mp > M) SSB_CUDA(cudaMemsetAsync(d_y + M, 0, (size_t)(Mp - M) * sizeof(T), st));
        SSB_CUDA(cudaMemcpyAsync(d_y, y, (size_t)M * sizeof(T), y_on_host ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice, st));
    }
    SingleArgs<T> a{};
    a.N = N; a.Npad = Npad; a.Mp = Mp; a.Kmax = kmax; a.ks = ks;
    a.At = At_; a.G = G_; a.ldg = Npad;
    a.y = d_y; a.x = x_on_host ? d_x : x; a.P = d_P; a.R = d_R; a.exact = exact_residual() ? 1 : 0; a.Cg = d_Cg; a.part_t = d_pt; a.part_i = d_pi; a.part_max = d_pm;
    a.iters = reports_on_host ? d_it : iters; a.errs = reports_on_host ? d_err : errs;
    if (zero_copy) {
        a.y = reinterpret_cast<T*>(hp + 64); a.ywork = d_y; a.x = d_x; a.x_out = reinterpret_cast<T*>(hp + 64) + Mp;
        a.errs = reinterpret_cast<double*>(hp); a.iters = reinterpret_cast<uint32_t*>(hp + 8);
    }
    a.tol = tol; a.max_iter = max_iter; a.cnt = cnt;
    a.pub_d = reinterpret_cast<T*>(reinterpret_cast<unsigned char*>(w) + pub_off); a.pub_x = a.pub_d + (kmax + 2);
    a.pub_gam = reinterpret_cast<uint32_t*>(a.pub_x + (kmax + 2));
    // Gram iterations as the lockstep loop has them (solve_chunk): the first iterations of a signal are the same arithmetic on both paths
    uint32_t gram_iters = 16;
    if (const char* e = getenv("SSB200_GRAM_ITERS")) gram_iters = (uint32_t)std::max(0, std::min(32, atoi(e)));   // tuning knob
    if (coherent()) gram_iters = 0;
    a.gram_iters = std::min<uint32_t>(gram_iters, (uint32_t)kmax);
    void* params[] = { &a };
    const int span = L.begin(timing, 1);
    nvtx_range nv("ssb200 single-signal solve");
    if (!launch_single_cluster(a, 1, st))
    SSB_CUDA(cudaLaunchCooperativeKernel((const void*)homotopy_single_kernel<T, THREADS, false>, dim3((unsigned)grid), dim3(THREADS), params, smem, st));
    L.end(span);
    L.counters[0]++;
    if (zero_copy) {
        SSB_CUDA(cudaStreamSynchronize(st));
        if (*reinterpret_cast<uint32_t*>(hp + 8) == 0xffffffffu) { merge(L); return false; }   // out of slots: the lockstep path takes it
        std::memcpy(x, hp + 64 + (size_t)Mp * sizeof(T), (size_t)N * sizeof(T));
        *errs = *reinterpret_cast<double*>(hp);
        *iters = *reinterpret_cast<uint32_t*>(hp + 8);
    } else {
    if (x_on_host) SSB_CUDA(cudaMemcpyAsync(x, d_x, (size_t)N * sizeof(T), cudaMemcpyDeviceToHost, st));
    if (reports_on_host) {
        unsigned char* h = single_rep_.template as<unsigned char>();
        SSB_CUDA(cudaMemcpyAsync(h, d_err, 16, cudaMemcpyDeviceToHost, st));
        SSB_CUDA(cudaStreamSynchronize(st));
        if (*reinterpret_cast<uint32_t*>(h + 8) == 0xffffffffu) { merge(L); return false; }
        *errs = *reinterpret_cast<double*>(h);
        *iters = *reinterpret_cast<uint32_t*>(h + 8);
    } else {
        SSB_CUDA(cudaStreamSynchronize(st));
    }
    }
    L.counters[3] += reports_on_host ? *iters : 0;
    merge(L);
    return true;
}

template <typename T>
void Homotopy<T>::solve_batch_dev(const T* dY, size_t batch, size_t ldy, T tol, uint32_t max_iter, T* dX, size_t ldx,
                                  uint32_t* d_iters, double* d_errs, cudaStream_t user_stream)
{
    SSB_REQUIRE(M > 0, "no dictionary");
    SSB_REQUIRE(max_iter > 0, "max_iterations must be positive");   // homotopy-cpu.cpp:199
    if (batch == 0) return;
    SSB_REQUIRE(ldy >= (size_t)M && ldx >= (size_t)N, "leading dimensions too small");
    // inputs produced on the caller's stream must be complete before the lanes read them
    SSB_CUDA(cudaStreamSynchronize(user_stream));
    const int kmax = active_set_capacity(max_iter, N);
    check_capacity(kmax);                                   // before anything is queued
    if (batch == 1 && solve_single(dY, false, tol, max_iter, dX, false, d_iters, d_errs, false)) return;
    size_t chunk; int nlanes;
    plan_batch(batch, kmax, false, chunk, nlanes);
    const size_t nchunks = (batch + chunk - 1) / chunk;
    nlanes = (int)std::min<size_t>(nlanes, nchunks);
    for (int li = 0; li < nlanes; ++li) reserve_workspace(lane(li), chunk, kmax);
    run_lanes(nlanes, nchunks, [&](Lane& L, size_t c) {
        const size_t b0 = c * chunk, rows = std::min(chunk, batch - b0);
        solve_chunk(L, dY + b0 * ldy, rows, ldy, tol, max_iter, dX + b0 * ldx, ldx, d_iters + b0, d_errs + b0);
    });
}

template <typename T>
void Homotopy<T>::solve_batch_host(const T* Y, size_t batch, size_t ldy, T tol, uint32_t max_iter, T* X, size_t ldx,
                                   uint32_t* iters, double* errs, const SparseOut* sp)
{
    SSB_REQUIRE(M > 0, "no dictionary");
    SSB_REQUIRE(max_iter > 0, "max_iterations must be positive");
    if (batch == 0) return;
    SSB_REQUIRE(ldy >= (size_t)M && (sp || ldx >= (size_t)N), "leading dimensions too small");
    const int kmax = active_set_capacity(max_iter, N);
    check_capacity(kmax);                                   // before anything is queued
    SSB_REQUIRE(!sp || (sp->idx && sp->val && sp->count && sp->cap >= (size_t)kmax), "sparse result: cap must be at least min(max_iterations + 1, N)");
    if (!sp && batch == 1 && solve_single(Y, true, tol, max_iter, X, true, iters, errs, true)) return;
    size_t cap; int nlanes;
    plan_batch(batch, kmax, true, cap, nlanes);

    // The batch is cut into tapering pieces dealt to two lanes (host threads with their own stream and workspace): one
    // lane's copy-in, thin tail iterations and launch gaps hide under the other's kernels, and each lane's copy-out runs
    // on its own stream under that lane's next piece (double-buffered staging).  Exposed: the first copy-in and the last
    // two, small, copy-outs.  C2 (tools/e2e_probe.py): one piece 47.9 ms/step, of which 2.5 copy-in and 7 copy-out.
    std::vector<size_t> sizes;
    {
        std::vector<double> taper = { 0.25, 0.25, 0.1875, 0.1875, 0.0625, 0.0625 };
        if (const char* e = getenv("SSB200_TAPER")) {                                                  // tuning knob: "f1,f2,..."
            taper.clear();
            for (const char* q = e; *q;) { char* end; const double f = strtod(q, &end); if (end == q) break; taper.push_back(f); q = *end ? end + 1 : end; }
        }
        size_t left = batch;
        if (const char* e = getenv("SSB200_HOST_PIECES")) {                                            // tuning knob: equal pieces
            const size_t n = (size_t)std::max(1, atoi(e)), per = std::min(cap, round_up((batch + n - 1) / n, 256));
            while (left) { sizes.push_back(std::min(per, left)); left -= sizes.back(); }
        } else if (batch >= 16384) {
            for (double f : taper) {
                const size_t want = std::min(std::min(cap, round_up((size_t)(f * (double)batch), 256)), left);
                if (want) { sizes.push_back(want); left -= want; }
            }
            while (left) { sizes.push_back(std::min(cap, left)); left -= sizes.back(); }
        } else {
            while (left) { sizes.push_back(std::min(cap, left)); left -= sizes.back(); }
        }
    }
    // Sparse return: a solution has at most kmax non-zeros, so when the rows are much wider than that the coefficients cross
    // PCIe compacted (compress_rows_kernel) and are scattered into X on the host, while X itself is zero-filled by helper
    // threads under the first pieces' kernels.  The caller sees the same dense X.
    bool sparse_out = batch >= 4096 && (size_t)N >= 8 * (size_t)kmax;
    if (const char* e = getenv("SSB200_HOST_SPARSE")) sparse_out = sparse_out && atoi(e) != 0;             // tuning knob
    if (sp) sparse_out = true;                               // the caller takes the sparse form itself: no dense rows at all
    const bool dense_rows = !sp;
    std::vector<std::thread> zero_threads;
    std::mutex zero_mu;
    bool zero_joined = false;
    auto ensure_zeroed = [&] {
        std::lock_guard<std::mutex> g(zero_mu);
        if (!zero_joined) { for (auto& t : zero_threads) t.join(); zero_joined = true; }
    };
    if (sparse_out && dense_rows) {
        // helper threads: half the cores this process may run on (its affinity mask: ranks pinned to disjoint core sets each
        // size their own pool), else the box's cores shared between its GPUs (one process per GPU each runs this); 2..8
        int ngpu = 1;
        if (cudaGetDeviceCount(&ngpu) != cudaSuccess || ngpu < 1) { ngpu = 1; cudaGetLastError(); }
        unsigned cores = std::thread::hardware_concurrency();
        bool pinned = false;
        cpu_set_t set;
        if (sched_getaffinity(0, sizeof(set), &set) == 0) {
            const unsigned mine = (unsigned)CPU_COUNT(&set);
            if (mine > 0 && mine < cores) { cores = mine; pinned = true; }
        }
        int nz = (int)std::max(2u, std::min(8u, cores / (2u * (pinned ? 1u : (unsigned)ngpu))));
        if (const char* e = getenv("SSB200_ZERO_THREADS")) nz = std::max(1, std::min(64, atoi(e)));      // tuning knob
        const size_t per = (batch + nz - 1) / nz;
        for (int t = 0; t < nz; ++t) {
            const size_t r0 = std::min(batch, (size_t)t * per), r1 = std::min(batch, r0 + per);
            if (r0 == r1) continue;
            zero_threads.emplace_back([=] {
                if (ldx == (size_t)N) std::memset(X + r0 * ldx, 0, (r1 - r0) * (size_t)N * sizeof(T));
                else for (size_t r = r0; r < r1; ++r) std::memset(X + r * ldx, 0, (size_t)N * sizeof(T));
            });
        }
    }
    struct join_guard { std::function<void()> f; ~join_guard() { f(); } } zero_guard{ [&] {   // also on the error paths
        for (auto& l : lanes_) for (auto& f : l->fin) if (f.valid()) f.wait();
        ensure_zeroed();
    } };
    const size_t big = *std::max_element(sizes.begin(), sizes.end());
    std::vector<size_t> start(sizes.size() + 1, 0);
    for (size_t c = 0; c < sizes.size(); ++c) start[c + 1] = start[c] + sizes[c];
    nlanes = (int)std::min<size_t>(nlanes, sizes.size());
    for (int li = 0; li < nlanes; ++li) {
        Lane& L = lane(li);
        reserve_workspace(L, big, kmax);
        if (!L.copy_out) {
            SSB_CUDA(cudaStreamCreateWithFlags(&L.copy_out, cudaStreamNonBlocking));
            for (int i = 0; i < 2; ++i) SSB_CUDA(cudaEventCreateWithFlags(&L.ev_out[i], cudaEventDisableTiming));
        }
        for (int i = 0; i < 2; ++i) {
            L.stY[i].reserve(big * M * sizeof(T));
            L.stX[i].reserve(big * N * sizeof(T));
            L.stIt[i].reserve(big * sizeof(uint32_t));
            L.stErr[i].reserve(big * sizeof(double));
            if (sparse_out) {
                L.stIdx[i].reserve(big * kmax * sizeof(int)); L.stVal[i].reserve(big * kmax * sizeof(T)); L.stCnt[i].reserve(big * sizeof(int));
                if (dense_rows) { L.hIdx[i].reserve(big * kmax * sizeof(int)); L.hVal[i].reserve(big * kmax * sizeof(T)); L.hCnt[i].reserve(big * sizeof(int)); }
            }
        }
        L.toggle = 0;
    }
    const bool nocopy = getenv("SSB200_HOST_NOCOPY") != nullptr;                                       // timing experiment
    run_lanes(nlanes, sizes.size(), [&](Lane& L, size_t c) {
        const int b = L.toggle; L.toggle ^= 1;
        const size_t b0 = start[c], rows = sizes[c];
        if (L.fin[b].valid()) L.fin[b].get();                      // ... and its host-side scatter is done with the pinned pair
        SSB_CUDA(cudaStreamWaitEvent(L.stream, L.ev_out[b], 0));   // the piece before last has left this staging pair
        SSB_CUDA(cudaMemcpy2DAsync(L.stY[b].p, M * sizeof(T), Y + b0 * ldy, ldy * sizeof(T), M * sizeof(T), rows,
                                   cudaMemcpyHostToDevice, L.stream));
        h2d_bytes += rows * (size_t)M * sizeof(T);
        d2h_bytes += rows * (sizeof(uint32_t) + sizeof(double)) +
                     (sparse_out ? rows * ((size_t)kmax * (sizeof(int) + sizeof(T)) + sizeof(int)) : rows * (size_t)N * sizeof(T));
        solve_chunk(L, L.stY[b].template as<T>(), rows, M, tol, max_iter, L.stX[b].template as<T>(), N,
                    L.stIt[b].template as<uint32_t>(), L.stErr[b].template as<double>());   // returns with the stream drained
        if (nocopy) return;
        if (sparse_out) {
            int* d_idx = L.stIdx[b].template as<int>(); T* d_val = L.stVal[b].template as<T>(); int* d_cnt = L.stCnt[b].template as<int>();
            compress_rows_kernel<T><<<(unsigned)((rows + 3) / 4), 128, 0, L.copy_out>>>(L.stX[b].template as<T>(), (size_t)N, N, rows, kmax,
                                                                                      d_idx, d_val, d_cnt);
            SSB_CUDA(cudaGetLastError());
            L.counters[0]++;
            if (!dense_rows) {   // straight into the caller's arrays (row stride cap >= kmax): nothing to zero, nothing to scatter
                SSB_CUDA(cudaMemcpyAsync(sp->count + b0, d_cnt, rows * sizeof(int), cudaMemcpyDeviceToHost, L.copy_out));
                SSB_CUDA(cudaMemcpy2DAsync(sp->idx + b0 * sp->cap, sp->cap * sizeof(int), d_idx, kmax * sizeof(int), kmax * sizeof(int), rows,
                                           cudaMemcpyDeviceToHost, L.copy_out));
                SSB_CUDA(cudaMemcpy2DAsync(sp->val + b0 * sp->cap, sp->cap * sizeof(T), d_val, kmax * sizeof(T), kmax * sizeof(T), rows,
                                           cudaMemcpyDeviceToHost, L.copy_out));
                SSB_CUDA(cudaMemcpyAsync(iters + b0, L.stIt[b].p, rows * sizeof(uint32_t), cudaMemcpyDeviceToHost, L.copy_out));
                SSB_CUDA(cudaMemcpyAsync(errs + b0, L.stErr[b].p, rows * sizeof(double), cudaMemcpyDeviceToHost, L.copy_out));
                SSB_CUDA(cudaEventRecord(L.ev_out[b], L.copy_out));
                return;
            }
            int* h_idx = L.hIdx[b].template as<int>();  T* h_val = L.hVal[b].template as<T>();  int* h_cnt = L.hCnt[b].template as<int>();
            SSB_CUDA(cudaMemcpyAsync(h_cnt, d_cnt, rows * sizeof(int), cudaMemcpyDeviceToHost, L.copy_out));
            SSB_CUDA(cudaMemcpyAsync(h_idx, d_idx, rows * kmax * sizeof(int), cudaMemcpyDeviceToHost, L.copy_out));
            SSB_CUDA(cudaMemcpyAsync(h_val, d_val, rows * kmax * sizeof(T), cudaMemcpyDeviceToHost, L.copy_out));
            SSB_CUDA(cudaMemcpyAsync(iters + b0, L.stIt[b].p, rows * sizeof(uint32_t), cudaMemcpyDeviceToHost, L.copy_out));
            SSB_CUDA(cudaMemcpyAsync(errs + b0, L.stErr[b].p, rows * sizeof(double), cudaMemcpyDeviceToHost, L.copy_out));
            SSB_CUDA(cudaEventRecord(L.ev_out[b], L.copy_out));
            cudaEvent_t ev = L.ev_out[b];
            const int dev = device, cap_k = kmax;
            L.fin[b] = std::async(std::launch::async, [=, &ensure_zeroed] {
                cudaSetDevice(dev);
                SSB_CUDA(cudaEventSynchronize(ev));
                ensure_zeroed();
                for (size_t r = 0; r < rows; ++r) {
                    const int n = std::min(h_cnt[r], cap_k);
                    T* xr = X + (b0 + r) * ldx;
                    const int* ir = h_idx + r * cap_k;
                    const T* vr = h_val + r * cap_k;
                    for (int k = 0; k < n; ++k) xr[ir[k]] = vr[k];
                }
            });
            return;
        }
        SSB_CUDA(cudaMemcpy2DAsync(X + b0 * ldx, ldx * sizeof(T), L.stX[b].p, N * sizeof(T), N * sizeof(T), rows,
                                   cudaMemcpyDeviceToHost, L.copy_out));
        SSB_CUDA(cudaMemcpyAsync(iters + b0, L.stIt[b].p, rows * sizeof(uint32_t), cudaMemcpyDeviceToHost, L.copy_out));
        SSB_CUDA(cudaMemcpyAsync(errs + b0, L.stErr[b].p, rows * sizeof(double), cudaMemcpyDeviceToHost, L.copy_out));
        SSB_CUDA(cudaEventRecord(L.ev_out[b], L.copy_out));
    });
    for (int li = 0; li < nlanes; ++li) SSB_CUDA(cudaStreamSynchronize(lane(li).copy_out));
    for (int li = 0; li < nlanes; ++li) for (auto& f : lane(li).fin) if (f.valid()) f.get();
}

template <typename T>
double Homotopy<T>::debug_contract(const T* P, size_t rows, T* Q, int path, int reps)
{
    SSB_REQUIRE(M > 0 && rows > 0 && reps > 0, "bad arguments");
    Lane& L = lane(0);
    reserve_workspace(L, rows, 2);
    const bool tc = path == 2 || path == 3;
    SSB_REQUIRE(!tc || (sizeof(T) == 4 && tc_supported(Npad, Mp)), "tcgen05 contraction unavailable for this dtype/shape");
    struct restore { int& ref; int v; ~restore() { ref = v; } } guard{ contraction_path, contraction_path };
    contraction_path = path;                                 // 2: fp16 containers, 3: tf32 containers
    const int split = tc ? (use_f16() ? 2 : 1) : 0;
    dev_buffer dP;
    dP.reserve(rows * M * sizeof(T));
    SSB_CUDA(cudaMemcpyAsync(dP.p, P, rows * M * sizeof(T), cudaMemcpyHostToDevice, L.stream));
    {
        void* hi = split == 2 ? L.Plo[0].p : L.P[0].p;
        void* lo = split == 2 ? static_cast<void*>(L.Plo[0].template as<__half>() + L.ws_rows * Mp) : split == 1 ? L.Plo[0].p : nullptr;
        const int blocks = (int)std::min<size_t>(8192, (rows + 7) / 8);
        stage_signals_kernel<T><<<blocks, 256, 0, L.stream>>>(dP.as<T>(), (size_t)M, M, (int)rows, hi, lo,
                                                             L.Pinv[0].template as<float>(), Mp, split, nullptr);
    }
    SSB_CUDA(cudaGetLastError());
    contract(L, tc, 0, rows, nullptr, true);   // warm-up + the result
    cudaEvent_t e0 = L.event(), e1 = L.event();
    SSB_CUDA(cudaEventRecord(e0, L.stream));
    for (int i = 0; i < reps; ++i) contract(L, tc, 0, rows, nullptr, true);
    SSB_CUDA(cudaEventRecord(e1, L.stream));
    SSB_CUDA(cudaMemcpy2DAsync(Q, N * sizeof(T), L.Q.p, Npad * sizeof(T), N * sizeof(T), rows, cudaMemcpyDeviceToHost, L.stream));
    SSB_CUDA(cudaStreamSynchronize(L.stream));
    float t = 0.f;
    SSB_CUDA(cudaEventElapsedTime(&t, e0, e1));
    L.pool.push_back(e0); L.pool.push_back(e1);
    merge(L);
    return (double)t / reps;
}

template <typename T>
void Homotopy<T>::reconstruct_host(const T* X, size_t batch, size_t ldx, T* Y, size_t ldy)
{
    SSB_REQUIRE(M > 0, "no dictionary");
    if (batch == 0) return;
    dev_buffer dX, dY;
    dX.reserve(batch * N * sizeof(T));
    dY.reserve(batch * M * sizeof(T));
    SSB_CUDA(cudaMemcpy2DAsync(dX.p, N * sizeof(T), X, ldx * sizeof(T), N * sizeof(T), batch, cudaMemcpyHostToDevice, stream));
    reconstruct_kernel<T><<<(unsigned)batch, 128, 0, stream>>>(At_, Mp, M, N, dX.as<T>(), N, dY.as<T>(), M);
    SSB_CUDA(cudaGetLastError());
    counters[0]++;
    SSB_CUDA(cudaMemcpy2DAsync(Y, ldy * sizeof(T), dY.p, M * sizeof(T), M * sizeof(T), batch, cudaMemcpyDeviceToHost, stream));
    SSB_CUDA(cudaStreamSynchronize(stream));
}

template <typename T>
void Homotopy<T>::reconstruct_dev(const T* dX, size_t batch, size_t ldx, T* dY, size_t ldy, cudaStream_t st)
{
    SSB_REQUIRE(M > 0, "no dictionary");
    SSB_REQUIRE(ldx >= (size_t)N && ldy >= (size_t)M, "leading dimension too small");
    if (batch == 0) return;
    reconstruct_kernel<T><<<(unsigned)batch, 128, 0, st>>>(At_, Mp, M, N, dX, ldx, dY, ldy);
    SSB_CUDA(cudaGetLastError());
    counters[0]++;
    SSB_CUDA(cudaStreamSynchronize(st));
}

}  // namespace ssb
