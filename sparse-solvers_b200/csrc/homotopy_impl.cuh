// Member definitions of ssb::Homotopy<T> (included by homotopy.cuh).
#pragma once

namespace ssb {

template <typename T>
void Homotopy<T>::reserve_workspace(size_t rows, int kmax)
{
    const size_t rows_alloc = round_up(rows, 256);
    if (rows_alloc <= ws_rows_ && kmax <= ws_kmax_) return;
    const size_t r = std::max(rows_alloc, ws_rows_);
    const int k = std::max(kmax, ws_kmax_);
    drop_plans();
    C_.reserve(r * Npad * sizeof(T));
    Q_.reserve(r * Npad * sizeof(T));
    for (int i = 0; i < 2; ++i) {
        const bool grew = P_[i].bytes < r * Mp * sizeof(T);
        P_[i].reserve(r * Mp * sizeof(T));
        if (grew) SSB_CUDA(cudaMemsetAsync(P_[i].p, 0, P_[i].bytes, stream));   // rows past `count` feed the tensor tiles
        if (sizeof(T) == 4) {
            const bool grew_lo = Plo_[i].bytes < r * Mp * sizeof(T);
            Plo_[i].reserve(r * Mp * sizeof(T));
            if (grew_lo) SSB_CUDA(cudaMemsetAsync(Plo_[i].p, 0, Plo_[i].bytes, stream));
        }
        act_[i].reserve(r * sizeof(int));
    }
    gam_.reserve(r * k * sizeof(uint32_t));
    dG_.reserve(r * k * sizeof(T));
    Ginv_.reserve(r * (size_t)k * k * sizeof(T));
    lam_.reserve(r * sizeof(T));
    K_.reserve(r * sizeof(int));
    counts_.reserve(2 * sizeof(int));
    h_count_.reserve(sizeof(int));
    SSB_CUDA(cudaStreamSynchronize(stream));
    ws_rows_ = r;
    ws_kmax_ = k;
}

// Q[r][:] = A^T P_[which][r][:] for r < rows
template <typename T>
void Homotopy<T>::contract(bool tc, int which, size_t rows, cudaStream_t st)
{
    const int span = begin_span(0, st);
    if constexpr (sizeof(T) == 4) {
        if (tc) {
            if (!plan_[which])
                plan_[which] = tc_plan_create(At_hi_.as<float>(), At_lo_.as<float>(), Npad, Mp,
                                              P_[which].template as<float>(), Plo_[which].template as<float>(), (int)ws_rows_);
            tc_contract(plan_[which], (int)rows, Q_.as<float>(), Npad, st);
            counters[0]++;
        }
    }
    if (!tc)
        counters[0] += launch_contract_simt<T>(At_.as<T>(), Mp, Npad, Mp, P_[which].template as<T>(), Mp, (int)rows,
                                               Q_.as<T>(), Npad, st);
    end_span(span, st);
    counters[1]++;
    counters[2] += rows;
}

template <typename T, int THREADS, int NCH, bool GS>
static void launch_step_variant(unsigned grid, size_t smem, cudaStream_t st, const HomArgs<T>& a, int kb, int ks)
{
    static size_t configured = 0;   // per process; one device per process in every supported deployment
    if (smem > 48 * 1024 && smem > configured) {
        SSB_CUDA(cudaFuncSetAttribute(homotopy_step_kernel<T, THREADS, NCH, GS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        configured = 200 * 1024;
    }
    homotopy_step_kernel<T, THREADS, NCH, GS><<<grid, THREADS, smem, st>>>(a, kb, ks);
    SSB_CUDA(cudaGetLastError());
}

template <typename T>
static void launch_step(int threads, int nch, bool gs, unsigned grid, size_t smem, cudaStream_t st, const HomArgs<T>& a, int kb, int ks)
{
#define SSB_STEP(TH, NC)                                                                          \
    if (threads == TH && nch == NC) {                                                             \
        if (gs) launch_step_variant<T, TH, NC, true>(grid, smem, st, a, kb, ks);                  \
        else    launch_step_variant<T, TH, NC, false>(grid, smem, st, a, kb, ks);                 \
        return;                                                                                   \
    }
    SSB_STEP(128, 1) SSB_STEP(128, 2) SSB_STEP(128, 4) SSB_STEP(256, 4) SSB_STEP(256, 0)
#undef SSB_STEP
    throw invalid_argument{ "no step-kernel variant for this shape" };
}

template <typename T>
void Homotopy<T>::solve_chunk(const T* dY, size_t rows, size_t ldy, T tol, uint32_t max_iter, int kmax, T* dX,
                              size_t ldx, uint32_t* d_iters, double* d_errs, cudaStream_t st)
{
    const bool tc = use_tc(rows);
    HomArgs<T> a{};
    a.N = N; a.ldc = Npad; a.Mp = Mp; a.ldp = Mp; a.Kmax = ws_kmax_;
    a.At = At_.as<T>(); a.G = G_.as<T>(); a.ldg = Npad;
    a.C = C_.as<T>(); a.Q = Q_.as<T>(); a.X = dX; a.ldx = ldx;
    a.gam = gam_.as<uint32_t>(); a.dG = dG_.as<T>(); a.Ginv = Ginv_.as<T>(); a.lam = lam_.as<T>(); a.Kcur = K_.as<int>();
    a.iters = d_iters; a.errs = d_errs;
    a.split = tc ? 1 : 0;
    a.tol = tol; a.max_iter = max_iter;
    (void)kmax;

    // c0 = A^T y for the whole chunk (homotopy-cpu.cpp:215)
    {
        const int blocks = (int)std::min<size_t>(4096, (rows * (Mp / (16 / sizeof(T))) + 255) / 256);
        stage_signals_kernel<T><<<blocks, 256, 0, st>>>(dY, ldy, M, (int)rows, P_[0].template as<T>(),
                                                       tc ? Plo_[0].template as<T>() : nullptr, Mp, a.split);
        SSB_CUDA(cudaGetLastError());
        counters[0]++;
    }
    contract(tc, 0, rows, st);
    {
        a.act = nullptr; a.act_next = nullptr; a.next_count = nullptr;
        a.Pn = P_[1].template as<T>(); a.Pn_lo = tc ? Plo_[1].template as<T>() : nullptr;
        a.it = 0;
        const int span = begin_span(1, st);
        homotopy_init_kernel<T, STEP_THREADS><<<(unsigned)rows, STEP_THREADS, 0, st>>>(a);
        SSB_CUDA(cudaGetLastError());
        end_span(span, st);
        counters[0]++;
    }

    // step-kernel variant: c,q register-resident when the dictionary is narrow enough, inverse on chip when it fits
    constexpr int V = 16 / sizeof(T);
    const int chunks128 = (Npad + 128 * V - 1) / (128 * V), chunks256 = (Npad + 256 * V - 1) / (256 * V);
    const int threads = chunks128 <= 4 ? 128 : 256;
    const int nch = chunks128 <= 1 ? 1 : chunks128 <= 2 ? 2 : chunks128 <= 4 ? 4 : chunks256 <= 4 ? 4 : 0;
    const size_t g_budget = 96 * 1024;
    size_t count = rows;
    int cur = 1;                 // P_[cur] holds the right-hand sides of this iteration
    int* d_counts = counts_.as<int>();
    int* h_count = h_count_.as<int>();
    for (uint32_t it = 1; it <= max_iter && count > 0; ++it) {
        contract(tc, cur, count, st);
        int* next_count = d_counts + (it & 1);
        SSB_CUDA(cudaMemsetAsync(next_count, 0, sizeof(int), st));
        a.act = (it == 1) ? nullptr : act_[cur].template as<int>();
        a.act_next = act_[cur ^ 1].template as<int>();
        a.next_count = next_count;
        a.Pn = P_[cur ^ 1].template as<T>();
        a.Pn_lo = tc ? Plo_[cur ^ 1].template as<T>() : nullptr;
        a.it = it;
        const int kb = (int)std::min<uint32_t>(it, (uint32_t)ws_kmax_);      // K <= kb before iteration `it`
        const int ks = (kb + 1) | 1;                                          // odd row stride: conflict-free column reads
        const bool gs = (size_t)(kb + 1) * ks * sizeof(T) <= g_budget;
        const size_t smem = homotopy_step_smem<T>(ws_kmax_, Npad, gs ? kb + 1 : 0, ks);
        SSB_REQUIRE(smem <= 200 * 1024, "max_iterations too large for the step kernel's shared memory");
        const int span = begin_span(1, st);
        launch_step(threads, nch, gs, (unsigned)count, smem, st, a, kb, ks);
        end_span(span, st);
        counters[0]++;
        counters[3]++;
        SSB_CUDA(cudaMemcpyAsync(h_count, next_count, sizeof(int), cudaMemcpyDeviceToHost, st));
        SSB_CUDA(cudaStreamSynchronize(st));
        count = (size_t)*h_count;
        cur ^= 1;
    }
    SSB_CUDA(cudaStreamSynchronize(st));
    collect_spans();
}

template <typename T>
void Homotopy<T>::solve_batch_dev(const T* dY, size_t batch, size_t ldy, T tol, uint32_t max_iter, T* dX, size_t ldx,
                                  uint32_t* d_iters, double* d_errs, cudaStream_t st)
{
    SSB_REQUIRE(M > 0, "no dictionary");
    SSB_REQUIRE(max_iter > 0, "max_iterations must be positive");   // homotopy-cpu.cpp:199
    if (batch == 0) return;
    SSB_REQUIRE(ldy >= (size_t)M && ldx >= (size_t)N, "leading dimensions too small");
    if (!st) st = stream;
    const int kmax = (int)std::min<uint64_t>((uint64_t)max_iter + 1, (uint64_t)N);
    const size_t chunk = choose_chunk(batch, kmax);
    reserve_workspace(chunk, kmax);
    for (size_t b0 = 0; b0 < batch; b0 += chunk) {
        const size_t rows = std::min(chunk, batch - b0);
        solve_chunk(dY + b0 * ldy, rows, ldy, tol, max_iter, kmax, dX + b0 * ldx, ldx, d_iters + b0, d_errs + b0, st);
    }
}

template <typename T>
void Homotopy<T>::solve_batch_host(const T* Y, size_t batch, size_t ldy, T tol, uint32_t max_iter, T* X, size_t ldx,
                                   uint32_t* iters, double* errs)
{
    SSB_REQUIRE(M > 0, "no dictionary");
    SSB_REQUIRE(max_iter > 0, "max_iterations must be positive");
    if (batch == 0) return;
    SSB_REQUIRE(ldy >= (size_t)M && ldx >= (size_t)N, "leading dimensions too small");
    const int kmax = (int)std::min<uint64_t>((uint64_t)max_iter + 1, (uint64_t)N);
    const size_t chunk = choose_chunk(batch, kmax);
    reserve_workspace(chunk, kmax);
    stY_.reserve(chunk * M * sizeof(T));
    stX_.reserve(chunk * N * sizeof(T));
    stIt_.reserve(chunk * sizeof(uint32_t));
    stErr_.reserve(chunk * sizeof(double));
    for (size_t b0 = 0; b0 < batch; b0 += chunk) {
        const size_t rows = std::min(chunk, batch - b0);
        SSB_CUDA(cudaMemcpy2DAsync(stY_.p, M * sizeof(T), Y + b0 * ldy, ldy * sizeof(T), M * sizeof(T), rows,
                                   cudaMemcpyHostToDevice, stream));
        solve_chunk(stY_.as<T>(), rows, M, tol, max_iter, kmax, stX_.as<T>(), N, stIt_.as<uint32_t>(),
                    stErr_.as<double>(), stream);
        SSB_CUDA(cudaMemcpy2DAsync(X + b0 * ldx, ldx * sizeof(T), stX_.p, N * sizeof(T), N * sizeof(T), rows,
                                   cudaMemcpyDeviceToHost, stream));
        SSB_CUDA(cudaMemcpyAsync(iters + b0, stIt_.p, rows * sizeof(uint32_t), cudaMemcpyDeviceToHost, stream));
        SSB_CUDA(cudaMemcpyAsync(errs + b0, stErr_.p, rows * sizeof(double), cudaMemcpyDeviceToHost, stream));
        SSB_CUDA(cudaStreamSynchronize(stream));
    }
}

template <typename T>
double Homotopy<T>::debug_contract(const T* P, size_t rows, T* Q, int path, int reps)
{
    SSB_REQUIRE(M > 0 && rows > 0 && reps > 0, "bad arguments");
    reserve_workspace(rows, 2);
    const bool tc = path == 2;
    SSB_REQUIRE(!tc || (sizeof(T) == 4 && tc_supported(Npad, Mp)), "tcgen05 contraction unavailable for this dtype/shape");
    dev_buffer dP;
    dP.reserve(rows * M * sizeof(T));
    SSB_CUDA(cudaMemcpyAsync(dP.p, P, rows * M * sizeof(T), cudaMemcpyHostToDevice, stream));
    const int blocks = (int)std::min<size_t>(4096, (rows * (Mp / (16 / sizeof(T))) + 255) / 256);
    stage_signals_kernel<T><<<blocks, 256, 0, stream>>>(dP.as<T>(), (size_t)M, M, (int)rows, P_[0].template as<T>(),
                                                       tc ? Plo_[0].template as<T>() : nullptr, Mp, tc ? 1 : 0);
    SSB_CUDA(cudaGetLastError());
    contract(tc, 0, rows, stream);   // warm-up + the result
    cudaEvent_t e0 = get_event(), e1 = get_event();
    SSB_CUDA(cudaEventRecord(e0, stream));
    for (int i = 0; i < reps; ++i) contract(tc, 0, rows, stream);
    SSB_CUDA(cudaEventRecord(e1, stream));
    SSB_CUDA(cudaMemcpy2DAsync(Q, N * sizeof(T), Q_.p, Npad * sizeof(T), N * sizeof(T), rows, cudaMemcpyDeviceToHost, stream));
    SSB_CUDA(cudaStreamSynchronize(stream));
    float ms = 0.f;
    SSB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    event_pool.push_back(e0); event_pool.push_back(e1);
    collect_spans();
    return (double)ms / reps;
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
    reconstruct_kernel<T><<<(unsigned)batch, 128, 0, stream>>>(At_.as<T>(), Mp, M, N, dX.as<T>(), N, dY.as<T>(), M);
    SSB_CUDA(cudaGetLastError());
    counters[0]++;
    SSB_CUDA(cudaMemcpy2DAsync(Y, ldy * sizeof(T), dY.p, M * sizeof(T), M * sizeof(T), batch, cudaMemcpyDeviceToHost, stream));
    SSB_CUDA(cudaStreamSynchronize(stream));
}

}  // namespace ssb
