/* oracle/_ref C entry points: a thin extern "C" skin over the UNMODIFIED reference
 * (compiled from /root/reference where it lies, see Makefile).  Only the public
 * C++ API is used: ss::homotopy<T> / ss::irls<T> (include/ss/ss.h:27-64) and
 * ss::as_span (include/ss/ndspan.h:105-118).
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under sparse-solvers_b200/ may link or load
 * this library; it exists so tests/, smoke() and bench.py's CPU leg can check
 * and time the reference itself.
 */
#include <ss/ss.h>

#include <cstddef>
#include <cstdint>
#include <cstring>
#include <string>
#include <thread>
#include <vector>
#include <atomic>

namespace {

template <typename T>
ss::ndspan<T, 2> mat_span(const T* A, size_t m, size_t n, ptrdiff_t rs, ptrdiff_t cs)
{
    return ss::as_span<2>(const_cast<T*>(A), std::array<size_t, 2>{ m, n },
                          std::array<size_t, 2>{ size_t(rs), size_t(cs) });
}

/* splits [0, batch) over `threads` host threads; each runs fn(b) */
template <typename F>
void for_each_signal(size_t batch, int threads, F&& fn)
{
    if (threads <= 1 || batch < 2) {
        for (size_t b = 0; b < batch; ++b) fn(b);
        return;
    }
    /* first signal on the calling thread: the reference's lazy BLAS
       singleton is not thread-safe on first use (src/linalg/blas_wrapper.cpp:33-37) */
    fn(0);
    std::atomic<size_t> next{ 1 };
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; ++t)
        pool.emplace_back([&] {
            for (;;) {
                size_t b = next.fetch_add(1);
                if (b >= batch) break;
                fn(b);
            }
        });
    for (auto& th : pool) th.join();
}

template <typename T>
int homotopy_batch(const T* A, size_t m, size_t n, ptrdiff_t rs, ptrdiff_t cs,
                   const T* Y, size_t batch, size_t ldy, T tol, uint32_t max_iter,
                   T* X, size_t ldx, uint32_t* iters, double* errs, int threads)
{
    ss::homotopy<T> solver(mat_span(A, m, n, rs, cs));
    std::atomic<int> failed{ 0 };
    for_each_signal(batch, threads, [&](size_t b) {
        auto r = solver.solve(ss::as_span(Y + b * ldy, m), tol, max_iter,
                              ss::as_span(X + b * ldx, n));
        if (r.template is<ss::homotopy_report>()) {
            auto rep = r.template get<ss::homotopy_report>();
            iters[b] = rep.iter;
            errs[b]  = rep.solution_error;
        } else {
            failed = 1;
        }
    });
    return failed.load();
}

template <typename T>
int irls_batch(const T* A, size_t m, size_t n, ptrdiff_t rs, ptrdiff_t cs,
               const T* Y, size_t batch, size_t ldy, T tol, uint32_t max_iter,
               T* X, size_t ldx, uint32_t* iters, double* errs, uint8_t* spd, int threads)
{
    ss::irls<T> solver(mat_span(A, m, n, rs, cs));
    std::atomic<int> failed{ 0 };
    for_each_signal(batch, threads, [&](size_t b) {
        auto r = solver.solve(ss::as_span(Y + b * ldy, m), tol, max_iter,
                              ss::as_span(X + b * ldx, n));
        if (r.template is<ss::irls_report>()) {
            auto rep = r.template get<ss::irls_report>();
            iters[b] = rep.iter;
            errs[b]  = rep.solution_error;
            spd[b]   = rep.spd_failure ? 1 : 0;
        } else {
            failed = 1;
        }
    });
    return failed.load();
}

} // namespace

extern "C" {

int ssref_homotopy_f32(const float* A, size_t m, size_t n, ptrdiff_t rs, ptrdiff_t cs,
                       const float* Y, size_t batch, size_t ldy, float tol, uint32_t max_iter,
                       float* X, size_t ldx, uint32_t* iters, double* errs, int threads)
{ return homotopy_batch<float>(A, m, n, rs, cs, Y, batch, ldy, tol, max_iter, X, ldx, iters, errs, threads); }

int ssref_homotopy_f64(const double* A, size_t m, size_t n, ptrdiff_t rs, ptrdiff_t cs,
                       const double* Y, size_t batch, size_t ldy, double tol, uint32_t max_iter,
                       double* X, size_t ldx, uint32_t* iters, double* errs, int threads)
{ return homotopy_batch<double>(A, m, n, rs, cs, Y, batch, ldy, tol, max_iter, X, ldx, iters, errs, threads); }

int ssref_irls_f32(const float* A, size_t m, size_t n, ptrdiff_t rs, ptrdiff_t cs,
                   const float* Y, size_t batch, size_t ldy, float tol, uint32_t max_iter,
                   float* X, size_t ldx, uint32_t* iters, double* errs, uint8_t* spd, int threads)
{ return irls_batch<float>(A, m, n, rs, cs, Y, batch, ldy, tol, max_iter, X, ldx, iters, errs, spd, threads); }

int ssref_irls_f64(const double* A, size_t m, size_t n, ptrdiff_t rs, ptrdiff_t cs,
                   const double* Y, size_t batch, size_t ldy, double tol, uint32_t max_iter,
                   double* X, size_t ldx, uint32_t* iters, double* errs, uint8_t* spd, int threads)
{ return irls_batch<double>(A, m, n, rs, cs, Y, batch, ldy, tol, max_iter, X, ldx, iters, errs, spd, threads); }

/* utilities of the same public header (include/ss/ss.h:79-93) */
void ssref_reconstruct_f32(const float* A, size_t m, size_t n, const float* x, float* y)
{ ss::reconstruct_signal(ss::as_span<2>(A, std::array<size_t,2>{ m, n }), ss::as_span(x, n), ss::as_span(y, m)); }
void ssref_reconstruct_f64(const double* A, size_t m, size_t n, const double* x, double* y)
{ ss::reconstruct_signal(ss::as_span<2>(A, std::array<size_t,2>{ m, n }), ss::as_span(x, n), ss::as_span(y, m)); }
void ssref_norm_l1_f32(float* A, size_t m, size_t n)  { ss::norm_l1(ss::as_span<2>(A, std::array<size_t,2>{ m, n })); }
void ssref_norm_l1_f64(double* A, size_t m, size_t n) { ss::norm_l1(ss::as_span<2>(A, std::array<size_t,2>{ m, n })); }

} // extern "C"
