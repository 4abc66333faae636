// Drop-in replacement for the policy section of the reference's src/lib.cpp (lines 26-112): the four
// Policy::run overloads, the IRLS state, reconstruct_signal and norm_l1, implemented over the C ABI of
// libssb200 (include/ssb200.h) instead of kernelpp::run<solve_*> + OpenBLAS.  Compiled against the
// reference's UNCHANGED public headers (include/ss/*.h); nothing else of the reference is linked.
//
// The reference's homotopy state is a non-owning view of A (include/ss/policies.h:42) and `run` is static,
// so the device handle lives in a small cache keyed by the view (pointer, shape, strides, dtype) and is
// refreshed when the bytes behind the view change (the reference bench edits A between solves,
// src/solvers/homotopy_bench.cpp:38-48).
#include <ss/ss.h>

#include <cstdint>
#include <cstring>
#include <immintrin.h>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <tuple>

#include "ssb200.h"

namespace {

template <typename R>
kernelpp::maybe<R> fail(int status)
{
    // same wording a kernelpp dispatch failure produced (kernelpp/kernel.h:122-133)
    return kernelpp::maybe<R>{ kernelpp::error{ ssb200_status_string(status) } };
}

template <typename T> struct abi;
template <> struct abi<float> {
    static constexpr auto create = ssb200_homotopy_create_f32;
    static constexpr auto refresh = ssb200_homotopy_refresh_f32;
    static constexpr auto solve = ssb200_homotopy_solve_f32;
    static constexpr auto irls_create = ssb200_irls_create_f32;
    static constexpr auto irls_solve = ssb200_irls_solve_f32;
    static constexpr auto reconstruct = ssb200_reconstruct_batch_f32;
    static constexpr auto norm_l1 = ssb200_norm_l1_f32;
};
template <> struct abi<double> {
    static constexpr auto create = ssb200_homotopy_create_f64;
    static constexpr auto refresh = ssb200_homotopy_refresh_f64;
    static constexpr auto solve = ssb200_homotopy_solve_f64;
    static constexpr auto irls_create = ssb200_irls_create_f64;
    static constexpr auto irls_solve = ssb200_irls_solve_f64;
    static constexpr auto reconstruct = ssb200_reconstruct_batch_f64;
    static constexpr auto norm_l1 = ssb200_norm_l1_f64;
};

template <typename Span> const typename Span::value_type* base(const Span& s) { return s.raw_data() + s.raw_data_offset(); }
template <typename Span> ptrdiff_t stride(const Span& s, size_t d) { return (ptrdiff_t)s.strides()[d]; }

// Hash of the elements the view covers -- computed before EVERY solve (the reference state is a non-owning view and its bench edits
// A in place between solves), so it has to cost a small fraction of a solve: rows that are contiguous in memory go through eight
// independent multiply-xor lanes over 64-bit words (the lanes pipeline: ~10 bytes per cycle, 4 MB in ~0.15 ms), strided views fall
// back to one element at a time.
inline uint64_t mix(uint64_t h, uint64_t w) { h ^= w; h *= 0x9E3779B97F4A7C15ull; return h ^ (h >> 29); }

// AVX2 form of the same idea for long contiguous runs (where the CPU has it: checked at run time): four 256-bit accumulators, per
// 32 bytes one xor with a key, one 32 x 32 -> 64 multiply of each word's halves and two adds (the accumulate step of XXH3) -- the
// stream runs at cache bandwidth instead of at the latency of a 64-bit multiply chain (1 MB: ~85 us -> ~25 us per solve at C1).
// The tail and the fold go through the scalar lanes.
__attribute__((target("avx2"))) inline const unsigned char* hash_stream_avx2(const unsigned char* p, size_t& nbytes, uint64_t (&lane)[8])
{
    if (nbytes < 512) return p;
    const __m256i key[4] = { _mm256_set1_epi64x((long long)0x9E3779B97F4A7C15ull), _mm256_set1_epi64x((long long)0xC2B2AE3D27D4EB4Full),
                             _mm256_set1_epi64x((long long)0x165667B19E3779F9ull), _mm256_set1_epi64x((long long)0x85EBCA77C2B2AE63ull) };
    __m256i acc[4];
    for (int i = 0; i < 4; ++i) acc[i] = _mm256_set_epi64x((long long)lane[(2 * i + 1) & 7], (long long)lane[(2 * i) & 7], (long long)(lane[i] ^ key[i][0]), (long long)lane[7 - i]);
    while (nbytes >= 128) {
        for (int i = 0; i < 4; ++i) {
            const __m256i d = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(p + 32 * i));
            const __m256i dk = _mm256_xor_si256(d, key[i]);
            const __m256i prod = _mm256_mul_epu32(dk, _mm256_shuffle_epi32(dk, 0x31));     // low half x high half of every 64-bit word
            acc[i] = _mm256_add_epi64(acc[i], _mm256_shuffle_epi32(d, 0x4E));              // the word itself, halves swapped
            acc[i] = _mm256_add_epi64(acc[i], prod);
        }
        p += 128; nbytes -= 128;
    }
    alignas(32) uint64_t out[16];
    for (int i = 0; i < 4; ++i) _mm256_store_si256(reinterpret_cast<__m256i*>(out + 4 * i), acc[i]);
    for (int i = 0; i < 16; ++i) { lane[i & 7] ^= out[i]; lane[i & 7] *= 0x9E3779B97F4A7C15ull; lane[i & 7] ^= lane[i & 7] >> 29; }
    return p;
}

inline void hash_row(const unsigned char* p, size_t nbytes, uint64_t (&lane)[8])
{
    static const bool avx2 = __builtin_cpu_supports("avx2");
    if (avx2) p = hash_stream_avx2(p, nbytes, lane);
    while (nbytes >= 64) {
        uint64_t w[8];
        std::memcpy(w, p, 64);
        for (int i = 0; i < 8; ++i) lane[i] = mix(lane[i], w[i]);
        p += 64; nbytes -= 64;
    }
    int i = 0;
    while (nbytes >= 8) { uint64_t w; std::memcpy(&w, p, 8); lane[i] = mix(lane[i], w); ++i; p += 8; nbytes -= 8; }
    if (nbytes) { uint64_t w = 0; std::memcpy(&w, p, nbytes); lane[7] = mix(lane[7], w ^ (uint64_t)nbytes << 56); }
}

template <typename T>
uint64_t view_hash(const ss::ndspan<T, 2>& A)
{
    const T* p = base(A);
    const size_t m = A.shape()[0], n = A.shape()[1];
    const ptrdiff_t rs = stride(A, 0), cs = stride(A, 1);
    uint64_t lane[8];
    for (int i = 0; i < 8; ++i) lane[i] = 1469598103934665603ull + 0x100000001B3ull * (uint64_t)i;
    if (cs == 1 && rs == (ptrdiff_t)n) hash_row(reinterpret_cast<const unsigned char*>(p), m * n * sizeof(T), lane);       // one block
    else if (cs == 1) for (size_t i = 0; i < m; ++i) hash_row(reinterpret_cast<const unsigned char*>(p + (ptrdiff_t)i * rs), n * sizeof(T), lane);
    else if (rs == 1) for (size_t j = 0; j < n; ++j) hash_row(reinterpret_cast<const unsigned char*>(p + (ptrdiff_t)j * cs), m * sizeof(T), lane);
    else
        for (size_t i = 0; i < m; ++i)
            for (size_t j = 0; j < n; ++j) {
                uint64_t bits = 0;
                std::memcpy(&bits, p + (ptrdiff_t)i * rs + (ptrdiff_t)j * cs, sizeof(T));
                lane[j & 7] = mix(lane[j & 7], bits);
            }
    uint64_t h = 0;
    for (int i = 0; i < 8; ++i) h = mix(h, lane[i]);
    return h;
}

struct cached { void* handle = nullptr; uint64_t hash = 0; };
using key_t = std::tuple<const void*, size_t, size_t, ptrdiff_t, ptrdiff_t, int>;
std::map<key_t, cached> g_cache;
std::mutex g_mu;

template <typename T>
int homotopy_handle(const ss::ndspan<T, 2>& A, void** out)
{
    std::lock_guard<std::mutex> g(g_mu);
    const key_t key{ base(A), A.shape()[0], A.shape()[1], stride(A, 0), stride(A, 1), (int)sizeof(T) };
    cached& c = g_cache[key];
    const uint64_t h = view_hash(A);
    if (!c.handle) {
        if (g_cache.size() > 16) {   // keep the cache small: drop everything else
            for (auto it = g_cache.begin(); it != g_cache.end();) {
                if (it->first != key) { ssb200_destroy(it->second.handle); it = g_cache.erase(it); } else ++it;
            }
        }
        int st = abi<T>::create(base(A), A.shape()[0], A.shape()[1], stride(A, 0), stride(A, 1), 0, &g_cache[key].handle);
        if (st) { g_cache.erase(key); return st; }
        g_cache[key].hash = h;
    } else if (c.hash != h) {
        int st = abi<T>::refresh(c.handle, base(A), stride(A, 0), stride(A, 1));
        if (st) return st;
        c.hash = h;
    }
    *out = g_cache[key].handle;
    return 0;
}

template <typename T>
kernelpp::maybe<ss::homotopy_report> run_homotopy(const ss::ndspan<T, 2>& A, const ss::ndspan<T> y, T tol,
                                                  uint32_t max_iter, ss::ndspan<T> x)
{
    void* h = nullptr;
    if (int st = homotopy_handle(A, &h)) return fail<ss::homotopy_report>(st);
    ss::homotopy_report r{};
    if (int st = abi<T>::solve(h, base(y), stride(y, 0), tol, max_iter, const_cast<T*>(base(x)), stride(x, 0),
                               &r.iter, &r.solution_error))
        return fail<ss::homotopy_report>(st);
    return r;
}

struct irls_handle {
    void* h = nullptr;
    int status = 0;
    ~irls_handle() { ssb200_destroy(h); }
};

template <typename T>
std::shared_ptr<irls_handle> make_irls(const ss::ndspan<T, 2> A)
{
    auto p = std::make_shared<irls_handle>();
    p->status = abi<T>::irls_create(base(A), A.shape()[0], A.shape()[1], stride(A, 0), stride(A, 1), 0, &p->h);
    return p;
}

template <typename T>
kernelpp::maybe<ss::irls_report> run_irls(ss::irls_state& s, const ss::ndspan<T> y, T tol, uint32_t max_iter, ss::ndspan<T> x)
{
    auto& p = xtl::any_cast<std::shared_ptr<irls_handle>&>(s.QR);
    if (p->status) return fail<ss::irls_report>(p->status);
    ss::irls_report r{};
    uint8_t spd = 0;
    if (int st = abi<T>::irls_solve(p->h, base(y), stride(y, 0), tol, max_iter, const_cast<T*>(base(x)), stride(x, 0),
                                    &r.iter, &r.solution_error, &spd))
        return fail<ss::irls_report>(st);
    r.spd_failure = spd != 0;
    return r;
}

template <typename T>
void reconstruct(const ss::ndspan<T, 2> A, const ss::ndspan<T> x, ss::ndspan<T> y)
{
    void* h = nullptr;
    if (homotopy_handle(A, &h)) return;
    std::vector<T> xx(x.size()), yy(y.size());
    for (size_t j = 0; j < xx.size(); ++j) xx[j] = base(x)[(ptrdiff_t)j * stride(x, 0)];
    abi<T>::reconstruct(h, xx.data(), 1, xx.size(), yy.data(), yy.size());
    for (size_t i = 0; i < yy.size(); ++i) const_cast<T*>(base(y))[(ptrdiff_t)i * stride(y, 0)] = yy[i];
}

}  // namespace

namespace ss {

kernelpp::maybe<homotopy_report> homotopy_policy::run(state_type<float>& A, const ndspan<float> y, float tol,
                                                      uint32_t max_iter, ndspan<float> x)
{ return run_homotopy<float>(A, y, tol, max_iter, x); }
kernelpp::maybe<homotopy_report> homotopy_policy::run(state_type<double>& A, const ndspan<double> y, double tol,
                                                      uint32_t max_iter, ndspan<double> x)
{ return run_homotopy<double>(A, y, tol, max_iter, x); }
homotopy_policy::homotopy_policy() = default;
homotopy_policy::homotopy_policy(homotopy_policy&&) = default;
homotopy_policy::~homotopy_policy() = default;

irls_state::irls_state(const ndspan<float, 2> A) : QR{ make_irls<float>(A) } {}
irls_state::irls_state(const ndspan<double, 2> A) : QR{ make_irls<double>(A) } {}
irls_state::~irls_state() = default;

kernelpp::maybe<irls_report> irls_policy::run(state_type<float>& s, const ndspan<float> y, float tol, uint32_t max_iter,
                                              ndspan<float> x)
{ return run_irls<float>(s, y, tol, max_iter, x); }
kernelpp::maybe<irls_report> irls_policy::run(state_type<double>& s, const ndspan<double> y, double tol, uint32_t max_iter,
                                              ndspan<double> x)
{ return run_irls<double>(s, y, tol, max_iter, x); }

void reconstruct_signal(const ndspan<float, 2> A, const ndspan<float> x, ndspan<float> y) { reconstruct<float>(A, x, y); }
void reconstruct_signal(const ndspan<double, 2> A, const ndspan<double> x, ndspan<double> y) { reconstruct<double>(A, x, y); }
void norm_l1(ndspan<float, 2> A)
{ abi<float>::norm_l1(const_cast<float*>(base(A)), A.shape()[0], A.shape()[1], stride(A, 0), stride(A, 1), 0); }
void norm_l1(ndspan<double, 2> A)
{ abi<double>::norm_l1(const_cast<double*>(base(A)), A.shape()[0], A.shape()[1], stride(A, 0), stride(A, 1), 0); }

}  // namespace ss
