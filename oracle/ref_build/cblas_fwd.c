/* Forwarding shim: the scipy wheel's OpenBLAS (0.3.31.dev, LP64) exports its
 * cblas entry points with a `scipy_` prefix; the reference dlsym()s the plain
 * names (src/linalg/blas_wrapper.h:39-56).  This .so re-exports the 18 symbols.
 * Test infrastructure for the oracle only. */
#include "shim/cblas.h"
#define FWD(P, T) \
    extern T scipy_cblas_##P##nrm2(blasint, const T*, blasint); \
    T cblas_##P##nrm2(blasint n, const T* x, blasint ix) { return scipy_cblas_##P##nrm2(n, x, ix); } \
    extern T scipy_cblas_##P##dot(blasint, const T*, blasint, const T*, blasint); \
    T cblas_##P##dot(blasint n, const T* x, blasint ix, const T* y, blasint iy) { return scipy_cblas_##P##dot(n, x, ix, y, iy); } \
    extern void scipy_cblas_##P##scal(blasint, T, T*, blasint); \
    void cblas_##P##scal(blasint n, T a, T* x, blasint ix) { scipy_cblas_##P##scal(n, a, x, ix); } \
    extern size_t scipy_cblas_i##P##amax(blasint, const T*, blasint); \
    size_t cblas_i##P##amax(blasint n, const T* x, blasint ix) { return scipy_cblas_i##P##amax(n, x, ix); } \
    extern void scipy_cblas_##P##gemv(enum CBLAS_ORDER, enum CBLAS_TRANSPOSE, blasint, blasint, T, const T*, blasint, const T*, blasint, T, T*, blasint); \
    void cblas_##P##gemv(enum CBLAS_ORDER o, enum CBLAS_TRANSPOSE t, blasint m, blasint n, T al, const T* a, blasint lda, const T* x, blasint ix, T be, T* y, blasint iy) \
        { scipy_cblas_##P##gemv(o, t, m, n, al, a, lda, x, ix, be, y, iy); } \
    extern void scipy_cblas_##P##ger(enum CBLAS_ORDER, blasint, blasint, T, const T*, blasint, const T*, blasint, T*, blasint); \
    void cblas_##P##ger(enum CBLAS_ORDER o, blasint m, blasint n, T al, const T* x, blasint ix, const T* y, blasint iy, T* a, blasint lda) \
        { scipy_cblas_##P##ger(o, m, n, al, x, ix, y, iy, a, lda); } \
    extern void scipy_cblas_##P##gemm(enum CBLAS_ORDER, enum CBLAS_TRANSPOSE, enum CBLAS_TRANSPOSE, blasint, blasint, blasint, T, const T*, blasint, const T*, blasint, T, T*, blasint); \
    void cblas_##P##gemm(enum CBLAS_ORDER o, enum CBLAS_TRANSPOSE ta, enum CBLAS_TRANSPOSE tb, blasint m, blasint n, blasint k, T al, const T* a, blasint lda, const T* b, blasint ldb, T be, T* c, blasint ldc) \
        { scipy_cblas_##P##gemm(o, ta, tb, m, n, k, al, a, lda, b, ldb, be, c, ldc); } \
    extern void scipy_cblas_##P##trsm(enum CBLAS_ORDER, enum CBLAS_SIDE, enum CBLAS_UPLO, enum CBLAS_TRANSPOSE, enum CBLAS_DIAG, blasint, blasint, T, const T*, blasint, T*, blasint); \
    void cblas_##P##trsm(enum CBLAS_ORDER o, enum CBLAS_SIDE s, enum CBLAS_UPLO u, enum CBLAS_TRANSPOSE t, enum CBLAS_DIAG d, blasint m, blasint n, T al, const T* a, blasint lda, T* b, blasint ldb) \
        { scipy_cblas_##P##trsm(o, s, u, t, d, m, n, al, a, lda, b, ldb); } \
    extern void scipy_cblas_##P##trsv(enum CBLAS_ORDER, enum CBLAS_UPLO, enum CBLAS_TRANSPOSE, enum CBLAS_DIAG, blasint, const T*, blasint, T*, blasint); \
    void cblas_##P##trsv(enum CBLAS_ORDER o, enum CBLAS_UPLO u, enum CBLAS_TRANSPOSE t, enum CBLAS_DIAG d, blasint n, const T* a, blasint lda, T* x, blasint ix) \
        { scipy_cblas_##P##trsv(o, u, t, d, n, a, lda, x, ix); }
FWD(s, float)
FWD(d, double)

/* thread-count control + version string, passed straight through */
extern void  scipy_openblas_set_num_threads(int);
extern int   scipy_openblas_get_num_threads(void);
extern char* scipy_openblas_get_config(void);
void  ssref_blas_set_num_threads(int n) { scipy_openblas_set_num_threads(n); }
int   ssref_blas_get_num_threads(void)  { return scipy_openblas_get_num_threads(); }
const char* ssref_blas_config(void)     { return scipy_openblas_get_config(); }
