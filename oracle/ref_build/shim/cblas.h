/* Declaration-only cblas.h: the 18 cblas symbols the reference late-binds via
 * dlopen/dlsym (src/linalg/blas_wrapper.h:39-56), nothing else.  Written for
 * the oracle build; not reference code. */
#pragma once
#include "openblas_config.h"
#ifdef __cplusplus
extern "C" {
#endif
typedef enum CBLAS_ORDER     {CblasRowMajor=101, CblasColMajor=102} CBLAS_ORDER;
typedef enum CBLAS_TRANSPOSE {CblasNoTrans=111, CblasTrans=112, CblasConjTrans=113, CblasConjNoTrans=114} CBLAS_TRANSPOSE;
typedef enum CBLAS_UPLO      {CblasUpper=121, CblasLower=122} CBLAS_UPLO;
typedef enum CBLAS_DIAG      {CblasNonUnit=131, CblasUnit=132} CBLAS_DIAG;
typedef enum CBLAS_SIDE      {CblasLeft=141, CblasRight=142} CBLAS_SIDE;

#define SS_CBLAS_DECL(P, T) \
    T      cblas_##P##nrm2(blasint, const T*, blasint); \
    T      cblas_##P##dot (blasint, const T*, blasint, const T*, blasint); \
    void   cblas_##P##scal(blasint, T, T*, blasint); \
    size_t cblas_i##P##amax(blasint, const T*, blasint); \
    void   cblas_##P##gemv(enum CBLAS_ORDER, enum CBLAS_TRANSPOSE, blasint, blasint, T, const T*, blasint, const T*, blasint, T, T*, blasint); \
    void   cblas_##P##ger (enum CBLAS_ORDER, blasint, blasint, T, const T*, blasint, const T*, blasint, T*, blasint); \
    void   cblas_##P##gemm(enum CBLAS_ORDER, enum CBLAS_TRANSPOSE, enum CBLAS_TRANSPOSE, blasint, blasint, blasint, T, const T*, blasint, const T*, blasint, T, T*, blasint); \
    void   cblas_##P##trsm(enum CBLAS_ORDER, enum CBLAS_SIDE, enum CBLAS_UPLO, enum CBLAS_TRANSPOSE, enum CBLAS_DIAG, blasint, blasint, T, const T*, blasint, T*, blasint); \
    void   cblas_##P##trsv(enum CBLAS_ORDER, enum CBLAS_UPLO, enum CBLAS_TRANSPOSE, enum CBLAS_DIAG, blasint, const T*, blasint, T*, blasint);
SS_CBLAS_DECL(s, float)
SS_CBLAS_DECL(d, double)
#undef SS_CBLAS_DECL
#ifdef __cplusplus
}
#endif
