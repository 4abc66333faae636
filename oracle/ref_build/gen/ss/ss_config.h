/* Hand-written result of configuring include/ss/ss_config.h.in:8-19
 * (version from CMakeLists.txt:4).  BLAS_RUNTIME_FILE is passed by the Makefile. */
#pragma once
#define LIB_NAME "sparsesolvers"
#define ss_VERSION_MAJOR 0
#define ss_VERSION_MINOR 8
#define ss_VERSION_PATCH 8
#define BLAS_OpenBLAS
#ifndef BLAS_RUNTIME_FILE
# error "pass -DBLAS_RUNTIME_FILE=\"/abs/path/to/libcblas_fwd.so\""
#endif
#ifndef BLAS_AVX_RUNTIME_FILE
# define BLAS_AVX_RUNTIME_FILE BLAS_RUNTIME_FILE
#endif
