/* Declaration-only stand-in for OpenBLAS' openblas_config.h (the reference
 * downloads OpenBLAS at configure time, cmake/BlasUtils.cmake:20-50; there is
 * no network here).  LP64 interface: blasint is a plain int. */
#pragma once
#include <stddef.h>
typedef int blasint;
