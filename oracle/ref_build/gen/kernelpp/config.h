/* Hand-written result of configuring third_party/kernelpp/include/kernelpp/config.h.in:8-9
 * (AVX dispatch slot on, CUDA off: the reference never enables it). */
#pragma once
#define kernelpp_WITH_AVX
