// Decoder kernels, one translation unit per family (scl_i8.cu, scl_f32.cu, fastssc_i8.cu, fastssc_f32.cu): the host
// runtime (polar_cuda.cu) only sees these getters.
#pragma once
#include "decode_params.h"
#include "schedule.h"

// launch bound of the middle register budget of scl_i8_kernel (768 threads = 80 registers; A/B builds override it)
#ifndef POLAR_SCL_MID_THREADS
#define POLAR_SCL_MID_THREADS 768
#endif

// launch bound of the wide Fast-SSC kernels (512 threads = 16 warps at 128 registers; A/B builds override it)
#ifndef POLAR_SC_WIDE_THREADS
#define POLAR_SC_WIDE_THREADS 512
#endif

namespace polar {

typedef void (*DecodeKernel)(const DecodeParams, const Schedule);

DecodeKernel scl_i8_kernel_for(int listSize, int maxThreads); // SclFipChar; maxThreads = 512 / 768 / 1024 (register budget 128 / 80 / 64)
DecodeKernel scl_f32_kernel_for(int listSize);       // SclAvxFloat
DecodeKernel fastssc_i8_kernel_for(bool wide);       // FastSscFipChar   (wide: 16-warp blocks at 128 registers)
DecodeKernel fastssc_f32_kernel_for(bool wide);      // FastSscAvxFloat

} // namespace polar
