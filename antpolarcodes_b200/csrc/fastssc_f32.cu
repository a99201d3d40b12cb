#include "kernels.h"
#include "fastssc_f32.cuh"

namespace polar {
DecodeKernel fastssc_f32_kernel_for(bool wide)
{
    return wide ? (DecodeKernel)fastssc_f32_kernel<true> : (DecodeKernel)fastssc_f32_kernel<false>;
}
} // namespace polar
