#include "kernels.h"
#include "fastssc_i8.cuh"

namespace polar {
DecodeKernel fastssc_i8_kernel_for(bool wide)
{
    return wide ? (DecodeKernel)fastssc_i8_kernel<true> : (DecodeKernel)fastssc_i8_kernel<false>;
}
} // namespace polar
