#include "kernels.h"
#include "scl_f32.cuh"

namespace polar {
DecodeKernel scl_f32_kernel_for(int L)
{
    switch (L) {
    case 1: return scl_f32_kernel<1>;
    case 2: return scl_f32_kernel<2>;
    case 4: return scl_f32_kernel<4>;
    case 8: return scl_f32_kernel<8>;
    case 16: return scl_f32_kernel<16>;
    case 32: return scl_f32_kernel<32>;
    }
    return nullptr;
}
} // namespace polar
