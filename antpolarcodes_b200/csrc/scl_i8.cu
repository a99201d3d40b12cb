#include "kernels.h"
#include "scl_i8.cuh"

namespace polar {
DecodeKernel scl_i8_kernel_for(int L)
{
    switch (L) {
    case 1: return scl_i8_kernel<1>;
    case 2: return scl_i8_kernel<2>;
    case 4: return scl_i8_kernel<4>;
    case 8: return scl_i8_kernel<8>;
    case 16: return scl_i8_kernel<16>;
    case 32: return scl_i8_kernel<32>;
    }
    return nullptr;
}
} // namespace polar
