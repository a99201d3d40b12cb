#include "kernels.h"
#include "scl_i8.cuh"

namespace polar {
template <int MAXT>
static DecodeKernel pick(int L)
{
    switch (L) {
    case 1: return scl_i8_kernel<1, MAXT>;
    case 2: return scl_i8_kernel<2, MAXT>;
    case 4: return scl_i8_kernel<4, MAXT>;
    case 8: return scl_i8_kernel<8, MAXT>;
    case 16: return scl_i8_kernel<16, MAXT>;
    case 32: return scl_i8_kernel<32, MAXT>;
    }
    return nullptr;
}
DecodeKernel scl_i8_kernel_for(int L, int maxThreads)
{
    if (maxThreads <= 512) return pick<512>(L);
    if (maxThreads <= POLAR_SCL_MID_THREADS) return pick<POLAR_SCL_MID_THREADS>(L);
    return pick<1024>(L);
}
} // namespace polar
