// Batch-level helper kernels of the C ABI (compiled into polar_cuda.cu only): LLR quantiser pre-pass and the
// compaction / gather / scatter steps of the adaptive decoders.
#pragma once
#include "decode_common.cuh"

namespace polar {

__global__ void __launch_bounds__(256) quantize_llr_kernel(const float4* __restrict__ in, uint32_t* __restrict__ out,
                                                           long long nWords, int largeRule)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nWords; i += stride)
        out[i] = quantize_llr4(__ldg(in + i), largeRule != 0);
}

} // namespace polar

// ---------------------------------------------------------------- adaptive decoding (AdaptiveChar / Float / Mixed)
// The reference runs the list decoder only on frames whose Fast-SSC result fails the detector
// (adaptive_char.cpp:34-46).  Batched: compact the failed frame numbers, gather their LLR rows into a dense
// batch for the list kernel, scatter its results back over the Fast-SSC ones.
namespace polar {

// frame numbers with ok == 0, in ascending order inside every warp's 32-frame window (the order across windows is
// that of the atomics; every frame is decoded independently, so it does not show in the results)
__global__ void __launch_bounds__(256) adaptive_compact_kernel(const uint8_t* __restrict__ ok, long long nFrames,
                                                               uint32_t* __restrict__ idx, uint32_t* __restrict__ count)
{
    const int lane = threadIdx.x & 31;
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long nPad = (nFrames + 31) & ~31LL;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nPad; i += stride) {
        const bool bad = i < nFrames && ok[i] == 0;
        const uint32_t m = __ballot_sync(0xFFFFFFFFu, bad);
        if (!m) continue;
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(count, (uint32_t)__popc(m));
        base = __shfl_sync(0xFFFFFFFFu, base, 0);
        if (bad) idx[base + __popc(m & ((1u << lane) - 1u))] = (uint32_t)i;
    }
}

// dst[r][:] = src[idx[r]][:], rows of rowVecs 16-byte (VEC = uint4) or 4-byte (VEC = uint32_t) vectors
template <typename VEC>
__global__ void __launch_bounds__(256) adaptive_gather_kernel(const VEC* __restrict__ src, const uint32_t* __restrict__ idx,
                                                              long long nRows, int rowVecs, VEC* __restrict__ dst)
{
    const long long total = nRows * rowVecs;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const long long r = i / rowVecs;
        const int c = (int)(i - r * rowVecs);
        dst[i] = __ldg(src + (long long)__ldg(idx + r) * rowVecs + c);
    }
}

// out[idx[r]][:] = outC[r][:], ok[idx[r]] = okC[r]   (kb bytes per row)
__global__ void __launch_bounds__(256) adaptive_scatter_kernel(const uint8_t* __restrict__ outC, const uint8_t* __restrict__ okC,
                                                               const uint32_t* __restrict__ idx, long long nRows, int kb,
                                                               uint8_t* __restrict__ out, uint8_t* __restrict__ ok)
{
    const long long total = nRows * kb;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const long long r = i / kb;
        const int c = (int)(i - r * kb);
        const long long f = __ldg(idx + r);
        out[f * kb + c] = outC[i];
        if (c == 0) ok[f] = okC[r];
    }
}

} // namespace polar
