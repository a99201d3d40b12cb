// Rate matching around the encoder / decoder: batched PolarCode::Puncturer (puncturer.cpp:45-85, puncturer.h:71-100).
// All three are index copies, HBM bound: rows are walked by consecutive threads so that the writes are coalesced and
// the (ascending) gathers touch consecutive lines.
#pragma once
#include <cstdint>

namespace polar {

// puncture<T>: out[f][j] = in[f][pos[j]]           (nPos elements out of rows of `parent`)
template <typename T>
__global__ void __launch_bounds__(256) puncture_kernel(const T* __restrict__ in, const uint32_t* __restrict__ pos, int nPos,
                                                       int parent, long long nFrames, T* __restrict__ out)
{
    const long long total = nFrames * nPos;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const long long f = i / nPos;
        const int j = (int)(i - f * nPos);
        out[i] = in[f * parent + __ldg(pos + j)];
    }
}

// depuncture<T>: out[f][p] = inv[p] < 0 ? 0 : in[f][inv[p]]   (zero fill + scatter of puncturer.h:93-100 as a gather)
template <typename T>
__global__ void __launch_bounds__(256) depuncture_kernel(const T* __restrict__ in, const int32_t* __restrict__ inv, int nPos,
                                                         int parent, long long nFrames, T* __restrict__ out)
{
    const long long total = nFrames * parent;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const long long f = i / parent;
        const int p = (int)(i - f * parent);
        const int j = __ldg(inv + p);
        out[i] = j < 0 ? T(0) : in[f * nPos + j];
    }
}

// int8 rows, four parent positions per thread (one 32-bit store)
__global__ void __launch_bounds__(256) depuncture_i8x4_kernel(const int8_t* __restrict__ in, const int32_t* __restrict__ inv,
                                                              int nPos, int parent, long long nFrames, uint32_t* __restrict__ out)
{
    const int pw = parent >> 2;
    const long long total = nFrames * pw;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const long long f = i / pw;
        const int p = (int)(i - f * pw) << 2;
        const int4 j = __ldg(reinterpret_cast<const int4*>(inv + p));
        const int8_t* row = in + f * nPos;
        uint32_t w = 0;
        if (j.x >= 0) w |= (uint32_t)(uint8_t)row[j.x];
        if (j.y >= 0) w |= (uint32_t)(uint8_t)row[j.y] << 8;
        if (j.z >= 0) w |= (uint32_t)(uint8_t)row[j.z] << 16;
        if (j.w >= 0) w |= (uint32_t)(uint8_t)row[j.w] << 24;
        out[i] = w;
    }
}

// puncturePacked: MSB-first bit j of out[f] = MSB-first bit pos[j] of in[f]; one output byte per thread
__global__ void __launch_bounds__(256) puncture_packed_kernel(const uint8_t* __restrict__ in, const uint32_t* __restrict__ pos,
                                                              int nPos, int parent, long long nFrames, uint8_t* __restrict__ out)
{
    const int ob = nPos >> 3, ib = parent >> 3;
    const long long total = nFrames * ob;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const long long f = i / ob;
        const int b = (int)(i - f * ob);
        const uint8_t* row = in + f * ib;
        const uint4 p0 = __ldg(reinterpret_cast<const uint4*>(pos + 8 * b));
        const uint4 p1 = __ldg(reinterpret_cast<const uint4*>(pos + 8 * b + 4));
        const uint32_t p[8] = { p0.x, p0.y, p0.z, p0.w, p1.x, p1.y, p1.z, p1.w };
        uint32_t v = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) v |= ((row[p[k] >> 3] >> (7 - (p[k] & 7))) & 1u) << (7 - k);
        out[i] = (uint8_t)v;
    }
}

} // namespace polar
