// Pieces shared by the decoder kernels: LLR quantiser pre-pass, warp-cooperative re-encoding,
// information-bit gather with fused CRC syndrome, output store.
#pragma once
#include "decode_params.h"
#include "primitives.cuh"

namespace polar {

// fp32 LLR -> int8 like CharContainer::insertLlr (bitcontainer.cpp:468-516).
//   N >= 32: cvtps_epi32 (round to nearest even; NaN and x >= 2^31 give INT_MIN, the "integer indefinite")
//            followed by two signed-saturating packs  => clamp to [-128,127], indefinite -> -128
//   N <  32: max_ps/min_ps clamp first (NaN -> -128), then round to nearest even
POLAR_DEV uint32_t quantize_llr4(float4 v, bool largeRule)
{
    const float f[4] = { v.x, v.y, v.z, v.w };
    uint32_t w = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        int q;
        if (f[k] != f[k]) q = -128;
        else if (largeRule && f[k] >= 2147483648.f) q = -128;
        else q = __float2int_rn(fminf(fmaxf(f[k], -128.f), 127.f));
        w |= (uint32_t)(uint8_t)(int8_t)q << (8 * k);
    }
    return w;
}

__global__ void __launch_bounds__(256) quantize_llr_kernel(const float4* __restrict__ in, uint32_t* __restrict__ out,
                                                           long long nWords, int largeRule)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nWords; i += stride)
        out[i] = quantize_llr4(__ldg(in + i), largeRule != 0);
}

// in-place x <- x * F^{(x)n} on a packed (LSB-first) codeword estimate of N bits held in shared memory
// (re-encoding of the non-systematic mode: fastssc_fip_char.cpp:619-622, butterfly_fip.cpp:31-63)
POLAR_DEV void warp_polar_transform(uint32_t* x, int N, int lane)
{
    const int W = N >> 5;
    for (int w = lane; w < max(W, 1); w += 32) {
        uint32_t v = x[w];
        if (N < 32) v &= (1u << N) - 1u; // bits beyond a short block are not ours
        v ^= (v >> 1) & 0x55555555u;
        v ^= (v >> 2) & 0x33333333u;
        v ^= (v >> 4) & 0x0F0F0F0Fu;
        v ^= (v >> 8) & 0x00FF00FFu;
        v ^= (v >> 16) & 0x0000FFFFu;
        x[w] = v;
    }
    __syncwarp();
    for (int d = 1; d < W; d <<= 1) {
        for (int w = lane; w < W; w += 32)
            if (!(w & d)) x[w] ^= x[w + d];
        __syncwarp();
    }
}

// Gather the K information bits of a codeword estimate (packed LSB-first in `cw`), pack them MSB-first
// into outw[] (one 32-bit word per 32 info bits, already in memory byte order) and return the CRC
// syndrome (0 = detector passes).  bitcontainer.cpp:536-564 + Detector::check.
// The CRC is affine over GF(2): S(info) = S0 ^ XOR_k bit_k * T[k] (tables from crc_host.cpp).
POLAR_DEV uint32_t gather_info_and_syndrome(const DecodeParams& p, const uint32_t* cw, uint32_t* outw, int lane)
{
    uint32_t syn = 0;
    const int rounds = (p.K + 31) >> 5;
    for (int r = 0; r < rounds; ++r) {
        const int k = 32 * r + lane;
        uint32_t bit = 0;
        if (k < p.K) {
            const uint32_t pos = __ldg(p.infoPos + k);
            bit = (cw[pos >> 5] >> (pos & 31u)) & 1u;
            if (bit && p.crcBits) syn ^= __ldg(p.crcTab + k);
        }
        const uint32_t word = __ballot_sync(FULL_MASK, bit);
        if (lane == 0) outw[r] = __byte_perm(__brev(word), 0u, 0x0123u);
    }
    syn = __reduce_xor_sync(FULL_MASK, syn);
    __syncwarp();
    return p.crcBits ? (syn ^ p.crcInit) : 0u;
}

POLAR_DEV void store_info(const DecodeParams& p, long long frame, const uint32_t* outw, int lane)
{
    uint8_t* dst = p.out + frame * p.kb;
    if ((p.kb & 3) == 0) {
        uint32_t* d4 = reinterpret_cast<uint32_t*>(dst);
        for (int i = lane; i < (p.kb >> 2); i += 32) d4[i] = outw[i];
    } else {
        const uint8_t* s = reinterpret_cast<const uint8_t*>(outw);
        for (int i = lane; i < p.kb; i += 32) dst[i] = s[i];
    }
}

} // namespace polar
