// On-device link simulator: the worker loop of the reference's pcsim (src/simulation/simulator.cpp:632-672) as kernels.
//   generateData (:850-868)  -> sim_payload_kernel      random payload bytes, Philox4x32-10 keyed by (seed, point, frame)
//   encode       (:870-879)  -> the library's encoder kernel (CRC fused)
//   modulate / transmit / demodulate / amplify (:881-931, bpsk.cpp:68-80, awgn.cpp:30-43, scale)
//                             -> sim_channel_kernel      BPSK 0 -> +1, y = x + sigma n, LLR = amplification * y,
//                                                        int8 decoders: the quantiser of CharContainer::insertLlr fused
//   decode       (:919-935)  -> the library's decoder kernel
//   countErrors  (:937-964)  -> sim_count_kernel        runs / block errors / bit errors over all K info bits /
//                                                        reported (detector) errors, 64-bit device counters
// Nothing returns to the host between the stages; the random streams are functions of the absolute frame number, so a
// point gives the same frames whatever the batch size, the number of GPUs or the order of the batches.
#pragma once
#include "decode_common.cuh"

namespace polar {

// Philox4x32-10 (Salmon et al., "Parallel random numbers: as easy as 1, 2, 3", SC'11)
POLAR_DEV uint4 philox4x32_10(uint4 c, uint2 k)
{
    constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(M0, c.x), lo0 = M0 * c.x;
        const uint32_t hi1 = __umulhi(M1, c.z), lo1 = M1 * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += W0;
        k.y += W1;
    }
    return c;
}

enum { SIM_STREAM_PAYLOAD = 0, SIM_STREAM_NOISE = 1 };

struct SimParams {
    uint8_t* info;      // nFrames x kb   (payload; the encoder then writes the CRC trailer in place)
    const uint8_t* code; // nFrames x nb
    void* llr;          // nFrames x N int8 or float
    unsigned long long firstFrame; // absolute number of frame 0 of this batch
    long long nFrames;
    uint32_t seedLo, seedHi; // Philox key: seedHi is mixed with the sweep point
    float sigma, amplification;
    int N, K, kb, nb, llrIsFloat;
};

// 16 payload bytes per thread
__global__ void __launch_bounds__(256) sim_payload_kernel(const SimParams p)
{
    const int chunks = (p.kb + 15) >> 4;
    const long long total = p.nFrames * chunks;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const long long f = i / chunks;
        const int j = (int)(i - f * chunks);
        const unsigned long long frame = p.firstFrame + (unsigned long long)f;
        const uint4 rnd = philox4x32_10(make_uint4((uint32_t)frame, (uint32_t)(frame >> 32), (uint32_t)j, SIM_STREAM_PAYLOAD),
                                        make_uint2(p.seedLo, p.seedHi));
        uint8_t* dst = p.info + f * p.kb + 16 * j;
        const int n = min(16, p.kb - 16 * j);
        const uint32_t w[4] = { rnd.x, rnd.y, rnd.z, rnd.w };
        if (n == 16 && (p.kb & 3) == 0) {
            uint32_t* d4 = reinterpret_cast<uint32_t*>(dst);
#pragma unroll
            for (int t = 0; t < 4; ++t) d4[t] = w[t];
        } else {
            for (int b = 0; b < n; ++b) dst[b] = (uint8_t)(w[b >> 2] >> (8 * (b & 3)));
        }
        if ((p.K & 7) && 16 * j + n == p.kb) dst[n - 1] &= (uint8_t)(0xFF << (8 - (p.K & 7))); // pad bits of the last byte are zero
    }
}

// four code positions per thread: BPSK, AWGN (Box-Muller on two Philox pairs), amplification, quantiser
__global__ void __launch_bounds__(256) sim_channel_kernel(const SimParams p)
{
    const int groups = p.N >> 2, lgGroups = 31 - __clz(groups); // N is a power of two: shifts, not a 64-bit division per thread
    const long long total = p.nFrames * groups;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const long long f = i >> lgGroups;
        const int j = (int)(i & (groups - 1));
        const unsigned long long frame = p.firstFrame + (unsigned long long)f;
        const uint4 rnd = philox4x32_10(make_uint4((uint32_t)frame, (uint32_t)(frame >> 32), (uint32_t)j, SIM_STREAM_NOISE),
                                        make_uint2(p.seedLo, p.seedHi));
        // u in (0, 1]: (x + 1) * 2^-32, computed so that it never rounds to 0
        const float u0 = ((float)(rnd.x >> 8) + 1.0f) * (1.0f / 16777216.0f), u1 = (float)(rnd.y >> 8) * (1.0f / 16777216.0f);
        const float u2 = ((float)(rnd.z >> 8) + 1.0f) * (1.0f / 16777216.0f), u3 = (float)(rnd.w >> 8) * (1.0f / 16777216.0f);
        const float r0 = sqrtf(-2.0f * __logf(u0)), r1 = sqrtf(-2.0f * __logf(u2));
        float s0, c0, s1, c1;
        __sincosf(6.283185307179586f * u1, &s0, &c0);
        __sincosf(6.283185307179586f * u3, &s1, &c1);
        const float nz[4] = { r0 * c0, r0 * s0, r1 * c1, r1 * s1 };
        const uint32_t byte = p.code[f * p.nb + (j >> 1)];
        const uint32_t nib = (j & 1) ? (byte & 15u) : (byte >> 4); // MSB first: position 4j is the top bit of the nibble
        float y[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const float x = ((nib >> (3 - e)) & 1u) ? -1.0f : 1.0f;   // bpsk.cpp:68-80: 0 -> +1
            y[e] = p.amplification * (x + p.sigma * nz[e]);
        }
        if (p.llrIsFloat) {
            reinterpret_cast<float4*>(p.llr)[i] = make_float4(y[0], y[1], y[2], y[3]);
        } else {
            reinterpret_cast<uint32_t*>(p.llr)[i] = quantize_llr4(make_float4(y[0], y[1], y[2], y[3]), p.N >= 32);
        }
    }
}

// counters: [0] runs, [1] block errors, [2] bit errors (all K info bits, CRC included, like simulator.cpp:937-964),
//           [3] reported errors (decode() returned false), [4] undetected errors (wrong block, detector passed)
__global__ void __launch_bounds__(256) sim_count_kernel(const uint8_t* __restrict__ sent, const uint8_t* __restrict__ dec,
                                                        const uint8_t* __restrict__ ok, long long nFrames, int kb,
                                                        unsigned long long* __restrict__ counters)
{
    __shared__ unsigned int sh[4][8];
    unsigned int blk = 0, bits = 0, rep = 0, und = 0;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long f = (long long)blockIdx.x * blockDim.x + threadIdx.x; f < nFrames; f += stride) {
        unsigned int be = 0;
        const uint8_t* a = sent + f * kb;
        const uint8_t* b = dec + f * kb;
        if ((kb & 3) == 0) {
            const uint32_t* a4 = reinterpret_cast<const uint32_t*>(a);
            const uint32_t* b4 = reinterpret_cast<const uint32_t*>(b);
            for (int i = 0; i < (kb >> 2); ++i) be += __popc(a4[i] ^ b4[i]);
        } else {
            for (int i = 0; i < kb; ++i) be += __popc((uint32_t)(a[i] ^ b[i]));
        }
        const bool passed = ok[f] != 0;
        bits += be;
        blk += be ? 1u : 0u;
        rep += passed ? 0u : 1u;
        und += (be && passed) ? 1u : 0u;
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    blk = __reduce_add_sync(FULL_MASK, blk);
    bits = __reduce_add_sync(FULL_MASK, bits);
    rep = __reduce_add_sync(FULL_MASK, rep);
    und = __reduce_add_sync(FULL_MASK, und);
    if (lane == 0) { sh[0][warp] = blk; sh[1][warp] = bits; sh[2][warp] = rep; sh[3][warp] = und; }
    __syncthreads();
    if (threadIdx.x < 4) {
        unsigned long long t = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += sh[threadIdx.x][w];
        if (t) atomicAdd(counters + 1 + threadIdx.x, t);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(counters, (unsigned long long)nFrames);
}

} // namespace polar
