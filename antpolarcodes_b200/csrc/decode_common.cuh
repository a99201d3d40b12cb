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
    if ((p.kb & 15) == 0 && (reinterpret_cast<uintptr_t>(dst) & 15u) == 0) { // 16-byte packed stores (outw is 16-byte aligned)
        const uint4* s16 = reinterpret_cast<const uint4*>(outw);
        uint4* d16 = reinterpret_cast<uint4*>(dst);
        for (int i = lane; i < (p.kb >> 4); i += 32) d16[i] = s16[i];
    } else if ((p.kb & 3) == 0 && (reinterpret_cast<uintptr_t>(dst) & 3u) == 0) {
        uint32_t* d4 = reinterpret_cast<uint32_t*>(dst);
        for (int i = lane; i < (p.kb >> 2); i += 32) d4[i] = outw[i];
    } else {
        const uint8_t* s = reinterpret_cast<const uint8_t*>(outw);
        for (int i = lane; i < p.kb; i += 32) dst[i] = s[i];
    }
}

// ---------------------------------------------------------------- one-frame-per-thread epilogue
// (Fast-SSC kernels: the lane's packed codeword estimate sits in rows bitRow.. of the warp's [row][lane] stack)
#define TF_ROW(w) sm[((bitRow + (w)) << 5) + lane]
// x <- x * F^{(x)n} on the lane's packed codeword estimate (non-systematic mode:
// fastssc_fip_char.cpp:619-622, butterfly_fip.cpp:31-63)
POLAR_DEV void thread_polar_transform(uint32_t* sm, int lane, int bitRow, int N)
{
    const int W = max(N >> 5, 1);
    for (int w = 0; w < W; ++w) {
        uint32_t v = TF_ROW(w);
        if (N < 32) v &= (1u << N) - 1u;
        v ^= (v >> 1) & 0x55555555u;
        v ^= (v >> 2) & 0x33333333u;
        v ^= (v >> 4) & 0x0F0F0F0Fu;
        v ^= (v >> 8) & 0x00FF00FFu;
        v ^= (v >> 16) & 0x0000FFFFu;
        TF_ROW(w) = v;
    }
    for (int d = 1; d < W; d <<= 1)
        for (int w = 0; w < W; ++w)
            if (!(w & d)) TF_ROW(w) ^= TF_ROW(w + d);
}

POLAR_DEV void thread_frame_epilogue(const DecodeParams& p, uint32_t* sm, int lane, int bitRow, long long frame0)
{
    // systematic: gather; non-systematic: re-encode first (fastssc_fip_char.cpp:616-630), then
    // getPackedInformationBits (bitcontainer.cpp:536-564) + Detector::check, all per lane.
    if (!p.systematic) thread_polar_transform(sm, lane, bitRow, p.N);
    __syncwarp(); // the stage rows are dead: reuse them as the image of the warp's 32 output rows
    uint8_t* img = reinterpret_cast<uint8_t*>(sm);
    uint32_t syn = 0, acc = 0;
    int curW = -1;
    uint32_t curV = 0;
    const bool wordOut = (p.kb & 3) == 0;
    const bool byRuns = wordOut && !p.crcBits && p.nRuns > 0;
    if (byRuns) { // information bits come in runs of consecutive positions: move a run at a time
        int outBits = 0;
        for (int r = 0; r < p.nRuns; ++r) {
            const uint32_t run = __ldg(p.runs + r);
            const int pos = (int)(run & 0xFFFFu), len = (int)(run >> 16);
            if ((pos >> 5) != curW) {
                curW = pos >> 5;
                curV = TF_ROW(curW);
            }
            const uint32_t rev = __brev(curV >> (pos & 31)) >> (32 - len); // first position -> highest bit
            acc = len == 32 ? rev : ((acc << len) | rev);
            outBits += len;
            if ((outBits & 31) == 0)
                reinterpret_cast<uint32_t*>(img)[lane * (p.kb >> 2) + (outBits >> 5) - 1] = __byte_perm(acc, 0u, 0x0123u);
        }
    }
    for (int k = byRuns ? p.K : 0; k < p.K; ++k) {
        const int pos = (int)__ldg(p.infoPos + k);
        if ((pos >> 5) != curW) { // uniform: positions ascend
            curW = pos >> 5;
            curV = TF_ROW(curW);
        }
        const uint32_t bit = (curV >> (pos & 31)) & 1u;
        acc = (acc << 1) | bit;
        if (p.crcBits) syn ^= __ldg(p.crcTab + k) & (0u - bit);
        if (wordOut) {
            if ((k & 31) == 31)
                reinterpret_cast<uint32_t*>(img)[lane * (p.kb >> 2) + (k >> 5)] = __byte_perm(acc, 0u, 0x0123u);
        } else if ((k & 7) == 7) {
            img[lane * p.kb + (k >> 3)] = (uint8_t)acc;
        }
    }
    if (wordOut) { // the last, partly filled unit is zero padded (MSB first)
        if (p.K & 31)
            reinterpret_cast<uint32_t*>(img)[lane * (p.kb >> 2) + (p.K >> 5)] =
                __byte_perm(acc << (32 - (p.K & 31)), 0u, 0x0123u);
    } else if (p.K & 7) {
        img[lane * p.kb + (p.K >> 3)] = (uint8_t)(acc << (8 - (p.K & 7)));
    }
    const bool ok = !p.crcBits || (syn ^ p.crcInit) == 0u;
    __syncwarp();
    // bit rows and the image may overlap only if the image is larger than the stage rows: never
    // (8*kb <= N words <= 32 * llrRows).  Coalesced copy of the image.
    const long long nHere = min(32LL, p.nFrames - frame0);
    uint8_t* dst = p.out + frame0 * p.kb;
    const long long bytes = nHere * p.kb;
    // the image is one contiguous run of bytes (16-byte aligned in shared memory): the widest store the caller's buffer allows
    if ((reinterpret_cast<uintptr_t>(dst) & 15u) == 0) {
        uint4* d16 = reinterpret_cast<uint4*>(dst);
        const uint4* s16 = reinterpret_cast<const uint4*>(img);
        for (long long i = lane; i < (bytes >> 4); i += 32) d16[i] = s16[i];
        for (long long i = (bytes & ~15LL) + lane; i < bytes; i += 32) dst[i] = img[i];
    } else if ((reinterpret_cast<uintptr_t>(dst) & 3u) == 0) {
        uint32_t* d4 = reinterpret_cast<uint32_t*>(dst);
        const uint32_t* s4 = reinterpret_cast<const uint32_t*>(img);
        for (long long i = lane; i < (bytes >> 2); i += 32) d4[i] = s4[i];
        for (long long i = (bytes & ~3LL) + lane; i < bytes; i += 32) dst[i] = img[i];
    } else {
        for (long long i = lane; i < bytes; i += 32) dst[i] = img[i];
    }
    if (frame0 + lane < p.nFrames) p.ok[frame0 + lane] = ok ? 1 : 0;
    __syncwarp();
}
#undef TF_ROW

} // namespace polar

