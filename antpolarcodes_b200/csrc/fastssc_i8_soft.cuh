// Soft-output variant of the int8 Fast-SSC decoder: Decoder::getSoftCodeword after FastSscFipChar::decode
// (src/polarcode/decoding/decoder.cpp:147, bitcontainer.cpp:566-569), i.e. the N BYTE partial sums the root node leaves in
// the decoder's bit container.  The throughput kernel (fastssc_i8.cuh) keeps partial sums as packed sign bits -- all the
// decisions need -- so the byte values are produced here, by the same node program, with the reference's byte semantics:
//   R0 -> 127 (fastssc_fip_char.cpp:202-208), R1 -> the LLR (:210-215), REP / REP2 -> the stepwise-saturating sums
//   (:225-272), SPC -> the LLR with the least reliable byte negated when the parity is odd (:274-319), 0-SPC / 0-1 ->
//   sat(l + r), decided, duplicated (:321-399); inner nodes: left ^= right as RAW BYTES (CombineInPlace, fip_char.h:164-171),
//   nodes of at most 32 elements clamp both halves to -127 first (CombineBitsShort, :185-205); right-rate-1 nodes XOR the
//   g result into the left half unclamped (fastssc_fip_char.cpp:436-474).
// Only the sign of such a byte is a decision; the lower bits are what the reference's byte arithmetic happens to leave.
//
// Mapping: one frame per thread, 32 frames per warp; LLR stages and the byte partial sums of a frame live in a global
// scratch laid out [word][lane] (every access of a warp is one 128-byte line).  This is the receiver-side companion of the
// decoder, not the hot path: no shared-memory staging, no register stages.
#pragma once
#include "decode_params.h"
#include "primitives.cuh"
#include "schedule.h"

namespace polar {

struct SoftParams {
    const int8_t* llr;   // nFrames x N int8
    int8_t* soft;        // nFrames x N: the soft codeword
    int8_t* softInfo;    // nFrames x K: soft codeword at the information positions (getSoftInformation), or null
    const uint16_t* infoPos;
    uint32_t* scratch;   // per resident warp: warpW words
    long long nFrames;
    int N, lgN, K, warpW, bitsOff; // bitsOff: word offset of the byte partial sums inside a lane's scratch column
};

// scalar int8 primitives for blocks below one word (fip_char.h:44-73)
POLAR_DEV int sf_f(int a, int b)
{
    const int sgn = (a ^ b) | 1;
    int aa = abs(max(a, -127)), bb = abs(max(b, -127));
    aa = max(aa, 1);
    bb = max(bb, 1);
    const int m = min(aa, bb);
    return (sgn & 0x80) ? -m : m;
}
POLAR_DEV int sf_g(int a, int b, int bit) { return (bit & 0x80) ? sat8i(b - a) : sat8i(b + a); }

// byte i of a [word][lane] column
POLAR_DEV int sf_ldb(const uint32_t* col, int i) { return (int)(int8_t)(col[(i >> 2) << 5] >> (8 * (i & 3))); }
POLAR_DEV void sf_stb(uint32_t* col, int i, int v)
{
    uint32_t* w = col + ((i >> 2) << 5);
    const int sh = 8 * (i & 3);
    *w = (*w & ~(0xFFu << sh)) | ((uint32_t)(uint8_t)v << sh);
}

// SPC decision on n = 2^s bytes starting at byte 0 of column `x` (fastssc_fip_char.cpp:274-319): index of the byte to negate
// or -1.  Long form (n > 32): per 32-byte vector first minimum of |x| as an unsigned byte, accepted when strictly below the
// running minimum (start 127, compared as signed char: |-128| = 128 reads as -128); short form: first minimum with +127 pads.
POLAR_DEV int sf_spc_index(const uint32_t* x, int n, bool longForm)
{
    int parity = 0;
    for (int i = 0; i < n; ++i) parity ^= sf_ldb(x, i) & 0x80;
    if (!parity) return -1;
    if (!longForm) {
        int best = 256, bp = 0;
        for (int i = 0; i < n; ++i) {
            const int a = abs(sf_ldb(x, i));
            if (a < best) { best = a; bp = i; }
        }
        if (best > 127 && n < 32) return -1; // a +127 pad wins: the negate lands outside the block
        return bp;
    }
    int minIdx = 0, minAbs = 127;
    for (int v = 0; v < n; v += 32) {
        if (minAbs > 0) {
            int best = 256, bp = 0;
            for (int i = 0; i < 32; ++i) {
                const int a = abs(sf_ldb(x, v + i));
                if (a < best) { best = a; bp = i; }
            }
            const int testAbs = (int)(int8_t)best;
            if (testAbs < minAbs) { minIdx = v + bp; minAbs = testAbs; }
        }
    }
    return minIdx;
}

// stepwise-saturating REP sums in the reference's 32-lane order (fastssc_fip_char.cpp:225-240, avxconvenience.h:92-101):
// acc[j] accumulates x[j], x[32 + j], ...; then the 16 / 8 / 4 / 2 / 1 tree.  stopAt = 1: the sum (REP), 2: the two
// interleaved sums of DoubleRepetition (half_reduce_adds_epi8), returned in bytes 0 and 1.
POLAR_DEV uint32_t sf_rep_sum(const uint32_t* x, int n, int stopAt)
{
    int acc[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) acc[j] = 0;
    for (int v = 0; v < n; v += 32) {
#pragma unroll
        for (int j = 0; j < 32; ++j)
            if (v + j < n) acc[j] = sat8i(acc[j] + sf_ldb(x, v + j));
    }
    if (stopAt == 1) {
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1)
#pragma unroll
            for (int j = 0; j < 16; ++j)
                if (j < d) acc[j] = sat8i(acc[j] + acc[j + d]);
        return (uint32_t)(uint8_t)acc[0];
    }
#pragma unroll
    for (int d = 16; d >= 2; d >>= 1)
#pragma unroll
        for (int j = 0; j < 16; ++j)
            if (j < d) acc[j] = sat8i(acc[j] + acc[j + d]);
    return (uint32_t)(uint8_t)acc[0] | ((uint32_t)(uint8_t)acc[1] << 8);
}

__global__ void __launch_bounds__(256) fastssc_i8_soft_kernel(const __grid_constant__ SoftParams p,
                                                              const __grid_constant__ Schedule sched)
{
    const int lane = threadIdx.x & 31, warpsPerBlock = blockDim.x >> 5;
    const long long warpId = (long long)blockIdx.x * warpsPerBlock + (threadIdx.x >> 5);
    const long long totalWarps = (long long)gridDim.x * warpsPerBlock;
    uint32_t* col = p.scratch + (size_t)warpId * p.warpW + lane; // this lane's column: word w at col[w << 5]
    uint32_t* bitsCol = col + ((size_t)p.bitsOff << 5);
    const int nW = max(p.N >> 2, 1);
    for (long long grp = warpId; grp * 32 < p.nFrames; grp += totalWarps) {
        const long long frame = grp * 32 + lane;
        const bool live = frame < p.nFrames;
        const uint32_t* chan = reinterpret_cast<const uint32_t*>(p.llr + (live ? frame : p.nFrames - 1) * p.N);
        uint32_t* top = col + (llr_woff(p.lgN) << 5); // the channel row is stage lgN of the column
        for (int w = 0; w < nW; ++w) top[w << 5] = __ldg(chan + w);
        for (int pc = 0;; ++pc) {
            const uint32_t op = sched.ops[pc];
            const uint32_t code = POLAR_OP_CODE(op);
            if (code == OP_END) break;
            const int s = (int)POLAR_OP_STAGE(op), pos = (int)POLAR_OP_POS(op), n = 1 << s;
            const bool longForm = (POLAR_OP_FLAGS(op) & OPF_LONG) != 0;
            uint32_t* dst = col + (llr_woff(s) << 5);            // stage s block (F / G / G-0R / G-R1: the child block)
            const uint32_t* par = col + (llr_woff(s + 1) << 5);  // its parent
            uint32_t* bits = bitsCol + ((pos >> 2) << 5);         // byte partial sums from position pos (word aligned for n >= 4)
            const int bo = pos & 3;                               // byte offset inside that word (n < 4 only)
            switch (code) {
            case OP_F:
            case OP_G:
            case OP_G_0R:
            case OP_G_R1:
                if (s >= 2) {
                    const int words = n >> 2;
                    for (int w = 0; w < words; ++w) {
                        const uint32_t a = par[w << 5], b = par[(w + words) << 5];
                        if (code == OP_F) dst[w << 5] = f4(a, b);
                        else if (code == OP_G_0R) dst[w << 5] = adds4(a, b);
                        else {
                            const uint32_t lw = bits[w << 5];      // byte k's sign bit = left bit of element k
                            const uint32_t r = g4(a, b, lw);
                            if (code == OP_G) dst[w << 5] = r;
                            else { bits[w << 5] = lw ^ r; bits[(w + words) << 5] = r; }
                        }
                    }
                } else { // blocks of 1 or 2 elements: scalar, both halves inside one parent word
                    for (int i = 0; i < n; ++i) {
                        const int a = sf_ldb(par, i), b = sf_ldb(par, i + n);
                        if (code == OP_F) sf_stb(dst, i, sf_f(a, b));
                        else if (code == OP_G_0R) sf_stb(dst, i, sat8i(a + b));
                        else {
                            const int lb = sf_ldb(bits, bo + i), r = sf_g(a, b, lb);
                            if (code == OP_G) sf_stb(dst, i, r);
                            else { sf_stb(bits, bo + i, lb ^ r); sf_stb(bits, bo + n + i, r); }
                        }
                    }
                }
                break;
            case OP_COMBINE: { // children of stage s: short form iff the node has at most 32 elements
                const bool shortForm = 2 * n <= 32;
                if (s >= 2) {
                    const int words = n >> 2;
                    for (int w = 0; w < words; ++w) {
                        uint32_t l = bits[w << 5], r = bits[(w + words) << 5];
                        if (shortForm) { l = __vmaxs4(l, 0x81818181u); r = __vmaxs4(r, 0x81818181u); bits[(w + words) << 5] = r; }
                        bits[w << 5] = l ^ r;
                    }
                } else {
                    for (int i = 0; i < n; ++i) { // always short here
                        const int l = max(sf_ldb(bits, bo + i), -127), r = max(sf_ldb(bits, bo + n + i), -127);
                        sf_stb(bits, bo + i, l ^ r);
                        sf_stb(bits, bo + n + i, r);
                    }
                }
                break;
            }
            case OP_COMBINE_0R:
                if (s >= 2) {
                    const int words = n >> 2;
                    for (int w = 0; w < words; ++w) bits[w << 5] = bits[(w + words) << 5];
                } else {
                    for (int i = 0; i < n; ++i) sf_stb(bits, bo + i, sf_ldb(bits, bo + n + i));
                }
                break;
            case OP_SC_R0:
                if (s >= 2) for (int w = 0; w < (n >> 2); ++w) bits[w << 5] = 0x7F7F7F7Fu;
                else for (int i = 0; i < n; ++i) sf_stb(bits, bo + i, 127);
                break;
            case OP_SC_R1:
                if (s >= 2) for (int w = 0; w < (n >> 2); ++w) bits[w << 5] = dst[w << 5];
                else for (int i = 0; i < n; ++i) sf_stb(bits, bo + i, sf_ldb(dst, i));
                break;
            case OP_SC_REP: {
                const uint32_t sum = sf_rep_sum(dst, n, 1);
                if (s >= 2) for (int w = 0; w < (n >> 2); ++w) bits[w << 5] = sum * 0x01010101u;
                else for (int i = 0; i < n; ++i) sf_stb(bits, bo + i, (int)(int8_t)sum);
                break;
            }
            case OP_SC_REP2: { // n >= 32: even positions get the sum of the even ones, odd positions that of the odd ones
                const uint32_t two = sf_rep_sum(dst, n, 2);
                for (int w = 0; w < (n >> 2); ++w) bits[w << 5] = two * 0x00010001u;
                break;
            }
            case OP_SC_SPC: {
                const int idx = sf_spc_index(dst, n, longForm);
                if (s >= 2) for (int w = 0; w < (n >> 2); ++w) bits[w << 5] = dst[w << 5];
                else for (int i = 0; i < n; ++i) sf_stb(bits, bo + i, sf_ldb(dst, i));
                if (idx >= 0) sf_stb(bits, bo + idx, -sf_ldb(bits, bo + idx)); // (char) -x: 0 and -128 stay
                break;
            }
            case OP_SC_ZERO_SPC:
            case OP_SC_ZERO_ONE: { // y = sat(l + r) over the halves -> the (free) child stage, decided, duplicated
                const int h = n >> 1, hs = s - 1;
                uint32_t* y = col + (llr_woff(hs) << 5);
                for (int i = 0; i < h; ++i) sf_stb(y, i, sat8i(sf_ldb(dst, i) + sf_ldb(dst, i + h)));
                if (code == OP_SC_ZERO_SPC) {
                    const int idx = sf_spc_index(y, h, longForm); // the form follows the NODE's length (ZeroSpc vs ShortZeroSpc)
                    if (idx >= 0) sf_stb(y, idx, -sf_ldb(y, idx));
                }
                for (int i = 0; i < h; ++i) {
                    const int v = sf_ldb(y, i);
                    sf_stb(bits, bo + i, v);
                    sf_stb(bits, bo + h + i, v);
                }
                break;
            }
            default: break;
            }
        }
        if (live) {
            uint32_t* out = reinterpret_cast<uint32_t*>(p.soft + frame * p.N);
            if (p.N >= 4) for (int w = 0; w < nW; ++w) out[w] = bitsCol[w << 5];
            if (p.softInfo)
                for (int k = 0; k < p.K; ++k) p.softInfo[frame * p.K + k] = (int8_t)sf_ldb(bitsCol, (int)__ldg(p.infoPos + k));
        }
    }
}

} // namespace polar
