// Fast-SSC int8 decoder for sm_100a: one frame per warp, the whole node program in the constant bank.
//
// Restates PolarCode::Decoding::FastSscFipChar (src/polarcode/decoding/fastssc_fip_char.cpp) bit-exactly,
// including the stepwise-saturating REP sums (avxconvenience.h:92-101,202-210), the "first minimum,
// threshold 127" SPC search (:274-319, avxconvenience.cpp:13-56) and the arithmetic-negate flip that
// silently fails for 0 and -128 (:301).  Partial sums are packed sign bits at absolute code positions.
#pragma once
#include "decode_params.h"
#include "primitives.cuh"
#include "schedule.h"
#include "scl_i8.cuh" // load_channel_llrs / gather_info_and_syndrome / store_info / warp_polar_transform

namespace polar {

struct ScCtx {
    uint32_t* chan;
    uint32_t* llr;
    uint32_t* bits;
    uint32_t* scr;
    int lane, lgN;
};

POLAR_DEV const uint32_t* sc_llr(const ScCtx& c, int s) { return s == c.lgN ? c.chan : c.llr + llr_woff(s); }

POLAR_DEV uint32_t or8(uint32_t v)
{
    v |= __shfl_xor_sync(FULL_MASK, v, 1);
    v |= __shfl_xor_sync(FULL_MASK, v, 2);
    v |= __shfl_xor_sync(FULL_MASK, v, 4);
    return v;
}

// Write (or XOR) the sign bits of a block of n = 2^s values into bits[pos .. pos+n).  getWord(w) returns
// LLR word w of the block.  dup: also copy to [pos+n, pos+2n) (Zero-SPC / Zero-One leaves).
template <class GetWord>
POLAR_DEV void sc_put_signs(ScCtx& c, int s, int pos, bool doXor, bool dup, uint32_t flipBit, bool doFlip,
                            GetWord getWord)
{
    const int lane = c.lane;
    if (s >= 5) {
        const int words = 1 << (s - 2);
        for (int w0 = 0; w0 < words; w0 += 32) {
            const int w = w0 + lane; // words is a multiple of 8, w0 of 32: groups of 8 lanes share a bit word
            uint32_t v = (w < words) ? sign_nibble(getWord(w)) << (4 * (lane & 7)) : 0u;
            v = or8(v);
            if ((lane & 7) == 0 && w < words) {
                const int bw = w >> 3;
                if (doFlip && (int)(flipBit >> 5) == bw) v ^= 1u << (flipBit & 31u);
                uint32_t* d = c.bits + (pos >> 5) + bw;
                if (doXor) *d ^= v; else *d = v;
                if (dup) d[1 << (s - 5)] = v;
            }
        }
    } else {
        const int n = 1 << s, words = n >= 4 ? n >> 2 : 1;
        uint32_t v = 0;
        if (lane < words) {
            uint32_t nib = sign_nibble(getWord(lane));
            if (n < 4) nib &= (1u << n) - 1u;
            v = nib << (4 * lane);
        }
        v = or8(v);
        if (lane == 0) {
            if (doFlip) v ^= 1u << flipBit;
            const uint32_t mask = (1u << n) - 1u; // n <= 16
            uint32_t* d = c.bits + (pos >> 5);
            const int sh = pos & 31;
            uint32_t wd = *d;
            if (doXor) wd ^= v << sh; else wd = (wd & ~(mask << sh)) | (v << sh);
            if (dup) wd = (wd & ~(mask << (sh + n))) | (v << (sh + n));
            *d = wd;
        }
    }
}

// bits[pos .. pos+n) = value (0 or 1) for every position; pattern word for REP2
POLAR_DEV void sc_fill_bits(ScCtx& c, int s, int pos, uint32_t pattern)
{
    if (s >= 5) {
        for (int j = c.lane; j < (1 << (s - 5)); j += 32) c.bits[(pos >> 5) + j] = pattern;
    } else if (c.lane == 0) {
        const uint32_t mask = (1u << (1 << s)) - 1u;
        const int sh = pos & 31;
        uint32_t* d = c.bits + (pos >> 5);
        *d = (*d & ~(mask << sh)) | ((pattern & mask) << sh);
    }
}

// Lane b < 8 accumulates "vector position block" b of the reference's 32-byte AVX lanes:
// acc[j] = sat(..sat(sat(0 + x[j]) + x[32+j]).. ) for j in 4b..4b+3 (fastssc_fip_char.cpp:227-232)
POLAR_DEV uint32_t sc_rep_lane_acc(const uint32_t* src, int s, int lane)
{
    uint32_t acc = 0;
    if (s >= 2) {
        const int words = 1 << (s - 2);
        if (lane < 8)
            for (int w = lane; w < words; w += 8) acc = adds4(acc, src[w]);
    } else if (lane == 0) {
        acc = src[0] & 0xFFFFu;
    }
    return acc;
}

// First position of the minimum of |x| (unsigned, |-128| = 128) over n = 2^s values; key = |x| << 16 | idx
template <class GetWord>
POLAR_DEV uint32_t sc_min_uabs(int s, int lane, GetWord getWord)
{
    uint32_t best = 0xFFFFFFFFu;
    const int n = 1 << s, words = n >= 4 ? n >> 2 : 1;
    for (int w = lane; w < words; w += 32) {
        const uint32_t m = uabs4(getWord(w));
#pragma unroll
        for (int e = 0; e < 4; ++e)
            if (4 * w + e < n) best = min(best, (((m >> (8 * e)) & 0xFFu) << 16) | (uint32_t)(4 * w + e));
    }
    return __reduce_min_sync(FULL_MASK, best);
}

// SPC decision on a block of n = 2^s values (fastssc_fip_char.cpp:274-319): returns whether a bit flips
// and which.  Long form = threshold search (start 127, strict <), short form = plain first minimum with
// +127 pads.  The flip is an arithmetic negate, so it is void for 0 and -128.
template <class GetWord>
POLAR_DEV bool sc_spc_decide(int s, int lane, bool longForm, GetWord getWord, uint32_t& flipIdx)
{
    const int n = 1 << s, words = n >= 4 ? n >> 2 : 1;
    uint32_t px = 0;
    for (int w = lane; w < words; w += 32) px ^= getWord(w);
    px = __reduce_xor_sync(FULL_MASK, px);
    const bool odd = (__popc(px & 0x80808080u) & 1) != 0;
    const uint32_t key = sc_min_uabs(s, lane, getWord);
    uint32_t mv = key >> 16, mi = key & 0xFFFFu;
    bool flip = odd;
    if (longForm) {
        if (mv >= 127u) mi = 0u;
    } else if (mv > 127u && n < 32) {
        flip = false; // a +127 pad wins: the negate lands outside the block
    }
    const uint32_t wv = getWord((int)(mi >> 2)); // uniform address: broadcast
    const int v = (int)(int8_t)(wv >> (8 * (mi & 3u)));
    if (v == 0 || v == -128) flip = false;
    flipIdx = mi;
    return flip;
}

// ---------------------------------------------------------------- the kernel
__global__ void __launch_bounds__(256) fastssc_i8_kernel(const __grid_constant__ DecodeParams p,
                                                         const __grid_constant__ Schedule sched)
{
    extern __shared__ __align__(16) uint32_t smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, warpsPerBlock = blockDim.x >> 5;
    ScCtx c;
    c.chan = smem + warp * p.warpW;
    c.llr = c.chan + p.chanW;
    c.bits = c.llr + p.slotW;
    c.scr = c.bits + p.bitSlotW;
    c.lane = lane;
    c.lgN = p.lgN;

    const long long totalWarps = (long long)gridDim.x * warpsPerBlock;
    for (long long frame = (long long)blockIdx.x * warpsPerBlock + warp; frame < p.nFrames; frame += totalWarps) {
        load_channel_llrs(p, frame, c.chan, lane);
        __syncwarp();
        for (int pc = 0;; ++pc) {
            const uint32_t op = sched.ops[pc];
            const uint32_t code = POLAR_OP_CODE(op);
            if (code == OP_END) break;
            const int s = (int)POLAR_OP_STAGE(op), pos = (int)POLAR_OP_POS(op);
            const bool longForm = (POLAR_OP_FLAGS(op) & OPF_LONG) != 0;
            switch (code) {
            case OP_F: { // fip_char.h:76-91
                const uint32_t* src = sc_llr(c, s + 1);
                uint32_t* dst = c.llr + llr_woff(s);
                if (s >= 2) {
                    const int words = 1 << (s - 2);
                    for (int w = lane; w < words; w += 32) dst[w] = f4(src[w], src[w + words]);
                } else if (lane == 0) {
                    const uint32_t x = src[0];
                    dst[0] = f4(x, x >> 16);
                }
                break;
            }
            case OP_G: { // fip_char.h:93-110
                const uint32_t* src = sc_llr(c, s + 1);
                uint32_t* dst = c.llr + llr_woff(s);
                if (s >= 2) {
                    const int words = 1 << (s - 2);
                    for (int w = lane; w < words; w += 32) {
                        const int bp = pos + 4 * w;
                        const uint32_t nib = c.bits[bp >> 5] >> (bp & 31);
                        dst[w] = g4(src[w], src[w + words], nibble_to_bytemask(nib));
                    }
                } else if (lane == 0) {
                    const uint32_t x = src[0];
                    const uint32_t nib = (c.bits[pos >> 5] >> (pos & 31)) & 3u;
                    dst[0] = g4(x, x >> 16, nibble_to_bytemask(nib));
                }
                break;
            }
            case OP_G_0R: { // fip_char.h:112-131
                const uint32_t* src = sc_llr(c, s + 1);
                uint32_t* dst = c.llr + llr_woff(s);
                if (s >= 2) {
                    const int words = 1 << (s - 2);
                    for (int w = lane; w < words; w += 32) dst[w] = adds4(src[w], src[w + words]);
                } else if (lane == 0) {
                    const uint32_t x = src[0];
                    dst[0] = adds4(x, x >> 16);
                }
                break;
            }
            case OP_G_R1: { // ROneNode::simplifiedRightRateOneDecode, fastssc_fip_char.cpp:436-474
                const uint32_t* src = sc_llr(c, s + 1);
                const int h = 1 << s;
                auto right = [&](int w) -> uint32_t {
                    if (s >= 2) {
                        const int bp = pos + 4 * w;
                        const uint32_t nib = c.bits[bp >> 5] >> (bp & 31);
                        return g4(src[w], src[w + (1 << (s - 2))], nibble_to_bytemask(nib));
                    }
                    const uint32_t x = src[0];
                    return g4(x, x >> 16, nibble_to_bytemask((c.bits[pos >> 5] >> (pos & 31)) & 3u));
                };
                // stash the right LLR words so that both bit updates see the same values
                uint32_t* tmp = c.llr + llr_woff(s);
                const int words = s >= 2 ? 1 << (s - 2) : 1;
                for (int w = lane; w < words; w += 32) tmp[w] = right(w);
                __syncwarp();
                auto get = [&](int w) -> uint32_t { return tmp[w]; };
                sc_put_signs(c, s, pos + h, false, false, 0u, false, get); // right bits = sign(g)
                __syncwarp();
                sc_put_signs(c, s, pos, true, false, 0u, false, get);      // left ^= right
                break;
            }
            case OP_COMBINE: { // fip_char.h:164-207
                if (s >= 5) {
                    const int hw = 1 << (s - 5);
                    for (int j = lane; j < hw; j += 32) c.bits[(pos >> 5) + j] ^= c.bits[(pos >> 5) + hw + j];
                } else if (lane == 0) {
                    const int h = 1 << s, sh = pos & 31;
                    uint32_t wd = c.bits[pos >> 5];
                    wd ^= ((wd >> (sh + h)) & ((1u << h) - 1u)) << sh;
                    c.bits[pos >> 5] = wd;
                }
                break;
            }
            case OP_COMBINE_0R: { // fip_char.h:209-220
                if (s >= 5) {
                    const int hw = 1 << (s - 5);
                    for (int j = lane; j < hw; j += 32) c.bits[(pos >> 5) + j] = c.bits[(pos >> 5) + hw + j];
                } else if (lane == 0) {
                    const int h = 1 << s, sh = pos & 31;
                    const uint32_t mask = (1u << h) - 1u;
                    uint32_t wd = c.bits[pos >> 5];
                    wd = (wd & ~(mask << sh)) | (((wd >> (sh + h)) & mask) << sh);
                    c.bits[pos >> 5] = wd;
                }
                break;
            }
            case OP_SC_R0: sc_fill_bits(c, s, pos, 0u); break;
            case OP_SC_R1: {
                const uint32_t* src = sc_llr(c, s);
                sc_put_signs(c, s, pos, false, false, 0u, false, [&](int w) -> uint32_t { return src[w]; });
                break;
            }
            case OP_SC_REP: { // fastssc_fip_char.cpp:225-240,265-272 + reduce_adds_epi8
                uint32_t t = sc_rep_lane_acc(sc_llr(c, s), s, lane);
                t = adds4(t, __shfl_down_sync(FULL_MASK, t, 4));
                t = adds4(t, __shfl_down_sync(FULL_MASK, t, 2));
                t = adds4(t, __shfl_down_sync(FULL_MASK, t, 1));
                t = adds4(t, t >> 16);
                t = adds4(t, t >> 8);
                t = __shfl_sync(FULL_MASK, t, 0);
                sc_fill_bits(c, s, pos, (t & 0x80u) ? 0xFFFFFFFFu : 0u);
                break;
            }
            case OP_SC_REP2: { // fastssc_fip_char.cpp:249-263 + half_reduce_adds_epi8 (n >= 32)
                uint32_t t = sc_rep_lane_acc(sc_llr(c, s), s, lane);
                t = adds4(t, __shfl_xor_sync(FULL_MASK, t, 4));
                t = adds4(t, __shfl_xor_sync(FULL_MASK, t, 2));
                t = adds4(t, __shfl_xor_sync(FULL_MASK, t, 1));
                t = adds4(t, __byte_perm(t, 0u, 0x1032u));
                t = __shfl_sync(FULL_MASK, t, 0);
                const uint32_t pat = ((t & 0x80u) ? 0x55555555u : 0u) | ((t & 0x8000u) ? 0xAAAAAAAAu : 0u);
                sc_fill_bits(c, s, pos, pat);
                break;
            }
            case OP_SC_SPC: { // fastssc_fip_char.cpp:274-319
                const uint32_t* src = sc_llr(c, s);
                auto get = [&](int w) -> uint32_t {
                    const uint32_t x = src[w];
                    return s >= 2 ? x : (x & 0xFFFFu);
                };
                uint32_t fi;
                const bool flip = sc_spc_decide(s, lane, longForm, get, fi);
                sc_put_signs(c, s, pos, false, false, fi, flip, get);
                break;
            }
            case OP_SC_ZERO_SPC:   // fastssc_fip_char.cpp:321-388
            case OP_SC_ZERO_ONE: { // fastssc_fip_char.cpp:390-399
                const uint32_t* src = sc_llr(c, s);
                const int hs = s - 1; // half block
                uint32_t* y = c.llr + llr_woff(hs);
                if (hs >= 2) {
                    const int words = 1 << (hs - 2);
                    for (int w = lane; w < words; w += 32) y[w] = adds4(src[w], src[w + words]);
                } else if (lane == 0) {
                    const uint32_t x = src[0];
                    y[0] = adds4(x, x >> 16) & 0xFFFFu;
                }
                __syncwarp();
                auto get = [&](int w) -> uint32_t { return y[w]; };
                uint32_t fi = 0;
                bool flip = false;
                if (code == OP_SC_ZERO_SPC) flip = sc_spc_decide(hs, lane, longForm, get, fi);
                sc_put_signs(c, hs, pos, false, true, fi, flip, get);
                break;
            }
            default: break;
            }
            __syncwarp();
        }
        // systematic: gather; non-systematic: re-encode first (fastssc_fip_char.cpp:616-630)
        uint32_t* outw = c.scr;
        const uint32_t* cw = c.bits;
        if (!p.systematic) {
            warp_polar_transform(c.bits, p.N, lane);
        }
        const uint32_t syn = gather_info_and_syndrome(p, cw, outw, lane);
        store_info(p, frame, outw, lane);
        if (lane == 0) p.ok[frame] = syn == 0u ? 1 : 0;
        __syncwarp();
    }
}

} // namespace polar
