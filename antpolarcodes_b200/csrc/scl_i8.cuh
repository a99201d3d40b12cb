// SCL int8 list decoder for sm_100a: one list (up to 32 paths) per warp, thousands of frames per launch.
//
// Restates PolarCode::Decoding::SclFipChar (src/polarcode/decoding/scl_fip_char.cpp) bit-exactly:
//   * F/G/combine per live path (:315-385) on packed int8x4 words (primitives.cuh)
//   * leaves R0 / R1 / REP / SPC (:387-726) with exact integer path metrics (higher = better)
//   * candidate pruning = the reference's selection sort WITH swaps (arrayfuncs.h:161-183), emulated
//     with one warp-wide REDUX.MAX per kept path, so ties resolve exactly like the CPU code
//   * findWeakLlrs (arrayfuncs.h:209-231) emulated including the swap displacement of elements 0..3
//   * CRC-first path pick (:816-855) fused as a syndrome computed while gathering the info bits
//
// B200 mapping (not the reference's design):
//   * lane q holds path q's state: metric and two 64-bit "lazy copy" vectors (5 bits per stage) that
//     say which physical slot holds the path's LLR block / left partial sums of each stage -- the
//     analogue of DataPool::lazyDuplicate (datapool.txx:100-104).  A fork is two 64-bit shuffles.
//     Because all paths run the same schedule in lockstep, every stage block is dead whenever it is
//     rewritten, so no copy-on-write is ever needed: writers always write their own slot.
//   * bulk work (F/G/combine/leaf scans) is spread over all 32 lanes: with P live paths each path
//     gets 32/P lanes, word-interleaved, so shared-memory accesses are conflict free.
//   * partial sums are packed sign bits (1 bit per code bit) -- the reference's byte soft-bits only
//     ever contribute their sign (fip_char.h:67-73,164-220).
#pragma once
#include "decode_params.h"
#include "primitives.cuh"
#include "schedule.h"

namespace polar {

static constexpr int SCL_METRIC_BIAS = 1 << 22;

struct SclCtx {
    uint32_t* chan;
    uint32_t* llr;
    uint32_t* bits;
    uint32_t* scr;
    int lane;
    int slotW, bitW, bitSlotW, lgN;
    int metric;
    unsigned long long llrPtr, lbPtr;
};

POLAR_DEV uint32_t scl_slot_of(unsigned long long vec, int stage, int path)
{
    return __shfl_sync(FULL_MASK, (uint32_t)(vec >> (5 * stage)), path) & 31u;
}
POLAR_DEV void scl_set_own(unsigned long long& vec, int stage, int lane)
{
    vec = (vec & ~(31ull << (5 * stage))) | ((unsigned long long)lane << (5 * stage));
}
POLAR_DEV unsigned long long shfl64(unsigned long long v, int src)
{
    uint32_t lo = __shfl_sync(FULL_MASK, (uint32_t)v, src);
    uint32_t hi = __shfl_sync(FULL_MASK, (uint32_t)(v >> 32), src);
    return ((unsigned long long)hi << 32) | lo;
}
// LLR block of `path` at `stage` (the channel block is shared by all paths)
POLAR_DEV const uint32_t* scl_llr_src(const SclCtx& c, int stage, int path)
{
    const uint32_t slot = scl_slot_of(c.llrPtr, stage, path); // all lanes shuffle (convergent)
    return stage == c.lgN ? c.chan : c.llr + slot * c.slotW + llr_woff(stage);
}
POLAR_DEV int group_add(int v, int lgG)
{
    for (int d = 1; d < (1 << lgG); d <<= 1) v += __shfl_xor_sync(FULL_MASK, v, d);
    return v;
}
POLAR_DEV uint32_t group_min(uint32_t v, int lgG)
{
    for (int d = 1; d < (1 << lgG); d <<= 1) v = min(v, __shfl_xor_sync(FULL_MASK, v, d));
    return v;
}
POLAR_DEV uint32_t group_xor(uint32_t v, int lgG)
{
    for (int d = 1; d < (1 << lgG); d <<= 1) v ^= __shfl_xor_sync(FULL_MASK, v, d);
    return v;
}

// ---------------------------------------------------------------- inner-node steps
POLAR_DEV void scl_op_f(SclCtx& c, int s, int lgP)
{
    const int lgG = 5 - lgP, G = 1 << lgG, q = c.lane >> lgG, g = c.lane & (G - 1);
    const uint32_t* src = scl_llr_src(c, s + 1, q);
    uint32_t* dst = c.llr + q * c.slotW + llr_woff(s);
    if (s >= 2) {
        const int words = 1 << (s - 2);
        for (int w = g; w < words; w += G) dst[w] = f4(src[w], src[w + words]);
    } else if (g == 0) { // block of 2: both halves sit in one word
        const uint32_t x = src[0];
        dst[0] = f4(x, x >> 16);
    }
    scl_set_own(c.llrPtr, s, c.lane);
    __syncwarp();
}

POLAR_DEV void scl_op_g(SclCtx& c, int s, int lgP)
{
    const int lgG = 5 - lgP, G = 1 << lgG, q = c.lane >> lgG, g = c.lane & (G - 1);
    const uint32_t* src = scl_llr_src(c, s + 1, q);
    const uint32_t* lb = c.bits + scl_slot_of(c.lbPtr, s, q) * c.bitSlotW + bit_woff(s);
    uint32_t* dst = c.llr + q * c.slotW + llr_woff(s);
    if (s >= 2) {
        const int words = 1 << (s - 2);
        for (int w = g; w < words; w += G) {
            const uint32_t nib = lb[w >> 3] >> ((w & 7) * 4);
            dst[w] = g4(src[w], src[w + words], nibble_to_bytemask(nib));
        }
    } else if (g == 0) {
        const uint32_t x = src[0];
        dst[0] = g4(x, x >> 16, nibble_to_bytemask(lb[0] & 3u));
    }
    scl_set_own(c.llrPtr, s, c.lane);
    __syncwarp();
}

POLAR_DEV void scl_op_combine(SclCtx& c, int s, int lgP, bool isRight)
{
    const int lgG = 5 - lgP, G = 1 << lgG, q = c.lane >> lgG, g = c.lane & (G - 1);
    const uint32_t* left = c.bits + scl_slot_of(c.lbPtr, s, q) * c.bitSlotW + bit_woff(s);
    const uint32_t* right = c.bits + q * c.bitSlotW + c.bitW + bit_woff(s);
    uint32_t* out = c.bits + q * c.bitSlotW + (isRight ? c.bitW : 0) + bit_woff(s + 1);
    if (s >= 5) {
        const int hw = 1 << (s - 5);
        for (int j = g; j < hw; j += G) {
            const uint32_t l = left[j], r = right[j];
            out[j] = l ^ r;
            out[j + hw] = r;
        }
    } else if (g == 0) {
        const uint32_t l = left[0], r = right[0];
        out[0] = (l ^ r) | (r << (1 << s));
    }
    if (!isRight) scl_set_own(c.lbPtr, s + 1, c.lane);
    __syncwarp();
}

// ---------------------------------------------------------------- leaf helpers
// findWeakLlrs(.., n, K): the K smallest |max(llr,-127)| with the reference's tie rule.
// Results are replicated over the lanes of the path's group.
template <int K>
POLAR_DEV void scl_weak_search(const uint32_t* src, int s, int lgG, int g, int (&val)[K], int (&idx)[K])
{
    const int G = 1 << lgG, n = 1 << s;
    // tail: elements >= 4, each lane keeps its K smallest keys (|llr| << 16 | position), ascending
    uint32_t T[K];
#pragma unroll
    for (int r = 0; r < K; ++r) T[r] = 0xFFFFFFFFu;
    if (s > 2) {
        const int words = 1 << (s - 2);
        for (int w = g; w < words; w += G) {
            if (w == 0) continue;
            const uint32_t x = src[w];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                uint32_t key = ((uint32_t)abs127_byte(x, e) << 16) | (uint32_t)(4 * w + e);
#pragma unroll
                for (int r = 0; r < K; ++r) {
                    const uint32_t lo = min(T[r], key);
                    key = max(T[r], key);
                    T[r] = lo;
                }
            }
        }
    }
    uint32_t tail[K];
#pragma unroll
    for (int r = 0; r < K; ++r) {
        const uint32_t gm = group_min(T[0], lgG);
        tail[r] = gm;
        if (T[0] == gm) { // keys are unique: exactly one lane pops (or all pop the sentinel)
#pragma unroll
            for (int t = 0; t + 1 < K; ++t) T[t] = T[t + 1];
            T[K - 1] = 0xFFFFFFFFu;
        }
    }
    // head: elements 0..3 can be displaced by the swaps; simulate the selection sort on them
    const uint32_t w0 = src[0];
    int ha[4], cur[4];
    bool alive[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        ha[e] = abs127_byte(w0, e);
        cur[e] = e;
        alive[e] = e < n;
    }
    int t = 0;
#pragma unroll
    for (int i = 0; i < K; ++i) {
        if (i < n) {
            uint32_t best = 0xFFFFFFFFu;
#pragma unroll
            for (int r = 0; r < K; ++r)
                if (r == t) best = tail[r];
            int bw = -1;
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const uint32_t key = ((uint32_t)ha[e] << 16) | (uint32_t)cur[e];
                if (alive[e] && key < best) { best = key; bw = e; }
            }
            const int bpos = (int)(best & 0xFFFFu);
#pragma unroll
            for (int e = 0; e < 4; ++e)
                if (alive[e] && cur[e] == i) cur[e] = bpos; // the element at position i swaps out
            if (bw >= 0) {
#pragma unroll
                for (int e = 0; e < 4; ++e)
                    if (e == bw) alive[e] = false;
                idx[i] = bw;
            } else {
                idx[i] = bpos;
                ++t;
            }
            val[i] = (int)(best >> 16);
        } else {
            idx[i] = 0;
            val[i] = 0;
        }
    }
}

// Keep the nOut best of M = P*F candidates in the reference's order.  In: cand[k] of path `lane`.
// Out (lane j < nOut): metric and original candidate index of list entry j.
template <int LGF, int R>
POLAR_DEV void scl_select(SclCtx& c, const int (&cand)[1 << LGF], int lgP, int nOut, int& outMet, int& outIdx)
{
    constexpr int F = 1 << LGF;
    const int P = 1 << lgP, M = P << LGF, lane = c.lane;
    if (lane < P) {
#pragma unroll
        for (int k = 0; k < F; ++k) {
            const int pos = lane * F + k;
            c.scr[pos] = ((uint32_t)(cand[k] + SCL_METRIC_BIAS) << 8) | (uint32_t)(255 - pos);
        }
    }
    __syncwarp();
    uint32_t key[R];
    int idx[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int pos = lane + 32 * r;
        key[r] = pos < M ? c.scr[pos] : 0u;
        idx[r] = pos;
    }
    __syncwarp();
    outMet = 0;
    outIdx = 0;
    for (int i = 0; i < nOut; ++i) {
        uint32_t m = key[0];
#pragma unroll
        for (int r = 1; r < R; ++r) m = max(m, key[r]);
        const uint32_t best = __reduce_max_sync(FULL_MASK, m);
        const int pos = 255 - (int)(best & 255u), wr = pos >> 5, wl = pos & 31;
        int wi = idx[0];
#pragma unroll
        for (int r = 1; r < R; ++r)
            if (r == wr) wi = idx[r];
        const int widx = __shfl_sync(FULL_MASK, wi, wl);
        const uint32_t ikey = __shfl_sync(FULL_MASK, key[0], i); // position i lives in lane i, register 0
        const int iidx = __shfl_sync(FULL_MASK, idx[0], i);
        if (lane == wl) {
#pragma unroll
            for (int r = 0; r < R; ++r)
                if (r == wr) {
                    key[r] = (ikey & ~255u) | (uint32_t)(255 - pos);
                    idx[r] = iidx;
                }
        }
        if (lane == i) {
            key[0] = 0u; // finalised
            outMet = (int)(best >> 8) - SCL_METRIC_BIAS;
            outIdx = widx;
        }
    }
}

// After a fork: adopt the parent's lazy-copy vectors and metric.
POLAR_DEV void scl_adopt(SclCtx& c, int outMet, int srcPath)
{
    c.metric = outMet;
    c.llrPtr = shfl64(c.llrPtr, srcPath & 31);
    c.lbPtr = shfl64(c.lbPtr, srcPath & 31);
}

// Sign bits of an LLR block (n = 2^s int8) packed LSB-first into bit word `bw`
POLAR_DEV uint32_t scl_sign_word(const uint32_t* src, int s, int bw)
{
    if (s < 2) return sign_nibble(src[0]) & 3u;
    const int words = 1 << (s - 2);
    uint32_t w = 0;
#pragma unroll
    for (int t = 0; t < 8; ++t) {
        const int lw = 8 * bw + t;
        if (lw < words) w |= sign_nibble(src[lw]) << (4 * t);
    }
    return w;
}

// ---------------------------------------------------------------- leaves
POLAR_DEV void scl_finish_leaf(SclCtx& c, int s, bool isRight)
{
    if (!isRight) scl_set_own(c.lbPtr, s, c.lane);
    __syncwarp();
}

// RateZero: scl_fip_char.cpp:387-421
POLAR_DEV void scl_leaf_r0(SclCtx& c, int s, int lgP, bool isRight)
{
    const int lgG = 5 - lgP, G = 1 << lgG, q = c.lane >> lgG, g = c.lane & (G - 1);
    const uint32_t* src = scl_llr_src(c, s, q);
    int acc = 0;
    if (s >= 2) {
        const int words = 1 << (s - 2);
        for (int w = g; w < words; w += G) acc = sum_neg4(src[w], acc);
    } else if (g == 0) {
        acc = sum_neg4(src[0] & (s == 1 ? 0xFFFFu : 0xFFu), 0);
    }
    acc = group_add(acc, lgG);
    const int pun = __shfl_sync(FULL_MASK, acc, (c.lane << lgG) & 31);
    if (c.lane < (1 << lgP)) c.metric += pun;
    uint32_t* out = c.bits + q * c.bitSlotW + (isRight ? c.bitW : 0) + bit_woff(s);
    const int nbw = s >= 5 ? 1 << (s - 5) : 1;
    for (int j = g; j < nbw; j += G) out[j] = 0u;
    scl_finish_leaf(c, s, isRight);
}

// Repetition: scl_fip_char.cpp:508-581.  Candidate 0 decides bit 0, candidate 1 bit 1; only the
// sign of mResults survives downstream, so the saturating sum itself is not needed here.
template <int L>
POLAR_DEV void scl_leaf_rep(SclCtx& c, int s, int lgP, int lgL, bool isRight)
{
    const int lgG = 5 - lgP, G = 1 << lgG, q = c.lane >> lgG, g = c.lane & (G - 1);
    const uint32_t* src = scl_llr_src(c, s, q);
    int neg = 0, pos = 0;
    if (s >= 2) {
        const int words = 1 << (s - 2);
        for (int w = g; w < words; w += G) {
            const uint32_t x = src[w];
            neg = sum_neg4(x, neg);
            pos = sum_pos4(x, pos);
        }
    } else if (g == 0) {
        const uint32_t x = src[0] & (s == 1 ? 0xFFFFu : 0xFFu);
        neg = sum_neg4(x, 0);
        pos = sum_pos4(x, 0);
    }
    neg = group_add(neg, lgG);
    pos = group_add(pos, lgG);
    const int srcLane = (c.lane << lgG) & 31;
    neg = __shfl_sync(FULL_MASK, neg, srcLane);
    pos = __shfl_sync(FULL_MASK, pos, srcLane);
    const int cand[2] = { c.metric + neg, c.metric - pos };
    const int lgPo = min(lgL, lgP + 1);
    int outMet, outIdx;
    scl_select<1, (L * 2 + 31) / 32>(c, cand, lgP, 1 << lgPo, outMet, outIdx);
    scl_adopt(c, outMet, outIdx >> 1);
    const int lgGo = 5 - lgPo, Go = 1 << lgGo, qo = c.lane >> lgGo, go = c.lane & (Go - 1);
    const int kq = __shfl_sync(FULL_MASK, outIdx & 1, qo);
    uint32_t* out = c.bits + qo * c.bitSlotW + (isRight ? c.bitW : 0) + bit_woff(s);
    const int nbw = s >= 5 ? 1 << (s - 5) : 1;
    const uint32_t full = s >= 5 ? 0xFFFFFFFFu : (1u << (1 << s)) - 1u;
    for (int j = go; j < nbw; j += Go) out[j] = kq ? full : 0u;
    scl_finish_leaf(c, s, isRight);
}

// Shared tail of the R1 / SPC leaves: write sign(llr) with the chosen flips as the new path's bits.
POLAR_DEV void scl_write_flipped(SclCtx& c, int s, int lgPo, bool isRight, uint32_t flipMask, uint32_t pk01,
                                 uint32_t pk23)
{
    const int lgGo = 5 - lgPo, Go = 1 << lgGo, qo = c.lane >> lgGo, go = c.lane & (Go - 1);
    const uint32_t fm = __shfl_sync(FULL_MASK, flipMask, qo);
    const uint32_t p01 = __shfl_sync(FULL_MASK, pk01, qo), p23 = __shfl_sync(FULL_MASK, pk23, qo);
    const uint32_t* srcn = scl_llr_src(c, s, qo); // through the adopted lazy-copy vector
    uint32_t* out = c.bits + qo * c.bitSlotW + (isRight ? c.bitW : 0) + bit_woff(s);
    const int nbw = s >= 5 ? 1 << (s - 5) : 1;
    const uint32_t f[4] = { p01 & 0xFFFFu, p01 >> 16, p23 & 0xFFFFu, p23 >> 16 };
    for (int j = go; j < nbw; j += Go) {
        uint32_t w = scl_sign_word(srcn, s, j);
#pragma unroll
        for (int t = 0; t < 4; ++t)
            if (((fm >> t) & 1u) && (int)(f[t] >> 5) == j) w ^= 1u << (f[t] & 31u);
        out[j] = w;
    }
    scl_finish_leaf(c, s, isRight);
}

// RateOne: scl_fip_char.cpp:423-505
template <int L>
POLAR_DEV void scl_leaf_r1(SclCtx& c, int s, int lgP, int lgL, bool isRight)
{
    const int lgG = 5 - lgP, G = 1 << lgG, q = c.lane >> lgG, g = c.lane & (G - 1);
    const uint32_t* src = scl_llr_src(c, s, q);
    int val[2], idx[2];
    scl_weak_search<2>(src, s, lgG, g, val, idx);
    const int srcLane = (c.lane << lgG) & 31;
    const int a0 = __shfl_sync(FULL_MASK, val[0], srcLane), a1 = __shfl_sync(FULL_MASK, val[1], srcLane);
    const uint32_t pk = __shfl_sync(FULL_MASK, (uint32_t)idx[0] | ((uint32_t)idx[1] << 16), srcLane);
    const int m = c.metric;
    const int cand[4] = { m, m - a0, m - a1, m - a0 - a1 };
    const int lgPo = min(lgL, lgP + 2);
    int outMet, outIdx;
    scl_select<2, (L * 4 + 31) / 32>(c, cand, lgP, 1 << lgPo, outMet, outIdx);
    const int srcPath = (outIdx >> 2) & 31;
    const uint32_t pkSrc = __shfl_sync(FULL_MASK, pk, srcPath);
    scl_adopt(c, outMet, srcPath);
    scl_write_flipped(c, s, lgPo, isRight, (uint32_t)(outIdx & 3), pkSrc, 0u);
}

// SPC: scl_fip_char.cpp:583-726
template <int L>
POLAR_DEV void scl_leaf_spc(SclCtx& c, int s, int lgP, int lgL, bool isRight)
{
    const int lgG = 5 - lgP, G = 1 << lgG, q = c.lane >> lgG, g = c.lane & (G - 1);
    const uint32_t* src = scl_llr_src(c, s, q);
    uint32_t px = 0;
    {
        const int words = 1 << (s - 2); // an SPC leaf has at least 4 elements
        for (int w = g; w < words; w += G) px ^= src[w];
    }
    px = group_xor(px, lgG);
    int par = __popc(px & 0x80808080u) & 1;
    int val[4], idx[4];
    scl_weak_search<4>(src, s, lgG, g, val, idx);
    const int srcLane = (c.lane << lgG) & 31;
    const int a0 = __shfl_sync(FULL_MASK, val[0], srcLane), a1 = __shfl_sync(FULL_MASK, val[1], srcLane);
    const int a2 = __shfl_sync(FULL_MASK, val[2], srcLane), a3 = __shfl_sync(FULL_MASK, val[3], srcLane);
    const uint32_t pk01 = __shfl_sync(FULL_MASK, (uint32_t)idx[0] | ((uint32_t)idx[1] << 16), srcLane);
    const uint32_t pk23 = __shfl_sync(FULL_MASK, (uint32_t)idx[2] | ((uint32_t)idx[3] << 16), srcLane);
    par = __shfl_sync(FULL_MASK, par, srcLane);
    int m = c.metric, w = 0;
    if (par) m -= a0; else w = a0;
    const int cand[8] = { m,           m - w - a1,  m - w - a2,  m - w - a3,
                          m - a1 - a2, m - a1 - a3, m - a2 - a3, m - w - a1 - a2 - a3 };
    const int lgPo = min(lgL, lgP + 3);
    int outMet, outIdx;
    scl_select<3, (L * 8 + 31) / 32>(c, cand, lgP, 1 << lgPo, outMet, outIdx);
    const int srcPath = (outIdx >> 3) & 31, kind = outIdx & 7;
    const uint32_t s01 = __shfl_sync(FULL_MASK, pk01, srcPath), s23 = __shfl_sync(FULL_MASK, pk23, srcPath);
    const int sp = __shfl_sync(FULL_MASK, par, srcPath);
    // flip sets over (i0,i1,i2,i3), one nibble per candidate kind (scl_fip_char.cpp:641-692)
    const uint32_t fm = ((sp ? 0xEDB78421u : 0xFCA69530u) >> (4 * kind)) & 15u;
    scl_adopt(c, outMet, srcPath);
    scl_write_flipped(c, s, lgPo, isRight, fm, s01, s23);
}

// ---------------------------------------------------------------- prologue / epilogue (shared with Fast-SSC)
// Stage one frame's channel LLRs into shared memory.  int8 input: 16-byte coalesced loads.
// float input: quantised like CharContainer::insertLlr (bitcontainer.cpp:468-516): rne + saturate.
POLAR_DEV void load_channel_llrs(const DecodeParams& p, long long frame, uint32_t* chan, int lane)
{
    if (!p.llrIsFloat) {
        const int8_t* src = static_cast<const int8_t*>(p.llr) + frame * p.N;
        if ((p.N & 15) == 0) {
            const uint4* s4 = reinterpret_cast<const uint4*>(src);
            uint4* d4 = reinterpret_cast<uint4*>(chan);
            for (int i = lane; i < (p.N >> 4); i += 32) d4[i] = __ldg(s4 + i);
        } else {
            const uint32_t* s1 = reinterpret_cast<const uint32_t*>(src);
            for (int i = lane; i < (p.N >> 2); i += 32) chan[i] = __ldg(s1 + i);
        }
    } else {
        const float4* src = reinterpret_cast<const float4*>(static_cast<const float*>(p.llr) + frame * p.N);
        for (int i = lane; i < (p.N >> 2); i += 32) {
            const float4 v = __ldg(src + i);
            const float f[4] = { v.x, v.y, v.z, v.w };
            uint32_t w = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                // cvtps_epi32 gives INT_MIN for NaN / out of range, then signed saturation -> -128
                int q = (f[k] != f[k]) ? -128 : __float2int_rn(fminf(fmaxf(f[k], -2147483648.f), 2147483520.f));
                if (f[k] >= 2147483648.f && p.N >= 32) q = -128; // "integer indefinite" of the N >= 32 path
                w |= (uint32_t)(uint8_t)(int8_t)sat8i(q) << (8 * k);
            }
            chan[i] = w;
        }
    }
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

// ---------------------------------------------------------------- the kernel
template <int L>
__global__ void __launch_bounds__(256) scl_i8_kernel(const __grid_constant__ DecodeParams p,
                                                     const __grid_constant__ Schedule sched)
{
    extern __shared__ __align__(16) uint32_t smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, warpsPerBlock = blockDim.x >> 5;
    uint32_t* base = smem + warp * p.warpW;
    SclCtx c;
    c.chan = base;
    c.llr = c.chan + p.chanW;
    c.bits = c.llr + L * p.slotW;
    c.scr = c.bits + L * p.bitSlotW;
    c.lane = lane;
    c.slotW = p.slotW;
    c.bitW = p.bitW;
    c.bitSlotW = p.bitSlotW;
    c.lgN = p.lgN;
    int lgL = 0;
    while ((1 << lgL) < L) ++lgL;

    const long long totalWarps = (long long)gridDim.x * warpsPerBlock;
    for (long long frame = (long long)blockIdx.x * warpsPerBlock + warp; frame < p.nFrames; frame += totalWarps) {
        load_channel_llrs(p, frame, c.chan, lane);
        c.metric = 0;
        c.llrPtr = 0ull;
        c.lbPtr = 0ull;
        __syncwarp();
        int lgPfinal = 0;
        for (int pc = 0;; ++pc) {
            const uint32_t op = sched.ops[pc];
            const uint32_t code = POLAR_OP_CODE(op);
            if (code == OP_END) break;
            const int s = (int)POLAR_OP_STAGE(op), lgP = (int)POLAR_OP_LGP(op);
            const bool isRight = (POLAR_OP_FLAGS(op) & OPF_RIGHT) != 0;
            switch (code) {
            case OP_F: scl_op_f(c, s, lgP); break;
            case OP_G: scl_op_g(c, s, lgP); break;
            case OP_COMBINE: scl_op_combine(c, s, lgP, isRight); break;
            case OP_SCL_R0: scl_leaf_r0(c, s, lgP, isRight); lgPfinal = lgP; break;
            case OP_SCL_REP: scl_leaf_rep<L>(c, s, lgP, lgL, isRight); lgPfinal = min(lgL, lgP + 1); break;
            case OP_SCL_R1: scl_leaf_r1<L>(c, s, lgP, lgL, isRight); lgPfinal = min(lgL, lgP + 2); break;
            case OP_SCL_SPC: scl_leaf_spc<L>(c, s, lgP, lgL, isRight); lgPfinal = min(lgL, lgP + 3); break;
            default: break;
            }
        }
        // extractBestPath (scl_fip_char.cpp:816-855): first path in list order that passes the detector
        const int P = 1 << lgPfinal;
        uint32_t* outw = c.scr;
        uint32_t* tmp = c.scr + ((p.K + 31) >> 5);
        bool okFlag = false;
        for (int path = 0; path < P; ++path) {
            const uint32_t* cw = c.bits + path * c.bitSlotW + bit_woff(c.lgN); // root bits: array A, own slot
            if (!p.systematic) {
                for (int w = lane; w < max(p.N >> 5, 1); w += 32) tmp[w] = cw[w];
                __syncwarp();
                warp_polar_transform(tmp, p.N, lane);
                cw = tmp;
            }
            const uint32_t syn = gather_info_and_syndrome(p, cw, outw, lane);
            if (syn == 0u) { okFlag = true; break; }
            if (!p.crcBits) break;
        }
        if (!okFlag && P > 1) { // fall back to the most likely path
            const uint32_t* cw = c.bits + bit_woff(c.lgN);
            if (!p.systematic) {
                for (int w = lane; w < max(p.N >> 5, 1); w += 32) tmp[w] = cw[w];
                __syncwarp();
                warp_polar_transform(tmp, p.N, lane);
                cw = tmp;
            }
            (void)gather_info_and_syndrome(p, cw, outw, lane);
        }
        store_info(p, frame, outw, lane);
        if (lane == 0) p.ok[frame] = okFlag ? 1 : 0;
        __syncwarp();
    }
}

} // namespace polar
