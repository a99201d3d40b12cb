// SCL int8 list decoder for sm_100a: one list of L paths per group of L lanes, 32/L frames per warp,
// thousands of frames per launch.
//
// Restates PolarCode::Decoding::SclFipChar (src/polarcode/decoding/scl_fip_char.cpp) bit-exactly:
//   * F/G/combine per live path (:315-385) on packed int8x4 words (primitives.cuh)
//   * leaves R0 / R1 / REP / SPC (:387-726) with exact integer path metrics (higher = better)
//   * candidate pruning = the reference's selection sort WITH swaps (arrayfuncs.h:161-183), emulated
//     in registers (every lane holds its own path's candidates), so ties resolve exactly like the CPU code
//   * findWeakLlrs (arrayfuncs.h:209-231) emulated including the swap displacement of elements 0..3
//   * CRC-first path pick (:816-855) fused as a syndrome computed while gathering the info bits
//
// B200 mapping (not the reference's design):
//   * the warp is split into 32/L groups of L lanes, one frame per group; lane q of a group holds path q's
//     state: metric and two 64-bit "lazy copy" vectors (5 bits per stage) that say which physical slot holds
//     the path's LLR block / left partial sums of each stage -- the analogue of DataPool::lazyDuplicate
//     (datapool.txx:100-104).  A fork is two 64-bit shuffles.  Because all paths (of all frames) run the same
//     schedule in lockstep, every stage block is dead whenever it is rewritten, so no copy-on-write is ever
//     needed: writers always write their own slot.
//   * with P live paths each path is worked on by L/P lanes (word-interleaved); once the list is full a
//     lane owns a path outright and walks its words sequentially -- every warp instruction then serves
//     32/L frames, and no lane idles on the small nodes that dominate a polar tree.
//   * the channel LLRs are read straight from global memory by the two top-level steps (L1/L2 resident the
//     second time); the top LLR stage can live in an L2-resident scratch ("spill", [word][lane] layout,
//     coalesced) so that the shared-memory footprint -- and with it the resident warp count -- is set by
//     the lower stages only.
//   * partial sums are packed sign bits (1 bit per code bit) -- the reference's byte soft-bits only
//     ever contribute their sign (fip_char.h:67-73,164-220).
#pragma once
#include "decode_common.cuh"
#include "decode_params.h"
#include "primitives.cuh"
#include "schedule.h"

namespace polar {

static constexpr int SCL_METRIC_BIAS = 1 << 22;

template <int L> struct Log2 { static constexpr int v = 1 + Log2<L / 2>::v; };
template <> struct Log2<1> { static constexpr int v = 0; };

// A block of 32-bit words reachable as p[w * stride]: shared memory (stride 1), a channel row in global
// memory (stride 1) or a column of the spill scratch (stride 32).
struct Words {
    uint32_t* p;
    int stride;
    POLAR_DEV uint32_t operator()(int w) const { return p[w * stride]; }
    POLAR_DEV void put(int w, uint32_t v) const { p[w * stride] = v; }
};

struct SclCtx {
    uint32_t* chan;   // this lane's frame: channel LLR row (global memory, read only)
    uint32_t* llr;    // this lane's frame: LLR slots (shared memory)
    uint32_t* bits;   // this lane's frame: bit slots (shared memory)
    uint32_t* spill;  // this lane's frame: first column of the frame in the warp's spill scratch
    int gl, fb;       // lane within the frame's group, first lane of the group
    int slotW, bitBW, bitSlotW, lgN, spillStage;
    int metric;
    unsigned long long llrPtr, lbPtr;
};

POLAR_DEV void scl_set_own(unsigned long long& vec, int stage, int gl)
{
    vec = (vec & ~(31ull << (5 * stage))) | ((unsigned long long)gl << (5 * stage));
}
POLAR_DEV unsigned long long shfl64(unsigned long long v, int src)
{
    uint32_t lo = __shfl_sync(FULL_MASK, (uint32_t)v, src);
    uint32_t hi = __shfl_sync(FULL_MASK, (uint32_t)(v >> 32), src);
    return ((unsigned long long)hi << 32) | lo;
}
// physical slot of `path`'s block at `stage` (path state lives in lane fb + path)
POLAR_DEV uint32_t scl_slot_of(const SclCtx& c, unsigned long long vec, int stage, int path, int lgG)
{
    uint32_t v = (uint32_t)(vec >> (5 * stage)) & 31u;
    if (lgG) v = __shfl_sync(FULL_MASK, v, c.fb + path); // a lane that owns its path reads its own register
    return v;
}
POLAR_DEV Words scl_llr_block(const SclCtx& c, int stage, uint32_t slot)
{
    if (stage == c.lgN) return Words{ c.chan, 1 };
    if (stage == c.spillStage) return Words{ c.spill + slot, 32 };
    return Words{ c.llr + slot * c.slotW + llr_woff(stage), 1 };
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
template <int L>
POLAR_DEV uint32_t list_max(uint32_t v)
{
    if (L == 32) return __reduce_max_sync(FULL_MASK, v);
#pragma unroll
    for (int d = 1; d < L; d <<= 1) v = max(v, __shfl_xor_sync(FULL_MASK, v, d));
    return v;
}

// ---------------------------------------------------------------- inner-node steps
template <int L>
POLAR_DEV void scl_op_f(SclCtx& c, int s, int lgP)
{
    const int lgG = Log2<L>::v - lgP, G = 1 << lgG, q = c.gl >> lgG, g = c.gl & (G - 1);
    const Words src = scl_llr_block(c, s + 1, scl_slot_of(c, c.llrPtr, s + 1, q, lgG));
    const Words dst = scl_llr_block(c, s, (uint32_t)q);
    if (s >= 2) {
        const int words = 1 << (s - 2);
        int w = g;
        for (; w + 3 * G < words; w += 4 * G) { // four independent butterflies in flight
            const uint32_t a0 = src(w), a1 = src(w + G), a2 = src(w + 2 * G), a3 = src(w + 3 * G);
            const uint32_t b0 = src(w + words), b1 = src(w + G + words), b2 = src(w + 2 * G + words),
                           b3 = src(w + 3 * G + words);
            dst.put(w, f4(a0, b0));
            dst.put(w + G, f4(a1, b1));
            dst.put(w + 2 * G, f4(a2, b2));
            dst.put(w + 3 * G, f4(a3, b3));
        }
        for (; w < words; w += G) dst.put(w, f4(src(w), src(w + words)));
    } else if (g == 0) { // block of 2: both halves sit in one word
        const uint32_t x = src(0);
        dst.put(0, f4(x, x >> 16));
    }
    scl_set_own(c.llrPtr, s, c.gl);
    __syncwarp();
}

template <int L>
POLAR_DEV void scl_op_g(SclCtx& c, int s, int lgP)
{
    const int lgG = Log2<L>::v - lgP, G = 1 << lgG, q = c.gl >> lgG, g = c.gl & (G - 1);
    const Words src = scl_llr_block(c, s + 1, scl_slot_of(c, c.llrPtr, s + 1, q, lgG));
    const uint32_t* lb = c.bits + scl_slot_of(c, c.lbPtr, s, q, lgG) * c.bitSlotW + bit_woff(s);
    const Words dst = scl_llr_block(c, s, (uint32_t)q);
    if (s >= 2) {
        const int words = 1 << (s - 2);
        int w = g;
        for (; w + 3 * G < words; w += 4 * G) {
            const uint32_t a0 = src(w), a1 = src(w + G), a2 = src(w + 2 * G), a3 = src(w + 3 * G);
            const uint32_t b0 = src(w + words), b1 = src(w + G + words), b2 = src(w + 2 * G + words),
                           b3 = src(w + 3 * G + words);
            const uint32_t n0 = lb[w >> 3] >> ((w & 7) * 4), n1 = lb[(w + G) >> 3] >> (((w + G) & 7) * 4);
            const uint32_t n2 = lb[(w + 2 * G) >> 3] >> (((w + 2 * G) & 7) * 4);
            const uint32_t n3 = lb[(w + 3 * G) >> 3] >> (((w + 3 * G) & 7) * 4);
            dst.put(w, g4(a0, b0, nibble_to_bytemask(n0)));
            dst.put(w + G, g4(a1, b1, nibble_to_bytemask(n1)));
            dst.put(w + 2 * G, g4(a2, b2, nibble_to_bytemask(n2)));
            dst.put(w + 3 * G, g4(a3, b3, nibble_to_bytemask(n3)));
        }
        for (; w < words; w += G) {
            const uint32_t nib = lb[w >> 3] >> ((w & 7) * 4);
            dst.put(w, g4(src(w), src(w + words), nibble_to_bytemask(nib)));
        }
    } else if (g == 0) {
        const uint32_t x = src(0);
        dst.put(0, g4(x, x >> 16, nibble_to_bytemask(lb[0] & 3u)));
    }
    scl_set_own(c.llrPtr, s, c.gl);
    __syncwarp();
}

template <int L>
POLAR_DEV void scl_op_combine(SclCtx& c, int s, int lgP, bool isRight)
{
    const int lgG = Log2<L>::v - lgP, G = 1 << lgG, q = c.gl >> lgG, g = c.gl & (G - 1);
    const uint32_t* left = c.bits + scl_slot_of(c, c.lbPtr, s, q, lgG) * c.bitSlotW + bit_woff(s);
    const uint32_t* right = c.bits + q * c.bitSlotW + c.bitBW + bit_woff(s);
    uint32_t* out = c.bits + q * c.bitSlotW + (isRight ? c.bitBW : 0) + bit_woff(s + 1);
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
    if (!isRight) scl_set_own(c.lbPtr, s + 1, c.gl);
    __syncwarp();
}

// ---------------------------------------------------------------- leaf helpers
// findWeakLlrs(.., n, K): the K smallest |max(llr,-127)| with the reference's tie rule.
// Results are replicated over the lanes of the path's group.
template <int K>
POLAR_DEV void scl_weak_search(const Words& src, int s, int lgG, int g, int (&val)[K], int (&idx)[K])
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
            const uint32_t x = src(w);
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
    if (lgG == 0) {
#pragma unroll
        for (int r = 0; r < K; ++r) tail[r] = T[r];
    } else {
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
    }
    // head: elements 0..3 can be displaced by the swaps; simulate the selection sort on them
    const uint32_t w0 = src(0);
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

// Keep the nOut best of M = P*F candidates in the reference's order (simplePartialSortDescending,
// arrayfuncs.h:161-183: selection sort that SWAPS the displaced element to the winner's place, strict >,
// so ties go to the lower current position).  Lane q < P of the group holds the candidates at list
// positions q*F .. q*F+F-1.  Out (lane j < nOut of the group): metric and original candidate index of
// list entry j.
template <int L, int LGF>
POLAR_DEV void scl_select(const SclCtx& c, const int (&cand)[1 << LGF], int lgP, int nOut, int& outMet, int& outIdx)
{
    constexpr int F = 1 << LGF;
    const int gl = c.gl;
    uint32_t key[F];
    int orig[F];
#pragma unroll
    for (int k = 0; k < F; ++k) {
        const int pos = gl * F + k;
        key[k] = gl < (1 << lgP) ? (((uint32_t)(cand[k] + SCL_METRIC_BIAS) << 8) | (uint32_t)(255 - pos)) : 0u;
        orig[k] = pos;
    }
    outMet = 0;
    outIdx = 0;
#pragma unroll(L <= 8 ? L : 1)
    for (int i = 0; i < L; ++i) {
        if (i >= nOut) break;
        uint32_t m = key[0];
#pragma unroll
        for (int k = 1; k < F; ++k) m = max(m, key[k]);
        const uint32_t best = list_max<L>(m);
        const int pos = 255 - (int)(best & 255u), wl = pos >> LGF, wk = pos & (F - 1);
        const int il = i >> LGF, ik = i & (F - 1);
        int wo = orig[0], io = orig[0];
        uint32_t ikl = key[0];
#pragma unroll
        for (int k = 1; k < F; ++k) {
            if (k == wk) wo = orig[k];
            if (k == ik) { io = orig[k]; ikl = key[k]; }
        }
        const int widx = __shfl_sync(FULL_MASK, wo, c.fb + wl);
        const uint32_t ikey = __shfl_sync(FULL_MASK, ikl, c.fb + il); // the element now at position i
        const int iidx = __shfl_sync(FULL_MASK, io, c.fb + il);
        if (gl == wl) { // it moves to the winner's place ...
#pragma unroll
            for (int k = 0; k < F; ++k)
                if (k == wk) {
                    key[k] = (ikey & ~255u) | (uint32_t)(255 - pos);
                    orig[k] = iidx;
                }
        }
        if (gl == il) { // ... and position i is final
#pragma unroll
            for (int k = 0; k < F; ++k)
                if (k == ik) key[k] = 0u;
        }
        if (gl == i) {
            outMet = (int)(best >> 8) - SCL_METRIC_BIAS;
            outIdx = widx;
        }
    }
}

// After a fork: adopt the parent's lazy-copy vectors and metric.
POLAR_DEV void scl_adopt(SclCtx& c, int outMet, int srcPath)
{
    c.metric = outMet;
    c.llrPtr = shfl64(c.llrPtr, c.fb + srcPath);
    c.lbPtr = shfl64(c.lbPtr, c.fb + srcPath);
}

// Sign bits of an LLR block (n = 2^s int8) packed LSB-first into bit word `bw`
POLAR_DEV uint32_t scl_sign_word(const Words& src, int s, int bw)
{
    if (s < 2) return sign_nibble(src(0)) & ((1u << (1 << s)) - 1u);
    const int words = 1 << (s - 2);
    uint32_t w = 0;
#pragma unroll
    for (int t = 0; t < 8; ++t) {
        const int lw = 8 * bw + t;
        if (lw < words) w |= sign_nibble(src(lw)) << (4 * t);
    }
    return w;
}

// ---------------------------------------------------------------- leaves
POLAR_DEV void scl_finish_leaf(SclCtx& c, int s, bool isRight)
{
    if (!isRight) scl_set_own(c.lbPtr, s, c.gl);
    __syncwarp();
}

// RateZero: scl_fip_char.cpp:387-421
template <int L>
POLAR_DEV void scl_leaf_r0(SclCtx& c, int s, int lgP, bool isRight)
{
    const int lgG = Log2<L>::v - lgP, G = 1 << lgG, q = c.gl >> lgG, g = c.gl & (G - 1);
    const Words src = scl_llr_block(c, s, scl_slot_of(c, c.llrPtr, s, q, lgG));
    int acc = 0;
    if (s >= 2) {
        const int words = 1 << (s - 2);
        for (int w = g; w < words; w += G) acc = sum_neg4(src(w), acc);
    } else if (g == 0) {
        acc = sum_neg4(src(0) & (s == 1 ? 0xFFFFu : 0xFFu), 0);
    }
    acc = group_add(acc, lgG);
    const int pun = __shfl_sync(FULL_MASK, acc, c.fb + ((c.gl << lgG) & (L - 1)));
    if (c.gl < (1 << lgP)) c.metric += pun;
    uint32_t* out = c.bits + q * c.bitSlotW + (isRight ? c.bitBW : 0) + bit_woff(s);
    const int nbw = s >= 5 ? 1 << (s - 5) : 1;
    for (int j = g; j < nbw; j += G) out[j] = 0u;
    scl_finish_leaf(c, s, isRight);
}

// Repetition: scl_fip_char.cpp:508-581.  Candidate 0 decides bit 0, candidate 1 bit 1; only the
// sign of mResults survives downstream, so the saturating sum itself is not needed here.
template <int L>
POLAR_DEV void scl_leaf_rep(SclCtx& c, int s, int lgP, bool isRight)
{
    constexpr int lgL = Log2<L>::v;
    const int lgG = lgL - lgP, G = 1 << lgG, q = c.gl >> lgG, g = c.gl & (G - 1);
    const Words src = scl_llr_block(c, s, scl_slot_of(c, c.llrPtr, s, q, lgG));
    int neg = 0, pos = 0;
    if (s >= 2) {
        const int words = 1 << (s - 2);
        for (int w = g; w < words; w += G) {
            const uint32_t x = src(w);
            neg = sum_neg4(x, neg);
            pos = sum_pos4(x, pos);
        }
    } else if (g == 0) {
        const uint32_t x = src(0) & (s == 1 ? 0xFFFFu : 0xFFu);
        neg = sum_neg4(x, 0);
        pos = sum_pos4(x, 0);
    }
    neg = group_add(neg, lgG);
    pos = group_add(pos, lgG);
    const int srcLane = c.fb + ((c.gl << lgG) & (L - 1));
    neg = __shfl_sync(FULL_MASK, neg, srcLane);
    pos = __shfl_sync(FULL_MASK, pos, srcLane);
    const int cand[2] = { c.metric + neg, c.metric - pos };
    const int lgPo = min(lgL, lgP + 1);
    int outMet, outIdx;
    scl_select<L, 1>(c, cand, lgP, 1 << lgPo, outMet, outIdx);
    scl_adopt(c, outMet, (outIdx >> 1) & (L - 1));
    const int lgGo = lgL - lgPo, Go = 1 << lgGo, qo = c.gl >> lgGo, go = c.gl & (Go - 1);
    const int kq = __shfl_sync(FULL_MASK, outIdx & 1, c.fb + qo);
    uint32_t* out = c.bits + qo * c.bitSlotW + (isRight ? c.bitBW : 0) + bit_woff(s);
    const int nbw = s >= 5 ? 1 << (s - 5) : 1;
    const uint32_t full = s >= 5 ? 0xFFFFFFFFu : (1u << (1 << s)) - 1u;
    for (int j = go; j < nbw; j += Go) out[j] = kq ? full : 0u;
    scl_finish_leaf(c, s, isRight);
}

// Shared tail of the R1 / SPC leaves: write sign(llr) with the chosen flips as the new path's bits.
template <int L>
POLAR_DEV void scl_write_flipped(SclCtx& c, int s, int lgPo, bool isRight, uint32_t flipMask, uint32_t pk01,
                                 uint32_t pk23)
{
    const int lgGo = Log2<L>::v - lgPo, Go = 1 << lgGo, qo = c.gl >> lgGo, go = c.gl & (Go - 1);
    uint32_t fm = flipMask, p01 = pk01, p23 = pk23;
    if (lgGo) {
        fm = __shfl_sync(FULL_MASK, fm, c.fb + qo);
        p01 = __shfl_sync(FULL_MASK, p01, c.fb + qo);
        p23 = __shfl_sync(FULL_MASK, p23, c.fb + qo);
    }
    const Words srcn = scl_llr_block(c, s, scl_slot_of(c, c.llrPtr, s, qo, lgGo)); // through the adopted vector
    uint32_t* out = c.bits + qo * c.bitSlotW + (isRight ? c.bitBW : 0) + bit_woff(s);
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
POLAR_DEV void scl_leaf_r1(SclCtx& c, int s, int lgP, bool isRight)
{
    constexpr int lgL = Log2<L>::v;
    const int lgG = lgL - lgP, G = 1 << lgG, q = c.gl >> lgG, g = c.gl & (G - 1);
    const Words src = scl_llr_block(c, s, scl_slot_of(c, c.llrPtr, s, q, lgG));
    int val[2], idx[2];
    scl_weak_search<2>(src, s, lgG, g, val, idx);
    const int srcLane = c.fb + ((c.gl << lgG) & (L - 1));
    const int a0 = __shfl_sync(FULL_MASK, val[0], srcLane), a1 = __shfl_sync(FULL_MASK, val[1], srcLane);
    const uint32_t pk = __shfl_sync(FULL_MASK, (uint32_t)idx[0] | ((uint32_t)idx[1] << 16), srcLane);
    const int m = c.metric;
    const int cand[4] = { m, m - a0, m - a1, m - a0 - a1 };
    const int lgPo = min(lgL, lgP + 2);
    int outMet, outIdx;
    scl_select<L, 2>(c, cand, lgP, 1 << lgPo, outMet, outIdx);
    const int srcPath = (outIdx >> 2) & (L - 1);
    const uint32_t pkSrc = __shfl_sync(FULL_MASK, pk, c.fb + srcPath);
    scl_adopt(c, outMet, srcPath);
    scl_write_flipped<L>(c, s, lgPo, isRight, (uint32_t)(outIdx & 3), pkSrc, 0u);
}

// SPC: scl_fip_char.cpp:583-726
template <int L>
POLAR_DEV void scl_leaf_spc(SclCtx& c, int s, int lgP, bool isRight)
{
    constexpr int lgL = Log2<L>::v;
    const int lgG = lgL - lgP, G = 1 << lgG, q = c.gl >> lgG, g = c.gl & (G - 1);
    const Words src = scl_llr_block(c, s, scl_slot_of(c, c.llrPtr, s, q, lgG));
    uint32_t px = 0;
    {
        const int words = 1 << (s - 2); // an SPC leaf has at least 4 elements
        for (int w = g; w < words; w += G) px ^= src(w);
    }
    px = group_xor(px, lgG);
    int par = __popc(px & 0x80808080u) & 1;
    int val[4], idx[4];
    scl_weak_search<4>(src, s, lgG, g, val, idx);
    const int srcLane = c.fb + ((c.gl << lgG) & (L - 1));
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
    scl_select<L, 3>(c, cand, lgP, 1 << lgPo, outMet, outIdx);
    const int srcPath = (outIdx >> 3) & (L - 1), kind = outIdx & 7;
    const uint32_t s01 = __shfl_sync(FULL_MASK, pk01, c.fb + srcPath), s23 = __shfl_sync(FULL_MASK, pk23, c.fb + srcPath);
    const int sp = __shfl_sync(FULL_MASK, par, c.fb + srcPath);
    // flip sets over (i0,i1,i2,i3), one nibble per candidate kind (scl_fip_char.cpp:641-692)
    const uint32_t fm = ((sp ? 0xEDB78421u : 0xFCA69530u) >> (4 * kind)) & 15u;
    scl_adopt(c, outMet, srcPath);
    scl_write_flipped<L>(c, s, lgPo, isRight, fm, s01, s23);
}

// ---------------------------------------------------------------- the kernel
template <int L>
__global__ void __launch_bounds__(256) scl_i8_kernel(const __grid_constant__ DecodeParams p,
                                                     const __grid_constant__ Schedule sched)
{
    constexpr int lgL = Log2<L>::v, FPW = 32 / L;
    extern __shared__ __align__(16) uint32_t smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, warpsPerBlock = blockDim.x >> 5;
    uint32_t* base = smem + warp * p.warpW;
    uint32_t* scr = base + FPW * p.frameW;
    SclCtx c;
    c.gl = lane & (L - 1);
    c.fb = lane & ~(L - 1);
    const int fi = lane >> lgL;
    c.llr = base + fi * p.frameW;
    c.bits = c.llr + L * p.slotW;
    c.spill = p.spill ? p.spill + (size_t)(blockIdx.x * warpsPerBlock + warp) * p.spillWarpW + c.fb : nullptr;
    c.slotW = p.slotW;
    c.bitBW = p.bitBW;
    c.bitSlotW = p.bitSlotW;
    c.lgN = p.lgN;
    c.spillStage = p.spillStage;

    const long long totalWarps = (long long)gridDim.x * warpsPerBlock;
    const long long nQuads = (p.nFrames + FPW - 1) / FPW;
    for (long long quad = (long long)blockIdx.x * warpsPerBlock + warp; quad < nQuads; quad += totalWarps) {
        const long long myFrame = min(quad * FPW + fi, p.nFrames - 1); // tail lanes redo the last frame
        c.chan = reinterpret_cast<uint32_t*>(const_cast<int8_t*>(static_cast<const int8_t*>(p.llr) + myFrame * p.N));
        c.metric = 0;
        c.llrPtr = 0ull;
        c.lbPtr = 0ull;
        int lgPfinal = 0;
        for (int pc = 0;; ++pc) {
            const uint32_t op = sched.ops[pc];
            const uint32_t code = POLAR_OP_CODE(op);
            if (code == OP_END) break;
            const int s = (int)POLAR_OP_STAGE(op), lgP = (int)POLAR_OP_LGP(op);
            const bool isRight = (POLAR_OP_FLAGS(op) & OPF_RIGHT) != 0;
            switch (code) {
            case OP_F: scl_op_f<L>(c, s, lgP); break;
            case OP_G: scl_op_g<L>(c, s, lgP); break;
            case OP_COMBINE: scl_op_combine<L>(c, s, lgP, isRight); break;
            case OP_SCL_R0: scl_leaf_r0<L>(c, s, lgP, isRight); lgPfinal = lgP; break;
            case OP_SCL_REP: scl_leaf_rep<L>(c, s, lgP, isRight); lgPfinal = min(lgL, lgP + 1); break;
            case OP_SCL_R1: scl_leaf_r1<L>(c, s, lgP, isRight); lgPfinal = min(lgL, lgP + 2); break;
            case OP_SCL_SPC: scl_leaf_spc<L>(c, s, lgP, isRight); lgPfinal = min(lgL, lgP + 3); break;
            default: break;
            }
        }
        // extractBestPath (scl_fip_char.cpp:816-855): first path in list order that passes the detector.
        // The whole warp works on one frame of the quad at a time.
        const int P = 1 << lgPfinal;
        uint32_t* outw = scr;
        uint32_t* tmp = scr + ((p.K + 31) >> 5);
        for (int f = 0; f < FPW; ++f) {
            const long long frame = quad * FPW + f;
            if (frame >= p.nFrames) break;
            const uint32_t* fbits = base + f * p.frameW + L * p.slotW;
            bool okFlag = false;
            for (int path = 0; path < P; ++path) {
                const uint32_t* cw = fbits + path * p.bitSlotW + bit_woff(p.lgN); // root bits: array A, own slot
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
                const uint32_t* cw = fbits + bit_woff(p.lgN);
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
}

} // namespace polar
