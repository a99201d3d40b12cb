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
    const uint32_t* p;
    int stride;
    bool cg; // spill column: read at L2 (the writers use st.cg, L1 is never involved)
    POLAR_DEV uint32_t operator()(int w) const { return cg ? __ldcg(p + w * stride) : p[w * stride]; }
};

struct SclCtx {
    const DecodeParams& P; // geometry is read where it is used: kernel parameters are constant-bank operands, not registers
    const uint32_t* chan; // this lane's frame: channel LLR row (global memory, read only)
    uint32_t* frame;  // this lane's frame in shared memory: LLR slots, bit slots, select scratch
    uint32_t* spill;  // this lane's frame: first column of the frame in the warp's spill scratch
    int gl, fb;       // lane within the frame's group, first lane of the group
    int metric;
    unsigned long long llrPtr, lbPtr;
    // register stages (stages <= P.regStages = min(4, lgN - 2) live in the owning lane's registers): LLR blocks of stage
    // 4 / 3 / 2 / 1 in r[4..7] / r[2..3] / r[1] / r[0]; left-child bits of stage s in bits [2^s, 2^(s+1)) of ba;
    // right-child bits of stage s in bb[s & 1]
    uint32_t r[8];
    uint32_t ba, bb[2];
    // LLR slots (only the stages above the register stages: word llr_woff(s) - llrBase of a slot), bit slots, select scratch
    POLAR_DEV uint32_t* llrp() const { return frame - P.llrBase; }
    POLAR_DEV uint32_t* bitsp() const { return frame + P.bitsOff; }
    POLAR_DEV uint32_t* selp() const { return frame + P.selOff; }
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
// lgG = log2 of the lanes per path: with a full list (lgG == 0, warp uniform) the path is the lane's own
POLAR_DEV uint32_t scl_slot_of(const SclCtx& c, unsigned long long vec, int stage, int path, int lgG)
{
    const uint32_t mine = (uint32_t)(vec >> (5 * stage)) & 31u;
    if (lgG == 0) return mine;
    return __shfl_sync(FULL_MASK, mine, c.fb + path);
}
// word offset of a spilled stage inside the warp's scratch ([word][lane] layout, 32 columns)
POLAR_DEV int scl_spill_woff(const SclCtx& c, int stage) { return ((1 << (stage - 2)) - (1 << (c.P.smemStages - 2))) << 5; }
// any LLR block through a generic pointer (leaves; the F/G steps use typed pointers)
POLAR_DEV Words scl_llr_block(const SclCtx& c, int stage, uint32_t slot)
{
    if (stage == c.P.lgN) return Words{ c.chan, 1, false };
    if (stage >= c.P.smemStages) return Words{ c.spill + scl_spill_woff(c, stage) + slot, 32, true };
    return Words{ c.llrp() + slot * c.P.slotW + llr_woff(stage), 1, false };
}
// bit blocks: region A keeps every stage (left-child results wait there for the combine), region B is a
// ping-pong pair (a right-child result is consumed by the very next combine)
// A block of packed partial sums: words of a shared-memory slot, or a column of the L2 scratch (stride 32, L2 only).
struct BitBlk {
    uint32_t* p;
    bool g;
    POLAR_DEV uint32_t ld(int j) const { return g ? __ldcg(p + (j << 5)) : p[j]; }
    POLAR_DEV void st(int j, uint32_t v) const
    {
        if (g) __stcg(p + (j << 5), v);
        else p[j] = v;
    }
};
// (bits / spill = the frame's bit slots / first scratch column; the epilogue walks the frames of the warp with these)
POLAR_DEV BitBlk scl_bits_a_at(const SclCtx& c, uint32_t* bits, uint32_t* spill, uint32_t slot, int stage)
{
    if (stage >= c.P.bitSpill) return BitBlk{ spill + ((c.P.spA + (1 << (stage - 5)) - (1 << (c.P.bitSpill - 5))) << 5) + slot, true };
    return BitBlk{ bits + slot * c.P.bitSlotW + bit_woff(stage) - c.P.bitBase, false };
}
POLAR_DEV BitBlk scl_bits_b_at(const SclCtx& c, uint32_t* bits, uint32_t* spill, uint32_t slot, int stage)
{
    if (stage >= c.P.bitSpill) return BitBlk{ spill + (((stage & 1) ? c.P.spB1 : c.P.spB0) << 5) + slot, true };
    return BitBlk{ bits + slot * c.P.bitSlotW + c.P.bitBW + ((stage & 1) ? c.P.bitB1 : 0), false };
}
POLAR_DEV BitBlk scl_bits_a(const SclCtx& c, uint32_t slot, int stage) { return scl_bits_a_at(c, c.bitsp(), c.spill, slot, stage); }
POLAR_DEV BitBlk scl_bits_b(const SclCtx& c, uint32_t slot, int stage) { return scl_bits_b_at(c, c.bitsp(), c.spill, slot, stage); }
POLAR_DEV BitBlk scl_bits_out(const SclCtx& c, uint32_t slot, int stage, bool isRight)
{
    return isRight ? scl_bits_b(c, slot, stage) : scl_bits_a(c, slot, stage);
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
#ifdef POLAR_SCL_REDUX_GROUPS // one REDUX per list group instead of log2(L) shuffle + max pairs (A/B: profiles/r2_experiments.txt)
    if (L >= POLAR_SCL_REDUX_GROUPS) {
        const uint32_t lane = threadIdx.x & 31u;
        return __reduce_max_sync((L == 32 ? 0xFFFFFFFFu : ((1u << L) - 1u)) << (lane & ~(uint32_t)(L - 1)), v);
    }
#endif
#pragma unroll
    for (int d = 1; d < L; d <<= 1) v = max(v, __shfl_xor_sync(FULL_MASK, v, d));
    return v;
}

// ---------------------------------------------------------------- inner-node steps
// Parent block in global memory: the channel row (read only, L1 allocating) or a spill column (L2 only).
template <bool SG>
POLAR_DEV uint32_t scl_ld(const uint32_t* p, bool chan)
{
    if (!SG) return *p;
    return chan ? __ldcs(p) : __ldcg(p);
}
template <bool DG>
POLAR_DEV void scl_st(uint32_t* p, uint32_t v)
{
    if (DG) __stcg(p, v); else *p = v;
}

template <bool SG, bool DG>
POLAR_DEV void scl_f_words(const uint32_t* src, int ss, bool chan, uint32_t* dst, int ds, int s, int g, int G)
{
    if (s >= 2) {
        const int words = 1 << (s - 2);
        const uint32_t* srcb = src + words * ss;
        int w = g;
        if (SG) { // parent block in the L2 scratch / global memory: eight butterflies (16 loads) in flight
            for (; w + 7 * G < words; w += 8 * G) {
                uint32_t a[8], b[8];
#pragma unroll
                for (int t = 0; t < 8; ++t) {
                    a[t] = scl_ld<SG>(src + (w + t * G) * ss, chan);
                    b[t] = scl_ld<SG>(srcb + (w + t * G) * ss, chan);
                }
#pragma unroll
                for (int t = 0; t < 8; ++t) scl_st<DG>(dst + (w + t * G) * ds, f4(a[t], b[t]));
            }
        }
        for (; w + 3 * G < words; w += 4 * G) { // four independent butterflies in flight
            uint32_t a[4], b[4];
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                a[t] = scl_ld<SG>(src + (w + t * G) * ss, chan);
                b[t] = scl_ld<SG>(srcb + (w + t * G) * ss, chan);
            }
#pragma unroll
            for (int t = 0; t < 4; ++t) scl_st<DG>(dst + (w + t * G) * ds, f4(a[t], b[t]));
        }
        for (; w < words; w += G) scl_st<DG>(dst + w * ds, f4(scl_ld<SG>(src + w * ss, chan), scl_ld<SG>(srcb + w * ss, chan)));
    } else if (g == 0) { // block of 2: both halves sit in one word
        const uint32_t x = scl_ld<SG>(src, chan);
        scl_st<DG>(dst, f4(x, x >> 16));
    }
}

template <bool SG, bool DG>
POLAR_DEV void scl_g_words(const uint32_t* src, int ss, bool chan, const BitBlk& lb, uint32_t* dst, int ds, int s,
                           int g, int G)
{
    if (s >= 2) {
        const int words = 1 << (s - 2);
        const uint32_t* srcb = src + words * ss;
        int w = g;
        if (G == 1 && SG) { // parent block in the L2 scratch / global memory: 16 loads in flight
            for (; w + 7 < words; w += 8) {
                uint32_t a[8], b[8];
                const uint32_t lw = lb.ld(w >> 3);
#pragma unroll
                for (int t = 0; t < 8; ++t) {
                    a[t] = scl_ld<SG>(src + (w + t) * ss, chan);
                    b[t] = scl_ld<SG>(srcb + (w + t) * ss, chan);
                }
#pragma unroll
                for (int t = 0; t < 8; ++t) scl_st<DG>(dst + (w + t) * ds, g4(a[t], b[t], nibble_to_msb(lw >> (4 * t))));
            }
        }
        if (G == 1) { // a lane owns the path: the four words of a batch share one word of left bits
            for (; w + 3 < words; w += 4) {
                uint32_t a[4], b[4];
                const uint32_t lw = lb.ld(w >> 3) >> ((w & 4) * 4);
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    a[t] = scl_ld<SG>(src + (w + t) * ss, chan);
                    b[t] = scl_ld<SG>(srcb + (w + t) * ss, chan);
                }
#pragma unroll
                for (int t = 0; t < 4; ++t) scl_st<DG>(dst + (w + t) * ds, g4(a[t], b[t], nibble_to_msb(lw >> (4 * t))));
            }
        }
        for (; w + 3 * G < words; w += 4 * G) {
            uint32_t a[4], b[4], n[4];
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                const int wt = w + t * G;
                a[t] = scl_ld<SG>(src + wt * ss, chan);
                b[t] = scl_ld<SG>(srcb + wt * ss, chan);
                n[t] = lb.ld(wt >> 3) >> ((wt & 7) * 4);
            }
#pragma unroll
            for (int t = 0; t < 4; ++t) scl_st<DG>(dst + (w + t * G) * ds, g4(a[t], b[t], nibble_to_msb(n[t])));
        }
        for (; w < words; w += G) {
            const uint32_t nib = lb.ld(w >> 3) >> ((w & 7) * 4);
            scl_st<DG>(dst + w * ds, g4(scl_ld<SG>(src + w * ss, chan), scl_ld<SG>(srcb + w * ss, chan), nibble_to_msb(nib)));
        }
    } else if (g == 0) {
        const uint32_t x = scl_ld<SG>(src, chan);
        scl_st<DG>(dst, g4(x, x >> 16, nibble_to_msb(lb.ld(0) & 3u)));
    }
}

// F (lb == nullptr) or G step producing the stage-s block of every live path
template <int L, bool ISG>
POLAR_DEV void scl_op_fg(SclCtx& c, int s, int lgP)
{
    const int lgG = Log2<L>::v - lgP, G = 1 << lgG, q = c.gl >> lgG, g = c.gl & (G - 1);
    const uint32_t slot = scl_slot_of(c, c.llrPtr, s + 1, q, lgG);
    BitBlk lb{ nullptr, false };
    if (ISG) lb = scl_bits_a(c, scl_slot_of(c, c.lbPtr, s, q, lgG), s);
    if (s + 1 < c.P.smemStages) { // shared -> shared: the bulk of the steps
        const uint32_t* src = c.llrp() + slot * c.P.slotW + llr_woff(s + 1);
        uint32_t* dst = c.llrp() + q * c.P.slotW + llr_woff(s);
        if (ISG) scl_g_words<false, false>(src, 1, false, lb, dst, 1, s, g, G);
        else scl_f_words<false, false>(src, 1, false, dst, 1, s, g, G);
    } else {
        const bool chan = (s + 1 == c.P.lgN);
        const uint32_t* src = chan ? c.chan : c.spill + scl_spill_woff(c, s + 1) + slot;
        const int ss = chan ? 1 : 32;
        if (s < c.P.smemStages) {
            uint32_t* dst = c.llrp() + q * c.P.slotW + llr_woff(s);
            if (ISG) scl_g_words<true, false>(src, ss, chan, lb, dst, 1, s, g, G);
            else scl_f_words<true, false>(src, ss, chan, dst, 1, s, g, G);
        } else {
            uint32_t* dst = c.spill + scl_spill_woff(c, s) + q;
            if (ISG) scl_g_words<true, true>(src, ss, chan, lb, dst, 32, s, g, G);
            else scl_f_words<true, true>(src, ss, chan, dst, 32, s, g, G);
        }
    }
    scl_set_own(c.llrPtr, s, c.gl);
    __syncwarp();
}

template <int L>
POLAR_DEV void scl_op_combine(SclCtx& c, int s, int lgP, bool isRight)
{
    if (s + 1 == c.P.lgN) return; // the root is assembled by the epilogue, straight from its two halves
    const int lgG = Log2<L>::v - lgP, G = 1 << lgG, q = c.gl >> lgG, g = c.gl & (G - 1);
    const BitBlk left = scl_bits_a(c, scl_slot_of(c, c.lbPtr, s, q, lgG), s);
    const BitBlk right = scl_bits_b(c, q, s);
    const BitBlk out = scl_bits_out(c, q, s + 1, isRight);
    if (s >= 5) {
        const int hw = 1 << (s - 5);
        for (int j = g; j < hw; j += G) {
            const uint32_t l = left.ld(j), r = right.ld(j);
            out.st(j, l ^ r);
            out.st(j + hw, r);
        }
    } else if (g == 0) {
        const uint32_t l = left.ld(0), r = right.ld(0);
        out.st(0, (l ^ r) | (r << (1 << s)));
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
            uint32_t m = uabs4(src(w));
            m -= (m >> 7) & 0x01010101u; // |max(x,-127)| per byte
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                uint32_t key = prmt(m, (uint32_t)(4 * w + e), 0x7054u | (e << 8)); // (|x_e| << 16) | position
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
    uint32_t m0 = uabs4(src(0));
    m0 -= (m0 >> 7) & 0x01010101u;
    int ha[4], cur[4];
    bool alive[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        ha[e] = (int)((m0 >> (8 * e)) & 0xFFu);
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
// so ties go to the lower current position).  Lane q < P of the group contributes the candidates at list
// positions q*F .. q*F+F-1, F = 2^lgF <= FMAX.  Out (lane j < nOut of the group): metric and original
// candidate index of list entry j.
// The keys (metric, 255 - current position) live in the frame's select scratch, indexed by position (a lane's F
// keys are one or two 16-byte rows); every lane only carries the maximum of its own F keys.  Step i: group maximum -> winner position w; the element at position i moves to w (key rewritten with
// its new position), position i is final; the two owner lanes refresh their maxima.  The swap history
// w_0 .. w_i then tells every output lane which original candidate it received.
template <int L, int FMAX>
POLAR_DEV void scl_select(const SclCtx& c, const int (&cand)[FMAX], int lgF, int lgP, int nOut, int& outMet, int& outIdx)
{
    const int gl = c.gl, F = 1 << lgF;
    uint32_t* ks = c.selp();          // L * 8 key words, ks[position]
    uint32_t* hs = c.selp() + 8 * L;  // L history words
    uint32_t* mine = ks + (gl << lgF);
    constexpr bool V4 = (FMAX >= 4 && L >= 4), V2 = (FMAX == 2 && L >= 2); // the lane's row is 16 / 8 byte aligned
    const bool act = gl < (1 << lgP);
    uint32_t key[FMAX];
    uint32_t lmax = 0u;
#pragma unroll
    for (int k = 0; k < FMAX; ++k) {
        const int pos = (gl << lgF) + k;
        key[k] = (act && k < F) ? (((uint32_t)(cand[k] + SCL_METRIC_BIAS) << 8) | (uint32_t)(255 - pos)) : 0u;
        lmax = max(lmax, key[k]);
    }
    if (V4) {
#pragma unroll
        for (int k = 0; k < FMAX; k += 4)
            if (k < F) *reinterpret_cast<uint4*>(mine + k) = make_uint4(key[k], key[k + 1], key[k + 2], key[k + 3]);
    } else if (V2) {
        *reinterpret_cast<uint2*>(mine) = make_uint2(key[0], key[1]);
    } else {
#pragma unroll
        for (int k = 0; k < FMAX; ++k)
            if (k < F) mine[k] = key[k];
    }
    __syncwarp();
    outMet = 0;
    int myW = 0;
#pragma unroll 1
    for (int i = 0; i < nOut; ++i) {
        const uint32_t best = list_max<L>(lmax);
        const int pos = 255 - (int)(best & 255u), wl = pos >> lgF, il = i >> lgF;
        if (gl == i) {
            outMet = (int)(best >> 8) - SCL_METRIC_BIAS;
            myW = pos;
        }
        const uint32_t ikey = ks[i]; // the element now at position i
        __syncwarp();
        if (gl == wl) ks[pos] = (ikey & ~255u) | (uint32_t)(255 - pos); // ... moves to the winner's place
        if (gl == il) ks[i] = 0u;                                       // position i is final
        if (gl == 0) hs[i] = (uint32_t)pos;
        __syncwarp();
        if (gl == wl || gl == il) {
            lmax = 0u;
            if (V4) {
#pragma unroll
                for (int k = 0; k < FMAX; k += 4) {
                    if (k < F) {
                        const uint4 v = *reinterpret_cast<const uint4*>(mine + k);
                        lmax = max(max(lmax, v.x), max(max(v.y, v.z), v.w));
                    }
                }
            } else if (V2) {
                const uint2 v = *reinterpret_cast<const uint2*>(mine);
                lmax = max(v.x, v.y);
            } else {
#pragma unroll
                for (int k = 0; k < FMAX; ++k)
                    if (k < F) lmax = max(lmax, mine[k]);
            }
        }
    }
    // which original candidate sits at position gl: follow the swaps backwards (the element found at w_t at step
    // t came from position t' if an earlier step t' displaced its element to w_t)
    int p = myW;
    for (int t = nOut - 2; t >= 0; --t) {
        const int wt = (int)hs[t];
        if (t < gl && wt == p) p = t;
    }
    outIdx = p;
    __syncwarp();
}

// Lists of at most 8 paths (M = P * F <= 64 candidates): the 32-bit key has room for the metric (|m| <= 128 N <= 2^19), the
// CURRENT position (6 bits, inverted so that ties go to the lower position) and the ORIGINAL candidate index (6 bits), so
// the winner of a step directly names the candidate it came from: no swap history, no walk back.  Steps are unrolled
// (compile-time scratch addresses), the routine is compiled once per (L, fan-out) and called from every leaf handler.
// Returns the key of list entry gl: metric = (key >> 12) - SCL_K32_BIAS, original candidate = key & 63.
static constexpr int SCL_K32_BIAS = (1 << 20) - 1;
template <int L, int LGF>
__device__ __noinline__ uint32_t scl_select_k32(uint32_t* ks, int gl, int lgP, int nOut, int c0, int c1, int c2, int c3, int c4,
                                                int c5, int c6, int c7)
{
    static_assert(L <= 8, "32-bit keys carry 6 + 6 position bits");
    constexpr int F = 1 << LGF;
    const int cand[8] = { c0, c1, c2, c3, c4, c5, c6, c7 };
    uint32_t* mine = ks + (gl << LGF);
    const bool act = gl < (1 << lgP);
    uint32_t key[F];
    uint32_t lmax = 0u;
#pragma unroll
    for (int k = 0; k < F; ++k) {
        const uint32_t pos = (uint32_t)(gl << LGF) + k;
        key[k] = act ? (((uint32_t)(cand[k] + SCL_K32_BIAS) << 12) | ((63u - pos) << 6) | pos) : 0u;
        lmax = max(lmax, key[k]);
    }
    // the frame's scratch is aligned to L words (see setup_decoder_geometry), a lane's row to min(L, F) words
    constexpr int VW = (L < F ? L : F) < 4 ? (L < F ? L : F) : 4;
    if constexpr (VW == 4) {
#pragma unroll
        for (int k = 0; k < F; k += 4) *reinterpret_cast<uint4*>(mine + k) = make_uint4(key[k], key[k + 1], key[k + 2], key[k + 3]);
    } else if constexpr (VW == 2) {
#pragma unroll
        for (int k = 0; k < F; k += 2) *reinterpret_cast<uint2*>(mine + k) = make_uint2(key[k], key[k + 1]);
    } else {
#pragma unroll
        for (int k = 0; k < F; ++k) mine[k] = key[k];
    }
    __syncwarp();
    uint32_t outKey = 0u;
#pragma unroll
    for (int i = 0; i < L; ++i) {
        if (i >= nOut) break;
        const uint32_t best = list_max<L>(lmax);
        if (gl == i) outKey = best;
        const uint32_t pf = best & 0xFC0u; // (63 - position of the winner) << 6
        const int pos = 63 - (int)(pf >> 6), wl = pos >> LGF;
        const int il = i >> LGF;
        const uint32_t ikey = ks[i]; // the element now at position i ...
        __syncwarp();
        if (gl == wl) ks[pos] = (ikey & ~0xFC0u) | pf; // ... moves to the winner's place
        if (gl == il) ks[i] = 0u;                       // position i is final
        __syncwarp();
        if (gl == wl || gl == il) {
            lmax = 0u;
            if constexpr (VW == 4) {
#pragma unroll
                for (int k = 0; k < F; k += 4) {
                    const uint4 a = *reinterpret_cast<const uint4*>(mine + k);
                    lmax = max(max(lmax, a.x), max(max(a.y, a.z), a.w));
                }
            } else if constexpr (VW == 2) {
#pragma unroll
                for (int k = 0; k < F; k += 2) {
                    const uint2 a = *reinterpret_cast<const uint2*>(mine + k);
                    lmax = max(lmax, max(a.x, a.y));
                }
            } else {
#pragma unroll
                for (int k = 0; k < F; ++k) lmax = max(lmax, mine[k]);
            }
        }
    }
    __syncwarp();
    return outKey;
}

// Candidate pruning for any list size: 32-bit keys with the origin inside for L <= 8, the swap-history version above else.
template <int L, int FMAX>
POLAR_DEV void scl_prune(const SclCtx& c, const int (&cand)[FMAX], int lgF, int lgP, int nOut, int& outMet, int& outIdx)
{
    if constexpr (L <= 8) {
        uint32_t k;
        if constexpr (FMAX == 2) {
            k = scl_select_k32<L, 1>(c.selp(), c.gl, lgP, nOut, cand[0], cand[1], 0, 0, 0, 0, 0, 0);
        } else {
            if (lgF == 3) k = scl_select_k32<L, 3>(c.selp(), c.gl, lgP, nOut, cand[0], cand[1], cand[2], cand[3], cand[4], cand[5], cand[6], cand[7]);
            else k = scl_select_k32<L, 2>(c.selp(), c.gl, lgP, nOut, cand[0], cand[1], cand[2], cand[3], 0, 0, 0, 0);
        }
        outMet = (int)(k >> 12) - SCL_K32_BIAS;
        outIdx = (int)(k & 63u);
    } else {
        scl_select<L, FMAX>(c, cand, lgF, lgP, nOut, outMet, outIdx);
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
    const int pun = lgG ? __shfl_sync(FULL_MASK, acc, c.fb + ((c.gl << lgG) & (L - 1))) : acc;
    if (c.gl < (1 << lgP)) c.metric += pun;
    const BitBlk out = scl_bits_out(c, q, s, isRight);
    const int nbw = s >= 5 ? 1 << (s - 5) : 1;
    for (int j = g; j < nbw; j += G) out.st(j, 0u);
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
    if (lgG) { // lane q (the path's state) fetches what the path's lane group computed
        const int srcLane = c.fb + ((c.gl << lgG) & (L - 1));
        neg = __shfl_sync(FULL_MASK, neg, srcLane);
        pos = __shfl_sync(FULL_MASK, pos, srcLane);
    }
    const int cand[2] = { c.metric + neg, c.metric - pos };
    const int lgPo = min(lgL, lgP + 1);
    int outMet, outIdx;
    scl_prune<L, 2>(c, cand, 1, lgP, 1 << lgPo, outMet, outIdx);
    scl_adopt(c, outMet, (outIdx >> 1) & (L - 1));
    const int lgGo = lgL - lgPo, Go = 1 << lgGo, qo = c.gl >> lgGo, go = c.gl & (Go - 1);
    const int kq = lgGo ? __shfl_sync(FULL_MASK, outIdx & 1, c.fb + qo) : (outIdx & 1);
    const BitBlk out = scl_bits_out(c, qo, s, isRight);
    const int nbw = s >= 5 ? 1 << (s - 5) : 1;
    const uint32_t full = s >= 5 ? 0xFFFFFFFFu : (1u << (1 << s)) - 1u;
    for (int j = go; j < nbw; j += Go) out.st(j, kq ? full : 0u);
    scl_finish_leaf(c, s, isRight);
}

// Shared tail of the R1 / SPC leaves: write sign(llr) with the chosen flips as the new path's bits.
template <int L>
POLAR_DEV void scl_write_flipped(SclCtx& c, int s, int lgPo, bool isRight, uint32_t flipMask, uint32_t pk01,
                                 uint32_t pk23)
{
    const int lgGo = Log2<L>::v - lgPo, Go = 1 << lgGo, qo = c.gl >> lgGo, go = c.gl & (Go - 1);
    uint32_t fm = flipMask, p01 = pk01, p23 = pk23;
    if (lgGo) { // the path's lane group fetches the decision from lane qo
        fm = __shfl_sync(FULL_MASK, fm, c.fb + qo);
        p01 = __shfl_sync(FULL_MASK, p01, c.fb + qo);
        p23 = __shfl_sync(FULL_MASK, p23, c.fb + qo);
    }
    const Words srcn = scl_llr_block(c, s, scl_slot_of(c, c.llrPtr, s, qo, lgGo)); // through the adopted vector
    const BitBlk out = scl_bits_out(c, qo, s, isRight);
    const int nbw = s >= 5 ? 1 << (s - 5) : 1;
    const uint32_t f[4] = { p01 & 0xFFFFu, p01 >> 16, p23 & 0xFFFFu, p23 >> 16 };
    for (int j = go; j < nbw; j += Go) {
        uint32_t w = scl_sign_word(srcn, s, j);
#pragma unroll
        for (int t = 0; t < 4; ++t)
            if (((fm >> t) & 1u) && (int)(f[t] >> 5) == j) w ^= 1u << (f[t] & 31u);
        out.st(j, w);
    }
    scl_finish_leaf(c, s, isRight);
}

// RateOne (scl_fip_char.cpp:423-505) and SPC (:583-726) share one routine: both look for the weakest LLRs,
// fork into 4 / 8 candidates that flip subsets of them, prune, and write sign(llr) with the chosen flips.
template <int L>
POLAR_DEV void scl_leaf_flip(SclCtx& c, int s, int lgP, bool isRight, bool isSpc)
{
    constexpr int lgL = Log2<L>::v;
    const int lgG = lgL - lgP, G = 1 << lgG, q = c.gl >> lgG, g = c.gl & (G - 1);
    const Words src = scl_llr_block(c, s, scl_slot_of(c, c.llrPtr, s, q, lgG));
    uint32_t px = 0;
    if (isSpc) {
        const int words = 1 << (s - 2); // an SPC leaf has at least 4 elements
        for (int w = g; w < words; w += G) px ^= src(w);
    }
    px = group_xor(px, lgG);
    int par = __popc(px & 0x80808080u) & 1;
    int val[4] = { 0, 0, 0, 0 }, idx[4] = { 0, 0, 0, 0 };
    if (isSpc) {
        scl_weak_search<4>(src, s, lgG, g, val, idx);
    } else { // a rate-1 leaf only needs the two weakest
        int v2[2], i2[2];
        scl_weak_search<2>(src, s, lgG, g, v2, i2);
        val[0] = v2[0]; val[1] = v2[1];
        idx[0] = i2[0]; idx[1] = i2[1];
    }
    int a0 = val[0], a1 = val[1], a2 = val[2], a3 = val[3];
    uint32_t pk01 = (uint32_t)idx[0] | ((uint32_t)idx[1] << 16), pk23 = (uint32_t)idx[2] | ((uint32_t)idx[3] << 16);
    if (lgG) { // lane q (the path's state) fetches what the path's lane group computed
        const int srcLane = c.fb + ((c.gl << lgG) & (L - 1));
        a0 = __shfl_sync(FULL_MASK, a0, srcLane);
        a1 = __shfl_sync(FULL_MASK, a1, srcLane);
        a2 = __shfl_sync(FULL_MASK, a2, srcLane);
        a3 = __shfl_sync(FULL_MASK, a3, srcLane);
        pk01 = __shfl_sync(FULL_MASK, pk01, srcLane);
        pk23 = __shfl_sync(FULL_MASK, pk23, srcLane);
        par = __shfl_sync(FULL_MASK, par, srcLane);
    }
    int m = c.metric, w = 0;
    int cand[8];
    if (isSpc) {
        if (par) m -= a0; else w = a0;
        cand[0] = m;           cand[1] = m - w - a1;  cand[2] = m - w - a2;  cand[3] = m - w - a3;
        cand[4] = m - a1 - a2; cand[5] = m - a1 - a3; cand[6] = m - a2 - a3; cand[7] = m - w - a1 - a2 - a3;
    } else {
        cand[0] = m; cand[1] = m - a0; cand[2] = m - a1; cand[3] = m - a0 - a1;
        cand[4] = cand[5] = cand[6] = cand[7] = 0;
    }
    const int lgF = isSpc ? 3 : 2;
    const int lgPo = min(lgL, lgP + lgF);
    int outMet, outIdx;
    scl_prune<L, 8>(c, cand, lgF, lgP, 1 << lgPo, outMet, outIdx);
    const int srcPath = (outIdx >> lgF) & (L - 1), kind = outIdx & ((1 << lgF) - 1);
    const uint32_t s01 = __shfl_sync(FULL_MASK, pk01, c.fb + srcPath), s23 = __shfl_sync(FULL_MASK, pk23, c.fb + srcPath);
    const int sp = __shfl_sync(FULL_MASK, par, c.fb + srcPath);
    // flip sets over (i0,i1,i2,i3), one nibble per candidate kind (scl_fip_char.cpp:641-692; :451-470 for rate 1)
    const uint32_t table = isSpc ? (sp ? 0xEDB78421u : 0xFCA69530u) : 0x3210u;
    const uint32_t fm = (table >> (4 * kind)) & 15u;
    scl_adopt(c, outMet, srcPath);
    scl_write_flipped<L>(c, s, lgPo, isRight, fm, s01, s23);
}

// extractBestPath (scl_fip_char.cpp:816-855, scl_avx_float.cpp:711-750): first path in list order that passes
// the detector, else the most likely one.  The whole warp works on one frame of the quad at a time.
template <int L>
POLAR_DEV void scl_extract_best_path(const DecodeParams& p, const SclCtx& c, uint32_t* base, uint32_t* scr, int lane,
                                     bool live, long long quad, int lgPfinal)
{
    constexpr int FPW = 32 / L;
    const int P = 1 << lgPfinal;
    uint32_t* outw = scr;
    uint32_t* cw = scr + ((p.K + 31) >> 5);
    const int top = p.lgN - 1, nW = max(p.N >> 5, 1);
    const uint32_t slotA = (uint32_t)(c.lbPtr >> (5 * top)) & 31u; // where this lane's path keeps its left half
    for (int f = 0; f < FPW && live; ++f) {
        const long long frame = quad * FPW + f;
        if (frame >= p.nFrames) break;
        uint32_t* fbits = base + f * p.frameW + L * p.slotW;
        uint32_t* fspill = c.spill - c.fb + f * L; // first scratch column of frame f
        bool okFlag = false;
        const bool handBack = p.cand != nullptr; // every surviving path goes to the host, which owns the detector
        for (int path = 0; path < P; ++path) {
            // codeword estimate of this path: the root leaf's bits, or [left ^ right | right] of the two halves
            const uint32_t sa = __shfl_sync(FULL_MASK, slotA, f * L + path);
            if (p.rootLeaf) {
                const BitBlk root = scl_bits_a_at(c, fbits, fspill, path, p.lgN);
                for (int w = lane; w < nW; w += 32) cw[w] = root.ld(w);
            } else {
                const BitBlk left = scl_bits_a_at(c, fbits, fspill, sa, top);
                const BitBlk right = scl_bits_b_at(c, fbits, fspill, path, top);
                if (top >= 5) {
                    const int hw = 1 << (top - 5);
                    for (int w = lane; w < nW; w += 32) cw[w] = w < hw ? (left.ld(w) ^ right.ld(w)) : right.ld(w - hw);
                } else if (lane == 0) {
                    cw[0] = (left.ld(0) ^ right.ld(0)) | (right.ld(0) << (1 << top));
                }
            }
            __syncwarp();
            if (!p.systematic) warp_polar_transform(cw, p.N, lane);
            const uint32_t syn = gather_info_and_syndrome(p, cw, outw, lane);
            if (handBack) {
                uint8_t* dst = p.cand + ((size_t)frame * L + path) * p.kb;
                const uint8_t* srcb = reinterpret_cast<const uint8_t*>(outw);
                for (int i = lane; i < p.kb; i += 32) dst[i] = srcb[i];
                __syncwarp();
                continue;
            }
            if (syn == 0u) { okFlag = true; break; }
            if (!p.crcBits || (path == 0 && P == 1)) break;
            // remember path 0's decision in case nobody passes (scl_fip_char.cpp:848-853)
            if (path == 0)
                for (int w = lane; w < ((p.K + 31) >> 5); w += 32) cw[nW + w] = outw[w];
            __syncwarp();
        }
        if (!okFlag && P > 1 && p.crcBits) { // fall back to the most likely path
            for (int w = lane; w < ((p.K + 31) >> 5); w += 32) outw[w] = cw[nW + w];
            __syncwarp();
        }
        if (handBack) {
            if (lane == 0) p.ok[frame] = (uint8_t)P;
            __syncwarp();
            continue;
        }
        store_info(p, frame, outw, lane);
        if (lane == 0) p.ok[frame] = okFlag ? 1 : 0;
        __syncwarp();
    }
}

// ---------------------------------------------------------------- register stages
// Most nodes of a polar tree are small: at N = 1024, L = 8 (5G set) 171 of the 246 F / G / combine steps and 76 of the
// 83 leaves work on blocks of at most 16 elements, always with lane = path.  Those blocks never touch shared memory:
// a stage-s block (s <= 4) is 2^s int8 = at most four words of the owning lane's registers, its partial sums a bit
// field of one register.  Every handler below is compiled per stage (compile-time word counts and register indices),
// so a small step is its butterflies and nothing else; a fork moves the live register blocks with the lazy-copy
// vectors (one shuffle per word).  Semantics: scl_fip_char.cpp:315-385 (inner nodes), :387-726 (leaves).
template <int S> struct RegStage {
    static_assert(S >= 1 && S <= 4, "register stages are 1..4");
    static constexpr int W = S >= 2 ? 1 << (S - 2) : 1;             // words of the block
    static constexpr int OFF = S == 4 ? 4 : S == 3 ? 2 : S == 2 ? 1 : 0;
    static constexpr int NB = 1 << S;                                // elements = bits
    static constexpr uint32_t MASK = (1u << NB) - 1u;
    static constexpr uint32_t EMASK = S == 1 ? 0xFFFFu : 0xFFFFFFFFu; // valid bytes of word 0
};

// parent block (stage S + 1) of every path, split in halves: registers, or shared / spill / channel through the path's slot
template <int S>
POLAR_DEV void rs_load_parent(const SclCtx& c, uint32_t (&a)[RegStage<S>::W], uint32_t (&b)[RegStage<S>::W])
{
    constexpr int W = RegStage<S>::W;
    bool inRegs = false;
    if constexpr (S < 4) inRegs = S + 1 <= c.P.regStages;
    if (inRegs) {
        if constexpr (S == 1) {
            a[0] = c.r[1];
            b[0] = c.r[1] >> 16;
        } else if constexpr (S < 4) {
            constexpr int PO = RegStage<S + 1>::OFF;
#pragma unroll
            for (int w = 0; w < W; ++w) { a[w] = c.r[PO + w]; b[w] = c.r[PO + W + w]; }
        }
    } else {
        const Words src = scl_llr_block(c, S + 1, (uint32_t)(c.llrPtr >> (5 * (S + 1))) & 31u);
        if constexpr (S == 1) {
            a[0] = src(0);
            b[0] = a[0] >> 16;
        } else {
#pragma unroll
            for (int w = 0; w < W; ++w) { a[w] = src(w); b[w] = src(w + W); }
        }
    }
}

template <int S, bool ISG>
POLAR_DEV void rs_op_fg(SclCtx& c)
{
    constexpr int W = RegStage<S>::W, OFF = RegStage<S>::OFF;
    uint32_t a[W], b[W];
    rs_load_parent<S>(c, a, b);
    if constexpr (ISG) {
        const uint32_t lb = c.ba >> RegStage<S>::NB;
#pragma unroll
        for (int w = 0; w < W; ++w) c.r[OFF + w] = g4(a[w], b[w], nibble_to_msb(S == 1 ? (lb & 3u) : (lb >> (4 * w))));
    } else {
#pragma unroll
        for (int w = 0; w < W; ++w) c.r[OFF + w] = f4(a[w], b[w]);
    }
}

// bits of a finished stage-S node: left children wait in ba, right children go to the ping-pong pair
template <int S>
POLAR_DEV void rs_put_bits(SclCtx& c, uint32_t v, bool isRight)
{
    constexpr uint32_t M = RegStage<S>::MASK;
    constexpr int NB = RegStage<S>::NB;
    if (isRight) c.bb[S & 1] = v & M;
    else c.ba = (c.ba & ~(M << NB)) | ((v & M) << NB);
}

template <int S>
POLAR_DEV void rs_op_combine(SclCtx& c, bool isRight)
{
    constexpr int NB = RegStage<S>::NB;
    const uint32_t l = (c.ba >> NB) & RegStage<S>::MASK, r = c.bb[S & 1];
    const uint32_t out = (l ^ r) | (r << NB);
    bool toRegs = false;
    if constexpr (S < 4) toRegs = S + 1 <= c.P.regStages;
    if (toRegs) {
        if constexpr (S < 4) rs_put_bits<S + 1>(c, out, isRight);
    } else { // the parent stage lives in shared memory (one word: S + 1 <= 5)
        scl_bits_out(c, (uint32_t)c.gl, S + 1, isRight).st(0, out);
        if (!isRight) scl_set_own(c.lbPtr, S + 1, c.gl);
        __syncwarp();
    }
}

// After a fork: adopt the parent's lazy-copy vectors, metric and the register blocks that are still live (stages > S).
template <int S>
POLAR_DEV void rs_adopt(SclCtx& c, int outMet, int srcPath)
{
    const int src = c.fb + srcPath;
    c.metric = outMet;
    c.llrPtr = shfl64(c.llrPtr, src);
    c.lbPtr = shfl64(c.lbPtr, src);
    c.ba = __shfl_sync(FULL_MASK, c.ba, src);
#pragma unroll
    for (int i = RegStage<S>::OFF + RegStage<S>::W; i < 8; ++i) // blocks of the stages above S sit behind it
        if (i >= (S == 1 ? 1 : 0)) c.r[i] = __shfl_sync(FULL_MASK, c.r[i], src);
}

template <int L, int S>
POLAR_DEV void rs_leaf_r0(SclCtx& c, bool isRight) // scl_fip_char.cpp:387-421
{
    constexpr int W = RegStage<S>::W, OFF = RegStage<S>::OFF;
    int acc = 0;
#pragma unroll
    for (int w = 0; w < W; ++w) acc = sum_neg4(c.r[OFF + w] & RegStage<S>::EMASK, acc);
    c.metric += acc;
    rs_put_bits<S>(c, 0u, isRight);
}

template <int L, int S>
POLAR_DEV void rs_leaf_rep(SclCtx& c, int lgP, bool isRight) // scl_fip_char.cpp:508-581
{
    constexpr int lgL = Log2<L>::v, W = RegStage<S>::W, OFF = RegStage<S>::OFF;
    int neg = 0, pos = 0;
#pragma unroll
    for (int w = 0; w < W; ++w) {
        const uint32_t x = c.r[OFF + w] & RegStage<S>::EMASK;
        neg = sum_neg4(x, neg);
        pos = sum_pos4(x, pos);
    }
    const int cand[2] = { c.metric + neg, c.metric - pos };
    const int lgPo = min(lgL, lgP + 1);
    int outMet, outIdx;
    scl_prune<L, 2>(c, cand, 1, lgP, 1 << lgPo, outMet, outIdx);
    rs_adopt<S>(c, outMet, (outIdx >> 1) & (L - 1));
    rs_put_bits<S>(c, (outIdx & 1) ? 0xFFFFFFFFu : 0u, isRight);
}

// findWeakLlrs (arrayfuncs.h:209-231) on a block held in registers: the K smallest |max(llr,-127)| in the reference's
// order, with their original positions.  Same emulation as scl_weak_search (only elements 0..3 can be displaced by
// the swaps), fully unrolled.
template <int S, int K>
POLAR_DEV void rs_weak_search(const uint32_t (&x)[RegStage<S>::W], int (&val)[K], int (&idx)[K])
{
    constexpr int W = RegStage<S>::W, n = RegStage<S>::NB;
    uint32_t m[W];
#pragma unroll
    for (int w = 0; w < W; ++w) {
        m[w] = uabs4(x[w]);
        m[w] -= (m[w] >> 7) & 0x01010101u; // |max(x,-127)| per byte
    }
    uint32_t tail[K];
#pragma unroll
    for (int r = 0; r < K; ++r) tail[r] = 0xFFFFFFFFu;
#pragma unroll
    for (int w = 1; w < W; ++w) {
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            uint32_t key = prmt(m[w], (uint32_t)(4 * w + e), 0x7054u | (e << 8)); // (|x_e| << 16) | position
#pragma unroll
            for (int r = 0; r < K; ++r) {
                const uint32_t lo = min(tail[r], key);
                key = max(tail[r], key);
                tail[r] = lo;
            }
        }
    }
    int ha[4], cur[4];
    bool alive[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        ha[e] = (int)((m[0] >> (8 * e)) & 0xFFu);
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

// RateOne (scl_fip_char.cpp:423-505) and SPC (:583-726) on a register block
template <int L, int S, bool SPC>
POLAR_DEV void rs_leaf_flip(SclCtx& c, int lgP, bool isRight)
{
    constexpr int lgL = Log2<L>::v, W = RegStage<S>::W, OFF = RegStage<S>::OFF, lgF = SPC ? 3 : 2;
    uint32_t x[W];
    uint32_t px = 0, sgn = 0;
#pragma unroll
    for (int w = 0; w < W; ++w) {
        x[w] = c.r[OFF + w];
        if (S == 1) x[w] = (x[w] & 0xFFFFu) | 0x7F7F0000u; // the two missing elements never win the search
        px ^= x[w];
        sgn |= sign_nibble(x[w]) << (4 * w);
    }
    int par = __popc(px & 0x80808080u) & 1;
    int val[4] = { 0, 0, 0, 0 }, idx[4] = { 0, 0, 0, 0 };
    if (SPC) {
        rs_weak_search<S, 4>(x, val, idx);
    } else {
        int v2[2], i2[2];
        rs_weak_search<S, 2>(x, v2, i2);
        val[0] = v2[0]; val[1] = v2[1];
        idx[0] = i2[0]; idx[1] = i2[1];
    }
    const int a0 = val[0], a1 = val[1], a2 = val[2], a3 = val[3];
    int m = c.metric, w = 0;
    int cand[8];
    if (SPC) {
        if (par) m -= a0; else w = a0;
        cand[0] = m;           cand[1] = m - w - a1;  cand[2] = m - w - a2;  cand[3] = m - w - a3;
        cand[4] = m - a1 - a2; cand[5] = m - a1 - a3; cand[6] = m - a2 - a3; cand[7] = m - w - a1 - a2 - a3;
    } else {
        cand[0] = m; cand[1] = m - a0; cand[2] = m - a1; cand[3] = m - a0 - a1;
        cand[4] = cand[5] = cand[6] = cand[7] = 0;
    }
    const int lgPo = min(lgL, lgP + lgF);
    int outMet, outIdx;
    scl_prune<L, 8>(c, cand, lgF, lgP, 1 << lgPo, outMet, outIdx);
    const int srcPath = (outIdx >> lgF) & (L - 1), kind = outIdx & ((1 << lgF) - 1);
    // what the new path needs from its parent: the four weak positions (4 bits each), the parity and the sign bits
    const uint32_t mine = (uint32_t)idx[0] | ((uint32_t)idx[1] << 4) | ((uint32_t)idx[2] << 8) | ((uint32_t)idx[3] << 12) |
                          ((uint32_t)par << 15 << 1) | (sgn << 17);
    // (sgn has at most 16 bits, but 17 + 16 > 32: a 16-element block sends its top sign bit separately)
    const uint32_t got = __shfl_sync(FULL_MASK, mine, c.fb + srcPath);
    uint32_t v = got >> 17;
    if constexpr (S == 4) v |= (__shfl_sync(FULL_MASK, sgn >> 15, c.fb + srcPath) & 1u) << 15;
    const int sp = (got >> 16) & 1;
    // flip sets over (i0,i1,i2,i3), one nibble per candidate kind (scl_fip_char.cpp:641-692; :451-470 for rate 1)
    const uint32_t table = SPC ? (sp ? 0xEDB78421u : 0xFCA69530u) : 0x3210u;
    const uint32_t fm = (table >> (4 * kind)) & 15u;
#pragma unroll
    for (int t = 0; t < 4; ++t)
        if ((fm >> t) & 1u) v ^= 1u << ((got >> (4 * t)) & 15u);
    rs_adopt<S>(c, outMet, srcPath);
    rs_put_bits<S>(c, v, isRight);
}

// ---------------------------------------------------------------- the kernel
// MAXT: launch bound = register budget (512 threads: 128 registers, 768: 80, 1024: 64); the host picks the variant that
// gives the most resident warps for the code's shared-memory footprint
template <int L, int MAXT>
__global__ void __launch_bounds__(MAXT, 1) scl_i8_kernel(const __grid_constant__ DecodeParams p,
                                                     const __grid_constant__ Schedule sched)
{
    constexpr int lgL = Log2<L>::v, FPW = 32 / L;
    extern __shared__ __align__(16) uint32_t smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, warpsPerBlock = blockDim.x >> 5;
    uint32_t* base = smem + warp * p.warpW;
    uint32_t* scr = base + FPW * p.frameW;
    SclCtx c{ p };
    c.gl = lane & (L - 1);
    c.fb = lane & ~(L - 1);
    const int fi = lane >> lgL;
    c.frame = base + fi * p.frameW;
    c.spill = p.spill ? p.spill + (size_t)(blockIdx.x * warpsPerBlock + warp) * p.spillWarpW + c.fb : nullptr;
#pragma unroll
    for (int i = 0; i < 8; ++i) c.r[i] = 0u;
    c.ba = 0u;
    c.bb[0] = c.bb[1] = 0u;

    // Every warp of a block walks the same schedule, so the block iterates together (same trip count for all
    // its warps); with p.lockstep the warps also meet at a barrier before every step: they then fetch the
    // same instruction lines at the same time, which is what keeps this branchy interpreter out of
    // instruction-fetch stalls (ncu: stall_no_instruction).
    const long long nQuads = (p.nFrames + FPW - 1) / FPW;
    for (long long qb = blockIdx.x; qb * warpsPerBlock < nQuads; qb += gridDim.x) {
        const bool live = qb * warpsPerBlock + warp < nQuads; // a spare warp redoes the last quad and drops it
        const long long quad = live ? qb * warpsPerBlock + warp : nQuads - 1;
        const long long myFrame = min(quad * FPW + fi, p.nFrames - 1); // tail lanes redo the last frame
        c.chan = reinterpret_cast<const uint32_t*>(static_cast<const int8_t*>(p.llr) + myFrame * p.N);
        c.metric = 0;
        c.llrPtr = 0ull;
        c.lbPtr = 0ull;
        int lgPfinal = 0;
        const uint32_t lockMask = p.lockstep ? (uint32_t)p.lockstep - 1u : 0x7FFFFFFFu;
        for (int pc = 0;; ++pc) {
            // op word of the list decoders: bits 31..16 hold a dense handler id written by the host (scl_handler_id): one
            // jump table instead of nested switches on (code, stage)
            const uint32_t op = sched.ops[pc];
            const uint32_t hid = op >> 16;
            if (hid == SCL_H_END) break;
            const int s = (int)POLAR_OP_STAGE(op), lgP = (int)POLAR_OP_LGP(op);
            const bool isRight = (POLAR_OP_FLAGS(op) & OPF_RIGHT) != 0;
            if ((pc & lockMask) == 0) __syncthreads(); // lockstep = 2^k: a barrier every 2^k steps
#define POLAR_RS_CASES(BASE, CALL)                \
    case BASE + 0: { constexpr int S = 1; CALL; } break; \
    case BASE + 1: { constexpr int S = 2; CALL; } break; \
    case BASE + 2: { constexpr int S = 3; CALL; } break; \
    case BASE + 3: { constexpr int S = 4; CALL; } break;
            switch (hid) {
            // register stages: handlers compiled per stage, lane = path
            POLAR_RS_CASES(SCL_H_RS_F, (rs_op_fg<S, false>(c)))
            POLAR_RS_CASES(SCL_H_RS_G, (rs_op_fg<S, true>(c)))
            POLAR_RS_CASES(SCL_H_RS_C, (rs_op_combine<S>(c, isRight)))
            POLAR_RS_CASES(SCL_H_RS_R0, (rs_leaf_r0<L, S>(c, isRight)); lgPfinal = lgP)
            POLAR_RS_CASES(SCL_H_RS_REP, (rs_leaf_rep<L, S>(c, lgP, isRight)); lgPfinal = min(lgL, lgP + 1))
            POLAR_RS_CASES(SCL_H_RS_R1, (rs_leaf_flip<L, S, false>(c, lgP, isRight)); lgPfinal = min(lgL, lgP + 2))
            case SCL_H_RS_SPC + 1: rs_leaf_flip<L, 2, true>(c, lgP, isRight); lgPfinal = min(lgL, lgP + 3); break; // an SPC leaf has >= 4 elements
            case SCL_H_RS_SPC + 2: rs_leaf_flip<L, 3, true>(c, lgP, isRight); lgPfinal = min(lgL, lgP + 3); break;
            case SCL_H_RS_SPC + 3: rs_leaf_flip<L, 4, true>(c, lgP, isRight); lgPfinal = min(lgL, lgP + 3); break;
            // stages in shared memory / the L2 scratch / the channel row
            case SCL_H_F: scl_op_fg<L, false>(c, s, lgP); break;
            case SCL_H_G: scl_op_fg<L, true>(c, s, lgP); break;
            case SCL_H_C: scl_op_combine<L>(c, s, lgP, isRight); break;
            case SCL_H_R0: scl_leaf_r0<L>(c, s, lgP, isRight); lgPfinal = lgP; break;
            case SCL_H_REP: scl_leaf_rep<L>(c, s, lgP, isRight); lgPfinal = min(lgL, lgP + 1); break;
            case SCL_H_R1: scl_leaf_flip<L>(c, s, lgP, isRight, false); lgPfinal = min(lgL, lgP + 2); break;
            case SCL_H_SPC: scl_leaf_flip<L>(c, s, lgP, isRight, true); lgPfinal = min(lgL, lgP + 3); break;
            default: break;
            }
#undef POLAR_RS_CASES
        }
        scl_extract_best_path<L>(p, c, base, scr, lane, live, quad, lgPfinal);
    }
}

} // namespace polar
