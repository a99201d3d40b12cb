// Fast-SSC int8 decoder for sm_100a: one frame per THREAD, 32 frames per warp in lockstep, the whole node
// program in the constant bank.
//
// Restates PolarCode::Decoding::FastSscFipChar (src/polarcode/decoding/fastssc_fip_char.cpp) bit-exactly,
// including the stepwise-saturating REP sums (avxconvenience.h:92-101,202-210), the "first minimum,
// threshold 127" SPC search (:274-319, avxconvenience.cpp:13-56) and the arithmetic-negate flip that
// silently fails for 0 and -128 (:301).  Partial sums are packed sign bits at absolute code positions.
//
// B200 mapping: the schedule is the same for every frame, so the 32 lanes of a warp never diverge: each
// lane walks its own frame through the node program.  A lane's data sits in shared memory as "rows" of 32
// words (word r of lane l at [r][l]): every access is bank-conflict free whatever the row index.  The channel
// LLRs are never staged: the two top-level steps pull them through a padded 32x33 transpose tile (coalesced
// 128-byte row reads), the second time out of L2.  Decoded bytes leave through a shared-memory image of the
// warp's 32 output rows, i.e. as fully coalesced stores.
#pragma once
#include "decode_common.cuh"
#include "decode_params.h"
#include "primitives.cuh"
#include "schedule.h"

namespace polar {

struct ScCtx {
    const DecodeParams& P; // geometry is read where it is used: kernel parameters are constant-bank operands, not registers
    uint32_t* sm;          // this warp's shared memory
    const uint32_t* chan;  // channel rows (global), N/4 words per frame
    long long frame0, lastFrame;
    uint32_t* spill;       // this warp's rows of the L2-resident scratch (stages >= smemStages), or null
    int lane;
    uint32_t* smLane;      // sm + lane
    uint32_t* spillLane;   // spill + lane, rebased so that row llr_woff(s) is the row of spilled stage s
    // register stages (stages <= P.regStages = min(4, lgN - 2)): LLR blocks of stage 4 / 3 / 2 / 1 in r[4..7] / r[2..3] /
    // r[1] / r[0]; the 32-bit word of partial sums the small nodes are working in sits in `bw` (word index curW, -1: none)
    uint32_t r[8];
    uint32_t bw;
    int curW;
};

#define SC_ROW(c, row) (c).sm[((row) << 5) + (c).lane]

// LLR block of stage s as a pointer to the lane's word 0 (word w at p[w << 5]): rows of the shared-memory
// stage buffers, rows of the spill scratch (same [row][lane] layout, coalesced 128-byte lines), or the staged
// channel (degenerate one-leaf codes).
POLAR_DEV uint32_t* sc_stage(const ScCtx& c, int s)
{
    if (s == c.P.lgN) return c.sm + (c.P.chanRow << 5) + c.lane;
    uint32_t* base = s >= c.P.smemStages ? c.spillLane : c.smLane; // both already offset to the lane's column
    return base + (llr_woff(s) << 5);
}

// Source views of a parent block split in halves a = [0, words), b = [words, 2*words)
struct ScRows { // both halves in the lane's own rows (shared memory or spill)
    const uint32_t* p;
    int words;
    POLAR_DEV uint32_t a(int w) const { return p[w << 5]; }
    POLAR_DEV uint32_t b(int w) const { return p[(w + words) << 5]; }
};
struct ScTiles { // a chunk [w0, w0+cw) of both halves in the transpose tiles
    const uint32_t* ta;
    const uint32_t* tb;
    int w0;
    POLAR_DEV uint32_t a(int w) const { return ta[w - w0]; }
    POLAR_DEV uint32_t b(int w) const { return tb[w - w0]; }
};

// tile[r*33 + i] = channel word (w0 + i) of frame (frame0 + r), i < cw <= 32
POLAR_DEV void sc_load_tile(const ScCtx& c, uint32_t* tile, int w0, int cw)
{
    // asynchronous 4-byte copies straight into the padded tile: all 32 rows of the lane's column in flight at
    // once; the caller waits (cp_async_wait_all) before the __syncwarp that publishes the tile
    if (c.lane < cw) {
        if (c.frame0 + 31 <= c.lastFrame) {
            const uint32_t* src = c.chan + c.frame0 * (c.P.N >> 2) + w0 + c.lane;
#pragma unroll 8
            for (int r = 0; r < 32; ++r) cp_async4(tile + r * 33 + c.lane, src + (long long)r * (c.P.N >> 2));
        } else {
            for (int r = 0; r < 32; ++r) {
                const long long f = min(c.frame0 + r, c.lastFrame);
                cp_async4(tile + r * 33 + c.lane, c.chan + f * (c.P.N >> 2) + w0 + c.lane);
            }
        }
    }
}

// ---- 16-byte variant of the channel tiles (rows of 32 words, 16-byte aligned channel rows): tile row r = frame frame0 + r, its
// 16-byte chunk q stored at chunk position q ^ (r & 7).  A lane copies one chunk of four rows per instruction (8 instead of 32
// copies per tile) and reads its own row back with 128-bit loads: the eight lanes of a quarter warp hit eight different chunk
// positions, i.e. all 32 banks -- no conflicts without the padding that would break the 16-byte alignment.
POLAR_DEV void sc_load_tile_v(const ScCtx& c, uint32_t* tile, int w0)
{
    const int rsub = c.lane >> 3, q = c.lane & 7;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int r = rsub + 4 * k;
        const long long f = min(c.frame0 + r, c.lastFrame);
        cp_async16(tile + r * 32 + ((q ^ (r & 7)) << 2), c.chan + f * (c.P.N >> 2) + w0 + 4 * q);
    }
}
// the 32 words [w0, w0 + 32) of a top-level step out of the 16-byte tiles
template <int CODE>
POLAR_DEV void sc_top_chunk_v(ScCtx& c, const uint32_t* tileA, const uint32_t* tileB, uint32_t* dst, int pos, int w0)
{
    const uint32_t* ra = tileA + c.lane * 32;
    const uint32_t* rb = tileB + c.lane * 32;
    const int sw = c.lane & 7;
    const uint32_t* bits = c.smLane + ((c.P.llrRows + (pos >> 5)) << 5);
#pragma unroll 2
    for (int q = 0; q < 8; ++q) {
        const uint4 a = *reinterpret_cast<const uint4*>(ra + ((q ^ sw) << 2));
        const uint4 b = *reinterpret_cast<const uint4*>(rb + ((q ^ sw) << 2));
        uint32_t nb = 0u; // left bits of the four words (four bits per word)
        if constexpr (CODE == OP_G) nb = bits[((w0 + 4 * q) >> 3) << 5] >> ((q & 1) * 16);
        const uint32_t av[4] = { a.x, a.y, a.z, a.w }, bv[4] = { b.x, b.y, b.z, b.w };
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            uint32_t v;
            if constexpr (CODE == OP_F) v = f4(av[e], bv[e]);
            else if constexpr (CODE == OP_G_0R) v = adds4(av[e], bv[e]);
            else v = g4(av[e], bv[e], nibble_to_msb(nb >> (4 * e)));
            dst[(w0 + 4 * q + e) << 5] = v;
        }
    }
}

// ---- TMA variant of the channel tiles (N >= 256): tile[r * 32 + i] = channel word (w0 + i) of frame (frame0 + r), filled by
// 32 bulk copies of 128 bytes (lane r copies the row segment of frame r; SASS UBLKCP) that complete on the warp's mbarrier.
// Rows are 128-byte aligned, so the transposing read is done ROTATED: at step i lane l takes word (i + l) & 31 of its row,
// which puts the 32 lanes of a step on 32 different banks (a padded row stride would break the 16-byte alignment TMA needs).
POLAR_DEV void sc_tma_fill(const ScCtx& c, uint32_t* tileA, uint32_t* tileB, int w0, int halfW, uint64_t* bar)
{
    fence_proxy_async(); // the tiles alias stage rows written through the generic proxy
    if (c.lane == 0) mbar_expect_tx(bar, 2u * 32u * 128u);
    __syncwarp();
    const long long f = min(c.frame0 + c.lane, c.lastFrame); // tail lanes re-read the last frame
    const uint32_t* row = c.chan + f * (c.P.N >> 2) + w0;
    tma_bulk_g2s(tileA + c.lane * 32, row, 128u, bar);
    tma_bulk_g2s(tileB + c.lane * 32, row + halfW, 128u, bar);
}

template <int CODE>
POLAR_DEV void sc_top_rot(ScCtx& c, const uint32_t* tileA, const uint32_t* tileB, uint32_t* dst, int pos, int w0)
{
    const uint32_t* ra = tileA + c.lane * 32;
    const uint32_t* rb = tileB + c.lane * 32;
    const uint32_t* bits = c.smLane + ((c.P.llrRows + (pos >> 5)) << 5);
#pragma unroll 1
    for (int i0 = 0; i0 < 32; i0 += 8) {
        uint32_t a[8], b[8];
#pragma unroll
        for (int t = 0; t < 8; ++t) {
            const int i = (i0 + t + c.lane) & 31;
            a[t] = ra[i];
            b[t] = rb[i];
        }
#pragma unroll
        for (int t = 0; t < 8; ++t) {
            const int w = w0 + ((i0 + t + c.lane) & 31);
            uint32_t v;
            if constexpr (CODE == OP_F) v = f4(a[t], b[t]);
            else if constexpr (CODE == OP_G_0R) v = adds4(a[t], b[t]);
            else v = g4(a[t], b[t], nibble_to_msb(bits[(w >> 3) << 5] >> ((w & 7) * 4)));
            dst[w << 5] = v;
        }
    }
}

// ---- per-thread steps over the word range [wb, we) of a block of `words` words per half
template <class Src>
POLAR_DEV void sc_f_range(const Src& src, uint32_t* dst, int wb, int we)
{
    int w = wb;
    for (; w + 7 < we; w += 8) { // 16 loads in flight: the parent block may sit in the L2 scratch
        uint32_t a[8], b[8];
#pragma unroll
        for (int t = 0; t < 8; ++t) { a[t] = src.a(w + t); b[t] = src.b(w + t); }
#pragma unroll
        for (int t = 0; t < 8; ++t) dst[(w + t) << 5] = f4(a[t], b[t]);
    }
    for (; w + 1 < we; w += 2) {
        const uint32_t a0 = src.a(w), a1 = src.a(w + 1), b0 = src.b(w), b1 = src.b(w + 1);
        dst[(w) << 5] = f4(a0, b0);
        dst[(w + 1) << 5] = f4(a1, b1);
    }
    for (; w < we; ++w) dst[(w) << 5] = f4(src.a(w), src.b(w));
}

template <class Src>
POLAR_DEV void sc_adds_range(const Src& src, uint32_t* dst, int wb, int we)
{
    int w = wb;
    for (; w + 7 < we; w += 8) {
        uint32_t a[8], b[8];
#pragma unroll
        for (int t = 0; t < 8; ++t) { a[t] = src.a(w + t); b[t] = src.b(w + t); }
#pragma unroll
        for (int t = 0; t < 8; ++t) dst[(w + t) << 5] = adds4(a[t], b[t]);
    }
    for (; w < we; ++w) dst[(w) << 5] = adds4(src.a(w), src.b(w));
}

// left bits of the block start at absolute bit position pos
template <class Src>
POLAR_DEV void sc_g_range(ScCtx& c, const Src& src, uint32_t* dst, int pos, int wb, int we)
{
    int w = wb;
    if (((we - wb) & 7) == 0 && (pos & 31) == 0) { // whole bit words: one load per 8 LLR words
        for (; w < we; w += 8) {
            const uint32_t lw = SC_ROW(c, c.P.llrRows + ((pos + 4 * w) >> 5));
            uint32_t a[8], b[8];
#pragma unroll
            for (int t = 0; t < 8; ++t) { a[t] = src.a(w + t); b[t] = src.b(w + t); }
#pragma unroll
            for (int t = 0; t < 8; ++t) dst[(w + t) << 5] = g4(a[t], b[t], nibble_to_msb(lw >> (4 * t)));
        }
        return;
    }
    for (; w < we; ++w) {
        const int bp = pos + 4 * w;
        const uint32_t nib = SC_ROW(c, c.P.llrRows + (bp >> 5)) >> (bp & 31);
        dst[(w) << 5] = g4(src.a(w), src.b(w), nibble_to_msb(nib));
    }
}

// ROneNode (fastssc_fip_char.cpp:436-474): right = g(a, b, left bits); right bits = sign(right);
// left bits ^= right bits.  h = 2^s >= 32 elements: whole bit words, range in LLR words (multiple of 8).
template <class Src>
POLAR_DEV void sc_g_r1_range(ScCtx& c, const Src& src, int s, int pos, int wb, int we)
{
    const int hw = 1 << (s - 5);
    for (int w = wb; w < we; w += 8) {
        const int j = w >> 3;
        const uint32_t lw = SC_ROW(c, c.P.llrRows + (pos >> 5) + j);
        uint32_t a[8], b[8];
#pragma unroll
        for (int t = 0; t < 8; ++t) { a[t] = src.a(w + t); b[t] = src.b(w + t); }
        uint32_t rw = 0;
#pragma unroll
        for (int t = 0; t < 8; ++t) rw |= sign_nibble(g4(a[t], b[t], nibble_to_msb(lw >> (4 * t)))) << (4 * t);
        SC_ROW(c, c.P.llrRows + (pos >> 5) + hw + j) = rw;
        SC_ROW(c, c.P.llrRows + (pos >> 5) + j) = lw ^ rw;
    }
}
// h < 32: everything inside one bit word
template <class Src>
POLAR_DEV void sc_g_r1_short(ScCtx& c, const Src& src, int s, int pos)
{
    const int h = 1 << s, sh = pos & 31;
    const uint32_t mask = (1u << h) - 1u;
    const uint32_t word = SC_ROW(c, c.P.llrRows + (pos >> 5));
    const uint32_t lbits = (word >> sh) & mask;
    uint32_t rbits = 0;
    if (s >= 2) {
        const int words = 1 << (s - 2); // <= 4
        for (int w = 0; w < words; ++w)
            rbits |= sign_nibble(g4(src.a(w), src.b(w), nibble_to_msb(lbits >> (4 * w)))) << (4 * w);
    } else { // h == 2: both halves in one word
        const uint32_t x = src.a(0);
        rbits = sign_nibble(g4(x, x >> 16, nibble_to_msb(lbits & 3u))) & 3u;
    }
    const uint32_t both = (lbits ^ rbits) | (rbits << h);
    const uint32_t clr = (mask | (mask << h)) << sh;
    SC_ROW(c, c.P.llrRows + (pos >> 5)) = (word & ~clr) | (both << sh);
}

// ---- bit helpers (block of n = 2^s bits at absolute position pos, pos % n == 0)
// sign bits of LLR rows [row, ..) of a block of n values, bit word j
POLAR_DEV uint32_t sc_sign_word(const uint32_t* src, int s, int j)
{
    if (s < 2) return sign_nibble(src[0]) & ((1u << (1 << s)) - 1u);
    const int words = 1 << (s - 2);
    uint32_t w = 0;
#pragma unroll
    for (int t = 0; t < 8; ++t) {
        const int lw = 8 * j + t;
        if (lw < words) w |= sign_nibble(src[(lw) << 5]) << (4 * t);
    }
    return w;
}
// store bit word j of a block (n >= 32: whole word; n < 32: j == 0, field inside a word)
POLAR_DEV void sc_put_bits(ScCtx& c, int s, int pos, int j, uint32_t v)
{
    if (s >= 5) {
        SC_ROW(c, c.P.llrRows + (pos >> 5) + j) = v;
    } else {
        const uint32_t mask = (1u << (1 << s)) - 1u;
        const int sh = pos & 31;
        const uint32_t wd = SC_ROW(c, c.P.llrRows + (pos >> 5));
        SC_ROW(c, c.P.llrRows + (pos >> 5)) = (wd & ~(mask << sh)) | ((v & mask) << sh);
    }
}
POLAR_DEV void sc_fill_bits(ScCtx& c, int s, int pos, uint32_t pattern)
{
    const int nbw = s >= 5 ? 1 << (s - 5) : 1;
    for (int j = 0; j < nbw; ++j) sc_put_bits(c, s, pos, j, pattern);
}
// bits = sign(LLR rows) with an optional flipped position; dup: the same bits again in [pos+n, pos+2n)
POLAR_DEV void sc_put_signs(ScCtx& c, const uint32_t* src, int s, int pos, bool dup, uint32_t flipIdx, bool doFlip)
{
    const int nbw = s >= 5 ? 1 << (s - 5) : 1;
    for (int j = 0; j < nbw; ++j) {
        uint32_t v = sc_sign_word(src, s, j);
        if (doFlip && (int)(flipIdx >> 5) == j) v ^= 1u << (flipIdx & 31u);
        sc_put_bits(c, s, pos, j, v);
        if (dup) sc_put_bits(c, s, pos + (1 << s), j, v);
    }
}

// ---- leaves
// acc[t] (t < 8) = byte lanes 4t..4t+3 of the reference's 32-byte accumulator:
// acc[j] = sat(..sat(sat(0 + x[j]) + x[32+j]).. ) (fastssc_fip_char.cpp:227-232)
POLAR_DEV void sc_rep_acc(const uint32_t* src, int s, uint32_t (&acc)[8])
{
    const int words = s >= 2 ? 1 << (s - 2) : 1;
#pragma unroll
    for (int t = 0; t < 8; ++t) acc[t] = 0u;
    if (s < 2) {
        acc[0] = src[0] & (s == 1 ? 0xFFFFu : 0xFFu);
    } else if (words < 8) {
#pragma unroll
        for (int t = 0; t < 4; ++t)
            if (t < words) acc[t] = src[(t) << 5]; // sat(0 + x) = x
    } else {
#pragma unroll
        for (int t = 0; t < 8; ++t) acc[t] = src[(t) << 5];
        for (int v = 8; v < words; v += 8) {
#pragma unroll
            for (int t = 0; t < 8; ++t) acc[t] = adds4(acc[t], src[(v + t) << 5]);
        }
    }
}

// SPC decision on rows [row, ..) holding n = 2^s values (fastssc_fip_char.cpp:274-319): returns whether a
// bit flips and which.  Long form = threshold search (start 127, strict <), short form = plain first
// minimum with +127 pads.  The flip is an arithmetic negate, so it is void for 0 and -128.
POLAR_DEV bool sc_spc_decide(const uint32_t* src, int s, bool longForm, uint32_t& flipIdx)
{
    const int n = 1 << s, words = n >= 4 ? n >> 2 : 1;
    uint32_t px = 0, best = 0xFFFFFFFFu;
    for (int w = 0; w < words; ++w) {
        uint32_t x = src[(w) << 5];
        if (s < 2) x &= 0xFFFFu;
        px ^= x;
        const uint32_t m = uabs4(x); // |x| as unsigned, |-128| = 128
#pragma unroll
        for (int e = 0; e < 4; ++e)
            if (4 * w + e < n) best = min(best, (((m >> (8 * e)) & 0xFFu) << 16) | (uint32_t)(4 * w + e));
    }
    const bool odd = (__popc(px & 0x80808080u) & 1) != 0;
    uint32_t mv = best >> 16, mi = best & 0xFFFFu;
    bool flip = odd;
    if (longForm) {
        if (mv >= 127u) mi = 0u;
    } else if (mv > 127u && n < 32) {
        flip = false; // a +127 pad wins: the negate lands outside the block
    }
    const uint32_t wv = src[(int)(mi >> 2) << 5];
    const int v = (int)(int8_t)(wv >> (8 * (mi & 3u)));
    if (v == 0 || v == -128) flip = false;
    flipIdx = mi;
    return flip;
}

// ---------------------------------------------------------------- register stages
// 170 of the 234 steps of the N = 1024 K = 512 program work on blocks of at most 16 elements.  Those blocks live in the
// thread's registers (a stage-s block is 2^s int8 = at most four words), their partial sums in one cached word of the
// codeword estimate: the tree is walked left to right, so the small nodes enter a 32-bit word, fill it, and leave it --
// one load and one store per word.  Every handler is compiled per stage (compile-time word counts and register indices).
template <int S> struct ScReg {
    static_assert(S >= 1 && S <= 4, "register stages are 1..4");
    static constexpr int W = S >= 2 ? 1 << (S - 2) : 1;
    static constexpr int OFF = S == 4 ? 4 : S == 3 ? 2 : S == 2 ? 1 : 0;
    static constexpr int NB = 1 << S;
    static constexpr uint32_t MASK = (1u << NB) - 1u;
    static constexpr uint32_t EMASK = S == 1 ? 0xFFFFu : 0xFFFFFFFFu;
};

// write the cached word back (before anything reads the bit rows of shared memory)
POLAR_DEV void sc_flush_bits(ScCtx& c)
{
    if (c.curW >= 0) {
        SC_ROW(c, c.P.llrRows + c.curW) = c.bw;
        c.curW = -1;
    }
}
// make the word of bit position pos the cached one
POLAR_DEV void sc_enter_word(ScCtx& c, int pos)
{
    const int w = pos >> 5;
    if (w != c.curW) {
        sc_flush_bits(c);
        c.curW = w;
        c.bw = SC_ROW(c, c.P.llrRows + w);
    }
}
template <int NBITS>
POLAR_DEV uint32_t sc_get_field(const ScCtx& c, int pos)
{
    return (c.bw >> (pos & 31)) & (NBITS >= 32 ? 0xFFFFFFFFu : (1u << NBITS) - 1u);
}
template <int NBITS>
POLAR_DEV void sc_put_field(ScCtx& c, int pos, uint32_t v)
{
    if constexpr (NBITS >= 32) {
        c.bw = v;
    } else {
        constexpr uint32_t M = (1u << NBITS) - 1u;
        const int sh = pos & 31;
        c.bw = (c.bw & ~(M << sh)) | ((v & M) << sh);
    }
}

template <int S>
POLAR_DEV void scr_load_parent(const ScCtx& c, uint32_t (&a)[ScReg<S>::W], uint32_t (&b)[ScReg<S>::W])
{
    constexpr int W = ScReg<S>::W;
    bool inRegs = false;
    if constexpr (S < 4) inRegs = S + 1 <= c.P.regStages;
    if (inRegs) {
        if constexpr (S == 1) {
            a[0] = c.r[1];
            b[0] = c.r[1] >> 16;
        } else if constexpr (S < 4) {
            constexpr int PO = ScReg<S + 1>::OFF;
#pragma unroll
            for (int w = 0; w < W; ++w) { a[w] = c.r[PO + w]; b[w] = c.r[PO + W + w]; }
        }
    } else {
        const uint32_t* src = sc_stage(c, S + 1);
        if constexpr (S == 1) {
            a[0] = src[0];
            b[0] = a[0] >> 16;
        } else {
#pragma unroll
            for (int w = 0; w < W; ++w) { a[w] = src[w << 5]; b[w] = src[(w + W) << 5]; }
        }
    }
}

// F / G / G-0R / G-R1 producing (or, for G-R1, consuming) a register block
template <int S, int CODE>
POLAR_DEV void scr_op(ScCtx& c, int pos)
{
    constexpr int W = ScReg<S>::W, OFF = ScReg<S>::OFF, NB = ScReg<S>::NB;
    uint32_t a[W], b[W];
    scr_load_parent<S>(c, a, b);
    if constexpr (CODE == OP_F) {
#pragma unroll
        for (int w = 0; w < W; ++w) c.r[OFF + w] = f4(a[w], b[w]);
    } else if constexpr (CODE == OP_G_0R) {
#pragma unroll
        for (int w = 0; w < W; ++w) c.r[OFF + w] = adds4(a[w], b[w]);
    } else if constexpr (CODE == OP_G) {
        const uint32_t lb = sc_get_field<NB>(c, pos);
#pragma unroll
        for (int w = 0; w < W; ++w) c.r[OFF + w] = g4(a[w], b[w], nibble_to_msb(lb >> (4 * w)));
    } else { // OP_G_R1 (ShortROneNode, fastssc_fip_char.cpp:451-474): right bits = sign(g), left bits ^= right bits
        const uint32_t lb = sc_get_field<NB>(c, pos);
        uint32_t rb = 0;
#pragma unroll
        for (int w = 0; w < W; ++w) rb |= sign_nibble(g4(a[w], b[w], nibble_to_msb(lb >> (4 * w)))) << (4 * w);
        rb &= ScReg<S>::MASK;
        sc_put_field<2 * NB>(c, pos, (lb ^ rb) | (rb << NB));
    }
}

template <int S, bool ZERO_LEFT>
POLAR_DEV void scr_combine(ScCtx& c, int pos) // fip_char.h:164-220 on sign bits
{
    constexpr int NB = ScReg<S>::NB;
    const uint32_t l = sc_get_field<NB>(c, pos), r = sc_get_field<NB>(c, pos + NB);
    sc_put_field<NB>(c, pos, ZERO_LEFT ? r : (l ^ r));
}

// sign bits of a register block
template <int S>
POLAR_DEV uint32_t scr_signs(const uint32_t* x)
{
    uint32_t v = 0;
#pragma unroll
    for (int w = 0; w < ScReg<S>::W; ++w) v |= sign_nibble(x[w]) << (4 * w);
    return v & ScReg<S>::MASK;
}

// short-form SPC decision (fastssc_fip_char.cpp:305-319: first minimum of |x| over the 32-byte vector whose pads are +127):
// returns the sign bits with the flip applied
template <int S>
POLAR_DEV uint32_t scr_spc(const uint32_t* x)
{
    constexpr int W = ScReg<S>::W, n = ScReg<S>::NB;
    uint32_t px = 0, best = 0xFFFFFFFFu, bits = 0;
#pragma unroll
    for (int w = 0; w < W; ++w) {
        const uint32_t xv = x[w] & ScReg<S>::EMASK;
        px ^= xv;
        bits |= sign_nibble(xv) << (4 * w);
        const uint32_t m = uabs4(xv); // |x| as unsigned, |-128| = 128
#pragma unroll
        for (int e = 0; e < 4; ++e)
            if (4 * w + e < n) best = min(best, prmt(m, (uint32_t)(4 * w + e), 0x7054u | (e << 8))); // (|x_e| << 16) | index
    }
    bits &= ScReg<S>::MASK;
    const bool odd = (__popc(px & 0x80808080u) & 1) != 0;
    const uint32_t mv = best >> 16, mi = best & 0xFFFFu;
    // a +127 pad wins when every |x| exceeds 127 (only -128 does); the negate is void for 0 (and -128)
    const bool flip = odd && mv != 0u && mv <= 127u;
    return flip ? bits ^ (1u << mi) : bits;
}

template <int S, int CODE>
POLAR_DEV void scr_leaf(ScCtx& c, int pos)
{
    constexpr int W = ScReg<S>::W, OFF = ScReg<S>::OFF, NB = ScReg<S>::NB;
    const uint32_t* x = c.r + OFF;
    if constexpr (CODE == OP_SC_R0) {
        sc_put_field<NB>(c, pos, 0u);
    } else if constexpr (CODE == OP_SC_R1) {
        sc_put_field<NB>(c, pos, scr_signs<S>(x));
    } else if constexpr (CODE == OP_SC_REP) { // ShortRepetition (fastssc_fip_char.cpp:265-272): zero pads, 32-lane reduce tree
        uint32_t a0 = x[0] & ScReg<S>::EMASK, a1 = 0u;
        if constexpr (W >= 2) a1 = x[1];
        if constexpr (W == 4) { a0 = adds4(a0, x[2]); a1 = adds4(a1, x[3]); }
        uint32_t t = W >= 2 ? adds4(a0, a1) : a0;
        t = adds4(t, t >> 16);
        t = adds4(t, t >> 8);
        sc_put_field<NB>(c, pos, (t & 0x80u) ? 0xFFFFFFFFu : 0u);
    } else if constexpr (CODE == OP_SC_SPC) {
        sc_put_field<NB>(c, pos, scr_spc<S>(x));
    } else { // OP_SC_ZERO_SPC (:361-388) / OP_SC_ZERO_ONE (:390-399): y = sat(l + r) over the halves, decided, duplicated
        static_assert(S >= 2, "a zero-left leaf has at least four elements");
        constexpr int H = S - 1, HW = ScReg<H>::W, HB = ScReg<H>::NB;
        uint32_t y[HW];
        if constexpr (H == 1) {
            y[0] = adds4(x[0], x[0] >> 16) & 0xFFFFu;
        } else {
#pragma unroll
            for (int w = 0; w < HW; ++w) y[w] = adds4(x[w], x[w + HW]);
        }
        uint32_t v;
        if constexpr (CODE == OP_SC_ZERO_SPC) v = scr_spc<H>(y);
        else v = scr_signs<H>(y);
        sc_put_field<NB>(c, pos, v | (v << HB));
    }
}

// ---------------------------------------------------------------- inner steps of the stages 5 .. lgN-2, compiled per stage
// (compile-time trip counts and row offsets; the parent block sits in the thread's rows of shared memory or of the L2
// scratch, never in the channel tiles: the top-level steps keep the generic path)
template <int S>
POLAR_DEV uint32_t* sc_stage_t(const ScCtx& c)
{
    return (S >= c.P.smemStages ? c.spillLane : c.smLane) + (llr_woff(S) << 5);
}

template <int S, int CODE>
POLAR_DEV void scb_fg(ScCtx& c, int pos)
{
    static_assert(S >= 5, "whole bit words");
    constexpr int words = 1 << (S - 2), hw = 1 << (S - 5);
    const uint32_t* src = sc_stage_t<S + 1>(c);
    uint32_t* dst = sc_stage_t<S>(c);
    uint32_t* bits = c.smLane + ((c.P.llrRows + (pos >> 5)) << 5); // the block's first bit row
#pragma unroll 2
    for (int w = 0; w < words; w += 8) {
        uint32_t a[8], b[8];
#pragma unroll
        for (int t = 0; t < 8; ++t) { a[t] = src[(w + t) << 5]; b[t] = src[(w + t + words) << 5]; }
        if constexpr (CODE == OP_F) {
#pragma unroll
            for (int t = 0; t < 8; ++t) dst[(w + t) << 5] = f4(a[t], b[t]);
        } else if constexpr (CODE == OP_G_0R) {
#pragma unroll
            for (int t = 0; t < 8; ++t) dst[(w + t) << 5] = adds4(a[t], b[t]);
        } else if constexpr (CODE == OP_G) {
            const uint32_t lw = bits[(w >> 3) << 5];
#pragma unroll
            for (int t = 0; t < 8; ++t) dst[(w + t) << 5] = g4(a[t], b[t], nibble_to_msb(lw >> (4 * t)));
        } else { // OP_G_R1 (ROneNode, fastssc_fip_char.cpp:427-449)
            const uint32_t lw = bits[(w >> 3) << 5];
            uint32_t rw = 0;
#pragma unroll
            for (int t = 0; t < 8; ++t) rw |= sign_nibble(g4(a[t], b[t], nibble_to_msb(lw >> (4 * t)))) << (4 * t);
            bits[((w >> 3) + hw) << 5] = rw;
            bits[(w >> 3) << 5] = lw ^ rw;
        }
    }
}

// ---------------------------------------------------------------- the kernel
// WIDE: blocks of up to 16 warps at 128 registers (the lockstep barrier then spans every resident warp of the
// SM); otherwise blocks of up to 8 warps with the full register budget.
template <bool WIDE>
__global__ void __launch_bounds__(WIDE ? POLAR_SC_WIDE_THREADS : 256, 1) fastssc_i8_kernel(const __grid_constant__ DecodeParams p,
                                                         const __grid_constant__ Schedule sched)
{
    extern __shared__ __align__(16) uint32_t smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, warpsPerBlock = blockDim.x >> 5;
    ScCtx c{ p };
    c.sm = smem + warp * p.warpW;
    c.chan = static_cast<const uint32_t*>(p.llr);
    c.lane = lane;
    c.spill = p.spill ? p.spill + (size_t)(blockIdx.x * warpsPerBlock + warp) * p.spillWarpW : nullptr;
    c.lastFrame = p.nFrames - 1;
#pragma unroll
    for (int i = 0; i < 8; ++i) c.r[i] = 0u;
    c.bw = 0u;
    c.curW = -1;
    c.smLane = c.sm + lane;
    c.spillLane = c.spill ? c.spill - (llr_woff(p.smemStages) << 5) + lane : c.smLane;
    uint32_t* tileA = c.sm + (p.tileRow << 5);
    uint32_t* tileB = tileA + 33 * 32;
    const uint32_t* myA = tileA + lane * 33;
    const uint32_t* myB = tileB + lane * 33;
    const int halfW = p.N >> 3, cw = min(32, halfW); // words per half of the channel row, chunk width
    uint64_t* mbar = reinterpret_cast<uint64_t*>(c.sm + (p.mbarRow << 5));
    unsigned mbarPhase = 0u;
    if (p.useTma) {
        if (lane == 0) mbar_init(mbar, 1);
        fence_proxy_async();
        __syncwarp();
    }

    const long long totalWarps = (long long)gridDim.x * warpsPerBlock;
    const long long nGroups = (p.nFrames + 31) >> 5;
    // block-uniform trip count: with p.lockstep the warps of a block meet at a barrier before every schedule
    // step, so they fetch the same instruction lines together (see scl_i8_kernel)
    for (long long g0 = (long long)blockIdx.x * warpsPerBlock; g0 < nGroups; g0 += totalWarps) {
        const bool live = g0 + warp < nGroups; // a spare warp redoes the last group and drops it
        const long long grp = live ? g0 + warp : nGroups - 1;
        c.frame0 = grp << 5;
        if (p.prefetchNext && g0 + totalWarps + warp < nGroups) { // the next group of this warp: its channel rows into L2
            const long long nf0 = (g0 + totalWarps + warp) << 5;
            warp_prefetch_l2(static_cast<const int8_t*>(p.llr) + nf0 * p.N, min(32LL, p.nFrames - nf0) * p.N, lane);
        }
        if (c.P.chanRow >= 0) { // one-leaf codes: stage the channel rows (transposed) once
            for (int w0 = 0; w0 < (c.P.N >> 2); w0 += 32) {
                const int n = min(32, (c.P.N >> 2) - w0);
                sc_load_tile(c, tileA, w0, n);
                cp_async_wait_all();
                __syncwarp();
                for (int i = 0; i < n; ++i) SC_ROW(c, c.P.chanRow + w0 + i) = myA[i];
                __syncwarp();
            }
        }
        c.curW = -1;
        const uint32_t lockMask = p.lockstep ? (uint32_t)p.lockstep - 1u : 0x7FFFFFFFu;
        for (int pc = 0;; ++pc) {
            // op word of this kernel: the host replaced the op code (bits 5..0) by a dense handler id (sc_handler_id)
            const uint32_t op = sched.ops[pc];
            const uint32_t hid = (op & 63u) | ((op >> 4) & 64u);
            if (hid == SC_H_END) break;
            if ((pc & lockMask) == 0) __syncthreads(); // lockstep = 2^k: a barrier every 2^k steps
            const int s = (int)POLAR_OP_STAGE(op), pos = (int)POLAR_OP_POS(op);
            if (hid < SC_H_GENERIC) { // register stages: handlers compiled per stage
#define POLAR_SCR_CASES(IDX, FIRST, CALL)                                                   \
    case 1 + 4 * IDX + 0: if constexpr (FIRST <= 1) { constexpr int S = 1; CALL; } break;      \
    case 1 + 4 * IDX + 1: if constexpr (FIRST <= 2) { constexpr int S = 2; CALL; } break;      \
    case 1 + 4 * IDX + 2: if constexpr (FIRST <= 3) { constexpr int S = FIRST <= 3 ? 3 : 4; CALL; } break; \
    case 1 + 4 * IDX + 3: { constexpr int S = 4; CALL; } break;
                switch (hid) {
                POLAR_SCR_CASES(0, 1, (scr_op<S, OP_F>(c, pos)))
                POLAR_SCR_CASES(1, 1, sc_enter_word(c, pos); (scr_op<S, OP_G>(c, pos)))
                POLAR_SCR_CASES(2, 1, (scr_op<S, OP_G_0R>(c, pos)))
                POLAR_SCR_CASES(3, 1, sc_enter_word(c, pos); (scr_op<S, OP_G_R1>(c, pos)))
                POLAR_SCR_CASES(4, 1, sc_enter_word(c, pos); (scr_combine<S, false>(c, pos)))
                POLAR_SCR_CASES(5, 1, sc_enter_word(c, pos); (scr_combine<S, true>(c, pos)))
                POLAR_SCR_CASES(6, 1, sc_enter_word(c, pos); (scr_leaf<S, OP_SC_R0>(c, pos)))
                POLAR_SCR_CASES(7, 1, sc_enter_word(c, pos); (scr_leaf<S, OP_SC_R1>(c, pos)))
                POLAR_SCR_CASES(8, 1, sc_enter_word(c, pos); (scr_leaf<S, OP_SC_REP>(c, pos)))
                POLAR_SCR_CASES(9, 2, sc_enter_word(c, pos); (scr_leaf<S, OP_SC_SPC>(c, pos)))
                POLAR_SCR_CASES(10, 3, sc_enter_word(c, pos); (scr_leaf<S, OP_SC_ZERO_SPC>(c, pos)))
                POLAR_SCR_CASES(11, 2, sc_enter_word(c, pos); (scr_leaf<S, OP_SC_ZERO_ONE>(c, pos)))
                default: break;
                }
#undef POLAR_SCR_CASES
                continue;
            }
            // stages in shared memory / the L2 scratch / the channel tiles: the cached bit word goes back first
            sc_flush_bits(c);
            if (hid >= SC_H_BIG) { // F / G / G-0R / G-R1 of the stages 5 .. lgN-2, compiled per stage
#define POLAR_SCB_CASES(S)                                                          \
    case SC_H_BIG + 4 * (S - 5) + 0: scb_fg<S, OP_F>(c, pos); break;                  \
    case SC_H_BIG + 4 * (S - 5) + 1: scb_fg<S, OP_G>(c, pos); break;                  \
    case SC_H_BIG + 4 * (S - 5) + 2: scb_fg<S, OP_G_0R>(c, pos); break;               \
    case SC_H_BIG + 4 * (S - 5) + 3: scb_fg<S, OP_G_R1>(c, pos); break;
                switch (hid) {
                POLAR_SCB_CASES(5) POLAR_SCB_CASES(6) POLAR_SCB_CASES(7) POLAR_SCB_CASES(8) POLAR_SCB_CASES(9) POLAR_SCB_CASES(10)
                default: break;
                }
#undef POLAR_SCB_CASES
                continue;
            }
            const uint32_t code = hid - SC_H_GENERIC + 1u;
            const bool longForm = (POLAR_OP_FLAGS(op) & OPF_LONG) != 0;
            const bool top = (s + 1 == c.P.lgN) && c.P.chanRow < 0; // parent block = channel row in global memory
            switch (code) {
            case OP_F:      // fip_char.h:76-91
            case OP_G:      // fip_char.h:93-110
            case OP_G_0R:   // fip_char.h:112-131
            case OP_G_R1: { // fastssc_fip_char.cpp:436-474
                uint32_t* dst = sc_stage(c, s);
                if (s < 2) { // block of 2: both halves in one word (never the channel: N >= 8)
                    const ScRows src{ sc_stage(c, s + 1), 1 };
                    const uint32_t x = src.a(0);
                    if (code == OP_F) dst[0] = f4(x, x >> 16);
                    else if (code == OP_G_0R) dst[0] = adds4(x, x >> 16);
                    else if (code == OP_G) {
                        const uint32_t nib = (SC_ROW(c, c.P.llrRows + (pos >> 5)) >> (pos & 31)) & 3u;
                        dst[0] = g4(x, x >> 16, nibble_to_msb(nib));
                    } else sc_g_r1_short(c, src, s, pos);
                    break;
                }
                const int words = 1 << (s - 2);
                if (!top) {
                    const ScRows src{ sc_stage(c, s + 1), words };
                    if (code == OP_F) sc_f_range(src, dst, 0, words);
                    else if (code == OP_G_0R) sc_adds_range(src, dst, 0, words);
                    else if (code == OP_G) sc_g_range(c, src, dst, pos, 0, words);
                    else if (s >= 5) sc_g_r1_range(c, src, s, pos, 0, words);
                    else sc_g_r1_short(c, src, s, pos);
                } else if (p.useTma && code != OP_G_R1) { // words == halfW >= 32: TMA-filled tiles, rotated transposing read
                    for (int w0 = 0; w0 < halfW; w0 += 32) {
                        sc_tma_fill(c, tileA, tileB, w0, halfW, mbar);
                        mbar_wait(mbar, mbarPhase);
                        mbarPhase ^= 1u;
                        if (code == OP_F) sc_top_rot<OP_F>(c, tileA, tileB, dst, pos, w0);
                        else if (code == OP_G_0R) sc_top_rot<OP_G_0R>(c, tileA, tileB, dst, pos, w0);
                        else sc_top_rot<OP_G>(c, tileA, tileB, dst, pos, w0);
                        __syncwarp();
                    }
                } else if (p.chanVec && cw == 32 && code != OP_G_R1) { // words == halfW >= 32, 16-byte aligned rows: 16-byte tiles
                    for (int w0 = 0; w0 < halfW; w0 += 32) {
                        sc_load_tile_v(c, tileA, w0);
                        sc_load_tile_v(c, tileB, halfW + w0);
                        cp_async_wait_all();
                        __syncwarp();
                        if (code == OP_F) sc_top_chunk_v<OP_F>(c, tileA, tileB, dst, pos, w0);
                        else if (code == OP_G_0R) sc_top_chunk_v<OP_G_0R>(c, tileA, tileB, dst, pos, w0);
                        else sc_top_chunk_v<OP_G>(c, tileA, tileB, dst, pos, w0);
                        __syncwarp();
                    }
                } else { // words == halfW
                    for (int w0 = 0; w0 < halfW; w0 += cw) {
                        sc_load_tile(c, tileA, w0, cw);
                        sc_load_tile(c, tileB, halfW + w0, cw);
                        cp_async_wait_all();
                        __syncwarp();
                        const ScTiles src{ myA, myB, w0 };
                        if (code == OP_F) sc_f_range(src, dst, w0, w0 + cw);
                        else if (code == OP_G_0R) sc_adds_range(src, dst, w0, w0 + cw);
                        else if (code == OP_G) sc_g_range(c, src, dst, pos, w0, w0 + cw);
                        else if (s >= 5) sc_g_r1_range(c, src, s, pos, w0, w0 + cw);
                        else sc_g_r1_short(c, src, s, pos); // halfW <= 4: a single chunk
                        __syncwarp();
                    }
                }
                break;
            }
            case OP_COMBINE: { // fip_char.h:164-207
                if (s >= 5) {
                    const int hw = 1 << (s - 5), r0 = c.P.llrRows + (pos >> 5);
                    for (int j = 0; j < hw; ++j) SC_ROW(c, r0 + j) ^= SC_ROW(c, r0 + hw + j);
                } else {
                    const int h = 1 << s, sh = pos & 31;
                    uint32_t wd = SC_ROW(c, c.P.llrRows + (pos >> 5));
                    wd ^= ((wd >> (sh + h)) & ((1u << h) - 1u)) << sh;
                    SC_ROW(c, c.P.llrRows + (pos >> 5)) = wd;
                }
                break;
            }
            case OP_COMBINE_0R: { // fip_char.h:209-220
                if (s >= 5) {
                    const int hw = 1 << (s - 5), r0 = c.P.llrRows + (pos >> 5);
                    for (int j = 0; j < hw; ++j) SC_ROW(c, r0 + j) = SC_ROW(c, r0 + hw + j);
                } else {
                    const int h = 1 << s, sh = pos & 31;
                    const uint32_t mask = (1u << h) - 1u;
                    uint32_t wd = SC_ROW(c, c.P.llrRows + (pos >> 5));
                    wd = (wd & ~(mask << sh)) | (((wd >> (sh + h)) & mask) << sh);
                    SC_ROW(c, c.P.llrRows + (pos >> 5)) = wd;
                }
                break;
            }
            case OP_SC_R0: sc_fill_bits(c, s, pos, 0u); break;
            case OP_SC_R1: sc_put_signs(c, sc_stage(c, s), s, pos, false, 0u, false); break;
            case OP_SC_REP: { // fastssc_fip_char.cpp:225-240,265-272 + reduce_adds_epi8
                uint32_t a[8];
                sc_rep_acc(sc_stage(c, s), s, a);
#pragma unroll
                for (int t = 0; t < 4; ++t) a[t] = adds4(a[t], a[t + 4]);
                a[0] = adds4(a[0], a[2]);
                a[1] = adds4(a[1], a[3]);
                uint32_t t = adds4(a[0], a[1]);
                t = adds4(t, t >> 16);
                t = adds4(t, t >> 8);
                sc_fill_bits(c, s, pos, (t & 0x80u) ? 0xFFFFFFFFu : 0u);
                break;
            }
            case OP_SC_REP2: { // fastssc_fip_char.cpp:249-263 + half_reduce_adds_epi8 (n >= 32)
                uint32_t a[8];
                sc_rep_acc(sc_stage(c, s), s, a);
#pragma unroll
                for (int t = 0; t < 4; ++t) a[t] = adds4(a[t], a[t + 4]);
                a[0] = adds4(a[0], a[2]);
                a[1] = adds4(a[1], a[3]);
                uint32_t t = adds4(a[0], a[1]);
                t = adds4(t, __byte_perm(t, 0u, 0x1032u));
                const uint32_t pat = ((t & 0x80u) ? 0x55555555u : 0u) | ((t & 0x8000u) ? 0xAAAAAAAAu : 0u);
                sc_fill_bits(c, s, pos, pat);
                break;
            }
            case OP_SC_SPC: { // fastssc_fip_char.cpp:274-319
                const uint32_t* src = sc_stage(c, s);
                uint32_t fi;
                const bool flip = sc_spc_decide(src, s, longForm, fi);
                sc_put_signs(c, src, s, pos, false, fi, flip);
                break;
            }
            case OP_SC_ZERO_SPC:   // fastssc_fip_char.cpp:321-388
            case OP_SC_ZERO_ONE: { // fastssc_fip_char.cpp:390-399
                // y = sat(l + r) goes to the (free) child stage, or to row 0 when that stage is spilled
                const int hs = s - 1;
                uint32_t* y = hs < c.P.smemStages ? sc_stage(c, hs) : c.sm + c.lane;
                const uint32_t* src = sc_stage(c, s);
                if (hs >= 2) {
                    const int words = 1 << (hs - 2);
                    sc_adds_range(ScRows{ src, words }, y, 0, words);
                } else {
                    const uint32_t x = src[0];
                    y[0] = adds4(x, x >> 16) & 0xFFFFu;
                }
                uint32_t fi = 0;
                bool flip = false;
                if (code == OP_SC_ZERO_SPC) flip = sc_spc_decide(y, hs, longForm, fi);
                sc_put_signs(c, y, hs, pos, true, fi, flip);
                break;
            }
            default: break;
            }
        }
        sc_flush_bits(c);
        if (live) thread_frame_epilogue(p, c.sm, lane, c.P.llrRows, c.frame0);
    }
}

} // namespace polar
