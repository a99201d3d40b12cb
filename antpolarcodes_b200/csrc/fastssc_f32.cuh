// Fast-SSC fp32 decoder for sm_100a: one frame per THREAD, 32 frames per warp in lockstep.
//
// Restates PolarCode::Decoding::FastSscAvxFloat (src/polarcode/decoding/fastssc_avx_float.cpp,
// include/polarcode/decoding/avx_float.h) bit-exactly: the reference's "bits" are floats whose sign bit is the
// bit, so only signs travel here (packed, 1 bit per code bit); every sum is evaluated in the reference's
// order -- 8 AVX lanes accumulated vector by vector from +0.0, then the reduction the node uses
// (reduce_add_ps = left to right, hadd_ps = pairwise) -- with IEEE round-to-nearest adds (__fadd_rn, never
// contracted), so the decisions are the reference's, including its quirks: ZeroSpcDecoder takes its bits
// from the RIGHT llr half (:508-517), the 4-element SPC kernels flip every tied minimum (avx_float.h:300-314),
// DoubleSpcDecoder keeps the LAST minimum of a lane (avx_float.h:257-266).
//
// Same B200 mapping as fastssc_i8.cuh: rows of 32 words ([row][lane], conflict free), channel LLRs through a
// padded transpose tile, top stages optionally in the L2-resident scratch, output through a coalesced image.
#pragma once
#include "decode_common.cuh"
#include "decode_params.h"
#include "primitives.cuh"
#include "schedule.h"

namespace polar {

struct ScfCtx {
    uint32_t* sm;
    const uint32_t* chan; // channel rows (global), N words per frame
    long long frame0, lastFrame;
    uint32_t* spill;
    int lane, lgN, nw;    // nw = N
    int bitRow, chanRow, smemStages;
    uint32_t* smLane;     // sm + lane
    uint32_t* spillLane;  // spill + lane, rebased so that row f32_woff(s) is the row of spilled stage s
};

#define SCF_ROW(c, row) (c).sm[((row) << 5) + (c).lane]


POLAR_DEV float* scf_stage(const ScfCtx& c, int s)
{
    if (s == c.lgN) return reinterpret_cast<float*>(c.sm + (c.chanRow << 5) + c.lane);
    uint32_t* base = s >= c.smemStages ? c.spillLane : c.smLane; // both already offset to the lane's column
    return reinterpret_cast<float*>(base + (f32_woff(s) << 5));
}

POLAR_DEV uint32_t fbits(float x) { return __float_as_uint(x); }
POLAR_DEV float bitsf(uint32_t u) { return __uint_as_float(u); }
POLAR_DEV float fabs_bits(float x) { return bitsf(fbits(x) & 0x7FFFFFFFu); }
// _mm256_polarf_ps (avx_float.h:55-63)
POLAR_DEV float f_f32(float a, float b)
{
    const float aa = fabs_bits(a), bb = fabs_bits(b);
    const float m = aa < bb ? aa : bb;
    return bitsf(((fbits(a) ^ fbits(b)) & 0x80000000u) | fbits(m));
}
// _mm256_polarg_ps (avx_float.h:71-81): (sign(bit) ^ a) + b; `neg` = bit in the sign position
POLAR_DEV float g_f32(float a, float b, uint32_t neg) { return __fadd_rn(bitsf(fbits(a) ^ neg), b); }

struct ScfRows {
    const float* p;
    int half; // elements per half
    POLAR_DEV float a(int i) const { return p[i << 5]; }
    POLAR_DEV float b(int i) const { return p[(i + half) << 5]; }
};
struct ScfTiles {
    const float* ta;
    const float* tb;
    int i0;
    POLAR_DEV float a(int i) const { return ta[i - i0]; }
    POLAR_DEV float b(int i) const { return tb[i - i0]; }
};

// tile[r*33 + i] = channel word (w0 + i) of frame (frame0 + r), i < cw <= 32
POLAR_DEV void scf_load_tile(const ScfCtx& c, uint32_t* tile, int w0, int cw)
{
    // asynchronous 4-byte copies straight into the padded tile: all 32 rows of the lane's column in flight at
    // once; the caller waits (cp_async_wait_all) before the __syncwarp that publishes the tile
    if (c.lane < cw) {
        if (c.frame0 + 31 <= c.lastFrame) {
            const uint32_t* src = c.chan + c.frame0 * c.nw + w0 + c.lane;
#pragma unroll 8
            for (int r = 0; r < 32; ++r) cp_async4(tile + r * 33 + c.lane, src + (long long)r * c.nw);
        } else {
            for (int r = 0; r < 32; ++r) {
                const long long f = min(c.frame0 + r, c.lastFrame);
                cp_async4(tile + r * 33 + c.lane, c.chan + f * c.nw + w0 + c.lane);
            }
        }
    }
}

// ---- 16-byte variant of the channel tiles (rows of 32 words, 16-byte aligned channel rows): tile row r = frame frame0 + r, its
// 16-byte chunk q stored at chunk position q ^ (r & 7).  A lane copies one chunk of four rows per instruction (8 instead of 32
// copies per tile) and reads its own row back with 128-bit loads: the eight lanes of a quarter warp hit eight different chunk
// positions, i.e. all 32 banks -- no conflicts without the padding that would break the 16-byte alignment.
POLAR_DEV void scf_load_tile_v(const ScfCtx& c, uint32_t* tile, int w0)
{
    const int rsub = c.lane >> 3, q = c.lane & 7;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int r = rsub + 4 * k;
        const long long f = min(c.frame0 + r, c.lastFrame);
        cp_async16(tile + r * 32 + ((q ^ (r & 7)) << 2), c.chan + f * c.nw + w0 + 4 * q);
    }
}
struct ScfTilesV { // the lane's rows of both tiles
    const uint32_t* ra;
    const uint32_t* rb;
    int sw; // lane & 7
    POLAR_DEV float4 a4(int q) const { return *reinterpret_cast<const float4*>(ra + ((q ^ sw) << 2)); }
    POLAR_DEV float4 b4(int q) const { return *reinterpret_cast<const float4*>(rb + ((q ^ sw) << 2)); }
};
// the 32 elements [i0, i0 + 32) of a top-level step out of the 16-byte tiles; dst = the lane's rows of the child stage
template <int CODE>
POLAR_DEV void scf_top_chunk_v(ScfCtx& c, const ScfTilesV& t, float* dst, int s, int pos, int i0)
{
    uint32_t lw = 0u, rw = 0u;
    if (CODE == OP_G || CODE == OP_G_R1) lw = SCF_ROW(c, c.bitRow + ((pos + i0) >> 5)); // pos and i0 are multiples of 32 here
#pragma unroll 2
    for (int q = 0; q < 8; ++q) {
        const float4 a = t.a4(q), b = t.b4(q);
        const float av[4] = { a.x, a.y, a.z, a.w }, bv[4] = { b.x, b.y, b.z, b.w };
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const int j = 4 * q + e;
            float r;
            if (CODE == OP_F) r = f_f32(av[e], bv[e]);
            else if (CODE == OP_G_0R) r = __fadd_rn(av[e], bv[e]);
            else r = g_f32(av[e], bv[e], (lw >> j) << 31);
            if (CODE == OP_G_R1) rw |= (fbits(r) >> 31) << j;
            else dst[(i0 + j) << 5] = r;
        }
    }
    if (CODE == OP_G_R1) { // fastssc_avx_float.cpp:204-219: right bits = sign(right), left bits ^= right bits
        const int h = 1 << s;
        SCF_ROW(c, c.bitRow + ((pos + h + i0) >> 5)) = rw;
        SCF_ROW(c, c.bitRow + ((pos + i0) >> 5)) = lw ^ rw;
    }
}

// ---- F / G / G-0R steps whose parent (stage S + 1 <= 5) and child blocks both sit in the lane's shared-memory rows, compiled
// per stage: compile-time trip counts and row offsets, one load of the left bits (most steps of a tree are this small)
template <int S, int CODE>
POLAR_DEV void scf_fg_small(ScfCtx& c, int pos)
{
    constexpr int n = 1 << S, CH = n < 8 ? n : 8;
    const float* src = reinterpret_cast<const float*>(c.smLane + (f32_woff(S + 1) << 5));
    float* dst = reinterpret_cast<float*>(c.smLane + (f32_woff(S) << 5));
    uint32_t lw = 0u;
    if (CODE == OP_G) lw = SCF_ROW(c, c.bitRow + (pos >> 5)) >> (pos & 31); // the block's bits sit inside one word (n <= 16)
#pragma unroll
    for (int i0 = 0; i0 < n; i0 += CH) {
        float a[CH], b[CH];
#pragma unroll
        for (int t = 0; t < CH; ++t) { a[t] = src[(i0 + t) << 5]; b[t] = src[(i0 + t + n) << 5]; }
#pragma unroll
        for (int t = 0; t < CH; ++t) {
            float r;
            if (CODE == OP_F) r = f_f32(a[t], b[t]);
            else if (CODE == OP_G_0R) r = __fadd_rn(a[t], b[t]);
            else r = g_f32(a[t], b[t], (lw >> (i0 + t)) << 31);
            dst[(i0 + t) << 5] = r;
        }
    }
}
template <int CODE>
POLAR_DEV void scf_fg_small_s(ScfCtx& c, int s, int pos)
{
    switch (s) {
    case 0: scf_fg_small<0, CODE>(c, pos); break;
    case 1: scf_fg_small<1, CODE>(c, pos); break;
    case 2: scf_fg_small<2, CODE>(c, pos); break;
    case 3: scf_fg_small<3, CODE>(c, pos); break;
    default: scf_fg_small<4, CODE>(c, pos); break;
    }
}

// bit `i` of the block starting at absolute position pos, moved to the sign position
POLAR_DEV uint32_t scf_bit_neg(const ScfCtx& c, int pos, int i)
{
    const int bp = pos + i;
    return (SCF_ROW(c, c.bitRow + (bp >> 5)) >> (bp & 31)) << 31;
}

// ---- per-thread steps over the element range [ib, ie)
template <class Src>
POLAR_DEV void scf_f_range(const Src& src, float* dst, int ib, int ie)
{
    int i = ib;
    for (; i + 7 < ie; i += 8) { // 16 loads in flight: the parent block may sit in the L2 scratch
        float a[8], b[8];
#pragma unroll
        for (int t = 0; t < 8; ++t) { a[t] = src.a(i + t); b[t] = src.b(i + t); }
#pragma unroll
        for (int t = 0; t < 8; ++t) dst[(i + t) << 5] = f_f32(a[t], b[t]);
    }
    for (; i < ie; ++i) dst[i << 5] = f_f32(src.a(i), src.b(i));
}
template <class Src>
POLAR_DEV void scf_add_range(const Src& src, float* dst, int ib, int ie) // G_function_0R (avx_float.h:164-173)
{
    int i = ib;
    for (; i + 7 < ie; i += 8) {
        float a[8], b[8];
#pragma unroll
        for (int t = 0; t < 8; ++t) { a[t] = src.a(i + t); b[t] = src.b(i + t); }
#pragma unroll
        for (int t = 0; t < 8; ++t) dst[(i + t) << 5] = __fadd_rn(a[t], b[t]);
    }
    for (; i < ie; ++i) dst[i << 5] = __fadd_rn(src.a(i), src.b(i));
}
template <class Src>
POLAR_DEV void scf_g_range(ScfCtx& c, const Src& src, float* dst, int pos, int ib, int ie)
{
    int i = ib;
    if (((ie - ib) & 31) == 0 && (pos & 31) == 0) { // whole bit words
        for (; i < ie; i += 32) {
            const uint32_t lw = SCF_ROW(c, c.bitRow + ((pos + i) >> 5));
#pragma unroll 1
            for (int t0 = 0; t0 < 32; t0 += 8) {
                float a[8], b[8];
#pragma unroll
                for (int t = 0; t < 8; ++t) { a[t] = src.a(i + t0 + t); b[t] = src.b(i + t0 + t); }
#pragma unroll
                for (int t = 0; t < 8; ++t) dst[(i + t0 + t) << 5] = g_f32(a[t], b[t], (lw >> (t0 + t)) << 31);
            }
        }
        return;
    }
    for (; i < ie; ++i) dst[i << 5] = g_f32(src.a(i), src.b(i), scf_bit_neg(c, pos, i));
}
// ROneNode::rightDecode (fastssc_avx_float.cpp:204-219): right = g(a, b, left bits); right bits = sign(right);
// left bits ^= right bits.  h = 2^s elements.
template <class Src>
POLAR_DEV void scf_g_r1_range(ScfCtx& c, const Src& src, int s, int pos, int ib, int ie)
{
    const int h = 1 << s;
    if (h >= 32) {
        for (int i = ib; i < ie; i += 32) {
            const int wl = c.bitRow + ((pos + i) >> 5), wr = c.bitRow + ((pos + h + i) >> 5);
            const uint32_t lw = SCF_ROW(c, wl);
            uint32_t rw = 0;
#pragma unroll 1
            for (int t0 = 0; t0 < 32; t0 += 8) {
                float a[8], b[8];
#pragma unroll
                for (int t = 0; t < 8; ++t) { a[t] = src.a(i + t0 + t); b[t] = src.b(i + t0 + t); }
#pragma unroll
                for (int t = 0; t < 8; ++t) rw |= (fbits(g_f32(a[t], b[t], (lw >> (t0 + t)) << 31)) >> 31) << (t0 + t);
            }
            SCF_ROW(c, wr) = rw;
            SCF_ROW(c, wl) = lw ^ rw;
        }
    } else { // everything inside one bit word (n = 2h <= 32, pos % n == 0)
        const int sh = pos & 31;
        const uint32_t mask = (1u << h) - 1u;
        const uint32_t word = SCF_ROW(c, c.bitRow + (pos >> 5));
        const uint32_t lbits = (word >> sh) & mask;
        uint32_t rbits = 0;
        for (int i = 0; i < h; ++i) rbits |= (fbits(g_f32(src.a(i), src.b(i), (lbits >> i) << 31)) >> 31) << i;
        const uint32_t both = (lbits ^ rbits) | (rbits << h);
        const uint32_t clr = (mask | (mask << h)) << sh;
        SCF_ROW(c, c.bitRow + (pos >> 5)) = (word & ~clr) | (both << sh);
    }
}

// ---- bit helpers (block of n = 2^s bits at absolute position pos, pos % n == 0)
POLAR_DEV void scf_put_bits(ScfCtx& c, int s, int pos, int j, uint32_t v)
{
    if (s >= 5) {
        SCF_ROW(c, c.bitRow + (pos >> 5) + j) = v;
    } else {
        const uint32_t mask = (1u << (1 << s)) - 1u;
        const int sh = pos & 31;
        const uint32_t wd = SCF_ROW(c, c.bitRow + (pos >> 5));
        SCF_ROW(c, c.bitRow + (pos >> 5)) = (wd & ~(mask << sh)) | ((v & mask) << sh);
    }
}
POLAR_DEV void scf_fill_bits(ScfCtx& c, int s, int pos, uint32_t pattern)
{
    const int nbw = s >= 5 ? 1 << (s - 5) : 1;
    for (int j = 0; j < nbw; ++j) scf_put_bits(c, s, pos, j, pattern);
}
// bits = sign(src[i]) with up to two flipped positions (flip index >= n: none)
POLAR_DEV void scf_put_signs(ScfCtx& c, const float* src, int s, int pos, uint32_t flipA, uint32_t flipB)
{
    const int n = 1 << s, nbw = s >= 5 ? n >> 5 : 1;
    for (int j = 0; j < nbw; ++j) {
        uint32_t v = 0;
        const int cnt = min(32, n);
#pragma unroll 4
        for (int t = 0; t < cnt; ++t) v |= (fbits(src[(32 * j + t) << 5]) >> 31) << t;
        if ((int)(flipA >> 5) == j && flipA < (uint32_t)n) v ^= 1u << (flipA & 31u);
        if ((int)(flipB >> 5) == j && flipB < (uint32_t)n) v ^= 1u << (flipB & 31u);
        scf_put_bits(c, s, pos, j, v);
    }
}

// ---- sums in the reference's order
// lanes[j] = ((0 + x[j]) + x[8 + j]) + ...; a block shorter than 8 is zero padded (RepetitionPrepare)
POLAR_DEV void scf_lane_sums(const float* src, int s, float (&l)[8])
{
    const int n = 1 << s;
#pragma unroll
    for (int j = 0; j < 8; ++j) l[j] = 0.0f;
    if (n < 8) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (j < n) l[j] = __fadd_rn(0.0f, src[j << 5]);
    } else {
        for (int v = 0; v < n; v += 8) {
#pragma unroll
            for (int j = 0; j < 8; ++j) l[j] = __fadd_rn(l[j], src[(v + j) << 5]);
        }
    }
}
POLAR_DEV float scf_reduce_add8(const float (&l)[8]) // reduce_add_ps: left to right (avxconvenience.h:256-272)
{
    float t = __fadd_rn(l[0], l[1]);
#pragma unroll
    for (int j = 2; j < 8; ++j) t = __fadd_rn(t, l[j]);
    return t;
}
// _mm256_spc_right4_ps (avx_float.h:300-314): 4 values, every tied minimum flips when the parity is odd.
// Returns the 4 result sign bits.
POLAR_DEV uint32_t scf_spc4(const float (&x)[4])
{
    float a[4];
    uint32_t par = 0, sg = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        a[i] = fabs_bits(x[i]);
        par ^= fbits(x[i]) >> 31;
        sg |= (fbits(x[i]) >> 31) << i;
    }
    const float m = fminf(fminf(a[0], a[2]), fminf(a[1], a[3]));
    if (par) {
#pragma unroll
        for (int i = 0; i < 4; ++i)
            if (a[i] == m) sg ^= 1u << i;
    }
    return sg;
}
// TypeFiveDecoder (:764-792) / RepetitionRateOneDecoderShort8 (:720-741) on 8 lane values: returns the 8 bits
POLAR_DEV uint32_t scf_type_five(const float (&l)[8], bool spcRight)
{
    float r[4], g[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) r[j] = f_f32(l[j], l[j + 4]);
    const float rep = __fadd_rn(__fadd_rn(r[0], r[1]), __fadd_rn(r[2], r[3])); // two hadd_ps
    const uint32_t neg = fbits(rep) & 0x80000000u;
#pragma unroll
    for (int j = 0; j < 4; ++j) g[j] = g_f32(l[j], l[j + 4], neg);
    uint32_t right;
    if (spcRight) right = scf_spc4(g);
    else right = (fbits(g[0]) >> 31) | ((fbits(g[1]) >> 31) << 1) | ((fbits(g[2]) >> 31) << 2) | ((fbits(g[3]) >> 31) << 3);
    const uint32_t left = right ^ (neg ? 15u : 0u);
    return left | (right << 4);
}
// first position of the minimum of |x| over n values, strict <, starting from +inf at index 0
// (SpcDecoder: per-vector _mm256_minidx_ps + strict < across vectors, fastssc_avx_float.cpp:342-375); also
// returns the sign parity
POLAR_DEV uint32_t scf_first_min(const float* src, int n, uint32_t& parity)
{
    float best = __int_as_float(0x7F800000);
    uint32_t idx = 0, par = 0;
    for (int i = 0; i < n; ++i) {
        const float x = src[i << 5];
        par ^= fbits(x);
        const float a = fabs_bits(x);
        if (a < best) { best = a; idx = (uint32_t)i; }
    }
    parity = par >> 31;
    return idx;
}

// ---------------------------------------------------------------- the kernel
// WIDE: blocks of up to 16 warps at 128 registers (the lockstep barrier then spans every resident warp of the
// SM); otherwise blocks of up to 8 warps with the full register budget.
template <bool WIDE>
__global__ void __launch_bounds__(WIDE ? POLAR_SC_WIDE_THREADS : 256, 1) fastssc_f32_kernel(const __grid_constant__ DecodeParams p,
                                                            const __grid_constant__ Schedule sched)
{
    extern __shared__ __align__(16) uint32_t smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, warpsPerBlock = blockDim.x >> 5;
    ScfCtx c;
    c.sm = smem + warp * p.warpW;
    c.chan = static_cast<const uint32_t*>(p.llr);
    c.lane = lane;
    c.lgN = p.lgN;
    c.nw = p.N;
    c.bitRow = p.llrRows;
    c.chanRow = p.chanRow;
    c.smemStages = p.smemStages;
    c.spill = p.spill ? p.spill + (size_t)(blockIdx.x * warpsPerBlock + warp) * p.spillWarpW : nullptr;
    c.lastFrame = p.nFrames - 1;
    c.smLane = c.sm + lane;
    c.spillLane = c.spill ? c.spill - (f32_woff(p.smemStages) << 5) + lane : c.smLane;
    uint32_t* tileA = c.sm + (p.tileRow << 5);
    uint32_t* tileB = tileA + 33 * 32;
    const float* myA = reinterpret_cast<const float*>(tileA + lane * 33);
    const float* myB = reinterpret_cast<const float*>(tileB + lane * 33);
    const int halfN = p.N >> 1, cw = min(32, halfN);

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
            warp_prefetch_l2(static_cast<const float*>(p.llr) + nf0 * p.N, min(32LL, p.nFrames - nf0) * p.N * 4, lane);
        }
        if (c.chanRow >= 0) { // one-leaf codes: stage the channel rows (transposed) once
            for (int w0 = 0; w0 < c.nw; w0 += 32) {
                const int n = min(32, c.nw - w0);
                scf_load_tile(c, tileA, w0, n);
                cp_async_wait_all();
                __syncwarp();
                for (int i = 0; i < n; ++i) SCF_ROW(c, c.chanRow + w0 + i) = tileA[lane * 33 + i];
                __syncwarp();
            }
        }
        for (int pc = 0;; ++pc) {
            const uint32_t op = sched.ops[pc];
            const uint32_t code = POLAR_OP_CODE(op);
            if (code == OP_END) break;
            if (p.lockstep && (pc & (p.lockstep - 1)) == 0) __syncthreads(); // lockstep = 2^k: a barrier every 2^k steps
            const int s = (int)POLAR_OP_STAGE(op), pos = (int)POLAR_OP_POS(op), n = 1 << s;
            const bool top = (s + 1 == c.lgN) && c.chanRow < 0; // parent block = channel row in global memory
            switch (code) {
            case OP_F:      // avx_float.h:98-125
            case OP_G:      // avx_float.h:146-162
            case OP_G_0R:   // avx_float.h:164-173
            case OP_G_R1: { // fastssc_avx_float.cpp:204-219
                if (!top && s <= 4 && s + 1 < c.smemStages && code != OP_G_R1) { // small step inside shared memory
                    if (code == OP_F) scf_fg_small_s<OP_F>(c, s, pos);
                    else if (code == OP_G) scf_fg_small_s<OP_G>(c, s, pos);
                    else scf_fg_small_s<OP_G_0R>(c, s, pos);
                    break;
                }
                float* dst = scf_stage(c, s);
                if (!top) {
                    const ScfRows src{ scf_stage(c, s + 1), n };
                    if (code == OP_F) scf_f_range(src, dst, 0, n);
                    else if (code == OP_G_0R) scf_add_range(src, dst, 0, n);
                    else if (code == OP_G) scf_g_range(c, src, dst, pos, 0, n);
                    else scf_g_r1_range(c, src, s, pos, 0, n);
                } else if (p.chanVec && cw == 32) { // n == halfN >= 32, 16-byte aligned channel rows: 16-byte tiles
                    const ScfTilesV tv{ tileA + lane * 32, tileB + lane * 32, lane & 7 };
                    for (int i0 = 0; i0 < halfN; i0 += 32) {
                        scf_load_tile_v(c, tileA, i0);
                        scf_load_tile_v(c, tileB, halfN + i0);
                        cp_async_wait_all();
                        __syncwarp();
                        if (code == OP_F) scf_top_chunk_v<OP_F>(c, tv, dst, s, pos, i0);
                        else if (code == OP_G_0R) scf_top_chunk_v<OP_G_0R>(c, tv, dst, s, pos, i0);
                        else if (code == OP_G) scf_top_chunk_v<OP_G>(c, tv, dst, s, pos, i0);
                        else scf_top_chunk_v<OP_G_R1>(c, tv, dst, s, pos, i0);
                        __syncwarp();
                    }
                } else { // n == halfN
                    for (int i0 = 0; i0 < halfN; i0 += cw) {
                        scf_load_tile(c, tileA, i0, cw);
                        scf_load_tile(c, tileB, halfN + i0, cw);
                        cp_async_wait_all();
                        __syncwarp();
                        const ScfTiles src{ myA, myB, i0 };
                        if (code == OP_F) scf_f_range(src, dst, i0, i0 + cw);
                        else if (code == OP_G_0R) scf_add_range(src, dst, i0, i0 + cw);
                        else if (code == OP_G) scf_g_range(c, src, dst, pos, i0, i0 + cw);
                        else scf_g_r1_range(c, src, s, pos, i0, i0 + cw); // h < 32: a single chunk
                        __syncwarp();
                    }
                }
                break;
            }
            case OP_COMBINE: { // avx_float.h:188-195 / CombineBitsShort
                if (s >= 5) {
                    const int hw = 1 << (s - 5), r0 = c.bitRow + (pos >> 5);
                    for (int j = 0; j < hw; ++j) SCF_ROW(c, r0 + j) ^= SCF_ROW(c, r0 + hw + j);
                } else {
                    const int sh = pos & 31;
                    uint32_t wd = SCF_ROW(c, c.bitRow + (pos >> 5));
                    wd ^= ((wd >> (sh + n)) & ((1u << n) - 1u)) << sh;
                    SCF_ROW(c, c.bitRow + (pos >> 5)) = wd;
                }
                break;
            }
            case OP_COMBINE_0R: { // avx_float.h:197-202
                if (s >= 5) {
                    const int hw = 1 << (s - 5), r0 = c.bitRow + (pos >> 5);
                    for (int j = 0; j < hw; ++j) SCF_ROW(c, r0 + j) = SCF_ROW(c, r0 + hw + j);
                } else {
                    const int sh = pos & 31;
                    const uint32_t mask = (1u << n) - 1u;
                    uint32_t wd = SCF_ROW(c, c.bitRow + (pos >> 5));
                    wd = (wd & ~(mask << sh)) | (((wd >> (sh + n)) & mask) << sh);
                    SCF_ROW(c, c.bitRow + (pos >> 5)) = wd;
                }
                break;
            }
            case OP_SC_R0: scf_fill_bits(c, s, pos, 0u); break;                            // +inf
            case OP_SC_R1: scf_put_signs(c, scf_stage(c, s), s, pos, 0xFFFFFFFFu, 0xFFFFFFFFu); break;
            case OP_SC_REP: { // RepetitionDecoder :284-298
                float l[8];
                scf_lane_sums(scf_stage(c, s), s, l);
                const float t = scf_reduce_add8(l);
                scf_fill_bits(c, s, pos, (fbits(t) >> 31) ? 0xFFFFFFFFu : 0u);
                break;
            }
            case OP_SC_REP2: { // DoubleRepetitionDecoder :314-339
                float l[8], even, odd;
                scf_lane_sums(scf_stage(c, s), s, l);
                if (n >= 8) {
                    even = __fadd_rn(__fadd_rn(l[0], l[4]), __fadd_rn(l[2], l[6]));
                    odd = __fadd_rn(__fadd_rn(l[1], l[5]), __fadd_rn(l[3], l[7]));
                } else {
                    even = __fadd_rn(__fadd_rn(__fadd_rn(l[0], l[2]), l[4]), l[6]);
                    odd = __fadd_rn(__fadd_rn(__fadd_rn(l[1], l[3]), l[5]), l[7]);
                }
                scf_fill_bits(c, s, pos, ((fbits(even) >> 31) ? 0x55555555u : 0u) | ((fbits(odd) >> 31) ? 0xAAAAAAAAu : 0u));
                break;
            }
            case OP_SC_SPC: { // SpcDecoder :342-375
                const float* src = scf_stage(c, s);
                uint32_t par;
                const uint32_t idx = scf_first_min(src, n, par);
                scf_put_signs(c, src, s, pos, par ? idx : 0xFFFFFFFFu, 0xFFFFFFFFu);
                break;
            }
            case OP_SC_ZERO_SPC: { // ZeroSpcDecoder :497-539: bits = sign(right half), flipped at argmin |l + r|
                const float* src = scf_stage(c, s);
                const int h = n >> 1;
                float best = __int_as_float(0x7F800000);
                uint32_t idx = 0, par = 0;
                for (int i = 0; i < h; ++i) {
                    const float x = __fadd_rn(src[i << 5], src[(i + h) << 5]);
                    par ^= fbits(x);
                    const float a = fabs_bits(x);
                    if (a < best) { best = a; idx = (uint32_t)i; }
                }
                const uint32_t f = (par >> 31) ? idx : 0xFFFFFFFFu;
                scf_put_signs(c, src + (h << 5), s - 1, pos, f, 0xFFFFFFFFu);
                scf_put_signs(c, src + (h << 5), s - 1, pos + h, f, 0xFFFFFFFFu);
                break;
            }
            case OP_F32_SPC2: { // DoubleSpcDecoder :421-461 (n >= 16)
                const float* src = scf_stage(c, s);
                float minv[8];
                uint32_t mini[8], parE = 0, parO = 0;
#pragma unroll
                for (int j = 0; j < 8; ++j) { minv[j] = 3.402823466e+38f; mini[j] = 0; }
                for (int v = 0; v < n; v += 8) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const float x = src[(v + j) << 5];
                        if (j & 1) parO ^= fbits(x); else parE ^= fbits(x);
                        const float a = fabs_bits(x);
                        if (!(a > minv[j])) { minv[j] = a; mini[j] = (uint32_t)(v + j); } // the LAST minimum of a lane
                    }
                }
                uint32_t flip[2];
#pragma unroll
                for (int cls = 0; cls < 2; ++cls) {
                    const float m = fminf(fminf(minv[cls], minv[cls + 4]), fminf(minv[cls + 2], minv[cls + 6]));
                    uint32_t f = mini[cls + 6];
                    if (minv[cls + 4] == m) f = mini[cls + 4];
                    if (minv[cls + 2] == m) f = mini[cls + 2];
                    if (minv[cls] == m) f = mini[cls]; // the first lane of the class that holds the minimum
                    flip[cls] = ((cls ? parO : parE) >> 31) ? f : 0xFFFFFFFFu;
                }
                scf_put_signs(c, src, s, pos, flip[0], flip[1]);
                break;
            }
            case OP_F32_SPC2_8: { // DoubleSpcDecoderShort8 :466-482: every tied minimum of a class flips
                const float* src = scf_stage(c, s);
                float x[8];
                uint32_t bitsOut = 0;
#pragma unroll
                for (int j = 0; j < 8; ++j) { x[j] = src[j << 5]; bitsOut |= (fbits(x[j]) >> 31) << j; }
#pragma unroll
                for (int cls = 0; cls < 2; ++cls) {
                    const float m = fminf(fminf(fabs_bits(x[cls]), fabs_bits(x[cls + 4])), fminf(fabs_bits(x[cls + 2]), fabs_bits(x[cls + 6])));
                    const uint32_t par = (fbits(x[cls]) ^ fbits(x[cls + 2]) ^ fbits(x[cls + 4]) ^ fbits(x[cls + 6])) >> 31;
#pragma unroll
                    for (int j = cls; j < 8; j += 2)
                        if (par && fabs_bits(x[j]) == m) bitsOut ^= 1u << j;
                }
                scf_put_bits(c, s, pos, 0, bitsOut);
                break;
            }
            case OP_F32_REP3:     // TripleRepetitionDecoder :566-585
            case OP_F32_ZSPC_8: { // ZeroSpcDecoderShort8 :546-556
                const float* src = scf_stage(c, s);
                float l[8], in4[4];
                if (code == OP_F32_REP3) scf_lane_sums(src, s, l);
                else {
#pragma unroll
                    for (int j = 0; j < 8; ++j) l[j] = src[j << 5];
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) in4[j] = __fadd_rn(l[j + 4], l[j]);
                const uint32_t b4 = scf_spc4(in4);
                scf_fill_bits(c, s, pos, b4 * 0x11111111u);
                break;
            }
            case OP_F32_TYPE5:      // TypeFiveDecoder :764-792
            case OP_F32_REPR1_8: {  // RepetitionRateOneDecoderShort8 :720-741
                const float* src = scf_stage(c, s);
                float l[8];
                if (code == OP_F32_TYPE5) scf_lane_sums(src, s, l);
                else {
#pragma unroll
                    for (int j = 0; j < 8; ++j) l[j] = src[j << 5];
                }
                const uint32_t b8 = scf_type_five(l, code == OP_F32_TYPE5);
                scf_fill_bits(c, s, pos, b8 * 0x01010101u);
                break;
            }
            default: break;
            }
        }
        if (live) thread_frame_epilogue(p, c.sm, lane, c.bitRow, c.frame0);
    }
}

} // namespace polar
