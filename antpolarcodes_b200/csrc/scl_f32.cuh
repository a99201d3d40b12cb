// SCL fp32 list decoder for sm_100a: the fp32 sibling of scl_i8.cuh (same mapping: one list per group of L
// lanes, 32/L frames per warp, lazy-copy slot vectors, lockstep warps, top stages in the L2 scratch).
//
// Restates PolarCode::Decoding::SclAvxFloat (src/polarcode/decoding/scl_avx_float.cpp):
//   * F / G per live path on floats (avx_float.h:55-81), bits = packed signs
//   * leaves RateZero (:320-340), RateOne (:356-419), Repetition (:432-481, only below 8 elements: :637-639)
//     and SPC (:498-617) with float path metrics.  Every float sum follows the reference's order: 8 AVX lanes
//     accumulated vector by vector from +0.0 (short blocks zero padded), reduce_add_ps left to right, the
//     candidate metrics as written (metric - inv*t0 - t1 ...), all with __fadd_rn / __fmul_rn.
//   * candidate pruning = simplePartialSortDescending (arrayfuncs.h:161-183) on 64-bit keys
//     (ordered float metric, 255 - position), so exact float ties resolve like the CPU code.
// A frame always starts from metric 0: the reference object forgets to reset path 0's metric between frames
// (PathList::clear / setFirstPath, scl_avx_float.cpp:46-56,102-108), which makes ITS roundings depend on the
// frames decoded before; parity is therefore defined against a fresh reference object per frame.
// Leaves are evaluated by the first lane of each path's lane group (sequential float orders); F / G / combine
// use all lanes.
#pragma once
#include "fastssc_f32.cuh" // f_f32 / g_f32 / fbits / fabs_bits
#include "scl_i8.cuh"      // SclCtx, slot vectors, bit regions, group helpers, scl_adopt

namespace polar {

struct FWords {
    const float* p;
    int stride;
    bool cg;
    POLAR_DEV float operator()(int i) const { return cg ? __ldcg(p + i * stride) : p[i * stride]; }
};

POLAR_DEV int sclf_spill_woff(const SclCtx& c, int stage) { return ((1 << stage) - (1 << c.smemStages)) << 5; }
POLAR_DEV FWords sclf_llr_block(const SclCtx& c, int stage, uint32_t slot)
{
    if (stage == c.lgN) return FWords{ reinterpret_cast<const float*>(c.chan), 1, false };
    if (stage >= c.smemStages) return FWords{ reinterpret_cast<const float*>(c.spill + sclf_spill_woff(c, stage) + slot), 32, true };
    return FWords{ reinterpret_cast<const float*>(c.llr + slot * c.slotW + f32_woff(stage)), 1, false };
}
POLAR_DEV void sclf_store(const SclCtx& c, int stage, uint32_t slot, int i, float v)
{
    if (stage >= c.smemStages) __stcg(reinterpret_cast<float*>(c.spill + sclf_spill_woff(c, stage) + slot) + i * 32, v);
    else reinterpret_cast<float*>(c.llr + slot * c.slotW + f32_woff(stage))[i] = v;
}

// monotone map float -> uint32 (no NaNs; -0.0 is canonicalised by the caller)
POLAR_DEV uint32_t f32_ordered(float x)
{
    const uint32_t u = fbits(x);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
POLAR_DEV float f32_unordered(uint32_t o) { return bitsf((o & 0x80000000u) ? (o & 0x7FFFFFFFu) : ~o); }

template <int L>
POLAR_DEV unsigned long long list_max64(unsigned long long v)
{
#pragma unroll
    for (int d = 1; d < L; d <<= 1) {
        const uint32_t lo = __shfl_xor_sync(FULL_MASK, (uint32_t)v, d), hi = __shfl_xor_sync(FULL_MASK, (uint32_t)(v >> 32), d);
        const unsigned long long o = ((unsigned long long)hi << 32) | lo;
        v = o > v ? o : v;
    }
    return v;
}

// ---------------------------------------------------------------- inner-node steps (all lanes)
template <int L, bool ISG>
POLAR_DEV void sclf_op_fg(SclCtx& c, int s, int lgP)
{
    const int lgG = Log2<L>::v - lgP, G = 1 << lgG, q = c.gl >> lgG, g = c.gl & (G - 1), n = 1 << s;
    const FWords src = sclf_llr_block(c, s + 1, scl_slot_of(c, c.llrPtr, s + 1, q));
    const uint32_t* lb = scl_bits_a(c, scl_slot_of(c, c.lbPtr, s, q), s);
    int i = g;
    for (; i + 3 * G < n; i += 4 * G) {
        float a[4], b[4], r[4];
#pragma unroll
        for (int t = 0; t < 4; ++t) { a[t] = src(i + t * G); b[t] = src(i + t * G + n); }
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const int it = i + t * G;
            r[t] = ISG ? g_f32(a[t], b[t], (lb[it >> 5] >> (it & 31)) << 31) : f_f32(a[t], b[t]);
        }
#pragma unroll
        for (int t = 0; t < 4; ++t) sclf_store(c, s, (uint32_t)q, i + t * G, r[t]);
    }
    for (; i < n; i += G) {
        const float a = src(i), b = src(i + n);
        sclf_store(c, s, (uint32_t)q, i, ISG ? g_f32(a, b, (lb[i >> 5] >> (i & 31)) << 31) : f_f32(a, b));
    }
    scl_set_own(c.llrPtr, s, c.gl);
    __syncwarp();
}

// ---------------------------------------------------------------- leaf helpers (one lane per path)
// reduce_add_ps of the 8 AVX lane sums of t(x) over the block; short blocks are zero padded first
template <class T>
POLAR_DEV float sclf_lane_reduce(const FWords& src, int n, T t)
{
    float l[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) l[j] = 0.0f;
    if (n < 8) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (j < n) l[j] = __fadd_rn(0.0f, t(src(j)));
    } else {
        for (int v = 0; v < n; v += 8) {
#pragma unroll
            for (int j = 0; j < 8; ++j) l[j] = __fadd_rn(l[j], t(src(v + j)));
        }
    }
    float r = __fadd_rn(l[0], l[1]);
#pragma unroll
    for (int j = 2; j < 8; ++j) r = __fadd_rn(r, l[j]);
    return r;
}

// findWeakLlrs(.., n, 4) on |llr| (arrayfuncs.h:209-231): values and original indices of the 4 smallest,
// with the reference's swap displacement of elements 0..3 (see scl_weak_search in scl_i8.cuh)
POLAR_DEV void sclf_weak_search(const FWords& src, int n, float (&val)[4], int (&idx)[4])
{
    unsigned long long T[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) T[r] = ~0ull;
    for (int i = 4; i < n; ++i) { // the tail never moves: keep its 4 smallest (|x|, position) keys
        unsigned long long key = ((unsigned long long)(fbits(src(i)) & 0x7FFFFFFFu) << 16) | (unsigned)i;
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const unsigned long long lo = key < T[r] ? key : T[r];
            key = key < T[r] ? T[r] : key;
            T[r] = lo;
        }
    }
    uint32_t ha[4];
    int cur[4];
    bool alive[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        ha[e] = e < n ? (fbits(src(e)) & 0x7FFFFFFFu) : 0u;
        cur[e] = e;
        alive[e] = e < n;
    }
    int t = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        if (i < n) {
            unsigned long long best = ~0ull;
#pragma unroll
            for (int r = 0; r < 4; ++r)
                if (r == t) best = T[r];
            int bw = -1;
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const unsigned long long key = ((unsigned long long)ha[e] << 16) | (unsigned)cur[e];
                if (alive[e] && key < best) { best = key; bw = e; }
            }
            const int bpos = (int)(best & 0xFFFFull);
#pragma unroll
            for (int e = 0; e < 4; ++e)
                if (alive[e] && cur[e] == i) cur[e] = bpos;
            if (bw >= 0) {
#pragma unroll
                for (int e = 0; e < 4; ++e)
                    if (e == bw) alive[e] = false;
                idx[i] = bw;
            } else {
                idx[i] = bpos;
                ++t;
            }
            val[i] = bitsf((uint32_t)(best >> 16));
        } else {
            idx[i] = 0;
            val[i] = __int_as_float(0x7F800000); // the +inf pad of a short block (scl_avx_float.cpp:370-373,527-530)
        }
    }
}
