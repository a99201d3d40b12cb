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

// L2 scratch of the spilled stages: [group of 4 elements][32 columns][4] -- column = lane of the slot's owner, so a lane reads
// four consecutive elements of its block with one 128-bit access and the 32 lanes of a warp touch one contiguous 512-byte row.
// Element i of a block sits at p + (i >> 2) * 128 + (i & 3) floats (c.spill already points at the frame's first column).
struct FWords {
    const float* p;
    int stride; // 1: contiguous (shared-memory slot, channel row); 32: spilled block
    bool cg;    // spilled block: read at L2 (the writers use st.cg, L1 is never involved)
    POLAR_DEV float operator()(int i) const { return cg ? __ldcg(p + ((i >> 2) << 7) + (i & 3)) : p[i]; }
};

POLAR_DEV int sclf_spill_woff(const SclCtx& c, int stage) { return ((1 << stage) - (1 << c.P.smemStages)) << 5; }
POLAR_DEV float* sclf_spill_block(const SclCtx& c, int stage, uint32_t slot)
{
    return reinterpret_cast<float*>(c.spill + sclf_spill_woff(c, stage) + 4 * slot);
}
POLAR_DEV FWords sclf_llr_block(const SclCtx& c, int stage, uint32_t slot)
{
    if (stage == c.P.lgN) return FWords{ reinterpret_cast<const float*>(c.chan), 1, false };
    if (stage >= c.P.smemStages) return FWords{ sclf_spill_block(c, stage, slot), 32, true };
    return FWords{ reinterpret_cast<const float*>(c.llrp() + slot * c.P.slotW + f32_woff(stage)), 1, false };
}
POLAR_DEV void sclf_store(const SclCtx& c, int stage, uint32_t slot, int i, float v)
{
    if (stage >= c.P.smemStages) __stcg(sclf_spill_block(c, stage, slot) + ((i >> 2) << 7) + (i & 3), v);
    else reinterpret_cast<float*>(c.llrp() + slot * c.P.slotW + f32_woff(stage))[i] = v;
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
    const FWords src = sclf_llr_block(c, s + 1, scl_slot_of(c, c.llrPtr, s + 1, q, lgG));
    const BitBlk lb = scl_bits_a(c, scl_slot_of(c, c.lbPtr, s, q, lgG), s);
    if (lgG == 0 && s >= 2 && s + 1 >= c.P.smemStages && (s + 1 < c.P.lgN || c.P.chanVec)) {
        // a lane owns the path and the parent block sits in the L2 scratch or in the channel row: 128-bit accesses, four
        // elements per step, eight loads in flight
        const bool chan = (s + 1 == c.P.lgN);
        const int vs = chan ? 1 : 32; // float4 stride between consecutive groups of four elements
        const float4* A = reinterpret_cast<const float4*>(src.p);
        const float4* B = A + (n >> 2) * vs;
        const bool toSpill = s >= c.P.smemStages;
        float4* dv = toSpill ? reinterpret_cast<float4*>(sclf_spill_block(c, s, (uint32_t)q)) : nullptr;
        float* ds = reinterpret_cast<float*>(c.llrp() + q * c.P.slotW + f32_woff(s));
        auto ld4 = [&](const float4* p4) { return chan ? __ldcs(p4) : __ldcg(p4); };
        auto one = [&](const float4& a, const float4& b, int e0) { // elements e0 .. e0+3
            float4 r;
            if (ISG) {
                const uint32_t lw = lb.ld(e0 >> 5) >> (e0 & 31);
                r.x = g_f32(a.x, b.x, lw << 31);
                r.y = g_f32(a.y, b.y, (lw >> 1) << 31);
                r.z = g_f32(a.z, b.z, (lw >> 2) << 31);
                r.w = g_f32(a.w, b.w, (lw >> 3) << 31);
            } else {
                r.x = f_f32(a.x, b.x); r.y = f_f32(a.y, b.y); r.z = f_f32(a.z, b.z); r.w = f_f32(a.w, b.w);
            }
            if (toSpill) {
                __stcg(dv + (e0 >> 2) * 32, r);
            } else {
                ds[e0] = r.x; ds[e0 + 1] = r.y; ds[e0 + 2] = r.z; ds[e0 + 3] = r.w;
            }
        };
        const int nv = n >> 2;
        int v = 0;
        for (; v + 3 < nv; v += 4) {
            float4 a[4], b[4];
#pragma unroll
            for (int t = 0; t < 4; ++t) { a[t] = ld4(A + (v + t) * vs); b[t] = ld4(B + (v + t) * vs); }
#pragma unroll
            for (int t = 0; t < 4; ++t) one(a[t], b[t], 4 * (v + t));
        }
        for (; v < nv; ++v) one(ld4(A + v * vs), ld4(B + v * vs), 4 * v);
        scl_set_own(c.llrPtr, s, c.gl);
        __syncwarp();
        return;
    }
    int i = g;
    if (s + 1 >= c.P.smemStages) { // parent block in the L2 scratch / global memory: 16 loads in flight
        for (; i + 7 * G < n; i += 8 * G) {
            float a[8], b[8], r[8];
#pragma unroll
            for (int t = 0; t < 8; ++t) { a[t] = src(i + t * G); b[t] = src(i + t * G + n); }
#pragma unroll
            for (int t = 0; t < 8; ++t) {
                const int it = i + t * G;
                r[t] = ISG ? g_f32(a[t], b[t], (lb.ld(it >> 5) >> (it & 31)) << 31) : f_f32(a[t], b[t]);
            }
#pragma unroll
            for (int t = 0; t < 8; ++t) sclf_store(c, s, (uint32_t)q, i + t * G, r[t]);
        }
    }
    for (; i + 3 * G < n; i += 4 * G) {
        float a[4], b[4], r[4];
#pragma unroll
        for (int t = 0; t < 4; ++t) { a[t] = src(i + t * G); b[t] = src(i + t * G + n); }
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const int it = i + t * G;
            r[t] = ISG ? g_f32(a[t], b[t], (lb.ld(it >> 5) >> (it & 31)) << 31) : f_f32(a[t], b[t]);
        }
#pragma unroll
        for (int t = 0; t < 4; ++t) sclf_store(c, s, (uint32_t)q, i + t * G, r[t]);
    }
    for (; i < n; i += G) {
        const float a = src(i), b = src(i + n);
        sclf_store(c, s, (uint32_t)q, i, ISG ? g_f32(a, b, (lb.ld(i >> 5) >> (i & 31)) << 31) : f_f32(a, b));
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
POLAR_DEV void sclf_weak_search(const FWords& src, int n, float (&val)[4], int (&idx)[4], uint32_t& signXor)
{
    unsigned long long T[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) T[r] = ~0ull;
    uint32_t px = 0u, thr = 0xFFFFFFFFu; // thr = |x| of the 4th smallest tail key so far: a key that is not below it changes nothing
    auto take = [&](uint32_t u, int i) { // (positions come in ascending order, so on equal |x| the earlier one stays: strict <)
        px ^= u;
        const uint32_t a = u & 0x7FFFFFFFu;
        if (a < thr) {
            unsigned long long key = ((unsigned long long)a << 16) | (unsigned)i;
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const unsigned long long lo = key < T[r] ? key : T[r];
                key = key < T[r] ? T[r] : key;
                T[r] = lo;
            }
            thr = (uint32_t)(T[3] >> 16);
        }
    };
    // the tail never moves: keep its 4 smallest (|x|, position) keys; eight loads in flight (the block may sit in the L2 scratch)
    int i = 4;
    for (; i < n && (i & 7); ++i) take(fbits(src(i)), i);
    for (; i + 7 < n; i += 8) {
        uint32_t u[8];
#pragma unroll
        for (int t = 0; t < 8; ++t) u[t] = fbits(src(i + t));
#pragma unroll
        for (int t = 0; t < 8; ++t) take(u[t], i + t);
    }
    for (; i < n; ++i) take(fbits(src(i)), i);
#pragma unroll
    for (int e = 0; e < 4; ++e)
        if (e < n) px ^= fbits(src(e));
    signXor = px;
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

// simplePartialSortDescending on float metrics (see scl_select in scl_i8.cuh): 64-bit keys
// (ordered metric << 8 | 255 - current position) in the frame's select scratch (low words, high words, swap
// history), every lane carries the maximum of its own F keys; key 0 marks a final / unused position.
template <int L, int FMAX>
POLAR_DEV void sclf_select(const SclCtx& c, const float (&cand)[FMAX], int lgF, int lgP, int nOut, float& outMet, int& outIdx)
{
    const int gl = c.gl, F = 1 << lgF;
    uint32_t* klo = c.selp();
    uint32_t* khi = c.selp() + 8 * L;
    uint32_t* hs = c.selp() + 16 * L;
    const int row = gl << lgF;
    const bool act = gl < (1 << lgP);
    unsigned long long lmax = 0ull;
#pragma unroll
    for (int k = 0; k < FMAX; ++k) {
        if (k < F) {
            const uint32_t o = f32_ordered(__fadd_rn(cand[k], 0.0f)); // -0.0 -> +0.0: equal floats, equal keys
            const unsigned long long key = act ? (((unsigned long long)o << 8) | (unsigned)(255 - (row + k))) : 0ull;
            klo[row + k] = (uint32_t)key;
            khi[row + k] = (uint32_t)(key >> 32);
            lmax = key > lmax ? key : lmax;
        }
    }
    __syncwarp();
    outMet = 0.0f;
    int myW = 0;
#pragma unroll 1
    for (int i = 0; i < nOut; ++i) {
        const unsigned long long best = list_max64<L>(lmax);
        const int pos = 255 - (int)(best & 255u), wl = pos >> lgF, il = i >> lgF;
        if (gl == i) {
            outMet = f32_unordered((uint32_t)(best >> 8));
            myW = pos;
        }
        const uint32_t ilo = klo[i], ihi = khi[i]; // the element now at position i
        __syncwarp();
        if (gl == wl) { // ... moves to the winner's place
            klo[pos] = (ilo & ~255u) | (uint32_t)(255 - pos);
            khi[pos] = ihi;
        }
        if (gl == il) { // position i is final
            klo[i] = 0u;
            khi[i] = 0u;
        }
        if (gl == 0) hs[i] = (uint32_t)pos;
        __syncwarp();
        if (gl == wl || gl == il) {
            lmax = 0ull;
#pragma unroll
            for (int k = 0; k < FMAX; ++k) {
                if (k < F) {
                    const unsigned long long key = ((unsigned long long)khi[row + k] << 32) | klo[row + k];
                    lmax = key > lmax ? key : lmax;
                }
            }
        }
    }
    int p = myW; // follow the swaps backwards to the original candidate (see scl_select)
    for (int t = nOut - 2; t >= 0; --t) {
        const int wt = (int)hs[t];
        if (t < gl && wt == p) p = t;
    }
    outIdx = p;
    __syncwarp();
}

// Lists of at most 8 paths (M = P * F <= 64 candidates), as scl_select_k32 of the int8 kernel: the 64-bit key is
// (ordered metric << 32) | (63 - CURRENT position) << 6 | ORIGINAL candidate index, so the winner of a step names the candidate
// it came from: no swap history, no walk back.  Compiled once per (L, fan-out), steps unrolled, rows moved with vector accesses.
// Returns the key of list entry gl: metric = f32_unordered(key >> 32), original candidate = key & 63.
template <int L, int LGF>
__device__ __noinline__ unsigned long long sclf_select_k64(uint32_t* sel, int gl, int lgP, int nOut, float c0, float c1, float c2,
                                                           float c3, float c4, float c5, float c6, float c7)
{
    static_assert(L <= 8, "6 + 6 position bits");
    constexpr int F = 1 << LGF;
    constexpr int VW = (L < F ? L : F) < 4 ? (L < F ? L : F) : 4; // the frame's scratch is aligned to L words, a lane's row to min(L, F)
    const float cand[8] = { c0, c1, c2, c3, c4, c5, c6, c7 };
    uint32_t* klo = sel + (gl << LGF);
    uint32_t* khi = klo + 8 * L;
    const bool act = gl < (1 << lgP);
    uint32_t lo[F], hi[F];
    unsigned long long lmax = 0ull;
#pragma unroll
    for (int k = 0; k < F; ++k) {
        const uint32_t pos = (uint32_t)(gl << LGF) + k;
        hi[k] = act ? f32_ordered(__fadd_rn(cand[k], 0.0f)) : 0u; // -0.0 -> +0.0: equal floats, equal keys
        lo[k] = act ? (((63u - pos) << 6) | pos) : 0u;
        const unsigned long long key = ((unsigned long long)hi[k] << 32) | lo[k];
        lmax = key > lmax ? key : lmax;
    }
    auto put = [&](uint32_t* dst, const uint32_t (&v)[F]) {
        if constexpr (VW == 4) {
#pragma unroll
            for (int k = 0; k < F; k += 4) *reinterpret_cast<uint4*>(dst + k) = make_uint4(v[k], v[k + 1], v[k + 2], v[k + 3]);
        } else if constexpr (VW == 2) {
#pragma unroll
            for (int k = 0; k < F; k += 2) *reinterpret_cast<uint2*>(dst + k) = make_uint2(v[k], v[k + 1]);
        } else {
#pragma unroll
            for (int k = 0; k < F; ++k) dst[k] = v[k];
        }
    };
    auto get = [&](const uint32_t* src, uint32_t (&v)[F]) {
        if constexpr (VW == 4) {
#pragma unroll
            for (int k = 0; k < F; k += 4) {
                const uint4 q = *reinterpret_cast<const uint4*>(src + k);
                v[k] = q.x; v[k + 1] = q.y; v[k + 2] = q.z; v[k + 3] = q.w;
            }
        } else if constexpr (VW == 2) {
#pragma unroll
            for (int k = 0; k < F; k += 2) {
                const uint2 q = *reinterpret_cast<const uint2*>(src + k);
                v[k] = q.x; v[k + 1] = q.y;
            }
        } else {
#pragma unroll
            for (int k = 0; k < F; ++k) v[k] = src[k];
        }
    };
    put(klo, lo);
    put(khi, hi);
    __syncwarp();
    unsigned long long outKey = 0ull;
    uint32_t* k0lo = sel;          // position-indexed views
    uint32_t* k0hi = sel + 8 * L;
#pragma unroll
    for (int i = 0; i < L; ++i) {
        if (i >= nOut) break;
        const unsigned long long best = list_max64<L>(lmax);
        if (gl == i) outKey = best;
        const uint32_t pf = (uint32_t)best & 0xFC0u; // (63 - position of the winner) << 6
        const int pos = 63 - (int)(pf >> 6), wl = pos >> LGF;
        const int il = i >> LGF;
        const uint32_t ilo = k0lo[i], ihi = k0hi[i]; // the element now at position i ...
        __syncwarp();
        if (gl == wl) { // ... moves to the winner's place
            k0lo[pos] = (ilo & ~0xFC0u) | pf;
            k0hi[pos] = ihi;
        }
        if (gl == il) { // position i is final
            k0lo[i] = 0u;
            k0hi[i] = 0u;
        }
        __syncwarp();
        if (gl == wl || gl == il) {
            get(klo, lo);
            get(khi, hi);
            lmax = 0ull;
#pragma unroll
            for (int k = 0; k < F; ++k) {
                const unsigned long long key = ((unsigned long long)hi[k] << 32) | lo[k];
                lmax = key > lmax ? key : lmax;
            }
        }
    }
    __syncwarp();
    return outKey;
}

// Candidate pruning for any list size: keys with the origin inside for L <= 8, the swap-history version above else.
template <int L, int FMAX>
POLAR_DEV void sclf_prune(const SclCtx& c, const float (&cand)[FMAX], int lgF, int lgP, int nOut, float& outMet, int& outIdx)
{
    if constexpr (L <= 8) {
        unsigned long long k;
        if constexpr (FMAX == 2) {
            k = sclf_select_k64<L, 1>(c.selp(), c.gl, lgP, nOut, cand[0], cand[1], 0.f, 0.f, 0.f, 0.f, 0.f, 0.f);
        } else {
            if (lgF == 3) k = sclf_select_k64<L, 3>(c.selp(), c.gl, lgP, nOut, cand[0], cand[1], cand[2], cand[3], cand[4], cand[5], cand[6], cand[7]);
            else k = sclf_select_k64<L, 2>(c.selp(), c.gl, lgP, nOut, cand[0], cand[1], cand[2], cand[3], 0.f, 0.f, 0.f, 0.f);
        }
        outMet = f32_unordered((uint32_t)(k >> 32));
        outIdx = (int)(k & 63ull);
    } else {
        sclf_select<L, FMAX>(c, cand, lgF, lgP, nOut, outMet, outIdx);
    }
}

POLAR_DEV float sclf_metric(const SclCtx& c) { return __int_as_float(c.metric); }

// sign bits of 32 consecutive LLRs (or the whole block when shorter), LSB first
POLAR_DEV uint32_t sclf_sign_word(const FWords& src, int n, int bw)
{
    uint32_t w = 0;
    const int cnt = min(32, n);
    for (int t = 0; t < cnt; ++t) w |= (fbits(src(32 * bw + t)) >> 31) << t;
    return w;
}

// ---------------------------------------------------------------- leaves
// RateZero: scl_avx_float.cpp:320-340
template <int L>
POLAR_DEV void sclf_leaf_r0(SclCtx& c, int s, int lgP, bool isRight)
{
    const int lgG = Log2<L>::v - lgP, G = 1 << lgG, q = c.gl >> lgG, g = c.gl & (G - 1), n = 1 << s;
    const FWords src = sclf_llr_block(c, s, scl_slot_of(c, c.llrPtr, s, q, lgG));
    float pun = 0.0f;
    if (g == 0) pun = sclf_lane_reduce(src, n, [](float x) { return x < 0.0f ? x : 0.0f; });
    pun = __shfl_sync(FULL_MASK, pun, c.fb + ((c.gl << lgG) & (L - 1)));
    if (c.gl < (1 << lgP)) c.metric = __float_as_int(__fadd_rn(sclf_metric(c), pun));
    const BitBlk out = scl_bits_out(c, q, s, isRight);
    const int nbw = s >= 5 ? 1 << (s - 5) : 1;
    for (int j = g; j < nbw; j += G) out.st(j, 0u);
    scl_finish_leaf(c, s, isRight);
}

// Repetition (blocks of 2 and 4 only): scl_avx_float.cpp:432-481
template <int L>
POLAR_DEV void sclf_leaf_rep(SclCtx& c, int s, int lgP, bool isRight)
{
    constexpr int lgL = Log2<L>::v;
    const int lgG = lgL - lgP, G = 1 << lgG, q = c.gl >> lgG, g = c.gl & (G - 1), n = 1 << s;
    const FWords src = sclf_llr_block(c, s, scl_slot_of(c, c.llrPtr, s, q, lgG));
    float z = 0.0f, o = 0.0f;
    if (g == 0) {
        z = sclf_lane_reduce(src, n, [](float x) { return x < 0.0f ? x : 0.0f; });
        o = sclf_lane_reduce(src, n, [](float x) { return x > 0.0f ? x : 0.0f; });
    }
    const int srcLane = c.fb + ((c.gl << lgG) & (L - 1));
    z = __shfl_sync(FULL_MASK, z, srcLane);
    o = __shfl_sync(FULL_MASK, o, srcLane);
    const float m = sclf_metric(c);
    const float cand[2] = { __fadd_rn(m, z), __fadd_rn(m, -o) };
    const int lgPo = min(lgL, lgP + 1);
    float outMet;
    int outIdx;
    sclf_prune<L, 2>(c, cand, 1, lgP, 1 << lgPo, outMet, outIdx);
    scl_adopt(c, __float_as_int(outMet), (outIdx >> 1) & (L - 1));
    const int lgGo = lgL - lgPo, Go = 1 << lgGo, qo = c.gl >> lgGo, go = c.gl & (Go - 1);
    const int kq = __shfl_sync(FULL_MASK, outIdx & 1, c.fb + qo);
    const BitBlk out = scl_bits_out(c, qo, s, isRight);
    if (go == 0) out.st(0, kq ? (1u << n) - 1u : 0u);
    scl_finish_leaf(c, s, isRight);
}

// RateOne (scl_avx_float.cpp:356-419) and SPC (:498-617)
template <int L>
POLAR_DEV void sclf_leaf_flip(SclCtx& c, int s, int lgP, bool isRight, bool isSpc)
{
    constexpr int lgL = Log2<L>::v;
    const int lgG = lgL - lgP, G = 1 << lgG, q = c.gl >> lgG, g = c.gl & (G - 1), n = 1 << s;
    const FWords src = sclf_llr_block(c, s, scl_slot_of(c, c.llrPtr, s, q, lgG));
    float a[4] = { 0.0f, 0.0f, 0.0f, 0.0f };
    int idx[4] = { 0, 0, 0, 0 };
    int par = 0;
    if (g == 0) {
        uint32_t px = 0;
        sclf_weak_search(src, n, a, idx, px); // one pass: the weak positions and the XOR of the signs
        if (isSpc) par = (int)(px >> 31);
    }
    const int srcLane = c.fb + ((c.gl << lgG) & (L - 1));
#pragma unroll
    for (int t = 0; t < 4; ++t) a[t] = __shfl_sync(FULL_MASK, a[t], srcLane);
    const uint32_t pk01 = __shfl_sync(FULL_MASK, (uint32_t)idx[0] | ((uint32_t)idx[1] << 16), srcLane);
    const uint32_t pk23 = __shfl_sync(FULL_MASK, (uint32_t)idx[2] | ((uint32_t)idx[3] << 16), srcLane);
    par = __shfl_sync(FULL_MASK, par, srcLane);
    float m = sclf_metric(c);
    float cand[8];
    auto sub = [](float x, float y) { return __fadd_rn(x, -y); };
    if (isSpc) {
        const float inv = par ? 0.0f : 1.0f;
        if (par) m = sub(m, a[0]);
        const float mw = sub(m, __fmul_rn(inv, a[0]));
        cand[0] = m;
        cand[1] = sub(mw, a[1]);
        cand[2] = sub(mw, a[2]);
        cand[3] = sub(mw, a[3]);
        cand[4] = sub(sub(m, a[1]), a[2]);
        cand[5] = sub(sub(m, a[1]), a[3]);
        cand[6] = sub(sub(m, a[2]), a[3]);
        cand[7] = sub(sub(sub(mw, a[1]), a[2]), a[3]);
    } else {
        cand[0] = m;
        cand[1] = sub(m, a[0]);
        cand[2] = sub(m, a[1]);
        cand[3] = sub(sub(m, a[0]), a[1]);
        cand[4] = cand[5] = cand[6] = cand[7] = 0.0f;
    }
    const int lgF = isSpc ? 3 : 2;
    const int lgPo = min(lgL, lgP + lgF);
    float outMet;
    int outIdx;
    sclf_prune<L, 8>(c, cand, lgF, lgP, 1 << lgPo, outMet, outIdx);
    const int srcPath = (outIdx >> lgF) & (L - 1), kind = outIdx & ((1 << lgF) - 1);
    const uint32_t s01 = __shfl_sync(FULL_MASK, pk01, c.fb + srcPath), s23 = __shfl_sync(FULL_MASK, pk23, c.fb + srcPath);
    const int sp = __shfl_sync(FULL_MASK, par, c.fb + srcPath);
    const uint32_t table = isSpc ? (sp ? 0xEDB78421u : 0xFCA69530u) : 0x3210u;
    const uint32_t fmask = (table >> (4 * kind)) & 15u;
    scl_adopt(c, __float_as_int(outMet), srcPath);
    // the new path's bits: sign(llr) with the chosen flips (an index past a short block flips nothing)
    const int lgGo = lgL - lgPo, Go = 1 << lgGo, qo = c.gl >> lgGo, go = c.gl & (Go - 1);
    const uint32_t fm = __shfl_sync(FULL_MASK, fmask, c.fb + qo);
    const uint32_t p01 = __shfl_sync(FULL_MASK, s01, c.fb + qo), p23 = __shfl_sync(FULL_MASK, s23, c.fb + qo);
    const FWords srcn = sclf_llr_block(c, s, scl_slot_of(c, c.llrPtr, s, qo, lgGo));
    const BitBlk out = scl_bits_out(c, qo, s, isRight);
    const int nbw = s >= 5 ? 1 << (s - 5) : 1;
    const uint32_t f[4] = { p01 & 0xFFFFu, p01 >> 16, p23 & 0xFFFFu, p23 >> 16 };
    for (int j = go; j < nbw; j += Go) {
        uint32_t w = sclf_sign_word(srcn, n, j);
#pragma unroll
        for (int t = 0; t < 4; ++t)
            if (((fm >> t) & 1u) && (int)f[t] < n && (int)(f[t] >> 5) == j) w ^= 1u << (f[t] & 31u);
        out.st(j, w);
    }
    scl_finish_leaf(c, s, isRight);
}

// ---------------------------------------------------------------- register stages (the fp32 version of scl_i8.cuh's)
// Most steps of the tree work on blocks of at most 16 elements, always with lane = path.  A stage-s block (s <= 4) is
// 2^s floats of the owning lane's registers (block of stage s at x[2^s - 1 ..]); its partial sums are the bit fields
// ba / bb of the int8 kernel (rs_put_bits / rs_op_combine are shared).  Handlers are compiled per stage; a fork moves
// the live register blocks with the lazy-copy vectors.  Float orders as in the generic handlers above.
struct SclfRegs {
    float x[31];
};

template <int S>
POLAR_DEV void rsf_load_parent(const SclCtx& c, const SclfRegs& R, float (&a)[1 << S], float (&b)[1 << S])
{
    constexpr int n = 1 << S;
    bool inRegs = false;
    if constexpr (S < 4) inRegs = S + 1 <= c.P.regStages;
    if (inRegs) {
        if constexpr (S < 4) {
            constexpr int PO = (2 << S) - 1;
#pragma unroll
            for (int i = 0; i < n; ++i) { a[i] = R.x[PO + i]; b[i] = R.x[PO + n + i]; }
        }
    } else {
        const FWords src = sclf_llr_block(c, S + 1, (uint32_t)(c.llrPtr >> (5 * (S + 1))) & 31u);
#pragma unroll
        for (int i = 0; i < n; ++i) { a[i] = src(i); b[i] = src(i + n); }
    }
}

template <int S, bool ISG>
POLAR_DEV void rsf_op_fg(SclCtx& c, SclfRegs& R)
{
    constexpr int n = 1 << S, OFF = n - 1;
    float a[n], b[n];
    rsf_load_parent<S>(c, R, a, b);
    if constexpr (ISG) {
        const uint32_t lb = c.ba >> n; // left-child bits of stage S
#pragma unroll
        for (int i = 0; i < n; ++i) R.x[OFF + i] = g_f32(a[i], b[i], ((lb >> i) & 1u) << 31);
    } else {
#pragma unroll
        for (int i = 0; i < n; ++i) R.x[OFF + i] = f_f32(a[i], b[i]);
    }
}

template <int S>
POLAR_DEV void rsf_adopt(SclCtx& c, SclfRegs& R, int outMet, int srcPath)
{
    const int src = c.fb + srcPath;
    c.metric = outMet;
    c.llrPtr = shfl64(c.llrPtr, src);
    c.lbPtr = shfl64(c.lbPtr, src);
    c.ba = __shfl_sync(FULL_MASK, c.ba, src);
#pragma unroll
    for (int i = (2 << S) - 1; i < 31; ++i) R.x[i] = __shfl_sync(FULL_MASK, R.x[i], src); // the blocks of the stages above S
}

// sclf_lane_reduce on a register block
template <int S, class T>
POLAR_DEV float rsf_lane_reduce(const float* x, T t)
{
    constexpr int n = 1 << S;
    float l[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) l[j] = 0.0f;
    if constexpr (n < 8) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (j < n) l[j] = __fadd_rn(0.0f, t(x[j]));
    } else {
#pragma unroll
        for (int v = 0; v < n; v += 8) {
#pragma unroll
            for (int j = 0; j < 8; ++j) l[j] = __fadd_rn(l[j], t(x[v + j]));
        }
    }
    float r = __fadd_rn(l[0], l[1]);
#pragma unroll
    for (int j = 2; j < 8; ++j) r = __fadd_rn(r, l[j]);
    return r;
}

template <int L, int S>
POLAR_DEV void rsf_leaf_r0(SclCtx& c, SclfRegs& R, bool isRight) // scl_avx_float.cpp:320-340
{
    const float pun = rsf_lane_reduce<S>(&R.x[(1 << S) - 1], [](float v) { return v < 0.0f ? v : 0.0f; });
    c.metric = __float_as_int(__fadd_rn(sclf_metric(c), pun));
    rs_put_bits<S>(c, 0u, isRight);
}

template <int L, int S>
POLAR_DEV void rsf_leaf_rep(SclCtx& c, SclfRegs& R, int lgP, bool isRight) // scl_avx_float.cpp:432-481
{
    constexpr int lgL = Log2<L>::v;
    const float* x = &R.x[(1 << S) - 1];
    const float z = rsf_lane_reduce<S>(x, [](float v) { return v < 0.0f ? v : 0.0f; });
    const float o = rsf_lane_reduce<S>(x, [](float v) { return v > 0.0f ? v : 0.0f; });
    const float m = sclf_metric(c);
    const float cand[2] = { __fadd_rn(m, z), __fadd_rn(m, -o) };
    const int lgPo = min(lgL, lgP + 1);
    float outMet;
    int outIdx;
    sclf_prune<L, 2>(c, cand, 1, lgP, 1 << lgPo, outMet, outIdx);
    rsf_adopt<S>(c, R, __float_as_int(outMet), (outIdx >> 1) & (L - 1));
    rs_put_bits<S>(c, (outIdx & 1) ? 0xFFFFFFFFu : 0u, isRight);
}

// sclf_weak_search on a register block: the K smallest |x| with their original positions, in the reference's order
template <int S, int K>
POLAR_DEV void rsf_weak_search(const float* x, float (&val)[K], int (&idx)[K])
{
    constexpr int n = 1 << S;
    unsigned long long T[K];
#pragma unroll
    for (int r = 0; r < K; ++r) T[r] = ~0ull;
#pragma unroll
    for (int i = 4; i < n; ++i) { // the tail never moves: keep its K smallest (|x|, position) keys
        unsigned long long key = ((unsigned long long)(fbits(x[i]) & 0x7FFFFFFFu) << 16) | (unsigned)i;
#pragma unroll
        for (int r = 0; r < K; ++r) {
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
        ha[e] = e < n ? (fbits(x[e < n ? e : 0]) & 0x7FFFFFFFu) : 0u;
        cur[e] = e;
        alive[e] = e < n;
    }
    int t = 0;
#pragma unroll
    for (int i = 0; i < K; ++i) {
        if (i < n) {
            unsigned long long best = ~0ull;
#pragma unroll
            for (int r = 0; r < K; ++r)
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
            val[i] = __int_as_float(0x7F800000);
        }
    }
}

// RateOne (scl_avx_float.cpp:356-419) and SPC (:498-617) on a register block
template <int L, int S, bool SPC>
POLAR_DEV void rsf_leaf_flip(SclCtx& c, SclfRegs& R, int lgP, bool isRight)
{
    constexpr int lgL = Log2<L>::v, n = 1 << S, lgF = SPC ? 3 : 2;
    const float* x = &R.x[n - 1];
    uint32_t sgn = 0;
#pragma unroll
    for (int i = 0; i < n; ++i) sgn |= (fbits(x[i]) >> 31) << i;
    const int par = SPC ? (__popc(sgn) & 1) : 0;
    float a[4] = { 0.0f, 0.0f, 0.0f, 0.0f };
    int idx[4] = { 0, 0, 0, 0 };
    if constexpr (SPC) {
        rsf_weak_search<S, 4>(x, a, idx);
    } else {
        float v2[2];
        int i2[2];
        rsf_weak_search<S, 2>(x, v2, i2);
        a[0] = v2[0]; a[1] = v2[1];
        idx[0] = i2[0]; idx[1] = i2[1];
    }
    float m = sclf_metric(c);
    float cand[8];
    auto sub = [](float p, float q) { return __fadd_rn(p, -q); };
    if constexpr (SPC) {
        const float inv = par ? 0.0f : 1.0f;
        if (par) m = sub(m, a[0]);
        const float mw = sub(m, __fmul_rn(inv, a[0]));
        cand[0] = m;
        cand[1] = sub(mw, a[1]);
        cand[2] = sub(mw, a[2]);
        cand[3] = sub(mw, a[3]);
        cand[4] = sub(sub(m, a[1]), a[2]);
        cand[5] = sub(sub(m, a[1]), a[3]);
        cand[6] = sub(sub(m, a[2]), a[3]);
        cand[7] = sub(sub(sub(mw, a[1]), a[2]), a[3]);
    } else {
        cand[0] = m;
        cand[1] = sub(m, a[0]);
        cand[2] = sub(m, a[1]);
        cand[3] = sub(sub(m, a[0]), a[1]);
        cand[4] = cand[5] = cand[6] = cand[7] = 0.0f;
    }
    const int lgPo = min(lgL, lgP + lgF);
    float outMet;
    int outIdx;
    sclf_prune<L, 8>(c, cand, lgF, lgP, 1 << lgPo, outMet, outIdx);
    const int srcPath = (outIdx >> lgF) & (L - 1), kind = outIdx & ((1 << lgF) - 1);
    // what the new path needs from its parent: the four weak positions (4 bits each), the parity and the sign bits
    const uint32_t mine = (uint32_t)idx[0] | ((uint32_t)idx[1] << 4) | ((uint32_t)idx[2] << 8) | ((uint32_t)idx[3] << 12) |
                          ((uint32_t)par << 16) | (sgn << 17);
    const uint32_t got = __shfl_sync(FULL_MASK, mine, c.fb + srcPath);
    uint32_t v = got >> 17;
    if constexpr (S == 4) v |= (__shfl_sync(FULL_MASK, sgn >> 15, c.fb + srcPath) & 1u) << 15;
    const int sp = (got >> 16) & 1;
    const uint32_t table = SPC ? (sp ? 0xEDB78421u : 0xFCA69530u) : 0x3210u;
    const uint32_t fm = (table >> (4 * kind)) & 15u;
#pragma unroll
    for (int t = 0; t < 4; ++t)
        if ((fm >> t) & 1u) v ^= 1u << ((got >> (4 * t)) & 15u);
    rsf_adopt<S>(c, R, __float_as_int(outMet), srcPath);
    rs_put_bits<S>(c, v, isRight);
}

// ---------------------------------------------------------------- the kernel
#ifndef POLAR_SCLF_MAXT
#define POLAR_SCLF_MAXT 512 // launch bound = register budget (512 threads: 128 registers); A/B builds override it
#endif
template <int L>
__global__ void __launch_bounds__(POLAR_SCLF_MAXT, 1) scl_f32_kernel(const __grid_constant__ DecodeParams p,
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
    c.spill = p.spill ? p.spill + (size_t)(blockIdx.x * warpsPerBlock + warp) * p.spillWarpW + 4 * c.fb : nullptr; // column = lane
    SclfRegs R;
#pragma unroll
    for (int i = 0; i < 31; ++i) R.x[i] = 0.0f;

    const long long nQuads = (p.nFrames + FPW - 1) / FPW;
    for (long long qb = blockIdx.x; qb * warpsPerBlock < nQuads; qb += gridDim.x) {
        const bool live = qb * warpsPerBlock + warp < nQuads;
        const long long quad = live ? qb * warpsPerBlock + warp : nQuads - 1;
        const long long myFrame = min(quad * FPW + fi, p.nFrames - 1);
        c.chan = reinterpret_cast<const uint32_t*>(static_cast<const float*>(p.llr) + myFrame * p.N);
        c.metric = __float_as_int(0.0f); // every frame starts a fresh list (see the header)
        c.llrPtr = 0ull;
        c.lbPtr = 0ull;
        int lgPfinal = 0;
        c.ba = 0u;
        c.bb[0] = c.bb[1] = 0u;
        for (int pc = 0;; ++pc) {
            // op word: bits 31..16 hold the dense handler id written by the host (scl_handler_id), as in scl_i8_kernel
            const uint32_t op = sched.ops[pc];
            const uint32_t hid = op >> 16;
            if (hid == SCL_H_END) break;
            const int s = (int)POLAR_OP_STAGE(op), lgP = (int)POLAR_OP_LGP(op);
            const bool isRight = (POLAR_OP_FLAGS(op) & OPF_RIGHT) != 0;
            if (p.lockstep && (pc & (p.lockstep - 1)) == 0) __syncthreads(); // lockstep = 2^k: a barrier every 2^k steps
#define POLAR_RSF_CASES(BASE, CALL)                \
    case BASE + 0: { constexpr int S = 1; CALL; } break; \
    case BASE + 1: { constexpr int S = 2; CALL; } break; \
    case BASE + 2: { constexpr int S = 3; CALL; } break; \
    case BASE + 3: { constexpr int S = 4; CALL; } break;
            switch (hid) {
            POLAR_RSF_CASES(SCL_H_RS_F, (rsf_op_fg<S, false>(c, R)))
            POLAR_RSF_CASES(SCL_H_RS_G, (rsf_op_fg<S, true>(c, R)))
            POLAR_RSF_CASES(SCL_H_RS_C, (rs_op_combine<S>(c, isRight)))
            POLAR_RSF_CASES(SCL_H_RS_R0, (rsf_leaf_r0<L, S>(c, R, isRight)); lgPfinal = lgP)
            case SCL_H_RS_REP + 0: rsf_leaf_rep<L, 1>(c, R, lgP, isRight); lgPfinal = min(lgL, lgP + 1); break; // blocks of 2 and 4 only
            case SCL_H_RS_REP + 1: rsf_leaf_rep<L, 2>(c, R, lgP, isRight); lgPfinal = min(lgL, lgP + 1); break;
            POLAR_RSF_CASES(SCL_H_RS_R1, (rsf_leaf_flip<L, S, false>(c, R, lgP, isRight)); lgPfinal = min(lgL, lgP + 2))
            case SCL_H_RS_SPC + 1: rsf_leaf_flip<L, 2, true>(c, R, lgP, isRight); lgPfinal = min(lgL, lgP + 3); break; // an SPC leaf has >= 4 elements
            case SCL_H_RS_SPC + 2: rsf_leaf_flip<L, 3, true>(c, R, lgP, isRight); lgPfinal = min(lgL, lgP + 3); break;
            case SCL_H_RS_SPC + 3: rsf_leaf_flip<L, 4, true>(c, R, lgP, isRight); lgPfinal = min(lgL, lgP + 3); break;
            case SCL_H_F: sclf_op_fg<L, false>(c, s, lgP); break;
            case SCL_H_G: sclf_op_fg<L, true>(c, s, lgP); break;
            case SCL_H_C: scl_op_combine<L>(c, s, lgP, isRight); break;
            case SCL_H_R0: sclf_leaf_r0<L>(c, s, lgP, isRight); lgPfinal = lgP; break;
            case SCL_H_REP: sclf_leaf_rep<L>(c, s, lgP, isRight); lgPfinal = min(lgL, lgP + 1); break;
            case SCL_H_R1: sclf_leaf_flip<L>(c, s, lgP, isRight, false); lgPfinal = min(lgL, lgP + 2); break;
            case SCL_H_SPC: sclf_leaf_flip<L>(c, s, lgP, isRight, true); lgPfinal = min(lgL, lgP + 3); break;
            default: break;
            }
#undef POLAR_RSF_CASES
        }
        scl_extract_best_path<L>(p, c, base, scr, lane, live, quad, lgPfinal);
    }
}

} // namespace polar
