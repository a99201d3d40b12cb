/* TEST INFRASTRUCTURE ONLY -- scalar restatement of the reference's fp32 decoders:
 *   FastSscAvxFloat  src/polarcode/decoding/fastssc_avx_float.cpp, include/polarcode/decoding/avx_float.h
 *   SclAvxFloat      src/polarcode/decoding/scl_avx_float.cpp
 * "Bits" are floats whose SIGN BIT is the bit (the reference XORs / ANDs float bit patterns); every sum is
 * written in the reference's evaluation order (8 AVX lanes first, then the order of the reduction used),
 * so IEEE-754 single precision gives the same bits here, there and in the CUDA kernels.
 * Compile without -ffast-math and with -ffp-contract=off.
 */
#include "oracle.h"
#include "oracle_common.h"
#include <algorithm>
#include <cmath>
#include <limits>

namespace oracle {

static inline uint32_t fbits(float x) { uint32_t u; memcpy(&u, &x, 4); return u; }
static inline float bitsf(uint32_t u) { float x; memcpy(&x, &u, 4); return x; }
static inline bool sgn(float x) { return (fbits(x) >> 31) != 0; }
static inline float fabs_bits(float x) { return bitsf(fbits(x) & 0x7FFFFFFFu); }
static inline float flip_if(float x, bool f) { return f ? bitsf(fbits(x) ^ 0x80000000u) : x; }

// _mm256_polarf_ps (avx_float.h:55-63): sign = sa ^ sb, magnitude = min(|a|, |b|)  (minps: second operand on NaN)
static inline float f_f32(float a, float b)
{
    float aa = fabs_bits(a), bb = fabs_bits(b);
    float m = aa < bb ? aa : bb;
    return bitsf(((fbits(a) ^ fbits(b)) & 0x80000000u) | fbits(m));
}
// _mm256_polarg_ps (avx_float.h:71-81): (sign(bit) ^ a) + b
static inline float g_f32(float a, float b, float bit) { return flip_if(a, sgn(bit)) + b; }

// 8 AVX lanes accumulate x[8v + j] in order of v, starting from +0.0; blocks shorter than 8 are padded
// with `pad` first (RepetitionPrepare / SpcPrepare, avx_float.h:238-250)
static inline void lane_sums(const float* x, int n, float lanes[8])
{
    for (int j = 0; j < 8; ++j) lanes[j] = 0.0f;
    for (int v = 0; v < (n + 7) / 8; ++v)
        for (int j = 0; j < 8; ++j) {
            int i = 8 * v + j;
            lanes[j] = lanes[j] + (i < n ? x[i] : 0.0f);
        }
}
// reduce_add_ps (avxconvenience.h:256-272): x[0] + x[1] + ... + x[7], left to right
static inline float reduce_add8(const float l[8]) { return ((((((l[0] + l[1]) + l[2]) + l[3]) + l[4]) + l[5]) + l[6]) + l[7]; }

// _mm256_spc_right4_ps (avx_float.h:300-314) on 4 values: every element whose |x| equals the minimum
// flips when the parity is odd (ties flip together).
static inline void spc4_all_ties(const float in[4], float out[4])
{
    float a[4];
    bool par = false;
    for (int i = 0; i < 4; ++i) { a[i] = fabs_bits(in[i]); par ^= sgn(in[i]); }
    float m = std::min(std::min(a[0], a[2]), std::min(a[1], a[3]));
    for (int i = 0; i < 4; ++i) out[i] = flip_if(in[i], par && a[i] == m);
}

struct F32Tree {
    int err = 0;

    // frozen pattern helpers on the node's mask
    static int last_frozen(const uint8_t* fm, int n, int k) // k-th frozen index from the end (k = 1: last)
    {
        for (int i = n - 1; i >= 0; --i)
            if (fm[i] && --k == 0) return i;
        return -1;
    }
    static bool prefix_frozen(const uint8_t* fm, int cnt)
    {
        for (int i = 0; i < cnt; ++i)
            if (!fm[i]) return false;
        return true;
    }

    // SpcDecoder::decode (fastssc_avx_float.cpp:342-375): first global minimum of |x| (per 8-vector first
    // index, strict < across vectors, start +inf), flipped when the sign parity is odd.
    static void spc(const float* in, float* out, int n)
    {
        bool par = false;
        int minIdx = 0;
        float minAbs = INFINITY;
        for (int i = 0; i < n; ++i) {
            out[i] = in[i];
            par ^= sgn(in[i]);
        }
        for (int v = 0; v < (n + 7) / 8; ++v) {
            float best = INFINITY; // lanes beyond a short block are +inf
            int bi = 0;
            for (int j = 0; j < 8 && 8 * v + j < n; ++j) {
                float a = fabs_bits(in[8 * v + j]);
                if (a < best) { best = a; bi = j; }
            }
            // _mm256_minidx_ps: index of the first lane equal to the vector minimum
            if (best == INFINITY) bi = 0;
            if (best < minAbs) { minIdx = 8 * v + bi; minAbs = best; }
        }
        out[minIdx] = flip_if(out[minIdx], par);
    }

    void node(const float* llr, float* bits, const uint8_t* fm, int n)
    {
        const int fc = count_frozen(fm, n);
        if (fc == n) { // RateZeroDecoder :258
            for (int i = 0; i < n; ++i) bits[i] = INFINITY;
            return;
        }
        if (fc == 0) { // RateOneDecoder :268-274
            for (int i = 0; i < n; ++i) bits[i] = llr[i];
            return;
        }
        if (fc == n - 1) { // RepetitionDecoder :284-298
            float l[8];
            lane_sums(llr, n, l);
            float s = reduce_add8(l);
            for (int i = 0; i < n; ++i) bits[i] = s;
            return;
        }
        if (fc == 1) { spc(llr, bits, n); return; }
        if (fc == n - 2) { // DoubleRepetitionDecoder :314-339
            if (!prefix_frozen(fm, n - 2)) { err = -4; return; }
            float l[8], even, odd;
            lane_sums(llr, n, l);
            if (n >= 8) {
                even = (l[0] + l[4]) + (l[2] + l[6]);
                odd = (l[1] + l[5]) + (l[3] + l[7]);
            } else {
                even = ((l[0] + l[2]) + l[4]) + l[6];
                odd = ((l[1] + l[3]) + l[5]) + l[7];
            }
            for (int i = 0; i < n; i += 2) { bits[i] = even; bits[i + 1] = odd; }
            return;
        }
        if (fc == 2 && fm[0] && fm[1]) {
            if (n == 8) { // DoubleSpcDecoderShort8 :466-482: all ties of a parity class flip together
                for (int cls = 0; cls < 2; ++cls) {
                    float m = INFINITY;
                    bool par = false;
                    for (int j = cls; j < 8; j += 2) { m = std::min(m, fabs_bits(llr[j])); par ^= sgn(llr[j]); }
                    for (int j = cls; j < 8; j += 2) bits[j] = flip_if(llr[j], par && fabs_bits(llr[j]) == m);
                }
            } else { // DoubleSpcDecoder :421-461
                float minv[8];
                int mini[8];
                bool par[2] = { false, false };
                for (int j = 0; j < 8; ++j) { minv[j] = std::numeric_limits<float>::max(); mini[j] = 0; }
                for (int i = 0; i < n; ++i) {
                    bits[i] = llr[i];
                    par[i & 1] ^= sgn(llr[i]);
                    float a = fabs_bits(llr[i]);
                    int j = i & 7;
                    if (!(a > minv[j])) { minv[j] = a; mini[j] = i; } // _mm256_argabsmin_ps: the LAST minimum of a lane
                }
                for (int cls = 0; cls < 2; ++cls) {
                    float m = std::min(std::min(minv[cls], minv[cls + 4]), std::min(minv[cls + 2], minv[cls + 6]));
                    int lane = -1;
                    for (int j = cls; j < 8 && lane < 0; j += 2)
                        if (minv[j] == m) lane = j; // first lane of the class holding the class minimum
                    if (lane < 0) { err = -5; return; }
                    bits[mini[lane]] = flip_if(bits[mini[lane]], par[cls]);
                }
            }
            return;
        }
        if (fc == n - 3 && n > 8 && last_frozen(fm, n, 1) == n - 4) { // TripleRepetitionDecoder :566-585
            float l[8], in4[4], out4[4];
            lane_sums(llr, n, l);
            for (int j = 0; j < 4; ++j) in4[j] = l[j + 4] + l[j]; // high lane of input + swaplane
            spc4_all_ties(in4, out4);
            for (int i = 0; i < n; ++i) bits[i] = out4[i & 3];
            return;
        }
        if (fc == n - 4 && last_frozen(fm, n, 1) == n - 4 && last_frozen(fm, n, 2) == n - 6) {
            type_five(llr, bits, n, true); // TypeFiveDecoder :764-792
            return;
        }
        if (n == 8 && fc == 3 && fm[0] && fm[1] && fm[2]) { // RepetitionRateOneDecoderShort8 :720-741
            type_five(llr, bits, n, false);
            return;
        }
        if (n == 8 && fc == 5 && last_frozen(fm, n, 1) == 4 && last_frozen(fm, n, 2) == 3) {
            // ZeroSpcDecoderShort8 :546-556
            float in4[4], out4[4];
            for (int j = 0; j < 4; ++j) in4[j] = llr[j + 4] + llr[j];
            spc4_all_ties(in4, out4);
            for (int i = 0; i < 8; ++i) bits[i] = out4[i & 3];
            return;
        }
        const int h = n / 2;
        const int lf = count_frozen(fm, h), rf = count_frozen(fm + h, h);
        std::vector<float> child(h);
        if (n > 8 && lf == h && rf == 1) { // ZeroSpcDecoder :497-539 (the bits are a copy of the RIGHT llr half)
            bool par = false;
            int minIdx = 0;
            float minAbs = INFINITY;
            for (int i = 0; i < h; ++i) {
                child[i] = llr[i] + llr[i + h];
                bits[i] = llr[i + h];
                bits[i + h] = llr[i + h];
                par ^= sgn(child[i]);
            }
            for (int v = 0; v < h / 8; ++v) {
                float best = INFINITY;
                int bi = 0;
                for (int j = 0; j < 8; ++j) {
                    float a = fabs_bits(child[8 * v + j]);
                    if (a < best) { best = a; bi = j; }
                }
                if (minAbs > 0 && best < minAbs) { minIdx = 8 * v + bi; minAbs = best; }
            }
            bits[minIdx] = flip_if(bits[minIdx], par);
            bits[minIdx + h] = flip_if(bits[minIdx + h], par);
            return;
        }
        if (n > 8 && rf == 0) { // ROneNode :193-219
            for (int i = 0; i < h; ++i) child[i] = f_f32(llr[i], llr[i + h]);
            node(child.data(), bits, fm, h);
            for (int i = 0; i < h; ++i) {
                float o = g_f32(llr[i], llr[i + h], bits[i]);
                bits[i] = bitsf(fbits(bits[i]) ^ fbits(o));
                bits[i + h] = o;
            }
            return;
        }
        if (n > 8 && lf == h) { // ZeroRNode :230-236
            for (int i = 0; i < h; ++i) child[i] = llr[i] + llr[i + h];
            node(child.data(), bits + h, fm + h, h);
            for (int i = 0; i < h; ++i) bits[i] = bits[i + h];
            return;
        }
        // RateRNode :147-154 / ShortRateRNode :178-185
        for (int i = 0; i < h; ++i) child[i] = f_f32(llr[i], llr[i + h]);
        node(child.data(), bits, fm, h);
        for (int i = 0; i < h; ++i) child[i] = g_f32(llr[i], llr[i + h], bits[i]);
        node(child.data(), bits + h, fm + h, h);
        for (int i = 0; i < h; ++i) bits[i] = bitsf(fbits(bits[i]) ^ fbits(bits[i + h]));
    }

    // TypeFiveDecoder (spcRight = true) / RepetitionRateOneDecoderShort8 (false): the block is folded to 8
    // lane sums s; r_j = f(s_j, s_{j+4}); rep = (r0 + r1) + (r2 + r3) (two hadd_ps); right_j = g(s_j, s_{j+4}, rep)
    // decoded as SPC(4) or rate-1; left bits = sign(rep) ^ sign(right_j).
    void type_five(const float* llr, float* bits, int n, bool spcRight)
    {
        float l[8];
        if (spcRight) lane_sums(llr, n, l);
        else for (int j = 0; j < 8; ++j) l[j] = llr[j];
        float r[4], g[4], o[4];
        for (int j = 0; j < 4; ++j) r[j] = f_f32(l[j + 4], l[j]);
        float rep = (r[0] + r[1]) + (r[2] + r[3]);
        for (int j = 0; j < 4; ++j) g[j] = g_f32(l[j], l[j + 4], rep);
        if (spcRight) spc4_all_ties(g, o);
        else for (int j = 0; j < 4; ++j) o[j] = g[j];
        float out8[8];
        for (int j = 0; j < 4; ++j) {
            out8[j] = (sgn(rep) ^ sgn(o[j])) ? -1.0f : 1.0f;
            out8[j + 4] = o[j];
        }
        for (int i = 0; i < n; ++i) bits[i] = out8[i & 7];
    }
};

int fastssc_f32_decode(int N, const uint8_t* fm, int systematic, int crcBits, const float* llr, uint8_t* out)
{
    std::vector<float> bits(N);
    F32Tree t;
    t.node(llr, bits.data(), fm, N);
    if (t.err) return t.err;
    std::vector<uint8_t> hard(N);
    for (int i = 0; i < N; ++i) hard[i] = sgn(bits[i]) ? 1 : 0;
    extract_info(hard.data(), fm, N, systematic, out);
    int K = N - count_frozen(fm, N);
    return crc_check(crcBits, out, (K + 7) / 8); // fastssc_avx_float.cpp:953-955
}

// ---------------------------------------------------------------- SclAvxFloat
struct FPath {
    std::vector<std::vector<float>> llr, bit, leftBit; // indexed by stage
    float metric = 0.0f;
};

// arrayfuncs.h:161-183 on float metrics
static void partial_sort_desc_f(std::vector<unsigned>& idx, std::vector<float>& val, unsigned n, unsigned size)
{
    for (unsigned i = 0; i < size; ++i) idx[i] = i;
    unsigned lim = std::min(size - 1, n);
    for (unsigned i = 0; i < lim; ++i) {
        unsigned index = i;
        for (unsigned j = i + 1; j < size; ++j)
            if (val[j] > val[index]) index = j;
        std::swap(val[i], val[index]);
        std::swap(idx[i], idx[index]);
    }
}
// arrayfuncs.h:209-231 on |llr|
static void find_weak_f(std::vector<unsigned>& idx, std::vector<float>& val, unsigned size, unsigned n)
{
    for (unsigned i = 0; i < size; ++i) idx[i] = i;
    unsigned lim = std::min(size - 1, n);
    for (unsigned i = 0; i < lim; ++i) {
        unsigned index = i;
        for (unsigned j = i + 1; j < size; ++j)
            if (val[j] < val[index]) index = j;
        std::swap(val[i], val[index]);
        std::swap(idx[i], idx[index]);
    }
}

struct SclF {
    int L;
    std::vector<FPath> paths;
    struct Cand { float metric; int src; int nflip; unsigned flip[4]; float repValue; };

    void select(std::vector<Cand>& cands, int stage, int n, bool isRep)
    {
        unsigned size = (unsigned)cands.size(), keep = std::min(size, (unsigned)L);
        std::vector<unsigned> idx(size);
        std::vector<float> val(size);
        for (unsigned i = 0; i < size; ++i) val[i] = cands[i].metric;
        partial_sort_desc_f(idx, val, keep, size);
        std::vector<FPath> next(keep);
        for (unsigned p = 0; p < keep; ++p) {
            const Cand& c = cands[idx[p]];
            next[p] = paths[c.src];
            next[p].metric = val[p];
            std::vector<float>& b = next[p].bit[stage];
            b.assign(n, 0.0f);
            if (isRep) {
                for (int i = 0; i < n; ++i) b[i] = c.repValue;
            } else {
                for (int i = 0; i < n; ++i) b[i] = next[p].llr[stage][i];
                for (int k = 0; k < c.nflip; ++k)
                    if ((int)c.flip[k] < n) b[c.flip[k]] = flip_if(b[c.flip[k]], true);
            }
        }
        paths.swap(next);
    }

    // padded copy of a short block: lanes beyond n read `pad` (the reference writes the pad into the block)
    static void padded(const std::vector<float>& x, int n, float pad, std::vector<float>& out)
    {
        out.assign(std::max(n, 8), pad);
        for (int i = 0; i < n; ++i) out[i] = x[i];
    }

    void decode(const uint8_t* fm, int stage)
    {
        const int n = 1 << stage, fc = count_frozen(fm, n), P = (int)paths.size();
        std::vector<float> x;
        if (fc == 0) { // RateOneDecoder scl_avx_float.cpp:356-419
            std::vector<Cand> cands(P * 4);
            for (int p = 0; p < P; ++p) {
                padded(paths[p].llr[stage], n, INFINITY, x);
                std::vector<float> a(x.size());
                for (size_t i = 0; i < x.size(); ++i) a[i] = fabs_bits(x[i]);
                std::vector<unsigned> idx(std::max(n, 4));
                find_weak_f(idx, a, n, 2);
                float m = paths[p].metric;
                cands[p * 4] = Cand{ m, p, 0, {0, 0, 0, 0}, 0 };
                cands[p * 4 + 1] = Cand{ m - a[0], p, 1, { idx[0], 0, 0, 0 }, 0 };
                cands[p * 4 + 2] = Cand{ m - a[1], p, 1, { idx[1], 0, 0, 0 }, 0 };
                cands[p * 4 + 3] = Cand{ m - a[0] - a[1], p, 2, { idx[0], idx[1], 0, 0 }, 0 };
            }
            select(cands, stage, n, false);
            return;
        }
        if (fc == n) { // RateZeroDecoder :320-340
            for (auto& p : paths) {
                padded(p.llr[stage], n, 0.0f, x);
                float lanes[8];
                for (int j = 0; j < 8; ++j) lanes[j] = 0.0f;
                for (size_t v = 0; v < x.size() / 8; ++v)
                    for (int j = 0; j < 8; ++j) {
                        float t = x[8 * v + j];
                        lanes[j] = lanes[j] + (t < 0.0f ? t : 0.0f); // minps(llr, zero)
                    }
                p.metric += reduce_add8(lanes);
                p.bit[stage].assign(n, INFINITY);
            }
            return;
        }
        if (fc == n - 1 && n < 8) { // RepetitionDecoder :432-481
            std::vector<Cand> cands(P * 2);
            for (int p = 0; p < P; ++p) {
                padded(paths[p].llr[stage], n, 0.0f, x);
                float z[8], o[8], r[8];
                for (int j = 0; j < 8; ++j) {
                    z[j] = 0.0f + (x[j] < 0.0f ? x[j] : 0.0f);
                    o[j] = 0.0f + (x[j] > 0.0f ? x[j] : 0.0f);
                    r[j] = 0.0f + x[j];
                }
                float result = std::fabs(reduce_add8(r)), m = paths[p].metric;
                cands[p * 2] = Cand{ m + reduce_add8(z), p, 0, {0, 0, 0, 0}, result };
                cands[p * 2 + 1] = Cand{ m - reduce_add8(o), p, 0, {0, 0, 0, 0}, -result };
            }
            select(cands, stage, n, true);
            return;
        }
        if (fc == 1) { // SpcDecoder :498-617
            std::vector<Cand> cands(P * 8);
            for (int p = 0; p < P; ++p) {
                padded(paths[p].llr[stage], n, INFINITY, x);
                std::vector<float> a(x.size());
                bool parity = false;
                for (size_t i = 0; i < x.size(); ++i) { a[i] = fabs_bits(x[i]); parity ^= sgn(x[i]); }
                std::vector<unsigned> idx(std::max(n, 4));
                find_weak_f(idx, a, n, 4);
                float m = paths[p].metric, inv;
                unsigned i0 = idx[0], i1 = idx[1], i2 = idx[2], i3 = idx[3];
                Cand c[8];
                for (int k = 0; k < 8; ++k) c[k] = Cand{ 0, p, 0, {0, 0, 0, 0}, 0 };
                auto add = [](Cand& q, unsigned i) { q.flip[q.nflip++] = i; };
                if (parity) {
                    inv = 0.0f;
                    m -= a[0];
                    add(c[0], i0); add(c[4], i0); add(c[5], i0); add(c[6], i0);
                } else {
                    inv = 1.0f;
                    add(c[1], i0); add(c[2], i0); add(c[3], i0); add(c[7], i0);
                }
                c[0].metric = m;
                c[1].metric = m - inv * a[0] - a[1];
                c[2].metric = m - inv * a[0] - a[2];
                c[3].metric = m - inv * a[0] - a[3];
                c[4].metric = m - a[1] - a[2];
                c[5].metric = m - a[1] - a[3];
                c[6].metric = m - a[2] - a[3];
                c[7].metric = m - inv * a[0] - a[1] - a[2] - a[3];
                add(c[1], i1); add(c[2], i2); add(c[3], i3);
                add(c[4], i1); add(c[4], i2);
                add(c[5], i1); add(c[5], i3);
                add(c[6], i2); add(c[6], i3);
                add(c[7], i1); add(c[7], i2); add(c[7], i3);
                for (int k = 0; k < 8; ++k) cands[p * 8 + k] = c[k];
            }
            select(cands, stage, n, false);
            return;
        }
        // RateRNode / ShortRateRNode :229-310
        const int cs = stage - 1, h = n / 2;
        for (auto& p : paths) {
            p.llr[cs].resize(h);
            for (int i = 0; i < h; ++i) p.llr[cs][i] = f_f32(p.llr[stage][i], p.llr[stage][i + h]);
        }
        decode(fm, cs);
        for (auto& p : paths) {
            p.leftBit[cs] = p.bit[cs];
            p.llr[cs].resize(h);
            for (int i = 0; i < h; ++i) p.llr[cs][i] = g_f32(p.llr[stage][i], p.llr[stage][i + h], p.leftBit[cs][i]);
        }
        decode(fm + h, cs);
        for (auto& p : paths) {
            p.bit[stage].resize(n);
            for (int i = 0; i < h; ++i) {
                p.bit[stage][i] = bitsf(fbits(p.leftBit[cs][i]) ^ fbits(p.bit[cs][i]));
                p.bit[stage][i + h] = p.bit[cs][i];
            }
        }
    }
};

int scl_f32_decode(int N, int L, const uint8_t* fm, int systematic, int crcBits, const float* llr, uint8_t* out)
{
    int stages = 0;
    while ((1 << stages) < N) ++stages;
    SclF s;
    s.L = L;
    s.paths.resize(1);
    FPath& p0 = s.paths[0];
    p0.llr.resize(stages + 1); p0.bit.resize(stages + 1); p0.leftBit.resize(stages + 1);
    p0.llr[stages].assign(llr, llr + N);
    s.decode(fm, stages);
    // extractBestPath: scl_avx_float.cpp:711-750
    int K = N - count_frozen(fm, N), kb = (K + 7) / 8;
    std::vector<uint8_t> hard(N);
    for (size_t p = 0; p < s.paths.size(); ++p) {
        for (int i = 0; i < N; ++i) hard[i] = sgn(s.paths[p].bit[stages][i]) ? 1 : 0;
        extract_info(hard.data(), fm, N, systematic, out);
        if (crc_check(crcBits, out, kb)) return 1;
    }
    for (int i = 0; i < N; ++i) hard[i] = sgn(s.paths[0].bit[stages][i]) ? 1 : 0;
    extract_info(hard.data(), fm, N, systematic, out);
    return 0;
}

} // namespace oracle
