/* TEST INFRASTRUCTURE ONLY -- scalar restatement of SclFipChar
 * (reference: src/polarcode/decoding/scl_fip_char.cpp, include/polarcode/arrayfuncs.h:161-231).
 * Paths are deep-copied on every fork (the reference's lazy copy / DataPool is bookkeeping only).
 */
#include "oracle.h"
#include "oracle_common.h"
#include <algorithm>

namespace oracle {

struct Stats { long* counts; long* sizes; };
static inline void tally(Stats* s, int id, long size)
{
    if (s && s->counts) { s->counts[id]++; s->sizes[id] += size; }
}

struct Path {
    std::vector<std::vector<int8_t>> llr, bit, leftBit; // indexed by stage (block length 1<<stage)
    int64_t metric = 0;
};

// arrayfuncs.h:161-183 : selection sort WITH swaps, strict '>' => first position among equals in
// the current (already swapped) arrangement.
static void partial_sort_desc(std::vector<unsigned>& idx, std::vector<int64_t>& val, unsigned n, unsigned size)
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

// arrayfuncs.h:209-231 : same with '<' over the |llr| scratch block
static void find_weak(std::vector<unsigned>& idx, std::vector<int>& val, unsigned size, unsigned n)
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

struct Scl {
    int L;
    std::vector<Path> paths;
    Stats* st = nullptr;

    struct Cand { int64_t metric; int src; int nflip; unsigned flip[4]; int8_t repValue; };

    // keep the best min(L, cands) candidates in the reference's order and rebuild the list
    // (scl_fip_char.cpp:475-504, 560-580, 696-725)
    void select(std::vector<Cand>& cands, int stage, int n, bool isRep)
    {
        unsigned size = (unsigned)cands.size();
        unsigned keep = std::min(size, (unsigned)L);
        std::vector<unsigned> idx(size);
        std::vector<int64_t> val(size);
        for (unsigned i = 0; i < size; ++i) val[i] = cands[i].metric;
        partial_sort_desc(idx, val, keep, size);
        std::vector<Path> next(keep);
        for (unsigned p = 0; p < keep; ++p) {
            const Cand& c = cands[idx[p]];
            next[p] = paths[c.src];
            next[p].metric = val[p];
            std::vector<int8_t>& b = next[p].bit[stage];
            b.assign(n, 0);
            if (isRep) {
                for (int i = 0; i < n; ++i) b[i] = c.repValue;
            } else {
                for (int i = 0; i < n; ++i) b[i] = next[p].llr[stage][i];
                for (int k = 0; k < c.nflip; ++k) b[c.flip[k]] = (int8_t)~b[c.flip[k]];
            }
        }
        paths.swap(next);
    }

    void decode(const uint8_t* fm, int stage)
    {
        int n = 1 << stage;
        int fc = count_frozen(fm, n);
        int P = (int)paths.size();
        // createDecoder: scl_fip_char.cpp:729-755
        if (fc == n) { // RateZero :387-421
            tally(st, ORACLE_NODE_R0, (long)n);
            for (auto& p : paths) {
                int64_t pun = 0;
                for (int i = 0; i < n; ++i) pun += std::min<int>(p.llr[stage][i], 0);
                p.metric += pun;
                p.bit[stage].assign(n, 127);
            }
            return;
        }
        if (fc == 0) { // RateOne :423-505
            tally(st, ORACLE_NODE_R1, (long)n);
            std::vector<Cand> cands(P * 4);
            for (int p = 0; p < P; ++p) {
                std::vector<int> a(n);
                std::vector<unsigned> idx(std::max(n, 2));
                for (int i = 0; i < n; ++i) a[i] = abs127(paths[p].llr[stage][i]);
                find_weak(idx, a, n, 2);
                int64_t m = paths[p].metric;
                int64_t a0 = a[0], a1 = n > 1 ? a[1] : 0;
                Cand c0{ m, p, 0, {0, 0, 0, 0}, 0 };
                Cand c1{ m - a0, p, 1, { idx[0], 0, 0, 0 }, 0 };
                Cand c2{ m - a1, p, 1, { idx[1], 0, 0, 0 }, 0 };
                Cand c3{ m - a0 - a1, p, 2, { idx[0], idx[1], 0, 0 }, 0 };
                if (n == 1) { c2.metric = -0x100000000000LL; c3.metric = -0x100000000000LL; }
                cands[p * 4] = c0; cands[p * 4 + 1] = c1; cands[p * 4 + 2] = c2; cands[p * 4 + 3] = c3;
            }
            select(cands, stage, n, false);
            return;
        }
        if (fc == n - 1) { // Repetition :508-581
            tally(st, ORACLE_NODE_REP, (long)n);
            std::vector<Cand> cands(P * 2);
            for (int p = 0; p < P; ++p) {
                const std::vector<int8_t>& l = paths[p].llr[stage];
                int64_t neg = 0, pos = 0;
                for (int i = 0; i < n; ++i) { neg += std::min<int>(l[i], 0); pos += std::max<int>(l[i], 0); }
                int8_t r = rep_sum_i8(l.data(), n);
                if (r < -127) r = -127;
                int8_t r0 = r < 0 ? (int8_t)~r : r, r1 = r < 0 ? r : (int8_t)~r;
                int64_t m = paths[p].metric;
                cands[p * 2] = Cand{ m + neg, p, 0, {0, 0, 0, 0}, r0 };
                cands[p * 2 + 1] = Cand{ m - pos, p, 0, {0, 0, 0, 0}, r1 };
            }
            select(cands, stage, n, true);
            return;
        }
        if (fc == 1) { // SPC :583-726
            tally(st, ORACLE_NODE_SPC, (long)n);
            std::vector<Cand> cands(P * 8);
            for (int p = 0; p < P; ++p) {
                const std::vector<int8_t>& l = paths[p].llr[stage];
                std::vector<int> a(n);
                std::vector<unsigned> idx(std::max(n, 4));
                int parity = 0;
                for (int i = 0; i < n; ++i) { a[i] = abs127(l[i]); parity ^= (l[i] & 0x80); }
                find_weak(idx, a, n, 4);
                int64_t m = paths[p].metric, w = 0;
                int64_t a0 = a[0], a1 = a[1], a2 = a[2], a3 = a[3];
                unsigned i0 = idx[0], i1 = idx[1], i2 = idx[2], i3 = idx[3];
                Cand c[8];
                for (int k = 0; k < 8; ++k) c[k] = Cand{ 0, p, 0, {0, 0, 0, 0}, 0 };
                auto add = [](Cand& x, unsigned i) { x.flip[x.nflip++] = i; };
                if (parity) {
                    m -= a0;
                    add(c[0], i0); add(c[4], i0); add(c[5], i0); add(c[6], i0);
                } else {
                    add(c[1], i0); add(c[2], i0); add(c[3], i0); add(c[7], i0);
                    w = a0;
                }
                c[0].metric = m;
                c[1].metric = m - w - a1;
                c[2].metric = m - w - a2;
                c[3].metric = m - w - a3;
                c[4].metric = m - a1 - a2;
                c[5].metric = m - a1 - a3;
                c[6].metric = m - a2 - a3;
                c[7].metric = m - w - a1 - a2 - a3;
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
        // RateRNode / ShortRateRNode :315-385
        int cs = stage - 1, h = n / 2;
        tally(st, ORACLE_NODE_F, (long)h);
        for (auto& p : paths) {
            p.llr[cs].resize(h);
            for (int i = 0; i < h; ++i) p.llr[cs][i] = f_i8(p.llr[stage][i], p.llr[stage][i + h]);
        }
        decode(fm, cs);
        tally(st, ORACLE_NODE_G, (long)h);
        for (auto& p : paths) {
            p.leftBit[cs] = p.bit[cs]; // prepareRightDecoding: Bit <-> LeftBit
            p.llr[cs].resize(h);
            for (int i = 0; i < h; ++i)
                p.llr[cs][i] = g_i8(p.llr[stage][i], p.llr[stage][i + h], p.leftBit[cs][i]);
        }
        decode(fm + h, cs);
        tally(st, ORACLE_NODE_COMBINE, (long)h);
        for (auto& p : paths) {
            p.bit[stage].resize(n);
            for (int i = 0; i < h; ++i) {
                p.bit[stage][i] = (int8_t)(p.leftBit[cs][i] ^ p.bit[cs][i]);
                p.bit[stage][i + h] = p.bit[cs][i];
            }
        }
    }
};

int scl_i8_decode(int N, int L, const uint8_t* fm, int systematic, int crcBits, const int8_t* llr,
                  uint8_t* out)
{
    int stages = 0;
    while ((1 << stages) < N) ++stages;
    Scl s;
    s.L = L;
    s.paths.resize(1);
    Path& p0 = s.paths[0];
    p0.llr.resize(stages + 1); p0.bit.resize(stages + 1); p0.leftBit.resize(stages + 1);
    p0.llr[stages].assign(llr, llr + N);
    s.decode(fm, stages);
    // extractBestPath: scl_fip_char.cpp:816-855
    int K = N - count_frozen(fm, N), kb = (K + 7) / 8;
    std::vector<uint8_t> hard(N);
    for (size_t p = 0; p < s.paths.size(); ++p) {
        for (int i = 0; i < N; ++i) hard[i] = (s.paths[p].bit[stages][i] & 0x80) ? 1 : 0;
        extract_info(hard.data(), fm, N, systematic, out);
        if (crc_check(crcBits, out, kb)) return 1;
    }
    for (int i = 0; i < N; ++i) hard[i] = (s.paths[0].bit[stages][i] & 0x80) ? 1 : 0;
    extract_info(hard.data(), fm, N, systematic, out);
    return 0;
}

void scl_i8_stats(int N, const uint8_t* fm, long* counts, long* sizes)
{
    int stages = 0;
    while ((1 << stages) < N) ++stages;
    Scl s;
    s.L = 1;
    Stats st{ counts, sizes };
    s.st = &st;
    s.paths.resize(1);
    Path& p0 = s.paths[0];
    p0.llr.resize(stages + 1); p0.bit.resize(stages + 1); p0.leftBit.resize(stages + 1);
    p0.llr[stages].assign(N, 1);
    s.decode(fm, stages);
}

} // namespace oracle
