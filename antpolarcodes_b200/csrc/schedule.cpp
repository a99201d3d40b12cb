// ScheduleCompiler: frozen set -> flat node program (see schedule.h).
// Node-selection rules restate the reference's createDecoder() functions:
//   int8 Fast-SSC : src/polarcode/decoding/fastssc_fip_char.cpp:496-569
//   int8 / fp32 SCL: src/polarcode/decoding/scl_fip_char.cpp:729-755, scl_avx_float.cpp:624-651
// (splitFrozenBits, src/polarcode/polarcode.cpp:14-34, becomes a prefix-sum lookup.)
#include "schedule.h"

#include <algorithm>
#include <cmath>

namespace polar {
namespace {

struct Compiler {
    TreeKind kind;
    int N, lgN, lgL;
    std::vector<int> prefix; // prefix[i] = number of frozen indices < i
    std::vector<uint32_t>* ops;
    ScheduleStats* st;
    const char* err = nullptr;
    int lgP = 0; // live paths (log2) while walking the SCL tree

    int frozenIn(int pos, int n) const { return prefix[pos + n] - prefix[pos]; }

    void emit(uint32_t code, int stage, uint32_t flags, int pos, double elemOps)
    {
        ops->push_back(make_op(code, (uint32_t)stage, (uint32_t)lgP, flags, (uint32_t)pos));
        st->count[code]++;
        st->size[code] += 1L << stage;
        const double paths = (double)(1 << lgP);
        const bool f32 = (kind == TREE_FASTSSC_F32 || kind == TREE_SCL_F32);
        st->laneOps += elemOps * paths / (f32 ? 1.0 : 4.0); // 4 int8 or 1 fp32 per lane-op
        if (code == OP_F || code == OP_G || code == OP_COMBINE) st->pathElems += (double)(1L << stage) * paths;
    }

    void sortCost(int lgFanout)
    {
        int lgM = lgP + lgFanout; // candidates = paths * fanout (already a power of two)
        double M = (double)(1 << lgM), S = lgM * (lgM + 1) / 2.0;
        st->laneOps += 4.0 * M * S;
    }

    // ---- list decoders: only R0 / R1 / REP / SPC leaves, else split
    void scl(int pos, int stage, bool right)
    {
        const int n = 1 << stage, fc = frozenIn(pos, n);
        const uint32_t fl = right ? OPF_RIGHT : 0u;
        const bool isFloat = (kind == TREE_SCL_F32);
        if (fc == n) { emit(OP_SCL_R0, stage, fl, pos, 1.0 * n); return; }
        if (fc == 0) {
            if (n < 2) { err = "rate-1 leaf of length 1 is not supported"; return; }
            emit(OP_SCL_R1, stage, fl, pos, 3.0 * n);
            sortCost(2);
            lgP = std::min(lgL, lgP + 2);
            return;
        }
        // scl_avx_float.cpp:637-639: the fp32 list decoder only has a REP leaf for n < 8
        if (fc == n - 1 && (!isFloat || n < 8)) {
            emit(OP_SCL_REP, stage, fl, pos, 3.0 * n);
            sortCost(1);
            lgP = std::min(lgL, lgP + 1);
            return;
        }
        if (fc == 1) {
            emit(OP_SCL_SPC, stage, fl, pos, 4.0 * n);
            sortCost(3);
            lgP = std::min(lgL, lgP + 3);
            return;
        }
        if (stage == 0) { err = "unexpected length-1 inner node"; return; }
        const int h = n / 2, cs = stage - 1;
        emit(OP_F, cs, 0, pos, 5.0 * h);
        scl(pos, cs, false);
        emit(OP_G, cs, 0, pos, 2.0 * h);
        scl(pos + h, cs, true);
        emit(OP_COMBINE, cs, fl, pos, h / 8.0);
    }

    // ---- int8 Fast-SSC (BYTESPERVECTOR = 32 decides short/long forms)
    void fastssc(int pos, int stage)
    {
        const int n = 1 << stage, fc = frozenIn(pos, n);
        const int VEC = 32;
        if (fc == n) { emit(OP_SC_R0, stage, 0, pos, 0); return; }
        if (fc == 0) { emit(OP_SC_R1, stage, 0, pos, 0); return; }
        if (fc == n - 1) { emit(OP_SC_REP, stage, 0, pos, 1.0 * n); return; }
        if (fc == 1) { emit(OP_SC_SPC, stage, n > VEC ? OPF_LONG : 0u, pos, 3.0 * n); return; }
        if (fc == n - 2 && n >= VEC) { emit(OP_SC_REP2, stage, 0, pos, 1.0 * n); return; }
        const int h = n / 2, cs = stage - 1;
        const int lf = frozenIn(pos, h), rf = frozenIn(pos + h, h);
        if (lf == h && rf == 0 && n <= VEC) { emit(OP_SC_ZERO_ONE, stage, 0, pos, 0.5 * n); return; }
        if (lf == h && rf == 1) { emit(OP_SC_ZERO_SPC, stage, n > VEC ? OPF_LONG : 0u, pos, 2.0 * n); return; }
        if (rf == 0) { // ROneNode / ShortROneNode
            emit(OP_F, cs, 0, pos, 5.0 * h);
            fastssc(pos, cs);
            emit(OP_G_R1, cs, 0, pos, 2.0 * h);
            return;
        }
        if (lf == h) { // ZeroRNode / ShortZeroRNode
            emit(OP_G_0R, cs, 0, pos, 1.0 * h);
            fastssc(pos + h, cs);
            emit(OP_COMBINE_0R, cs, 0, pos, 0);
            return;
        }
        emit(OP_F, cs, 0, pos, 5.0 * h);
        fastssc(pos, cs);
        emit(OP_G, cs, 0, pos, 2.0 * h);
        fastssc(pos + h, cs);
        emit(OP_COMBINE, cs, 0, pos, h / 8.0);
    }

    // ---- fp32 Fast-SSC (fastssc_avx_float.cpp:797-896): more leaf kinds, no 32-byte vector rules
    int lastFrozen(int pos, int n, int k) const // k-th frozen index from the end of the block, relative
    {
        for (int i = n - 1; i >= 0; --i)
            if (frozenIn(pos + i, 1) && --k == 0) return i;
        return -1;
    }
    void fastsscF32(int pos, int stage)
    {
        const int n = 1 << stage, fc = frozenIn(pos, n);
        if (fc == n) { emit(OP_SC_R0, stage, 0, pos, 0); return; }
        if (fc == 0) { emit(OP_SC_R1, stage, 0, pos, 0); return; }
        if (fc == n - 1) { emit(OP_SC_REP, stage, 0, pos, 1.0 * n); return; }
        if (fc == 1) { emit(OP_SC_SPC, stage, 0, pos, 3.0 * n); return; }
        if (fc == n - 2) {
            // DoubleRepetitionDecoder: the reference throws unless the frozen bits are the first n-2 (:816-823)
            if (frozenIn(pos, n - 2) != n - 2) { err = "fp32 Fast-SSC: double-repetition pattern with scattered frozen bits"; return; }
            emit(OP_SC_REP2, stage, 0, pos, 1.0 * n);
            return;
        }
        if (fc == 2 && frozenIn(pos, 2) == 2) { emit(n == 8 ? OP_F32_SPC2_8 : OP_F32_SPC2, stage, 0, pos, 3.0 * n); return; }
        if (fc == n - 3 && n > 8 && lastFrozen(pos, n, 1) == n - 4) { emit(OP_F32_REP3, stage, 0, pos, 1.0 * n); return; }
        if (fc == n - 4 && lastFrozen(pos, n, 1) == n - 4 && lastFrozen(pos, n, 2) == n - 6) {
            emit(OP_F32_TYPE5, stage, 0, pos, 1.0 * n);
            return;
        }
        if (n == 8 && fc == 3 && frozenIn(pos, 3) == 3) { emit(OP_F32_REPR1_8, stage, 0, pos, 2.0 * n); return; }
        if (n == 8 && fc == 5 && lastFrozen(pos, n, 1) == 4 && lastFrozen(pos, n, 2) == 3) {
            emit(OP_F32_ZSPC_8, stage, 0, pos, 2.0 * n);
            return;
        }
        if (stage == 0) { err = "unexpected length-1 inner node"; return; }
        const int h = n / 2, cs = stage - 1;
        const int lf = frozenIn(pos, h), rf = frozenIn(pos + h, h);
        if (n > 8) {
            if (lf == h && rf == 1) { emit(OP_SC_ZERO_SPC, stage, 0, pos, 2.0 * n); return; }
            if (rf == 0) {
                emit(OP_F, cs, 0, pos, 5.0 * h);
                fastsscF32(pos, cs);
                emit(OP_G_R1, cs, 0, pos, 2.0 * h);
                return;
            }
            if (lf == h) {
                emit(OP_G_0R, cs, 0, pos, 1.0 * h);
                fastsscF32(pos + h, cs);
                emit(OP_COMBINE_0R, cs, 0, pos, 0);
                return;
            }
        }
        emit(OP_F, cs, 0, pos, 5.0 * h);
        fastsscF32(pos, cs);
        emit(OP_G, cs, 0, pos, 2.0 * h);
        fastsscF32(pos + h, cs);
        emit(OP_COMBINE, cs, 0, pos, h / 32.0);
    }
};

} // namespace

bool compile_schedule(TreeKind kind, int N, const std::vector<uint32_t>& frozen, int listSize,
                      std::vector<uint32_t>& ops, ScheduleStats& stats, const char** err)
{
    static const char* e_n = "block length must be a power of two in [8, 4096]";
    static const char* e_l = "list size must be a power of two in [1, 32]";
    static const char* e_f = "frozen indices must be sorted, unique and < N";
    static const char* e_ops = "schedule too long";
    static const char* e_kind = "decoder kind not implemented";
    if (N < 8 || N > 4096 || (N & (N - 1))) { *err = e_n; return false; }
    if (listSize < 1 || listSize > 32 || (listSize & (listSize - 1))) { *err = e_l; return false; }
    for (size_t i = 0; i < frozen.size(); ++i)
        if (frozen[i] >= (uint32_t)N || (i && frozen[i] <= frozen[i - 1])) { *err = e_f; return false; }
    Compiler c;
    c.kind = kind;
    c.N = N;
    c.lgN = 0;
    while ((1 << c.lgN) < N) ++c.lgN;
    c.lgL = 0;
    while ((1 << c.lgL) < listSize) ++c.lgL;
    c.prefix.assign(N + 1, 0);
    {
        std::vector<uint8_t> fm(N, 0);
        for (uint32_t f : frozen) fm[f] = 1;
        for (int i = 0; i < N; ++i) c.prefix[i + 1] = c.prefix[i] + fm[i];
    }
    ops.clear();
    stats = ScheduleStats();
    c.ops = &ops;
    c.st = &stats;
    switch (kind) {
    case TREE_SCL_I8:
    case TREE_SCL_F32: c.scl(0, c.lgN, false); break;
    case TREE_FASTSSC_I8: c.fastssc(0, c.lgN); break;
    case TREE_FASTSSC_F32: c.fastsscF32(0, c.lgN); break;
    default: *err = e_kind; return false;
    }
    if (c.err) { *err = c.err; return false; }
    ops.push_back(make_op(OP_END, 0, 0, 0, 0));
    if ((int)ops.size() > MAX_OPS) { *err = e_ops; return false; }
    // epilogue: info gather + CRC over the surviving paths
    const int K = N - (int)frozen.size();
    stats.laneOps += 8.0 * (K / 8.0) * (double)(1 << c.lgP);
    return true;
}

} // namespace polar
