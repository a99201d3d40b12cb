// Flat decoder schedules: the host compiles a frozen set into a linear program of node ops
// (the reference builds a tree of virtual node objects instead: fastssc_fip_char.cpp:496-569,
// scl_fip_char.cpp:729-755).  The program travels to the kernel as a __grid_constant__ kernel
// parameter, i.e. it sits in the constant bank and every warp reads it with uniform loads.
#pragma once
#include <cstdint>
#include <vector>

namespace polar {

enum OpCode : uint32_t {
    OP_END = 0,
    // inner-node steps (every decoder kind)
    OP_F,           // child llr  = f(parent first half, second half)
    OP_G,           // child llr  = g(parent halves, left bits)
    OP_COMBINE,     // parent bits = [left ^ right | right]
    // Fast-SSC shortcuts (L = 1 only)
    OP_G_R1,        // right child is rate-1: fused g + hard decision + combine (ROneNode)
    OP_G_0R,        // left child is rate-0: child llr = sat(a + b)              (ZeroRNode)
    OP_COMBINE_0R,  // parent bits = [right | right]
    OP_SC_R0, OP_SC_R1, OP_SC_REP, OP_SC_REP2, OP_SC_SPC, OP_SC_ZERO_SPC, OP_SC_ZERO_ONE,
    // list-decoder leaves
    OP_SCL_R0, OP_SCL_R1, OP_SCL_REP, OP_SCL_SPC,
    // extra leaves of the fp32 Fast-SSC tree (fastssc_avx_float.cpp:797-896)
    OP_F32_SPC2,      // DoubleSpcDecoder (n >= 16)
    OP_F32_SPC2_8,    // DoubleSpcDecoderShort8
    OP_F32_REP3,      // TripleRepetitionDecoder
    OP_F32_TYPE5,     // TypeFiveDecoder
    OP_F32_REPR1_8,   // RepetitionRateOneDecoderShort8
    OP_F32_ZSPC_8,    // ZeroSpcDecoderShort8
    OP_COUNT
};

// op word: code[5:0] | stage[9:6] | lgPaths[12:10] | flags[15:13] | pos[31:16]
//   stage   : log2 of the block the op produces / decodes (children stage for F/G/COMBINE)
//   lgPaths : log2 of the live path count entering the op (list decoders)
//   flags   : bit0 = the node is a right child (its bits go to bit array B), bit1 = "long form"
//   pos     : absolute bit position of the block within the codeword (Fast-SSC, in units of 1 bit)
static constexpr uint32_t OPF_RIGHT = 1u, OPF_LONG = 2u;
inline constexpr uint32_t make_op(uint32_t code, uint32_t stage, uint32_t lgPaths, uint32_t flags, uint32_t pos)
{
    return code | (stage << 6) | (lgPaths << 10) | (flags << 13) | (pos << 16);
}
#define POLAR_OP_CODE(w) ((w) & 63u)
#define POLAR_OP_STAGE(w) (((w) >> 6) & 15u)
#define POLAR_OP_LGP(w) (((w) >> 10) & 7u)
#define POLAR_OP_FLAGS(w) (((w) >> 13) & 7u)
#define POLAR_OP_POS(w) ((w) >> 16)

// List decoders (int8): the host stores a dense handler id in bits 31..16 of every op word (`pos` is a Fast-SSC field), so
// the kernel dispatches with one jump table.  regTop = highest stage held in registers (-1: none).
enum SclHandler : uint32_t {
    SCL_H_END = 0,
    SCL_H_RS_F = 1, SCL_H_RS_G = 5, SCL_H_RS_C = 9, SCL_H_RS_R0 = 13, SCL_H_RS_REP = 17, SCL_H_RS_R1 = 21, SCL_H_RS_SPC = 25, // + (stage - 1)
    SCL_H_F = 29, SCL_H_G, SCL_H_C, SCL_H_R0, SCL_H_REP, SCL_H_R1, SCL_H_SPC
};
inline uint32_t scl_handler_id(uint32_t op, int regTop)
{
    const uint32_t code = op & 63u, stage = (op >> 6) & 15u;
    const bool rs = (int)stage <= regTop && stage >= 1;
    switch (code) {
    case OP_F: return rs ? SCL_H_RS_F + stage - 1 : SCL_H_F;
    case OP_G: return rs ? SCL_H_RS_G + stage - 1 : SCL_H_G;
    case OP_COMBINE: return rs ? SCL_H_RS_C + stage - 1 : SCL_H_C;
    case OP_SCL_R0: return rs ? SCL_H_RS_R0 + stage - 1 : SCL_H_R0;
    case OP_SCL_REP: return rs ? SCL_H_RS_REP + stage - 1 : SCL_H_REP;
    case OP_SCL_R1: return rs ? SCL_H_RS_R1 + stage - 1 : SCL_H_R1;
    case OP_SCL_SPC: return rs ? SCL_H_RS_SPC + stage - 1 : SCL_H_SPC;
    }
    return SCL_H_END;
}

// Fast-SSC int8 kernel: the host replaces the op code (bits 5..0) by a dense handler id.  1 + 4 * idx + (stage - 1) for the
// steps on register stages (idx: F, G, G-0R, G-R1, combine, combine-0R, R0, R1, REP, SPC, 0-SPC, 0-1), SC_H_GENERIC + code - 1
// for everything else.
// SC_H_BIG + 4 * (stage - 5) + {F, G, G-0R, G-R1} for the inner steps of the stages 5 .. lgN-2 (7-bit ids: bit 6 travels
// in bit 10 of the op word, a list-decoder field).
enum ScHandler : uint32_t { SC_H_END = 0, SC_H_GENERIC = 49, SC_H_BIG = 64 };
inline uint32_t sc_handler_id(uint32_t op, int regTop, int lgN)
{
    const uint32_t code = op & 63u, stage = (op >> 6) & 15u;
    if (code == OP_END) return SC_H_END;
    if ((int)stage >= 5 && (int)stage > regTop && (int)stage <= lgN - 2 && (int)stage <= 10 && regTop >= 0) {
        const int k = code == OP_F ? 0 : code == OP_G ? 1 : code == OP_G_0R ? 2 : code == OP_G_R1 ? 3 : -1;
        if (k >= 0) return SC_H_BIG + 4u * (stage - 5u) + (uint32_t)k;
    }
    static const int idx[16] = { -1, 0, 1, 4, 3, 2, 5, 6, 7, 8, -1, 9, 10, 11, -1, -1 }; // by OpCode (REP2 has >= 32 elements)
    const int k = code < 16 ? idx[code] : -1;
    if ((int)stage <= regTop && stage >= 1 && k >= 0) return 1u + 4u * (uint32_t)k + (stage - 1u);
    return SC_H_GENERIC + code - 1u;
}

static constexpr int MAX_OPS = 3072;
struct Schedule {
    uint32_t ops[MAX_OPS];
};

struct ScheduleStats {
    long count[OP_COUNT] = {};
    long size[OP_COUNT] = {};   // sum of block sizes (elements) per op kind
    double laneOps = 0;         // algorithmic lane-ops per frame (SURVEY 8d weights)
    double pathElems = 0;       // sum over F/G/C ops of h * paths
};

// tree flavours (node-selection rules differ per reference decoder)
enum TreeKind { TREE_FASTSSC_I8 = 0, TREE_FASTSSC_F32 = 1, TREE_SCL_I8 = 2, TREE_SCL_F32 = 3 };

// frozen: sorted ascending frozen indices.  listSize must be a power of two <= 32.
// Returns false (and sets err) if the code cannot be scheduled.
bool compile_schedule(TreeKind kind, int N, const std::vector<uint32_t>& frozen, int listSize,
                      std::vector<uint32_t>& ops, ScheduleStats& stats, const char** err);

} // namespace polar
