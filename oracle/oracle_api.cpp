/* TEST INFRASTRUCTURE ONLY -- extern "C" entry points of the CPU oracle (see oracle.h). */
#include "oracle.h"
#include "oracle_common.h"

namespace oracle {
int fastssc_i8_decode(int N, const uint8_t* fm, int systematic, int crcBits, const int8_t* llr, uint8_t* out);
void fastssc_i8_soft(int N, const uint8_t* fm, const int8_t* llr, int8_t* soft);
void fastssc_i8_stats(int N, const uint8_t* fm, long* counts, long* sizes);
int scl_i8_decode(int N, int L, const uint8_t* fm, int systematic, int crcBits, const int8_t* llr, uint8_t* out);
void scl_i8_stats(int N, const uint8_t* fm, long* counts, long* sizes);
int fastssc_f32_decode(int N, const uint8_t* fm, int systematic, int crcBits, const float* llr, uint8_t* out);
int scl_f32_decode(int N, int L, const uint8_t* fm, int systematic, int crcBits, const float* llr, uint8_t* out);
}
using namespace oracle;

extern "C" int oracle_decode(int kind, int N, const uint32_t* frozen, int nFrozen, int listSize,
                             int systematic, int crcBits, const void* llr, long nFrames,
                             uint8_t* out, uint8_t* ok)
{
    if (N < 2 || (N & (N - 1))) return -1;
    std::vector<uint8_t> fm = frozen_mask(N, frozen, nFrozen);
    int K = N - nFrozen, kb = (K + 7) / 8;
    for (long f = 0; f < nFrames; ++f) {
        int r;
        switch (kind) {
        case ORACLE_FASTSSC_I8:
            r = fastssc_i8_decode(N, fm.data(), systematic, crcBits, (const int8_t*)llr + f * N, out + f * kb);
            break;
        case ORACLE_SCL_I8:
            r = scl_i8_decode(N, listSize, fm.data(), systematic, crcBits, (const int8_t*)llr + f * N, out + f * kb);
            break;
        case ORACLE_FASTSSC_F32:
            r = fastssc_f32_decode(N, fm.data(), systematic, crcBits, (const float*)llr + f * N, out + f * kb);
            break;
        case ORACLE_SCL_F32:
            r = scl_f32_decode(N, listSize, fm.data(), systematic, crcBits, (const float*)llr + f * N, out + f * kb);
            break;
        default:
            return -2;
        }
        if (r < 0) return r;
        if (ok) ok[f] = (uint8_t)r;
    }
    return 0;
}

extern "C" int oracle_tree_stats(int kind, int N, const uint32_t* frozen, int nFrozen, long* counts, long* sizes)
{
    std::vector<uint8_t> fm = frozen_mask(N, frozen, nFrozen);
    for (int i = 0; i < ORACLE_NODE_COUNT; ++i) { counts[i] = 0; sizes[i] = 0; }
    if (kind == ORACLE_FASTSSC_I8) fastssc_i8_stats(N, fm.data(), counts, sizes);
    else if (kind == ORACLE_SCL_I8) scl_i8_stats(N, fm.data(), counts, sizes);
    else return -2;
    return 0;
}

// Soft codeword of the int8 Fast-SSC decoder (Decoder::getSoftCodeword, decoder.cpp:147): nFrames x N bytes
extern "C" int oracle_fastssc_i8_soft(int N, const uint32_t* frozen, int nFrozen, const int8_t* llr, long nFrames, int8_t* soft)
{
    std::vector<uint8_t> fm = frozen_mask(N, frozen, nFrozen);
    for (long f = 0; f < nFrames; ++f) fastssc_i8_soft(N, fm.data(), llr + f * N, soft + f * N);
    return 0;
}
