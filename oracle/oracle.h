/* TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's polar-code hot path.
 *
 * Plain scalar C++ (no intrinsics), one frame at a time, written from the reference's
 * algorithm (david13pod/antPolarCodes; file:line citations at each function).  It is
 * pinned against (a) the reference's own golden vectors / known answers and (b) the
 * unmodified reference compiled into oracle/_ref/libpolar_ref.so (tests/test_oracle_*.py).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load this library.  Nothing in antpolarcodes_b200/ links or calls it.
 */
#ifndef POLAR_ORACLE_H
#define POLAR_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* kinds: same numbering as include/polar_cuda.h */
enum { ORACLE_ENC = 0, ORACLE_FASTSSC_I8 = 1, ORACLE_FASTSSC_F32 = 2, ORACLE_SCL_I8 = 3, ORACLE_SCL_F32 = 4 };

/* CRC detectors (crc8.cpp:18-56, crc16.cpp:21-43, crc32.cpp:28-66). bits in {0,8,16,32}. */
void oracle_crc_generate(int bits, uint8_t* data, int bytes);
int oracle_crc_check(int bits, const uint8_t* data, int bytes);

/* float -> int8 quantiser (bitcontainer.cpp:27-39,449-516) */
void oracle_quantize(const float* llr, int8_t* out, int N);

/* int8 primitives, exposed for the golden-vector tests (fip_char.h:44-73) */
int8_t oracle_f_i8(int8_t a, int8_t b);
int8_t oracle_g_i8(int8_t a, int8_t b, int8_t bit);
int8_t oracle_rep_sum_i8(const int8_t* x, int n);

/* Rate matching: Puncturer (puncturer.cpp:45-85, puncturer.h:71-100).  pos = kept positions of the parent block. */
int oracle_puncturer_positions(int blockLength, const uint32_t* frozen, int nFrozen, uint32_t* pos);
void oracle_puncture(const uint32_t* pos, int nPos, int elem, const uint8_t* in, uint8_t* out);
void oracle_depuncture(const uint32_t* pos, int nPos, int parent, int elem, const uint8_t* in, uint8_t* out);
void oracle_puncture_packed(const uint32_t* pos, int nPos, const uint8_t* in, uint8_t* out);

/* Encoder (butterfly_fip_packed.cpp:45-70).  info: nFrames x K/8 (CRC written in place),
 * code: nFrames x N/8, both MSB-first packed. */
int oracle_encode(int N, const uint32_t* frozen, int nFrozen, int systematic, int crcBits,
                  uint8_t* info, uint8_t* code, long nFrames);

/* Decoders.  llr: nFrames x N (int8 or float per kind); out: nFrames x ceil(K/8); ok: nFrames.
 * Returns 0 or a negative error code. */
int oracle_decode(int kind, int N, const uint32_t* frozen, int nFrozen, int listSize,
                  int systematic, int crcBits, const void* llr, long nFrames, uint8_t* out,
                  uint8_t* ok);

/* Node histogram of the decoder tree the reference would build (for schedule-compiler tests).
 * counts[16], sizes[16] indexed by the ORACLE_NODE_* ids below. */
enum {
    ORACLE_NODE_F = 0, ORACLE_NODE_G, ORACLE_NODE_COMBINE, ORACLE_NODE_G_R1, ORACLE_NODE_G_0R,
    ORACLE_NODE_R0, ORACLE_NODE_R1, ORACLE_NODE_REP, ORACLE_NODE_REP2, ORACLE_NODE_SPC,
    ORACLE_NODE_ZERO_SPC, ORACLE_NODE_ZERO_ONE, ORACLE_NODE_COUNT
};
int oracle_tree_stats(int kind, int N, const uint32_t* frozen, int nFrozen, long* counts, long* sizes);

#ifdef __cplusplus
}
#endif
#endif
