// Kernel parameter blocks shared by the host launcher and the decoder kernels.
#pragma once
#include <cstdint>

namespace polar {

struct DecodeParams {
    const void* llr;          // nFrames x N int8 (float inputs are quantised by a pre-pass)
    uint8_t* out;             // nFrames x kb
    uint8_t* ok;              // nFrames
    uint8_t* cand;            // list decoders, candidate hand-back: nFrames x listSize x kb (all surviving paths in list
                              // order, no detector on the device; ok[] then holds the number of paths), else null
    const uint16_t* infoPos;  // K code positions of the information bits, ascending
    const uint32_t* crcTab;   // K syndrome contributions (see crc_host.h)
    const uint32_t* runs;     // Fast-SSC: runs of consecutive information positions (pos | len << 16)
    uint32_t* spill;          // L2-resident scratch for the top LLR stage (list decoders), or null
    long long nFrames;
    uint32_t crcInit;         // syndrome of the all-zero info block
    int N, lgN, K, kb;        // kb = ceil(K/8)
    int crcBits, systematic;
    int listSize;
    // shared-memory geometry, in 32-bit words
    int slotW;                // list decoders: LLR slot stride (one path, stages below smemStages)
    int bitBW;                // list decoders: offset of bit region "B" inside a bit slot (= words of region A)
    int bitB1;                // list decoders: offset of the odd-stage half of region B (B is a ping-pong pair)
    int bitSlotW;             // list decoders: bit slot stride (regions A and B)
    int rootLeaf;             // the whole code is one leaf: region A also holds stage lgN
    int frameW;               // list decoders: stride between the frames of one warp
    int scratchW;             // per-warp scratch
    int warpW;                // total words per warp
    int smemStages;           // LLR stages [0, smemStages) live in shared memory, [smemStages, lgN) in `spill`
    int spillWarpW;           // words of `spill` per resident warp
    int lockstep;             // warps of a block meet at a barrier before every schedule step
    int bitsOff, selOff;      // list decoders: word offsets of the bit slots / the select scratch inside a frame
    int llrBase;              // int8 list decoder: llr_woff of the lowest stage kept in the shared-memory slots
    int bitBase, bitSpill;    // list decoders: bit_woff of the lowest bit stage in shared memory; first bit stage in the L2 scratch
    int spA, spB0, spB1;      // word offsets of the spilled bit regions inside the warp's scratch
    int regStages;            // int8 list decoder: stages <= regStages live in registers (min(4, lgN - 2), or -1: none)
    // Fast-SSC (one frame per thread): rows of 32 words
    int llrRows, bitRows, tileRow, outRow, chanRow, nRuns;
    int chanVec;              // the channel rows are 16-byte aligned (fp32 list decoder: 128-bit channel loads)
    int prefetchNext;         // Fast-SSC kernels: a warp prefetches the channel rows of its next frame group into L2
    int useTma, mbarRow;      // Fast-SSC int8: channel tiles filled by TMA bulk copies (one 128-byte row segment per lane)
};

// word offset of stage s inside an llr slot (stage s block = 2^s int8, at least one word)
__host__ __device__ inline int llr_woff(int s) { return s < 2 ? s : (1 << (s - 2)) + 1; }
// word offset of stage s inside a bit array (stage s block = 2^s bits, at least one word)
__host__ __device__ inline int bit_woff(int s) { return s <= 5 ? s : (1 << (s - 5)) + 4; }

// fp32 decoders: word (row) offset of stage s (2^s floats)
__host__ __device__ inline int f32_woff(int s) { return (1 << s) - 1; }

} // namespace polar
