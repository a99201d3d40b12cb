// Kernel parameter blocks shared by the host launcher and the decoder kernels.
#pragma once
#include <cstdint>

namespace polar {

struct DecodeParams {
    const void* llr;          // nFrames x N, int8 or float
    uint8_t* out;             // nFrames x kb
    uint8_t* ok;              // nFrames
    const uint16_t* infoPos;  // K code positions of the information bits, ascending
    const uint32_t* crcTab;   // K syndrome contributions (see crc_tables.h)
    long long nFrames;
    uint32_t crcInit;         // syndrome of the all-zero info block
    int N, lgN, K, kb;        // kb = ceil(K/8)
    int crcBits, systematic, llrIsFloat;
    int listSize;
    // shared-memory geometry per warp, in 32-bit words
    int chanW;                // N/4
    int slotW;                // llr slot stride
    int bitW;                 // words of one bit array (all stages)
    int bitSlotW;             // bit slot stride (two arrays + pad)
    int scratchW;
    int warpW;                // total words per warp
};

// word offset of stage s inside an llr slot (stage s block = 2^s int8, at least one word)
__host__ __device__ inline int llr_woff(int s) { return s < 2 ? s : (1 << (s - 2)) + 1; }
// word offset of stage s inside a bit array (stage s block = 2^s bits, at least one word)
__host__ __device__ inline int bit_woff(int s) { return s <= 5 ? s : (1 << (s - 5)) + 4; }

} // namespace polar
