// Host-side CRC detectors and the linear "syndrome tables" the kernels use.
//
// Same arithmetic as the reference detectors:
//   CRC8  : poly 0x07, init 0, MSB first, 1 trailing byte            (errordetection/crc8.cpp:18-56)
//   CRC16 : CRC-16/CCITT-FALSE poly 0x1021 init 0xFFFF, trailer hi,lo (errordetection/crc16.cpp:21-43)
//   CRC32 : CRC-32C as the x86 crc32 instruction over LE u32 words, init 0, no final xor, native u32
//           trailer in word (bytes/4 - 1)                             (errordetection/crc32.cpp:28-66)
// A CRC is affine over GF(2), so check(info) == [ S0 ^ XOR_{k : info bit k = 1} T[k] == 0 ] where
// S is "computed checksum XOR received trailer".  The device evaluates that with one table lookup per
// set bit while it gathers the information bits, and the encoder uses the same trick to generate.
#pragma once
#include <cstdint>
#include <vector>

namespace polar {

uint32_t crc_compute(int crcBits, const uint8_t* data, int bytes);        // checksum over the payload part
void crc_generate(int crcBits, uint8_t* data, int bytes);                  // Detector::generate
bool crc_check(int crcBits, const uint8_t* data, int bytes);               // Detector::check
uint32_t crc_syndrome(int crcBits, const uint8_t* data, int bytes);        // computed ^ received

struct CrcTables {
    uint32_t synInit = 0;            // syndrome of the all-zero block
    std::vector<uint32_t> syn;       // per info bit k: syndrome contribution
    uint32_t genInit = 0;            // checksum of the all-zero payload
    std::vector<uint32_t> gen;       // per info bit k: checksum contribution (0 for trailer bits)
};
// K info bits; syn covers the ceil(K/8) zero-padded bytes a decoder checks, gen the floor(K/8) bytes an encoder signs
CrcTables build_crc_tables(int crcBits, int K);

} // namespace polar
