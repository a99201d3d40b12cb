#include "crc_host.h"

namespace polar {
namespace {
uint8_t crc8_bytes(const uint8_t* d, int n)
{
    uint8_t c = 0;
    for (int i = 0; i < n; ++i) {
        c ^= d[i];
        for (int j = 0; j < 8; ++j) c = (uint8_t)((c << 1) ^ ((c & 0x80) ? 0x07 : 0));
    }
    return c;
}
uint16_t crc16_bytes(const uint8_t* d, int n)
{
    uint16_t c = 0xFFFF;
    for (int i = 0; i < n; ++i) {
        c ^= (uint16_t)(d[i] << 8);
        for (int j = 0; j < 8; ++j) c = (uint16_t)((c << 1) ^ ((c & 0x8000) ? 0x1021 : 0));
    }
    return c;
}
uint32_t crc32c_bytes(const uint8_t* d, int n)
{
    uint32_t c = 0;
    for (int i = 0; i < n; ++i) {
        c ^= d[i];
        for (int j = 0; j < 8; ++j) c = (c >> 1) ^ ((c & 1u) ? 0x82F63B78u : 0u);
    }
    return c;
}
// index of the first trailer byte
int trailer_offset(int crcBits, int bytes)
{
    if (crcBits == 8) return bytes - 1;
    if (crcBits == 16) return bytes - 2;
    if (crcBits == 32) return 4 * ((bytes >> 2) - 1);
    return bytes;
}
} // namespace

uint32_t crc_compute(int crcBits, const uint8_t* data, int bytes)
{
    const int n = trailer_offset(crcBits, bytes);
    if (crcBits == 8) return crc8_bytes(data, n);
    if (crcBits == 16) return crc16_bytes(data, n);
    if (crcBits == 32) return crc32c_bytes(data, n);
    return 0;
}

static uint32_t received(int crcBits, const uint8_t* data, int bytes)
{
    const int t = trailer_offset(crcBits, bytes);
    if (crcBits == 8) return data[t];
    if (crcBits == 16) return ((uint32_t)data[t] << 8) | data[t + 1];
    if (crcBits == 32)
        return (uint32_t)data[t] | ((uint32_t)data[t + 1] << 8) | ((uint32_t)data[t + 2] << 16) |
               ((uint32_t)data[t + 3] << 24);
    return 0;
}

void crc_generate(int crcBits, uint8_t* data, int bytes)
{
    const uint32_t c = crc_compute(crcBits, data, bytes);
    const int t = trailer_offset(crcBits, bytes);
    if (crcBits == 8) data[t] = (uint8_t)c;
    if (crcBits == 16) { data[t] = (uint8_t)(c >> 8); data[t + 1] = (uint8_t)c; }
    if (crcBits == 32)
        for (int i = 0; i < 4; ++i) data[t + i] = (uint8_t)(c >> (8 * i));
}

uint32_t crc_syndrome(int crcBits, const uint8_t* data, int bytes)
{
    if (!crcBits) return 0;
    return crc_compute(crcBits, data, bytes) ^ received(crcBits, data, bytes);
}

bool crc_check(int crcBits, const uint8_t* data, int bytes) { return crc_syndrome(crcBits, data, bytes) == 0; }

CrcTables build_crc_tables(int crcBits, int K)
{
    CrcTables t;
    t.syn.assign(K, 0);
    t.gen.assign(K, 0);
    if (!crcBits) return t;
    // decoders check ceil(K/8) bytes (zero padded: fastssc_fip_char.cpp:627, scl_fip_char.cpp:808), the encoder
    // generates over floor(K/8) (butterfly_fip_packed.cpp:48): the two differ when K is not a multiple of 8
    const int cb = (K + 7) / 8, gb = K / 8;
    std::vector<uint8_t> buf(cb, 0);
    const bool canCheck = cb >= crcBits / 8 && (crcBits != 32 || cb >= 4);
    const bool canGen = gb >= crcBits / 8 && (crcBits != 32 || gb >= 4);
    if (canCheck) t.synInit = crc_syndrome(crcBits, buf.data(), cb);
    if (canGen) t.genInit = crc_compute(crcBits, buf.data(), gb);
    const int tr = trailer_offset(crcBits, gb);
    for (int k = 0; k < K; ++k) {
        buf[k >> 3] = (uint8_t)(0x80 >> (k & 7));
        if (canCheck) t.syn[k] = crc_syndrome(crcBits, buf.data(), cb) ^ t.synInit;
        if (canGen) t.gen[k] = ((k >> 3) < tr) ? (crc_compute(crcBits, buf.data(), gb) ^ t.genInit) : 0u;
        buf[k >> 3] = 0;
    }
    return t;
}

} // namespace polar
