/* TEST INFRASTRUCTURE ONLY -- CRC detectors, quantiser and encoder of the CPU oracle. */
#include "oracle.h"
#include "oracle_common.h"
#include <cmath>

namespace oracle {

// crc8.cpp:18-42 : poly 0x07, init 0, MSB first, no reflection, no xor-out
static uint8_t crc8(const uint8_t* d, int n)
{
    uint8_t c = 0;
    for (int i = 0; i < n; ++i) {
        c ^= d[i];
        for (int j = 0; j < 8; ++j) c = (uint8_t)((c << 1) ^ ((c & 0x80) ? 0x07 : 0));
    }
    return c;
}

// crc16.cpp:21-24 -> CRC.h CRC_16_CCITTFALSE (CRC.h:1337-1343): poly 0x1021, init 0xFFFF
static uint16_t crc16(const uint8_t* d, int n)
{
    uint16_t c = 0xFFFF;
    for (int i = 0; i < n; ++i) {
        c ^= (uint16_t)(d[i] << 8);
        for (int j = 0; j < 8; ++j) c = (uint16_t)((c << 1) ^ ((c & 0x8000) ? 0x1021 : 0));
    }
    return c;
}

// crc32.cpp:28-36 : x86 crc32 instruction = CRC-32C (Castagnoli) reflected, init 0, no xor-out,
// over consecutive little-endian u32 words == byte-wise over the same bytes.
static uint32_t crc32c_words(const uint8_t* d, int words)
{
    uint32_t c = 0;
    for (int i = 0; i < words * 4; ++i) {
        c ^= d[i];
        for (int j = 0; j < 8; ++j) c = (c >> 1) ^ ((c & 1) ? 0x82F63B78u : 0);
    }
    return c;
}

void crc_generate(int bits, uint8_t* data, int bytes)
{
    if (bits == 8) {
        data[bytes - 1] = crc8(data, bytes - 1); // crc8.cpp:51-56
    } else if (bits == 16) {
        uint16_t c = crc16(data, bytes - 2); // crc16.cpp:35-43 : big-endian trailer
        data[bytes - 2] = (uint8_t)(c >> 8);
        data[bytes - 1] = (uint8_t)(c & 0xFF);
    } else if (bits == 32) {
        int w = bytes / 4 - 1; // crc32.cpp:55-66 : native (little-endian) u32 trailer
        uint32_t c = crc32c_words(data, w);
        for (int i = 0; i < 4; ++i) data[4 * w + i] = (uint8_t)(c >> (8 * i));
    }
}

int crc_check(int bits, const uint8_t* data, int bytes)
{
    if (bits == 8) return crc8(data, bytes - 1) == data[bytes - 1];
    if (bits == 16) {
        uint16_t rx = (uint16_t)((data[bytes - 2] << 8) | data[bytes - 1]);
        return crc16(data, bytes - 2) == rx;
    }
    if (bits == 32) {
        int w = (bytes >> 2) - 1;
        uint32_t rx = 0;
        for (int i = 0; i < 4; ++i) rx |= (uint32_t)data[4 * w + i] << (8 * i);
        return crc32c_words(data, w) == rx;
    }
    return 1; // Dummy detector: dummy.cpp:20-33
}

} // namespace oracle

using namespace oracle;

extern "C" void oracle_crc_generate(int bits, uint8_t* data, int bytes) { crc_generate(bits, data, bytes); }
extern "C" int oracle_crc_check(int bits, const uint8_t* data, int bytes) { return crc_check(bits, data, bytes); }

extern "C" int8_t oracle_f_i8(int8_t a, int8_t b) { return f_i8(a, b); }
extern "C" int8_t oracle_g_i8(int8_t a, int8_t b, int8_t bit) { return g_i8(a, b, bit); }
extern "C" int8_t oracle_rep_sum_i8(const int8_t* x, int n) { return rep_sum_i8(x, n); }

// bitcontainer.cpp:468-516.  N >= 8: round-to-nearest-even then saturate (NaN -> -128 via the
// INT_MIN "integer indefinite" of cvtps_epi32 for N >= 32); N < 8: clamp then round() half away.
extern "C" void oracle_quantize(const float* llr, int8_t* out, int N)
{
    for (int i = 0; i < N; ++i) {
        float x = llr[i];
        if (N >= 8) {
            // N >= 32: cvtps_epi32 first (NaN and |x| >= 2^31 give INT_MIN, the "integer indefinite"),
            // then signed saturation -> -128.  8 <= N < 32: clamp first (max_ps/min_ps send NaN to -128).
            if (std::isnan(x)) { out[i] = -128; continue; }
            if (N >= 32 && x >= 2147483648.0f) { out[i] = -128; continue; }
            float c = x < -128.f ? -128.f : (x > 127.f ? 127.f : x);
            out[i] = (int8_t)std::nearbyintf(c); // default rounding mode = nearest-even
        } else {
            float c = std::fmin(std::fmax(x, -128.0f), 127.0f);
            out[i] = (int8_t)std::round(c);
        }
    }
}

// butterfly_fip_packed.cpp:45-59 + bitcontainer.cpp:886-898 (scatter) + :1004-1010 (reset frozen)
extern "C" int oracle_encode(int N, const uint32_t* frozen, int nFrozen, int systematic, int crcBits,
                             uint8_t* info, uint8_t* code, long nFrames)
{
    std::vector<uint8_t> fm = frozen_mask(N, frozen, nFrozen);
    int K = N - nFrozen;
    int kb = (K + 7) / 8, nb = (N + 7) / 8;
    std::vector<uint8_t> u(N);
    for (long f = 0; f < nFrames; ++f) {
        uint8_t* in = info + f * kb;
        if (crcBits) crc_generate(crcBits, in, K / 8); // (N - |frozen|) / 8 truncates
        int k = 0;
        for (int i = 0; i < N; ++i) u[i] = fm[i] ? 0 : (uint8_t)get_bit(in, k++);
        polar_transform(u.data(), N);
        if (systematic) {
            for (int i = 0; i < N; ++i)
                if (fm[i]) u[i] = 0;
            polar_transform(u.data(), N);
        }
        uint8_t* out = code + f * nb;
        memset(out, 0, nb);
        for (int i = 0; i < N; ++i) set_bit(out, i, u[i]);
    }
    return 0;
}

// ---------------------------------------------------------------- rate matching (TEST INFRASTRUCTURE, like the rest)
// Puncturer (puncturer.cpp:45-61): parent length = next power of two; the first (parent - blockLength) frozen
// positions are punctured, every other position of the parent code word is kept, in ascending order.
// Returns the number of kept positions written to `pos` (= blockLength) or -1 (std::out_of_range upstream).
extern "C" int oracle_puncturer_positions(int blockLength, const uint32_t* frozen, int nFrozen, uint32_t* pos)
{
    uint32_t parent = (uint32_t)blockLength - 1; // round_up_power_of_two, puncturer.cpp:18-28
    parent |= parent >> 1; parent |= parent >> 2; parent |= parent >> 4; parent |= parent >> 8; parent |= parent >> 16;
    ++parent;
    const int nPunct = (int)parent - blockLength;
    if (nPunct > nFrozen) return -1;
    int n = 0, fi = 0;
    // inverse_set_difference (puncturer.cpp:30-43): 0..parent-1 minus the SORTED prefix of the frozen set
    for (uint32_t i = 0; i < parent; ++i) {
        while (fi < nPunct && frozen[fi] < i) ++fi;
        if (fi < nPunct && frozen[fi] == i) continue;
        pos[n++] = i;
    }
    return n;
}

// puncture<T> (puncturer.h:71-79): out[j] = in[pos[j]]; elem = bytes per element
extern "C" void oracle_puncture(const uint32_t* pos, int nPos, int elem, const uint8_t* in, uint8_t* out)
{
    for (int j = 0; j < nPos; ++j) std::memcpy(out + (size_t)j * elem, in + (size_t)pos[j] * elem, elem);
}

// depuncture<T> (puncturer.h:93-100): zero fill the parent block, then out[pos[j]] = in[j]
extern "C" void oracle_depuncture(const uint32_t* pos, int nPos, int parent, int elem, const uint8_t* in, uint8_t* out)
{
    std::memset(out, 0, (size_t)parent * elem);
    for (int j = 0; j < nPos; ++j) std::memcpy(out + (size_t)pos[j] * elem, in + (size_t)j * elem, elem);
}

// puncturePacked (puncturer.cpp:65-85): MSB-first bit j of the output = MSB-first bit pos[j] of the input
extern "C" void oracle_puncture_packed(const uint32_t* pos, int nPos, const uint8_t* in, uint8_t* out)
{
    for (int b = 0; b < nPos / 8; ++b) {
        uint8_t v = 0;
        for (int i = 0; i < 8; ++i) {
            const uint32_t p = pos[8 * b + i];
            v |= (uint8_t)(((in[p / 8] >> (7 - p % 8)) & 1) << (7 - i));
        }
        out[b] = v;
    }
}
