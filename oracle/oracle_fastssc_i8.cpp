/* TEST INFRASTRUCTURE ONLY -- scalar restatement of FastSscFipChar
 * (reference: src/polarcode/decoding/fastssc_fip_char.cpp, include/polarcode/decoding/fip_char.h).
 * Soft bits are kept as int8 bytes exactly like the reference (sign bit = bit value).
 */
#include "oracle.h"
#include "oracle_common.h"

namespace oracle {

static const int VEC = 32; // BYTESPERVECTOR of the AVX2 build

struct Stats { long* counts; long* sizes; };
static inline void tally(Stats* s, int id, long size)
{
    if (s && s->counts) { s->counts[id]++; s->sizes[id] += size; }
}

// First position of the minimum of |x| read as unsigned byte (|-128| = 128):
// avxconvenience.cpp:13-56 (minpos_epu8) on pabsb output.
static inline void first_min_uabs(const int8_t* x, int n, int* pos, int* val)
{
    int best = 256, bp = 0;
    for (int i = 0; i < n; ++i) {
        int a = x[i] < 0 ? -(int)x[i] : x[i]; // -128 -> 128
        if (a < best) { best = a; bp = i; }
    }
    *pos = bp; *val = best;
}

// SPC on `n` values already placed in `bits` (copy of the LLRs): parity + flip.
// Long form fastssc_fip_char.cpp:274-303, short form :305-319.
static void spc_fix(int8_t* bits, int n, bool longForm)
{
    int parity = 0;
    for (int i = 0; i < n; ++i) parity ^= (bits[i] & 0x80);
    if (!parity) return;
    if (!longForm) {
        // pads are +127; minpos over the whole 32-byte vector, no threshold
        int pos, val;
        first_min_uabs(bits, n, &pos, &val);
        if (val > 127 && n < VEC) return;   // a +127 pad wins: the flip lands outside the block
        // (val == 127: the first 127 inside the block precedes the pads, pos is already right)
        bits[pos] = (int8_t)(-(int)bits[pos]);
    } else {
        int minIdx = 0, minAbs = 127;
        for (int v = 0; v < n / VEC; ++v) {
            if (minAbs > 0) {
                int pos, val;
                first_min_uabs(bits + v * VEC, VEC, &pos, &val);
                int8_t testAbs = (int8_t)val; // stored into a char (128 -> -128: undefined upstream)
                if (testAbs < minAbs) { minIdx = pos + v * VEC; minAbs = testAbs; }
            }
        }
        bits[minIdx] = (int8_t)(-(int)bits[minIdx]); // 0 stays 0, -128 stays -128
    }
}

static void decode_node(const int8_t* llr, int8_t* bits, const uint8_t* fm, int n, Stats* st)
{
    int fc = count_frozen(fm, n);
    // fastssc_fip_char.cpp:496-569 (createDecoder)
    if (fc == n) { // RateZero :202-208
        tally(st, ORACLE_NODE_R0, n);
        for (int i = 0; i < n; ++i) bits[i] = 127;
        return;
    }
    if (fc == 0) { // RateOne :210-215
        tally(st, ORACLE_NODE_R1, n);
        memcpy(bits, llr, n);
        return;
    }
    if (fc == n - 1) { // Repetition :225-240, :265-272
        tally(st, ORACLE_NODE_REP, n);
        int8_t s = rep_sum_i8(llr, n);
        for (int i = 0; i < n; ++i) bits[i] = s;
        return;
    }
    if (fc == 1) { // SPC
        tally(st, ORACLE_NODE_SPC, n);
        memcpy(bits, llr, n);
        spc_fix(bits, n, n > VEC);
        return;
    }
    if (fc == n - 2 && n >= VEC) { // DoubleRepetition :249-263 + half_reduce_adds_epi8
        tally(st, ORACLE_NODE_REP2, n);
        int8_t acc[32], t[32];
        rep_lane_acc(llr, n, acc);
        for (int d = 16; d >= 2; d >>= 1) {
            for (int j = 0; j < 32; ++j) t[j] = sat8((int)acc[j] + (int)acc[j ^ d]);
            memcpy(acc, t, 32);
        }
        for (int i = 0; i < n; ++i) bits[i] = acc[i % 32];
        return;
    }
    int h = n / 2;
    int lf = count_frozen(fm, h), rf = count_frozen(fm + h, h);
    if (lf == h && rf == 0 && n <= VEC) { // ShortZeroOne :390-399
        tally(st, ORACLE_NODE_ZERO_ONE, n);
        for (int i = 0; i < h; ++i) bits[i] = bits[i + h] = sat8((int)llr[i] + (int)llr[i + h]);
        return;
    }
    if (lf == h && rf == 1) { // ZeroSpc :321-359 / ShortZeroSpc :361-388
        tally(st, ORACLE_NODE_ZERO_SPC, n);
        std::vector<int8_t> y(h);
        for (int i = 0; i < h; ++i) y[i] = sat8((int)llr[i] + (int)llr[i + h]);
        if (n <= VEC) {
            // pads 127 over a full 32-byte vector, unconditional minpos when parity is odd
            int parity = 0;
            for (int i = 0; i < h; ++i) parity ^= (y[i] & 0x80);
            if (parity) {
                int pos, val;
                first_min_uabs(y.data(), h, &pos, &val);
                if (!(val > 127)) y[pos] = (int8_t)(-(int)y[pos]); // else a pad is hit (h < 32 always)
            }
        } else {
            // long form: same threshold search as SpcDecoder, over h >= 32 values (also for h == 32)
            spc_fix(y.data(), h, true);
        }
        for (int i = 0; i < h; ++i) bits[i] = bits[i + h] = y[i];
        return;
    }
    // inner nodes
    std::vector<int8_t> child(h);
    if (rf == 0) { // ROneNode :427-449 / ShortROneNode :451-474
        tally(st, ORACLE_NODE_F, h);
        for (int i = 0; i < h; ++i) child[i] = f_i8(llr[i], llr[i + h]);
        decode_node(child.data(), bits, fm, h, st);
        tally(st, ORACLE_NODE_G_R1, h);
        for (int i = 0; i < h; ++i) {
            int8_t r = g_i8(llr[i], llr[i + h], bits[i]);
            bits[i] = (int8_t)(bits[i] ^ r);
            bits[i + h] = r;
        }
        return;
    }
    if (lf == h) { // ZeroRNode :476-492
        tally(st, ORACLE_NODE_G_0R, h);
        for (int i = 0; i < h; ++i) child[i] = sat8((int)llr[i] + (int)llr[i + h]);
        decode_node(child.data(), bits + h, fm + h, h, st);
        memcpy(bits, bits + h, h);
        return;
    }
    // RateRNode :401-425
    tally(st, ORACLE_NODE_F, h);
    for (int i = 0; i < h; ++i) child[i] = f_i8(llr[i], llr[i + h]);
    decode_node(child.data(), bits, fm, h, st);
    tally(st, ORACLE_NODE_G, h);
    for (int i = 0; i < h; ++i) child[i] = g_i8(llr[i], llr[i + h], bits[i]);
    decode_node(child.data(), bits + h, fm + h, h, st);
    tally(st, ORACLE_NODE_COMBINE, h);
    if (n <= VEC) { // ShortRateRNode -> CombineBitsShort (fip_char.h:185-205): both halves are clamped to -127 first.  Only
                    // the byte values change (soft outputs), never a sign.
        for (int i = 0; i < h; ++i) {
            const int8_t l = bits[i] < -127 ? -127 : bits[i], r = bits[i + h] < -127 ? -127 : bits[i + h];
            bits[i] = (int8_t)(l ^ r);
            bits[i + h] = r;
        }
    } else { // RateRNode -> CombineInPlace (fip_char.h:164-171): raw byte XOR, right half untouched
        for (int i = 0; i < h; ++i) bits[i] = (int8_t)(bits[i] ^ bits[i + h]);
    }
}

int fastssc_i8_decode(int N, const uint8_t* fm, int systematic, int crcBits, const int8_t* llr,
                      uint8_t* out)
{
    std::vector<int8_t> bits(N);
    decode_node(llr, bits.data(), fm, N, nullptr);
    std::vector<uint8_t> hard(N);
    for (int i = 0; i < N; ++i) hard[i] = (bits[i] & 0x80) ? 1 : 0;
    extract_info(hard.data(), fm, N, systematic, out);
    int K = N - count_frozen(fm, N);
    return crc_check(crcBits, out, (K + 7) / 8); // fastssc_fip_char.cpp:627-628
}

// Decoder::getSoftCodeword after FastSscFipChar::decode (decoder.cpp:147, bitcontainer.cpp:566-569): the N byte "soft bits"
// the root node left in mBitContainer -- R0: 127, R1: the LLR, REP: the saturating sum, SPC: the LLR with one byte negated,
// inner nodes: raw byte XOR of the halves (sign = bit value, lower bits carry no meaning but are reproducible).
void fastssc_i8_soft(int N, const uint8_t* fm, const int8_t* llr, int8_t* soft)
{
    decode_node(llr, soft, fm, N, nullptr);
}

void fastssc_i8_stats(int N, const uint8_t* fm, long* counts, long* sizes)
{
    std::vector<int8_t> llr(N, 1), bits(N);
    Stats st{ counts, sizes };
    decode_node(llr.data(), bits.data(), fm, N, &st);
}

} // namespace oracle
