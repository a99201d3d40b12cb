/* TEST INFRASTRUCTURE ONLY -- shared scalar helpers of the CPU oracle (see oracle.h). */
#ifndef POLAR_ORACLE_COMMON_H
#define POLAR_ORACLE_COMMON_H
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace oracle {

static inline int8_t sat8(int v) { return (int8_t)(v > 127 ? 127 : (v < -128 ? -128 : v)); }

// min-sum f on int8: fip_char.h:44-65.  Output never 0 and never -128.
static inline int8_t f_i8(int8_t a, int8_t b)
{
    int8_t signProbe = (int8_t)((a ^ b) | 1);
    int aa = a < -127 ? -127 : a, bb = b < -127 ? -127 : b;
    aa = aa < 0 ? -aa : aa;
    bb = bb < 0 ? -bb : bb;
    if (aa < 1) aa = 1;
    if (bb < 1) bb = 1;
    int m = aa < bb ? aa : bb;
    return (int8_t)(signProbe < 0 ? -m : m);
}

// g on int8: fip_char.h:67-73 (a = left/first-half llr, b = right, bit = soft left bit)
static inline int8_t g_i8(int8_t a, int8_t b, int8_t bit)
{
    return (bit & 0x80) ? sat8((int)b - (int)a) : sat8((int)b + (int)a);
}

// |max(x,-127)| : scl_fip_char.cpp:449,618
static inline int abs127(int8_t x)
{
    int v = x < -127 ? -127 : x;
    return v < 0 ? -v : v;
}

// Stepwise-saturating sum of n int8 in the reference's 32-lane order:
// fastssc_fip_char.cpp:225-240 + avxconvenience.h:92-101 (reduce_adds_epi8).
static inline void rep_lane_acc(const int8_t* x, int n, int8_t acc[32])
{
    for (int j = 0; j < 32; ++j) acc[j] = 0;
    int vecs = (n + 31) / 32;
    for (int v = 0; v < vecs; ++v)
        for (int j = 0; j < 32; ++j) {
            int idx = v * 32 + j;
            acc[j] = sat8((int)acc[j] + (idx < n ? (int)x[idx] : 0));
        }
}
static inline int8_t rep_sum_i8(const int8_t* x, int n)
{
    int8_t acc[32];
    rep_lane_acc(x, n, acc);
    for (int d = 16; d >= 1; d >>= 1)
        for (int j = 0; j < d; ++j) acc[j] = sat8((int)acc[j] + (int)acc[j + d]);
    return acc[0];
}

// Sorted frozen index list -> per-position flag
static inline std::vector<uint8_t> frozen_mask(int N, const uint32_t* frozen, int nFrozen)
{
    std::vector<uint8_t> m(N, 0);
    for (int i = 0; i < nFrozen; ++i)
        if ((int)frozen[i] < N) m[frozen[i]] = 1;
    return m;
}

static inline int count_frozen(const uint8_t* fm, int n)
{
    int c = 0;
    for (int i = 0; i < n; ++i) c += fm[i];
    return c;
}

// MSB-first bit helpers (bitcontainer.cpp:520-534)
static inline int get_bit(const uint8_t* p, long i) { return (p[i >> 3] >> (7 - (i & 7))) & 1; }
static inline void set_bit(uint8_t* p, long i, int v)
{
    uint8_t m = (uint8_t)(0x80 >> (i & 7));
    if (v) p[i >> 3] |= m; else p[i >> 3] &= (uint8_t)~m;
}

// x <- x * F^{(x)n} over GF(2), natural order: butterfly_fip.cpp:31-63
static inline void polar_transform(uint8_t* u, int N)
{
    for (int s = 1; s < N; s <<= 1)
        for (int i = 0; i < N; ++i)
            if (!(i & s)) u[i] ^= u[i + s];
}

// Gather hard decisions at non-frozen positions, MSB first (bitcontainer.cpp:536-564).
// systematic: bits are the codeword estimate; else re-encode first (fastssc_fip_char.cpp:619-622).
static inline void extract_info(const uint8_t* hard, const uint8_t* fm, int N, int systematic,
                                uint8_t* out)
{
    std::vector<uint8_t> u(hard, hard + N);
    if (!systematic) polar_transform(u.data(), N);
    int K = N - count_frozen(fm, N);
    memset(out, 0, (K + 7) / 8);
    int k = 0;
    for (int i = 0; i < N; ++i)
        if (!fm[i]) { set_bit(out, k, u[i]); ++k; }
}

void crc_generate(int bits, uint8_t* data, int bytes);
int crc_check(int bits, const uint8_t* data, int bytes);

} // namespace oracle
#endif
