// Packed int8x4 primitives for the sm_100a polar decoders.
//
// Four int8 LLRs live in one 32-bit word (element i of a stage block = byte i%4 of word i/4).
// sm_100a has no byte-SIMD ALU except VABSDIFF4.U8, but it does have 16x2 SIMD min/max/add
// (VIMNMX.{S,U}16x2, VIMNMX3, VIADD.16x2, VIADDMNMX.S16x2 -- the DPX family), so the F/G
// butterflies below are written against exactly those instructions.  The semantics are the
// reference's int8 kernels, bit for bit:
//   f: include/polarcode/decoding/fip_char.h:44-65   (min-sum, |.| clamped to [1,127], 0 is "+")
//   g: include/polarcode/decoding/fip_char.h:67-73   (saturating b+a / b-a selected by bit sign)
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace polar {

#define POLAR_DEV __device__ __forceinline__

static constexpr uint32_t FULL_MASK = 0xffffffffu;

// prmt.b32 in default mode: selector nibble bit 3 replicates the sign of the selected byte.
POLAR_DEV uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel)
{
    uint32_t r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
    return r;
}

// bytes 0,2 (resp. 1,3) of x sign-extended into the two 16-bit halves
POLAR_DEV uint32_t sext_even(uint32_t x) { return prmt(x, 0u, 0xA280u); }
POLAR_DEV uint32_t sext_odd(uint32_t x) { return prmt(x, 0u, 0xB391u); }
// low bytes of the four halves of (even, odd) back into one word
POLAR_DEV uint32_t pack_even_odd(uint32_t e, uint32_t o) { return prmt(e, o, 0x6240u); }

// |x| per byte as unsigned (|-128| = 128): one LOP3 + one VABSDIFF4.U8
POLAR_DEV uint32_t uabs4(uint32_t x) { return __vabsdiffu4(x ^ 0x80808080u, 0x80808080u); }

// 4 sign bits of the bytes of x -> low nibble (bit k = sign of byte k)
POLAR_DEV uint32_t sign_nibble(uint32_t x) { return ((x & 0x80808080u) * 0x00204081u) >> 28; }

// low nibble (bit k) -> 0xFF in byte k
POLAR_DEV uint32_t nibble_to_bytemask(uint32_t nib)
{
    return (((nib & 0xFu) * 0x00204081u) & 0x01010101u) * 0xFFu;
}

// low nibble (bit k) -> most significant bit of byte k (all g4 needs: its PRMTs replicate that bit)
POLAR_DEV uint32_t nibble_to_msb(uint32_t nib) { return (nib & 0xFu) * 0x10204081u; }

// min-sum f on four int8 pairs.  16 ALU ops per 4 elements.
POLAR_DEV uint32_t f4(uint32_t a, uint32_t b)
{
    const uint32_t ma = uabs4(a), mb = uabs4(b); // 0..128
    // even bytes: clamp(min(ma,mb),1,127) in u16x2
    uint32_t e = __vimin3_u16x2(ma & 0x00FF00FFu, mb & 0x00FF00FFu, 0x007F007Fu);
    e = __vmaxu2(e, 0x00010001u);
    // odd bytes: the 16-bit unsigned order is decided by the high byte first, low-byte garbage is harmless
    uint32_t o = __vimin3_u16x2(ma, mb, 0x7F007F00u);
    o = __vmaxu2(o, 0x01000100u);
    const uint32_t m = (o & 0xFF00FF00u) | e; // e already has clean high bytes
    // apply sign(a)*sign(b): negate the bytes whose signs differ.  m in [1,127] so (m^0xFF)+1 never carries.
    const uint32_t sh = ((a ^ b) & 0x80808080u) >> 7;
    return (m ^ (sh * 0xFFu)) + sh;
}

// saturating a + b on four int8 pairs (g with an all-zero left block: fip_char.h:112-131; REP accumulate)
POLAR_DEV uint32_t adds4(uint32_t a, uint32_t b)
{
    uint32_t re = __vmaxs2(__viaddmin_s16x2(sext_even(a), sext_even(b), 0x007F007Fu), 0xFF80FF80u);
    uint32_t ro = __vmaxs2(__viaddmin_s16x2(sext_odd(a), sext_odd(b), 0x007F007Fu), 0xFF80FF80u);
    return pack_even_odd(re, ro);
}

// g on four int8 pairs: bit ? sat(b - a) : sat(b + a); `neg` carries the left bit of element k in the most
// significant bit of byte k (nibble_to_msb; the other bits are ignored).
// sat(b - a) = ~sat(~b + a) holds exactly on int8 (~x = -x - 1, and both saturation ends map onto each other), so the
// conditional subtraction is one saturating add between two XORs with the byte mask of the left bits.
POLAR_DEV uint32_t g4(uint32_t a, uint32_t b, uint32_t neg)
{
    const uint32_t m = prmt(neg, 0u, 0xBA98u); // 0xFF in byte k where the left bit is set (sign-replicating selectors)
    return adds4(b ^ m, a) ^ m;
}

// exact sums of min(x,0) / max(x,0) over the four bytes (IDP.4A)
POLAR_DEV int sum_neg4(uint32_t x, int acc)
{
    const uint32_t m = nibble_to_bytemask(sign_nibble(x));
    return __dp4a((int)(x & m), 0x01010101, acc);
}
POLAR_DEV int sum_pos4(uint32_t x, int acc)
{
    const uint32_t m = nibble_to_bytemask(sign_nibble(x));
    return __dp4a((int)(x & ~m), 0x01010101, acc);
}

// |max(x,-127)| of byte k of w (scl_fip_char.cpp:449,618)
POLAR_DEV int abs127_byte(uint32_t w, int k)
{
    int v = (int)(int8_t)(w >> (8 * k));
    v = max(v, -127);
    return abs(v);
}

// 4-byte asynchronous global -> shared copy (LDGSTS): no register staging, any number in flight
POLAR_DEV void cp_async4(uint32_t* smemDst, const uint32_t* globalSrc)
{
    const unsigned sa = (unsigned)__cvta_generic_to_shared(smemDst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sa), "l"(globalSrc) : "memory");
}
// 16-byte variant (L2 only: streaming data; both addresses 16-byte aligned)
POLAR_DEV void cp_async16(uint32_t* smemDst, const uint32_t* globalSrc)
{
    const unsigned sa = (unsigned)__cvta_generic_to_shared(smemDst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(globalSrc) : "memory");
}
POLAR_DEV void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// ---- L2 prefetch of a contiguous byte range by one warp (the next frame group's channel rows: the top-level steps of the
// group then find them in L2 instead of waiting for HBM)
POLAR_DEV void warp_prefetch_l2(const void* base, long long bytes, int lane)
{
    const char* p = reinterpret_cast<const char*>(reinterpret_cast<uintptr_t>(base) & ~(uintptr_t)127);
    const long long span = bytes + (long long)(reinterpret_cast<uintptr_t>(base) & 127u);
    for (long long off = (long long)lane * 128; off < span; off += 32 * 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(p + off));
}

// ---- TMA bulk copies (cp.async.bulk, SASS UBLKCP) completing on a shared-memory mbarrier
POLAR_DEV void mbar_init(uint64_t* bar, int count)
{
    const unsigned a = (unsigned)__cvta_generic_to_shared(bar);
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(a), "r"(count) : "memory");
}
POLAR_DEV void mbar_expect_tx(uint64_t* bar, unsigned bytes) // one arrival + the byte count the copies will complete
{
    const unsigned a = (unsigned)__cvta_generic_to_shared(bar);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(a), "r"(bytes) : "memory");
}
POLAR_DEV void mbar_wait(uint64_t* bar, unsigned parity)
{
    const unsigned a = (unsigned)__cvta_generic_to_shared(bar);
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(a), "r"(parity) : "memory");
}
// bytes: multiple of 16; both addresses 16-byte aligned
POLAR_DEV void tma_bulk_g2s(void* smemDst, const void* globalSrc, unsigned bytes, uint64_t* bar)
{
    const unsigned d = (unsigned)__cvta_generic_to_shared(smemDst), b = (unsigned)__cvta_generic_to_shared(bar);
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(d), "l"(globalSrc),
                 "r"(bytes), "r"(b) : "memory");
}
// order earlier generic-proxy accesses to shared memory before the async-proxy writes that follow
POLAR_DEV void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

POLAR_DEV int sat8i(int v) { return min(max(v, -128), 127); }

} // namespace polar
