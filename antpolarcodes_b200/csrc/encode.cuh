// Bit-packed polar encoder for sm_100a (replaces Encoding::ButterflyFipPacked::encode,
// src/polarcode/encoding/butterfly_fip_packed.cpp:45-70).
//
// One frame = N/32 words = one word per lane (N <= 1024: 32/(N/32) frames per warp; N > 1024: several
// words per lane).  Per frame: CRC generate in place (Detector::generate, via the linear table of
// crc_host.h) -> scatter the K info bits to the non-frozen positions (bitcontainer.cpp:886-898) ->
// F^{(x)n}: stages 0-4 are shift/mask XORs inside the word, stages >= 5 are __shfl_xor_sync exchanges
// (butterfly_fip.cpp:31-63) -> systematic: mask the frozen positions and transform again (:52-57) ->
// MSB-first bytes with coalesced 4-byte stores.
#pragma once
#include "primitives.cuh"
#include <algorithm>
#include <cstring>

namespace polar {

struct EncodeParams {
    uint8_t* info;          // nFrames x kb, CRC trailer written in place
    uint8_t* code;          // nFrames x nb
    const uint2* wordTab;   // per code word: x = mask of info positions, y = number of info bits before it
    const uint32_t* genTab; // per info bit: CRC contribution
    const uint32_t* genNib; // encode_reg_kernel: per info nibble (4 stream bits, MSB first) and value: 16 contributions
    const uint32_t* runTab; // [0..W]: first run of each code word; then the runs: b | len << 5 | first info bit << 11
    int runBase;            // offset of the run entries inside runTab
    uint32_t genInit;
    long long nFrames;
    int N, lgN, K, kb, nb, crcBits, systematic;
    int vecIn;              // info rows are 16-byte aligned (encode_reg_kernel: 128-bit loads)
    int byteOut;            // codeword rows are not 4-byte aligned (encode_kernel: byte stores)
};

POLAR_DEV uint32_t inword_transform(uint32_t v)
{
    v ^= (v >> 1) & 0x55555555u;
    v ^= (v >> 2) & 0x33333333u;
    v ^= (v >> 4) & 0x0F0F0F0Fu;
    v ^= (v >> 8) & 0x00FF00FFu;
    v ^= (v >> 16) & 0x0000FFFFu;
    return v;
}

// WPL = words per lane (1 for N <= 1024, N/1024 above)
template <int WPL>
__global__ void __launch_bounds__(256) encode_kernel(const __grid_constant__ EncodeParams p)
{
    __shared__ __align__(16) uint8_t sinfo[8][512 + 8]; // + slack for the 5-byte window reads of the scatter
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, warpsPerBlock = blockDim.x >> 5;
    const int W = p.N >= 32 ? p.N >> 5 : 1;          // words per frame
    const int lpf = W < 32 ? W : 32;                  // lanes per frame
    const int lgLpf = 31 - __clz(lpf);
    const int fpw = 32 / lpf;                         // frames per warp
    const int fi = lane >> lgLpf, g = lane & (lpf - 1);
    uint8_t* myInfo = &sinfo[warp][fi * p.kb];
    const int trailerBit = p.crcBits == 8 ? 8 * (p.kb - 1) : p.crcBits == 16 ? 8 * (p.kb - 2)
                         : p.crcBits == 32 ? 32 * ((p.kb >> 2) - 1) : 8 * p.kb;

    const long long totalGroups = (long long)gridDim.x * warpsPerBlock;
    const long long nGroups = (p.nFrames + fpw - 1) / fpw;
    for (long long grp = (long long)blockIdx.x * warpsPerBlock + warp; grp < nGroups; grp += totalGroups) {
        const long long frame = grp * fpw + fi;
        const bool live = frame < p.nFrames;
        uint8_t* ginfo = p.info + frame * p.kb;
        // stage the info bytes of this lane group's frame
        if (live)
            for (int i = g; i < p.kb; i += lpf) myInfo[i] = ginfo[i];
        __syncwarp();
        if (p.crcBits) {
            uint32_t crc = 0;
            if (live)
                for (int k = g; k < trailerBit; k += lpf)
                    if ((myInfo[k >> 3] >> (7 - (k & 7))) & 1) crc ^= __ldg(p.genTab + k);
            for (int d = 1; d < lpf; d <<= 1) crc ^= __shfl_xor_sync(FULL_MASK, crc, d);
            crc ^= p.genInit;
            __syncwarp();
            if (live && g == 0) {
                const int t = trailerBit >> 3;
                if (p.crcBits == 8) {
                    myInfo[t] = (uint8_t)crc;
                } else if (p.crcBits == 16) { // big-endian trailer: crc16.cpp:35-43
                    myInfo[t] = (uint8_t)(crc >> 8);
                    myInfo[t + 1] = (uint8_t)crc;
                } else {                       // native little-endian u32: crc32.cpp:55-66
                    for (int i = 0; i < 4; ++i) myInfo[t + i] = (uint8_t)(crc >> (8 * i));
                }
                for (int i = 0; i < p.crcBits / 8; ++i) ginfo[t + i] = myInfo[t + i];
            }
            __syncwarp();
        }
        // scatter: deposit the next popc(mask) info bits into the mask positions of each word
        uint32_t word[WPL], mask[WPL];
#pragma unroll
        for (int j = 0; j < WPL; ++j) {
            const uint2 t = __ldg(p.wordTab + (j * 32 + g));
            mask[j] = t.x;
            uint32_t m = t.x, v = 0;
            int k = (int)t.y;
            while (m) { // one run of consecutive info positions at a time (a polar info set has few runs per word)
                const int b = __ffs(m) - 1;
                const uint32_t rest = ~(m >> b);
                const int len = rest ? __ffs(rest) - 1 : 32;
                // stream bits k .. k+len-1 (MSB first inside the bytes) -> bits b .. b+len-1 of the word
                const uint8_t* q = myInfo + (k >> 3);
                const uint32_t hi = ((uint32_t)q[0] << 24) | ((uint32_t)q[1] << 16) | ((uint32_t)q[2] << 8) | q[3];
                const uint32_t x = __funnelshift_l((uint32_t)q[4] << 24, hi, k & 7); // len + (k & 7) <= 39 bits needed
                const uint32_t field = __brev(x) & (len == 32 ? 0xFFFFFFFFu : (1u << len) - 1u);
                v |= field << b;
                m = len + b >= 32 ? 0u : m & (0xFFFFFFFFu << (b + len));
                k += len;
            }
            word[j] = live ? v : 0u;
        }
        for (int pass = 0; pass < (p.systematic ? 2 : 1); ++pass) {
#pragma unroll
            for (int j = 0; j < WPL; ++j) {
                if (pass) word[j] &= mask[j];
                word[j] = inword_transform(word[j]);
            }
            for (int d = 1; d < lpf; d <<= 1) {
#pragma unroll
                for (int j = 0; j < WPL; ++j) {
                    const uint32_t other = __shfl_xor_sync(FULL_MASK, word[j], d);
                    if (!(g & d)) word[j] ^= other;
                }
            }
#pragma unroll
            for (int dj = 1; dj < WPL; dj <<= 1)
#pragma unroll
                for (int j = 0; j < WPL; ++j)
                    if (!(j & dj)) word[j] ^= word[j + dj];
        }
        if (live) {
            uint8_t* gcode = p.code + frame * p.nb;
#pragma unroll
            for (int j = 0; j < WPL; ++j) {
                const uint32_t o = __byte_perm(__brev(word[j]), 0u, 0x0123u);
                if (p.N >= 32 && !p.byteOut) {
                    reinterpret_cast<uint32_t*>(gcode)[j * 32 + g] = o;
                } else if (p.N >= 32) {
                    for (int i = 0; i < 4; ++i) gcode[4 * (j * 32 + g) + i] = (uint8_t)(o >> (8 * i));
                } else {
                    for (int i = 0; i < p.nb; ++i) gcode[i] = (uint8_t)(o >> (8 * i));
                }
            }
        }
        __syncwarp();
    }
}

// ---------------------------------------------------------------- one frame per thread (N = 32 W <= 1024, kb % 4 == 0)
// The whole codeword lives in W registers of the thread: no shuffles, no per-frame synchronisation.  The info
// bytes of the block's 128 frames are staged (coalesced) as big-endian words in a padded [frame][word] tile, so
// that a run of info positions is one funnel shift; codewords leave through the same kind of tile.
template <int W>
__global__ void __launch_bounds__(128) encode_thread_kernel(const __grid_constant__ EncodeParams p)
{
    extern __shared__ __align__(16) uint32_t esm[];
    const int tid = threadIdx.x, kbw = p.kb >> 2, istride = kbw + 1 + ((kbw & 1) ? 1 : 0); // odd stride: conflict free
    constexpr int ostride = W + 1;
    uint32_t* sin = esm;                          // 128 x istride big-endian info words
    uint32_t* sout = esm + 128 * istride;         // 128 x ostride codeword words
    const uint32_t kbwRecip = 0xFFFFFFFFu / (uint32_t)kbw + 1u; // ceil(2^32 / kbw)
    const long long nBlocks = (p.nFrames + 127) >> 7;
    for (long long blk = blockIdx.x; blk < nBlocks; blk += gridDim.x) {
        const long long f0 = blk << 7;
        const int nHere = (int)min(128LL, p.nFrames - f0);
        uint32_t* ginfo = reinterpret_cast<uint32_t*>(p.info + f0 * p.kb);
        for (int idx = tid; idx < nHere * kbw; idx += 128) {
            const int f = kbw == 1 ? idx : (int)__umulhi((uint32_t)idx, kbwRecip), i = idx - f * kbw; // idx < 2^13: exact
            sin[f * istride + i] = __byte_perm(ginfo[idx], 0u, 0x0123u);
        }
        __syncthreads();
        const uint32_t* mine = sin + tid * istride;
        if (p.crcBits) { // Detector::generate: the checksum is linear in the info bits (crc_host.h); trailer bits weigh 0
            uint32_t crc = p.genInit;
            for (int i = 0; i < kbw; ++i) {
                const uint32_t w = mine[i];
#pragma unroll 8
                for (int b = 0; b < 32; ++b) crc ^= __ldg(p.genTab + 32 * i + b) & (0u - ((w >> (31 - b)) & 1u));
            }
            uint32_t last = mine[kbw - 1];
            if (p.crcBits == 8) last = (last & ~0xFFu) | (crc & 0xFFu);
            else if (p.crcBits == 16) last = (last & ~0xFFFFu) | (crc & 0xFFFFu);   // big-endian trailer: crc16.cpp:35-43
            else last = __byte_perm(crc, 0u, 0x0123u);                             // native little-endian u32: crc32.cpp:55-66
            sin[tid * istride + kbw - 1] = last;
            if (tid < nHere) ginfo[tid * kbw + kbw - 1] = __byte_perm(last, 0u, 0x0123u);
        }
        uint32_t word[W];
#pragma unroll
        for (int j = 0; j < W; ++j) {
            uint32_t v = 0;
            const int r1 = (int)__ldg(p.runTab + j + 1);
            for (int r = (int)__ldg(p.runTab + j); r < r1; ++r) { // uniform: every thread of the grid walks the same runs
                const uint32_t e = __ldg(p.runTab + p.runBase + r);
                const int b = e & 31u, len = (e >> 5) & 63u, k = e >> 11, wi = k >> 5;
                const uint32_t x = __funnelshift_l(mine[wi + 1], mine[wi], k & 31); // stream bit k on top (pad word is masked off)
                const uint32_t field = __brev(x) & (0xFFFFFFFFu >> (32 - len));
                v |= field << b;
            }
            word[j] = v;
        }
        for (int pass = 0; pass < (p.systematic ? 2 : 1); ++pass) {
#pragma unroll
            for (int j = 0; j < W; ++j) {
                if (pass) word[j] &= __ldg(p.wordTab + j).x;
                word[j] = inword_transform(word[j]);
            }
#pragma unroll
            for (int d = 1; d < W; d <<= 1)
#pragma unroll
                for (int j = 0; j < W; ++j)
                    if (!(j & d)) word[j] ^= word[j + d];
        }
#pragma unroll
        for (int j = 0; j < W; ++j) sout[tid * ostride + j] = __byte_perm(__brev(word[j]), 0u, 0x0123u);
        __syncthreads();
        uint32_t* gcode = reinterpret_cast<uint32_t*>(p.code + f0 * p.nb);
        for (int idx = tid; idx < nHere * W; idx += 128) {
            const int f = idx / W, j = idx - f * W;
            gcode[idx] = sout[f * ostride + j];
        }
        __syncthreads();
    }
}

// ---------------------------------------------------------------- one frame per thread, no shared staging of the codeword
// (N = 32 W <= 1024, kb % 4 == 0).  Second generation of the kernel above:
//   * words are kept BIT-REVERSED (code position i of a word = bit 31 - i), the order of the MSB-first bytes on both sides:
//     the info words are the big-endian loads themselves, a codeword word is one byte permutation away from memory order, the
//     in-word butterflies shift left (v ^= (v << s) & m);
//   * the scatter program sits in the constant bank (kernel parameter): run r of code word j is one funnel shift of two
//     neighbouring info words and one masked OR -- indices, shift and mask are uniform-datapath values, the thread spends
//     two shared loads, SHF.L.W and LOP3;
//   * a thread loads its own info row (128- or 256-bit loads, rows of one warp are contiguous) and stores its own codeword
//     (STG.256 where the codeword has eight words or more): no block-wide synchronisation at all.
static constexpr int ENC_MAX_RUNS = 512; // 16 runs per word x 32 words: every frozen set of N <= 1024 fits
static constexpr int ENC_FEW_RUNS = 96;  // the usual case (N=1024 K=512: 55 - 63 runs): a 1.7 KB kernel parameter instead of 8.4 KB
template <int CAP>                       // (parameters above 4 KB take a slower launch path: +6 us per launch measured)
struct EncodeProgramT {
    uint32_t maskRev[32];             // info positions of code word j, bit-reversed
    uint16_t first[34];               // runs of word j: first[j] .. first[j+1]-1
    uint4 run[CAP];                   // x = byte offset of the row of the HIGH source word (row = info word index + 1; row 0 and
};                                    // row kbw+1 are zero pads), y = funnel shift, z = mask of the destination field (reversed)
using EncodeProgram = EncodeProgramT<ENC_MAX_RUNS>;

POLAR_DEV uint32_t inword_transform_rev(uint32_t v)
{
    v ^= (v << 1) & 0xAAAAAAAAu;
    v ^= (v << 2) & 0xCCCCCCCCu;
    v ^= (v << 4) & 0xF0F0F0F0u;
    v ^= (v << 8) & 0xFF00FF00u;
    v ^= (v << 16) & 0xFFFF0000u;
    return v;
}

POLAR_DEV void stg256(uint32_t* dst, const uint32_t* v)
{
    asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(dst), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]),
                 "r"(v[5]), "r"(v[6]), "r"(v[7]) : "memory");
}

static constexpr int ENC_REG_THREADS = 128;
// ENC_FPT frames per thread: the run entry, its address arithmetic and the loop control serve all of them (measured: 2 frames
// 5775, 3 -> 5501, 4 -> 5629 Gbit/s at N=256); batches too small to fill the GPU with two frames per thread take one
template <int W, int ENC_FPT, int CAP>
__global__ void __launch_bounds__(ENC_REG_THREADS) encode_reg_kernel(const __grid_constant__ EncodeParams p, const __grid_constant__ EncodeProgramT<CAP> prog)
{
    constexpr int ENC_ROW_WORDS = ENC_REG_THREADS * ENC_FPT; // words between two rows of a thread
    extern __shared__ __align__(16) uint32_t esm[]; // row r of frame f of thread t at esm[(r * FPT + f) * 128 + t]: no conflicts
    const int tid = threadIdx.x, kbw = p.kb >> 2;
    uint32_t* mine = esm + tid;
#pragma unroll
    for (int f = 0; f < ENC_FPT; ++f) {
        mine[f * ENC_REG_THREADS] = 0u;
        mine[(kbw + 1) * ENC_ROW_WORDS + f * ENC_REG_THREADS] = 0u;
    }
    // CRC generate as a table walk: the checksum is linear in the info bits (crc_host.h), so it is the XOR of one entry per
    // info nibble.  16 consecutive words per nibble: the 32 lanes of a load hit at most 16 different banks, equal values
    // broadcast -- no conflicts.  kb * 128 bytes of shared memory, copied once per block.
    const uint32_t* nibTab = esm + (kbw + 2) * ENC_ROW_WORDS;
    if (p.crcBits) {
        uint32_t* t = esm + (kbw + 2) * ENC_ROW_WORDS;
        for (int i = tid; i < kbw * 128; i += ENC_REG_THREADS) t[i] = __ldg(p.genNib + i);
        __syncthreads();
    }
    const long long stride = (long long)gridDim.x * ENC_ROW_WORDS;
    for (long long frame0 = (long long)blockIdx.x * ENC_ROW_WORDS + tid; frame0 < p.nFrames; frame0 += stride) {
        bool live[ENC_FPT];
#pragma unroll
        for (int f = 0; f < ENC_FPT; ++f) {
            const long long frame = frame0 + f * ENC_REG_THREADS;
            live[f] = frame < p.nFrames;
            if (!live[f]) continue;
            uint32_t* row = mine + f * ENC_REG_THREADS;
            uint32_t* ginfo = reinterpret_cast<uint32_t*>(p.info + frame * p.kb);
            if (p.vecIn) {
                for (int i = 0; i < kbw; i += 4) {
                    const uint4 q = *reinterpret_cast<const uint4*>(ginfo + i);
                    row[(i + 1) * ENC_ROW_WORDS] = __byte_perm(q.x, 0u, 0x0123u);
                    row[(i + 2) * ENC_ROW_WORDS] = __byte_perm(q.y, 0u, 0x0123u);
                    row[(i + 3) * ENC_ROW_WORDS] = __byte_perm(q.z, 0u, 0x0123u);
                    row[(i + 4) * ENC_ROW_WORDS] = __byte_perm(q.w, 0u, 0x0123u);
                }
            } else {
                for (int i = 0; i < kbw; ++i) row[(i + 1) * ENC_ROW_WORDS] = __byte_perm(ginfo[i], 0u, 0x0123u);
            }
            if (p.crcBits) { // Detector::generate: the checksum is linear in the info bits (crc_host.h); trailer bits weigh 0
                uint32_t crc = p.genInit;
                for (int i = 0; i < kbw; ++i) {
                    const uint32_t w = row[(i + 1) * ENC_ROW_WORDS];
                    const char* t = reinterpret_cast<const char*>(nibTab + i * 128);
#pragma unroll
                    for (int n = 0; n < 8; ++n) { // nibble n = stream bits 4n .. 4n+3 of the word = bits 31-4n .. 28-4n
                        const uint32_t off = n == 7 ? (w << 2) & 0x3Cu : (w >> (26 - 4 * n)) & 0x3Cu;
                        crc ^= *reinterpret_cast<const uint32_t*>(t + n * 64 + off);
                    }
                }
                uint32_t last = row[kbw * ENC_ROW_WORDS];
                if (p.crcBits == 8) last = (last & ~0xFFu) | (crc & 0xFFu);
                else if (p.crcBits == 16) last = (last & ~0xFFFFu) | (crc & 0xFFFFu);   // big-endian trailer: crc16.cpp:35-43
                else last = __byte_perm(crc, 0u, 0x0123u);                             // native little-endian u32: crc32.cpp:55-66
                row[kbw * ENC_ROW_WORDS] = last;
                ginfo[kbw - 1] = __byte_perm(last, 0u, 0x0123u);
            }
        }
        uint32_t word[ENC_FPT][W];
#pragma unroll
        for (int j = 0; j < W; ++j) {
            uint32_t v[ENC_FPT];
#pragma unroll
            for (int f = 0; f < ENC_FPT; ++f) v[f] = 0u;
            const int r1 = prog.first[j + 1];
            for (int r = prog.first[j]; r < r1; ++r) { // uniform: every thread of the grid walks the same runs
                const uint4 e = prog.run[r];
                const uint32_t* hi = reinterpret_cast<const uint32_t*>(reinterpret_cast<const char*>(mine) + e.x);
#pragma unroll
                for (int f = 0; f < ENC_FPT; ++f)
                    v[f] |= __funnelshift_l(hi[ENC_ROW_WORDS + f * ENC_REG_THREADS], hi[f * ENC_REG_THREADS], e.y) & e.z;
            }
#pragma unroll
            for (int f = 0; f < ENC_FPT; ++f) word[f][j] = v[f];
        }
#pragma unroll
        for (int f = 0; f < ENC_FPT; ++f) {
            for (int pass = 0; pass < (p.systematic ? 2 : 1); ++pass) {
#pragma unroll
                for (int j = 0; j < W; ++j) {
                    if (pass) word[f][j] &= prog.maskRev[j];
                    word[f][j] = inword_transform_rev(word[f][j]);
                }
#pragma unroll
                for (int d = 1; d < W; d <<= 1)
#pragma unroll
                    for (int j = 0; j < W; ++j)
                        if (!(j & d)) word[f][j] ^= word[f][j + d];
            }
            if (!live[f]) continue;
#pragma unroll
            for (int j = 0; j < W; ++j) word[f][j] = __byte_perm(word[f][j], 0u, 0x0123u);
            uint32_t* gcode = reinterpret_cast<uint32_t*>(p.code + (frame0 + f * ENC_REG_THREADS) * p.nb);
            if constexpr (W >= 8) {
#pragma unroll
                for (int j = 0; j < W; j += 8) stg256(gcode + j, word[f] + j);
            } else if constexpr (W == 4) {
                *reinterpret_cast<uint4*>(gcode) = make_uint4(word[f][0], word[f][1], word[f][2], word[f][3]);
            } else if constexpr (W == 2) {
                *reinterpret_cast<uint2*>(gcode) = make_uint2(word[f][0], word[f][1]);
            } else {
                gcode[0] = word[f][0];
            }
        }
    }
}

template <int W, int FPT, int CAP>
inline int launch_encode_reg_cap(const EncodeParams& p, const EncodeProgramT<CAP>& prog, int numSMs, cudaStream_t st)
{
    constexpr int rowWords = ENC_REG_THREADS * FPT;
    const int kbw = p.kb >> 2;
    const size_t smem = (size_t)rowWords * (kbw + 2) * 4 + (p.crcBits ? (size_t)kbw * 512 : 0);
    // the limit belongs to the kernel function, not to the handle: always the same value (the largest footprint of this
    // instantiation: kb = 128 with a CRC), so handles and host threads cannot lower it for each other
    constexpr int maxSmem = rowWords * (32 + 2) * 4 + 32 * 512;
    if (smem > 48 * 1024 &&
        cudaFuncSetAttribute((const void*)encode_reg_kernel<W, FPT, CAP>, cudaFuncAttributeMaxDynamicSharedMemorySize, maxSmem) != cudaSuccess)
        return -1;
    const long long nBlocks = (p.nFrames + rowWords - 1) / rowWords;
    const int grid = (int)std::min<long long>(nBlocks, (long long)numSMs * 16);
    encode_reg_kernel<W, FPT, CAP><<<grid, ENC_REG_THREADS, smem, st>>>(p, prog);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

template <int W, int FPT>
inline int launch_encode_reg_fpt(const EncodeParams& p, const EncodeProgram& prog, int numSMs, cudaStream_t st)
{
    if (prog.first[33] <= ENC_FEW_RUNS) { // first[33] = number of runs
        EncodeProgramT<ENC_FEW_RUNS> few;
        memcpy(few.maskRev, prog.maskRev, sizeof(few.maskRev));
        memcpy(few.first, prog.first, sizeof(few.first));
        memcpy(few.run, prog.run, sizeof(few.run));
        return launch_encode_reg_cap<W, FPT, ENC_FEW_RUNS>(p, few, numSMs, st);
    }
    return launch_encode_reg_cap<W, FPT, ENC_MAX_RUNS>(p, prog, numSMs, st);
}

// prog[0]: row offsets for one frame per thread, prog[1]: for two
template <int W>
inline int launch_encode_reg(const EncodeParams& p, const EncodeProgram* prog, int numSMs, cudaStream_t st)
{
    const bool two = p.nFrames >= (long long)numSMs * 8 * ENC_REG_THREADS * 2; // eight blocks per SM even at two frames per thread
    return two ? launch_encode_reg_fpt<W, 2>(p, prog[1], numSMs, st) : launch_encode_reg_fpt<W, 1>(p, prog[0], numSMs, st);
}

template <int W>
inline int launch_encode_thread(const EncodeParams& p, int numSMs, cudaStream_t st)
{
    const int kbw = p.kb >> 2, istride = kbw + 1 + ((kbw & 1) ? 1 : 0);
    const size_t smem = (size_t)128 * (istride + W + 1) * 4;
    if (smem > 48 * 1024) {
        if (cudaFuncSetAttribute((const void*)encode_thread_kernel<W>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
            return -1;
    }
    const long long nBlocks = (p.nFrames + 127) >> 7;
    const int grid = (int)std::min<long long>(nBlocks, (long long)numSMs * 16);
    encode_thread_kernel<W><<<grid, 128, smem, st>>>(p);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

inline int launch_encode(const EncodeParams& p0, const EncodeProgram* prog /* [2] or null */, int numSMs, cudaStream_t st)
{
    EncodeParams p = p0;
    const int W = p.N >= 32 ? p.N >> 5 : 1;
    // second-generation kernel: needs the scatter program and codeword rows aligned for its widest store
    if (prog && p.N >= 32 && p.N <= 1024 && (p.kb & 3) == 0 && p.kb >= 4 && ((uintptr_t)p.code & (size_t)(std::min(p.nb, 32) - 1)) == 0 &&
        ((uintptr_t)p.info & 3u) == 0) {
        p.vecIn = ((p.kb & 15) == 0 && ((uintptr_t)p.info & 15u) == 0) ? 1 : 0;
        switch (W) {
        case 1: return launch_encode_reg<1>(p, prog, numSMs, st);
        case 2: return launch_encode_reg<2>(p, prog, numSMs, st);
        case 4: return launch_encode_reg<4>(p, prog, numSMs, st);
        case 8: return launch_encode_reg<8>(p, prog, numSMs, st);
        case 16: return launch_encode_reg<16>(p, prog, numSMs, st);
        case 32: return launch_encode_reg<32>(p, prog, numSMs, st);
        }
    }
    // one frame per thread whenever the codeword fits the registers and the info block is whole (aligned) words
    p.byteOut = ((uintptr_t)p.code & 3u) ? 1 : 0;
    if (p.N >= 32 && p.N <= 1024 && (p.kb & 3) == 0 && p.kb >= 4 && (((uintptr_t)p.info | (uintptr_t)p.code) & 3u) == 0) {
        switch (W) {
        case 1: return launch_encode_thread<1>(p, numSMs, st);
        case 2: return launch_encode_thread<2>(p, numSMs, st);
        case 4: return launch_encode_thread<4>(p, numSMs, st);
        case 8: return launch_encode_thread<8>(p, numSMs, st);
        case 16: return launch_encode_thread<16>(p, numSMs, st);
        case 32: return launch_encode_thread<32>(p, numSMs, st);
        }
    }
    const int fpw = W < 32 ? 32 / W : 1;
    const long long groups = (p.nFrames + fpw - 1) / fpw;
    const int warpsPerBlock = 8;
    long long blocks = (groups + warpsPerBlock - 1) / warpsPerBlock;
    const long long cap = (long long)numSMs * 8;
    if (blocks > cap) blocks = cap;
    const int wpl = W <= 32 ? 1 : W / 32;
    switch (wpl) {
    case 1: encode_kernel<1><<<(int)blocks, warpsPerBlock * 32, 0, st>>>(p); break;
    case 2: encode_kernel<2><<<(int)blocks, warpsPerBlock * 32, 0, st>>>(p); break;
    case 4: encode_kernel<4><<<(int)blocks, warpsPerBlock * 32, 0, st>>>(p); break;
    default: return -1;
    }
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

} // namespace polar
