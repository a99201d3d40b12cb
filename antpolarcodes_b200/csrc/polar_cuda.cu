// libpolar_cuda: C ABI (include/polar_cuda.h) over the sm_100a kernels.
#include "../../include/polar_cuda.h"

#include "construct.h"
#include "crc_host.h"
#include "batch_kernels.cuh"
#include "decode_params.h"
#include "encode.cuh"
#include "fastssc_i8_soft.cuh"
#include "kernels.h"
#include "puncture.cuh"
#include "schedule.h"
#include "sim.cuh"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <list>
#include <map>
#include <mutex>
#include <string>
#include <vector>

using namespace polar;

namespace {

thread_local std::string g_lastError;

int fail(int code, const std::string& msg)
{
    g_lastError = msg;
    return code;
}

#define CUDA_TRY(expr)                                                                                \
    do {                                                                                              \
        cudaError_t e__ = (expr);                                                                     \
        if (e__ != cudaSuccess)                                                                       \
            return fail(POLAR_E_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e__));           \
    } while (0)

// Compiled-schedule cache: decoder objects are rebuilt whenever setSystematic / setErrorDetection
// change, one per host thread and per sweep point -- the node program only depends on (tree flavour, N, frozen set,
// list size), so it is compiled once.  Small LRU, most recent first; entries are compared on the full frozen set.
struct CachedSchedule {
    int tk, N, L;
    std::vector<uint32_t> frozen, ops;
    ScheduleStats stats;
};
std::mutex g_schedMu;
std::list<CachedSchedule> g_schedCache;
long g_schedHits = 0, g_schedMisses = 0;
constexpr size_t kSchedCacheEntries = 64;

bool compile_schedule_cached(TreeKind tk, int N, const std::vector<uint32_t>& frozen, int L, std::vector<uint32_t>& ops,
                             ScheduleStats& stats, const char** err)
{
    {
        std::lock_guard<std::mutex> lk(g_schedMu);
        for (auto it = g_schedCache.begin(); it != g_schedCache.end(); ++it)
            if (it->tk == (int)tk && it->N == N && it->L == L && it->frozen == frozen) {
                ops = it->ops;
                stats = it->stats;
                g_schedCache.splice(g_schedCache.begin(), g_schedCache, it);
                ++g_schedHits;
                return true;
            }
    }
    if (!compile_schedule(tk, N, frozen, L, ops, stats, err)) return false;
    std::lock_guard<std::mutex> lk(g_schedMu);
    ++g_schedMisses;
    g_schedCache.push_front(CachedSchedule{ (int)tk, N, L, frozen, ops, stats });
    if (g_schedCache.size() > kSchedCacheEntries) g_schedCache.pop_back();
    return true;
}

} // namespace

// ---------------------------------------------------------------- scratch of the allocation-free entry point
// polar_cuda_decode_batch_noalloc (end of the decode section) serves callers that may not allocate device memory: code
// under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.  Its kernels take the scratch that
// holds the top LLR / partial-sum stages from this file-scope array instead of the handle's cudaMalloc'ed buffer.
// Every GPU that loads the library holds the array until the process exits; every call on a device shares it.
// In HOST code the name below is only a stand-in for the symbol, not a device address: the address is looked up once per
// device, with that device current, by cudaGetSymbolAddress (noalloc_scratch_for) and that pointer goes to the kernels.
static __device__ __align__(256) unsigned char g_noallocScratch[POLAR_NOALLOC_SCRATCH_BYTES];

namespace {
std::mutex g_noallocMu;
constexpr int kNoallocMaxDevices = 64;
void* g_noallocPtr[kNoallocMaxDevices] = {};
cudaError_t g_noallocErr[kNoallocMaxDevices] = {};

// Device address of g_noallocScratch on `device` (which must be the current device).  Looked up at set-up time
// (polar_cuda_code_create) so that the entry point itself only reads the cache.
int noalloc_scratch_for(int device, void** ptr)
{
    if (device < 0 || device >= kNoallocMaxDevices) return fail(POLAR_E_ARG, "no-alloc scratch: device index out of range");
    std::lock_guard<std::mutex> lk(g_noallocMu);
    if (!g_noallocPtr[device] && g_noallocErr[device] == cudaSuccess) {
        void* d = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&d, g_noallocScratch);
        if (e != cudaSuccess) { g_noallocErr[device] = e; cudaGetLastError(); }
        else g_noallocPtr[device] = d;
    }
    if (!g_noallocPtr[device])
        return fail(POLAR_E_CUDA, std::string("cudaGetSymbolAddress(no-alloc scratch): ") + cudaGetErrorString(g_noallocErr[device]));
    *ptr = g_noallocPtr[device];
    return POLAR_OK;
}
} // namespace

struct polar_code {
    int device = 0, kind = 0, N = 0, lgN = 0, K = 0, kb = 0, L = 1, systematic = 1, crcBits = 0;
    std::vector<uint32_t> frozen;
    std::vector<uint32_t> ops;
    ScheduleStats stats;
    Schedule sched;
    // device tables
    uint16_t* dInfoPos = nullptr;  // K
    uint2* dWordTab = nullptr;     // per code word: info-position mask, info bits before it (encoder)
    uint32_t* dSynTab = nullptr;   // K
    uint32_t* dGenTab = nullptr;   // K
    uint32_t* dGenNib = nullptr;   // encoder with a CRC, kb % 4 == 0: kb * 2 nibbles x 16 values (encode_reg_kernel)
    uint32_t synInit = 0, genInit = 0;
    // launch geometry
    DecodeKernel kernel = nullptr;
    DecodeParams geo{};            // geometry fields only
    int warpsPerBlock = 0, blocksPerSM = 0, numSMs = 0;
    size_t smemPerBlock = 0;
    uint32_t* dRuns = nullptr;     // Fast-SSC: runs of consecutive information positions
    int nRuns = 0;
    polar::EncodeProgram* encProg = nullptr; // encoder, 32 <= N <= 1024: the scatter program of encode_reg_kernel (kernel parameter)
    uint32_t* dSpill = nullptr;    // list decoders: L2-resident scratch for the top LLR stage
    size_t spillBytes = 0, l2WindowBytes = 0;
    uint32_t* dSpill2 = nullptr;   // HOST pipeline: the second stream's scratch (allocated on first use)
    float l2Persist = 0.f;         // > 0: the scratch is launched as a persisting L2 access-policy window with this hit ratio
    int8_t* dQuant = nullptr;      // int8 copy of fp32 LLRs (quantiser pre-pass)
    size_t quantBytes = 0;
    // staging buffers for HOST-memory calls
    void* dLlr = nullptr;
    uint8_t* dOut = nullptr;
    uint8_t* dOk = nullptr;
    uint8_t* dInfo = nullptr;
    size_t stageFrames = 0;
    int stageLlrBytes = 0;
    cudaStream_t pipe[2] = { nullptr, nullptr }; // HOST-buffer calls: double-buffered copy / compute pipeline
    cudaEvent_t pipeDone[2] = { nullptr, nullptr };
    long launches = 0;
    // the spill / quantiser scratch belongs to the handle: a launch on another stream waits for the previous one
    cudaEvent_t scratchFree = nullptr;
    cudaStream_t scratchStream = nullptr;
    bool scratchBusy = false;
    uint32_t* dSoftScratch = nullptr; // soft-output kernel: LLR stages + byte partial sums per resident warp
    size_t softScratchWords = 0;
    // adaptive kinds: the two decoders and the compaction scratch
    polar_code* fast = nullptr;
    polar_code* list = nullptr;
    uint32_t* dFailIdx = nullptr;  // frame numbers the list decoder redoes
    uint32_t* dFailCount = nullptr;
    uint32_t* hFailCount = nullptr; // pinned
    void* dGather = nullptr;       // dense LLR rows of those frames
    uint8_t* dOutC = nullptr;
    uint8_t* dOkC = nullptr;
    size_t idxFrames = 0, gatherFrames = 0;
    int gatherRowBytes = 0;
    long fallbacks = 0;
};

namespace {

int env_int(const char* name, int dflt)
{
    const char* v = getenv(name);
    return v && *v ? atoi(v) : dflt;
}

// The scratch of the spilled stages as a persisting L2 access-policy window (launch attribute of the decoder kernel): the
// input that streams through the cache no longer evicts it.  Measured on B200 (profiles/r2_experiments.txt t12): Fast-SSC int8
// N=1024 (57 MB of scratch) 181.6 -> 190.6 Gbit/s; the list decoders, whose scratch (124 - 250 MB) exceeds what the device can
// set aside, lose 2 - 10 %.  So: on when the whole scratch fits the set-aside (POLAR_CUDA_L2_PERSIST unset or -1), 0 = never,
// 1 = always (hit ratio = the share that fits).  The set-aside is per device and only ever grows to the largest scratch seen.
void setup_l2_window(polar_code* c, const cudaDeviceProp& prop)
{
    static std::mutex mu;
    static std::map<int, size_t> asideOf;
    const int mode = env_int("POLAR_CUDA_L2_PERSIST", -1);
    if (mode == 0 || !c->spillBytes || prop.persistingL2CacheMaxSize <= 0) return;
    size_t cap = (size_t)prop.persistingL2CacheMaxSize;
    if (const int mb = env_int("POLAR_CUDA_L2_PERSIST_MB", 0)) cap = std::min(cap, (size_t)mb << 20);
    if (mode < 0 && c->spillBytes > cap) return;
    const size_t want = std::min(cap, c->spillBytes);
    {
        std::lock_guard<std::mutex> lk(mu);
        size_t& cur = asideOf[c->device];
        if (want > cur) {
            if (cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want) != cudaSuccess) { cudaGetLastError(); return; }
            cur = want;
        }
    }
    c->l2WindowBytes = std::min(c->spillBytes, (size_t)prop.accessPolicyMaxWindowSize);
    c->l2Persist = std::min(1.0f, (float)want / (float)c->l2WindowBytes);
}

int setup_decoder_geometry(polar_code* c)
{
    DecodeParams& g = c->geo;
    g.N = c->N;
    g.lgN = c->lgN;
    g.K = c->K;
    g.kb = c->kb;
    g.crcBits = c->crcBits;
    g.systematic = c->systematic;
    g.listSize = c->L;
    g.smemStages = c->lgN;
    // barrier every 2^k schedule steps: often enough that the warps of a block keep fetching the same instruction
    // lines, rarely enough that they do not wait for each other at every short step (measured per kernel on B200:
    // SCL int8 1 -> 64: 10.05 -> 10.42 Gbit/s, Fast-SSC int8 1 -> 128: 130 -> 145.5, Fast-SSC fp32 1 -> 64: 77 -> 88.2,
    // SCL fp32 1 -> 8: 4.11 -> 4.19; without any barrier the kernels stall on instruction fetch)
    g.lockstep = env_int("POLAR_CUDA_LOCKSTEP", c->kind == POLAR_KIND_FASTSSC_I8 ? 256 : c->kind == POLAR_KIND_SCL_F32 ? 8 : 64); // (Fast-SSC int8 after the 16-byte tiles: 32 -> 188.7, 64 -> 192.2, 128 -> 193.8, 256 -> 196.3 Gbit/s)
    if (g.lockstep & (g.lockstep - 1)) g.lockstep = 1; // the kernels test pc & (lockstep - 1)
    g.prefetchNext = env_int("POLAR_CUDA_PREFETCH", 0);
    g.rootLeaf = 0;
    g.llrBase = 0; g.bitBase = 0; g.bitSpill = c->lgN + 1; g.spA = g.spB0 = g.spB1 = 0;
    g.regStages = c->kind == POLAR_KIND_SCL_I8 ? std::min(env_int("POLAR_CUDA_REG_STAGES", 4), c->lgN - 2)
                : c->kind == POLAR_KIND_SCL_F32 ? std::min(env_int("POLAR_CUDA_F32_REG_STAGES", 4), c->lgN - 2) : -1;
    if (c->kind == POLAR_KIND_SCL_F32 && g.regStages < 1) g.regStages = 0; // 0 = every stage through the slots (round-1 layout)
    g.bitB1 = 0;
    g.spillWarpW = 0;
    g.chanRow = -1;
    const bool scl = (c->kind == POLAR_KIND_SCL_I8 || c->kind == POLAR_KIND_SCL_F32);
    const bool sclF32 = (c->kind == POLAR_KIND_SCL_F32);
    const int L = scl ? c->L : 1;
    bool wide = false; // Fast-SSC kernels: 16-warp blocks at 128 registers
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, c->device));
    c->numSMs = prop.multiProcessorCount;
    if (scl) {
        // One list per group of L lanes, 32/L frames per warp.  Per frame: L LLR slots (stages below
        // spillStage) and L bit slots (array A: stages 0..lgN, array B: stages 0..lgN-1).  Strides are
        // == 1 (mod L) and the frame stride == L (mod 32), so the 32 lanes of a warp -- lane = (frame, path)
        // -- hit 32 distinct banks when every path walks its own slot.
        // The top LLR stages live in an L2-resident scratch instead (coalesced [word][lane] layout), which is
        // what sets the resident warp count; region A of the bit slots holds stages 0..lgN-1 (the root is
        // assembled by the epilogue), region B is a ping-pong pair.
        // (fp32 lists: one float per word, stage s at word 2^s - 1 of the slot)
        const bool f32Regs = sclF32 && g.regStages >= 1; // fp32 lists with the stages <= regStages in registers
        int nSpill = env_int("POLAR_CUDA_SPILL", c->lgN >= 10 ? (sclF32 ? (f32Regs ? 4 : 5) : (c->lgN >= 11 && L >= 16 ? 4 : 3)) : // N=2048, L=32: 8 -> 10 resident warps, 1.56 -> 1.67 Gbit/s
                                                  c->lgN >= 9 ? (sclF32 ? 3 : 2) : (c->lgN >= 7 ? 1 : 0));
        nSpill = std::max(0, std::min(nSpill, c->lgN - 5));
        if (sclF32) // keep a warp's footprint within reach of a few resident warps per SM
            while (nSpill < c->lgN - 5 &&
                   (size_t)32 * (f32_woff(c->lgN - nSpill) - (f32Regs ? f32_woff(g.regStages + 1) : 0)) * 4 > 96 * 1024)
                ++nSpill;
        g.smemStages = c->lgN - nSpill;
        g.rootLeaf = (!c->ops.empty() && (int)POLAR_OP_STAGE(c->ops[0]) == c->lgN) ? 1 : 0;
        // int8 lists: the stages <= regStages never touch shared memory, so the slots only hold the stages above them
        const int firstSm = sclF32 ? (f32Regs ? g.regStages + 1 : 0) : std::max(0, g.regStages + 1);
        g.llrBase = sclF32 ? f32_woff(firstSm) : llr_woff(firstSm);
        int slot = sclF32 ? std::max(1, f32_woff(g.smemStages) - g.llrBase) : std::max(1, llr_woff(g.smemStages) - g.llrBase);
        while (slot % L != 1 % L) ++slot;
        g.slotW = slot;
        const int nW = std::max(1, c->N / 32);
        auto stageWords = [](int s) { return s >= 5 ? 1 << (s - 5) : 1; };
        // partial sums: the two top stages of a long int8 code (each used by one G step and one combine per frame) go to
        // the L2 scratch next to the spilled LLR stages; with them gone the footprint is set by stages 5 .. lgN-3
        const int nBitSpill = (sclF32 || g.rootLeaf || !nSpill) ? 0 : std::max(0, std::min(env_int("POLAR_CUDA_BIT_SPILL", c->lgN >= 10 ? 2 : 0), c->lgN - 6));
        g.bitSpill = nBitSpill ? c->lgN - nBitSpill : c->lgN + 1;
        const int topSm = std::min(c->lgN - 1, g.bitSpill - 1); // highest bit stage in shared memory
        g.bitBase = bit_woff(firstSm);
        const int bitA = bit_woff(topSm + 1) - g.bitBase + (g.rootLeaf ? nW : 0);
        const int topEven = topSm & ~1, topOdd = (topSm & 1) ? topSm : topSm - 1;
        g.bitBW = bitA;
        g.bitB1 = stageWords(topEven);
        int bslot = bitA + stageWords(topEven) + stageWords(std::max(topOdd, 1));
        while (bslot % L != 1 % L) ++bslot;
        g.bitSlotW = bslot;
        g.bitsOff = L * g.slotW;
        g.selOff = L * g.slotW + L * g.bitSlotW;
        int frame = L * g.slotW + L * g.bitSlotW + (sclF32 ? 17 : 9) * L; // + select scratch (candidate keys, swap history)
        while (frame % 32 != L % 32) ++frame;
        g.frameW = frame;
        g.scratchW = 2 * ((c->K + 31) / 32) + nW + 4;
        g.warpW = ((32 / L) * g.frameW + g.scratchW + 3) & ~3;
        if (nSpill) {
            int words = sclF32 ? ((1 << c->lgN) - (1 << g.smemStages)) : ((1 << (c->lgN - 2)) - (1 << (g.smemStages - 2)));
            if (nBitSpill) { // region A of the spilled stages, then the even / odd ping-pong buffers of region B
                const int top = c->lgN - 1, sEven = top & ~1, sOdd = (top & 1) ? top : top - 1;
                g.spA = words;
                words += (1 << (c->lgN - 5)) - (1 << (g.bitSpill - 5));
                g.spB0 = words;
                words += sEven >= g.bitSpill ? stageWords(sEven) : 0;
                g.spB1 = words;
                words += sOdd >= g.bitSpill ? stageWords(sOdd) : 0;
            }
            g.spillWarpW = 32 * words;
        }
        c->kernel = sclF32 ? scl_f32_kernel_for(L) : scl_i8_kernel_for(L, 512);
        // dense handler ids for the kernels' jump tables
        for (size_t i = 0; i < c->ops.size(); ++i)
            c->sched.ops[i] = (c->ops[i] & 0xFFFFu) | (scl_handler_id(c->ops[i], g.regStages) << 16);
    } else if (c->kind == POLAR_KIND_FASTSSC_F32) {
        // fp32 Fast-SSC: same row stack, one float per word; stage s sits at row 2^s - 1.  Stages from 6 up
        // (64 floats and more) go to the L2-resident scratch.
        const bool rootLeaf = !c->ops.empty() && (int)POLAR_OP_STAGE(c->ops[0]) == c->lgN;
        int keep = env_int("POLAR_CUDA_F32_SMEM_STAGES", 6);
        g.smemStages = rootLeaf ? c->lgN : std::max(3, std::min(keep, c->lgN));
        const int nSpill = c->lgN - g.smemStages;
        const int imgRows = (8 * c->kb + 31) / 32;
        g.llrRows = std::max(f32_woff(g.smemStages), imgRows);
        g.bitRows = std::max(1, c->N / 32);
        const int tileRows = 66;
        int rows;
        if (nSpill) { g.llrRows = std::max(g.llrRows, tileRows); g.tileRow = 0; rows = g.llrRows + g.bitRows; }
        else if (f32_woff(c->lgN - 1) >= tileRows) { g.tileRow = 0; rows = g.llrRows + g.bitRows; }
        else { rows = g.llrRows + g.bitRows; g.tileRow = rows; rows += tileRows; }
        if (rootLeaf) { g.chanRow = rows; rows += c->N; }
        g.outRow = 0;
        g.warpW = rows * 32;
        if (nSpill) g.spillWarpW = 32 * (f32_woff(c->lgN) - f32_woff(g.smemStages));
        wide = env_int("POLAR_CUDA_SC_WIDE", 1) != 0;
        c->kernel = fastssc_f32_kernel_for(wide);
    } else {
        // One frame per thread; a warp's shared memory is a stack of 32-word rows: LLR stages below smemStages,
        // then the packed partial sums.  The top stages live in an L2-resident scratch with the same
        // [row][lane] layout (every access a full 128-byte line), which sets the resident warp count.
        const bool rootLeaf = !c->ops.empty() && (int)POLAR_OP_STAGE(c->ops[0]) == c->lgN;
        int nSpill = env_int("POLAR_CUDA_SC_SPILL", c->lgN >= 9 ? 2 : (c->lgN >= 8 ? 1 : 0));
        nSpill = rootLeaf ? 0 : std::max(0, std::min(nSpill, c->lgN - 6));
        g.smemStages = c->lgN - nSpill;
        const int imgRows = (8 * c->kb + 31) / 32; // image of the warp's 32 output rows (reuses the stage rows)
        g.llrRows = std::max(llr_woff(g.smemStages), imgRows);
        g.bitRows = std::max(1, c->N / 32);
        int rows = g.llrRows + g.bitRows;
        const int tileRows = 66; // two 32x33 transpose tiles
        // the shared-memory stage rows below the top-level destination are dead at both top-level steps
        if (nSpill) { g.llrRows = std::max(g.llrRows, tileRows); g.tileRow = 0; rows = g.llrRows + g.bitRows; }
        else if (llr_woff(c->lgN - 1) >= tileRows) g.tileRow = 0;
        else { g.tileRow = rows; rows += tileRows; }
        if (rootLeaf) { g.chanRow = rows; rows += c->N / 4; }
        g.outRow = 0;
        // TMA bulk copies for the channel tiles (rows of whole 128-byte segments, N >= 256; one more row holds the mbarrier).
        // Measured on B200 (N=1024 K=512, 2^20 frames): 32 per-lane 128-byte cp.async.bulk copies per tile (UBLKCP takes
        // uniform operands, so the warp issues them one lane after the other) 146.9 Gbit/s against 181.7 for the 4-byte
        // cp.async tiles below -- off by default, POLAR_CUDA_TMA=1 selects it (profiles/r2_s17_tma_ab.txt).
        g.useTma = (!rootLeaf && c->N >= 256 && env_int("POLAR_CUDA_TMA", 0)) ? 1 : 0;
        g.mbarRow = rows;
        if (g.useTma) rows += 1;
        g.warpW = rows * 32;
        if (nSpill) g.spillWarpW = 32 * (llr_woff(c->lgN) - llr_woff(g.smemStages));
        wide = env_int("POLAR_CUDA_SC_WIDE", 1) != 0;
        c->kernel = fastssc_i8_kernel_for(wide);
        g.regStages = rootLeaf ? -1 : std::min(env_int("POLAR_CUDA_SC_REG_STAGES", 4), c->lgN - 2);
        for (size_t i = 0; i < c->ops.size(); ++i) // dense handler ids in place of the op codes
        {
            const uint32_t h = sc_handler_id(c->ops[i], g.regStages, c->lgN);
            c->sched.ops[i] = (c->ops[i] & ~(63u | (1u << 10))) | (h & 63u) | ((h & 64u) << 4);
        }
    }
    if (!c->kernel) return fail(POLAR_E_UNSUPPORTED, "no kernel for this list size");
    const size_t perWarp = (size_t)g.warpW * 4;
    if (c->kind == POLAR_KIND_SCL_I8) {
        // one block per SM (one lockstep domain), as many warps as the shared memory holds; the kernel exists in three
        // register budgets (launch bounds 512 / 768 / 1024 threads = 128 / 80 / 64 registers)
        if (perWarp > (size_t)prop.sharedMemPerBlockOptin)
            return fail(POLAR_E_UNSUPPORTED, "code does not fit in shared memory (N x listSize too large)");
        int w = (int)std::min<size_t>(prop.sharedMemPerBlockOptin / perWarp, 32);
        w = std::min(w, env_int("POLAR_CUDA_SCL_MAX_WARPS", POLAR_SCL_MID_THREADS / 32)); // 64 registers spill: 24 warps at 80 registers is the widest useful block
        if (const int forceW = env_int("POLAR_CUDA_WARPS_PER_BLOCK", 0)) w = std::min(w, forceW);
        const int maxT = w <= 16 ? 512 : w <= POLAR_SCL_MID_THREADS / 32 ? POLAR_SCL_MID_THREADS : 1024;
        c->kernel = scl_i8_kernel_for(L, maxT);
        CUDA_TRY(cudaFuncSetAttribute((const void*)c->kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)prop.sharedMemPerBlockOptin));
        int nb = 0;
        CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, (const void*)c->kernel, w * 32, perWarp * w));
        if (nb < 1) return fail(POLAR_E_UNSUPPORTED, "kernel cannot be resident with this shared memory footprint");
        // small codes: several such blocks fit; more than two lockstep domains per SM lose the shared instruction fetch
        nb = std::min(nb, env_int("POLAR_CUDA_MAX_BLOCKS_PER_SM", w >= 16 ? 1 : 2));
        c->warpsPerBlock = w;
        c->blocksPerSM = nb;
        c->smemPerBlock = perWarp * w;
        if (g.spillWarpW) {
            const size_t words = (size_t)g.spillWarpW * c->warpsPerBlock * c->blocksPerSM * c->numSMs;
            CUDA_TRY(cudaMalloc(&c->dSpill, words * sizeof(uint32_t)));
            c->spillBytes = words * sizeof(uint32_t);
            setup_l2_window(c, prop);
        }
        return POLAR_OK;
    }
    if (perWarp > (size_t)prop.sharedMemPerBlockOptin)
        return fail(POLAR_E_UNSUPPORTED, "code does not fit in shared memory (N x listSize too large)");
    // The dynamic shared-memory limit belongs to the kernel FUNCTION on the device, not to this handle: it is raised
    // to the device maximum once and never lowered, so handles with different geometries (and host threads that each
    // own a decoder) can share one template instance.
    CUDA_TRY(cudaFuncSetAttribute((const void*)c->kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)prop.sharedMemPerBlockOptin));
    int bestWarps = 0, bestW = 1, bestB = 1;
    const int forceW = env_int("POLAR_CUDA_WARPS_PER_BLOCK", 0);
    // equal resident warps: wide blocks first (one or two lockstep domains per SM keep the instruction fetch shared)
    // (the Fast-SSC kernels take the widest block instead: one barrier domain per SM)
    static const int sclOrder[] = { 16, 8, 12, 6, 4, 10, 5, 7, 3, 2, 1, 9 }; // one barrier domain per SM first (10.39 -> 10.76 Gbit/s at N=1024, L=8)
    static const int sclF32Order[] = { 8, 4, 6, 5, 7, 3, 2, 1, 12, 16, 10, 9 }; // measured (register stages): 8 x 2 4.64, 4 x 4 4.51, 16 x 1 4.13 Gbit/s
    static const int scOrder[] = { 16, 12, 8, 6, 4, 3, 2, 1, 14, 10, 7, 5 };
    for (int i = 0; i < 12; ++i) {
        const int w = sclF32 ? sclF32Order[i] : scl ? sclOrder[i] : scOrder[i];
        if (forceW && w != forceW) continue;
        if (!scl && w > (wide ? POLAR_SC_WIDE_THREADS / 32 : 8)) continue; // launch bounds of the Fast-SSC kernels
        const size_t smem = perWarp * w;
        if (smem > (size_t)prop.sharedMemPerBlockOptin) continue;
        int nb = 0;
        CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, (const void*)c->kernel, w * 32, smem));
        nb = std::min(nb, env_int("POLAR_CUDA_MAX_BLOCKS_PER_SM", 64)); // experiments: cap the occupancy
        if (nb * w > bestWarps) { bestWarps = nb * w; bestW = w; bestB = nb; }
    }
    if (!bestWarps) return fail(POLAR_E_UNSUPPORTED, "kernel cannot be resident with this shared memory footprint");
    c->warpsPerBlock = bestW;
    c->blocksPerSM = bestB;
    c->smemPerBlock = perWarp * bestW;
    if (g.spillWarpW) {
        const size_t words = (size_t)g.spillWarpW * c->warpsPerBlock * c->blocksPerSM * c->numSMs;
        CUDA_TRY(cudaMalloc(&c->dSpill, words * sizeof(uint32_t)));
        c->spillBytes = words * sizeof(uint32_t);
        setup_l2_window(c, prop);
    }
    return POLAR_OK;
}

int ensure_staging(polar_code* c, size_t nFrames, int llrBytesPerElem)
{
    if (nFrames <= c->stageFrames && llrBytesPerElem <= c->stageLlrBytes) return POLAR_OK;
    cudaFree(c->dLlr); cudaFree(c->dOut); cudaFree(c->dOk); cudaFree(c->dInfo);
    c->dLlr = nullptr; c->dOut = c->dOk = c->dInfo = nullptr;
    c->stageFrames = 0;
    const size_t f = std::max(nFrames, (size_t)1);
    const int eb = std::max(llrBytesPerElem, c->stageLlrBytes);
    if (c->kind == POLAR_KIND_ENC) {
        CUDA_TRY(cudaMalloc(&c->dInfo, f * c->kb));
        CUDA_TRY(cudaMalloc(&c->dOut, f * (size_t)std::max(1, c->N / 8)));
    } else {
        CUDA_TRY(cudaMalloc(&c->dLlr, f * c->N * (size_t)eb));
        CUDA_TRY(cudaMalloc(&c->dOut, f * c->kb));
        CUDA_TRY(cudaMalloc(&c->dOk, f));
    }
    c->stageFrames = f;
    c->stageLlrBytes = eb;
    return POLAR_OK;
}

} // namespace

// ---------------------------------------------------------------- C ABI
extern "C" {

const char* polar_cuda_last_error(void) { return g_lastError.c_str(); }

int polar_cuda_version(void) { return 100; }

int polar_cuda_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int polar_cuda_code_destroy(polar_code_t c)
{
    if (!c) return POLAR_OK;
    cudaSetDevice(c->device);
    cudaFree(c->dInfoPos); cudaFree(c->dWordTab); cudaFree(c->dSynTab); cudaFree(c->dGenTab); cudaFree(c->dGenNib);
    cudaFree(c->dLlr); cudaFree(c->dOut); cudaFree(c->dOk); cudaFree(c->dInfo);
    if (c->l2Persist > 0.f) cudaCtxResetPersistingL2Cache(); // the lines of the scratch window go back to normal before the memory is freed
    cudaFree(c->dSpill); cudaFree(c->dSpill2); cudaFree(c->dQuant); cudaFree(c->dRuns); cudaFree(c->dSoftScratch);
    cudaFree(c->dFailIdx); cudaFree(c->dFailCount); cudaFree(c->dGather); cudaFree(c->dOutC); cudaFree(c->dOkC);
    if (c->hFailCount) cudaFreeHost(c->hFailCount);
    if (c->scratchFree) cudaEventDestroy(c->scratchFree);
    delete[] c->encProg;
    polar_cuda_code_destroy(c->fast);
    polar_cuda_code_destroy(c->list);
    for (int i = 0; i < 2; ++i) {
        if (c->pipe[i]) cudaStreamDestroy(c->pipe[i]);
        if (c->pipeDone[i]) cudaEventDestroy(c->pipeDone[i]);
    }
    delete c;
    return POLAR_OK;
}

int polar_cuda_code_create(int device, int kind, int N, const uint32_t* frozen, int nFrozen, int listSize,
                           int systematic, int crcBits, polar_code_t* out)
{
    if (!out) return fail(POLAR_E_ARG, "out is null");
    *out = nullptr;
    if (kind < POLAR_KIND_ENC || kind > POLAR_KIND_ADAPTIVE_MIXED) return fail(POLAR_E_ARG, "unknown kind");
    if (kind >= POLAR_KIND_ADAPTIVE_I8) {
        // AdaptiveChar / AdaptiveFloat / AdaptiveMixed constructors (adaptive_char.cpp:15-26): one Fast-SSC and one
        // list decoder over the same code, detector and systematic flag
        const int fastKind = kind == POLAR_KIND_ADAPTIVE_F32 ? POLAR_KIND_FASTSSC_F32 : POLAR_KIND_FASTSSC_I8;
        const int listKind = kind == POLAR_KIND_ADAPTIVE_I8 ? POLAR_KIND_SCL_I8 : POLAR_KIND_SCL_F32;
        polar_code_t fast = nullptr, list = nullptr;
        int rc = polar_cuda_code_create(device, fastKind, N, frozen, nFrozen, 1, systematic, crcBits, &fast);
        if (rc) return rc;
        if (listSize > 1) {
            rc = polar_cuda_code_create(device, listKind, N, frozen, nFrozen, listSize, systematic, crcBits, &list);
            if (rc) { const std::string msg = g_lastError; polar_cuda_code_destroy(fast); return fail(rc, msg); }
        }
        polar_code* c = new polar_code;
        c->device = device; c->kind = kind; c->N = N; c->lgN = fast->lgN; c->K = fast->K; c->kb = fast->kb;
        c->L = listSize > 1 ? listSize : 1; c->systematic = fast->systematic; c->crcBits = crcBits;
        c->numSMs = fast->numSMs;
        c->frozen = fast->frozen;
        c->fast = fast; c->list = list;
        if (cudaMalloc(&c->dFailCount, sizeof(uint32_t)) != cudaSuccess ||
            cudaMallocHost(&c->hFailCount, sizeof(uint32_t)) != cudaSuccess) {
            polar_cuda_code_destroy(c);
            return fail(POLAR_E_CUDA, "adaptive decoder: counter allocation failed");
        }
        *out = c;
        return POLAR_OK;
    }
    if (N < 8 || N > 4096 || (N & (N - 1))) return fail(POLAR_E_ARG, "block length must be a power of two in [8, 4096]");
    if (nFrozen < 0 || nFrozen > N || (nFrozen && !frozen)) return fail(POLAR_E_ARG, "bad frozen set");
    if (crcBits != 0 && crcBits != 8 && crcBits != 16 && crcBits != 32) return fail(POLAR_E_ARG, "CRC INVALID SIZE!");
    const int K = N - nFrozen;
    // decoders hand ceil(K/8) zero-padded bytes to Detector::check (fastssc_fip_char.cpp:627, scl_fip_char.cpp:808),
    // so K need not be a multiple of 8 there; the encoder signs floor(K/8) bytes (butterfly_fip_packed.cpp:48)
    const int crcBytes = kind == POLAR_KIND_ENC ? K / 8 : (K + 7) / 8;
    if (crcBits && crcBytes < crcBits / 8) return fail(POLAR_E_ARG, "with a CRC, the info block must be at least as long as the CRC");
    if (crcBits && kind == POLAR_KIND_ENC && K % 8 != 0)
        return fail(POLAR_E_ARG, "encoder with a CRC: K must be a multiple of 8");
    if (crcBits == 32 && crcBytes < 4) return fail(POLAR_E_ARG, "CRC32 needs at least 4 info bytes");
    if (K < 1) return fail(POLAR_E_ARG, "no information bits");
    if (cudaSetDevice(device) != cudaSuccess) {
        cudaGetLastError();
        return fail(POLAR_E_CUDA, "cudaSetDevice failed (no CUDA device?)");
    }
    polar_code* c = new polar_code;
    c->device = device; c->kind = kind; c->N = N; c->K = K; c->kb = (K + 7) / 8;
    c->L = (kind == POLAR_KIND_SCL_I8 || kind == POLAR_KIND_SCL_F32) ? listSize : 1;
    c->systematic = systematic ? 1 : 0; c->crcBits = crcBits;
    while ((1 << c->lgN) < N) ++c->lgN;
    c->frozen.assign(frozen, frozen + nFrozen);
    int rc = POLAR_OK;
    do {
        const char* err = nullptr;
        if (kind != POLAR_KIND_ENC) {
            TreeKind tk = kind == POLAR_KIND_SCL_I8 ? TREE_SCL_I8 : kind == POLAR_KIND_SCL_F32 ? TREE_SCL_F32
                          : kind == POLAR_KIND_FASTSSC_F32 ? TREE_FASTSSC_F32 : TREE_FASTSSC_I8;
            if (!compile_schedule_cached(tk, N, c->frozen, c->L, c->ops, c->stats, &err)) { rc = fail(POLAR_E_ARG, err); break; }
            memset(&c->sched, 0, sizeof(c->sched));
            memcpy(c->sched.ops, c->ops.data(), c->ops.size() * sizeof(uint32_t));
        } else {
            for (size_t i = 0; i < c->frozen.size(); ++i)
                if (c->frozen[i] >= (uint32_t)N || (i && c->frozen[i] <= c->frozen[i - 1])) {
                    rc = fail(POLAR_E_ARG, "frozen indices must be sorted, unique and < N");
                    break;
                }
            if (rc) break;
        }
        // info-position LUT (bitcontainer.cpp:68-84) and its inverse
        std::vector<uint16_t> infoPos;
        const int nWords = std::max(1, N / 32);
        std::vector<uint2> wordTab(nWords, make_uint2(0u, 0u));
        {
            std::vector<uint8_t> fm(N, 0);
            for (uint32_t f : c->frozen) fm[f] = 1;
            for (int i = 0; i < N; ++i) {
                if ((i & 31) == 0) wordTab[i >> 5].y = (uint32_t)infoPos.size();
                if (!fm[i]) { wordTab[i >> 5].x |= 1u << (i & 31); infoPos.push_back((uint16_t)i); }
            }
        }
        CrcTables t = build_crc_tables(crcBits, K);
        c->synInit = t.synInit; c->genInit = t.genInit;
        cudaError_t e;
        if ((e = cudaMalloc(&c->dInfoPos, K * sizeof(uint16_t))) != cudaSuccess ||
            (e = cudaMalloc(&c->dWordTab, nWords * sizeof(uint2))) != cudaSuccess ||
            (e = cudaMalloc(&c->dSynTab, K * sizeof(uint32_t))) != cudaSuccess ||
            (e = cudaMalloc(&c->dGenTab, K * sizeof(uint32_t))) != cudaSuccess ||
            (e = cudaMemcpy(c->dInfoPos, infoPos.data(), K * sizeof(uint16_t), cudaMemcpyHostToDevice)) != cudaSuccess ||
            (e = cudaMemcpy(c->dWordTab, wordTab.data(), nWords * sizeof(uint2), cudaMemcpyHostToDevice)) != cudaSuccess ||
            (e = cudaMemcpy(c->dSynTab, t.syn.data(), K * sizeof(uint32_t), cudaMemcpyHostToDevice)) != cudaSuccess ||
            (e = cudaMemcpy(c->dGenTab, t.gen.data(), K * sizeof(uint32_t), cudaMemcpyHostToDevice)) != cudaSuccess) {
            rc = fail(POLAR_E_CUDA, std::string("table upload: ") + cudaGetErrorString(e));
            break;
        }
        if (kind == POLAR_KIND_FASTSSC_I8 || kind == POLAR_KIND_FASTSSC_F32) {
            // runs of consecutive information positions that stay inside one 32-bit word on both sides
            std::vector<uint32_t> runs;
            for (int k = 0; k < K;) {
                int len = 1;
                while (k + len < K && infoPos[k + len] == infoPos[k] + len && ((infoPos[k] + len) & 31) != 0 &&
                       ((k + len) & 31) != 0)
                    ++len;
                runs.push_back((uint32_t)infoPos[k] | ((uint32_t)len << 16));
                k += len;
            }
            c->nRuns = (int)runs.size();
            if ((e = cudaMalloc(&c->dRuns, runs.size() * sizeof(uint32_t))) != cudaSuccess ||
                (e = cudaMemcpy(c->dRuns, runs.data(), runs.size() * sizeof(uint32_t), cudaMemcpyHostToDevice)) != cudaSuccess) {
                rc = fail(POLAR_E_CUDA, std::string("run table upload: ") + cudaGetErrorString(e));
                break;
            }
        }
        if (kind == POLAR_KIND_ENC && crcBits && (c->kb & 3) == 0) {
            // checksum contribution of every value of every info nibble (stream bits 4n .. 4n+3, first bit = value bit 3)
            std::vector<uint32_t> nib((size_t)c->kb * 2 * 16, 0u);
            for (int n = 0; n < c->kb * 2; ++n)
                for (int v = 1; v < 16; ++v) {
                    uint32_t x = 0;
                    for (int b = 0; b < 4; ++b)
                        if (((v >> (3 - b)) & 1) && 4 * n + b < K) x ^= t.gen[4 * n + b];
                    nib[(size_t)n * 16 + v] = x;
                }
            if ((e = cudaMalloc(&c->dGenNib, nib.size() * sizeof(uint32_t))) != cudaSuccess ||
                (e = cudaMemcpy(c->dGenNib, nib.data(), nib.size() * sizeof(uint32_t), cudaMemcpyHostToDevice)) != cudaSuccess) {
                rc = fail(POLAR_E_CUDA, std::string("CRC nibble table upload: ") + cudaGetErrorString(e));
                break;
            }
        }
        if (kind == POLAR_KIND_ENC) {
            // encoder: per code word the runs of consecutive information positions (b | len << 5 | first info bit << 11)
            // behind a table of run offsets per word
            std::vector<uint32_t> tab(nWords + 1, 0u), runs;
            for (int w = 0; w < nWords; ++w) {
                tab[w] = (uint32_t)runs.size();
                uint32_t m = wordTab[w].x;
                int k = (int)wordTab[w].y;
                while (m) {
                    const int b = __builtin_ctz(m);
                    const uint32_t rest = ~(m >> b);
                    const int len = rest ? __builtin_ctz(rest) : 32;
                    runs.push_back((uint32_t)b | ((uint32_t)len << 5) | ((uint32_t)k << 11));
                    m = len + b >= 32 ? 0u : m & (0xFFFFFFFFu << (b + len));
                    k += len;
                }
            }
            tab[nWords] = (uint32_t)runs.size();
            if (N >= 32 && N <= 1024 && (int)runs.size() <= ENC_MAX_RUNS && !env_int("POLAR_CUDA_ENC_V1", 0)) {
                // the same runs for the bit-reversed words of encode_reg_kernel: source row, funnel shift, destination mask
                EncodeProgram* g2 = new EncodeProgram[2]; // [0]: one frame per thread, [1]: two (row strides differ)
                memset(g2, 0, 2 * sizeof(EncodeProgram));
                for (int fpt = 1; fpt <= 2; ++fpt) {
                    EncodeProgram* g = g2 + (fpt - 1);
                    for (int w = 0; w < nWords; ++w) {
                        g->first[w] = (uint16_t)tab[w];
                        uint32_t rev = 0;
                        for (int i = 0; i < 32; ++i) rev |= ((wordTab[w].x >> i) & 1u) << (31 - i);
                        g->maskRev[w] = rev;
                    }
                    for (int w = nWords; w < 34; ++w) g->first[w] = (uint16_t)runs.size();
                    for (size_t r = 0; r < runs.size(); ++r) {
                        const int b = runs[r] & 31u, len = (runs[r] >> 5) & 63u, k = runs[r] >> 11, s = k & 31, wi = k >> 5;
                        const uint32_t hiRow = s >= b ? wi + 1 : wi, sh = s >= b ? s - b : 32 - (b - s);
                        g->run[r].x = hiRow * (uint32_t)(ENC_REG_THREADS * fpt * 4);
                        g->run[r].y = sh;
                        g->run[r].z = (len == 32 ? 0xFFFFFFFFu : ((1u << len) - 1u)) << (32 - b - len);
                    }
                }
                c->encProg = g2;
            }
            tab.insert(tab.end(), runs.begin(), runs.end());
            c->nRuns = nWords + 1; // offset of the run entries inside the table
            if ((e = cudaMalloc(&c->dRuns, tab.size() * sizeof(uint32_t))) != cudaSuccess ||
                (e = cudaMemcpy(c->dRuns, tab.data(), tab.size() * sizeof(uint32_t), cudaMemcpyHostToDevice)) != cudaSuccess) {
                rc = fail(POLAR_E_CUDA, std::string("run table upload: ") + cudaGetErrorString(e));
                break;
            }
        }
        if (kind != POLAR_KIND_ENC) {
            rc = setup_decoder_geometry(c);
        } else {
            cudaDeviceProp prop;
            if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) { rc = fail(POLAR_E_CUDA, "cudaGetDeviceProperties"); break; }
            c->numSMs = prop.multiProcessorCount;
        }
    } while (0);
    if (rc) { polar_cuda_code_destroy(c); return rc; }
    if (kind != POLAR_KIND_ENC) { // set-up for polar_cuda_decode_batch_noalloc: the per-device address of its static scratch
        void* unused = nullptr;   // (a failure is kept and reported by that entry point; the other paths do not need it)
        const std::string keep = g_lastError;
        noalloc_scratch_for(device, &unused);
        g_lastError = keep;
    }
    *out = c;
    return POLAR_OK;
}

// One batch with DEVICE buffers on stream st: optional quantiser pre-pass + the decoder kernel.
// spillOverride: the HOST pipeline gives each of its two streams a scratch of its own, so the kernels of consecutive chunks may
// overlap (the caller then owns the ordering: the handle's cross-stream guard is skipped)
static int decode_on_device(polar_code* c, const void* llr, int llrType, size_t nFrames, uint8_t* out, uint8_t* ok,
                            cudaStream_t st, uint8_t* cand = nullptr, uint32_t* spillOverride = nullptr, bool ownOrdering = false)
{
    DecodeParams p = c->geo;
    p.cand = cand;
    p.infoPos = c->dInfoPos;
    p.crcTab = c->dSynTab;
    p.crcInit = c->synInit;
    p.nFrames = (long long)nFrames;
    p.llr = llr; p.out = out; p.ok = ok;
    p.spill = spillOverride ? spillOverride : c->dSpill;
    p.runs = c->dRuns;
    p.nRuns = c->nRuns;
    if ((reinterpret_cast<uintptr_t>(llr) & 3u) != 0) // the kernels read the rows as 32-bit words (float rows are aligned by type)
        return fail(POLAR_E_ARG, "DEVICE LLR buffer must be 4-byte aligned");
    if (!ownOrdering && c->scratchBusy && c->scratchStream != st) CUDA_TRY(cudaStreamWaitEvent(st, c->scratchFree, 0));
    const bool f32Kind = (c->kind == POLAR_KIND_FASTSSC_F32 || c->kind == POLAR_KIND_SCL_F32);
    if (llrType == POLAR_LLR_F32 && !f32Kind) { // CharContainer::insertLlr(const float*): quantise once, decode the int8 copy
        const size_t need = nFrames * (size_t)c->N;
        if (need > c->quantBytes) {
            cudaFree(c->dQuant);
            c->dQuant = nullptr;
            c->quantBytes = 0;
            CUDA_TRY(cudaMalloc(&c->dQuant, need));
            c->quantBytes = need;
        }
        const long long nWords = (long long)(need / 4);
        const int qgrid = (int)std::min<long long>((nWords + 255) / 256, (long long)c->numSMs * 16);
        quantize_llr_kernel<<<qgrid, 256, 0, st>>>(static_cast<const float4*>(p.llr), reinterpret_cast<uint32_t*>(c->dQuant),
                                                  nWords, c->N >= 32 ? 1 : 0);
        c->launches++;
        CUDA_TRY(cudaGetLastError());
        p.llr = c->dQuant;
    }
    if (p.useTma && ((uintptr_t)p.llr & 15u)) p.useTma = 0; // bulk copies need 16-byte aligned rows: 4-byte cp.async tiles instead
    p.chanVec = (((uintptr_t)p.llr & 15u) == 0 && ((size_t)c->N * (f32Kind ? 4 : 1)) % 16 == 0) ? 1 : 0;
    const int framesPerWarp = (c->kind == POLAR_KIND_SCL_I8 || c->kind == POLAR_KIND_SCL_F32) ? 32 / c->L : 32;
    const long long warpsNeeded = ((long long)nFrames + framesPerWarp - 1) / framesPerWarp;
    const long long blocksNeeded = (warpsNeeded + c->warpsPerBlock - 1) / c->warpsPerBlock;
    const int grid = (int)std::min<long long>(blocksNeeded, (long long)c->blocksPerSM * c->numSMs);
    if (p.spill && c->l2Persist > 0.f) {
        // the scratch of the spilled stages as a persisting L2 window: the input streaming through the cache does not evict it
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3(grid); cfg.blockDim = dim3(c->warpsPerBlock * 32); cfg.dynamicSmemBytes = c->smemPerBlock; cfg.stream = st;
        cudaLaunchAttribute at{};
        at.id = cudaLaunchAttributeAccessPolicyWindow;
        at.val.accessPolicyWindow.base_ptr = p.spill;
        at.val.accessPolicyWindow.num_bytes = c->l2WindowBytes;
        at.val.accessPolicyWindow.hitRatio = c->l2Persist;
        at.val.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
        at.val.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
        cfg.attrs = &at; cfg.numAttrs = 1;
        CUDA_TRY(cudaLaunchKernelEx(&cfg, c->kernel, p, c->sched));
    } else {
        c->kernel<<<grid, c->warpsPerBlock * 32, c->smemPerBlock, st>>>(p, c->sched);
    }
    c->launches++;
    CUDA_TRY(cudaGetLastError());
    if (!ownOrdering && (c->dSpill || c->dQuant)) { // batches of one handle on different streams must not share the scratch concurrently
        if (!c->scratchFree) CUDA_TRY(cudaEventCreateWithFlags(&c->scratchFree, cudaEventDisableTiming));
        CUDA_TRY(cudaEventRecord(c->scratchFree, st));
        c->scratchStream = st;
        c->scratchBusy = true;
    }
    return POLAR_OK;
}

// AdaptiveChar::decode (adaptive_char.cpp:34-46) over a batch with DEVICE buffers: Fast-SSC on every frame; the
// frames it fails are compacted, gathered into a dense batch for the list decoder and scattered back.
static int adaptive_on_device(polar_code* c, const void* llr, int llrType, size_t nFrames, uint8_t* out, uint8_t* ok,
                              cudaStream_t st)
{
    polar_code* fast = c->fast;
    int rc = decode_on_device(fast, llr, llrType, nFrames, out, ok, st);
    if (rc || !c->list) return rc;
    if (nFrames > c->idxFrames) {
        cudaFree(c->dFailIdx);
        c->dFailIdx = nullptr; c->idxFrames = 0;
        CUDA_TRY(cudaMalloc(&c->dFailIdx, nFrames * sizeof(uint32_t)));
        c->idxFrames = nFrames;
    }
    CUDA_TRY(cudaMemsetAsync(c->dFailCount, 0, sizeof(uint32_t), st));
    const int cgrid = (int)std::min<long long>(((long long)nFrames + 255) / 256, (long long)c->numSMs * 8);
    adaptive_compact_kernel<<<cgrid, 256, 0, st>>>(ok, (long long)nFrames, c->dFailIdx, c->dFailCount);
    c->launches++;
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaMemcpyAsync(c->hFailCount, c->dFailCount, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st)); // the list decoder's grid is sized by the count
    const size_t nFail = *c->hFailCount;
    if (!nFail) return POLAR_OK;
    c->fallbacks += (long)nFail;
    // the list decoder reads what its CPU twin would: the int8 copy for SclFipChar, the float LLRs for SclAvxFloat
    const bool listF32 = c->list->kind == POLAR_KIND_SCL_F32;
    const void* src = llr;
    if (!listF32 && llrType == POLAR_LLR_F32) src = fast->dQuant; // quantised once by the Fast-SSC pass
    const int rowBytes = c->N * (listF32 ? 4 : 1);
    if (nFail > c->gatherFrames || rowBytes > c->gatherRowBytes) {
        cudaFree(c->dGather); cudaFree(c->dOutC); cudaFree(c->dOkC);
        c->dGather = nullptr; c->dOutC = c->dOkC = nullptr; c->gatherFrames = 0;
        const size_t cap = std::max(nFail, std::min(nFrames, std::max<size_t>(1024, 2 * nFail)));
        const int rb = std::max(rowBytes, c->gatherRowBytes);
        CUDA_TRY(cudaMalloc(&c->dGather, cap * (size_t)rb));
        CUDA_TRY(cudaMalloc(&c->dOutC, cap * (size_t)c->kb));
        CUDA_TRY(cudaMalloc(&c->dOkC, cap));
        c->gatherFrames = cap; c->gatherRowBytes = rb;
    }
    const bool v16 = rowBytes % 16 == 0 && ((uintptr_t)src & 15) == 0;
    const long long vecs = (long long)nFail * (rowBytes / (v16 ? 16 : 4));
    const int ggrid = (int)std::min<long long>((vecs + 255) / 256, (long long)c->numSMs * 16);
    if (v16)
        adaptive_gather_kernel<uint4><<<ggrid, 256, 0, st>>>(static_cast<const uint4*>(src), c->dFailIdx, (long long)nFail,
                                                             rowBytes / 16, static_cast<uint4*>(c->dGather));
    else
        adaptive_gather_kernel<uint32_t><<<ggrid, 256, 0, st>>>(static_cast<const uint32_t*>(src), c->dFailIdx, (long long)nFail,
                                                                rowBytes / 4, static_cast<uint32_t*>(c->dGather));
    c->launches++;
    CUDA_TRY(cudaGetLastError());
    rc = decode_on_device(c->list, c->dGather, listF32 ? POLAR_LLR_F32 : POLAR_LLR_I8, nFail, c->dOutC, c->dOkC, st);
    if (rc) return rc;
    const long long bytes = (long long)nFail * c->kb;
    const int sgrid = (int)std::min<long long>((bytes + 255) / 256, (long long)c->numSMs * 16);
    adaptive_scatter_kernel<<<sgrid, 256, 0, st>>>(c->dOutC, c->dOkC, c->dFailIdx, (long long)nFail, c->kb, out, ok);
    c->launches++;
    CUDA_TRY(cudaGetLastError());
    return POLAR_OK;
}

static int adaptive_decode_batch(polar_code* c, const void* llr, int llrType, int mem, size_t nFrames, uint8_t* info_out,
                                 uint8_t* ok_out, cudaStream_t st)
{
    if (c->kind != POLAR_KIND_ADAPTIVE_I8 && llrType != POLAR_LLR_F32)
        return fail(POLAR_E_ARG, "AdaptiveFloat / AdaptiveMixed take float LLRs");
    const int eb = llrType == POLAR_LLR_F32 ? 4 : 1;
    if (mem == POLAR_MEM_DEVICE) {
        uint8_t* okDev = ok_out;
        if (!okDev) {
            int rc = ensure_staging(c, nFrames, 0);
            if (rc) return rc;
            okDev = c->dOk;
        }
        return adaptive_on_device(c, llr, llrType, nFrames, info_out, okDev, st);
    }
    // HOST buffers: chunks of about 128 MB of LLRs (large enough that the failed frames of a chunk still fill the list
    // kernel's grid), double buffered over two internal streams.  The H2D copy of chunk k+1 is queued before chunk k is
    // decoded, so it runs under both kernels of chunk k; the kernels of consecutive chunks are chained by events
    // (they share the compaction scratch and the decoders' spill buffers).
    size_t chunk = std::max<size_t>(1, ((size_t)env_int("POLAR_CUDA_ADAPTIVE_CHUNK_MB", 128) << 20) / ((size_t)c->N * eb));
    if (const char* v = getenv("POLAR_CUDA_CHUNK_MB")) // the knob of the plain decoders overrides, when set
        if (*v) chunk = std::max<size_t>(1, ((size_t)atoi(v) << 20) / ((size_t)c->N * eb));
    if (chunk * 2 > nFrames || !env_int("POLAR_CUDA_PIPELINE", 1)) chunk = nFrames;
    const size_t nChunks = (nFrames + chunk - 1) / chunk;
    const int depth = nChunks > 1 ? 2 : 1;
    int rc = ensure_staging(c, chunk * depth, eb);
    if (rc) return rc;
    if (depth == 2 && !c->pipe[0]) {
        for (int i = 0; i < 2; ++i) {
            CUDA_TRY(cudaStreamCreateWithFlags(&c->pipe[i], cudaStreamNonBlocking));
            CUDA_TRY(cudaEventCreateWithFlags(&c->pipeDone[i], cudaEventDisableTiming));
        }
    }
    CUDA_TRY(cudaStreamSynchronize(st));
    const uint8_t* src = static_cast<const uint8_t*>(llr);
    auto stageIn = [&](size_t k) -> cudaError_t {
        const size_t f0 = k * chunk, n = std::min(chunk, nFrames - f0);
        const int h = depth == 2 ? (int)(k & 1) : 0;
        uint8_t* dL = static_cast<uint8_t*>(c->dLlr) + (size_t)h * chunk * c->N * eb;
        return cudaMemcpyAsync(dL, src + f0 * c->N * eb, n * c->N * (size_t)eb, cudaMemcpyHostToDevice, depth == 2 ? c->pipe[h] : st);
    };
    CUDA_TRY(stageIn(0));
    for (size_t k = 0; k < nChunks; ++k) {
        const size_t f0 = k * chunk, n = std::min(chunk, nFrames - f0);
        const int h = depth == 2 ? (int)(k & 1) : 0;
        cudaStream_t s = depth == 2 ? c->pipe[h] : st;
        uint8_t* dL = static_cast<uint8_t*>(c->dLlr) + (size_t)h * chunk * c->N * eb;
        uint8_t* dO = c->dOut + (size_t)h * chunk * c->kb;
        uint8_t* dK = c->dOk + (size_t)h * chunk;
        if (k + 1 < nChunks) CUDA_TRY(stageIn(k + 1)); // behind the D2H of chunk k-1 on that stream
        if (depth == 2 && k > 0) CUDA_TRY(cudaStreamWaitEvent(s, c->pipeDone[h ^ 1], 0));
        rc = adaptive_on_device(c, dL, llrType, n, dO, dK, s);
        if (rc) return rc;
        if (depth == 2) CUDA_TRY(cudaEventRecord(c->pipeDone[h], s));
        CUDA_TRY(cudaMemcpyAsync(info_out + f0 * c->kb, dO, n * (size_t)c->kb, cudaMemcpyDeviceToHost, s));
        if (ok_out) CUDA_TRY(cudaMemcpyAsync(ok_out + f0, dK, n, cudaMemcpyDeviceToHost, s));
    }
    if (depth == 2) {
        CUDA_TRY(cudaStreamSynchronize(c->pipe[0]));
        CUDA_TRY(cudaStreamSynchronize(c->pipe[1]));
    } else {
        CUDA_TRY(cudaStreamSynchronize(st));
    }
    return POLAR_OK;
}

int polar_cuda_decode_batch(polar_code_t c, const void* llr, int llrType, int mem, size_t nFrames,
                            uint8_t* info_out, uint8_t* ok_out, void* stream)
{
    if (!c) return fail(POLAR_E_ARG, "null code handle");
    if (c->kind == POLAR_KIND_ENC) return fail(POLAR_E_ARG, "handle is an encoder");
    if (nFrames == 0) return POLAR_OK;
    if (!llr || !info_out) return fail(POLAR_E_ARG, "null buffer");
    if (llrType != POLAR_LLR_I8 && llrType != POLAR_LLR_F32) return fail(POLAR_E_ARG, "bad llr type");
    if ((c->kind == POLAR_KIND_FASTSSC_F32 || c->kind == POLAR_KIND_SCL_F32) && llrType != POLAR_LLR_F32)
        return fail(POLAR_E_ARG, "an fp32 decoder takes float LLRs");
    CUDA_TRY(cudaSetDevice(c->device));
    cudaStream_t st = (cudaStream_t)stream;
    if (c->kind >= POLAR_KIND_ADAPTIVE_I8) return adaptive_decode_batch(c, llr, llrType, mem, nFrames, info_out, ok_out, st);
    const int eb = llrType == POLAR_LLR_F32 ? 4 : 1;
    if (mem == POLAR_MEM_DEVICE) {
        uint8_t* okDev = ok_out;
        if (!okDev) { // the kernels always write the verdict
            int rc = ensure_staging(c, nFrames, 0);
            if (rc) return rc;
            okDev = c->dOk;
        }
        return decode_on_device(c, llr, llrType, nFrames, info_out, okDev, st);
    }
    // HOST buffers: chunks of whole grid waves, double buffered over two internal streams, so that the H2D copy
    // of chunk k+1 and the D2H copy of chunk k-1 run under the kernel of chunk k.  The kernels themselves are
    // chained by events (they share the spill / quantiser scratch of the handle).  A launch costs a fixed ~60 us
    // (measured: one-wave launches run at 13.0 Gbit/s, 16-wave launches at 14.6), a large first chunk leaves its copy
    // exposed: the chunks therefore GROW -- 1, 2, 4, ... waves up to POLAR_CUDA_CHUNK_MB -- so the pipeline fills
    // after one wave and runs on large launches from then on.
    const int framesPerWarp = (c->kind == POLAR_KIND_SCL_I8 || c->kind == POLAR_KIND_SCL_F32) ? 32 / c->L : 32;
    const size_t wave = (size_t)framesPerWarp * c->warpsPerBlock * c->blocksPerSM * c->numSMs;
    const size_t rowBytes = (size_t)c->N * eb;
    // with a scratch per stream (below) the kernels of consecutive chunks overlap, a launch costs nothing and the last chunk is what
    // stays exposed: chunks of about two waves (32 MB) are best then (headline e2e 14.25 -> 14.4, fp32 lists 5.19 -> 5.41 Gbit/s;
    // one-wave chunks 13.8); chained kernels keep the large launches
    const bool nativeType = llrType == ((c->kind == POLAR_KIND_FASTSSC_F32 || c->kind == POLAR_KIND_SCL_F32) ? POLAR_LLR_F32 : POLAR_LLR_I8);
    const bool mayOverlap = nativeType && c->l2Persist <= 0.f && env_int("POLAR_CUDA_OVERLAP", 1) != 0;
    const size_t maxChunk = std::max<size_t>(1, (((size_t)env_int("POLAR_CUDA_CHUNK_MB", mayOverlap ? 32 : 128) << 20) / rowBytes) / wave) * wave;
    const size_t firstChunk = std::max<size_t>(1, (((size_t)env_int("POLAR_CUDA_FIRST_CHUNK_MB", 8) << 20) / rowBytes) / wave) * wave;
    std::vector<std::pair<size_t, size_t>> chunks; // (first frame, frames)
    if (std::min(firstChunk, maxChunk) * 2 > nFrames || !env_int("POLAR_CUDA_PIPELINE", 1)) {
        chunks.emplace_back(0, nFrames); // too small to split
    } else {
        size_t f0 = 0, sz = std::min(firstChunk, maxChunk);
        while (f0 < nFrames) {
            size_t n = std::min(sz, nFrames - f0);
            if (nFrames - f0 - n < wave / 2) n = nFrames - f0; // do not leave a sliver for the last launch
            chunks.emplace_back(f0, n);
            f0 += n;
            sz = std::min(sz * 2, maxChunk);
        }
    }
    size_t largest = 0;
    for (auto& ch : chunks) largest = std::max(largest, ch.second);
    const size_t nChunks = chunks.size();
    const int depth = nChunks > 1 ? 2 : 1;
    int rc = ensure_staging(c, largest * depth, eb);
    if (rc) return rc;
    if (depth == 2 && !c->pipe[0]) {
        for (int i = 0; i < 2; ++i) {
            CUDA_TRY(cudaStreamCreateWithFlags(&c->pipe[i], cudaStreamNonBlocking));
            CUDA_TRY(cudaEventCreateWithFlags(&c->pipeDone[i], cudaEventDisableTiming));
        }
    }
    CUDA_TRY(cudaStreamSynchronize(st)); // work queued before this call on the caller's stream comes first
    // Two scratches (one per pipeline stream) let the kernel of chunk k+1 start on the SMs that the tail of chunk k has left;
    // with one shared scratch the kernels are chained by events.  Not with the quantiser pre-pass (its scratch is shared too) and
    // not while the scratch is pinned in L2 (two windows would fight for the set-aside).
    bool overlap = depth == 2 && mayOverlap;
    if (overlap && c->dSpill && !c->dSpill2 && cudaMalloc(&c->dSpill2, c->spillBytes) != cudaSuccess) {
        cudaGetLastError();
        c->dSpill2 = nullptr;
        overlap = false;
    }
    if (overlap && c->scratchBusy) CUDA_TRY(cudaEventSynchronize(c->scratchFree)); // a DEVICE-buffer batch on another stream still owns the scratch
    const uint8_t* src = static_cast<const uint8_t*>(llr);
    for (size_t k = 0; k < nChunks; ++k) {
        const size_t f0 = chunks[k].first, n = chunks[k].second;
        const int h = depth == 2 ? (int)(k & 1) : 0;
        cudaStream_t s = depth == 2 ? c->pipe[h] : st;
        uint8_t* dL = static_cast<uint8_t*>(c->dLlr) + (size_t)h * largest * rowBytes;
        uint8_t* dO = c->dOut + (size_t)h * largest * c->kb;
        uint8_t* dK = c->dOk + (size_t)h * largest;
        CUDA_TRY(cudaMemcpyAsync(dL, src + f0 * rowBytes, n * rowBytes, cudaMemcpyHostToDevice, s));
        if (depth == 2 && k > 0 && !overlap) CUDA_TRY(cudaStreamWaitEvent(s, c->pipeDone[h ^ 1], 0)); // kernel k-1 has finished
        rc = overlap ? decode_on_device(c, dL, llrType, n, dO, dK, s, nullptr, h ? c->dSpill2 : c->dSpill, true)
                     : decode_on_device(c, dL, llrType, n, dO, dK, s);
        if (rc) return rc;
        if (depth == 2) CUDA_TRY(cudaEventRecord(c->pipeDone[h], s));
        CUDA_TRY(cudaMemcpyAsync(info_out + f0 * c->kb, dO, n * (size_t)c->kb, cudaMemcpyDeviceToHost, s));
        if (ok_out) CUDA_TRY(cudaMemcpyAsync(ok_out + f0, dK, n, cudaMemcpyDeviceToHost, s));
    }
    if (depth == 2) {
        CUDA_TRY(cudaStreamSynchronize(c->pipe[0]));
        CUDA_TRY(cudaStreamSynchronize(c->pipe[1]));
    } else {
        CUDA_TRY(cudaStreamSynchronize(st));
    }
    return POLAR_OK;
}

// Candidate hand-back (a detector the kernels do not fuse is run by the host over all L candidates).
int polar_cuda_decode_candidates(polar_code_t c, const void* llr, int llrType, int mem, size_t nFrames, uint8_t* cand_out,
                                 uint8_t* count_out, void* stream)
{
    if (!c) return fail(POLAR_E_ARG, "null code handle");
    if (c->kind == POLAR_KIND_ENC || c->kind >= POLAR_KIND_ADAPTIVE_I8)
        return fail(POLAR_E_ARG, "candidate hand-back needs a Fast-SSC or list decoder handle");
    if (c->crcBits) return fail(POLAR_E_ARG, "candidate hand-back: create the handle without a fused CRC");
    if (nFrames == 0) return POLAR_OK;
    if (!llr || !cand_out || !count_out) return fail(POLAR_E_ARG, "null buffer");
    if (llrType != POLAR_LLR_I8 && llrType != POLAR_LLR_F32) return fail(POLAR_E_ARG, "bad llr type");
    const bool f32Kind = (c->kind == POLAR_KIND_FASTSSC_F32 || c->kind == POLAR_KIND_SCL_F32);
    if (f32Kind && llrType != POLAR_LLR_F32) return fail(POLAR_E_ARG, "an fp32 decoder takes float LLRs");
    CUDA_TRY(cudaSetDevice(c->device));
    cudaStream_t st = (cudaStream_t)stream;
    const bool list = (c->kind == POLAR_KIND_SCL_I8 || c->kind == POLAR_KIND_SCL_F32);
    const size_t candBytes = nFrames * (size_t)(list ? c->L : 1) * c->kb;
    if (mem == POLAR_MEM_DEVICE) {
        if (list) return decode_on_device(c, llr, llrType, nFrames, nullptr, count_out, st, cand_out);
        int rc = decode_on_device(c, llr, llrType, nFrames, cand_out, count_out, st); // Fast-SSC: one candidate, verdict 1
        return rc;
    }
    const size_t llrBytes = nFrames * (size_t)c->N * (llrType == POLAR_LLR_F32 ? 4 : 1);
    void* dL = nullptr; uint8_t *dC = nullptr, *dN = nullptr;
    cudaError_t e;
    if ((e = cudaMalloc(&dL, llrBytes)) != cudaSuccess || (e = cudaMalloc(&dC, candBytes)) != cudaSuccess ||
        (e = cudaMalloc(&dN, nFrames)) != cudaSuccess) {
        cudaFree(dL); cudaFree(dC); cudaFree(dN);
        cudaGetLastError();
        return fail(POLAR_E_NOMEM, std::string("candidate staging: ") + cudaGetErrorString(e));
    }
    int rc = POLAR_OK;
    e = cudaMemcpyAsync(dL, llr, llrBytes, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) rc = list ? decode_on_device(c, dL, llrType, nFrames, nullptr, dN, st, dC) : decode_on_device(c, dL, llrType, nFrames, dC, dN, st);
    if (e == cudaSuccess && !rc) e = cudaMemcpyAsync(cand_out, dC, candBytes, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess && !rc) e = cudaMemcpyAsync(count_out, dN, nFrames, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess && !rc) e = cudaStreamSynchronize(st);
    cudaFree(dL); cudaFree(dC); cudaFree(dN);
    if (rc) return rc;
    if (e != cudaSuccess) return fail(POLAR_E_CUDA, std::string("candidate hand-back: ") + cudaGetErrorString(e));
    return POLAR_OK;
}

// Decoder::getSoftCodeword / getSoftInformation after FastSscFipChar::decode (decoder.cpp:147-152) over a batch.
int polar_cuda_decode_soft_batch(polar_code_t c, const int8_t* llr, int mem, size_t nFrames, int8_t* soft_out,
                                 int8_t* soft_info_out, void* stream)
{
    if (!c) return fail(POLAR_E_ARG, "null code handle");
    if (c->kind != POLAR_KIND_FASTSSC_I8) return fail(POLAR_E_UNSUPPORTED, "soft outputs exist for the int8 Fast-SSC decoder only");
    if (nFrames == 0) return POLAR_OK;
    if (!llr || !soft_out) return fail(POLAR_E_ARG, "null buffer");
    CUDA_TRY(cudaSetDevice(c->device));
    cudaStream_t st = (cudaStream_t)stream;
    SoftParams p{};
    p.N = c->N; p.lgN = c->lgN; p.K = c->K;
    p.bitsOff = llr_woff(c->lgN + 1);
    p.warpW = 32 * (p.bitsOff + std::max(1, c->N / 4));
    p.infoPos = c->dInfoPos;
    p.nFrames = (long long)nFrames;
    const int warpsPerBlock = 8;
    const long long groups = ((long long)nFrames + 31) / 32;
    const int grid = (int)std::min<long long>((groups + warpsPerBlock - 1) / warpsPerBlock, (long long)c->numSMs * (c->N >= 2048 ? 1 : 2));
    const size_t words = (size_t)p.warpW * warpsPerBlock * grid;
    if (words > c->softScratchWords) {
        cudaFree(c->dSoftScratch);
        c->dSoftScratch = nullptr; c->softScratchWords = 0;
        CUDA_TRY(cudaMalloc(&c->dSoftScratch, words * sizeof(uint32_t)));
        c->softScratchWords = words;
    }
    p.scratch = c->dSoftScratch;
    Schedule sched; // the node program with its op codes (c->sched carries the throughput kernel's handler ids instead)
    memset(&sched, 0, sizeof(sched));
    memcpy(sched.ops, c->ops.data(), c->ops.size() * sizeof(uint32_t));
    const size_t llrBytes = nFrames * (size_t)c->N, infoBytes = nFrames * (size_t)c->K;
    int8_t *dL = nullptr, *dS = nullptr, *dI = nullptr;
    if (mem == POLAR_MEM_HOST) {
        cudaError_t e;
        if ((e = cudaMalloc(&dL, llrBytes)) != cudaSuccess || (e = cudaMalloc(&dS, llrBytes)) != cudaSuccess ||
            (soft_info_out && (e = cudaMalloc(&dI, infoBytes)) != cudaSuccess)) {
            cudaFree(dL); cudaFree(dS); cudaFree(dI);
            cudaGetLastError();
            return fail(POLAR_E_NOMEM, std::string("soft-output staging: ") + cudaGetErrorString(e));
        }
        p.llr = dL; p.soft = dS; p.softInfo = soft_info_out ? dI : nullptr;
    } else {
        p.llr = llr; p.soft = soft_out; p.softInfo = soft_info_out;
    }
    cudaError_t e = cudaSuccess;
    if (mem == POLAR_MEM_HOST) e = cudaMemcpyAsync(dL, llr, llrBytes, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) {
        fastssc_i8_soft_kernel<<<grid, warpsPerBlock * 32, 0, st>>>(p, sched);
        c->launches++;
        e = cudaGetLastError();
    }
    if (mem == POLAR_MEM_HOST) {
        if (e == cudaSuccess) e = cudaMemcpyAsync(soft_out, dS, llrBytes, cudaMemcpyDeviceToHost, st);
        if (e == cudaSuccess && soft_info_out) e = cudaMemcpyAsync(soft_info_out, dI, infoBytes, cudaMemcpyDeviceToHost, st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(st);
        cudaFree(dL); cudaFree(dS); cudaFree(dI);
    }
    if (e != cudaSuccess) return fail(POLAR_E_CUDA, std::string("soft-output decode: ") + cudaGetErrorString(e));
    return POLAR_OK;
}

// Decoder::decode_vector over a batch WITHOUT any allocation (see g_noallocScratch above).  Same kernels and launch
// geometry as polar_cuda_decode_batch with DEVICE buffers; only the scratch pointer differs.  Nothing here calls
// cudaMalloc / cudaFree / an allocating routine: no staging, no quantiser copy, no event creation.
int polar_cuda_decode_batch_noalloc(polar_code_t c, const void* llr, int llrType, size_t nFrames, uint8_t* info_out,
                                    uint8_t* ok_out, void* stream)
{
    if (!c) return fail(POLAR_E_ARG, "null code handle");
    if (c->kind != POLAR_KIND_FASTSSC_I8 && c->kind != POLAR_KIND_FASTSSC_F32 && c->kind != POLAR_KIND_SCL_I8 &&
        c->kind != POLAR_KIND_SCL_F32)
        return fail(POLAR_E_ARG, "no-alloc decode: the handle must be a Fast-SSC or list decoder (not an encoder or an adaptive kind)");
    if (nFrames == 0) return POLAR_OK;
    if (!llr || !info_out || !ok_out) return fail(POLAR_E_ARG, "no-alloc decode: llr, info_out and ok_out must be DEVICE buffers (none may be null)");
    const bool f32Kind = (c->kind == POLAR_KIND_FASTSSC_F32 || c->kind == POLAR_KIND_SCL_F32);
    if (llrType != (f32Kind ? POLAR_LLR_F32 : POLAR_LLR_I8))
        return fail(POLAR_E_ARG, "no-alloc decode: LLRs must have the decoder's own type (float LLRs into an int8 decoder need the "
                                 "quantiser's scratch copy: use polar_cuda_decode_batch)");
    // scratch this handle's full grid needs: one spillWarpW-word region per resident warp
    const size_t need = (size_t)c->geo.spillWarpW * sizeof(uint32_t) * c->warpsPerBlock * c->blocksPerSM * c->numSMs;
    if (need > (size_t)POLAR_NOALLOC_SCRATCH_BYTES)
        return fail(POLAR_E_UNSUPPORTED, "no-alloc decode: this code needs " + std::to_string(need) + " bytes of scratch, the static array holds " +
                                             std::to_string((size_t)POLAR_NOALLOC_SCRATCH_BYTES) + " (POLAR_NOALLOC_SCRATCH_BYTES): use polar_cuda_decode_batch");
    CUDA_TRY(cudaSetDevice(c->device));
    void* scratch = nullptr;
    int rc = noalloc_scratch_for(c->device, &scratch); // cached by polar_cuda_code_create
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    DecodeParams p = c->geo;
    p.cand = nullptr;
    p.infoPos = c->dInfoPos;
    p.crcTab = c->dSynTab;
    p.crcInit = c->synInit;
    p.nFrames = (long long)nFrames;
    p.llr = llr; p.out = info_out; p.ok = ok_out;
    p.spill = c->geo.spillWarpW ? static_cast<uint32_t*>(scratch) : nullptr; // every word is written before it is read: nothing to clear
    p.runs = c->dRuns;
    p.nRuns = c->nRuns;
    if (p.useTma && ((uintptr_t)p.llr & 15u)) p.useTma = 0;
    p.chanVec = (((uintptr_t)p.llr & 15u) == 0 && ((size_t)c->N * (f32Kind ? 4 : 1)) % 16 == 0) ? 1 : 0;
    const int framesPerWarp = (c->kind == POLAR_KIND_SCL_I8 || c->kind == POLAR_KIND_SCL_F32) ? 32 / c->L : 32;
    const long long warpsNeeded = ((long long)nFrames + framesPerWarp - 1) / framesPerWarp;
    const long long blocksNeeded = (warpsNeeded + c->warpsPerBlock - 1) / c->warpsPerBlock;
    const int grid = (int)std::min<long long>(blocksNeeded, (long long)c->blocksPerSM * c->numSMs);
    c->kernel<<<grid, c->warpsPerBlock * 32, c->smemPerBlock, st>>>(p, c->sched);
    c->launches++;
    CUDA_TRY(cudaGetLastError());
    return POLAR_OK;
}

int polar_cuda_encode_batch(polar_code_t c, uint8_t* info, int mem, size_t nFrames, uint8_t* code_out, void* stream)
{
    if (!c) return fail(POLAR_E_ARG, "null code handle");
    if (c->kind != POLAR_KIND_ENC) return fail(POLAR_E_ARG, "handle is not an encoder");
    if (nFrames == 0) return POLAR_OK;
    if (!info || !code_out) return fail(POLAR_E_ARG, "null buffer");
    CUDA_TRY(cudaSetDevice(c->device));
    cudaStream_t st = (cudaStream_t)stream;
    EncodeParams p{};
    p.N = c->N; p.lgN = c->lgN; p.K = c->K; p.kb = c->kb; p.nb = std::max(1, c->N / 8);
    p.crcBits = c->crcBits; p.systematic = c->systematic;
    p.wordTab = c->dWordTab; p.genTab = c->dGenTab; p.genNib = c->dGenNib; p.genInit = c->genInit;
    p.runTab = c->dRuns; p.runBase = c->nRuns;
    p.nFrames = (long long)nFrames;
    if (mem == POLAR_MEM_HOST) {
        int rc = ensure_staging(c, nFrames, 0);
        if (rc) return rc;
        CUDA_TRY(cudaMemcpyAsync(c->dInfo, info, nFrames * (size_t)c->kb, cudaMemcpyHostToDevice, st));
        p.info = c->dInfo; p.code = c->dOut;
    } else {
        p.info = info; p.code = code_out;
    }
    int rc = launch_encode(p, c->encProg, c->numSMs, st);
    c->launches++;
    if (rc) return fail(POLAR_E_CUDA, std::string("encode launch: ") + cudaGetErrorString(cudaGetLastError()));
    if (mem == POLAR_MEM_HOST) {
        CUDA_TRY(cudaMemcpyAsync(code_out, c->dOut, nFrames * (size_t)p.nb, cudaMemcpyDeviceToHost, st));
        if (c->crcBits) CUDA_TRY(cudaMemcpyAsync(info, c->dInfo, nFrames * (size_t)c->kb, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaStreamSynchronize(st));
    }
    return POLAR_OK;
}

int polar_cuda_crc_generate(int crcBits, uint8_t* data, int bytes)
{
    if (crcBits != 0 && crcBits != 8 && crcBits != 16 && crcBits != 32) return fail(POLAR_E_ARG, "CRC INVALID SIZE!");
    if (!data || bytes < crcBits / 8 || (crcBits == 32 && bytes < 4)) return fail(POLAR_E_ARG, "buffer too short for the CRC");
    crc_generate(crcBits, data, bytes);
    return POLAR_OK;
}

int polar_cuda_crc_check(int crcBits, const uint8_t* data, int bytes)
{
    if (crcBits != 0 && crcBits != 8 && crcBits != 16 && crcBits != 32) return fail(POLAR_E_ARG, "CRC INVALID SIZE!");
    if (!data || bytes < crcBits / 8 || (crcBits == 32 && bytes < 4)) return fail(POLAR_E_ARG, "buffer too short for the CRC");
    return crc_check(crcBits, data, bytes) ? 1 : 0;
}

int polar_cuda_frozen_bits(int N, int K, float designSnr, const char* type, uint32_t* out)
{
    if (!type || !out) return fail(POLAR_E_ARG, "null argument");
    std::vector<uint32_t> f;
    std::string err;
    if (!construct_frozen_bits(N, K, designSnr, type, f, err)) return fail(POLAR_E_ARG, err);
    for (size_t i = 0; i < f.size(); ++i) out[i] = f[i];
    return (int)f.size();
}

int polar_cuda_code_info(polar_code_t c, struct polar_code_info* info)
{
    if (!c || !info) return fail(POLAR_E_ARG, "null argument");
    if (c->fast) { // adaptive: geometry and algorithmic figures of the Fast-SSC pass every frame takes
        int rc = polar_cuda_code_info(c->fast, info);
        info->kind = c->kind; info->listSize = c->L;
        return rc;
    }
    memset(info, 0, sizeof(*info));
    info->kind = c->kind; info->N = c->N; info->K = c->K; info->listSize = c->L;
    info->systematic = c->systematic; info->crcBits = c->crcBits;
    info->nOps = (int)c->ops.size();
    info->warpsPerBlock = c->warpsPerBlock; info->blocksPerSM = c->blocksPerSM;
    info->smemPerWarp = c->geo.warpW * 4; info->gridBlocks = c->blocksPerSM * c->numSMs;
    info->laneOpsPerFrame = c->stats.laneOps;
    const bool f32 = (c->kind == POLAR_KIND_FASTSSC_F32 || c->kind == POLAR_KIND_SCL_F32);
    info->bytesPerFrame = c->kind == POLAR_KIND_ENC ? (double)(c->kb + std::max(1, c->N / 8))
                                                    : (double)((f32 ? 4 : 1) * c->N + c->kb + 1);
    for (int i = 0; i < 32 && i < (int)OP_COUNT; ++i) { info->opCount[i] = c->stats.count[i]; info->opSize[i] = c->stats.size[i]; }
    return POLAR_OK;
}

int polar_cuda_code_schedule(polar_code_t c, uint32_t* ops, int maxOps)
{
    if (!c) return fail(POLAR_E_ARG, "null code handle");
    if (c->fast) c = c->fast;
    const int n = (int)c->ops.size();
    if (ops) memcpy(ops, c->ops.data(), sizeof(uint32_t) * std::min(n, maxOps));
    return n;
}

long polar_cuda_code_launches(polar_code_t c)
{
    if (!c) return 0;
    return c->launches + (c->fast ? c->fast->launches : 0) + (c->list ? c->list->launches : 0);
}

long polar_cuda_code_fallbacks(polar_code_t c) { return c ? c->fallbacks : 0; }

int polar_cuda_schedule_cache_stats(long* hits, long* misses)
{
    std::lock_guard<std::mutex> lk(g_schedMu);
    if (hits) *hits = g_schedHits;
    if (misses) *misses = g_schedMisses;
    return (int)g_schedCache.size();
}

int polar_cuda_schedule_stats(int kind, int N, const uint32_t* frozen, int nFrozen, int listSize, long* counts,
                              long* sizes, double* laneOpsPerFrame)
{
    if (kind != POLAR_KIND_FASTSSC_I8 && kind != POLAR_KIND_SCL_I8 && kind != POLAR_KIND_FASTSSC_F32 &&
        kind != POLAR_KIND_SCL_F32)
        return fail(POLAR_E_ARG, "kind has no schedule");
    if (nFrozen < 0 || (nFrozen && !frozen)) return fail(POLAR_E_ARG, "bad frozen set");
    const TreeKind tk = kind == POLAR_KIND_FASTSSC_I8 ? TREE_FASTSSC_I8 : kind == POLAR_KIND_SCL_I8 ? TREE_SCL_I8
                      : kind == POLAR_KIND_FASTSSC_F32 ? TREE_FASTSSC_F32 : TREE_SCL_F32;
    std::vector<uint32_t> f(frozen, frozen + nFrozen), ops;
    ScheduleStats st;
    const char* err = nullptr;
    const bool isScl = (tk == TREE_SCL_I8 || tk == TREE_SCL_F32);
    if (!compile_schedule_cached(tk, N, f, isScl ? listSize : 1, ops, st, &err)) return fail(POLAR_E_ARG, err);
    for (int i = 0; i < 32; ++i) {
        if (counts) counts[i] = i < (int)OP_COUNT ? st.count[i] : 0;
        if (sizes) sizes[i] = i < (int)OP_COUNT ? st.size[i] : 0;
    }
    if (laneOpsPerFrame) *laneOpsPerFrame = st.laneOps;
    return (int)ops.size();
}

} // extern "C"

// ---------------------------------------------------------------- rate matching (Puncturer)
struct polar_puncturer {
    int device = 0, blockLength = 0, parent = 0;
    std::vector<uint32_t> pos;
    uint32_t* dPos = nullptr;  // blockLength kept positions (padded to a multiple of 8 entries)
    int32_t* dInv = nullptr;   // parent entries: index into the punctured block or -1
};

namespace {

int puncturer_positions(int blockLength, const uint32_t* frozen, int nFrozen, std::vector<uint32_t>& pos, int& parent)
{
    if (blockLength < 1 || blockLength > (1 << 20)) return fail(POLAR_E_ARG, "bad punctured block length");
    if (nFrozen < 0 || (nFrozen && !frozen)) return fail(POLAR_E_ARG, "bad frozen set");
    parent = 1;
    while (parent < blockLength) parent <<= 1; // round_up_power_of_two (puncturer.cpp:18-28)
    const int nPunct = parent - blockLength;
    if (nPunct > nFrozen)
        return fail(POLAR_E_RANGE, "Number of required puncturing positions exceeds frozen bit positions!");
    // inverse_set_difference (puncturer.cpp:30-43) of the first nPunct frozen positions (std::set_difference wants
    // them ascending, which frozen sets are)
    std::vector<uint8_t> cut(parent, 0);
    for (int i = 0; i < nPunct; ++i) {
        if (frozen[i] >= (uint32_t)parent || (i && frozen[i] <= frozen[i - 1]))
            return fail(POLAR_E_ARG, "frozen indices must be sorted, unique and < parent block length");
        cut[frozen[i]] = 1;
    }
    pos.clear();
    for (int i = 0; i < parent; ++i)
        if (!cut[i]) pos.push_back((uint32_t)i);
    return POLAR_OK;
}

// run `launch(dIn, dOut)` on device buffers; HOST buffers are staged through temporary device memory
template <typename F>
int with_device_buffers(int device, const void* in, size_t inBytes, void* out, size_t outBytes, int mem, cudaStream_t st, F launch)
{
    CUDA_TRY(cudaSetDevice(device));
    if (mem == POLAR_MEM_DEVICE) {
        launch(in, out);
        CUDA_TRY(cudaGetLastError());
        return POLAR_OK;
    }
    void *dIn = nullptr, *dOut = nullptr;
    if (cudaMalloc(&dIn, inBytes) != cudaSuccess || cudaMalloc(&dOut, outBytes) != cudaSuccess) {
        cudaFree(dIn);
        cudaGetLastError();
        return fail(POLAR_E_NOMEM, "staging buffers for the puncturer");
    }
    cudaError_t e = cudaMemcpyAsync(dIn, in, inBytes, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) { launch(dIn, dOut); e = cudaGetLastError(); }
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, dOut, outBytes, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    cudaFree(dIn); cudaFree(dOut);
    if (e != cudaSuccess) return fail(POLAR_E_CUDA, std::string("puncturer: ") + cudaGetErrorString(e));
    return POLAR_OK;
}

int copy_grid(long long items) { return (int)std::min<long long>((items + 255) / 256, 148LL * 32); }

} // namespace

extern "C" {

int polar_cuda_puncturer_positions(int blockLength, const uint32_t* frozen, int nFrozen, uint32_t* out)
{
    std::vector<uint32_t> pos;
    int parent = 0;
    int rc = puncturer_positions(blockLength, frozen, nFrozen, pos, parent);
    if (rc) return rc;
    if (out) memcpy(out, pos.data(), pos.size() * sizeof(uint32_t));
    return (int)pos.size();
}

int polar_cuda_puncturer_destroy(polar_puncturer_t p)
{
    if (!p) return POLAR_OK;
    cudaSetDevice(p->device);
    cudaFree(p->dPos); cudaFree(p->dInv);
    delete p;
    return POLAR_OK;
}

int polar_cuda_puncturer_create(int device, int blockLength, const uint32_t* frozen, int nFrozen, polar_puncturer_t* out)
{
    if (!out) return fail(POLAR_E_ARG, "out is null");
    *out = nullptr;
    std::vector<uint32_t> pos;
    int parent = 0;
    int rc = puncturer_positions(blockLength, frozen, nFrozen, pos, parent);
    if (rc) return rc;
    if (cudaSetDevice(device) != cudaSuccess) {
        cudaGetLastError();
        return fail(POLAR_E_CUDA, "cudaSetDevice failed (no CUDA device?)");
    }
    polar_puncturer* p = new polar_puncturer;
    p->device = device; p->blockLength = blockLength; p->parent = parent; p->pos = pos;
    std::vector<uint32_t> padded(pos);
    padded.resize((pos.size() + 7) & ~(size_t)7, 0u);
    std::vector<int32_t> inv((parent + 3) & ~3, -1);
    for (size_t j = 0; j < pos.size(); ++j) inv[pos[j]] = (int32_t)j;
    cudaError_t e;
    if ((e = cudaMalloc(&p->dPos, padded.size() * sizeof(uint32_t))) != cudaSuccess ||
        (e = cudaMalloc(&p->dInv, inv.size() * sizeof(int32_t))) != cudaSuccess ||
        (e = cudaMemcpy(p->dPos, padded.data(), padded.size() * sizeof(uint32_t), cudaMemcpyHostToDevice)) != cudaSuccess ||
        (e = cudaMemcpy(p->dInv, inv.data(), inv.size() * sizeof(int32_t), cudaMemcpyHostToDevice)) != cudaSuccess) {
        polar_cuda_puncturer_destroy(p);
        return fail(POLAR_E_CUDA, std::string("puncturer tables: ") + cudaGetErrorString(e));
    }
    *out = p;
    return POLAR_OK;
}

int polar_cuda_puncturer_parent_length(polar_puncturer_t p) { return p ? p->parent : fail(POLAR_E_ARG, "null puncturer"); }

int polar_cuda_puncture_batch(polar_puncturer_t p, const void* in, int elemBytes, int mem, size_t nFrames, void* out, void* stream)
{
    if (!p) return fail(POLAR_E_ARG, "null puncturer");
    if (elemBytes != 1 && elemBytes != 4 && elemBytes != 8) return fail(POLAR_E_ARG, "element size must be 1, 4 or 8 bytes");
    if (nFrames == 0) return POLAR_OK;
    if (!in || !out) return fail(POLAR_E_ARG, "null buffer");
    cudaStream_t st = (cudaStream_t)stream;
    const long long total = (long long)nFrames * p->blockLength;
    return with_device_buffers(p->device, in, nFrames * (size_t)p->parent * elemBytes, out, (size_t)total * elemBytes, mem, st,
                               [&](const void* dIn, void* dOut) {
        const int grid = copy_grid(total);
        if (elemBytes == 1)
            puncture_kernel<uint8_t><<<grid, 256, 0, st>>>((const uint8_t*)dIn, p->dPos, p->blockLength, p->parent, (long long)nFrames, (uint8_t*)dOut);
        else if (elemBytes == 4)
            puncture_kernel<uint32_t><<<grid, 256, 0, st>>>((const uint32_t*)dIn, p->dPos, p->blockLength, p->parent, (long long)nFrames, (uint32_t*)dOut);
        else
            puncture_kernel<unsigned long long><<<grid, 256, 0, st>>>((const unsigned long long*)dIn, p->dPos, p->blockLength, p->parent, (long long)nFrames, (unsigned long long*)dOut);
    });
}

int polar_cuda_depuncture_batch(polar_puncturer_t p, const void* in, int elemBytes, int mem, size_t nFrames, void* out, void* stream)
{
    if (!p) return fail(POLAR_E_ARG, "null puncturer");
    if (elemBytes != 1 && elemBytes != 4 && elemBytes != 8) return fail(POLAR_E_ARG, "element size must be 1, 4 or 8 bytes");
    if (nFrames == 0) return POLAR_OK;
    if (!in || !out) return fail(POLAR_E_ARG, "null buffer");
    cudaStream_t st = (cudaStream_t)stream;
    const long long total = (long long)nFrames * p->parent;
    return with_device_buffers(p->device, in, nFrames * (size_t)p->blockLength * elemBytes, out, (size_t)total * elemBytes, mem, st,
                               [&](const void* dIn, void* dOut) {
        if (elemBytes == 1 && p->parent % 4 == 0 && ((uintptr_t)dOut & 3) == 0)
            depuncture_i8x4_kernel<<<copy_grid(total / 4), 256, 0, st>>>((const int8_t*)dIn, p->dInv, p->blockLength, p->parent, (long long)nFrames, (uint32_t*)dOut);
        else if (elemBytes == 1)
            depuncture_kernel<uint8_t><<<copy_grid(total), 256, 0, st>>>((const uint8_t*)dIn, p->dInv, p->blockLength, p->parent, (long long)nFrames, (uint8_t*)dOut);
        else if (elemBytes == 4)
            depuncture_kernel<uint32_t><<<copy_grid(total), 256, 0, st>>>((const uint32_t*)dIn, p->dInv, p->blockLength, p->parent, (long long)nFrames, (uint32_t*)dOut);
        else
            depuncture_kernel<unsigned long long><<<copy_grid(total), 256, 0, st>>>((const unsigned long long*)dIn, p->dInv, p->blockLength, p->parent, (long long)nFrames, (unsigned long long*)dOut);
    });
}

int polar_cuda_puncture_packed_batch(polar_puncturer_t p, const uint8_t* in, int mem, size_t nFrames, uint8_t* out, void* stream)
{
    if (!p) return fail(POLAR_E_ARG, "null puncturer");
    if (p->blockLength % 8 || p->parent % 8) return fail(POLAR_E_ARG, "packed puncturing needs block lengths that are multiples of 8");
    if (nFrames == 0) return POLAR_OK;
    if (!in || !out) return fail(POLAR_E_ARG, "null buffer");
    cudaStream_t st = (cudaStream_t)stream;
    const long long total = (long long)nFrames * (p->blockLength / 8);
    return with_device_buffers(p->device, in, nFrames * (size_t)(p->parent / 8), out, (size_t)total, mem, st,
                               [&](const void* dIn, void* dOut) {
        puncture_packed_kernel<<<copy_grid(total), 256, 0, st>>>((const uint8_t*)dIn, p->dPos, p->blockLength, p->parent, (long long)nFrames, (uint8_t*)dOut);
    });
}

} // extern "C"

// ---------------------------------------------------------------- on-device link simulator (pcsim worker loop)
struct polar_sim {
    int device = 0, N = 0, K = 0, kb = 0, nb = 0, llrIsFloat = 0;
    size_t batch = 0;
    polar_code* enc = nullptr;
    polar_code* dec = nullptr;
    uint8_t *dInfo = nullptr, *dCode = nullptr, *dOut = nullptr, *dOk = nullptr;
    void* dLlr = nullptr;
    size_t lastFrames = 0;
    long launches = 0;
};

extern "C" {

int polar_cuda_sim_destroy(polar_sim_t s)
{
    if (!s) return POLAR_OK;
    cudaSetDevice(s->device);
    cudaFree(s->dInfo); cudaFree(s->dCode); cudaFree(s->dOut); cudaFree(s->dOk); cudaFree(s->dLlr);
    polar_cuda_code_destroy(s->enc);
    polar_cuda_code_destroy(s->dec);
    delete s;
    return POLAR_OK;
}

int polar_cuda_sim_create(int device, int N, const uint32_t* frozen, int nFrozen, int decoderKind, int listSize, int systematic,
                          int crcBits, size_t batchFrames, polar_sim_t* out)
{
    if (!out) return fail(POLAR_E_ARG, "out is null");
    *out = nullptr;
    if (decoderKind < POLAR_KIND_FASTSSC_I8 || decoderKind > POLAR_KIND_ADAPTIVE_MIXED) return fail(POLAR_E_ARG, "not a decoder kind");
    if (batchFrames < 1 || batchFrames > ((size_t)1 << 24)) return fail(POLAR_E_ARG, "batch must be 1 .. 2^24 frames");
    if (N < 4) return fail(POLAR_E_UNSUPPORTED, "the link simulator draws four code positions per thread: block length >= 4");
    polar_sim* s = new polar_sim;
    s->device = device; s->N = N; s->batch = batchFrames;
    int rc = polar_cuda_code_create(device, POLAR_KIND_ENC, N, frozen, nFrozen, 1, systematic, crcBits, &s->enc);
    if (!rc) rc = polar_cuda_code_create(device, decoderKind, N, frozen, nFrozen, listSize, systematic, crcBits, &s->dec);
    if (rc) { const std::string msg = g_lastError; polar_cuda_sim_destroy(s); return fail(rc, msg); }
    s->K = s->enc->K; s->kb = s->enc->kb; s->nb = std::max(1, N / 8);
    s->llrIsFloat = (decoderKind == POLAR_KIND_FASTSSC_F32 || decoderKind == POLAR_KIND_SCL_F32 ||
                     decoderKind == POLAR_KIND_ADAPTIVE_F32 || decoderKind == POLAR_KIND_ADAPTIVE_MIXED) ? 1 : 0;
    cudaError_t e;
    if ((e = cudaMalloc(&s->dInfo, batchFrames * (size_t)s->kb)) != cudaSuccess ||
        (e = cudaMalloc(&s->dCode, batchFrames * (size_t)s->nb)) != cudaSuccess ||
        (e = cudaMalloc(&s->dOut, batchFrames * (size_t)s->kb)) != cudaSuccess ||
        (e = cudaMalloc(&s->dOk, batchFrames)) != cudaSuccess ||
        (e = cudaMalloc(&s->dLlr, batchFrames * (size_t)N * (s->llrIsFloat ? 4 : 1))) != cudaSuccess) {
        polar_cuda_sim_destroy(s);
        cudaGetLastError();
        return fail(POLAR_E_NOMEM, std::string("simulator buffers: ") + cudaGetErrorString(e));
    }
    *out = s;
    return POLAR_OK;
}

int polar_cuda_sim_run(polar_sim_t s, double ebn0_db, double amplification, uint64_t seed, uint32_t point, uint64_t firstFrame,
                       size_t nFrames, unsigned long long* counters_dev, void* stream)
{
    if (!s) return fail(POLAR_E_ARG, "null simulator");
    if (!counters_dev) return fail(POLAR_E_ARG, "null counters");
    CUDA_TRY(cudaSetDevice(s->device));
    cudaStream_t st = (cudaStream_t)stream;
    SimParams p{};
    p.info = s->dInfo; p.code = s->dCode; p.llr = s->dLlr;
    p.seedLo = (uint32_t)seed; p.seedHi = (uint32_t)(seed >> 32) ^ (point * 0x9E3779B9u);
    // simulator.cpp:821-838: Es/N0 = Eb/N0 * K / N (one bit per symbol), awgn.cpp:30-43: sigma = 1 / sqrt(2 Es/N0)
    p.sigma = (float)(1.0 / std::sqrt(2.0 * std::pow(10.0, ebn0_db / 10.0) * (double)s->K / (double)s->N));
    p.amplification = (float)amplification;
    p.N = s->N; p.K = s->K; p.kb = s->kb; p.nb = s->nb; p.llrIsFloat = s->llrIsFloat;
    const int numSMs = s->enc->numSMs;
    for (size_t done = 0; done < nFrames; done += s->batch) {
        const size_t n = std::min(s->batch, nFrames - done);
        p.firstFrame = firstFrame + done;
        p.nFrames = (long long)n;
        const long long payloadThreads = (long long)n * ((s->kb + 15) / 16);
        sim_payload_kernel<<<(int)std::min<long long>((payloadThreads + 255) / 256, (long long)numSMs * 8), 256, 0, st>>>(p);
        CUDA_TRY(cudaGetLastError());
        int rc = polar_cuda_encode_batch(s->enc, s->dInfo, POLAR_MEM_DEVICE, n, s->dCode, st);
        if (rc) return rc;
        const long long chanThreads = (long long)n * (s->N / 4);
        sim_channel_kernel<<<(int)std::min<long long>((chanThreads + 255) / 256, (long long)numSMs * 16), 256, 0, st>>>(p);
        CUDA_TRY(cudaGetLastError());
        rc = polar_cuda_decode_batch(s->dec, s->dLlr, s->llrIsFloat ? POLAR_LLR_F32 : POLAR_LLR_I8, POLAR_MEM_DEVICE, n, s->dOut,
                                     s->dOk, st);
        if (rc) return rc;
        sim_count_kernel<<<(int)std::min<long long>(((long long)n + 255) / 256, (long long)numSMs * 4), 256, 0, st>>>(
            s->dInfo, s->dOut, s->dOk, (long long)n, s->kb, counters_dev);
        CUDA_TRY(cudaGetLastError());
        s->launches += 3;
        s->lastFrames = n;
    }
    return POLAR_OK;
}

int polar_cuda_sim_buffers(polar_sim_t s, void** llr, uint8_t** sent, uint8_t** decoded, uint8_t** ok, size_t* frames)
{
    if (!s) return fail(POLAR_E_ARG, "null simulator");
    if (llr) *llr = s->dLlr;
    if (sent) *sent = s->dInfo;
    if (decoded) *decoded = s->dOut;
    if (ok) *ok = s->dOk;
    if (frames) *frames = s->lastFrames;
    return s->llrIsFloat;
}

long polar_cuda_sim_launches(polar_sim_t s, long* decoderLaunches)
{
    if (!s) return 0;
    if (decoderLaunches) *decoderLaunches = polar_cuda_code_launches(s->dec);
    return s->launches + polar_cuda_code_launches(s->enc) + polar_cuda_code_launches(s->dec);
}

// Philox4x32-10 known-answer hook (tests): out = philox(ctr, key)
int polar_cuda_test_philox(const uint32_t* ctr, const uint32_t* key, uint32_t* out);

} // extern "C"

namespace {
__global__ void test_philox_kernel(uint4 c, uint2 k, uint4* out) { *out = philox4x32_10(c, k); }
} // namespace

extern "C" int polar_cuda_test_philox(const uint32_t* ctr, const uint32_t* key, uint32_t* out)
{
    uint4* d = nullptr;
    CUDA_TRY(cudaMalloc(&d, sizeof(uint4)));
    test_philox_kernel<<<1, 1>>>(make_uint4(ctr[0], ctr[1], ctr[2], ctr[3]), make_uint2(key[0], key[1]), d);
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaMemcpy(out, d, sizeof(uint4), cudaMemcpyDeviceToHost));
    cudaFree(d);
    return POLAR_OK;
}

// ---------------------------------------------------------------- primitive self-test
namespace {
__global__ void test_fg_kernel(const uint32_t* a, const uint32_t* b, const uint32_t* bits, int words, uint32_t* f,
                               uint32_t* g)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= words) return;
    f[i] = f4(a[i], b[i]);
    g[i] = g4(a[i], b[i], nibble_to_msb(sign_nibble(bits[i])));
}
} // namespace

extern "C" int polar_cuda_test_fg(const int8_t* a, const int8_t* b, const int8_t* bits, int n, int8_t* f_out,
                                  int8_t* g_out)
{
    if (n <= 0 || (n & 3)) return fail(POLAR_E_ARG, "n must be a positive multiple of 4");
    uint32_t* d[5] = {};
    for (auto& p : d) CUDA_TRY(cudaMalloc(&p, n));
    CUDA_TRY(cudaMemcpy(d[0], a, n, cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMemcpy(d[1], b, n, cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMemcpy(d[2], bits, n, cudaMemcpyHostToDevice));
    const int words = n / 4;
    test_fg_kernel<<<(words + 255) / 256, 256>>>(d[0], d[1], d[2], words, d[3], d[4]);
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaMemcpy(f_out, d[3], n, cudaMemcpyDeviceToHost));
    CUDA_TRY(cudaMemcpy(g_out, d[4], n, cudaMemcpyDeviceToHost));
    for (auto& p : d) cudaFree(p);
    return POLAR_OK;
}
