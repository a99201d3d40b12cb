/* libpolar_cuda -- C ABI of the B200 (sm_100a) batched polar-code engine.
 *
 * This is the drop-in boundary: the reference (david13pod/antPolarCodes) has no FFI of its own,
 * its hot path sits behind three C++ interfaces.  Each entry point below names the reference
 * interface it replaces; the C++ subclasses in antpolarcodes_b200/csrc/host/ (FastSscCudaChar,
 * SclCudaChar, ButterflyCudaPacked, ...) and the Python mirror call nothing but these.
 *
 *   Encoder::encode_vector(void* info, void* code)          src/polarcode/encoding/encoder.cpp:79-90
 *   Decoder::decode_vector(const char* or float* llr, void*) src/polarcode/decoding/decoder.cpp:154-181
 *   Detector::generate / check                              include/polarcode/errordetection/errordetector.h:24-70
 *   Construction::frozen_bits(N, K, dSNR, type)             src/polarcode/construction/constructor.cpp:41-63
 *
 * Conventions (same as the reference): LLR >= 0 <=> bit 0; every packed buffer is MSB first
 * (bit i in byte i/8 at mask 0x80 >> (i%8)); K = N - |frozen| includes the CRC bits; frozen
 * indices are sorted ascending.  All functions return 0 on success or a negative POLAR_E_* code;
 * polar_cuda_last_error() gives the message of the calling thread's last failure.
 */
#ifndef POLAR_CUDA_H
#define POLAR_CUDA_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct polar_code* polar_code_t;

enum polar_kind {
    POLAR_KIND_ENC = 0,         /* Encoding::ButterflyFipPacked   butterfly_fip_packed.h:25   */
    POLAR_KIND_FASTSSC_I8 = 1,  /* Decoding::FastSscFipChar       fastssc_fip_char.h:298      */
    POLAR_KIND_FASTSSC_F32 = 2, /* Decoding::FastSscAvxFloat      fastssc_avx_float.h:283     */
    POLAR_KIND_SCL_I8 = 3,      /* Decoding::SclFipChar           scl_fip_char.h:391          */
    POLAR_KIND_SCL_F32 = 4,     /* Decoding::SclAvxFloat          scl_avx_float.h:407         */
    /* adaptive decoders: Fast-SSC on every frame, the list decoder only on the frames whose Fast-SSC result
     * fails the detector (adaptive_char.cpp:34-46); listSize 1 = Fast-SSC only, like the reference */
    POLAR_KIND_ADAPTIVE_I8 = 5,    /* Decoding::AdaptiveChar   FastSscFipChar  + SclFipChar   adaptive_char.h:26  */
    POLAR_KIND_ADAPTIVE_F32 = 6,   /* Decoding::AdaptiveFloat  FastSscAvxFloat + SclAvxFloat  adaptive_float.h    */
    POLAR_KIND_ADAPTIVE_MIXED = 7  /* Decoding::AdaptiveMixed  FastSscFipChar  + SclAvxFloat  adaptive_mixed.h    */
};

enum polar_mem { POLAR_MEM_HOST = 0, POLAR_MEM_DEVICE = 1 };
enum polar_llr { POLAR_LLR_I8 = 0, POLAR_LLR_F32 = 1 };

enum polar_error {
    POLAR_OK = 0,
    POLAR_E_ARG = -1,      /* invalid argument (std::invalid_argument in the C++ layer) */
    POLAR_E_UNSUPPORTED = -2,
    POLAR_E_CUDA = -3,     /* CUDA runtime failure (std::runtime_error) */
    POLAR_E_NOMEM = -4,
    POLAR_E_RANGE = -5     /* std::out_of_range in the C++ layer (Puncturer, puncturer.cpp:51-54) */
};

const char* polar_cuda_last_error(void);
int polar_cuda_device_count(void);
int polar_cuda_version(void);

/* Compile a code for one device: builds the Fast-SSC / SCL node schedule, the info-position
 * LUT and the CRC syndrome table and uploads them.  Replaces the decoder / encoder constructors
 * (blockLength, [listSize,] frozenBits) + setSystematic + setErrorDetection.
 *   listSize : power of two in [1,32] (ignored for ENC / FASTSSC kinds)
 *   crcBits  : 0 (Dummy detector), 8, 16 or 32 (ErrorDetection::create(bits,"crc")) */
int polar_cuda_code_create(int device, int kind, int N, const uint32_t* frozen, int nFrozen, int listSize,
                           int systematic, int crcBits, polar_code_t* out);
int polar_cuda_code_destroy(polar_code_t code);

/* Decoder::decode_vector over nFrames independent frames.
 *   llr      : nFrames x N LLRs, int8 or float (float into an int8 decoder is quantised like
 *              CharContainer::insertLlr, bitcontainer.cpp:468-516)
 *   info_out : nFrames x ceil(K/8) bytes;  ok_out : nFrames bytes (the bool decode() returns)
 *   mem      : where llr / info_out / ok_out live (all three the same)
 *   stream   : cudaStream_t (0 = default stream).  HOST buffers are copied synchronously;
 *              DEVICE buffers are processed asynchronously on the stream (adaptive kinds wait once on the
 *              stream for the count of failed frames before they launch the list decoder).
 *              DEVICE llr must be 4-byte aligned (POLAR_E_ARG otherwise; 16-byte aligned rows take the 128-bit paths);
 *              DEVICE info_out / ok_out may have any alignment (16-byte packed stores when info_out allows).
 * Adaptive kinds: ADAPTIVE_I8 takes int8 or float LLRs (float is quantised once, both decoders see the int8
 * copy, like AdaptiveChar::setSignal, adaptive_char.cpp:60-64); ADAPTIVE_F32 and ADAPTIVE_MIXED take float. */
int polar_cuda_decode_batch(polar_code_t code, const void* llr, int llrType, int mem, size_t nFrames,
                            uint8_t* info_out, uint8_t* ok_out, void* stream);

/* Candidate hand-back for detectors the kernels do not fuse (anything but Dummy / CRC8 / CRC16 / CRC32, e.g. the reference's
 * cmac): the list decoders return the info blocks of ALL surviving paths in list order, so that the host runs
 * Detector::check over them exactly like extractBestPath does (scl_fip_char.cpp:816-855: first passing path wins, else
 * path 0 and decode() == false).  The handle must have been created with crcBits = 0 and must not be an adaptive kind.
 *   cand_out  : nFrames x listSize x ceil(K/8) bytes (Fast-SSC kinds: nFrames x ceil(K/8), one candidate)
 *   count_out : nFrames bytes, the number of candidates of the frame (paths alive at the end of the tree, <= listSize) */
int polar_cuda_decode_candidates(polar_code_t code, const void* llr, int llrType, int mem, size_t nFrames, uint8_t* cand_out,
                                 uint8_t* count_out, void* stream);

/* Decoder::getSoftCodeword / getSoftInformation (src/polarcode/decoding/decoder.cpp:147-152) after FastSscFipChar::decode, over
 * nFrames frames: the N byte partial sums ("soft bits", include/polarcode/decoding/fip_char.h:164-220) the root node leaves in
 * the decoder's bit container, byte for byte what the reference returns (sign = hard decision; R0 127, R1 the LLR, REP the
 * saturating sum, SPC the LLR with one byte negated, inner nodes the raw byte XOR of their halves), and the same bytes
 * gathered at the K information positions.  int8 Fast-SSC handles only (POLAR_E_UNSUPPORTED otherwise).
 *   llr : nFrames x N int8;  soft_out : nFrames x N int8;  soft_info_out : nFrames x K int8 or NULL;  mem as in decode_batch */
int polar_cuda_decode_soft_batch(polar_code_t code, const int8_t* llr, int mem, size_t nFrames, int8_t* soft_out,
                                 int8_t* soft_info_out, void* stream);

/* Allocation-free decode.  NO COUNTERPART IN THE REFERENCE: added for callers that cannot
 * call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same kernels and results as
 * polar_cuda_decode_batch with DEVICE buffers, but the kernel's scratch (the top LLR / partial-sum stages it keeps in L2)
 * lives in a file-scope static __device__ array of POLAR_NOALLOC_SCRATCH_BYTES instead of a cudaMalloc'ed buffer, and the
 * call allocates nothing: no cudaMalloc, cudaFree, staging copy or event creation.
 *   code     : a FASTSSC_I8 / FASTSSC_F32 / SCL_I8 / SCL_F32 handle (creating it is the set-up step and does allocate)
 *   llr      : DEVICE, nFrames x N, the decoder's own type (int8 for the int8 kinds, float for the fp32 kinds)
 *   info_out : DEVICE, nFrames x ceil(K/8);  ok_out : DEVICE, nFrames (required);  asynchronous on `stream`
 * Limit: the handle's launch geometry must need at most POLAR_NOALLOC_SCRATCH_BYTES of scratch (spill words per resident
 * warp x resident warps: e.g. Fast-SSC int8 N <= 1024 and SCL int8 L = 8 N <= 512 fit on a 148-SM B200, SCL L = 8 N = 1024
 * does not); larger codes are rejected with POLAR_E_UNSUPPORTED and a message that states both sizes.
 * Every GPU that loads the library holds the array until exit.  ALL calls on a device share it -- whatever the handle or the
 * stream -- and it keeps the last call's contents (the kernels write every scratch word before they read it, so nothing
 * has to be cleared): calls that overlap on the GPU race.  Serialise them (one stream, or events between streams). */
#define POLAR_NOALLOC_SCRATCH_BYTES (64u << 20)
int polar_cuda_decode_batch_noalloc(polar_code_t code, const void* llr, int llrType, size_t nFrames, uint8_t* info_out,
                                    uint8_t* ok_out, void* stream);

/* Encoder::encode_vector over nFrames frames.  info: nFrames x K/8 bytes; like the reference the
 * trailing CRC bytes of every info block are overwritten in place.  code_out: nFrames x N/8.
 * DEVICE buffers may have any alignment; the fast kernel (one thread per two frames, 128/256-bit row accesses) runs when
 * the info rows are 4-byte and the codeword rows min(N/8, 32)-byte aligned, which every cudaMalloc'ed [nFrames][...] array is. */
int polar_cuda_encode_batch(polar_code_t code, uint8_t* info, int mem, size_t nFrames, uint8_t* code_out,
                            void* stream);

/* Detector::generate / Detector::check on the host (same arithmetic the kernels fuse). */
int polar_cuda_crc_generate(int crcBits, uint8_t* data, int bytes);
int polar_cuda_crc_check(int crcBits, const uint8_t* data, int bytes); /* 1 = pass, 0 = fail */

/* Construction::frozen_bits: type "BB" (Bhattacharyya), "BE" (beta expansion) or "5G".
 * Writes N-K sorted indices to out; returns the count or a negative error. */
int polar_cuda_frozen_bits(int N, int K, float designSnr, const char* type, uint32_t* out);

/* Rate matching either side of the coder: PolarCode::Puncturer (include/polarcode/puncturer.h:29-101,
 * src/polarcode/puncturer.cpp:45-85).  parentBlockLength = next power of two >= blockLength; the first
 * (parent - blockLength) entries of the frozen set are punctured, every other position is kept in ascending order.
 *   positions : pure host; writes the blockLength kept positions (blockOutputPositions()), returns their count,
 *               POLAR_E_RANGE when more positions would have to be punctured than there are frozen bits
 *   puncture  : out[f][j] = in[f][pos[j]]                        in: nFrames x parent, out: nFrames x blockLength
 *   depuncture: out[f][pos[j]] = in[f][j], every other entry 0   in: nFrames x blockLength, out: nFrames x parent
 *   puncture_packed: the same gather on MSB-first packed bits    in: nFrames x parent/8, out: nFrames x blockLength/8
 * elemBytes = 1 (int8 / uint8), 4 (float) or 8 (double), the three types the pypolar binding instantiates
 * (puncturer_python.cc:45-145). */
typedef struct polar_puncturer* polar_puncturer_t;
int polar_cuda_puncturer_positions(int blockLength, const uint32_t* frozen, int nFrozen, uint32_t* out);
int polar_cuda_puncturer_create(int device, int blockLength, const uint32_t* frozen, int nFrozen, polar_puncturer_t* out);
int polar_cuda_puncturer_destroy(polar_puncturer_t p);
int polar_cuda_puncturer_parent_length(polar_puncturer_t p);
int polar_cuda_puncture_batch(polar_puncturer_t p, const void* in, int elemBytes, int mem, size_t nFrames, void* out,
                              void* stream);
int polar_cuda_depuncture_batch(polar_puncturer_t p, const void* in, int elemBytes, int mem, size_t nFrames, void* out,
                                void* stream);
int polar_cuda_puncture_packed_batch(polar_puncturer_t p, const uint8_t* in, int mem, size_t nFrames, uint8_t* out,
                                     void* stream);

/* On-device link simulator: the worker loop of the reference's pcsim (src/simulation/simulator.cpp:632-672) over batches of
 * frames, every stage a kernel -- generateData (:850-868), encode (:870-879), BPSK / AWGN / amplifier / quantiser
 * (:881-931, awgn.cpp:30-43, setChannel :821-841), decode (:919-935), countErrors (:937-964).  Nothing returns to the host
 * between the stages.  The random streams are Philox4x32-10 keyed by (seed, point) and counted by the ABSOLUTE frame number,
 * so frame f of a point is the same frame whatever the batch size or the number of GPUs sharing the point.
 *   decoderKind   : any decoder kind; int8 kinds receive int8 LLRs = quantise(amplification * y), fp32 kinds amplification * y
 *   amplification : pcsim's default is 10 for the 8-bit decoders (setup.cpp:58) and 4 * 10^(EbN0/10) for float (simulator.cpp:377)
 *   counters_dev  : DEVICE pointer to 5 x uint64, ADDED to: runs, block errors, bit errors (over all K info bits, CRC
 *                   included), reported errors (decode() == false), undetected errors (wrong block that passed the detector)
 * sim_run is asynchronous on `stream`; read the counters after synchronising it.  sim_buffers returns the device buffers of
 * the LAST batch (llr, sent info, decoded info, verdicts) for subsample checks; its return value is 1 if llr is float. */
typedef struct polar_sim* polar_sim_t;
int polar_cuda_sim_create(int device, int N, const uint32_t* frozen, int nFrozen, int decoderKind, int listSize, int systematic,
                          int crcBits, size_t batchFrames, polar_sim_t* out);
int polar_cuda_sim_destroy(polar_sim_t sim);
int polar_cuda_sim_run(polar_sim_t sim, double ebn0_db, double amplification, uint64_t seed, uint32_t point, uint64_t firstFrame,
                       size_t nFrames, unsigned long long* counters_dev, void* stream);
int polar_cuda_sim_buffers(polar_sim_t sim, void** llr, uint8_t** sent, uint8_t** decoded, uint8_t** ok, size_t* frames);
long polar_cuda_sim_launches(polar_sim_t sim, long* decoderLaunches);
int polar_cuda_test_philox(const uint32_t* ctr, const uint32_t* key, uint32_t* out); /* Philox4x32-10 known-answer hook */

/* Introspection for tests, bench and the roofline. */
struct polar_code_info {
    int kind, N, K, listSize, systematic, crcBits;
    int nOps;                 /* schedule length */
    int warpsPerBlock, blocksPerSM, smemPerWarp, gridBlocks;
    double laneOpsPerFrame;   /* algorithmic ALU lane-ops per frame */
    double bytesPerFrame;     /* algorithmic HBM bytes per frame */
    long opCount[32], opSize[32]; /* indexed by polar::OpCode */
};
int polar_cuda_code_info(polar_code_t code, struct polar_code_info* info);
int polar_cuda_code_schedule(polar_code_t code, uint32_t* ops, int maxOps); /* returns op count */
/* Pure-host schedule compiler (no device needed): node histogram of the program the kernels would run,
 * i.e. of the tree createDecoder() builds (fastssc_fip_char.cpp:496-569, scl_fip_char.cpp:729-755).
 * counts / sizes: 32 entries indexed by polar::OpCode; returns the op count or a negative error. */
int polar_cuda_schedule_stats(int kind, int N, const uint32_t* frozen, int nFrozen, int listSize, long* counts,
                              long* sizes, double* laneOpsPerFrame);
/* Compiled schedules are cached per process, keyed by (tree flavour, N, frozen set, list size): re-creating a code
 * after setSystematic / setErrorDetection, or one decoder object per host thread, compiles the node program once
 * (the reference rebuilds its node tree in every initialize(), fastssc_fip_char.cpp:592-614).  Returns the number
 * of cached entries; hits / misses count lookups since the library was loaded. */
int polar_cuda_schedule_cache_stats(long* hits, long* misses);
/* number of kernel launches issued through this handle so far (bench's gpu_launches) */
long polar_cuda_code_launches(polar_code_t code);
/* adaptive kinds: frames the list decoder had to redo, summed over all decode calls on this handle so far
 * (the `!success && mListSize > 1` branch of AdaptiveChar::decode, adaptive_char.cpp:39); 0 for other kinds */
long polar_cuda_code_fallbacks(polar_code_t code);

/* Primitive self-test hooks (golden vectors of decodingtest.cpp:504-737): element-wise F / G over
 * n int8 values on the device, through the very device functions the decoders use. */
int polar_cuda_test_fg(const int8_t* a, const int8_t* b, const int8_t* bits, int n, int8_t* f_out,
                       int8_t* g_out);

#ifdef __cplusplus
}
#endif
#endif
