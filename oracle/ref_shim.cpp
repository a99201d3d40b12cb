// TEST INFRASTRUCTURE ONLY -- not part of the product path.
//
// extern "C" shim over the UNMODIFIED reference library (david13pod/antPolarCodes),
// compiled from the sources where they lie under /root/reference by oracle/Makefile
// into oracle/_ref/libpolar_ref.so.  It lets tests/ and bench.py's cpu_baseline /
// --impl reference legs drive the reference's own classes:
//   PolarCode::Construction::frozen_bits      (src/polarcode/construction/constructor.cpp:41-63)
//   PolarCode::Encoding::ButterflyFipPacked   (src/polarcode/encoding/butterfly_fip_packed.cpp:45-59)
//   PolarCode::Decoding::FastSscFipChar       (src/polarcode/decoding/fastssc_fip_char.cpp:616-630)
//   PolarCode::Decoding::FastSscAvxFloat      (src/polarcode/decoding/fastssc_avx_float.cpp:940-955)
//   PolarCode::Decoding::SclFipChar           (src/polarcode/decoding/scl_fip_char.cpp:801-855)
//   PolarCode::Decoding::SclAvxFloat          (src/polarcode/decoding/scl_avx_float.cpp:696-750)
//   PolarCode::Decoding::AdaptiveChar / AdaptiveFloat / AdaptiveMixed  (src/polarcode/decoding/adaptive_char.cpp:34-46)
//   PolarCode::ErrorDetection::create         (src/polarcode/errordetection/errordetector.cpp:23-67)
// No reference source is copied here; this file only calls the public API.
#include <polarcode/bitcontainer.h>
#include <polarcode/construction/constructor.h>
#include <polarcode/decoding/adaptive_char.h>
#include <polarcode/decoding/adaptive_float.h>
#include <polarcode/decoding/adaptive_mixed.h>
#include <polarcode/decoding/decoder.h>
#include <polarcode/decoding/fastssc_avx_float.h>
#include <polarcode/decoding/fastssc_fip_char.h>
#include <polarcode/decoding/scl_avx_float.h>
#include <polarcode/decoding/scl_fip_char.h>
#include <polarcode/encoding/butterfly_fip_packed.h>
#include <polarcode/errordetection/errordetector.h>
#include <polarcode/puncturer.h>

#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

using namespace PolarCode;

namespace {

// kind values shared with include/polar_cuda.h (POLAR_KIND_*)
enum { KIND_ENC = 0, KIND_FASTSSC_I8 = 1, KIND_FASTSSC_F32 = 2, KIND_SCL_I8 = 3, KIND_SCL_F32 = 4,
       KIND_ADAPTIVE_I8 = 5, KIND_ADAPTIVE_F32 = 6, KIND_ADAPTIVE_MIXED = 7 };

struct RefCode {
    int kind = 0;
    size_t N = 0, K = 0, L = 1;
    int systematic = 1, crcBits = 0;
    std::vector<unsigned> frozen;
};

Decoding::Decoder* make_decoder(const RefCode& c)
{
    Decoding::Decoder* d = nullptr;
    switch (c.kind) {
    case KIND_FASTSSC_I8: d = new Decoding::FastSscFipChar(c.N, c.frozen); break;
    case KIND_FASTSSC_F32: d = new Decoding::FastSscAvxFloat(c.N, c.frozen); break;
    case KIND_SCL_I8: d = new Decoding::SclFipChar(c.N, c.L, c.frozen); break;
    case KIND_SCL_F32: d = new Decoding::SclAvxFloat(c.N, c.L, c.frozen); break;
    case KIND_ADAPTIVE_I8: d = new Decoding::AdaptiveChar(c.N, c.L, c.frozen); break;
    case KIND_ADAPTIVE_F32: d = new Decoding::AdaptiveFloat(c.N, c.L, c.frozen); break;
    case KIND_ADAPTIVE_MIXED: d = new Decoding::AdaptiveMixed(c.N, c.L, c.frozen); break;
    default: return nullptr;
    }
    d->setSystematic(c.systematic != 0);
    d->setErrorDetection(ErrorDetection::create(c.crcBits, "crc"));
    return d;
}

Encoding::Encoder* make_encoder(const RefCode& c)
{
    Encoding::Encoder* e = new Encoding::ButterflyFipPacked(c.N, c.frozen);
    e->setSystematic(c.systematic != 0);
    e->setErrorDetection(ErrorDetection::create(c.crcBits, "crc"));
    return e;
}

} // namespace

extern "C" {

// returns number of frozen indices written (N-K) or -1 on exception
int polar_ref_frozen_bits(int N, int K, float designSnr, const char* type, uint32_t* out)
{
    try {
        std::vector<unsigned> f = Construction::frozen_bits(N, K, designSnr, std::string(type));
        for (size_t i = 0; i < f.size(); ++i) out[i] = f[i];
        return (int)f.size();
    } catch (...) {
        return -1;
    }
}

void* polar_ref_code_create(int kind, int N, const uint32_t* frozen, int nFrozen, int listSize,
                            int systematic, int crcBits)
{
    RefCode* c = new RefCode;
    c->kind = kind;
    c->N = N;
    c->K = N - nFrozen;
    c->L = listSize;
    c->systematic = systematic;
    c->crcBits = crcBits;
    c->frozen.assign(frozen, frozen + nFrozen);
    return c;
}

void polar_ref_code_destroy(void* h) { delete static_cast<RefCode*>(h); }

// info: nFrames x K/8 bytes (CRC bytes are overwritten in place, like the reference does);
// code: nFrames x N/8 bytes.
int polar_ref_encode(void* h, uint8_t* info, uint8_t* code, long nFrames)
{
    RefCode* c = static_cast<RefCode*>(h);
    try {
        Encoding::Encoder* e = make_encoder(*c);
        const size_t kb = (c->K + 7) / 8, nb = c->N / 8 ? c->N / 8 : 1;
        for (long f = 0; f < nFrames; ++f) e->encode_vector(info + f * kb, code + f * nb);
        delete e;
        return 0;
    } catch (...) {
        return -1;
    }
}

// Decode nFrames with nThreads worker threads (one decoder object per thread, disjoint
// contiguous slices).  llr is int8 for *_I8 kinds, float for *_F32 kinds and for all three adaptive kinds (the
// adaptive classes only override setSignal(const float*), adaptive_char.cpp:60-64).  repeat > 1
// re-decodes the same pool (for timing).  Returns elapsed seconds (decode loop only) or < 0.
double polar_ref_decode(void* h, const void* llr, long nFrames, uint8_t* out, uint8_t* ok,
                        int nThreads, int repeat)
{
    RefCode* c = static_cast<RefCode*>(h);
    if (nThreads < 1) nThreads = 1;
    if (repeat < 1) repeat = 1;
    const size_t kb = (c->K + 7) / 8;
    const bool isFloat = (c->kind == KIND_FASTSSC_F32 || c->kind == KIND_SCL_F32 || c->kind >= KIND_ADAPTIVE_I8);
    std::vector<Decoding::Decoder*> decs(nThreads, nullptr);
    try {
        for (int t = 0; t < nThreads; ++t) decs[t] = make_decoder(*c);
    } catch (...) {
        return -1.0;
    }
    std::atomic<int> failed{ 0 };
    auto worker = [&](int t) {
        long lo = nFrames * t / nThreads, hi = nFrames * (t + 1) / nThreads;
        Decoding::Decoder* d = decs[t];
        try {
            for (int r = 0; r < repeat; ++r) {
                for (long f = lo; f < hi; ++f) {
                    bool res;
                    if (isFloat)
                        res = d->decode_vector(static_cast<const float*>(llr) + f * c->N, out + f * kb);
                    else
                        res = d->decode_vector(static_cast<const char*>(llr) + f * c->N, out + f * kb);
                    if (ok) ok[f] = res ? 1 : 0;
                }
            }
        } catch (...) {
            failed = 1;
        }
    };
    auto t0 = std::chrono::steady_clock::now();
    if (nThreads == 1) {
        worker(0);
    } else {
        std::vector<std::thread> th;
        for (int t = 0; t < nThreads; ++t) th.emplace_back(worker, t);
        for (auto& x : th) x.join();
    }
    auto t1 = std::chrono::steady_clock::now();
    for (auto d : decs) delete d;
    if (failed) return -2.0;
    return std::chrono::duration<double>(t1 - t0).count();
}

// polar_ref_decode with two more knobs:
//   llrIsFloat : 1 = float LLRs whatever the kind -- the int8 decoders then quantise them themselves, which is what
//                pcbench times (Decoder::decode_vector(const float*), benchmark/main_benchmark.cpp:202-241); 0 = int8;
//                -1 = by kind, like polar_ref_decode
//   fresh      : 1 = a NEW decoder object for every frame (SclAvxFloat keeps the winning metric of the previous
//                frame, scl_avx_float.cpp:46-56,102-108)
double polar_ref_decode_ex(void* h, const void* llr, int llrIsFloat, long nFrames, uint8_t* out, uint8_t* ok,
                           int nThreads, int repeat, int fresh)
{
    RefCode* c = static_cast<RefCode*>(h);
    if (nThreads < 1) nThreads = 1;
    if (repeat < 1) repeat = 1;
    const size_t kb = (c->K + 7) / 8;
    bool isFloat = (c->kind == KIND_FASTSSC_F32 || c->kind == KIND_SCL_F32 || c->kind >= KIND_ADAPTIVE_I8);
    if (llrIsFloat >= 0) isFloat = llrIsFloat != 0;
    std::atomic<int> failed{ 0 };
    auto worker = [&](int t) {
        long lo = nFrames * t / nThreads, hi = nFrames * (t + 1) / nThreads;
        Decoding::Decoder* d = nullptr;
        try {
            if (!fresh) d = make_decoder(*c);
            for (int r = 0; r < repeat; ++r) {
                for (long f = lo; f < hi; ++f) {
                    if (fresh) d = make_decoder(*c);
                    bool res;
                    if (isFloat)
                        res = d->decode_vector(static_cast<const float*>(llr) + f * c->N, out + f * kb);
                    else
                        res = d->decode_vector(static_cast<const char*>(llr) + f * c->N, out + f * kb);
                    if (ok) ok[f] = res ? 1 : 0;
                    if (fresh) { delete d; d = nullptr; }
                }
            }
            delete d;
        } catch (...) {
            failed = 1;
        }
    };
    auto t0 = std::chrono::steady_clock::now();
    if (nThreads == 1) {
        worker(0);
    } else {
        std::vector<std::thread> th;
        for (int t = 0; t < nThreads; ++t) th.emplace_back(worker, t);
        for (auto& x : th) x.join();
    }
    auto t1 = std::chrono::steady_clock::now();
    if (failed) return -2.0;
    return std::chrono::duration<double>(t1 - t0).count();
}

// mode 0: generate (writes trailing check bytes in place), returns 0
// mode 1: check, returns 1/0
int polar_ref_crc(int crcBits, uint8_t* data, int bytes, int mode)
{
    try {
        ErrorDetection::Detector* d = ErrorDetection::create(crcBits, "crc");
        int r = 0;
        if (mode == 0)
            d->generate(data, bytes);
        else
            r = d->check(data, bytes) ? 1 : 0;
        delete d;
        return r;
    } catch (...) {
        return -1;
    }
}

// CharContainer::insertLlr(const float*) (bitcontainer.cpp:505-516): the float -> int8 quantiser
void polar_ref_quantize(const float* llr, int8_t* out, int N)
{
    CharContainer c(N);
    c.insertLlr(llr);
    c.getSoftBits(out);
}

// PolarCode::Puncturer (src/polarcode/puncturer.cpp:45-85, include/polarcode/puncturer.h:71-100).
// mode 0: positions -> out (uint32, returns count); 1: puncture<T>; 2: depuncture<T>; 3: puncturePacked.  -1 = threw.
int polar_ref_puncturer(int blockLength, const uint32_t* frozen, int nFrozen, int mode, int elem, const void* in, void* out)
{
    try {
        Puncturer p(blockLength, std::vector<unsigned>(frozen, frozen + nFrozen));
        if (mode == 0) {
            std::vector<unsigned> pos = p.blockOutputPositions();
            for (size_t i = 0; i < pos.size(); ++i) static_cast<uint32_t*>(out)[i] = pos[i];
            return (int)pos.size();
        }
        if (mode == 3) { p.puncturePacked(static_cast<unsigned char*>(out), static_cast<const unsigned char*>(in)); return 0; }
        if (elem == 1) {
            if (mode == 1) p.puncture(static_cast<uint8_t*>(out), static_cast<const uint8_t*>(in));
            else p.depuncture(static_cast<uint8_t*>(out), static_cast<const uint8_t*>(in));
        } else if (elem == 4) {
            if (mode == 1) p.puncture(static_cast<float*>(out), static_cast<const float*>(in));
            else p.depuncture(static_cast<float*>(out), static_cast<const float*>(in));
        } else {
            if (mode == 1) p.puncture(static_cast<double*>(out), static_cast<const double*>(in));
            else p.depuncture(static_cast<double*>(out), static_cast<const double*>(in));
        }
        return (int)p.parentBlockLength();
    } catch (...) {
        return -1;
    }
}

// Decoder::getSoftCodeword / getSoftInformation (decoder.cpp:147-152) after decode_vector, for any decoder kind that
// takes the given LLR type.  soft: nFrames x N elements, softInfo: nFrames x K elements (int8 for the char decoders,
// float for the float decoders); either may be null.
int polar_ref_soft(void* h, const void* llr, long nFrames, void* soft, void* softInfo)
{
    RefCode* c = static_cast<RefCode*>(h);
    const bool isFloat = (c->kind == KIND_FASTSSC_F32 || c->kind == KIND_SCL_F32);
    const size_t es = isFloat ? 4 : 1, kb = (c->K + 7) / 8;
    try {
        Decoding::Decoder* d = make_decoder(*c);
        std::vector<uint8_t> out(kb);
        for (long f = 0; f < nFrames; ++f) {
            if (isFloat) d->decode_vector(static_cast<const float*>(llr) + f * c->N, out.data());
            else d->decode_vector(static_cast<const char*>(llr) + f * c->N, out.data());
            if (soft) d->getSoftCodeword(static_cast<char*>(soft) + f * c->N * es);
            if (softInfo) d->getSoftInformation(static_cast<char*>(softInfo) + f * c->K * es);
        }
        delete d;
        return 0;
    } catch (...) {
        return -1;
    }
}

int polar_ref_hw_threads() { return (int)std::thread::hardware_concurrency(); }

} // extern "C"
