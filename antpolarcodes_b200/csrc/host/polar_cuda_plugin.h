// Drop-in subclasses of the reference's Decoder / Encoder interfaces on top of the libpolar_cuda C ABI.
//
// This header is the code a maintainer of david13pod/antPolarCodes adds to the reference tree (see
// INTEGRATION.md): it is compiled against the reference's OWN headers (include/polarcode/...) and calls nothing
// but include/polar_cuda.h.  Nothing here restates an algorithm; the classes only adapt the call conventions:
//
//   PolarCode::Decoding::FastSscCudaChar   next to FastSscFipChar   (fastssc_fip_char.h:298)
//   PolarCode::Decoding::FastSscCudaFloat  next to FastSscAvxFloat  (fastssc_avx_float.h:283)
//   PolarCode::Decoding::SclCudaChar       next to SclFipChar       (scl_fip_char.h:391)
//   PolarCode::Decoding::SclCudaFloat      next to SclAvxFloat      (scl_avx_float.h:407)
//   PolarCode::Decoding::AdaptiveCudaChar / AdaptiveCudaFloat / AdaptiveCudaMixed
//                                          next to AdaptiveChar / AdaptiveFloat / AdaptiveMixed (adaptive_char.h:26)
//   PolarCode::Decoding::makeCudaDecoder   like makeDecoder         (decoder.cpp:50-87)
//   PolarCode::Encoding::ButterflyCudaPacked next to ButterflyFipPacked (butterfly_fip_packed.h:25)
//   PolarCode::CudaPuncturer               a Puncturer (puncturer.h:29) with batched device methods
//
// The one-frame virtual API (decode_vector / decode / encode_vector / encode) is served by a batch of one --
// correct, and as slow as a kernel launch; throughput comes from decode_batch / encode_batch, the entry points
// this header adds.  The base-class setters (setSystematic, setErrorDetection) stay usable at any time: the
// device-side code handle is rebuilt lazily when the configuration it was compiled for has changed.
#pragma once
#include <polar_cuda.h>
#include <polarcode/bitcontainer.h>
#include <polarcode/decoding/decoder.h>
#include <polarcode/encoding/encoder.h>
#include <polarcode/errordetection/crc8.h>
#include <polarcode/errordetection/errordetector.h>
#include <polarcode/puncturer.h>
#include <cstring>
#include <functional>
#include <stdexcept>
#include <string>
#include <vector>

namespace PolarCode {
namespace CudaDetail {

inline void check(int rc)
{
    if (rc >= POLAR_OK) return;
    const std::string msg = polar_cuda_last_error();
    if (rc == POLAR_E_RANGE) throw std::out_of_range(msg); // Puncturer's constructor (puncturer.cpp:51-54)
    if (rc == POLAR_E_ARG) throw std::invalid_argument(msg); // the reference's convention for bad parameters
    throw std::runtime_error(msg);
}

// CRC width the kernels fuse for this detector: Dummy -> 0, CRC8/16/32 -> 8/16/32.  Anything else (the reference's cmac,
// a user's own Detector subclass) returns -1: the decoders then hand all surviving candidates back and the detector
// runs on the host, in list order, exactly like extractBestPath (scl_fip_char.cpp:816-855).
inline int fusedCrcBits(ErrorDetection::Detector* d)
{
    const std::string type = d->getType();
    const unsigned bits = d->getCheckBitCount();
    if (bits == 0 && type == "DUMMY") return 0;
    if (type == "CRC" && (bits == 8 || bits == 16 || bits == 32)) return (int)bits;
    return -1;
}

// Decoder::getSoftCodeword / getSoftInformation read mBitContainer (decoder.cpp:147-152).  The kernels keep partial sums
// as packed sign bits and never fill a byte / float container, so a CUDA decoder's bit container refuses instead of
// returning stale memory.
// int8 Fast-SSC: the byte soft bits are produced on demand by polar_cuda_decode_soft_batch for the frame decode() saw last
// (the decoder hands its LLR snapshot over through `fill`), then served by the CharContainer accessors as usual.
class SoftOnDemandChar : public CharContainer
{
public:
    std::function<void(CharContainer&)> fill;
    bool stale = false;
    SoftOnDemandChar(size_t size, const std::vector<unsigned>& frozen) : CharContainer(size, frozen) {}
    void refresh()
    {
        if (stale && fill) { fill(*this); stale = false; }
    }
    void getSoftBits(void* pData) override { refresh(); CharContainer::getSoftBits(pData); }
    void getSoftInformation(void* pData) override { refresh(); CharContainer::getSoftInformation(pData); }
    void getPackedBits(void* pData) override { refresh(); CharContainer::getPackedBits(pData); }
    void getPackedInformationBits(void* pData) override { refresh(); CharContainer::getPackedInformationBits(pData); }
};

template <class Base>
class NoSoftBits : public Base
{
public:
    using Base::Base;
    void getSoftBits(void*) override { throw std::logic_error("polar_cuda: soft outputs are not produced by the GPU decoders"); }
    void getSoftInformation(void*) override { throw std::logic_error("polar_cuda: soft outputs are not produced by the GPU decoders"); }
};

// A compiled code handle plus the configuration it was compiled for.
struct Handle {
    polar_code_t code = nullptr;
    int sys = -1, crc = -1;
    ~Handle() { if (code) polar_cuda_code_destroy(code); }
    bool hostDetector() const { return crc < 0; }
    polar_code_t get(int device, int kind, size_t N, const std::vector<unsigned>& frozen, size_t L, bool systematic,
                     ErrorDetection::Detector* det)
    {
        const int crcBits = fusedCrcBits(det);
        if (code && sys == (int)systematic && crc == crcBits) return code;
        if (code) { polar_cuda_code_destroy(code); code = nullptr; }
        if (crcBits < 0 && (kind == POLAR_KIND_ENC || kind >= POLAR_KIND_ADAPTIVE_I8))
            throw std::runtime_error("polar_cuda: error detector " + det->getType() + "-" + std::to_string(det->getCheckBitCount()) +
                                     " is not fused on the GPU (encoders and adaptive decoders need CRC8/16/32 or none)");
        check(polar_cuda_code_create(device, kind, (int)N, frozen.data(), (int)frozen.size(), (int)L, systematic ? 1 : 0,
                                     crcBits < 0 ? 0 : crcBits, &code));
        sys = (int)systematic;
        crc = crcBits;
        return code;
    }
};

} // namespace CudaDetail

namespace Decoding {

// Common part of the four decoders.  FLOAT selects the arithmetic (and with it the input container type, so
// that the inherited setSignal() overloads convert exactly like the CPU decoder of the same type does:
// CharContainer::insertLlr quantises floats, FloatContainer::insertLlr widens chars).
template <bool FLOAT>
class CudaDecoder : public Decoder
{
    CudaDetail::Handle mHandle;
    size_t mListSize;
    int mDevice;
    int mAdaptiveKind; // POLAR_KIND_ADAPTIVE_* or -1
    std::vector<char> mLastLlr; // int8 Fast-SSC: the frame decode() saw last (soft outputs are computed from it on demand)

    int kind() const
    {
        if (mAdaptiveKind >= 0) return mAdaptiveKind;
        if (mListSize > 1) return FLOAT ? POLAR_KIND_SCL_F32 : POLAR_KIND_SCL_I8;
        return FLOAT ? POLAR_KIND_FASTSSC_F32 : POLAR_KIND_FASTSSC_I8;
    }
    polar_code_t code()
    {
        return mHandle.get(mDevice, kind(), mBlockLength, mFrozenBits, mListSize, mSystematic, mErrorDetector);
    }

public:
    CudaDecoder(size_t blockLength, size_t listSize, const std::vector<unsigned>& frozenBits, int device = 0,
                int adaptiveKind = -1)
        : mListSize(listSize), mDevice(device), mAdaptiveKind(adaptiveKind)
    {
        initialize(blockLength, frozenBits);
    }
    ~CudaDecoder() override {} // the base class deletes the containers

    void initialize(size_t blockLength, const std::vector<unsigned>& frozenBits) override
    {
        Decoder::initialize(blockLength, frozenBits);
        delete mLlrContainer;
        delete mBitContainer;
        delete[] mOutputContainer;
        if (FLOAT) {
            mLlrContainer = new FloatContainer(blockLength);
            mBitContainer = new CudaDetail::NoSoftBits<FloatContainer>(blockLength, mFrozenBits);
        } else if (mListSize == 1 && mAdaptiveKind < 0) { // FastSscCudaChar: getSoftCodeword / getSoftInformation work
            mLlrContainer = new CharContainer(blockLength);
            auto* soft = new CudaDetail::SoftOnDemandChar(blockLength, mFrozenBits);
            soft->fill = [this](CharContainer& dst) {
                std::vector<char> bytes(mBlockLength);
                CudaDetail::check(polar_cuda_decode_soft_batch(code(), reinterpret_cast<const int8_t*>(mLastLlr.data()), POLAR_MEM_HOST, 1,
                                                               reinterpret_cast<int8_t*>(bytes.data()), nullptr, nullptr));
                dst.insertLlr(bytes.data()); // N raw bytes into the container's data
            };
            mBitContainer = soft;
        } else {
            mLlrContainer = new CharContainer(blockLength);
            mBitContainer = new CudaDetail::NoSoftBits<CharContainer>(blockLength, mFrozenBits);
        }
        mOutputContainer = new unsigned char[(blockLength - frozenBits.size() + 7) / 8];
        mHandle.sys = -1; // force a rebuild
    }

    size_t getListSize() override { return mListSize; }

    // A detector the kernels do not fuse: all candidates of every frame come back, Detector::check runs here in list
    // order; the first passing candidate wins, else candidate 0 and verdict false (scl_fip_char.cpp:816-855).
    void decodeWithHostDetector(const void* llr, int llrType, size_t nFrames, unsigned char* info, unsigned char* ok)
    {
        const size_t kb = (mBlockLength - mFrozenBits.size() + 7) / 8, L = mListSize > 1 ? mListSize : 1;
        std::vector<unsigned char> cand(nFrames * L * kb), count(nFrames);
        CudaDetail::check(polar_cuda_decode_candidates(code(), llr, llrType, POLAR_MEM_HOST, nFrames, cand.data(), count.data(), nullptr));
        for (size_t f = 0; f < nFrames; ++f) {
            unsigned char* c0 = cand.data() + f * L * kb;
            size_t pick = 0;
            bool pass = false;
            for (size_t path = 0; path < count[f] && !pass; ++path)
                if (mErrorDetector->check(c0 + path * kb, (int)kb)) { pick = path; pass = true; }
            memcpy(info + f * kb, c0 + pick * kb, kb);
            if (ok) ok[f] = pass ? 1 : 0;
        }
    }

    // Decoder::decode(): the frame setSignal() left in mLlrContainer, a batch of one.
    bool decode() override
    {
        uint8_t ok = 0;
        const void* llr = FLOAT ? static_cast<const void*>(static_cast<FloatContainer*>(mLlrContainer)->data())
                                : static_cast<const void*>(static_cast<CharContainer*>(mLlrContainer)->data());
        polar_code_t h = code();
        if (auto* soft = dynamic_cast<CudaDetail::SoftOnDemandChar*>(mBitContainer)) {
            mLastLlr.assign(static_cast<const char*>(llr), static_cast<const char*>(llr) + mBlockLength);
            soft->stale = true;
        }
        if (mHandle.hostDetector()) {
            decodeWithHostDetector(llr, FLOAT ? POLAR_LLR_F32 : POLAR_LLR_I8, 1, mOutputContainer, &ok);
            return ok != 0;
        }
        CudaDetail::check(polar_cuda_decode_batch(h, llr, FLOAT ? POLAR_LLR_F32 : POLAR_LLR_I8, POLAR_MEM_HOST, 1,
                                                  mOutputContainer, &ok, nullptr));
        return ok != 0;
    }

    // adaptive kinds: frames the list decoder had to redo so far
    long fallbacks() { return polar_cuda_code_fallbacks(code()); }

    // nFrames x blockLength() LLRs in, nFrames x ceil(infoLength()/8) bytes and nFrames verdicts out.
    // Host buffers: returns when the results are in place.  Device buffers: asynchronous on `stream`.
    void decode_batch(const char* llr, size_t nFrames, unsigned char* info, unsigned char* ok, bool deviceBuffers = false,
                      void* stream = nullptr)
    {
        if (FLOAT) throw std::invalid_argument("a float decoder takes float LLRs");
        polar_code_t h = code();
        if (mHandle.hostDetector()) {
            if (deviceBuffers) throw std::invalid_argument("a host-side detector needs host buffers");
            return decodeWithHostDetector(llr, POLAR_LLR_I8, nFrames, info, ok);
        }
        CudaDetail::check(polar_cuda_decode_batch(h, llr, POLAR_LLR_I8, deviceBuffers ? POLAR_MEM_DEVICE : POLAR_MEM_HOST,
                                                  nFrames, info, ok, stream));
    }
    // float LLRs into a char decoder are quantised on the device like CharContainer::insertLlr(const float*)
    void decode_batch(const float* llr, size_t nFrames, unsigned char* info, unsigned char* ok, bool deviceBuffers = false,
                      void* stream = nullptr)
    {
        polar_code_t h = code();
        if (mHandle.hostDetector()) {
            if (deviceBuffers) throw std::invalid_argument("a host-side detector needs host buffers");
            return decodeWithHostDetector(llr, POLAR_LLR_F32, nFrames, info, ok);
        }
        CudaDetail::check(polar_cuda_decode_batch(h, llr, POLAR_LLR_F32, deviceBuffers ? POLAR_MEM_DEVICE : POLAR_MEM_HOST,
                                                  nFrames, info, ok, stream));
    }
};

class FastSscCudaChar : public CudaDecoder<false>
{
public:
    FastSscCudaChar(size_t blockLength, const std::vector<unsigned>& frozenBits, int device = 0)
        : CudaDecoder<false>(blockLength, 1, frozenBits, device) {}
};
class FastSscCudaFloat : public CudaDecoder<true>
{
public:
    FastSscCudaFloat(size_t blockLength, const std::vector<unsigned>& frozenBits, int device = 0)
        : CudaDecoder<true>(blockLength, 1, frozenBits, device) {}
};
class SclCudaChar : public CudaDecoder<false>
{
public:
    SclCudaChar(size_t blockLength, size_t listSize, const std::vector<unsigned>& frozenBits, int device = 0)
        : CudaDecoder<false>(blockLength, listSize, frozenBits, device) {}
};
class SclCudaFloat : public CudaDecoder<true>
{
public:
    SclCudaFloat(size_t blockLength, size_t listSize, const std::vector<unsigned>& frozenBits, int device = 0)
        : CudaDecoder<true>(blockLength, listSize, frozenBits, device) {}
};

// Adaptive decoders (adaptive_char.cpp:34-46): Fast-SSC on every frame, the list decoder on the detector failures.
// The float container of the Mixed flavour keeps the LLRs as floats: the int8 Fast-SSC pass quantises them on the
// device (like FastSscFipChar::setSignal would), the fp32 list decoder reads them as they are.
class AdaptiveCudaChar : public CudaDecoder<false>
{
public:
    AdaptiveCudaChar(size_t blockLength, size_t listSize, const std::vector<unsigned>& frozenBits, int device = 0)
        : CudaDecoder<false>(blockLength, listSize, frozenBits, device, POLAR_KIND_ADAPTIVE_I8) {}
};
class AdaptiveCudaFloat : public CudaDecoder<true>
{
public:
    AdaptiveCudaFloat(size_t blockLength, size_t listSize, const std::vector<unsigned>& frozenBits, int device = 0)
        : CudaDecoder<true>(blockLength, listSize, frozenBits, device, POLAR_KIND_ADAPTIVE_F32) {}
};
class AdaptiveCudaMixed : public CudaDecoder<true>
{
public:
    AdaptiveCudaMixed(size_t blockLength, size_t listSize, const std::vector<unsigned>& frozenBits, int device = 0)
        : CudaDecoder<true>(blockLength, listSize, frozenBits, device, POLAR_KIND_ADAPTIVE_MIXED) {}
};

// makeDecoder (decoder.cpp:50-87) for the GPU types: decoder_impl 0 = char, 1 = float, 2 = adaptive float (what
// Decoding::create maps "mixed" to); like the CPU factory a fresh decoder checks CRC8.
inline Decoder* makeCudaDecoder(size_t blockLength, size_t listSize, const std::vector<unsigned>& frozenBits,
                                int decoder_impl = 0, int device = 0)
{
    Decoder* dec;
    if (listSize == 1) {
        if (decoder_impl == 1) dec = new FastSscCudaFloat(blockLength, frozenBits, device);
        else dec = new FastSscCudaChar(blockLength, frozenBits, device);
    } else {
        if (decoder_impl == 1) dec = new SclCudaFloat(blockLength, listSize, frozenBits, device);
        else if (decoder_impl == 2) dec = new AdaptiveCudaFloat(blockLength, listSize, frozenBits, device);
        else dec = new SclCudaChar(blockLength, listSize, frozenBits, device);
    }
    dec->setErrorDetection(new ErrorDetection::CRC8());
    return dec;
}

// Decoding::create (decoder.cpp:26-48) with the GPU types added: "cuda-char", "cuda-float", "cuda-mixed" select the classes
// above, every other string goes to the reference's own factory.  This is the function the pybind11 module binds as
// PolarDecoder's constructor.
inline Decoder* createWithCuda(size_t blockLength, size_t listSize, const std::vector<unsigned>& frozenBits,
                               std::string decoderType)
{
    if (decoderType.rfind("cuda", 0) != 0) return create(blockLength, listSize, frozenBits, decoderType);
    int impl;
    if (decoderType.find("char") != std::string::npos) impl = 0;
    else if (decoderType.find("float") != std::string::npos) impl = 1;
    else if (decoderType.find("mixed") != std::string::npos) impl = 2;
    else throw std::invalid_argument("Unknown PolarDecoder type!");
    if (listSize < 2) listSize = 1;
    return makeCudaDecoder(blockLength, listSize, frozenBits, impl);
}

} // namespace Decoding

namespace Encoding {

class ButterflyCudaPacked : public Encoder
{
    CudaDetail::Handle mHandle;
    int mDevice;

    polar_code_t code()
    {
        return mHandle.get(mDevice, POLAR_KIND_ENC, mBlockLength, mFrozenBits, 1, mSystematic, mErrorDetector);
    }

public:
    ButterflyCudaPacked(size_t blockLength, const std::vector<unsigned>& frozenBits, int device = 0) : mDevice(device)
    {
        initialize(blockLength, frozenBits);
    }
    ~ButterflyCudaPacked() override {}

    void initialize(size_t blockLength, const std::vector<unsigned>& frozenBits) override
    {
        mBlockLength = blockLength;
        mFrozenBits.assign(frozenBits.begin(), frozenBits.end());
        delete mBitContainer;
        mBitContainer = new PackedContainer(mBlockLength, mFrozenBits); // getEncodedData() reads the codeword from here
        mHandle.sys = -1;
    }

    // ButterflyFipPacked::encode (butterfly_fip_packed.cpp:45-59): checksum into the caller's info block, then
    // the codeword into mBitContainer.  setCodeword()-style pre-loaded containers are outside the hot path.
    void encode() override
    {
        if (mCodewordReady) throw std::runtime_error("ButterflyCudaPacked: encode from a pre-loaded codeword is not supported");
        std::vector<unsigned char> cw(mBlockLength / 8);
        CudaDetail::check(polar_cuda_encode_batch(code(), xmInputData, POLAR_MEM_HOST, 1, cw.data(), nullptr));
        mBitContainer->insertPackedBits(cw.data());
    }

    // nFrames x infoLength()/8 bytes in (the checksum bytes of every block are overwritten in place, like
    // Detector::generate does), nFrames x blockLength()/8 bytes out.
    void encode_batch(unsigned char* info, size_t nFrames, unsigned char* codewords, bool deviceBuffers = false,
                      void* stream = nullptr)
    {
        CudaDetail::check(polar_cuda_encode_batch(code(), info, deviceBuffers ? POLAR_MEM_DEVICE : POLAR_MEM_HOST, nFrames,
                                                  codewords, stream));
    }
};

} // namespace Encoding

// Rate matching: the base class keeps the one-vector puncture<T> / depuncture<T> / puncturePacked (host loops over
// mOutputPositions); the *_batch methods run the same index copies over nFrames rows on the device.
class CudaPuncturer : public Puncturer
{
    polar_puncturer_t mHandle = nullptr;

public:
    CudaPuncturer(const size_t blockLength, const std::vector<unsigned> frozenBitPositions, int device = 0)
        : Puncturer(blockLength, frozenBitPositions) // throws std::out_of_range like the reference
    {
        CudaDetail::check(polar_cuda_puncturer_create(device, (int)blockLength, frozenBitPositions.data(),
                                                      (int)frozenBitPositions.size(), &mHandle));
    }
    ~CudaPuncturer() override { polar_cuda_puncturer_destroy(mHandle); }
    CudaPuncturer(const CudaPuncturer&) = delete;
    CudaPuncturer& operator=(const CudaPuncturer&) = delete;

    // nFrames x parentBlockLength() elements in, nFrames x blockLength() out
    template <typename T>
    void puncture_batch(T* pOutput, const T* pInput, size_t nFrames, bool deviceBuffers = false, void* stream = nullptr)
    {
        CudaDetail::check(polar_cuda_puncture_batch(mHandle, pInput, (int)sizeof(T), deviceBuffers ? POLAR_MEM_DEVICE : POLAR_MEM_HOST,
                                                    nFrames, pOutput, stream));
    }
    // nFrames x blockLength() in, nFrames x parentBlockLength() out, punctured positions 0
    template <typename T>
    void depuncture_batch(T* pOutput, const T* pInput, size_t nFrames, bool deviceBuffers = false, void* stream = nullptr)
    {
        CudaDetail::check(polar_cuda_depuncture_batch(mHandle, pInput, (int)sizeof(T), deviceBuffers ? POLAR_MEM_DEVICE : POLAR_MEM_HOST,
                                                      nFrames, pOutput, stream));
    }
    void puncturePacked_batch(unsigned char* pOutput, const unsigned char* pInput, size_t nFrames, bool deviceBuffers = false,
                              void* stream = nullptr)
    {
        CudaDetail::check(polar_cuda_puncture_packed_batch(mHandle, pInput, deviceBuffers ? POLAR_MEM_DEVICE : POLAR_MEM_HOST,
                                                           nFrames, pOutput, stream));
    }
};

} // namespace PolarCode
