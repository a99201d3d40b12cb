// Drop-in test: the reference's own CPU classes and the CUDA subclasses (polar_cuda_plugin.h) are driven through
// the SAME base-class pointers (PolarCode::Decoding::Decoder*, PolarCode::Encoding::Encoder*) on the same data,
// the way the reference's decodingtest.cpp / encodingtest.cpp drive theirs, and must agree bit for bit.
//
// Built in the dev container by tests/cpp/Makefile against /root/reference/include (headers only) and linked with
// oracle/_ref/libpolar_ref.so (the unmodified reference library = the host application) and libpolar_cuda.so.
// The binary travels to the GPU box; tests/test_plugin.py runs it there.
#include <polar_cuda_plugin.h>

#include <polarcode/construction/constructor.h>
#include <polarcode/decoding/adaptive_char.h>
#include <polarcode/decoding/adaptive_float.h>
#include <polarcode/decoding/adaptive_mixed.h>
#include <polarcode/decoding/fastssc_avx_float.h>
#include <polarcode/decoding/fastssc_fip_char.h>
#include <polarcode/decoding/scl_avx_float.h>
#include <polarcode/decoding/scl_fip_char.h>
#include <polarcode/encoding/butterfly_fip_packed.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <random>

using namespace PolarCode;

struct Case {
    int N, K, L, impl /*0 char, 1 float*/, crc, frames;
    const char* ctype;
};

static int g_fail = 0, g_checks = 0;
#define CHECK(cond, ...)                                                                                               \
    do {                                                                                                               \
        ++g_checks;                                                                                                    \
        if (!(cond)) {                                                                                                 \
            ++g_fail;                                                                                                  \
            std::printf("FAIL %s:%d: ", __FILE__, __LINE__);                                                           \
            std::printf(__VA_ARGS__);                                                                                  \
            std::printf("\n");                                                                                         \
        }                                                                                                              \
    } while (0)

static Decoding::Decoder* makeRef(const Case& c, const std::vector<unsigned>& frozen, bool sys)
{
    Decoding::Decoder* d = Decoding::makeDecoder(c.N, c.L, frozen, c.impl);
    d->setErrorDetection(ErrorDetection::create(c.crc, "crc"));
    d->setSystematic(sys);
    return d;
}

static void runCase(const Case& c, bool sys)
{
    const std::vector<unsigned> frozen = Construction::frozen_bits(c.N, c.K, 0.0f, c.ctype);
    const int kb = (c.K + 7) / 8, nb = c.N / 8;
    std::mt19937 rng(1234u + c.N * 31u + c.K * 7u + c.L);

    // ---- encoders through Encoder*
    std::unique_ptr<Encoding::Encoder> refEnc(new Encoding::ButterflyFipPacked(c.N, frozen));
    std::unique_ptr<Encoding::Encoder> gpuEnc(new Encoding::ButterflyCudaPacked(c.N, frozen));
    for (Encoding::Encoder* e : { refEnc.get(), gpuEnc.get() }) {
        e->setSystematic(sys);
        e->setErrorDetection(ErrorDetection::create(c.crc, "crc"));
    }
    CHECK(gpuEnc->blockLength() == refEnc->blockLength() && gpuEnc->infoLength() == refEnc->infoLength(), "encoder geometry");
    CHECK(gpuEnc->getErrorDetectionMode() == refEnc->getErrorDetectionMode(), "encoder detection mode");
    std::vector<unsigned char> info((size_t)c.frames * kb), code((size_t)c.frames * nb), infoB, codeB(code.size());
    for (auto& b : info) b = (unsigned char)(rng() & 0xFF);
    if (c.K % 8) // the reference packs K bits MSB first; keep the pad bits zero
        for (int f = 0; f < c.frames; ++f) info[(size_t)f * kb + kb - 1] &= (unsigned char)(0xFF << (8 - c.K % 8));
    infoB = info;
    int encBad = 0;
    for (int f = 0; f < c.frames; ++f) {
        std::vector<unsigned char> a(info.begin() + (size_t)f * kb, info.begin() + (size_t)(f + 1) * kb), b = a, ca(nb), cb(nb);
        refEnc->encode_vector(a.data(), ca.data());
        if (f < 8) { // the one-frame virtual API on the GPU object
            gpuEnc->encode_vector(b.data(), cb.data());
            if (a != b || ca != cb) ++encBad;
        }
        std::memcpy(&info[(size_t)f * kb], a.data(), kb); // with the checksum the encoder wrote
        std::memcpy(&code[(size_t)f * nb], ca.data(), nb);
    }
    static_cast<Encoding::ButterflyCudaPacked*>(gpuEnc.get())->encode_batch(infoB.data(), c.frames, codeB.data());
    CHECK(encBad == 0, "encode_vector N=%d K=%d crc=%d sys=%d: %d frames differ", c.N, c.K, c.crc, (int)sys, encBad);
    CHECK(infoB == info && codeB == code, "encode_batch N=%d K=%d crc=%d sys=%d", c.N, c.K, c.crc, (int)sys);

    // ---- channel: BPSK + AWGN at Eb/N0 = 2 dB (systematic codeword bits, MSB first)
    const float sigma = 1.0f / std::sqrt(2.0f * ((float)c.K / c.N) * std::pow(10.0f, 0.2f));
    std::normal_distribution<float> noise(0.0f, sigma);
    std::vector<float> fl((size_t)c.frames * c.N);
    std::vector<char> ch(fl.size());
    for (int f = 0; f < c.frames; ++f)
        for (int i = 0; i < c.N; ++i) {
            const int bit = (code[(size_t)f * nb + i / 8] >> (7 - i % 8)) & 1;
            const float y = (1.0f - 2.0f * bit) + noise(rng);
            fl[(size_t)f * c.N + i] = 2.0f * y / (sigma * sigma);
            ch[(size_t)f * c.N + i] = (char)std::max(-128.0f, std::min(127.0f, std::nearbyint(10.0f * y)));
        }

    // ---- decoders through Decoder*
    std::unique_ptr<Decoding::Decoder> refDec(makeRef(c, frozen, sys));
    std::unique_ptr<Decoding::Decoder> gpuDec(Decoding::makeCudaDecoder(c.N, c.L, frozen, c.impl));
    CHECK(gpuDec->getErrorDetectionMode() == "CRC-8", "a fresh decoder checks CRC8 like makeDecoder's");
    gpuDec->setErrorDetection(ErrorDetection::create(c.crc, "crc"));
    gpuDec->setSystematic(sys);
    CHECK(gpuDec->blockLength() == refDec->blockLength() && gpuDec->infoLength() == refDec->infoLength() &&
              gpuDec->getListSize() == refDec->getListSize() && gpuDec->isSystematic() == refDec->isSystematic() &&
              gpuDec->getErrorDetectionMode() == refDec->getErrorDetectionMode(),
          "decoder geometry");
    const bool freshRef = c.impl == 1 && c.L > 1; // SclAvxFloat carries path 0's metric from frame to frame
    // The reference's char decoders return garbage in non-systematic mode below N = 256
    // (PackedContainer::insertCharBits writes at byte 0, every reader uses the right-aligned offset:
    // bitcontainer.cpp:958-973 vs :1033-1066) -- nothing to compare against there.
    const bool refBroken = c.impl == 0 && !sys && c.N < 256;
    std::vector<unsigned char> want((size_t)c.frames * kb), wantOk(c.frames), got(want.size()), gotOk(c.frames);
    int oneBad = 0, okCount = 0;
    for (int f = 0; f < c.frames; ++f) {
        if (freshRef) refDec.reset(makeRef(c, frozen, sys));
        bool r;
        if (c.impl == 1) r = refDec->decode_vector(&fl[(size_t)f * c.N], &want[(size_t)f * kb]);
        else r = refDec->decode_vector(&ch[(size_t)f * c.N], &want[(size_t)f * kb]);
        wantOk[f] = r ? 1 : 0;
        okCount += r;
        if (f < 8) { // decode_vector on the GPU object: setSignal + decode + getDecodedInformationBits
            std::vector<unsigned char> o(kb);
            bool g;
            if (c.impl == 1) g = gpuDec->decode_vector(&fl[(size_t)f * c.N], o.data());
            else g = gpuDec->decode_vector(&ch[(size_t)f * c.N], o.data());
            if (g != r || std::memcmp(o.data(), &want[(size_t)f * kb], kb)) ++oneBad;
        }
    }
    auto* cd = gpuDec.get();
    if (c.impl == 1) static_cast<Decoding::CudaDecoder<true>*>(cd)->decode_batch(fl.data(), c.frames, got.data(), gotOk.data());
    else static_cast<Decoding::CudaDecoder<false>*>(cd)->decode_batch(ch.data(), c.frames, got.data(), gotOk.data());
    int batchBad = 0;
    for (int f = 0; f < c.frames; ++f)
        if (gotOk[f] != wantOk[f] || std::memcmp(&got[(size_t)f * kb], &want[(size_t)f * kb], kb)) ++batchBad;
    if (refBroken) { // compare with what was sent instead
        int same = 0;
        for (int f = 0; f < c.frames; ++f) same += !std::memcmp(&got[(size_t)f * kb], &info[(size_t)f * kb], kb);
        CHECK(2 * same >= c.frames, "non-systematic N=%d: only %d of %d frames equal the transmitted payload", c.N, same, c.frames);
        oneBad = batchBad = 0;
    }
    CHECK(oneBad == 0, "decode_vector N=%d K=%d L=%d impl=%d crc=%d sys=%d: %d of 8 frames differ", c.N, c.K, c.L, c.impl, c.crc,
          (int)sys, oneBad);
    CHECK(batchBad == 0, "decode_batch N=%d K=%d L=%d impl=%d crc=%d sys=%d: %d of %d frames differ", c.N, c.K, c.L, c.impl,
          c.crc, (int)sys, batchBad, c.frames);
    if (c.impl == 0 && !refBroken) { // a char decoder fed floats quantises like CharContainer::insertLlr
        std::vector<unsigned char> o(kb), w(kb);
        std::vector<float> ten(c.N);
        for (int i = 0; i < c.N; ++i) ten[i] = 0.37f * fl[i];
        const bool r = refDec->decode_vector(ten.data(), w.data()), g = gpuDec->decode_vector(ten.data(), o.data());
        CHECK(r == g && o == w, "float LLRs into the char decoder N=%d L=%d", c.N, c.L);
    }
    if (c.impl == 0 && c.L == 1) { // FastSscCudaChar: getSoftCodeword / getSoftInformation equal the reference's, byte for byte
        int bad = 0;
        std::vector<char> a(c.N), b(c.N), ai(c.K), bi(c.K);
        std::vector<unsigned char> o(kb);
        for (int f = 0; f < std::min(c.frames, 16); ++f) {
            refDec->decode_vector(&ch[(size_t)f * c.N], o.data());
            gpuDec->decode_vector(&ch[(size_t)f * c.N], o.data());
            refDec->getSoftCodeword(a.data());
            gpuDec->getSoftCodeword(b.data());
            refDec->getSoftInformation(ai.data());
            gpuDec->getSoftInformation(bi.data());
            if (a != b || ai != bi) ++bad;
        }
        CHECK(bad == 0, "soft outputs N=%d K=%d sys=%d: %d of 16 frames differ", c.N, c.K, (int)sys, bad);
    } else { // elsewhere soft outputs are not produced on the GPU: the accessors refuse instead of returning stale memory
        std::vector<float> soft(c.N);
        bool threw = false;
        try { gpuDec->getSoftCodeword(soft.data()); } catch (const std::logic_error&) { threw = true; }
        CHECK(threw, "getSoftCodeword on a CUDA decoder must throw");
        threw = false;
        try { gpuDec->getSoftInformation(soft.data()); } catch (const std::logic_error&) { threw = true; }
        CHECK(threw, "getSoftInformation on a CUDA decoder must throw");
    }
    if (c.impl == 0 && sys && c.K >= 64 && c.frames >= 32) { // a detector the kernels do not fuse (cmac): candidates come back, the host checks
        std::unique_ptr<Decoding::Decoder> r2(Decoding::makeDecoder(c.N, c.L, frozen, 0)), g2(Decoding::makeCudaDecoder(c.N, c.L, frozen, 0));
        r2->setErrorDetection(ErrorDetection::create(32, "cmac"));
        g2->setErrorDetection(ErrorDetection::create(32, "cmac"));
        int bad = 0;
        std::vector<unsigned char> a(kb), b(kb);
        for (int f = 0; f < 32; ++f) {
            const bool ra = r2->decode_vector(&ch[(size_t)f * c.N], a.data()), rb = g2->decode_vector(&ch[(size_t)f * c.N], b.data());
            if (ra != rb || a != b) ++bad;
        }
        CHECK(bad == 0, "host-side detector (cmac) N=%d L=%d: %d of 32 frames differ", c.N, c.L, bad);
    }
    std::printf("case N=%d K=%d L=%d %s crc=%d sys=%d %s: %d frames, %d pass the detector\n", c.N, c.K, c.L,
                c.impl ? "float" : "char", c.crc, (int)sys, c.ctype, c.frames, okCount);
}

// Adaptive decoders: the reference's AdaptiveChar / AdaptiveFloat / AdaptiveMixed (adaptive_char.cpp:34-46) against
// AdaptiveCudaChar / Float / Mixed through Decoder*, float LLRs (the only setSignal the reference classes override).
// flavour 0 char, 1 float, 2 mixed.  The fp32 list decoder inside Float / Mixed is history dependent: fresh object per frame.
static Decoding::Decoder* makeRefAdaptive(int flavour, int N, int L, const std::vector<unsigned>& frozen, int crc)
{
    Decoding::Decoder* d = flavour == 0   ? (Decoding::Decoder*)new Decoding::AdaptiveChar(N, L, frozen)
                           : flavour == 1 ? (Decoding::Decoder*)new Decoding::AdaptiveFloat(N, L, frozen)
                                          : (Decoding::Decoder*)new Decoding::AdaptiveMixed(N, L, frozen);
    d->setErrorDetection(ErrorDetection::create(crc, "crc"));
    return d;
}

static void runAdaptive(int flavour, int N, int K, int L, int crc, int frames, const char* ctype, float ebn0)
{
    const std::vector<unsigned> frozen = Construction::frozen_bits(N, K, 0.0f, ctype);
    const int kb = K / 8, nb = N / 8;
    std::mt19937 rng(77u + N + 3u * flavour);
    std::unique_ptr<Encoding::Encoder> enc(new Encoding::ButterflyFipPacked(N, frozen));
    enc->setErrorDetection(ErrorDetection::create(crc, "crc"));
    const float sigma = 1.0f / std::sqrt(2.0f * ((float)K / N) * std::pow(10.0f, ebn0 / 10.0f));
    std::normal_distribution<float> noise(0.0f, sigma);
    std::vector<float> fl((size_t)frames * N);
    for (int f = 0; f < frames; ++f) {
        std::vector<unsigned char> a(kb), cw(nb);
        for (auto& b : a) b = (unsigned char)(rng() & 0xFF);
        enc->encode_vector(a.data(), cw.data());
        for (int i = 0; i < N; ++i) {
            const int bit = (cw[i / 8] >> (7 - i % 8)) & 1;
            const float y = (1.0f - 2.0f * bit) + noise(rng);
            fl[(size_t)f * N + i] = flavour == 0 ? 10.0f * y : 2.0f * y / (sigma * sigma);
        }
    }
    std::unique_ptr<Decoding::Decoder> refDec(makeRefAdaptive(flavour, N, L, frozen, crc));
    std::unique_ptr<Decoding::Decoder> gpuDec(flavour == 0   ? (Decoding::Decoder*)new Decoding::AdaptiveCudaChar(N, L, frozen)
                                              : flavour == 1 ? (Decoding::Decoder*)new Decoding::AdaptiveCudaFloat(N, L, frozen)
                                                             : (Decoding::Decoder*)new Decoding::AdaptiveCudaMixed(N, L, frozen));
    gpuDec->setErrorDetection(ErrorDetection::create(crc, "crc"));
    std::vector<unsigned char> want((size_t)frames * kb), wantOk(frames), got(want.size()), gotOk(frames);
    int oneBad = 0, okCount = 0;
    for (int f = 0; f < frames; ++f) {
        if (flavour != 0) refDec.reset(makeRefAdaptive(flavour, N, L, frozen, crc));
        const bool r = refDec->decode_vector(&fl[(size_t)f * N], &want[(size_t)f * kb]);
        wantOk[f] = r ? 1 : 0;
        okCount += r;
        if (f < 12) {
            std::vector<unsigned char> o(kb);
            const bool g = gpuDec->decode_vector(&fl[(size_t)f * N], o.data());
            if (g != r || std::memcmp(o.data(), &want[(size_t)f * kb], kb)) ++oneBad;
        }
    }
    long redone;
    if (flavour == 0) {
        auto* d = static_cast<Decoding::CudaDecoder<false>*>(gpuDec.get());
        const long before = d->fallbacks();
        d->decode_batch(fl.data(), frames, got.data(), gotOk.data());
        redone = d->fallbacks() - before;
    } else {
        auto* d = static_cast<Decoding::CudaDecoder<true>*>(gpuDec.get());
        const long before = d->fallbacks();
        d->decode_batch(fl.data(), frames, got.data(), gotOk.data());
        redone = d->fallbacks() - before;
    }
    int batchBad = 0;
    for (int f = 0; f < frames; ++f)
        if (gotOk[f] != wantOk[f] || std::memcmp(&got[(size_t)f * kb], &want[(size_t)f * kb], kb)) ++batchBad;
    CHECK(oneBad == 0, "adaptive decode_vector flavour=%d N=%d L=%d: %d of 12 frames differ", flavour, N, L, oneBad);
    CHECK(batchBad == 0, "adaptive decode_batch flavour=%d N=%d L=%d: %d of %d frames differ", flavour, N, L, batchBad, frames);
    CHECK(redone > 0 && redone < frames, "adaptive flavour=%d: %ld of %d frames fell through to the list decoder", flavour, redone, frames);
    std::printf("adaptive %s N=%d K=%d L=%d crc=%d: %d frames, %ld redone by the list decoder, %d pass the detector\n",
                flavour == 0 ? "char" : flavour == 1 ? "float" : "mixed", N, K, L, crc, frames, redone, okCount);
}

// Puncturer (puncturer.h:29-101) vs CudaPuncturer: same positions, the batched device copies equal the base-class loops.
static void runPuncturer(int M, int N, int K, int frames)
{
    const std::vector<unsigned> frozen = Construction::frozen_bits(N, K, 0.0f, "BB");
    Puncturer ref(M, frozen);
    CudaPuncturer gpu(M, frozen);
    CHECK(gpu.blockLength() == ref.blockLength() && gpu.parentBlockLength() == ref.parentBlockLength() &&
              gpu.blockOutputPositions() == ref.blockOutputPositions(), "puncturer geometry M=%d", M);
    std::mt19937 rng(M);
    std::vector<float> x((size_t)frames * N), px((size_t)frames * M), wantP(px.size()), dx(x.size()), wantD(x.size());
    for (auto& v : x) v = (float)(int)(rng() % 255) - 127.0f;
    gpu.puncture_batch(px.data(), x.data(), frames);
    for (int f = 0; f < frames; ++f) ref.puncture(&wantP[(size_t)f * M], &x[(size_t)f * N]);
    CHECK(px == wantP, "puncture_batch<float> M=%d", M);
    gpu.depuncture_batch(dx.data(), px.data(), frames);
    for (int f = 0; f < frames; ++f) ref.depuncture(&wantD[(size_t)f * N], &wantP[(size_t)f * M]);
    CHECK(dx == wantD, "depuncture_batch<float> M=%d", M);
    std::vector<char> c((size_t)frames * M), dc((size_t)frames * N), wantC(dc.size());
    for (auto& v : c) v = (char)(rng() & 0xFF);
    gpu.depuncture_batch(dc.data(), c.data(), frames);
    for (int f = 0; f < frames; ++f) ref.depuncture(&wantC[(size_t)f * N], &c[(size_t)f * M]);
    CHECK(dc == wantC, "depuncture_batch<char> M=%d", M);
    if (M % 8 == 0) {
        std::vector<unsigned char> b((size_t)frames * N / 8), pb((size_t)frames * M / 8), wantB(pb.size());
        for (auto& v : b) v = (unsigned char)(rng() & 0xFF);
        gpu.puncturePacked_batch(pb.data(), b.data(), frames);
        for (int f = 0; f < frames; ++f) ref.puncturePacked(&wantB[(size_t)f * M / 8], &b[(size_t)f * N / 8]);
        CHECK(pb == wantB, "puncturePacked_batch M=%d", M);
    }
    bool threw = false;
    try { CudaPuncturer bad(N / 2 + 1, std::vector<unsigned>(frozen.begin(), frozen.begin() + 3)); } catch (const std::out_of_range&) { threw = true; }
    CHECK(threw, "too many punctured positions must throw std::out_of_range");
    std::printf("puncturer M=%d parent=%d: %d frames\n", M, N, frames);
}

int main(int argc, char** argv)
{
    if (polar_cuda_device_count() < 1) {
        std::printf("no CUDA device\n");
        return 77;
    }
    const Case cases[] = {
        { 1024, 512, 8, 0, 16, 64, "5G" }, { 1024, 512, 1, 0, 0, 128, "BB" },  { 256, 128, 1, 1, 0, 128, "BB" },
        { 1024, 512, 8, 1, 16, 24, "5G" }, { 2048, 1024, 32, 0, 32, 8, "BB" }, { 128, 64, 4, 0, 8, 64, "BB" },
        { 64, 32, 2, 1, 8, 64, "BB" },     { 512, 256, 1, 0, 8, 64, "BE" },    { 32, 16, 1, 1, 0, 64, "BB" },
    };
    const bool quick = argc > 1 && std::string(argv[1]) == "--quick";
    int n = 0;
    for (const Case& c : cases) {
        if (quick && n++ >= 3) break;
        runCase(c, true);
        runCase(c, false);
    }
    runPuncturer(1000, 1024, 512, 64);
    runPuncturer(203, 256, 100, 33);
    runAdaptive(0, 1024, 512, 8, 16, 96, "5G", 1.5f);
    if (!quick) {
        runAdaptive(1, 1024, 512, 8, 16, 32, "5G", 1.5f);
        runAdaptive(2, 1024, 512, 8, 16, 32, "5G", 1.5f);
        runAdaptive(0, 256, 128, 4, 8, 128, "BB", 1.0f);
        runAdaptive(2, 256, 128, 4, 8, 64, "BB", 1.0f);
    }
    // error behaviour: the reference throws std::logic_error("CRC INVALID SIZE!") for an unknown CRC width
    // (errordetector.cpp:33-35); a block length that is not a power of two is an invalid argument
    bool threw = false;
    try {
        std::vector<unsigned> half(500);
        for (unsigned i = 0; i < 500; ++i) half[i] = i;
        Decoding::FastSscCudaChar bad(1000, half);
        unsigned char o[128];
        std::vector<char> z(1000, 1);
        bad.decode_vector(z.data(), o);
    } catch (const std::invalid_argument&) { threw = true; }
    CHECK(threw, "block length 1000 must be rejected with std::invalid_argument");
    std::printf("%s: %d checks, %d failed\n", g_fail ? "FAILED" : "PASSED", g_checks, g_fail);
    return g_fail ? 1 : 0;
}
