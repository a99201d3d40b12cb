// bind_decoder() for the reference's pybind11 module `polarcode_python` WITH the GPU decoders: the replacement a maintainer
// drops in for python/bindings/decoder_python.cc.  python_bindings.cc, detector_python.cc,
// frozen_bits_python.cc and puncturer_python.cc of the reference are compiled unchanged next to it.
//
// Same surface as the reference's binding (decoder_python.cc:21-74): PolarDecoder(blockLength, listSize, frozenBitPositions,
// decoderType) with blockLength / infoLength / listSize / setSystematic / isSystematic / frozenBits / getErrorDetectionMode /
// setErrorDetection(size = 0, type = "crc") / decode_vector(float32[N] | int8[N]).  New:
//   * decoderType "cuda-char" | "cuda-float" | "cuda-mixed" -> the CUDA subclasses (createWithCuda, polar_cuda_plugin.h);
//     every other type string still goes to Decoding::create (decoder.cpp:26-48)
//   * decode_batch(float32[nFrames, N] | int8[nFrames, N]) -> (uint8[nFrames, K/8], bool[nFrames]): one kernel launch for a
//     CUDA decoder, a decode_vector loop for the CPU decoders (so scripts can switch types without changing calls)
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <polar_cuda_plugin.h>
#include <polarcode/decoding/decoder.h>
#include <polarcode/errordetection/errordetector.h>

namespace py = pybind11;

namespace {
using PolarCode::Decoding::CudaDecoder;
using PolarCode::Decoding::Decoder;

template <typename T>
py::tuple decode_batch(Decoder& self, const py::array_t<T, py::array::c_style | py::array::forcecast>& array)
{
    py::buffer_info inb = array.request();
    if (inb.ndim != 2) throw std::runtime_error("decode_batch wants a TWO-dimensional array [nFrames, blockLength]");
    if ((size_t)inb.shape[1] != self.blockLength()) throw std::runtime_error("Input row size != blockLength!");
    const size_t n = (size_t)inb.shape[0], kb = (self.infoLength() + 7) / 8;
    py::array_t<uint8_t> info({ n, kb });
    py::array_t<bool> ok(n);
    auto* out = static_cast<unsigned char*>(info.request().ptr);
    auto* okp = static_cast<unsigned char*>(ok.request().ptr);
    const T* llr = static_cast<const T*>(inb.ptr);
    using LlrT = typename std::conditional<std::is_same<T, float>::value, float, char>::type;
    const LlrT* in = reinterpret_cast<const LlrT*>(llr);
    {
        py::gil_scoped_release nogil; // the batch runs on the device (or in a C++ loop): other Python threads may run
        if (auto* c = dynamic_cast<CudaDecoder<false>*>(&self)) {
            c->decode_batch(in, n, out, okp);
        } else if (auto* f = dynamic_cast<CudaDecoder<true>*>(&self)) {
            if (!std::is_same<T, float>::value) throw std::runtime_error("a float decoder takes float32 LLRs");
            f->decode_batch(reinterpret_cast<const float*>(llr), n, out, okp);
        } else {
            for (size_t i = 0; i < n; ++i) okp[i] = self.decode_vector(in + i * self.blockLength(), out + i * kb) ? 1 : 0;
        }
    }
    return py::make_tuple(info, ok);
}

template <typename T, typename LlrT>
py::array_t<uint8_t> decode_vector(Decoder& self, const py::array_t<T, py::array::c_style | py::array::forcecast>& array)
{
    py::buffer_info inb = array.request();
    if (inb.ndim != 1) throw std::runtime_error("Only ONE-dimensional vectors allowed!");
    if ((size_t)inb.size != self.blockLength()) throw std::runtime_error("Input vector size != blockSize // 8!");
    auto result = py::array_t<uint8_t>((self.infoLength() + 7) / 8);
    self.decode_vector(reinterpret_cast<const LlrT*>(inb.ptr), result.request().ptr);
    return result;
}
} // namespace

void bind_decoder(py::module& m)
{
    py::class_<Decoder>(m, "PolarDecoder")
        .def(py::init(&PolarCode::Decoding::createWithCuda), py::arg("blockLength"), py::arg("listSize"),
             py::arg("frozenBitPositions"), py::arg("decoderType"))
        .def("blockLength", &Decoder::blockLength)
        .def("infoLength", &Decoder::infoLength)
        .def("listSize", &Decoder::getListSize)
        .def("setSystematic", &Decoder::setSystematic)
        .def("isSystematic", &Decoder::isSystematic)
        .def("frozenBits", &Decoder::frozenBits)
        .def("getErrorDetectionMode", &Decoder::getErrorDetectionMode)
        .def(
            "setErrorDetection",
            [](Decoder& self, unsigned size, std::string type) {
                self.setErrorDetection(PolarCode::ErrorDetection::create(size, type));
            },
            py::arg("size") = 0, py::arg("type") = "crc")
        .def("decode_vector", &decode_vector<float, float>)
        .def("decode_vector", &decode_vector<int8_t, char>)
        .def("decode_batch", &decode_batch<float>)
        .def("decode_batch", &decode_batch<int8_t>)
        .def("isCuda", [](Decoder& self) {
            return dynamic_cast<CudaDecoder<false>*>(&self) != nullptr || dynamic_cast<CudaDecoder<true>*>(&self) != nullptr;
        });
}
