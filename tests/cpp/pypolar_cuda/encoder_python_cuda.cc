// bind_encoder() for the reference's pybind11 module `polarcode_python` WITH the GPU encoder: the replacement for
// python/bindings/encoder_python.cc.  Same surface as the reference's binding (encoder_python.cc:28-61):
// PolarEncoder(blockLength, frozenBitPositions) with blockLength / infoLength / setSystematic / isSystematic / frozenBits /
// getErrorDetectionMode / setErrorDetection(size = 0, type = "crc") / encode_vector(uint8[K/8]) -> uint8[N/8].  New:
//   * a third constructor argument encoderType = "cpu" (ButterflyFipPacked, the default) | "cuda" (ButterflyCudaPacked)
//   * encode_batch(uint8[nFrames, K/8]) -> uint8[nFrames, N/8]; like Encoder::encode, the checksum bytes are written into
//     a COPY of the input (numpy arrays handed to pybind may be temporaries), which is returned as the second element
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <polar_cuda_plugin.h>
#include <polarcode/encoding/butterfly_fip_packed.h>
#include <polarcode/encoding/encoder.h>
#include <polarcode/errordetection/errordetector.h>

namespace py = pybind11;

void bind_encoder(py::module& m)
{
    using namespace PolarCode::Encoding;
    py::class_<Encoder>(m, "PolarEncoder")
        .def(py::init([](size_t blockLength, std::vector<unsigned> frozen, std::string type) -> Encoder* {
                 if (type.rfind("cuda", 0) == 0) return new ButterflyCudaPacked(blockLength, frozen);
                 return new ButterflyFipPacked(blockLength, frozen);
             }),
             py::arg("blockLength"), py::arg("frozenBitPositions"), py::arg("encoderType") = "cpu")
        .def("blockLength", &Encoder::blockLength)
        .def("infoLength", &Encoder::infoLength)
        .def("setSystematic", &Encoder::setSystematic)
        .def("isSystematic", &Encoder::isSystematic)
        .def("frozenBits", &Encoder::frozenBits)
        .def("getErrorDetectionMode", &Encoder::getErrorDetectionMode)
        .def(
            "setErrorDetection",
            [](Encoder& self, unsigned size, std::string type) {
                self.setErrorDetection(PolarCode::ErrorDetection::create(size, type));
            },
            py::arg("size") = 0, py::arg("type") = "crc")
        .def("encode_vector",
             [](Encoder& self, const py::array_t<uint8_t, py::array::c_style | py::array::forcecast>& array) {
                 py::buffer_info inb = array.request();
                 if (inb.ndim != 1) throw std::runtime_error("Only ONE-dimensional vectors allowed!");
                 if ((size_t)inb.size != self.infoLength() / 8) throw std::runtime_error("Input vector size != infoSize // 8!");
                 auto result = py::array_t<uint8_t>(self.blockLength() / 8);
                 self.encode_vector(inb.ptr, result.request().ptr);
                 return result;
             })
        .def("encode_batch",
             [](Encoder& self, const py::array_t<uint8_t, py::array::c_style | py::array::forcecast>& array) {
                 py::buffer_info inb = array.request();
                 if (inb.ndim != 2) throw std::runtime_error("encode_batch wants a TWO-dimensional array [nFrames, infoLength / 8]");
                 const size_t n = (size_t)inb.shape[0], kb = self.infoLength() / 8, nb = self.blockLength() / 8;
                 if ((size_t)inb.shape[1] != kb) throw std::runtime_error("Input row size != infoSize // 8!");
                 py::array_t<uint8_t> info({ n, kb }), code({ n, nb });
                 auto* ip = static_cast<unsigned char*>(info.request().ptr);
                 auto* cp = static_cast<unsigned char*>(code.request().ptr);
                 memcpy(ip, inb.ptr, n * kb);
                 {
                     py::gil_scoped_release nogil;
                     if (auto* c = dynamic_cast<ButterflyCudaPacked*>(&self)) c->encode_batch(ip, n, cp);
                     else
                         for (size_t i = 0; i < n; ++i) self.encode_vector(ip + i * kb, cp + i * nb);
                 }
                 return py::make_tuple(code, info);
             });
}
