"""antpolarcodes_b200 -- B200 (sm_100a) batched polar-code engine.

Python mirror of the reference's pypolar surface (python/bindings/*.cc): ``PolarEncoder``,
``PolarDecoder``, ``Detector`` and ``frozen_bits`` keep the reference's names, argument meaning and
error behaviour, and gain batched methods (``encode_batch`` / ``decode_batch``) that take
``[nFrames, ...]`` numpy arrays (host) or torch CUDA tensors (device resident, zero copy).

Everything computes in ``libpolar_cuda.so`` (hand-written CUDA, C ABI in include/polar_cuda.h).
There is no CPU fallback: importing works without a GPU, creating a code does not.
"""
import ctypes as _c
import os as _os

import numpy as _np

_HERE = _os.path.dirname(_os.path.abspath(__file__))
_LIB_PATH = _os.environ.get("POLAR_CUDA_LIB") or _os.path.join(_HERE, "libpolar_cuda.so")  # the override is for A/B builds

KIND_ENC, KIND_FASTSSC_I8, KIND_FASTSSC_F32, KIND_SCL_I8, KIND_SCL_F32 = range(5)
KIND_ADAPTIVE_I8, KIND_ADAPTIVE_F32, KIND_ADAPTIVE_MIXED = 5, 6, 7
MEM_HOST, MEM_DEVICE = 0, 1
LLR_I8, LLR_F32 = 0, 1

EXPORTS = [
    "polar_cuda_last_error", "polar_cuda_device_count", "polar_cuda_version", "polar_cuda_code_create",
    "polar_cuda_code_destroy", "polar_cuda_decode_batch", "polar_cuda_encode_batch",
    "polar_cuda_crc_generate", "polar_cuda_crc_check", "polar_cuda_frozen_bits", "polar_cuda_code_info",
    "polar_cuda_code_schedule", "polar_cuda_code_launches", "polar_cuda_test_fg", "polar_cuda_schedule_stats",
    "polar_cuda_code_fallbacks", "polar_cuda_schedule_cache_stats", "polar_cuda_puncturer_positions", "polar_cuda_puncturer_create",
    "polar_cuda_puncturer_destroy", "polar_cuda_puncturer_parent_length", "polar_cuda_puncture_batch",
    "polar_cuda_depuncture_batch", "polar_cuda_puncture_packed_batch",
    "polar_cuda_sim_create", "polar_cuda_sim_destroy", "polar_cuda_sim_run", "polar_cuda_sim_buffers", "polar_cuda_sim_launches",
    "polar_cuda_test_philox", "polar_cuda_decode_candidates", "polar_cuda_decode_batch_noalloc",
    "polar_cuda_decode_soft_batch",
]


class CodeInfo(_c.Structure):
    _fields_ = [("kind", _c.c_int), ("N", _c.c_int), ("K", _c.c_int), ("listSize", _c.c_int),
                ("systematic", _c.c_int), ("crcBits", _c.c_int), ("nOps", _c.c_int),
                ("warpsPerBlock", _c.c_int), ("blocksPerSM", _c.c_int), ("smemPerWarp", _c.c_int),
                ("gridBlocks", _c.c_int), ("laneOpsPerFrame", _c.c_double), ("bytesPerFrame", _c.c_double),
                ("opCount", _c.c_long * 32), ("opSize", _c.c_long * 32)]


_lib = None


def lib():
    """The loaded C-ABI library.  Raises (loudly) if the CUDA extension has not been built."""
    global _lib
    if _lib is None:
        if not _os.path.exists(_LIB_PATH):
            raise ImportError("libpolar_cuda.so is missing: run `python antpolarcodes_b200/build.py` "
                              "(or __graft_entry__.build()); there is no CPU fallback")
        l = _c.CDLL(_LIB_PATH)
        vp, i, sz = _c.c_void_p, _c.c_int, _c.c_size_t
        l.polar_cuda_last_error.restype = _c.c_char_p
        l.polar_cuda_code_create.argtypes = [i, i, i, vp, i, i, i, i, _c.POINTER(vp)]
        l.polar_cuda_code_destroy.argtypes = [vp]
        l.polar_cuda_decode_batch.argtypes = [vp, vp, i, i, sz, vp, vp, vp]
        l.polar_cuda_encode_batch.argtypes = [vp, vp, i, sz, vp, vp]
        l.polar_cuda_decode_candidates.argtypes = [vp, vp, i, i, sz, vp, vp, vp]
        l.polar_cuda_decode_batch_noalloc.argtypes = [vp, vp, i, sz, vp, vp, vp]
        l.polar_cuda_decode_soft_batch.argtypes = [vp, vp, i, sz, vp, vp, vp]
        l.polar_cuda_crc_generate.argtypes = [i, vp, i]
        l.polar_cuda_crc_check.argtypes = [i, vp, i]
        l.polar_cuda_frozen_bits.argtypes = [i, i, _c.c_float, _c.c_char_p, vp]
        l.polar_cuda_code_info.argtypes = [vp, _c.POINTER(CodeInfo)]
        l.polar_cuda_code_schedule.argtypes = [vp, vp, i]
        l.polar_cuda_code_launches.argtypes = [vp]
        l.polar_cuda_code_launches.restype = _c.c_long
        l.polar_cuda_code_fallbacks.argtypes = [vp]
        l.polar_cuda_code_fallbacks.restype = _c.c_long
        l.polar_cuda_test_fg.argtypes = [vp, vp, vp, i, vp, vp]
        l.polar_cuda_puncturer_positions.argtypes = [i, vp, i, vp]
        l.polar_cuda_puncturer_create.argtypes = [i, i, vp, i, _c.POINTER(vp)]
        l.polar_cuda_puncturer_destroy.argtypes = [vp]
        l.polar_cuda_puncturer_parent_length.argtypes = [vp]
        l.polar_cuda_puncture_batch.argtypes = [vp, vp, i, i, sz, vp, vp]
        l.polar_cuda_depuncture_batch.argtypes = [vp, vp, i, i, sz, vp, vp]
        l.polar_cuda_puncture_packed_batch.argtypes = [vp, vp, i, sz, vp, vp]
        l.polar_cuda_sim_create.argtypes = [i, i, vp, i, i, i, i, i, sz, _c.POINTER(vp)]
        l.polar_cuda_sim_destroy.argtypes = [vp]
        l.polar_cuda_sim_run.argtypes = [vp, _c.c_double, _c.c_double, _c.c_uint64, _c.c_uint32, _c.c_uint64, sz, vp, vp]
        l.polar_cuda_sim_buffers.argtypes = [vp, _c.POINTER(vp), _c.POINTER(vp), _c.POINTER(vp), _c.POINTER(vp), _c.POINTER(sz)]
        l.polar_cuda_sim_launches.argtypes = [vp, _c.POINTER(_c.c_long)]
        l.polar_cuda_sim_launches.restype = _c.c_long
        l.polar_cuda_test_philox.argtypes = [vp, vp, vp]
        _lib = l
    return _lib


def _check(rc):
    if rc < 0:
        msg = lib().polar_cuda_last_error().decode()
        if rc == -1:
            raise ValueError(msg)        # std::invalid_argument / std::logic_error in the reference
        if rc == -5:
            raise IndexError(msg)        # std::out_of_range (pybind11 translates it to IndexError)
        raise RuntimeError(msg)
    return rc


def _is_torch(x):
    return type(x).__module__.startswith("torch")


def _buf(x):
    """(pointer, mem) of a numpy array or torch tensor (must be contiguous)."""
    if _is_torch(x):
        if not x.is_contiguous():
            raise RuntimeError("tensor must be contiguous")
        return x.data_ptr(), (MEM_DEVICE if x.is_cuda else MEM_HOST)
    if not x.flags["C_CONTIGUOUS"]:
        raise RuntimeError("array must be C-contiguous")
    return x.ctypes.data, MEM_HOST


def frozen_bits(blockLength, infoLength, designSNR, constructorType="BB"):
    """Construction::frozen_bits (frozen_bits_python.cc:24-29): sorted frozen indices."""
    out = _np.zeros(max(int(blockLength), 1), _np.uint32)
    n = _check(lib().polar_cuda_frozen_bits(int(blockLength), int(infoLength), float(designSNR),
                                            str(constructorType).encode(), out.ctypes.data))
    return out[:n].copy()


class Detector:
    """ErrorDetection::Detector created by ErrorDetection::create(size, type) (detector_python.cc:26-56)."""

    def __init__(self, size=0, type="crc"):
        t = str(type).lower()
        if "crc" not in t:
            if "cmac" in t:
                raise RuntimeError("cmac detector is outside the GPU hot path")
            raise RuntimeError("Unknown Error detector requested!")
        if size not in (0, 8, 16, 32):
            raise ValueError("CRC INVALID SIZE!")
        self._bits = int(size)

    def getCheckBitCount(self):
        return self._bits

    def getType(self):
        return "CRC" if self._bits else "DUMMY"

    def generate(self, data):
        d = _np.ascontiguousarray(data, _np.uint8).copy()
        _check(lib().polar_cuda_crc_generate(self._bits, d.ctypes.data, d.size))
        return d

    def check(self, data):
        d = _np.ascontiguousarray(data, _np.uint8)
        return bool(_check(lib().polar_cuda_crc_check(self._bits, d.ctypes.data, d.size)))


class _Code:
    def __init__(self, kind, N, frozen, L, systematic, crc, device=0):
        self._h = _c.c_void_p()
        f = _np.ascontiguousarray(frozen, _np.uint32)
        _check(lib().polar_cuda_code_create(int(device), kind, int(N), f.ctypes.data, len(f), int(L),
                                            int(bool(systematic)), int(crc), _c.byref(self._h)))

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            lib().polar_cuda_code_destroy(self._h)
            self._h = _c.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:   # interpreter shutdown: the module globals may already be gone, the process frees the device memory
            pass

    def info(self):
        ci = CodeInfo()
        _check(lib().polar_cuda_code_info(self._h, _c.byref(ci)))
        return ci

    def schedule(self):
        n = _check(lib().polar_cuda_code_schedule(self._h, None, 0))
        ops = _np.zeros(n, _np.uint32)
        lib().polar_cuda_code_schedule(self._h, ops.ctypes.data, n)
        return ops

    def launches(self):
        return int(lib().polar_cuda_code_launches(self._h))

    def fallbacks(self):
        return int(lib().polar_cuda_code_fallbacks(self._h))


class PolarEncoder:
    """Encoding::ButterflyFipPacked behind the pypolar PolarEncoder surface (encoder_python.cc:28-61)."""

    def __init__(self, blockLength, frozenBitPositions, device=0):
        self._N = int(blockLength)
        self._frozen = _np.ascontiguousarray(frozenBitPositions, _np.uint32).copy()
        self._sys, self._crc, self._dev, self._code = True, 0, device, None

    def _get(self):
        if self._code is None:
            self._code = _Code(KIND_ENC, self._N, self._frozen, 1, self._sys, self._crc, self._dev)
        return self._code

    def _reset(self):
        if self._code is not None:
            self._code.close()
        self._code = None

    def blockLength(self):
        return self._N

    def infoLength(self):
        return self._N - len(self._frozen)

    def frozenBits(self):
        return self._frozen.copy()

    def setSystematic(self, flag):
        self._sys = bool(flag)
        self._reset()

    def isSystematic(self):
        return self._sys

    def setErrorDetection(self, size=0, type="crc"):
        self._crc = Detector(size, type).getCheckBitCount()
        self._reset()

    def getErrorDetectionMode(self):
        return "%s-%d" % ("CRC" if self._crc else "DUMMY", self._crc)

    def encode_vector(self, info):
        """uint8[K/8] -> uint8[N/8] (one frame).  The caller's array is NOT modified (the pybind wrapper
        of the reference mutates its argument in place; use encode_batch for that behaviour)."""
        a = _np.ascontiguousarray(info, _np.uint8)
        if a.ndim != 1:
            raise RuntimeError("Number of dimensions must be one")
        if a.size != (self.infoLength() + 7) // 8:
            raise RuntimeError("Input vector size mismatch!")
        return self.encode_batch(a.reshape(1, -1).copy())[0]

    def encode_batch(self, info, out=None, stream=0):
        """[nFrames, K/8] uint8 -> [nFrames, N/8] uint8.  CRC bytes are written into `info` in place
        (butterfly_fip_packed.cpp:47-50).  numpy arrays are staged through the device, torch CUDA tensors
        are used where they are."""
        kb, nb = (self.infoLength() + 7) // 8, max(self._N // 8, 1)
        if info.ndim != 2 or info.shape[1] != kb:
            raise RuntimeError("Input vector size mismatch!")
        n = info.shape[0]
        if out is None:
            if _is_torch(info):
                import torch
                out = torch.empty((n, nb), dtype=torch.uint8, device=info.device)
            else:
                out = _np.empty((n, nb), _np.uint8)
        ip, mem = _buf(info)
        op, mem2 = _buf(out)
        if mem != mem2:
            raise RuntimeError("info and out must live in the same memory space")
        _check(lib().polar_cuda_encode_batch(self._get()._h, ip, mem, n, op, stream))
        return out


class PolarDecoder:
    """Decoding::create(blockLength, listSize, frozenBits, decoderType) behind the pypolar PolarDecoder
    surface (decoder_python.cc:21-74).  listSize 1 -> Fast-SSC, listSize > 1 -> SCL; "char" = int8
    arithmetic (FastSscFipChar / SclFipChar), "float" = fp32, "mixed" with listSize > 1 = AdaptiveFloat
    (decoder.cpp:37-38,74-76: Fast-SSC first, the list decoder only on detector failures).  Like the
    reference's makeDecoder (decoder.cpp:85) a fresh decoder checks CRC8 until setErrorDetection is called.

    ``adaptive=True`` (not in pypolar, whose factory cannot reach them) selects the adaptive classes the
    reference's simulator builds directly: "char" -> AdaptiveChar, "float" -> AdaptiveFloat, "mixed" ->
    AdaptiveMixed (int8 Fast-SSC + fp32 list decoder)."""

    def __init__(self, blockLength, listSize, frozenBitPositions, decoderType="char", device=0, adaptive=False):
        t = str(decoderType).lower()
        if "char" in t:
            flavour = "char"
        elif "float" in t:
            flavour = "float"
        elif "mixed" in t:
            flavour = "mixed"
        elif "scan" in t:
            raise RuntimeError("decoder type '%s' is outside the GPU hot path" % decoderType)
        else:
            raise ValueError("Unknown PolarDecoder type!")
        self._N, self._L = int(blockLength), int(listSize)
        self._frozen = _np.ascontiguousarray(frozenBitPositions, _np.uint32).copy()
        if adaptive:
            self._kind = {"char": KIND_ADAPTIVE_I8, "float": KIND_ADAPTIVE_F32, "mixed": KIND_ADAPTIVE_MIXED}[flavour]
        elif self._L <= 1:  # decoder.cpp:44-46: below two paths everything but "char" is the float Fast-SSC
            self._kind = KIND_FASTSSC_I8 if flavour == "char" else KIND_FASTSSC_F32
        elif flavour == "mixed":
            self._kind = KIND_ADAPTIVE_F32
        else:
            self._kind = KIND_SCL_F32 if flavour == "float" else KIND_SCL_I8
        self._sys, self._crc, self._dev, self._code = True, 8, device, None

    def _get(self):
        if self._code is None:
            self._code = _Code(self._kind, self._N, self._frozen, max(self._L, 1), self._sys, self._crc, self._dev)
        return self._code

    def _reset(self):
        if self._code is not None:
            self._code.close()
        self._code = None

    def blockLength(self):
        return self._N

    def infoLength(self):
        return self._N - len(self._frozen)

    def listSize(self):
        return max(self._L, 1)

    def frozenBits(self):
        return self._frozen.copy()

    def setSystematic(self, flag):
        self._sys = bool(flag)
        self._reset()

    def isSystematic(self):
        return self._sys

    def setErrorDetection(self, size=0, type="crc"):
        self._crc = Detector(size, type).getCheckBitCount()
        self._reset()

    def getErrorDetectionMode(self):
        return "%s-%d" % ("CRC" if self._crc else "DUMMY", self._crc)

    def info(self):
        return self._get().info()

    def fallbacks(self):
        """adaptive decoders: frames redone by the list decoder so far"""
        return self._get().fallbacks()

    def soft_batch(self, llr, want_info=True):
        """Decoder::getSoftCodeword / getSoftInformation over a batch (int8 Fast-SSC only): [nFrames, N] int8 LLRs ->
        ([nFrames, N] int8 soft codeword, [nFrames, K] int8 soft information or None)."""
        if llr.ndim != 2 or llr.shape[1] != self._N:
            raise RuntimeError("Input vector size mismatch!")
        n, K = llr.shape[0], self.infoLength()
        if _is_torch(llr):
            import torch
            if llr.dtype != torch.int8:
                raise RuntimeError("soft outputs take int8 LLRs")
            soft = torch.empty((n, self._N), dtype=torch.int8, device=llr.device)
            info = torch.empty((n, K), dtype=torch.int8, device=llr.device) if want_info else None
        else:
            if llr.dtype != _np.int8:
                raise RuntimeError("soft outputs take int8 LLRs")
            llr = _np.ascontiguousarray(llr)
            soft = _np.empty((n, self._N), _np.int8)
            info = _np.empty((n, K), _np.int8) if want_info else None
        lp, mem = _buf(llr)
        sp, _ = _buf(soft)
        ip = _buf(info)[0] if info is not None else None
        _check(lib().polar_cuda_decode_soft_batch(self._get()._h, lp, mem, n, sp, ip, 0))
        return soft, info

    def decode_vector(self, llr):
        """float32[N] or int8[N] -> uint8[K/8]; the detector verdict is dropped, as in decoder_python.cc."""
        a = _np.ascontiguousarray(llr)
        if a.ndim != 1:
            raise RuntimeError("Number of dimensions must be one")
        if a.size != self._N:
            raise RuntimeError("Input vector size mismatch!")
        out, _ = self.decode_batch(a.reshape(1, -1))
        return out[0]

    def decode_batch(self, llr, out=None, ok=None, stream=0):
        """[nFrames, N] int8 / float32 LLRs -> ([nFrames, ceil(K/8)] uint8, [nFrames] uint8 verdicts)."""
        if llr.ndim != 2 or llr.shape[1] != self._N:
            raise RuntimeError("Input vector size mismatch!")
        n, kb = llr.shape[0], (self.infoLength() + 7) // 8
        if _is_torch(llr):
            import torch
            if llr.dtype == torch.int8:
                lt = LLR_I8
            elif llr.dtype == torch.float32:
                lt = LLR_F32
            else:
                raise RuntimeError("LLRs must be int8 or float32")
            if out is None:
                out = torch.empty((n, kb), dtype=torch.uint8, device=llr.device)
            if ok is None:
                ok = torch.empty((n,), dtype=torch.uint8, device=llr.device)
        else:
            if llr.dtype == _np.int8:
                lt = LLR_I8
            elif llr.dtype == _np.float32:
                lt = LLR_F32
            else:
                raise RuntimeError("LLRs must be int8 or float32")
            llr = _np.ascontiguousarray(llr)
            if out is None:
                out = _np.empty((n, kb), _np.uint8)
            if ok is None:
                ok = _np.empty((n,), _np.uint8)
        lp, mem = _buf(llr)
        op, m2 = _buf(out)
        kp, m3 = _buf(ok)
        if not (mem == m2 == m3):
            raise RuntimeError("llr, out and ok must live in the same memory space")
        _check(lib().polar_cuda_decode_batch(self._get()._h, lp, lt, mem, n, op, kp, stream))
        return out, ok


class Puncturer:
    """PolarCode::Puncturer behind the pypolar surface (puncturer_python.cc:19-150): rate matching by dropping the
    first parentBlockLength - blockLength frozen positions of the parent code word.  The one-vector methods
    (puncture / depuncture / puncturePacked) keep the binding's signatures; the *_batch methods take
    [nFrames, ...] numpy arrays (host) or torch CUDA tensors."""

    def __init__(self, blockLength, frozenBitPositions, device=0):
        self._M = int(blockLength)
        self._frozen = _np.ascontiguousarray(frozenBitPositions, _np.uint32).copy()
        out = _np.zeros(max(2 * self._M, 8), _np.uint32)
        n = _check(lib().polar_cuda_puncturer_positions(self._M, self._frozen.ctypes.data, len(self._frozen),
                                                        out.ctypes.data))
        self._pos = out[:n].copy()
        self._parent = 1
        while self._parent < self._M:
            self._parent <<= 1
        self._dev, self._h = device, None

    def _get(self):
        if self._h is None:
            h = _c.c_void_p()
            _check(lib().polar_cuda_puncturer_create(int(self._dev), self._M, self._frozen.ctypes.data,
                                                     len(self._frozen), _c.byref(h)))
            self._h = h
        return self._h

    def close(self):
        if getattr(self, "_h", None) is not None:
            lib().polar_cuda_puncturer_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:   # interpreter shutdown: the module globals may already be gone
            pass

    def blockLength(self):
        return self._M

    def parentBlockLength(self):
        return self._parent

    def blockOutputPositions(self):
        return [int(p) for p in self._pos]

    @staticmethod
    def _elem(x):
        size = x.element_size() if _is_torch(x) else x.dtype.itemsize
        if size not in (1, 4, 8):
            raise RuntimeError("elements must be 1, 4 or 8 bytes wide (uint8 / int8, float32, float64)")
        return size

    def _run(self, fn, x, in_cols, out_cols, packed=False, out=None, stream=0):
        if x.ndim != 2 or x.shape[1] != in_cols:
            raise RuntimeError("Input vector size mismatch!")
        n = x.shape[0]
        if out is None:
            if _is_torch(x):
                import torch
                out = torch.empty((n, out_cols), dtype=x.dtype, device=x.device)
            else:
                x = _np.ascontiguousarray(x)
                out = _np.empty((n, out_cols), x.dtype)
        ip, mem = _buf(x)
        op, mem2 = _buf(out)
        if mem != mem2:
            raise RuntimeError("input and output must live in the same memory space")
        if packed:
            _check(fn(self._get(), ip, mem, n, op, stream))
        else:
            _check(fn(self._get(), ip, self._elem(x), mem, n, op, stream))
        return out

    def puncture_batch(self, x, out=None, stream=0):
        """[nFrames, parentBlockLength] -> [nFrames, blockLength]"""
        return self._run(lib().polar_cuda_puncture_batch, x, self._parent, self._M, out=out, stream=stream)

    def depuncture_batch(self, x, out=None, stream=0):
        """[nFrames, blockLength] -> [nFrames, parentBlockLength], punctured positions 0 (an erasure as an LLR)"""
        return self._run(lib().polar_cuda_depuncture_batch, x, self._M, self._parent, out=out, stream=stream)

    def puncturePacked_batch(self, x, out=None, stream=0):
        """[nFrames, parentBlockLength/8] uint8 -> [nFrames, blockLength/8] uint8, bits MSB first"""
        if self._M % 8:
            raise ValueError("packed puncturing needs a block length that is a multiple of 8")
        return self._run(lib().polar_cuda_puncture_packed_batch, x, self._parent // 8, self._M // 8, packed=True,
                         out=out, stream=stream)

    def _one(self, batch_fn, vec, length, dtype=None):
        a = _np.ascontiguousarray(vec) if dtype is None else _np.ascontiguousarray(vec, dtype)
        if a.dtype not in (_np.uint8, _np.float32, _np.float64):
            a = a.astype(_np.float64)   # the binding's overload order: double first (puncturer_python.cc:45-60)
        if a.ndim != 1:
            raise RuntimeError("Number of dimensions must be one")
        if a.size != length:
            raise RuntimeError("Input vector size mismatch!")
        return batch_fn(a.reshape(1, -1))[0]

    def puncture(self, vec):
        return self._one(self.puncture_batch, vec, self._parent)

    def depuncture(self, vec):
        return self._one(self.depuncture_batch, vec, self._M)

    def puncturePacked(self, vec):
        return self._one(self.puncturePacked_batch, vec, self._parent // 8, _np.uint8)


class LinkSimulator:
    """The worker loop of the reference's pcsim (src/simulation/simulator.cpp:632-672) on one GPU: payload -> CRC -> encoder ->
    BPSK -> AWGN -> amplifier / quantiser -> decoder -> error counters, every stage a kernel, no host round trip in between.
    `counters` is a torch int64 CUDA tensor of 5 entries that run() ADDS to: runs, block errors, bit errors (over all K
    info bits), reported errors (detector failures), undetected errors."""

    COUNTERS = ("runs", "errors", "biterrors", "reported", "undetected")

    def __init__(self, blockLength, listSize, frozenBitPositions, decoderType="char", crc=0, systematic=True, adaptive=False,
                 device=0, batch=1 << 16):
        kind = PolarDecoder(blockLength, listSize, frozenBitPositions, decoderType, device, adaptive)._kind
        f = _np.ascontiguousarray(frozenBitPositions, _np.uint32)
        self._h = _c.c_void_p()
        self.N, self.K, self.batch, self.kind = int(blockLength), int(blockLength) - len(f), int(batch), kind
        self.kb = (self.K + 7) // 8
        _check(lib().polar_cuda_sim_create(int(device), self.N, f.ctypes.data, len(f), kind, max(int(listSize), 1),
                                           int(bool(systematic)), int(crc), self.batch, _c.byref(self._h)))
        self.is_float = kind in (KIND_FASTSSC_F32, KIND_SCL_F32, KIND_ADAPTIVE_F32, KIND_ADAPTIVE_MIXED)
        self.device = device

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            lib().polar_cuda_sim_destroy(self._h)
            self._h = _c.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:   # interpreter shutdown: the module globals may already be gone
            pass

    def default_amplification(self, ebn0_db):
        """pcsim's defaults: 10 ahead of the 8-bit quantiser (setup.cpp:58), 4 * 10^(Eb/N0 / 10) for float (simulator.cpp:377)"""
        return 4.0 * 10.0 ** (ebn0_db / 10.0) if self.is_float else 10.0

    def run(self, ebn0_db, seed, point, first_frame, n_frames, counters, amplification=None, stream=0):
        if not (_is_torch(counters) and counters.is_cuda and counters.numel() >= 5 and counters.element_size() == 8):
            raise RuntimeError("counters must be a CUDA int64 tensor with 5 entries")
        amp = self.default_amplification(ebn0_db) if amplification is None else float(amplification)
        _check(lib().polar_cuda_sim_run(self._h, float(ebn0_db), amp, int(seed), int(point), int(first_frame), int(n_frames),
                                        counters.data_ptr(), stream))

    def last_batch(self):
        """(llr, sent, decoded, ok) of the last batch as torch tensors that alias the simulator's device buffers."""
        import torch
        p = [_c.c_void_p() for _ in range(4)]
        n = _c.c_size_t()
        _check(lib().polar_cuda_sim_buffers(self._h, _c.byref(p[0]), _c.byref(p[1]), _c.byref(p[2]), _c.byref(p[3]), _c.byref(n)))
        n = n.value

        def view(ptr, shape, dtype, itemsize):
            nbytes = int(_np.prod(shape)) * itemsize

            class _Holder:  # __cuda_array_interface__ over a raw device pointer
                pass
            h = _Holder()
            h.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": dtype, "data": (ptr.value, False), "version": 2}
            h.owner = self
            return torch.as_tensor(h, device=torch.device("cuda", self.device))
        llr = view(p[0], (n, self.N), "<f4" if self.is_float else "|i1", 4 if self.is_float else 1)
        sent = view(p[1], (n, self.kb), "|u1", 1)
        dec = view(p[2], (n, self.kb), "|u1", 1)
        ok = view(p[3], (n,), "|u1", 1)
        return llr, sent, dec, ok

    def launches(self):
        d = _c.c_long(0)
        total = lib().polar_cuda_sim_launches(self._h, _c.byref(d))
        return int(total), int(d.value)


def test_philox(ctr, key):
    """Philox4x32-10 of the device (known-answer hook)."""
    c = _np.ascontiguousarray(ctr, _np.uint32)
    k = _np.ascontiguousarray(key, _np.uint32)
    out = _np.zeros(4, _np.uint32)
    _check(lib().polar_cuda_test_philox(c.ctypes.data, k.ctypes.data, out.ctypes.data))
    return out


def schedule_cache_stats():
    """(entries, hits, misses) of the per-process cache of compiled node programs."""
    h, m = _c.c_long(0), _c.c_long(0)
    n = lib().polar_cuda_schedule_cache_stats(_c.byref(h), _c.byref(m))
    return n, h.value, m.value


def test_fg(a, b, bits):
    """Element-wise device F / G (golden-vector hook)."""
    a = _np.ascontiguousarray(a, _np.int8)
    b = _np.ascontiguousarray(b, _np.int8)
    bits = _np.ascontiguousarray(bits, _np.int8)
    f = _np.empty_like(a)
    g = _np.empty_like(a)
    _check(lib().polar_cuda_test_fg(a.ctypes.data, b.ctypes.data, bits.ctypes.data, a.size, f.ctypes.data,
                                    g.ctypes.data))
    return f, g
