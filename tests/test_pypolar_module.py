"""The reference's pybind11 module `polarcode_python` built WITH the GPU classes (tests/cpp/Makefile: the reference's own
python_bindings.cc / detector / frozen_bits / puncturer bindings + tests/cpp/pypolar_cuda/{decoder,encoder}_python_cuda.cc).
The module is the drop-in a pypolar user sees: same classes and methods as python/bindings/*.cc, plus the type strings
"cuda-char" / "cuda-float" / "cuda-mixed", encoderType "cuda", and decode_batch / encode_batch.  On the GPU the CUDA types are
compared with the CPU types OF THE SAME MODULE, i.e. with the unmodified reference decoders behind the same Python calls."""
import importlib
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "tests", "cpp", "_build")


def _module():
    if os.path.isdir("/root/reference/python/bindings"):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "tests", "cpp")])
    if not os.path.isdir(os.path.join(BUILD, "pypolar")):
        pytest.skip("the pybind11 module was not built (needs the reference sources, dev container only)")
    if BUILD not in sys.path:
        sys.path.insert(0, BUILD)
    return importlib.import_module("pypolar")


def _frames(pp, N, K, crc, n, ebn0=2.0, seed=3, ctype="BB"):
    f = pp.frozen_bits(N, K, 0.0, ctype)
    enc = pp.PolarEncoder(N, f)
    enc.setErrorDetection(crc)
    rng = np.random.default_rng(seed)
    code, info = enc.encode_batch(rng.integers(0, 256, (n, K // 8), dtype=np.uint8))
    sigma = 1.0 / np.sqrt(2.0 * (K / N) * 10.0 ** (ebn0 / 10.0))
    y = (1.0 - 2.0 * np.unpackbits(code, axis=1).astype(np.float32)) + sigma * rng.standard_normal((n, N)).astype(np.float32)
    return f, info, y


def test_module_surface_and_cpu_types():
    """pypolar surface of SURVEY.md 8b + the batched methods; the CPU types work through them unchanged."""
    pp = _module()
    for name in ("PolarEncoder", "PolarDecoder", "Detector", "Puncturer", "frozen_bits"):
        assert hasattr(pp, name)
    N, K = 256, 128
    f, info, y = _frames(pp, N, K, 8, 16, ebn0=6.0)
    dec = pp.PolarDecoder(N, 4, f, "char")
    for m in ("blockLength", "infoLength", "listSize", "setSystematic", "isSystematic", "frozenBits", "getErrorDetectionMode",
              "setErrorDetection", "decode_vector", "decode_batch", "isCuda"):
        assert hasattr(dec, m)
    assert not dec.isCuda() and dec.getErrorDetectionMode() == "CRC-8" and dec.listSize() == 4
    llr = np.clip(np.rint(10 * y), -128, 127).astype(np.int8)
    out, ok = dec.decode_batch(llr)
    assert np.array_equal(out, info) and ok.all()
    assert np.array_equal(dec.decode_vector(llr[3]), info[3])
    with pytest.raises(RuntimeError):
        dec.decode_batch(llr[0])
    with pytest.raises(Exception):
        pp.PolarDecoder(N, 4, f, "cuda-quantum")


@pytest.mark.gpu
@pytest.mark.parametrize("N,K,L,crc,ctype", [(1024, 512, 8, 16, "5G"), (1024, 512, 1, 0, "BB"), (512, 256, 4, 32, "BB"), (256, 128, 2, 8, "BB")])
def test_cuda_char_equals_reference_char(gpu, N, K, L, crc, ctype):
    pp = _module()
    f, info, y = _frames(pp, N, K, crc, 512, ebn0=1.5, ctype=ctype)
    llr = np.clip(np.rint(10 * y), -128, 127).astype(np.int8)
    cpu, gpu_ = pp.PolarDecoder(N, L, f, "char"), pp.PolarDecoder(N, L, f, "cuda-char")
    assert gpu_.isCuda() and gpu_.listSize() == cpu.listSize() and gpu_.getErrorDetectionMode() == cpu.getErrorDetectionMode() == "CRC-8"
    for d in (cpu, gpu_):
        d.setErrorDetection(crc)
    (o1, k1), (o2, k2) = cpu.decode_batch(llr), gpu_.decode_batch(llr)
    assert np.array_equal(o1, o2) and np.array_equal(k1, k2) and 0 < k1.sum()
    # float LLRs into the int8 decoders: both quantise like CharContainer::insertLlr
    (o3, k3), (o4, k4) = cpu.decode_batch(10 * y), gpu_.decode_batch(10 * y)
    assert np.array_equal(o3, o4) and np.array_equal(k3, k4)
    assert np.array_equal(gpu_.decode_vector(llr[5]), cpu.decode_vector(llr[5]))
    for d in (cpu, gpu_):
        d.setSystematic(False)
    if N >= 256:  # the reference's char decoders are broken in non-systematic mode below N = 256
        (o5, k5), (o6, k6) = cpu.decode_batch(llr[:64]), gpu_.decode_batch(llr[:64])
        assert np.array_equal(o5, o6) and np.array_equal(k5, k6)


@pytest.mark.gpu
def test_cuda_float_and_encoder(gpu):
    pp = _module()
    N, K, crc = 256, 128, 8
    f, info, y = _frames(pp, N, K, crc, 256, ebn0=2.0)
    cpu, gpu_ = pp.PolarDecoder(N, 1, f, "float"), pp.PolarDecoder(N, 1, f, "cuda-float")
    for d in (cpu, gpu_):
        d.setErrorDetection(crc)
    llr = (4.0 * 10 ** 0.2 * y).astype(np.float32)
    (o1, k1), (o2, k2) = cpu.decode_batch(llr), gpu_.decode_batch(llr)
    assert np.array_equal(o1, o2) and np.array_equal(k1, k2)
    raw = np.random.default_rng(1).integers(0, 256, (1000, K // 8), dtype=np.uint8)
    for sysm in (True, False):
        e1, e2 = pp.PolarEncoder(N, f), pp.PolarEncoder(N, f, "cuda")
        for e in (e1, e2):
            e.setErrorDetection(crc)
            e.setSystematic(sysm)
        (c1, i1), (c2, i2) = e1.encode_batch(raw), e2.encode_batch(raw)
        assert np.array_equal(c1, c2) and np.array_equal(i1, i2)
        assert np.array_equal(e2.encode_vector(i1[0]), c1[0])


@pytest.mark.gpu
@pytest.mark.parametrize("L", [1, 8])
def test_host_side_detector_gets_all_candidates(gpu, L):
    """A detector the kernels do not fuse (the reference's cmac): the list decoder hands all surviving paths back and
    Detector::check runs on the host in list order -- same decisions as the reference's SclFipChar with the same detector."""
    pp = _module()
    N, K = 512, 256
    f = pp.frozen_bits(N, K, 0.0, "BB")
    enc = pp.PolarEncoder(N, f)
    enc.setErrorDetection(32, "cmac")
    rng = np.random.default_rng(9)
    code, info = enc.encode_batch(rng.integers(0, 256, (256, K // 8), dtype=np.uint8))
    sigma = 1.0 / np.sqrt(2.0 * (K / N) * 10.0 ** (1.5 / 10.0))
    y = (1.0 - 2.0 * np.unpackbits(code, axis=1).astype(np.float32)) + sigma * rng.standard_normal((256, N)).astype(np.float32)
    llr = np.clip(np.rint(10 * y), -128, 127).astype(np.int8)
    cpu, gpu_ = pp.PolarDecoder(N, L, f, "char"), pp.PolarDecoder(N, L, f, "cuda-char")
    for d in (cpu, gpu_):
        d.setErrorDetection(32, "cmac")
    assert gpu_.getErrorDetectionMode() == cpu.getErrorDetectionMode()
    (o1, k1), (o2, k2) = cpu.decode_batch(llr), gpu_.decode_batch(llr)
    assert np.array_equal(o1, o2) and np.array_equal(k1, k2) and k1.sum() > 100
    assert np.array_equal(gpu_.decode_vector(llr[0]), cpu.decode_vector(llr[0]))
