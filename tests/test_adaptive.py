"""Adaptive decoders: Fast-SSC on every frame, the list decoder only where the detector rejects
the Fast-SSC result (AdaptiveChar::decode, adaptive_char.cpp:34-46; AdaptiveFloat, AdaptiveMixed).

CPU part: the oracle composition (Fast-SSC oracle + list oracle on the failed frames) is pinned against committed
outputs of the reference's own Adaptive* objects (tests/golden/ref_adaptive.npz) and, where oracle/_ref exists,
against them live.  GPU part: POLAR_KIND_ADAPTIVE_* through the C ABI against that oracle, bit exact."""
import os

import numpy as np
import pytest

from polar_testlib import (KIND_ADAPTIVE_F32, KIND_ADAPTIVE_I8, KIND_ADAPTIVE_MIXED, awgn_llrs, have_ref,
                           oracle_decode_adaptive, ref_decode_adaptive, stress_llrs)

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

GOLDEN_CASES = [  # kind, N, K, construction, L, crc, frames, Eb/N0 (dB)
    (KIND_ADAPTIVE_I8, 1024, 512, "5G", 8, 16, 48, 1.5),
    (KIND_ADAPTIVE_I8, 256, 128, "BB", 4, 8, 96, 1.0),
    (KIND_ADAPTIVE_I8, 2048, 1024, "BB", 32, 32, 8, 1.0),
    (KIND_ADAPTIVE_I8, 64, 32, "BB", 1, 8, 64, 1.0),      # listSize 1: the Fast-SSC result stands
    (KIND_ADAPTIVE_F32, 1024, 512, "5G", 8, 16, 24, 1.5),
    (KIND_ADAPTIVE_F32, 256, 128, "BB", 4, 8, 64, 1.0),
    (KIND_ADAPTIVE_MIXED, 1024, 512, "5G", 8, 16, 24, 1.5),
    (KIND_ADAPTIVE_MIXED, 256, 128, "BB", 4, 8, 64, 1.0),
]
_FLAVOUR = {KIND_ADAPTIVE_I8: "char", KIND_ADAPTIVE_F32: "float", KIND_ADAPTIVE_MIXED: "mixed"}


def golden_key(case):
    return "ada_%d_%d_%d_%s_%d_%d_%d_%g" % case


def adaptive_llrs(kind, N, frozen, n, crc, ebn0, seed=0xADA):
    return awgn_llrs(N, frozen, n, ebn0, crc, seed=seed, fmt="i8" if kind == KIND_ADAPTIVE_I8 else "f32")[0]


def _frozen(N, K, t):
    import antpolarcodes_b200 as ap  # host-side construction, no device needed
    return ap.frozen_bits(N, K, 0.0, t)


@pytest.mark.parametrize("case", GOLDEN_CASES, ids=golden_key)
def test_oracle_composition_matches_reference_adaptive(case):
    kind, N, K, t, L, crc, n, ebn0 = case
    fz = _frozen(N, K, t)
    llr = adaptive_llrs(kind, N, fz, n, crc, ebn0)
    out, ok, redone = oracle_decode_adaptive(kind, N, fz, llr, L, 1, crc)
    gold = np.load(os.path.join(GOLD, "ref_adaptive.npz"))
    assert np.array_equal(out, gold[golden_key(case) + "_out"])
    assert np.array_equal(ok, gold[golden_key(case) + "_ok"])
    if L > 1:
        assert 0 < redone < n, "the workload must exercise both branches"
    if have_ref():
        ro, rk = ref_decode_adaptive(kind, N, fz, llr, L, 1, crc, fresh=(kind != KIND_ADAPTIVE_I8))
        assert np.array_equal(out, ro) and np.array_equal(ok, rk)


def test_mirror_factory_rules():
    """decoder.cpp:26-48: 'mixed' with two or more paths is the adaptive float decoder, with one path the float
    Fast-SSC; adaptive=True reaches the classes the simulator builds directly."""
    import antpolarcodes_b200 as ap
    fz = _frozen(64, 32, "BB")
    assert ap.PolarDecoder(64, 4, fz, "mixed")._kind == ap.KIND_ADAPTIVE_F32
    assert ap.PolarDecoder(64, 1, fz, "mixed")._kind == ap.KIND_FASTSSC_F32
    assert ap.PolarDecoder(64, 4, fz, "char", adaptive=True)._kind == ap.KIND_ADAPTIVE_I8
    assert ap.PolarDecoder(64, 4, fz, "float", adaptive=True)._kind == ap.KIND_ADAPTIVE_F32
    assert ap.PolarDecoder(64, 4, fz, "mixed", adaptive=True)._kind == ap.KIND_ADAPTIVE_MIXED
    with pytest.raises(RuntimeError):
        ap.PolarDecoder(64, 4, fz, "scan")


# ---------------------------------------------------------------- GPU
GPU_CASES = GOLDEN_CASES + [
    (KIND_ADAPTIVE_I8, 1024, 512, "BB", 8, 16, 256, 2.0),
    (KIND_ADAPTIVE_I8, 8, 4, "BB", 4, 0, 64, 1.0),        # no detector: nothing ever falls through
    (KIND_ADAPTIVE_I8, 32, 16, "BB", 8, 8, 200, 0.0),     # 32-byte rows: the 4-byte gather path
    (KIND_ADAPTIVE_I8, 16, 8, "BB", 2, 8, 333, 0.0),
    (KIND_ADAPTIVE_F32, 32, 16, "BB", 8, 8, 100, 0.0),
    (KIND_ADAPTIVE_MIXED, 512, 256, "BE", 2, 32, 48, 1.0),
    (KIND_ADAPTIVE_I8, 4096, 2048, "BB", 4, 16, 6, 1.0),
]


def _decoder(kind, N, L, fz, crc, sysm=1):
    import antpolarcodes_b200 as ap
    dec = ap.PolarDecoder(N, L, fz, _FLAVOUR[kind], adaptive=True)
    dec.setErrorDetection(crc)
    dec.setSystematic(bool(sysm))
    return dec


@pytest.mark.gpu
@pytest.mark.parametrize("case", GPU_CASES, ids=golden_key)
@pytest.mark.parametrize("sysm", [1, 0])
def test_adaptive_parity(gpu, case, sysm):
    kind, N, K, t, L, crc, n, ebn0 = case
    fz = _frozen(N, K, t)
    dec = _decoder(kind, N, L, fz, crc, sysm)
    total = 0
    for seed in (0xADA, 5):
        llr = adaptive_llrs(kind, N, fz, n, crc, ebn0, seed=seed)
        out, ok = dec.decode_batch(llr)
        eo, ek, redone = oracle_decode_adaptive(kind, N, fz, llr, L, sysm, crc)
        bad = np.where((out != eo).any(axis=1) | (ok != ek))[0]
        assert len(bad) == 0, "%d/%d frames differ from the oracle, first %s" % (len(bad), n, bad[:5])
        total += redone
        assert dec.fallbacks() == total
    if sysm and case in GOLDEN_CASES:  # the committed outputs of the reference's own Adaptive* objects
        gold = np.load(os.path.join(GOLD, "ref_adaptive.npz"))
        llr = adaptive_llrs(kind, N, fz, n, crc, ebn0)
        out, ok = dec.decode_batch(llr)
        assert np.array_equal(out, gold[golden_key(case) + "_out"]) and np.array_equal(ok, gold[golden_key(case) + "_ok"])


@pytest.mark.gpu
def test_adaptive_device_buffers_and_float_into_char(gpu):
    """torch CUDA tensors (zero copy) and float LLRs into AdaptiveChar: quantised once, both decoders read the int8
    copy (AdaptiveChar::setSignal, adaptive_char.cpp:60-64)."""
    import torch
    N, K, L, crc, n = 1024, 512, 8, 16, 2048
    fz = _frozen(N, K, "5G")
    lf = awgn_llrs(N, fz, n, 1.5, crc, seed=11, fmt="f32")[0] * 2.5
    dec = _decoder(KIND_ADAPTIVE_I8, N, L, fz, crc)
    eo, ek, redone = oracle_decode_adaptive(KIND_ADAPTIVE_I8, N, fz, lf, L, 1, crc)
    assert 0 < redone < n
    out, ok = dec.decode_batch(torch.from_numpy(lf).cuda())
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy(), eo) and np.array_equal(ok.cpu().numpy(), ek)
    assert dec.fallbacks() == redone
    o2, k2 = dec.decode_batch(lf)  # host buffers
    assert np.array_equal(o2, eo) and np.array_equal(k2, ek)
    with pytest.raises(ValueError):
        _decoder(KIND_ADAPTIVE_MIXED, N, L, fz, crc).decode_batch(np.zeros((4, N), np.int8))


@pytest.mark.gpu
def test_adaptive_all_fail_and_none_fail(gpu):
    """Every frame falls through (noise only) / none does (noiseless), and host calls larger than one staging chunk."""
    N, K, L, crc = 256, 128, 4, 16
    fz = _frozen(N, K, "BB")
    dec = _decoder(KIND_ADAPTIVE_I8, N, L, fz, crc)
    noise = stress_llrs(N, 300)
    out, ok = dec.decode_batch(noise)
    eo, ek, redone = oracle_decode_adaptive(KIND_ADAPTIVE_I8, N, fz, noise, L, 1, crc)
    assert redone >= 290 and np.array_equal(out, eo) and np.array_equal(ok, ek)
    clean = awgn_llrs(N, fz, 5000, 12.0, crc, seed=3)[0]
    before = dec.fallbacks()
    out, ok = dec.decode_batch(clean)
    eo, ek, redone = oracle_decode_adaptive(KIND_ADAPTIVE_I8, N, fz, clean, L, 1, crc)
    assert redone == 0 and dec.fallbacks() == before and ok.all()
    assert np.array_equal(out, eo) and np.array_equal(ok, ek)


@pytest.mark.gpu
@pytest.mark.parametrize("kind", [KIND_ADAPTIVE_I8, KIND_ADAPTIVE_MIXED])
def test_adaptive_host_pipeline(gpu, kind):
    """HOST-buffer calls split into double-buffered chunks (a ragged last one) give what one device call gives."""
    N, K, L, crc, n = 256, 128, 4, 16, 13001
    fz = _frozen(N, K, "BB")
    llr = adaptive_llrs(kind, N, fz, n, crc, 1.5, seed=21)
    dec = _decoder(kind, N, L, fz, crc)
    os.environ["POLAR_CUDA_CHUNK_MB"] = "1"   # 4096 int8 / 1024 float frames per chunk
    try:
        out, ok = dec.decode_batch(llr)
    finally:
        del os.environ["POLAR_CUDA_CHUNK_MB"]
    eo, ek, redone = oracle_decode_adaptive(kind, N, fz, llr, L, 1, crc)
    assert 0 < redone < n and dec.fallbacks() == redone
    assert np.array_equal(out, eo) and np.array_equal(ok, ek)
