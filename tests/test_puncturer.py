"""Rate matching (SURVEY.md 8b pypolar surface, 8f-4): PolarCode::Puncturer (puncturer.cpp:45-85, puncturer.h:71-100).
CPU part: position rule of the C ABI and of the oracle against the reference's own QA expectations
(python/qa_pypolar_puncturer-style checks restated here) and, where oracle/_ref exists, against the reference
object live.  GPU part: the batched kernels against the oracle, bit exact."""
import numpy as np
import pytest

import antpolarcodes_b200 as ap
from polar_testlib import (have_ref, oracle_depuncture, oracle_puncture, oracle_puncture_packed,
                           oracle_puncturer_positions, ref_puncturer)

CASES = [(96, 128, 64), (1000, 1024, 512), (768, 1024, 256), (200, 256, 100), (48, 64, 16), (1024, 1024, 512),
         (3000, 4096, 1000), (24, 32, 8)]  # punctured length M, parent N, K of the parent code


def _frozen(N, K):
    return ap.frozen_bits(N, K, 0.0, "BB")


@pytest.mark.parametrize("M,N,K", CASES)
def test_positions_rule(M, N, K):
    fz = _frozen(N, K)
    p = ap.Puncturer(M, fz)
    pos = np.array(p.blockOutputPositions(), np.uint32)
    assert p.blockLength() == M and p.parentBlockLength() == N and len(pos) == M
    # the first N - M frozen positions are gone, everything else is kept in ascending order
    assert np.array_equal(pos, np.setdiff1d(np.arange(N, dtype=np.uint32), fz[:N - M]))
    assert np.array_equal(pos, oracle_puncturer_positions(M, fz))
    if have_ref():
        assert np.array_equal(pos, ref_puncturer(M, fz, 0))


def test_too_many_punctured_positions():
    """puncturer.cpp:51-54 throws std::out_of_range (IndexError through pybind11)."""
    fz = _frozen(128, 100)        # 28 frozen bits, 65 -> 63 positions to puncture
    with pytest.raises(IndexError):
        ap.Puncturer(65, fz)
    assert oracle_puncturer_positions(65, fz) is None
    if have_ref():
        assert ref_puncturer(65, fz, 0) is None


@pytest.mark.parametrize("M,N,K", CASES[:5])
def test_oracle_matches_reference_object(M, N, K):
    if not have_ref():
        pytest.skip("oracle/_ref not built")
    fz = _frozen(N, K)
    pos = oracle_puncturer_positions(M, fz)
    rng = np.random.default_rng(M)
    for dt in (np.uint8, np.float32, np.float64):
        x = (rng.standard_normal((5, N)) * 40).astype(dt)
        assert np.array_equal(oracle_puncture(pos, x), ref_puncturer(M, fz, 1, x))
        y = (rng.standard_normal((5, M)) * 40).astype(dt)
        assert np.array_equal(oracle_depuncture(pos, N, y), ref_puncturer(M, fz, 2, y, parent=N))
    bits = rng.integers(0, 256, (5, N // 8), dtype=np.uint8)
    assert np.array_equal(oracle_puncture_packed(pos, bits), ref_puncturer(M, fz, 3, bits))


@pytest.mark.gpu
@pytest.mark.parametrize("M,N,K", CASES)
def test_puncturer_kernels(gpu, M, N, K):
    import torch
    fz = _frozen(N, K)
    p = ap.Puncturer(M, fz)
    pos = oracle_puncturer_positions(M, fz)
    rng = np.random.default_rng(N + M)
    for n in (1, 333):
        for dt in (np.int8, np.uint8, np.float32, np.float64):
            x = (rng.standard_normal((n, N)) * 40).astype(dt)
            want = oracle_puncture(pos, x)
            assert np.array_equal(p.puncture_batch(x), want)
            assert np.array_equal(p.puncture_batch(torch.from_numpy(x).cuda()).cpu().numpy(), want)
            y = (rng.standard_normal((n, M)) * 40).astype(dt)
            want = oracle_depuncture(pos, N, y)
            assert np.array_equal(p.depuncture_batch(y), want)
            assert np.array_equal(p.depuncture_batch(torch.from_numpy(y).cuda()).cpu().numpy(), want)
            assert np.array_equal(p.puncture_batch(p.depuncture_batch(y)), y)   # round trip
        bits = rng.integers(0, 256, (n, N // 8), dtype=np.uint8)
        want = oracle_puncture_packed(pos, bits)
        assert np.array_equal(p.puncturePacked_batch(bits), want)
        assert np.array_equal(p.puncturePacked_batch(torch.from_numpy(bits).cuda()).cpu().numpy(), want)
    # one-vector methods of the binding
    v = rng.standard_normal(N).astype(np.float32)
    assert np.array_equal(p.puncture(v), v[pos])
    assert np.array_equal(p.depuncture(v[pos])[pos], v[pos])
    with pytest.raises(RuntimeError):
        p.puncture(v[:-1])


@pytest.mark.gpu
def test_rate_matched_link(gpu):
    """encode -> puncturePacked -> BPSK/AWGN on the M kept bits -> depuncture (0 = erasure) -> decode: the punctured
    positions are frozen ones, so the decoder still recovers the payload at a comfortable SNR."""
    N, K, M, crc, n = 1024, 512, 900, 16, 512
    fz = _frozen(N, K)
    enc = ap.PolarEncoder(N, fz)
    enc.setSystematic(False)      # frozen positions of u are 0; puncturing them is only harmless for non-systematic
    enc.setErrorDetection(crc)
    dec = ap.PolarDecoder(N, 8, fz, "char")
    dec.setSystematic(False)
    dec.setErrorDetection(crc)
    p = ap.Puncturer(M, fz)
    rng = np.random.default_rng(9)
    info = rng.integers(0, 256, (n, K // 8), dtype=np.uint8)
    code = enc.encode_batch(info)
    tx = p.puncturePacked_batch(code[:, : N // 8]) if M % 8 == 0 else None
    pos = np.array(p.blockOutputPositions())
    bits = np.unpackbits(code, axis=1)[:, pos] if tx is None else np.unpackbits(tx, axis=1)
    y = (1.0 - 2.0 * bits) + 0.45 * rng.standard_normal(bits.shape)
    llr = np.clip(np.rint(10.0 * y), -128, 127).astype(np.int8)
    full = p.depuncture_batch(llr)
    assert full.shape == (n, N) and (np.delete(full, pos, axis=1) == 0).all()
    out, ok = dec.decode_batch(full)
    good = (out == info).all(axis=1)
    assert good.mean() > 0.9 and (ok[good] == 1).all()
