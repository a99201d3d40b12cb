"""Soft outputs of the int8 Fast-SSC decoder: Decoder::getSoftCodeword / getSoftInformation
(src/polarcode/decoding/decoder.cpp:147-152) = the N byte partial sums the root node leaves in the bit container.
oracle restatement <-> committed outputs of the compiled reference (tests/golden/ref_soft.npz, make_golden_soft.py) <-> live
reference; on the GPU: polar_cuda_decode_soft_batch <-> oracle and reference, byte for byte."""
import os

import numpy as np
import pytest

import antpolarcodes_b200 as ap
from polar_testlib import KIND_FASTSSC_I8, have_ref, oracle_fastssc_i8_soft, ref_soft, workload

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = np.load(os.path.join(ROOT, "tests", "golden", "ref_soft.npz"))
KEYS = sorted(k[:-4] for k in GOLD.files if k.endswith("_llr"))


@pytest.mark.parametrize("key", KEYS)
def test_oracle_soft_matches_reference_fixture(key):
    _, N, K, t, _ = key.split("_")
    N, K = int(N), int(K)
    fz = ap.frozen_bits(N, K, 0.0, t)
    llr, soft, info = GOLD[key + "_llr"], GOLD[key + "_soft"], GOLD[key + "_info"]
    mine = oracle_fastssc_i8_soft(N, fz, llr)
    assert np.array_equal(mine, soft)
    pos = np.setdiff1d(np.arange(N), fz)                       # getSoftInformation = the soft codeword at the info positions
    assert np.array_equal(mine[:, pos], info)
    # the sign of a soft bit is the hard decision of the decoder (systematic: the info bits themselves)
    from polar_testlib import oracle_decode, unpack_bits
    out, _ = oracle_decode(KIND_FASTSSC_I8, N, fz, llr, 1, 1, 0)
    assert np.array_equal((mine[:, pos] < 0).astype(np.uint8), unpack_bits(out, K))


@pytest.mark.skipif(not have_ref(), reason="oracle/_ref not built")
@pytest.mark.parametrize("N,K,t", [(512, 448, "BB"), (128, 16, "BB"), (4096, 1024, "BB"), (1024, 768, "BE"), (16, 8, "BB")])
def test_oracle_soft_matches_live_reference(N, K, t):
    fz = ap.frozen_bits(N, K, 0.0, t)
    for w in ("awgn", "awgn0", "pcbench", "stress"):
        llr = workload(w, KIND_FASTSSC_I8, N, fz, 48, 0)
        soft, info = ref_soft(KIND_FASTSSC_I8, N, fz, llr)
        assert np.array_equal(oracle_fastssc_i8_soft(N, fz, llr), soft), w


SOFT_GPU = [(1024, 512, "BB", 2048), (1024, 512, "5G", 512), (256, 128, "BB", 1000), (2048, 1024, "BB", 200), (4096, 1024, "BB", 70),
            (512, 448, "BB", 300), (128, 16, "BB", 300), (64, 40, "BB", 300), (32, 11, "BB", 300), (16, 8, "BB", 300), (8, 4, "BB", 300),
            (1024, 768, "BE", 200), (1024, 64, "5G", 200)]


@pytest.mark.gpu
@pytest.mark.parametrize("N,K,t,n", SOFT_GPU)
def test_gpu_soft_outputs(gpu, N, K, t, n):
    fz = ap.frozen_bits(N, K, 0.0, t)
    dec = ap.PolarDecoder(N, 1, fz, "char")
    dec.setErrorDetection(0)
    pos = np.setdiff1d(np.arange(N), fz)
    for w in ("awgn", "awgn0", "pcbench", "stress"):
        llr = workload(w, KIND_FASTSSC_I8, N, fz, n, 0)
        soft, info = dec.soft_batch(llr)
        want = oracle_fastssc_i8_soft(N, fz, llr)
        bad = np.where((soft != want).any(axis=1))[0]
        assert len(bad) == 0, "%s: %d/%d frames differ from the oracle, first %s" % (w, len(bad), n, bad[:5])
        assert np.array_equal(info, want[:, pos])
        if have_ref():
            rs, ri = ref_soft(KIND_FASTSSC_I8, N, fz, llr[:256])
            assert np.array_equal(soft[:256], rs) and np.array_equal(info[:256], ri), w
        out, _ = dec.decode_batch(llr)                           # the throughput kernel's decisions = the signs of the soft bits
        assert np.array_equal(np.packbits((info < 0).astype(np.uint8), axis=1, bitorder="big")[:, :out.shape[1]], out)


@pytest.mark.gpu
def test_gpu_soft_device_buffers_and_errors(gpu):
    import torch
    N, K = 1024, 512
    fz = ap.frozen_bits(N, K, 0.0, "BB")
    dec = ap.PolarDecoder(N, 1, fz, "char")
    dec.setErrorDetection(0)
    llr = workload("awgn", KIND_FASTSSC_I8, N, fz, 3000, 0)
    soft, info = dec.soft_batch(torch.from_numpy(llr).cuda())
    want = oracle_fastssc_i8_soft(N, fz, llr)
    assert np.array_equal(soft.cpu().numpy(), want) and info.shape == (3000, K)
    s2, none = dec.soft_batch(llr[:5], want_info=False)
    assert none is None and np.array_equal(s2, want[:5])
    lst = ap.PolarDecoder(N, 8, fz, "char")
    with pytest.raises(RuntimeError):
        lst.soft_batch(llr[:4])                                  # list decoders: not built, clear error
