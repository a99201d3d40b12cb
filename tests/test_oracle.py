"""CPU tests of the oracle (oracle/liboracle.so): pinned against the reference's own golden vectors / known
answers, against the committed outputs of the compiled reference (tests/golden/ref_outputs.npz)
and, where oracle/_ref is present, against the reference itself on fresh seeded inputs."""
import json
import os

import numpy as np
import pytest

from polar_testlib import (KIND_FASTSSC_F32, KIND_FASTSSC_I8, KIND_SCL_F32, KIND_SCL_I8, awgn_llrs, have_ref, oracle,
                           oracle_crc_check, oracle_crc_generate, oracle_decode, oracle_encode, oracle_quantize,
                           oracle_tree_stats, pcbench_llrs, ref, ref_decode, ref_decode_fresh, ref_encode,
                           stress_llrs, unpack_bits, workload)

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
REFOUT = np.load(os.path.join(GOLD, "ref_outputs.npz"))


def frozen(N, K, t, d=0.0):
    key = "frozen_%d_%d_%g_%s" % (N, K, d, t)
    if key in REFOUT.files:
        return REFOUT[key]
    import antpolarcodes_b200 as ap  # host-side construction (pinned against the committed sets in test_host.py)
    return ap.frozen_bits(N, K, d, t)


def test_fg_golden_vectors():
    """decodingtest.cpp:504-737: F never outputs 0, -128 -> -127; G saturates."""
    k = json.load(open(os.path.join(GOLD, "kat_fg.json")))
    o = oracle()
    f = [o.oracle_f_i8(a, b) for a, b in zip(k["f"]["left"], k["f"]["right"])]
    assert f == k["f"]["expected"]
    g = [o.oracle_g_i8(a, b, bit) for a, b, bit in zip(k["g"]["left"], k["g"]["right"], k["g"]["bits"])]
    assert g == k["g"]["expected"]


def test_combine_sign_table():
    """decodingtest.cpp:841-901: the combined left bit is negative iff exactly one input is negative."""
    a = np.arange(-128, 128, dtype=np.int8)
    x = (a[:, None] ^ a[None, :]).astype(np.int8)
    assert np.array_equal(x < 0, (a[:, None] < 0) != (a[None, :] < 0))


def test_rep_leaf_sums():
    """decodingtest.cpp:77-156: REP of an iota input equals the clamped true sum (no intermediate saturation)."""
    o = oracle()
    for n in (32, 64):
        x = (np.arange(n) - n // 2 + 1).astype(np.int8)
        assert o.oracle_rep_sum_i8(x.ctypes.data, n) == int(np.clip(int(x.astype(int).sum()), -128, 127))
    # the documented failure mode (fastssc_fip_char.cpp:218-224): 127+127+127+127-128 steps to -1, not 380
    x = np.zeros(32, np.int8)
    x[[0, 16, 8, 4]] = 127
    x[2] = -128
    assert o.oracle_rep_sum_i8(x.ctypes.data, 32) == -1


def test_kat_n8():
    """decodingtest.cpp:1022-1050,1128-1177: N=8, frozen {0,1,2,4}, systematic -> info nibble 0xF."""
    fz = np.array([0, 1, 2, 4], np.uint32)
    llr = np.array([[-5, -6, -4, 1, -4, -5, -7, 2]], np.int8)
    out, ok = oracle_decode(KIND_FASTSSC_I8, 8, fz, llr, 1, 1, 0)
    assert out[0, 0] & 0xF0 == 0xF0
    out, ok = oracle_decode(KIND_SCL_I8, 8, fz, llr, 4, 1, 0)
    assert out[0, 0] & 0xF0 == 0xF0
    for kind, L in ((KIND_SCL_F32, 4), (KIND_FASTSSC_F32, 1)):  # the fp32 list decoder has the same KAT (:1128-1177)
        out, ok = oracle_decode(kind, 8, fz, llr.astype(np.float32), L, 1, 0)
        assert out[0, 0] & 0xF0 == 0xF0


def test_encoder_kats():
    """encodingtest.cpp:26-67: info 0x80 with only the last bit unfrozen -> all-ones codeword, N = 8..1024;
    :69-90 systematic N=16 K=8 all-ones info; qa_pypolar_encoder.py:64-86: u G mod 2."""
    for lg in range(3, 11):
        N = 1 << lg
        fz = np.arange(N - 1, dtype=np.uint32)
        code, _ = oracle_encode(N, fz, np.array([[0x80]], np.uint8), 0, 0)
        assert (code == 0xFF).all()
    rng = np.random.default_rng(3)
    for N, K in ((32, 16), (64, 48), (128, 32), (256, 128), (1024, 512)):
        fz = np.sort(rng.choice(N, N - K, replace=False)).astype(np.uint32)
        info = rng.integers(0, 256, (10, K // 8), dtype=np.uint8)
        G = np.array([[1]], np.uint8)
        for _ in range(int(np.log2(N))):
            G = np.kron(G, np.array([[1, 0], [1, 1]], np.uint8))
        u = np.zeros((10, N), np.uint8)
        u[:, np.setdiff1d(np.arange(N), fz)] = unpack_bits(info, K)
        x = (u.astype(np.int64) @ G.astype(np.int64)) % 2
        code, _ = oracle_encode(N, fz, info, 0, 0)
        assert np.array_equal(unpack_bits(code, N), x)
        # systematic: encode, zero frozen, encode again; info bits reappear at the non-frozen positions
        x2 = x.copy()
        x2[:, fz] = 0
        x2 = (x2 @ G.astype(np.int64)) % 2
        code, _ = oracle_encode(N, fz, info, 1, 0)
        assert np.array_equal(unpack_bits(code, N), x2)
    # with a proper (Bhattacharyya) frozen set the two-pass encoder is systematic: info bits reappear
    for N, K in ((256, 128), (1024, 512), (64, 40)):
        fz = frozen(N, K, "BB")
        info = rng.integers(0, 256, (10, K // 8), dtype=np.uint8)
        code, _ = oracle_encode(N, fz, info, 1, 0)
        assert np.array_equal(unpack_bits(code, N)[:, np.setdiff1d(np.arange(N), fz)], unpack_bits(info, K))


CRC_KATS = [  # python/qa_pypolar_detector.py:33-181
    (8, b"TestFooB", bytes([0xC2])), (8, b"FooBarPolar", bytes([0xA1])),
    (32, b"Test", bytes([0x8C, 0x2D, 0xE2, 0x19])), (32, b"FooBarPolarT", bytes([0x38, 0xAC, 0x62, 0xC9])),
    (16, b"Test", bytes([0x28, 0x88])), (16, b"RIPloPTiger", bytes([0x69, 0x6F])),
]
CRC_CHECKS = [(8, b"ChaoticLama" + bytes([0x67])), (8, b"NeverListenToTheVoid!" + bytes([0x69])),
              (32, b"DisgustinRoastedWhip" + bytes([0xD0, 0x0B, 0xD6, 0xFE])),
              (16, b"DisgustinRoastedWhip" + bytes([0xA3, 0x2B]))]


def test_crc_known_answers():
    for bits, msg, want in CRC_KATS:
        buf = np.frombuffer(msg + bytes(bits // 8), np.uint8)
        got = oracle_crc_generate(bits, buf)
        assert bytes(got[-(bits // 8):]) == want, (bits, msg)
        assert oracle_crc_check(bits, got)
        bad = got.copy()
        bad[0] ^= 1
        assert not oracle_crc_check(bits, bad)
    for bits, msg in CRC_CHECKS:
        assert oracle_crc_check(bits, np.frombuffer(msg, np.uint8)), (bits, msg)
    assert oracle_crc_check(0, np.zeros(4, np.uint8))


def test_quantiser():
    f = np.array([[0.5, 1.5, 2.5, -0.5, -1.5, -2.5, 127.4, 127.6, -128.4, -129.0, 300.0, -300.0, 0.49, -0.51,
                   3.0, 4.0] + [0.0] * 16], np.float32)
    q = oracle_quantize(f)[0, :14]
    assert list(q) == [0, 2, 2, 0, -2, -2, 127, 127, -128, -128, 127, -128, 0, -1]
    if have_ref():
        import ctypes
        r = ref()
        r.polar_ref_quantize.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
        rng = np.random.default_rng(0)
        for N in (8, 16, 32, 256):
            x = (rng.standard_normal(N) * 60).astype(np.float32)
            x[:6] = [np.nan, np.inf, -np.inf, 3e9, -3e9, 2147483520.0]
            o = np.zeros(N, np.int8)
            r.polar_ref_quantize(x.ctypes.data, o.ctypes.data, N)
            assert np.array_equal(o, oracle_quantize(x.reshape(1, -1))[0]), N


def _cases():
    for key in REFOUT.files:
        if key.startswith("dec_") and key.endswith("_out"):
            p = key[:-4].split("_")
            yield (int(p[1]), int(p[2]), int(p[3]), p[4], int(p[5]), int(p[6]), p[7], int(p[8]))


@pytest.mark.parametrize("kind,N,K,t,L,crc,wname,sysm", list(_cases()))
def test_decoders_against_committed_reference_outputs(kind, N, K, t, L, crc, wname, sysm):
    fz = frozen(N, K, t)
    key = "dec_%d_%d_%d_%s_%d_%d_%s_%d" % (kind, N, K, t, L, crc, wname, sysm)
    n = REFOUT[key + "_ok"].shape[0]
    llr = workload(wname, kind, N, fz, n, crc)
    out, ok = oracle_decode(kind, N, fz, llr, L, sysm, crc)
    assert np.array_equal(out, REFOUT[key + "_out"]) and np.array_equal(ok, REFOUT[key + "_ok"])


def test_encoder_against_committed_reference_outputs():
    n = 0
    for key in REFOUT.files:
        if key.startswith("enc_") and key.endswith("_in"):
            p = key[:-3].split("_")
            N, K, t, sysm, crc = int(p[1]), int(p[2]), p[3], int(p[4]), int(p[5])
            code, info = oracle_encode(N, frozen(N, K, t), REFOUT[key], sysm, crc)
            assert np.array_equal(code, REFOUT[key[:-3] + "_code"]) and np.array_equal(info, REFOUT[key[:-3] + "_info"])
            n += 1
    assert n >= 5


@pytest.mark.skipif(not have_ref(), reason="oracle/_ref not built here")
@pytest.mark.parametrize("kind,N,K,t,L,crc,n", [
    (KIND_SCL_I8, 1024, 512, "5G", 8, 16, 40), (KIND_SCL_I8, 256, 128, "BB", 4, 8, 100),
    (KIND_SCL_I8, 64, 32, "BB", 16, 0, 100), (KIND_FASTSSC_I8, 1024, 512, "BB", 1, 0, 200),
    (KIND_FASTSSC_I8, 4096, 1024, "BB", 1, 0, 20), (KIND_FASTSSC_I8, 64, 40, "BB", 1, 8, 200),
    (KIND_FASTSSC_F32, 256, 128, "BB", 1, 0, 200), (KIND_FASTSSC_F32, 1024, 512, "5G", 1, 16, 50),
    (KIND_FASTSSC_F32, 32, 11, "BB", 1, 0, 200), (KIND_FASTSSC_F32, 16, 8, "BB", 1, 0, 200),
    (KIND_FASTSSC_F32, 8, 4, "BB", 1, 0, 200), (KIND_FASTSSC_F32, 512, 384, "BB", 1, 32, 50),
    (KIND_SCL_F32, 256, 128, "BB", 8, 8, 40), (KIND_SCL_F32, 128, 64, "BB", 2, 8, 60),
    (KIND_SCL_F32, 32, 16, "BB", 4, 8, 60)])
def test_oracle_against_live_reference(kind, N, K, t, L, crc, n):
    fz = frozen(N, K, t)
    for seed, wname in ((11, "awgn0"), (12, "pcbench"), (13, "stress")):
        llr = workload(wname, kind, N, fz, n, crc, seed=seed)
        if kind == KIND_SCL_F32:
            o1, k1 = ref_decode_fresh(kind, N, fz, llr, L, 1, crc)
            o2, k2 = oracle_decode(kind, N, fz, llr, L, 1, crc)
            assert np.array_equal(o1, o2) and np.array_equal(k1, k2), seed
            continue
        o1, k1, _ = ref_decode(kind, N, fz, llr, L, 1, crc)
        o2, k2 = oracle_decode(kind, N, fz, llr, L, 1, crc)
        assert np.array_equal(o1, o2) and np.array_equal(k1, k2), seed


def test_tree_histograms_match_survey():
    """SURVEY.md 8d node tables (computed from the reference's construction + node-selection rules)."""
    s = oracle_tree_stats(KIND_FASTSSC_I8, 1024, frozen(1024, 512, "BB"))
    assert s["F"] == (59, 2404) and s["G"] == (49, 2300) and s["C"] == (49, 2300)
    assert s["G_R1"] == (10, 104) and s["G_0R"] == (8, 208) and s["REP"] == (21, 224) and s["REP2"] == (1, 64)
    assert s["SPC"] == (23, 396) and s["ZERO_SPC"] == (2, 16) and s["ZERO_ONE"] == (3, 12)
    s = oracle_tree_stats(KIND_SCL_I8, 1024, frozen(1024, 512, "5G"))
    assert s["F"] == (82, 2742) and s["R0"] == (14, 118) and s["REP"] == (28, 396)
    assert s["R1"] == (15, 146) and s["SPC"] == (26, 364)
    s = oracle_tree_stats(KIND_SCL_I8, 2048, frozen(2048, 1024, "BB"))
    assert s["F"] == (136, 6072) and s["R0"] == (27, 304) and s["REP"] == (36, 712)
    assert s["R1"] == (30, 548) and s["SPC"] == (44, 484)
