"""GPU parity: the CUDA path, called through the C ABI, against the CPU oracle (and against the compiled
reference where oracle/_ref is present) on the same seeded inputs.  Bit-exact: info bytes and verdicts."""
import os

import numpy as np
import pytest

import antpolarcodes_b200 as ap
from polar_testlib import (KIND_FASTSSC_F32, KIND_FASTSSC_I8, KIND_SCL_F32, KIND_SCL_I8, awgn_llrs, have_ref,
                           oracle_decode, oracle_encode, pcbench_llrs, ref_decode, ref_decode_ex, ref_decode_fresh, ref_encode,
                           stress_llrs, workload)

pytestmark = pytest.mark.gpu
THREADS = os.cpu_count() or 1


def _workloads(N, frozen, n, crc):
    return [("awgn2dB", awgn_llrs(N, frozen, n, 2.0, crc)[0]),
            ("awgn0dB", awgn_llrs(N, frozen, n, 0.0, crc, seed=7)[0]),
            ("pcbench", pcbench_llrs(N, n)),
            ("stress", stress_llrs(N, n))]


def _compare(dec, kind, N, frozen, L, sysm, crc, llr, tag):
    out, ok = dec.decode_batch(llr)
    eo, ek = oracle_decode(kind, N, frozen, llr, L, sysm, crc)
    bad = np.where((out != eo).any(axis=1) | (ok != ek))[0]
    assert len(bad) == 0, "%s: %d/%d frames differ from the oracle, first %s" % (tag, len(bad), len(llr), bad[:5])
    if have_ref() and (sysm or N >= 256):
        # the reference's fp32 list decoder depends on its history: fresh object per frame; all host threads
        ro, rk, _ = ref_decode_ex(kind, N, frozen, llr, L, sysm, crc, THREADS, 1, fresh=(kind == KIND_SCL_F32))
        bad = np.where((out != ro).any(axis=1) | (ok != rk))[0]
        assert len(bad) == 0, "%s: %d frames differ from the reference, first %s" % (tag, len(bad), bad[:5])


SCL_CASES = [
    # N, K, construction, L, crc, frames
    (1024, 512, "5G", 8, 16, 96),
    (1024, 512, "BB", 8, 0, 48),
    (2048, 1024, "BB", 32, 32, 12),
    (256, 128, "BB", 4, 8, 128),
    (128, 64, "BB", 2, 8, 128),
    (64, 32, "BB", 16, 0, 128),
    (512, 384, "BB", 8, 32, 64),
    (1024, 128, "5G", 8, 8, 48),
    (32, 16, "BB", 8, 8, 128),
    (8, 4, "BB", 4, 0, 128),
    (4096, 2048, "BB", 4, 16, 8),
    (1024, 512, "BE", 1, 16, 32),
    (4096, 3072, "BB", 32, 32, 4),   # the largest code and list the library accepts
    (4096, 256, "BE", 32, 8, 4),
]


@pytest.mark.parametrize("N,K,ctype,L,crc,n", SCL_CASES)
@pytest.mark.parametrize("sysm", [1, 0])
def test_scl_i8_parity(gpu, N, K, ctype, L, crc, n, sysm):
    frozen = ap.frozen_bits(N, K, 0.0, ctype)
    dec = ap.PolarDecoder(N, L, frozen, "char")
    dec._kind = ap.KIND_SCL_I8  # also for L == 1 (Decoding::create would pick Fast-SSC)
    dec.setErrorDetection(crc)
    dec.setSystematic(bool(sysm))
    for name, llr in _workloads(N, frozen, n, crc):
        _compare(dec, KIND_SCL_I8, N, frozen, L, sysm, crc, llr, "scl N=%d L=%d %s sys=%d" % (N, L, name, sysm))


FASTSSC_CASES = [
    (1024, 512, "BB", 0, 256), (1024, 512, "5G", 16, 128), (256, 128, "BB", 8, 256), (2048, 1024, "BB", 32, 64),
    (4096, 1024, "BB", 0, 32), (512, 448, "BB", 0, 128), (128, 16, "BB", 8, 128), (64, 40, "BB", 8, 256),
    (32, 11, "BB", 0, 256), (16, 8, "BB", 0, 256), (8, 4, "BB", 0, 256), (1024, 768, "BE", 16, 64),
    (1024, 64, "5G", 0, 64),
]


@pytest.mark.parametrize("N,K,ctype,crc,n", FASTSSC_CASES)
@pytest.mark.parametrize("sysm", [1, 0])
def test_fastssc_i8_parity(gpu, N, K, ctype, crc, n, sysm):
    frozen = ap.frozen_bits(N, K, 0.0, ctype)
    dec = ap.PolarDecoder(N, 1, frozen, "char")
    dec.setErrorDetection(crc)
    dec.setSystematic(bool(sysm))
    for name, llr in _workloads(N, frozen, n, crc):
        _compare(dec, KIND_FASTSSC_I8, N, frozen, 1, sysm, crc, llr, "fastssc N=%d %s sys=%d" % (N, name, sysm))


FASTSSC_F32_CASES = [
    (256, 128, "BB", 0, 256), (1024, 512, "BB", 8, 128), (1024, 512, "5G", 16, 64), (2048, 1024, "BB", 0, 64),
    (4096, 1024, "BB", 0, 32), (512, 448, "BB", 0, 128), (128, 16, "BB", 8, 128), (64, 40, "BB", 8, 256),
    (32, 11, "BB", 0, 256), (16, 8, "BB", 0, 256), (8, 4, "BB", 0, 256), (1024, 768, "BE", 16, 64),
    (1024, 64, "5G", 0, 64), (256, 192, "BB", 32, 100), (128, 64, "BE", 8, 100),
]


def _float_workloads(kind, N, frozen, n, crc):
    return [(w, workload(w, kind, N, frozen, n, crc)) for w in ("awgn", "awgn0", "pcbench", "stress")]


@pytest.mark.parametrize("N,K,ctype,crc,n", FASTSSC_F32_CASES)
@pytest.mark.parametrize("sysm", [1, 0])
def test_fastssc_f32_parity(gpu, N, K, ctype, crc, n, sysm):
    """FastSscAvxFloat: bit-exact (every sum in the reference's evaluation order)."""
    frozen = ap.frozen_bits(N, K, 0.0, ctype)
    dec = ap.PolarDecoder(N, 1, frozen, "float")
    dec.setErrorDetection(crc)
    dec.setSystematic(bool(sysm))
    for name, llr in _float_workloads(KIND_FASTSSC_F32, N, frozen, n, crc):
        _compare(dec, KIND_FASTSSC_F32, N, frozen, 1, sysm, crc, llr, "fastssc f32 N=%d %s sys=%d" % (N, name, sysm))


SCL_F32_CASES = [
    # N, K, construction, L, crc, frames
    (1024, 512, "5G", 8, 16, 48), (1024, 512, "BB", 4, 0, 32), (256, 128, "BB", 8, 8, 64), (128, 64, "BB", 2, 8, 64),
    (64, 32, "BB", 16, 0, 64), (512, 384, "BB", 8, 32, 32), (32, 16, "BB", 8, 8, 64), (8, 4, "BB", 4, 0, 64),
    (2048, 1024, "BB", 32, 32, 8), (4096, 2048, "BB", 4, 16, 8), (1024, 512, "BE", 1, 16, 32), (16, 8, "BB", 32, 0, 64),
    (4096, 3072, "BB", 32, 32, 3),
]


@pytest.mark.parametrize("N,K,ctype,L,crc,n", SCL_F32_CASES)
@pytest.mark.parametrize("sysm", [1, 0])
def test_scl_f32_parity(gpu, N, K, ctype, L, crc, n, sysm):
    """SclAvxFloat with every frame decoded by a fresh list (the reference object carries path 0's metric over
    from frame to frame, scl_avx_float.cpp:46-56,102-108, so only a fresh object per frame is a function of
    the frame): float sums in the reference's order, ties through its selection sort."""
    frozen = ap.frozen_bits(N, K, 0.0, ctype)
    dec = ap.PolarDecoder(N, L, frozen, "float")
    dec._kind = ap.KIND_SCL_F32
    dec.setErrorDetection(crc)
    dec.setSystematic(bool(sysm))
    for name, llr in _float_workloads(KIND_SCL_F32, N, frozen, n, crc):
        _compare(dec, KIND_SCL_F32, N, frozen, L, sysm, crc, llr, "scl f32 N=%d L=%d %s sys=%d" % (N, L, name, sysm))


@pytest.mark.parametrize("N,K,L,n", [(64, 32, 1, 200_000), (128, 64, 4, 90_000)])
def test_host_buffer_pipeline(gpu, monkeypatch, N, K, L, n):
    """HOST-buffer calls are split into grid-wave chunks and double buffered (copy under compute): the result must
    not depend on the chunking, including a ragged last chunk."""
    frozen = ap.frozen_bits(N, K, 0.0, "BB")
    dec = ap.PolarDecoder(N, L, frozen, "char")
    dec.setErrorDetection(8)
    kind = KIND_SCL_I8 if L > 1 else KIND_FASTSSC_I8
    rng = np.random.default_rng(5)
    base, _ = awgn_llrs(N, frozen, 4096, 1.0, 8)
    llr = base[rng.integers(0, len(base), n)]
    llr[::7] = np.clip(llr[::7].astype(np.int16) + rng.integers(-20, 21, (len(llr[::7]), N)), -128, 127).astype(np.int8)
    monkeypatch.setenv("POLAR_CUDA_PIPELINE", "0")
    o0, k0 = dec.decode_batch(llr)
    monkeypatch.setenv("POLAR_CUDA_PIPELINE", "1")
    monkeypatch.setenv("POLAR_CUDA_CHUNK_MB", "1")
    monkeypatch.setenv("POLAR_CUDA_FIRST_CHUNK_MB", "1")
    launches = dec._get().launches()
    o1, k1 = dec.decode_batch(llr)
    assert dec._get().launches() - launches >= 2, "the batch was not split"
    assert np.array_equal(o0, o1) and np.array_equal(k0, k1)
    for sl in (slice(0, 256), slice(n - 256, n)):
        eo, ek = oracle_decode(kind, N, frozen, llr[sl], L, 1, 8)
        assert np.array_equal(o1[sl], eo) and np.array_equal(k1[sl], ek)


def test_float_llr_into_int8_decoder(gpu):
    """decode_vector(const float*) on a char decoder quantises like CharContainer::insertLlr."""
    N, K = 1024, 512
    frozen = ap.frozen_bits(N, K, 0.0, "BB")
    rng = np.random.default_rng(5)
    f = (rng.standard_normal((64, N)) * 30).astype(np.float32)
    f[0, :8] = [0.5, 1.5, 2.5, -0.5, -1.5, 127.5, -128.5, 1e9]
    f[1, :4] = [np.nan, np.inf, -np.inf, 126.5]
    dec = ap.PolarDecoder(N, 1, frozen, "char")
    dec.setErrorDetection(0)
    out, ok = dec.decode_batch(f)
    from polar_testlib import oracle_quantize
    q = oracle_quantize(f)
    eo, ek = oracle_decode(KIND_FASTSSC_I8, N, frozen, q, 1, 1, 0)
    assert np.array_equal(out, eo) and np.array_equal(ok, ek)


ENC_CASES = [(1024, 512, "BB"), (256, 128, "BB"), (2048, 1024, "BB"), (4096, 3072, "BB"), (64, 24, "BB"),
             (32, 16, "BB"), (16, 8, "BB"), (8, 8, "BB"), (1024, 40, "5G"), (128, 120, "BE")]


@pytest.mark.parametrize("N,K,ctype", ENC_CASES)
@pytest.mark.parametrize("sysm", [1, 0])
@pytest.mark.parametrize("crc", [0, 8, 16, 32])
def test_encoder_parity(gpu, N, K, ctype, sysm, crc):
    if crc and (K % 8 or K < crc + 8 or (crc == 32 and (K // 8) % 4)):
        pytest.skip("CRC does not fit this K")
    frozen = ap.frozen_bits(N, K, 0.0, ctype)
    enc = ap.PolarEncoder(N, frozen)
    enc.setSystematic(bool(sysm))
    enc.setErrorDetection(crc)
    rng = np.random.default_rng(N + K + crc)
    info = rng.integers(0, 256, (257, (K + 7) // 8), dtype=np.uint8)
    if K % 8:
        info[:, -1] &= (0xFF << (8 - K % 8)) & 0xFF
    mine = info.copy()
    code = enc.encode_batch(mine)
    ecode, einfo = oracle_encode(N, frozen, info, sysm, crc)
    assert np.array_equal(mine, einfo), "CRC bytes written in place differ"
    assert np.array_equal(code, ecode)
    if have_ref():
        rcode, rinfo = ref_encode(N, frozen, info, sysm, crc)
        assert np.array_equal(code, rcode) and np.array_equal(mine, rinfo)


@pytest.mark.parametrize("N,K,ctype", [(256, 128, "BB"), (1024, 512, "5G"), (1024, 992, "BB"), (512, 96, "BE"), (64, 32, "BB"), (1024, 44 * 8, "5G")])
@pytest.mark.parametrize("crc", [0, 16])
def test_encoder_device_buffers_aligned_and_not(gpu, N, K, ctype, crc):
    """DEVICE buffers: the one-frame-per-thread kernel with 128/256-bit accesses when the rows are aligned, the staged kernel when the
    caller's pointers are not (an offset view of a larger allocation); several thousand frames so that the grid-stride loop
    wraps.  Both against the oracle."""
    import torch
    frozen = ap.frozen_bits(N, K, 0.0, ctype)
    kb, nb, n = K // 8, N // 8, 40000 + 77
    rng = np.random.default_rng(N * 3 + K + crc)
    info = rng.integers(0, 256, (n, kb), dtype=np.uint8)
    ecode, einfo = oracle_encode(N, frozen, info, 1, crc)
    enc = ap.PolarEncoder(N, frozen)
    enc.setErrorDetection(crc)
    for off_in, off_out in ((0, 0), (4, 0), (0, 4), (1, 3)):
        raw_in = torch.zeros(n * kb + 64, dtype=torch.uint8, device="cuda")
        raw_out = torch.zeros(n * nb + 64, dtype=torch.uint8, device="cuda")
        d_info = raw_in[off_in:off_in + n * kb].view(n, kb)
        d_info.copy_(torch.from_numpy(info))
        d_code = raw_out[off_out:off_out + n * nb].view(n, nb)
        enc.encode_batch(d_info, out=d_code)
        torch.cuda.synchronize()
        assert np.array_equal(d_code.cpu().numpy(), ecode), (off_in, off_out)
        assert np.array_equal(d_info.cpu().numpy(), einfo), (off_in, off_out)
        assert int(raw_out[:off_out].sum()) == 0 and int(raw_out[off_out + n * nb:].sum()) == 0   # nothing outside the rows


def test_scl_f32_llr_rows_not_16_byte_aligned(gpu):
    """DEVICE float LLRs at an address that is only 4-byte aligned: the fp32 list decoder must not use its 128-bit channel loads."""
    import torch
    N, K, L, crc, n = 256, 128, 4, 8, 300
    frozen = ap.frozen_bits(N, K, 0.0, "BB")
    llr, _ = awgn_llrs(N, frozen, n, 2.0, crc, fmt="f32")
    dec = ap.PolarDecoder(N, L, frozen, "float")
    dec.setErrorDetection(crc)
    want, wok = dec.decode_batch(llr)
    raw = torch.zeros(n * N + 8, dtype=torch.float32, device="cuda")
    for off in (1, 2, 3):
        view = raw[off:off + n * N].view(n, N)
        view.copy_(torch.from_numpy(llr))
        got, gok = dec.decode_batch(view)
        assert np.array_equal(got.cpu().numpy(), want) and np.array_equal(gok.cpu().numpy(), wok), off
    eo, ek = oracle_decode(KIND_SCL_F32, N, frozen, llr, L, 1, crc)
    assert np.array_equal(want, eo) and np.array_equal(wok, ek)


@pytest.mark.parametrize("L,crc,ctype", [(1, 0, "BB"), (8, 16, "5G")])
def test_decode_device_out_any_alignment(gpu, L, crc, ctype):
    """DEVICE output rows at 16-, 4- and 1-byte aligned addresses: the epilogues pick the widest store the buffer allows and
    write nothing outside the rows."""
    import torch
    N, K, n = 1024, 512, 1000 + 7
    frozen = ap.frozen_bits(N, K, 0.0, ctype)
    llr, _ = awgn_llrs(N, frozen, n, 2.0, crc)
    dec = ap.PolarDecoder(N, L, frozen, "char")
    dec.setErrorDetection(crc)
    want, wok = dec.decode_batch(llr)
    d_llr = torch.from_numpy(llr).cuda()
    kb = K // 8
    for off in (0, 16, 4, 1, 7):
        raw = torch.full((n * kb + 64,), 0xEE, dtype=torch.uint8, device="cuda")
        out = raw[off:off + n * kb].view(n, kb)
        ok = torch.empty(n, dtype=torch.uint8, device="cuda")
        dec.decode_batch(d_llr, out=out, ok=ok)
        torch.cuda.synchronize()
        assert np.array_equal(out.cpu().numpy(), want) and np.array_equal(ok.cpu().numpy(), wok), off
        assert bool((raw[:off] == 0xEE).all()) and bool((raw[off + n * kb:] == 0xEE).all()), off


def test_primitives_exhaustive(gpu):
    """F and G over all 256 x 256 int8 pairs, both bit values (superset of decodingtest.cpp:504-737)."""
    from polar_testlib import oracle
    a = np.repeat(np.arange(-128, 128, dtype=np.int8), 256)
    b = np.tile(np.arange(-128, 128, dtype=np.int8), 256)
    o = oracle()
    for bitval in (0, -128, 127, -1):
        bits = np.full(a.shape, bitval, np.int8)
        f, g = ap.test_fg(a, b, bits)
        ef = np.array([o.oracle_f_i8(int(x), int(y)) for x, y in zip(a[::97], b[::97])], np.int8)
        assert np.array_equal(f[::97], ef)
        # full-width check against a numpy restatement of the same formulas
        ai, bi = a.astype(np.int32), b.astype(np.int32)
        m = np.minimum(np.clip(np.abs(np.maximum(ai, -127)), 1, 127), np.clip(np.abs(np.maximum(bi, -127)), 1, 127))
        nf = np.where((ai < 0) != (bi < 0), -m, m).astype(np.int8)
        assert np.array_equal(f, nf)
        ng = np.clip(bi - ai if bitval < 0 else bi + ai, -128, 127).astype(np.int8)
        assert np.array_equal(g, ng)


def test_empty_and_ragged_batches(gpu):
    """Zero frames is a no-op for every entry point; batch sizes that do not fill a warp / a block / a grid wave
    (1, 31, 33, one more than a wave) decode like the same frames inside a large batch."""
    N, K, crc = 256, 128, 8
    fz = ap.frozen_bits(N, K, 0.0, "BB")
    enc = ap.PolarEncoder(N, fz)
    enc.setErrorDetection(crc)
    assert enc.encode_batch(np.zeros((0, K // 8), np.uint8)).shape == (0, N // 8)
    for L, flavour, adaptive in ((1, "char", False), (4, "char", False), (1, "float", False), (4, "float", False),
                                 (4, "char", True)):
        dec = ap.PolarDecoder(N, L, fz, flavour, adaptive=adaptive)
        dec.setErrorDetection(crc)
        dt = np.float32 if flavour == "float" else np.int8
        out, ok = dec.decode_batch(np.zeros((0, N), dt))
        assert out.shape == (0, K // 8) and ok.shape == (0,)
        wave = dec.info().gridBlocks * dec.info().warpsPerBlock * (32 // L if L > 1 and not adaptive else 32)
        n = wave + 1
        llr = awgn_llrs(N, fz, 512, 1.0, crc, fmt="f32" if flavour == "float" else "i8")[0]
        llr = np.ascontiguousarray(np.tile(llr, ((n + 511) // 512, 1))[:n])
        big, bigk = dec.decode_batch(llr)
        for m in (1, 31, 33, 100):
            o, k = dec.decode_batch(llr[:m])
            assert np.array_equal(o, big[:m]) and np.array_equal(k, bigk[:m]), (L, flavour, adaptive, m)
        assert np.array_equal(big[n - 1], big[(n - 1) % 512]) and bigk[n - 1] == bigk[(n - 1) % 512]
    p = ap.Puncturer(200, fz)
    assert p.depuncture_batch(np.zeros((0, 200), np.int8)).shape == (0, N)


# Many distinct frames against the COMPILED REFERENCE (not the oracle restatement), all host threads: tie-break differences at
# the 1-in-10^4 level would show here (96 / 12 / 16 frames were too few).
MANY = [
    # kind, N, K, construction, L, crc, frames per workload
    (KIND_SCL_I8, 1024, 512, "5G", 8, 16, 32768),     # the headline code
    (KIND_SCL_I8, 2048, 1024, "BB", 32, 32, 1024),    # C4
    (KIND_SCL_F32, 1024, 512, "5G", 8, 16, 512),      # fp32 list decoder, fresh reference object per frame
    (KIND_FASTSSC_I8, 1024, 512, "BB", 1, 0, 65536),  # C1
    (KIND_FASTSSC_F32, 256, 128, "BB", 1, 0, 65536),  # C2
]


@pytest.mark.parametrize("kind,N,K,ctype,L,crc,n", MANY)
def test_many_frames_against_the_compiled_reference(gpu, kind, N, K, ctype, L, crc, n):
    if not have_ref():
        pytest.skip("oracle/_ref not built")
    frozen = ap.frozen_bits(N, K, 0.0, ctype)
    f32 = kind in (KIND_FASTSSC_F32, KIND_SCL_F32)
    dec = ap.PolarDecoder(N, L, frozen, "float" if f32 else "char")
    dec._kind = kind
    dec.setErrorDetection(crc)
    for name in ("awgn", "awgn0", "pcbench", "stress"):
        llr = workload(name, kind, N, frozen, n, crc, seed=1000 + len(name))
        out, ok = dec.decode_batch(llr)
        ro, rk, _ = ref_decode_ex(kind, N, frozen, llr, L, 1, crc, THREADS, 1, fresh=(kind == KIND_SCL_F32))
        bad = np.where((out != ro).any(axis=1) | (ok != rk))[0]
        assert len(bad) == 0, "%s: %d/%d frames differ from the reference, first %s" % (name, len(bad), n, bad[:5])
