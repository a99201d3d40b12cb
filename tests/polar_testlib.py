"""Shared test/bench helpers: ctypes bindings of the CPU oracle (oracle/liboracle.so), of the
compiled reference (oracle/_ref/libpolar_ref.so, optional) and seeded AWGN/BPSK workloads.

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product package never imports this module.
"""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

KIND_ENC, KIND_FASTSSC_I8, KIND_FASTSSC_F32, KIND_SCL_I8, KIND_SCL_F32 = range(5)
KIND_ADAPTIVE_I8, KIND_ADAPTIVE_F32, KIND_ADAPTIVE_MIXED = 5, 6, 7
# adaptive kind -> (Fast-SSC kind, list kind) (adaptive_char.cpp:24-25, adaptive_float.cpp, adaptive_mixed.cpp:24-25)
ADAPTIVE_PARTS = {KIND_ADAPTIVE_I8: (KIND_FASTSSC_I8, KIND_SCL_I8), KIND_ADAPTIVE_F32: (KIND_FASTSSC_F32, KIND_SCL_F32),
                  KIND_ADAPTIVE_MIXED: (KIND_FASTSSC_I8, KIND_SCL_F32)}
NODE_NAMES = ["F", "G", "C", "G_R1", "G_0R", "R0", "R1", "REP", "REP2", "SPC", "ZERO_SPC", "ZERO_ONE"]

_c = ctypes
_vp, _i, _l, _d = _c.c_void_p, _c.c_int, _c.c_long, _c.c_double


def _ptr(a):
    return None if a is None else a.ctypes.data


def build_oracle():
    subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "all"])


_oracle = None
_ref = None


def oracle():
    global _oracle
    if _oracle is None:
        p = os.path.join(ORACLE_DIR, "liboracle.so")
        if not os.path.exists(p):
            subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "liboracle.so"])
        lib = _c.CDLL(p)
        lib.oracle_crc_generate.argtypes = [_i, _vp, _i]
        lib.oracle_crc_check.argtypes = [_i, _vp, _i]
        lib.oracle_quantize.argtypes = [_vp, _vp, _i]
        lib.oracle_f_i8.argtypes = [_c.c_int8, _c.c_int8]
        lib.oracle_f_i8.restype = _c.c_int8
        lib.oracle_g_i8.argtypes = [_c.c_int8, _c.c_int8, _c.c_int8]
        lib.oracle_g_i8.restype = _c.c_int8
        lib.oracle_rep_sum_i8.argtypes = [_vp, _i]
        lib.oracle_rep_sum_i8.restype = _c.c_int8
        lib.oracle_encode.argtypes = [_i, _vp, _i, _i, _i, _vp, _vp, _l]
        lib.oracle_decode.argtypes = [_i, _i, _vp, _i, _i, _i, _i, _vp, _l, _vp, _vp]
        lib.oracle_tree_stats.argtypes = [_i, _i, _vp, _i, _vp, _vp]
        _oracle = lib
    return _oracle


def have_ref():
    return os.path.exists(os.path.join(ORACLE_DIR, "_ref", "libpolar_ref.so"))


def ref():
    global _ref
    if _ref is None:
        lib = _c.CDLL(os.path.join(ORACLE_DIR, "_ref", "libpolar_ref.so"))
        lib.polar_ref_frozen_bits.argtypes = [_i, _i, _c.c_float, _c.c_char_p, _vp]
        lib.polar_ref_code_create.argtypes = [_i, _i, _vp, _i, _i, _i, _i]
        lib.polar_ref_code_create.restype = _vp
        lib.polar_ref_code_destroy.argtypes = [_vp]
        lib.polar_ref_encode.argtypes = [_vp, _vp, _vp, _l]
        lib.polar_ref_decode.argtypes = [_vp, _vp, _l, _vp, _vp, _i, _i]
        lib.polar_ref_decode.restype = _d
        lib.polar_ref_crc.argtypes = [_i, _vp, _i, _i]
        _ref = lib
    return _ref


# ---------------------------------------------------------------- reference wrappers
def ref_frozen_bits(N, K, dsnr=0.0, ctype="BB"):
    out = np.zeros(N, np.uint32)
    n = ref().polar_ref_frozen_bits(N, K, dsnr, ctype.encode(), _ptr(out))
    if n < 0:
        raise RuntimeError("reference frozen_bits threw")
    return out[:n].copy()


def ref_encode(N, frozen, info, systematic=1, crc=0):
    frozen = np.ascontiguousarray(frozen, np.uint32)
    info = np.ascontiguousarray(info, np.uint8).copy()
    nF = info.shape[0]
    code = np.zeros((nF, max(N // 8, 1)), np.uint8)
    h = ref().polar_ref_code_create(KIND_ENC, N, _ptr(frozen), len(frozen), 1, systematic, crc)
    rc = ref().polar_ref_encode(h, _ptr(info), _ptr(code), nF)
    ref().polar_ref_code_destroy(h)
    if rc:
        raise RuntimeError("reference encode threw")
    return code, info


def ref_decode(kind, N, frozen, llr, L=1, systematic=1, crc=0, threads=1, repeat=1):
    frozen = np.ascontiguousarray(frozen, np.uint32)
    llr = np.ascontiguousarray(llr)
    nF = llr.shape[0]
    K = N - len(frozen)
    out = np.zeros((nF, (K + 7) // 8), np.uint8)
    ok = np.zeros(nF, np.uint8)
    h = ref().polar_ref_code_create(kind, N, _ptr(frozen), len(frozen), L, systematic, crc)
    t = ref().polar_ref_decode(h, _ptr(llr), nF, _ptr(out), _ptr(ok), threads, repeat)
    ref().polar_ref_code_destroy(h)
    if t < 0:
        raise RuntimeError("reference decode threw")
    return out, ok, t


def ref_decode_ex(kind, N, frozen, llr, L=1, systematic=1, crc=0, threads=1, repeat=1, fresh=False):
    """ref_decode that takes the LLR type from the array (float32 into an int8 decoder = the decoder's own quantiser,
    as pcbench feeds it) and can build a fresh decoder object per frame (see ref_decode_fresh), multi-threaded."""
    frozen = np.ascontiguousarray(frozen, np.uint32)
    llr = np.ascontiguousarray(llr)
    assert llr.dtype in (np.int8, np.float32)
    nF = llr.shape[0]
    K = N - len(frozen)
    out = np.zeros((nF, (K + 7) // 8), np.uint8)
    ok = np.zeros(nF, np.uint8)
    fn = ref().polar_ref_decode_ex
    fn.argtypes = [_vp, _vp, _i, _l, _vp, _vp, _i, _i, _i]
    fn.restype = _d
    h = ref().polar_ref_code_create(kind, N, _ptr(frozen), len(frozen), L, systematic, crc)
    t = fn(h, _ptr(llr), 1 if llr.dtype == np.float32 else 0, nF, _ptr(out), _ptr(ok), threads, repeat, 1 if fresh else 0)
    ref().polar_ref_code_destroy(h)
    if t < 0:
        raise RuntimeError("reference decode threw")
    return out, ok, t


def ref_decode_fresh(kind, N, frozen, llr, L=1, systematic=1, crc=0):
    """One NEW reference decoder object per frame.  Needed for SclAvxFloat: PathList never resets the metric
    of path 0 (scl_avx_float.cpp:20-56,102-108), so a reused object starts every frame from the previous
    frame's winning metric and its fp32 roundings -- hence, rarely, its decisions -- depend on history."""
    llr = np.ascontiguousarray(llr)
    outs, oks = [], []
    for f in range(llr.shape[0]):
        o, k, _ = ref_decode(kind, N, frozen, llr[f:f + 1], L, systematic, crc)
        outs.append(o)
        oks.append(k)
    return np.concatenate(outs), np.concatenate(oks)


def ref_soft(kind, N, frozen, llr, L=1, systematic=1, crc=0):
    """(soft codeword [nF, N], soft information [nF, K]) of the reference decoder: Decoder::getSoftCodeword /
    getSoftInformation (decoder.cpp:147-152) after decode_vector."""
    frozen = np.ascontiguousarray(frozen, np.uint32)
    llr = np.ascontiguousarray(llr)
    nF, K = llr.shape[0], N - len(frozen)
    soft, info = np.zeros((nF, N), llr.dtype), np.zeros((nF, K), llr.dtype)
    fn = ref().polar_ref_soft
    fn.argtypes = [_vp, _vp, _l, _vp, _vp]
    fn.restype = _i
    h = ref().polar_ref_code_create(kind, N, _ptr(frozen), len(frozen), L, systematic, crc)
    rc = fn(h, _ptr(llr), nF, _ptr(soft), _ptr(info))
    ref().polar_ref_code_destroy(h)
    if rc:
        raise RuntimeError("reference soft output threw")
    return soft, info


def oracle_fastssc_i8_soft(N, frozen, llr):
    """soft codeword of the int8 Fast-SSC oracle, [nF, N] int8"""
    frozen = np.ascontiguousarray(frozen, np.uint32)
    llr = np.ascontiguousarray(llr, np.int8)
    soft = np.zeros(llr.shape, np.int8)
    fn = oracle().oracle_fastssc_i8_soft
    fn.argtypes = [_i, _vp, _i, _vp, _l, _vp]
    assert fn(N, _ptr(frozen), len(frozen), _ptr(llr), llr.shape[0], _ptr(soft)) == 0
    return soft


def is_float_kind(kind):
    return kind in (KIND_FASTSSC_F32, KIND_SCL_F32, KIND_ADAPTIVE_F32, KIND_ADAPTIVE_MIXED)


def ref_decode_adaptive(kind, N, frozen, llr, L, systematic=1, crc=0, fresh=False):
    """The reference's AdaptiveChar / AdaptiveFloat / AdaptiveMixed objects.  They only take float LLRs
    (adaptive_char.cpp:60-64); int8 LLRs are widened, which CharContainer::insertLlr maps back unchanged.
    fresh = one new object per frame (the fp32 list decoder inside is history dependent, see ref_decode_fresh)."""
    f = np.ascontiguousarray(llr, np.float32)
    if fresh:
        return ref_decode_fresh(kind, N, frozen, f, L, systematic, crc)
    o, k, _ = ref_decode(kind, N, frozen, f, L, systematic, crc)
    return o, k


def oracle_decode_adaptive(kind, N, frozen, llr, L, systematic=1, crc=0):
    """AdaptiveChar::decode (adaptive_char.cpp:34-46) composed from the oracle's Fast-SSC and list decoders:
    the list decoder's result replaces the Fast-SSC one wherever the detector rejected it (and L > 1).
    Returns (info, ok, number of frames the list decoder redid)."""
    fastKind, listKind = ADAPTIVE_PARTS[kind]
    llr = np.ascontiguousarray(llr)
    as_i8 = (lambda x: x if x.dtype == np.int8 else oracle_quantize(x))
    fin = llr if is_float_kind(fastKind) else as_i8(llr)
    out, ok = oracle_decode(fastKind, N, frozen, fin, 1, systematic, crc)
    bad = np.flatnonzero(ok == 0)
    if L > 1 and len(bad):
        lin = llr if is_float_kind(listKind) else as_i8(llr)
        o2, k2 = oracle_decode(listKind, N, frozen, lin[bad], L, systematic, crc)
        out[bad], ok[bad] = o2, k2
    return out, ok, (len(bad) if L > 1 else 0)


def workload(name, kind, N, frozen, n, crc, seed=None):
    """The seeded LLR sets of the parity tests, int8 or fp32 according to the decoder kind."""
    fmt = "f32" if is_float_kind(kind) else "i8"
    kw = {} if seed is None else {"seed": seed}
    if name == "awgn":
        return awgn_llrs(N, frozen, n, 2.0, crc, fmt=fmt, **kw)[0]
    if name == "awgn0":
        return awgn_llrs(N, frozen, n, 0.0, crc, fmt=fmt, **({"seed": 7} if seed is None else kw))[0]
    if name == "pcbench":
        return pcbench_llrs(N, n, fmt=fmt, **kw)
    if name == "stress":
        if fmt == "f32":  # small integers as floats: exact ties in |llr| and in the path metrics, and zeros
            return np.rint(2.0 * pcbench_llrs(N, n, fmt="f32", **({"seed": 3} if seed is None else kw))).astype(np.float32)
        return stress_llrs(N, n, **kw)
    raise ValueError(name)


# ---------------------------------------------------------------- oracle wrappers
def oracle_encode(N, frozen, info, systematic=1, crc=0):
    frozen = np.ascontiguousarray(frozen, np.uint32)
    info = np.ascontiguousarray(info, np.uint8).copy()
    nF = info.shape[0]
    code = np.zeros((nF, (N + 7) // 8), np.uint8)
    oracle().oracle_encode(N, _ptr(frozen), len(frozen), systematic, crc, _ptr(info), _ptr(code), nF)
    return code, info


def oracle_decode(kind, N, frozen, llr, L=1, systematic=1, crc=0):
    frozen = np.ascontiguousarray(frozen, np.uint32)
    llr = np.ascontiguousarray(llr)
    nF = llr.shape[0]
    K = N - len(frozen)
    out = np.zeros((nF, (K + 7) // 8), np.uint8)
    ok = np.zeros(nF, np.uint8)
    rc = oracle().oracle_decode(kind, N, _ptr(frozen), len(frozen), L, systematic, crc, _ptr(llr), nF,
                                _ptr(out), _ptr(ok))
    if rc:
        raise RuntimeError("oracle decode failed rc=%d" % rc)
    return out, ok


def oracle_tree_stats(kind, N, frozen):
    frozen = np.ascontiguousarray(frozen, np.uint32)
    counts = np.zeros(16, np.int64)
    sizes = np.zeros(16, np.int64)
    rc = oracle().oracle_tree_stats(kind, N, _ptr(frozen), len(frozen), _ptr(counts), _ptr(sizes))
    assert rc == 0
    return {NODE_NAMES[i]: (int(counts[i]), int(sizes[i])) for i in range(len(NODE_NAMES))}


def oracle_crc_generate(bits, data):
    d = np.ascontiguousarray(data, np.uint8).copy()
    oracle().oracle_crc_generate(bits, _ptr(d), len(d))
    return d


def oracle_crc_check(bits, data):
    d = np.ascontiguousarray(data, np.uint8)
    return bool(oracle().oracle_crc_check(bits, _ptr(d), len(d)))


def oracle_quantize(llr):
    llr = np.ascontiguousarray(llr, np.float32)
    out = np.zeros(llr.shape, np.int8)
    flat_in, flat_out = llr.reshape(-1, llr.shape[-1]), out.reshape(-1, llr.shape[-1])
    for r in range(flat_in.shape[0]):
        oracle().oracle_quantize(flat_in[r].ctypes.data, flat_out[r].ctypes.data, llr.shape[-1])
    return out


# ---------------------------------------------------------------- workloads
def unpack_bits(packed, nbits):
    return np.unpackbits(np.ascontiguousarray(packed, np.uint8), axis=-1, bitorder="big")[..., :nbits]


def awgn_llrs(N, frozen, nFrames, ebn0_db=2.0, crc=0, seed=0xB200, systematic=1, encode=None,
              fmt="i8"):
    """payload -> CRC -> encode -> BPSK (0 -> +1) -> AWGN -> LLR.
    sigma = 1/sqrt(2 (K/N) 10^(EbN0/10)) (simulator.cpp:832-838); int8 LLR = clamp(rne(10 y))
    (setup.cpp:58, bitcontainer.cpp:487-497); fp32 LLR = 4 (K/N) 10^(EbN0/10) y ... the reference
    scales by 4*EsN0-ish factors (simulator.cpp:377-378); we use 2 y / sigma^2.
    Returns (llr, info_with_crc)."""
    K = N - len(frozen)
    rng = np.random.default_rng(seed)
    info = rng.integers(0, 256, size=(nFrames, (K + 7) // 8), dtype=np.uint8)
    if K % 8:
        info[:, -1] &= (0xFF << (8 - K % 8)) & 0xFF
    enc = encode or oracle_encode
    code, info = enc(N, frozen, info, systematic, crc)
    bits = unpack_bits(code, N).astype(np.float32)
    x = 1.0 - 2.0 * bits
    esn0 = (K / N) * 10.0 ** (ebn0_db / 10.0)
    sigma = 1.0 / np.sqrt(2.0 * esn0)
    y = x + sigma * rng.standard_normal(size=x.shape, dtype=np.float32)
    if fmt == "i8":
        llr = np.clip(np.rint(10.0 * y), -128, 127).astype(np.int8)
    else:
        llr = (2.0 / (sigma * sigma) * y).astype(np.float32)
    return llr, info


def pcbench_llrs(N, nFrames, seed=1, fmt="i8"):
    """pcbench-style inputs: iid N(0, 0.5) floats, not codewords (main_benchmark.cpp:224); after
    int8 rounding they are mostly {-1,0,1}: the tie-break stress case."""
    rng = np.random.default_rng(seed)
    y = (0.5 * rng.standard_normal(size=(nFrames, N))).astype(np.float32)
    if fmt == "i8":
        return np.clip(np.rint(y), -128, 127).astype(np.int8)
    return y


def stress_llrs(N, nFrames, seed=2, scale=40.0):
    """Large-magnitude iid int8 LLRs: exercises saturation (+-127/-128) in G / REP / SPC."""
    rng = np.random.default_rng(seed)
    y = scale * rng.standard_normal(size=(nFrames, N))
    return np.clip(np.rint(y), -128, 127).astype(np.int8)


# ---------------------------------------------------------------- rate matching (Puncturer)
def _punct_api():
    o = oracle()
    o.oracle_puncturer_positions.argtypes = [_i, _vp, _i, _vp]
    o.oracle_puncture.argtypes = [_vp, _i, _i, _vp, _vp]
    o.oracle_depuncture.argtypes = [_vp, _i, _i, _i, _vp, _vp]
    o.oracle_puncture_packed.argtypes = [_vp, _i, _vp, _vp]
    return o


def oracle_puncturer_positions(M, frozen):
    _punct_api()
    frozen = np.ascontiguousarray(frozen, np.uint32)
    pos = np.zeros(2 * M + 8, np.uint32)
    n = oracle().oracle_puncturer_positions(M, _ptr(frozen), len(frozen), _ptr(pos))
    return None if n < 0 else pos[:n].copy()


def oracle_puncture(pos, x):
    x = np.ascontiguousarray(x)
    out = np.zeros((x.shape[0], len(pos)), x.dtype)
    for r in range(x.shape[0]):
        _punct_api().oracle_puncture(_ptr(pos), len(pos), x.dtype.itemsize, ctypes.c_void_p(x[r].ctypes.data), ctypes.c_void_p(out[r].ctypes.data))
    return out


def oracle_depuncture(pos, parent, x):
    x = np.ascontiguousarray(x)
    out = np.full((x.shape[0], parent), 77, x.dtype)
    for r in range(x.shape[0]):
        _punct_api().oracle_depuncture(_ptr(pos), len(pos), parent, x.dtype.itemsize, ctypes.c_void_p(x[r].ctypes.data), ctypes.c_void_p(out[r].ctypes.data))
    return out


def oracle_puncture_packed(pos, x):
    x = np.ascontiguousarray(x, np.uint8)
    out = np.zeros((x.shape[0], len(pos) // 8), np.uint8)
    for r in range(x.shape[0]):
        _punct_api().oracle_puncture_packed(_ptr(pos), len(pos), ctypes.c_void_p(x[r].ctypes.data), ctypes.c_void_p(out[r].ctypes.data))
    return out


def ref_puncturer(M, frozen, mode, x=None, parent=None):
    """The reference's Puncturer object: mode 0 positions, 1 puncture, 2 depuncture, 3 puncturePacked (row by row)."""
    frozen = np.ascontiguousarray(frozen, np.uint32)
    fn = ref().polar_ref_puncturer
    fn.restype = ctypes.c_int
    fn.argtypes = [_i, _vp, _i, _i, _i, _vp, _vp]
    if mode == 0:
        pos = np.zeros(2 * M + 8, np.uint32)
        n = fn(M, _ptr(frozen), len(frozen), 0, 0, None, _ptr(pos))
        return None if n < 0 else pos[:n].copy()
    x = np.ascontiguousarray(x)
    cols = {1: M, 2: parent, 3: M // 8}[mode]
    out = np.zeros((x.shape[0], cols), x.dtype)
    for r in range(x.shape[0]):
        rc = fn(M, _ptr(frozen), len(frozen), mode, x.dtype.itemsize, ctypes.c_void_p(x[r].ctypes.data),
                ctypes.c_void_p(out[r].ctypes.data))
        assert rc >= 0
    return out
