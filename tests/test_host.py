"""CPU tests of the host side: the C ABI loads and exports every declared symbol, construction, CRC host
arithmetic, schedule compiler, argument / error behaviour, and the world_size-2 sharding logic on gloo."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest

import antpolarcodes_b200 as ap
from polar_testlib import (KIND_FASTSSC_I8, KIND_SCL_I8, oracle_crc_check, oracle_crc_generate, oracle_tree_stats)

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REFOUT = np.load(os.path.join(ROOT, "tests", "golden", "ref_outputs.npz"))


def test_abi_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "polar_cuda.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(polar_cuda_\w+)\s*\(", hdr))
    assert declared and declared == set(ap.EXPORTS)
    lib = ap.lib()
    for name in declared:
        assert getattr(lib, name) is not None
    assert lib.polar_cuda_version() >= 100


def test_frozen_bits_match_reference_fixtures():
    n = 0
    for key in REFOUT.files:
        if key.startswith("frozen_"):
            _, N, K, d, t = key.split("_")
            assert np.array_equal(ap.frozen_bits(int(N), int(K), float(d), t), REFOUT[key]), key
            n += 1
    assert n >= 8
    with pytest.raises(ValueError):
        ap.frozen_bits(2048, 1024, 0.0, "5G")   # fiveGList.cpp:21-23 throws for N > 1024
    with pytest.raises(ValueError):
        ap.frozen_bits(1000, 500, 0.0, "BB")    # constructor.cpp:24-30: not a power of two


def test_detector_known_answers_and_errors():
    from test_oracle import CRC_CHECKS, CRC_KATS
    for bits, msg, want in CRC_KATS:
        d = ap.Detector(bits, "CRC")
        got = d.generate(np.frombuffer(msg + bytes(bits // 8), np.uint8))
        assert bytes(got[-(bits // 8):]) == want
        assert d.check(got)
        got[0] ^= 0x40
        assert not d.check(got)
    for bits, msg in CRC_CHECKS:
        assert ap.Detector(bits).check(np.frombuffer(msg, np.uint8))
    assert ap.Detector(0).check(np.zeros(3, np.uint8)) and ap.Detector(0).getType() == "DUMMY"
    assert ap.Detector(16).getCheckBitCount() == 16 and ap.Detector(16).getType() == "CRC"
    with pytest.raises(ValueError):
        ap.Detector(12, "crc")                   # errordetector.cpp:33-35 "CRC INVALID SIZE!"
    with pytest.raises(RuntimeError):
        ap.Detector(8, "md5")                    # errordetector.cpp:64
    rng = np.random.default_rng(4)
    for bits in (8, 16, 32):
        for nbytes in (8, 12, 64, 128):
            m = rng.integers(0, 256, nbytes, dtype=np.uint8)
            assert np.array_equal(ap.Detector(bits).generate(m), oracle_crc_generate(bits, m))
            assert ap.Detector(bits).check(m) == oracle_crc_check(bits, m)


def _hist(kind, N, frozen, L):
    """Node histogram of the compiled schedule through the pure-host entry of the library."""
    lib = ap.lib()
    lib.polar_cuda_schedule_stats.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                                              ctypes.c_void_p, ctypes.c_void_p, ctypes.POINTER(ctypes.c_double)]
    counts, sizes = np.zeros(32, np.int64), np.zeros(32, np.int64)
    ops = ctypes.c_double()
    f = np.ascontiguousarray(frozen, np.uint32)
    rc = lib.polar_cuda_schedule_stats(kind, N, f.ctypes.data, len(f), L, counts.ctypes.data, sizes.ctypes.data,
                                       ctypes.byref(ops))
    assert rc >= 0, lib.polar_cuda_last_error()
    return counts, sizes, ops.value, rc


OP = dict(F=1, G=2, C=3, G_R1=4, G_0R=5, C_0R=6, SC_R0=7, SC_R1=8, SC_REP=9, SC_REP2=10, SC_SPC=11, SC_ZSPC=12,
          SC_Z1=13, R0=14, R1=15, REP=16, SPC=17)


@pytest.mark.parametrize("N,K,t", [(1024, 512, "BB"), (1024, 512, "5G"), (2048, 1024, "BB"), (256, 128, "BB"),
                                   (64, 40, "BB"), (4096, 1024, "BB")])
def test_schedule_compiler_matches_reference_tree(N, K, t):
    fz = REFOUT["frozen_%d_%d_0_%s" % (N, K, t)]
    c, s, ops, nops = _hist(ap.KIND_FASTSSC_I8, N, fz, 1)
    o = oracle_tree_stats(KIND_FASTSSC_I8, N, fz)
    for a, b in (("F", "F"), ("G", "G"), ("C", "C"), ("G_R1", "G_R1"), ("G_0R", "G_0R"), ("SC_REP", "REP"),
                 ("SC_REP2", "REP2"), ("SC_SPC", "SPC"), ("SC_ZSPC", "ZERO_SPC"), ("SC_Z1", "ZERO_ONE"),
                 ("SC_R0", "R0"), ("SC_R1", "R1")):
        size = o[b][1] if a not in ("F", "G", "C", "G_R1", "G_0R") else o[b][1]
        assert (int(c[OP[a]]), int(s[OP[a]])) == (o[b][0], size), (a, int(c[OP[a]]), int(s[OP[a]]), o[b])
    c, s, ops, nops = _hist(ap.KIND_SCL_I8, N, fz, 8)
    o = oracle_tree_stats(KIND_SCL_I8, N, fz)
    for a, b in (("F", "F"), ("G", "G"), ("C", "C"), ("R0", "R0"), ("R1", "R1"), ("REP", "REP"), ("SPC", "SPC")):
        assert (int(c[OP[a]]), int(s[OP[a]])) == o[b], (a, o[b])
    assert ops > 0 and nops == int(c.sum()) + 1


def test_schedule_rejects_bad_arguments():
    lib = ap.lib()
    fz = np.array([3, 1, 2], np.uint32)
    with pytest.raises(AssertionError):
        _hist(ap.KIND_SCL_I8, 64, fz, 8)          # unsorted frozen set
    with pytest.raises(AssertionError):
        _hist(ap.KIND_SCL_I8, 48, np.arange(8, dtype=np.uint32), 8)   # N not a power of two
    with pytest.raises(AssertionError):
        _hist(ap.KIND_SCL_I8, 64, np.arange(8, dtype=np.uint32), 3)   # list size not a power of two
    assert b"power of two" in lib.polar_cuda_last_error()


def test_python_surface_errors_without_gpu():
    fz = ap.frozen_bits(64, 32, 0.0)
    with pytest.raises(ValueError):
        ap.PolarDecoder(64, 4, fz, "quantum")      # decoder.cpp:45 "Unknown PolarDecoder type!"
    d = ap.PolarDecoder(64, 4, fz, "char")
    assert d.blockLength() == 64 and d.infoLength() == 32 and d.listSize() == 4 and d.isSystematic()
    assert d.getErrorDetectionMode() == "CRC-8"    # decoder.cpp:85: makeDecoder installs CRC8
    d.setErrorDetection(16)
    assert d.getErrorDetectionMode() == "CRC-16"
    with pytest.raises(RuntimeError):
        d.decode_vector(np.zeros(63, np.int8))     # decoder_python.cc:45-56 length check
    with pytest.raises(RuntimeError):
        d.decode_vector(np.zeros((2, 64), np.int8))
    e = ap.PolarEncoder(64, fz)
    assert e.getErrorDetectionMode() == "DUMMY-0" and e.infoLength() == 32
    with pytest.raises(RuntimeError):
        e.encode_vector(np.zeros(5, np.uint8))
    if ap.lib().polar_cuda_device_count() == 0:
        with pytest.raises(RuntimeError):          # no CPU fallback: fails loudly without a device
            d.decode_vector(np.zeros(64, np.int8))


def test_sharding_world2_gloo(tmp_path):
    """frame ranges tile [0, n) and the counter all-reduce sums over ranks (gloo, 2 processes)."""
    script = tmp_path / "w.py"
    script.write_text(
        "import os, sys, json\n"
        "sys.path.insert(0, %r)\n"
        "import torch.distributed as dist\n"
        "from antpolarcodes_b200.sharding import frame_range, reduce_counters, max_over_ranks\n"
        "dist.init_process_group('gloo')\n"
        "r, w = dist.get_rank(), dist.get_world_size()\n"
        "lo, hi = frame_range(1001, r, w)\n"
        "c = reduce_counters({'frames': hi - lo, 'block_errors': r + 1, 'bit_errors': 10 * (r + 1)})\n"
        "m = max_over_ranks(1.5 + r)\n"
        "if r == 0: print(json.dumps({'c': c, 'm': m, 'lo': lo, 'hi': hi}))\n"
        "dist.destroy_process_group()\n" % ROOT)
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", "29731", str(script)],
                         capture_output=True, text=True, timeout=240)
    assert out.returncode == 0, out.stderr[-2000:]
    import json
    line = [l for l in out.stdout.splitlines() if l.startswith("{")][-1]
    d = json.loads(line)
    assert d["c"] == {"frames": 1001, "block_errors": 3, "bit_errors": 30} and d["m"] == 2.5 and d["lo"] == 0 and d["hi"] == 500
    from antpolarcodes_b200.sharding import frame_range
    for n, w in ((7, 8), (1 << 20, 8), (1001, 3)):
        r = [frame_range(n, k, w) for k in range(w)]
        assert r[0][0] == 0 and r[-1][1] == n and all(r[i][1] == r[i + 1][0] for i in range(w - 1))


def test_schedule_cache():
    """the node program of a (tree flavour, N, frozen set, list size) is compiled once per process;
    a cached program is the same program (histogram, op count, lane-op total)."""
    fz = ap.frozen_bits(512, 200, 0.0, "BE")
    n0, h0, m0 = ap.schedule_cache_stats()
    first = _hist(ap.KIND_SCL_I8, 512, fz, 4)
    n1, h1, m1 = ap.schedule_cache_stats()
    assert (h1, m1) == (h0, m0 + 1) and n1 == min(n0 + 1, 64)
    again = _hist(ap.KIND_SCL_I8, 512, fz, 4)
    assert ap.schedule_cache_stats()[1:] == (h1 + 1, m1)
    assert np.array_equal(first[0], again[0]) and np.array_equal(first[1], again[1]) and first[2:] == again[2:]
    _hist(ap.KIND_SCL_I8, 512, fz, 8)                     # another list size is another program
    _hist(ap.KIND_FASTSSC_I8, 512, fz, 1)
    fz2 = fz.copy()
    fz2[-1] = 511 if fz[-1] != 511 else 510                # same length, different set
    fz2.sort()
    _hist(ap.KIND_SCL_I8, 512, fz2, 4)
    assert ap.schedule_cache_stats()[2] == m1 + 3
    for i in range(70):                                    # more distinct codes than entries: the LRU stays bounded
        _hist(ap.KIND_FASTSSC_I8, 256, ap.frozen_bits(256, 60 + i, 0.0, "BB"), 1)
    assert ap.schedule_cache_stats()[0] == 64
