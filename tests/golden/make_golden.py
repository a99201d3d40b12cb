#!/usr/bin/env python
"""Generates tests/golden/*.json|npz in the dev container (needs /root/reference and oracle/_ref).

1. kat_fg.json      : the int8 F / G golden vectors of the reference's own unit test
                      (test/polarcode/decodingtest.cpp:498-737), parsed from the source text.
2. ref_outputs.npz  : outputs of the UNMODIFIED compiled reference (oracle/_ref/libpolar_ref.so) on the
                      seeded workloads of tests/polar_testlib.py, for every decoder / encoder configuration
                      the GPU tests cover, plus the frozen sets of Construction::frozen_bits.
The GPU box has no /root/reference: tests there read only these fixtures (and the prebuilt .so files).
"""
import json
import os
import re
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from polar_testlib import (KIND_FASTSSC_F32, KIND_FASTSSC_I8, KIND_SCL_F32, KIND_SCL_I8, ref_decode,  # noqa: E402
                           ref_decode_fresh, ref_encode, ref_frozen_bits, workload)

REF = "/root/reference"


def parse_fg():
    src = open(os.path.join(REF, "test/polarcode/decodingtest.cpp")).read()
    body = src[src.index("void DecodingTest::testGeneralDecodingFunctionsAvx2()"):]
    vecs = []
    for m in re.finditer(r"(llr\[0\]|llr\[1\]|bits|expected)\s*=\s*_mm256_(setr?)_epi8\(([^;]*?)\);", body):
        vals = [int(x) for x in re.findall(r"-?\d+", m.group(3))]
        if len(vals) != 32:
            continue
        if m.group(2) == "set":          # _mm256_set_epi8 lists element 31 first
            vals = vals[::-1]
        vecs.append((m.group(1), vals))
        if len(vecs) == 7:
            break
    names = [v[0] for v in vecs]
    assert names == ["llr[0]", "llr[1]", "expected", "llr[1]", "llr[0]", "bits", "expected"], names
    return {"source": "test/polarcode/decodingtest.cpp:504-737",
            "f": {"left": vecs[0][1], "right": vecs[1][1], "expected": vecs[2][1]},
            "g": {"right": vecs[3][1], "left": vecs[4][1], "bits": vecs[5][1], "expected": vecs[6][1]}}


DEC_CASES = [  # kind, N, K, ctype, L, crc, frames
    (KIND_SCL_I8, 1024, 512, "5G", 8, 16, 32), (KIND_SCL_I8, 2048, 1024, "BB", 32, 32, 4),
    (KIND_SCL_I8, 256, 128, "BB", 4, 8, 32), (KIND_SCL_I8, 64, 32, "BB", 16, 0, 32),
    (KIND_FASTSSC_I8, 1024, 512, "BB", 1, 0, 32), (KIND_FASTSSC_I8, 256, 128, "BB", 1, 8, 32),
    (KIND_FASTSSC_I8, 4096, 1024, "BB", 1, 0, 8), (KIND_FASTSSC_I8, 64, 40, "BB", 1, 8, 32),
    # fp32 decoders (BASELINE configs[1]: Fast-SSC float N=256 K=128; the fp32 variant of configs[2])
    (KIND_FASTSSC_F32, 256, 128, "BB", 1, 0, 32), (KIND_FASTSSC_F32, 1024, 512, "BB", 1, 8, 32),
    (KIND_FASTSSC_F32, 2048, 1024, "BB", 1, 0, 8), (KIND_FASTSSC_F32, 64, 40, "BB", 1, 8, 32),
    (KIND_FASTSSC_F32, 512, 448, "BB", 1, 0, 16), (KIND_FASTSSC_F32, 128, 16, "BB", 1, 8, 16),
    (KIND_SCL_F32, 1024, 512, "5G", 8, 16, 16), (KIND_SCL_F32, 256, 128, "BB", 4, 8, 32),
    (KIND_SCL_F32, 64, 32, "BB", 16, 0, 32),
]
ENC_CASES = [(1024, 512, "5G", 1, 16), (256, 128, "BB", 1, 8), (256, 128, "BB", 0, 0), (2048, 1024, "BB", 1, 32),
             (64, 24, "BB", 1, 8)]
FROZEN_CASES = [(1024, 512, 0.0, "BB"), (1024, 512, 0.0, "5G"), (2048, 1024, 0.0, "BB"), (256, 128, 0.0, "BB"),
                (512, 256, 0.0, "BE"), (4096, 1024, 0.0, "BB"), (64, 40, 0.0, "BB"), (64, 32, 0.0, "BB"),
                (64, 24, 0.0, "BB"), (1024, 512, 2.0, "BB")]


def main():
    json.dump(parse_fg(), open(os.path.join(HERE, "kat_fg.json"), "w"), indent=1)
    out = {}
    for N, K, d, t in FROZEN_CASES:
        out["frozen_%d_%d_%g_%s" % (N, K, d, t)] = ref_frozen_bits(N, K, d, t)
    for kind, N, K, t, L, crc, n in DEC_CASES:
        fz = ref_frozen_bits(N, K, 0.0, t)
        for wname in ("awgn", "pcbench", "stress"):
            llr = workload(wname, kind, N, fz, n, crc)
            for sysm in (1, 0):
                if not sysm and N < 256:
                    continue  # the reference's non-systematic path is broken below 256
                if kind == KIND_SCL_F32:  # a fresh object per frame: see ref_decode_fresh
                    o, k = ref_decode_fresh(kind, N, fz, llr, L, sysm, crc)
                else:
                    o, k, _ = ref_decode(kind, N, fz, llr, L, sysm, crc)
                key = "dec_%d_%d_%d_%s_%d_%d_%s_%d" % (kind, N, K, t, L, crc, wname, sysm)
                out[key + "_out"] = o
                out[key + "_ok"] = k
    rng = np.random.default_rng(99)
    for N, K, t, sysm, crc in ENC_CASES:
        fz = ref_frozen_bits(N, K, 0.0, t)
        info = rng.integers(0, 256, (16, K // 8), dtype=np.uint8)
        code, info2 = ref_encode(N, fz, info, sysm, crc)
        key = "enc_%d_%d_%s_%d_%d" % (N, K, t, sysm, crc)
        out[key + "_in"] = info
        out[key + "_info"] = info2
        out[key + "_code"] = code
    np.savez_compressed(os.path.join(HERE, "ref_outputs.npz"), **out)
    print("wrote", len(out), "arrays")


if __name__ == "__main__":
    main()
