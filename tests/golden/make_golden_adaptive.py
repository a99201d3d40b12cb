#!/usr/bin/env python
"""Generates tests/golden/ref_adaptive.npz in the dev container (needs /root/reference and oracle/_ref):
outputs of the UNMODIFIED reference's AdaptiveChar / AdaptiveFloat / AdaptiveMixed objects
(src/polarcode/decoding/adaptive_char.cpp:34-46 and siblings) on the seeded workloads of tests/test_adaptive.py.
The fp32 list decoder inside AdaptiveFloat / AdaptiveMixed is history dependent, so those two
kinds are recorded with a fresh reference object per frame."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from polar_testlib import KIND_ADAPTIVE_I8, ref_decode_adaptive, ref_frozen_bits  # noqa: E402
from test_adaptive import GOLDEN_CASES, adaptive_llrs, golden_key  # noqa: E402


def main():
    out = {}
    for case in GOLDEN_CASES:
        kind, N, K, t, L, crc, n, ebn0 = case
        fz = ref_frozen_bits(N, K, 0.0, t)
        llr = adaptive_llrs(kind, N, fz, n, crc, ebn0)
        o, k = ref_decode_adaptive(kind, N, fz, llr, L, 1, crc, fresh=(kind != KIND_ADAPTIVE_I8))
        out[golden_key(case) + "_out"] = o
        out[golden_key(case) + "_ok"] = k
    np.savez_compressed(os.path.join(HERE, "ref_adaptive.npz"), **out)
    print("wrote", len(out), "arrays")


if __name__ == "__main__":
    main()
