#!/usr/bin/env python
"""Generates tests/golden/ref_soft.npz in the dev container: Decoder::getSoftCodeword / getSoftInformation of the UNMODIFIED
reference's FastSscFipChar (oracle/_ref, decoder.cpp:147-152) on seeded LLRs.  Run: python tests/golden/make_golden_soft.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from polar_testlib import KIND_FASTSSC_I8, ref_frozen_bits, ref_soft, workload  # noqa: E402

CASES = [(1024, 512, "BB", 12), (1024, 512, "5G", 8), (256, 128, "BB", 16), (64, 40, "BB", 16), (32, 11, "BB", 16), (8, 4, "BB", 16),
         (2048, 1024, "BB", 4)]

if __name__ == "__main__":
    out = {}
    for N, K, t, n in CASES:
        fz = ref_frozen_bits(N, K, 0.0, t)
        for w in ("awgn", "pcbench", "stress"):
            llr = workload(w, KIND_FASTSSC_I8, N, fz, n, 0)
            soft, info = ref_soft(KIND_FASTSSC_I8, N, fz, llr)
            key = "soft_%d_%d_%s_%s" % (N, K, t, w)
            out[key + "_llr"], out[key + "_soft"], out[key + "_info"] = llr, soft, info
    np.savez_compressed(os.path.join(HERE, "ref_soft.npz"), **out)
    print("wrote", len(out), "arrays")
