#!/usr/bin/env python
"""Encoder rate against batch size (device-resident), for the choice of frames per thread.  usage: tools/enc_sizes.py [N K ctype crc]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import antpolarcodes_b200 as ap
N, K, ct, crc = (int(sys.argv[1]), int(sys.argv[2]), sys.argv[3], int(sys.argv[4])) if len(sys.argv) > 4 else (1024, 512, "BB", 0)
f = ap.frozen_bits(N, K, 0.0, ct)
enc = ap.PolarEncoder(N, f); enc.setErrorDetection(crc)
for lg in (12, 14, 16, 17, 18, 19, 20, 21):
    n = 1 << lg
    info = torch.randint(0, 256, (n, K // 8), dtype=torch.uint8, device="cuda")
    out = torch.empty((n, N // 8), dtype=torch.uint8, device="cuda")
    for _ in range(3): enc.encode_batch(info, out=out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): enc.encode_batch(info, out=out)
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1000 / 20
    print("N=%d K=%d crc=%d frames=2^%d  %.1f us  %.0f Gbit/s info" % (N, K, crc, lg, us, n * K / us / 1e3))
