"""BASELINE.json's full sizes (2^20 frames per launch), checked through size-independent properties -- the oracle
finishes only small batches in seconds:
  * tiling: a small oracle-checked set of noisy frames repeated to the full batch must decode to the same bytes in
    every tile (position independence of every warp / block / grid wave),
  * noiseless encode -> decode round trips return the payload with verdict 1,
  * the encoder is linear over GF(2) and, in systematic mode, shows the payload at the information positions."""
import numpy as np
import pytest

import antpolarcodes_b200 as ap
from polar_testlib import (KIND_ADAPTIVE_I8, KIND_FASTSSC_F32, KIND_FASTSSC_I8, KIND_SCL_I8, awgn_llrs, oracle_decode,
                           oracle_decode_adaptive)

pytestmark = pytest.mark.gpu

FULL = [  # kind, N, K, construction, L, crc, oracle-checked tile, full batch, flavour, adaptive
    (KIND_SCL_I8, 1024, 512, "5G", 8, 16, 1024, 1 << 19, "char", False),        # configs[2], the headline
    (KIND_FASTSSC_I8, 1024, 512, "BB", 1, 0, 2048, 1 << 20, "char", False),     # configs[0]
    (KIND_FASTSSC_F32, 256, 128, "BB", 1, 0, 2048, 1 << 20, "float", False),    # configs[1]
    (KIND_SCL_I8, 2048, 1024, "BB", 32, 32, 128, 1 << 15, "char", False),        # configs[3]
    (KIND_ADAPTIVE_I8, 1024, 512, "5G", 8, 16, 1024, 1 << 20, "char", True),
]


@pytest.mark.parametrize("kind,N,K,t,L,crc,tile,full,flavour,adaptive", FULL)
def test_tiled_full_batch(gpu, kind, N, K, t, L, crc, tile, full, flavour, adaptive):
    import torch
    fz = ap.frozen_bits(N, K, 0.0, t)
    llr = awgn_llrs(N, fz, tile, 2.0, crc, seed=77, fmt="f32" if flavour == "float" else "i8")[0]
    if adaptive:
        eo, ek, redone = oracle_decode_adaptive(kind, N, fz, llr, L, 1, crc)
        assert 0 < redone < tile
    else:
        eo, ek = oracle_decode(kind, N, fz, llr, L, 1, crc)
    dec = ap.PolarDecoder(N, L, fz, flavour, adaptive=adaptive)
    dec.setErrorDetection(crc)
    d = torch.from_numpy(llr).cuda().repeat(full // tile, 1).contiguous()
    assert d.shape[0] == full
    out, ok = dec.decode_batch(d)
    torch.cuda.synchronize()
    out = out.view(full // tile, tile, -1)
    ok = ok.view(full // tile, tile)
    assert np.array_equal(out[0].cpu().numpy(), eo) and np.array_equal(ok[0].cpu().numpy(), ek)
    assert bool((out == out[0:1]).all().item()) and bool((ok == ok[0:1]).all().item())
    if adaptive:
        assert dec.fallbacks() == redone * (full // tile)


@pytest.mark.parametrize("N,K,t,L,crc,frames,sysm", [(1024, 512, "5G", 8, 16, 1 << 19, 1), (1024, 512, "BB", 1, 0, 1 << 20, 1),
                                                    (1024, 512, "BB", 1, 8, 1 << 20, 0), (256, 128, "BB", 1, 0, 1 << 20, 1)])
def test_noiseless_round_trip(gpu, N, K, t, L, crc, frames, sysm):
    import torch
    fz = ap.frozen_bits(N, K, 0.0, t)
    enc = ap.PolarEncoder(N, fz)
    enc.setSystematic(bool(sysm))
    enc.setErrorDetection(crc)
    dec = ap.PolarDecoder(N, L, fz, "char")
    dec.setSystematic(bool(sysm))
    dec.setErrorDetection(crc)
    g = torch.Generator(device="cuda")
    g.manual_seed(5)
    info = torch.randint(0, 256, (frames, K // 8), dtype=torch.uint8, device="cuda", generator=g)
    code = enc.encode_batch(info)                        # CRC bytes land in `info`
    shifts = torch.arange(7, -1, -1, device="cuda", dtype=torch.uint8)
    llr = torch.empty((frames, N), dtype=torch.int8, device="cuda")
    for lo in range(0, frames, 1 << 17):
        c = code[lo:lo + (1 << 17)]
        bits = ((c.unsqueeze(-1) >> shifts) & 1).reshape(c.shape[0], N)
        llr[lo:lo + (1 << 17)] = (20 - 40 * bits.to(torch.int16)).to(torch.int8)
    out, ok = dec.decode_batch(llr)
    torch.cuda.synchronize()
    assert bool((out == info).all().item()) and bool((ok == 1).all().item())


def test_encoder_linearity_and_systematic_positions(gpu):
    import torch
    N, K, frames = 1024, 512, 1 << 20
    fz = ap.frozen_bits(N, K, 0.0, "BB")
    g = torch.Generator(device="cuda")
    g.manual_seed(6)
    a = torch.randint(0, 256, (frames, K // 8), dtype=torch.uint8, device="cuda", generator=g)
    b = torch.randint(0, 256, (frames, K // 8), dtype=torch.uint8, device="cuda", generator=g)
    for sysm in (True, False):
        enc = ap.PolarEncoder(N, fz)
        enc.setSystematic(sysm)
        enc.setErrorDetection(0)
        ca, cb, cab = enc.encode_batch(a.clone()), enc.encode_batch(b.clone()), enc.encode_batch(a ^ b)
        torch.cuda.synchronize()
        assert bool(((ca ^ cb) == cab).all().item())
        if sysm:  # the payload bits sit at the non-frozen positions of the code word, in order
            info_pos = torch.from_numpy(np.setdiff1d(np.arange(N), fz).astype(np.int64)).cuda()
            shifts = torch.arange(7, -1, -1, device="cuda", dtype=torch.uint8)
            n = 1 << 16
            bits = ((ca[:n].unsqueeze(-1) >> shifts) & 1).reshape(n, N)[:, info_pos]
            want = ((a[:n].unsqueeze(-1) >> shifts) & 1).reshape(n, K)
            assert bool((bits == want).all().item())
