"""polar_cuda_decode_batch_noalloc: the allocation-free variant of the decode entry point (scratch in a file-scope static
__device__ array, include/polar_cuda.h).  No counterpart in the reference.

Called through the C ABI (ctypes) on torch CUDA tensors, twice in a row, and compared with polar_cuda_decode_batch on the same
DEVICE buffers: bit-exact (integer / byte results, the tolerance of every decoder test in this suite).  Collected last on
purpose (file name)."""
import ctypes

import numpy as np
import pytest

import antpolarcodes_b200 as ap
from polar_testlib import awgn_llrs

pytestmark = pytest.mark.gpu

NOALLOC_BYTES = 64 << 20  # POLAR_NOALLOC_SCRATCH_BYTES


def _noalloc(code, llr, llr_type, out, ok, stream=0):
    return ap.lib().polar_cuda_decode_batch_noalloc(code._h, llr.data_ptr(), llr_type, llr.shape[0], out.data_ptr(), ok.data_ptr(), stream)


# kind, N, K, construction, L, crc, frames: codes whose scratch fits the static array, with and without spilled stages
CASES = [
    (ap.KIND_FASTSSC_I8, 1024, 512, "BB", 1, 0, 20000),    # two spilled stages: 24 KB of scratch per warp
    (ap.KIND_SCL_I8, 512, 256, "BB", 8, 16, 6000),         # list decoder, spilled LLR stages
    (ap.KIND_FASTSSC_F32, 256, 128, "BB", 1, 8, 5000),     # fp32 kind, float LLRs
    (ap.KIND_SCL_I8, 64, 32, "BB", 4, 8, 3000),            # no spilled stage at all: the scratch is not touched
]


@pytest.mark.parametrize("kind,N,K,ctype,L,crc,n", CASES)
def test_noalloc_equals_decode_batch_twice(gpu, kind, N, K, ctype, L, crc, n):
    import torch
    frozen = ap.frozen_bits(N, K, 0.0, ctype)
    f32 = kind in (ap.KIND_FASTSSC_F32, ap.KIND_SCL_F32)
    code = ap._Code(kind, N, frozen, L, True, crc, 0)
    kb = (K + 7) // 8
    lt = ap.LLR_F32 if f32 else ap.LLR_I8
    for call, (seed, ebn0) in enumerate(((11, 1.5), (12, 3.0))):   # two calls in a row on different data: the array keeps the
        h, _ = awgn_llrs(N, frozen, n, ebn0, crc, seed=seed, fmt="f32" if f32 else "i8")   # first call's contents
        llr = torch.from_numpy(h).cuda()
        want, want_ok = torch.empty((n, kb), dtype=torch.uint8, device="cuda"), torch.empty(n, dtype=torch.uint8, device="cuda")
        rc = ap.lib().polar_cuda_decode_batch(code._h, llr.data_ptr(), lt, ap.MEM_DEVICE, n, want.data_ptr(), want_ok.data_ptr(), 0)
        assert rc == 0, ap.lib().polar_cuda_last_error()
        got = torch.full((n, kb), 0xA5, dtype=torch.uint8, device="cuda")
        got_ok = torch.full((n,), 7, dtype=torch.uint8, device="cuda")
        rc = _noalloc(code, llr, lt, got, got_ok)
        assert rc == 0, ap.lib().polar_cuda_last_error()
        torch.cuda.synchronize()
        assert torch.equal(got, want) and torch.equal(got_ok, want_ok), "call %d differs from polar_cuda_decode_batch" % call
        assert 0 < int(want_ok.sum()) or crc == 0
    code.close()


def test_noalloc_two_handles_share_the_array_in_stream_order(gpu):
    """Two handles, one stream: calls are ordered by the stream, so sharing the array is safe."""
    import torch
    fa, fb = ap.frozen_bits(1024, 512, 0.0, "BB"), ap.frozen_bits(512, 256, 0.0, "BB")
    a, b = ap._Code(ap.KIND_FASTSSC_I8, 1024, fa, 1, True, 0, 0), ap._Code(ap.KIND_SCL_I8, 512, fb, 8, True, 16, 0)
    la = torch.from_numpy(awgn_llrs(1024, fa, 8192, 2.0, 0, seed=3)[0]).cuda()
    lb = torch.from_numpy(awgn_llrs(512, fb, 4096, 2.0, 16, seed=4)[0]).cuda()
    res = []
    for code, llr, kb in ((a, la, 64), (b, lb, 32), (a, la, 64)):
        n = llr.shape[0]
        want, wok = torch.empty((n, kb), dtype=torch.uint8, device="cuda"), torch.empty(n, dtype=torch.uint8, device="cuda")
        assert ap.lib().polar_cuda_decode_batch(code._h, llr.data_ptr(), ap.LLR_I8, ap.MEM_DEVICE, n, want.data_ptr(), wok.data_ptr(), 0) == 0
        got, gok = torch.empty_like(want), torch.empty_like(wok)
        res.append((code, llr, got, gok, want, wok))
    for code, llr, got, gok, _, _ in res:            # the three no-alloc calls back to back, no synchronisation in between
        assert _noalloc(code, llr, ap.LLR_I8, got, gok) == 0, ap.lib().polar_cuda_last_error()
    torch.cuda.synchronize()
    for _, _, got, gok, want, wok in res:
        assert torch.equal(got, want) and torch.equal(gok, wok)


def test_noalloc_rejects_what_does_not_fit_or_would_allocate(gpu):
    import torch
    L = ap.lib()
    # SCL L=8 at N=1024 needs more scratch than the static array holds: clear error, both sizes in the message
    f = ap.frozen_bits(1024, 512, 0.0, "5G")
    big = ap._Code(ap.KIND_SCL_I8, 1024, f, 8, True, 16, 0)
    llr = torch.zeros((64, 1024), dtype=torch.int8, device="cuda")
    out, ok = torch.zeros((64, 64), dtype=torch.uint8, device="cuda"), torch.zeros(64, dtype=torch.uint8, device="cuda")
    assert _noalloc(big, llr, ap.LLR_I8, out, ok) == -2          # POLAR_E_UNSUPPORTED
    msg = L.polar_cuda_last_error().decode()
    assert str(NOALLOC_BYTES) in msg and "POLAR_NOALLOC_SCRATCH_BYTES" in msg
    # float LLRs into an int8 decoder would need the quantiser's scratch copy; a missing verdict buffer would need staging
    small = ap._Code(ap.KIND_FASTSSC_I8, 256, ap.frozen_bits(256, 128, 0.0, "BB"), 1, True, 0, 0)
    fl = torch.zeros((8, 256), dtype=torch.float32, device="cuda")
    o2, k2 = torch.zeros((8, 16), dtype=torch.uint8, device="cuda"), torch.zeros(8, dtype=torch.uint8, device="cuda")
    assert _noalloc(small, fl, ap.LLR_F32, o2, k2) == -1
    assert L.polar_cuda_decode_batch_noalloc(small._h, fl.data_ptr(), ap.LLR_I8, 8, o2.data_ptr(), None, 0) == -1
    enc = ap._Code(ap.KIND_ENC, 256, ap.frozen_bits(256, 128, 0.0, "BB"), 1, True, 0, 0)
    assert L.polar_cuda_decode_batch_noalloc(enc._h, fl.data_ptr(), ap.LLR_I8, 8, o2.data_ptr(), k2.data_ptr(), 0) == -1
    assert L.polar_cuda_decode_batch_noalloc(None, fl.data_ptr(), ap.LLR_I8, 8, o2.data_ptr(), k2.data_ptr(), 0) == -1
