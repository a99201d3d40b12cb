"""On-device link simulator (csrc/sim.cuh behind polar_cuda_sim_*) and the BLER sweep driver (antpolarcodes_b200/bler.py):
random streams, the chain payload -> CRC -> encoder -> channel -> decoder -> counters against the oracle, independence of
batch size / sharding, resume, the reference's CSV columns; the multi-rank round logic runs on gloo with a stand-in
simulator."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def philox4x32_10(ctr, key):
    """Philox4x32-10 (Salmon et al., SC'11), plain Python: the restatement the device generator is checked against."""
    M0, M1, W0, W1 = 0xD2511F53, 0xCD9E8D57, 0x9E3779B9, 0xBB67AE85
    c, k = [int(x) for x in ctr], [int(x) for x in key]
    for _ in range(10):
        p0, p1 = M0 * c[0], M1 * c[2]
        c = [(p1 >> 32) ^ c[1] ^ k[0], p1 & 0xFFFFFFFF, (p0 >> 32) ^ c[3] ^ k[1], p0 & 0xFFFFFFFF]
        k = [(k[0] + W0) & 0xFFFFFFFF, (k[1] + W1) & 0xFFFFFFFF]
    return c


# Random123's known-answer vectors for philox4x32-10
PHILOX_KAT = [([0, 0, 0, 0], [0, 0], [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]),
              ([0xffffffff] * 4, [0xffffffff] * 2, [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]),
              ([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0],
               [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1])]


def test_philox_restatement_kat():
    for ctr, key, want in PHILOX_KAT:
        assert philox4x32_10(ctr, key) == want


def test_sweep_helpers():
    from antpolarcodes_b200.bler import CSV_HEADER, _parse_sweep, point_key, wilson
    assert _parse_sweep("0:4:1") == [0.0, 1.0, 2.0, 3.0, 4.0]
    assert _parse_sweep("0:4:0.5")[-1] == 4.0 and len(_parse_sweep("0:4:0.5")) == 9
    assert _parse_sweep("1.5,2") == [1.5, 2.0]
    lo, hi = wilson(10, 1000)
    assert 0.0048 < lo < 0.0056 and 0.018 < hi < 0.0185      # textbook value: [0.0054, 0.0183]
    assert wilson(0, 100)[0] == 0.0 and wilson(0, 100)[1] < 0.04
    assert point_key(1024, 512, 8, 16, "char", False, "BB", 0.0, 2.0, 1) != point_key(1024, 512, 8, 16, "char", True, "BB", 0.0, 2.0, 1)
    # column names of the reference's result table (src/simulation/simulator.cpp:519-527)
    cols = [c.strip('"') for c in CSV_HEADER.split(",")]
    assert cols[:12] == ["N", "K", "dSNR", "C", "L", "Eb/N0", "BPS", "BLER", "BER", "RER", "Runs", "Errors"]
    assert cols[12:] == ["Time", "Blockspeed", "Coded Bitrate", "Payload Bitrate", "Effective Payload Bitrate", "Encoder Bitrate",
                         "Amplification", "time min", "time max", "time mean", "time deviation"]


def test_round_logic_world2_gloo(tmp_path):
    """simulate_point on two gloo ranks with a stand-in simulator whose frame f is in error iff f % 7 == 3: the ranks take
    disjoint consecutive frame ranges, the all-reduced totals equal the closed form, both ranks stop in the same round."""
    script = tmp_path / "w.py"
    script.write_text(
        "import os, sys, json\n"
        "sys.path.insert(0, %r)\n"
        "import torch, torch.distributed as dist\n"
        "from antpolarcodes_b200.bler import simulate_point\n"
        "class Fake:\n"
        "    batch = 64\n"
        "    def __init__(self): self.ranges = []\n"
        "    def run(self, eb, seed, point, first, n, counters, amp, stream):\n"
        "        self.ranges.append((first, n))\n"
        "        f = torch.arange(first, first + n)\n"
        "        err = int((f %% 7 == 3).sum())\n"
        "        counters += torch.tensor([n, err, 2 * err, err, 0])\n"
        "dist.init_process_group('gloo')\n"
        "r, w = dist.get_rank(), dist.get_world_size()\n"
        "sim = Fake()\n"
        "tot, rounds, sub = simulate_point(sim, 2.0, 1, 0, r, w, 'cpu', 100, 10**6, 128)\n"
        "open(os.path.join(%r, 'rank%%d.json' %% r), 'w').write(json.dumps({'rank': r, 'tot': tot, 'ranges': sim.ranges, 'rounds': len(rounds)}))\n"
        "dist.destroy_process_group()\n" % (ROOT, str(tmp_path)))
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", "29741", str(script)],
                         capture_output=True, text=True, timeout=240)
    assert out.returncode == 0, out.stderr[-2000:]
    recs = [json.loads((tmp_path / ("rank%d.json" % r)).read_text()) for r in range(2)]   # one file per rank: two ranks printing
    # to one pipe can interleave their lines
    assert len(recs) == 2 and recs[0]["tot"] == recs[1]["tot"] and recs[0]["rounds"] == recs[1]["rounds"]
    nr = recs[0]["rounds"]
    frames = nr * 2 * 128
    want_err = sum(1 for f in range(frames) if f % 7 == 3)
    assert recs[0]["tot"] == {"runs": frames, "errors": want_err, "biterrors": 2 * want_err, "reported": want_err, "undetected": 0}
    assert want_err >= 100 and sum(1 for f in range(frames - 256) if f % 7 == 3) < 100      # stopped in the first round that reached 100
    for k in range(nr):
        assert recs[0]["ranges"][k] == [k * 256, 128] and recs[1]["ranges"][k] == [k * 256 + 128, 128]


# ---------------------------------------------------------------- GPU
@pytest.mark.gpu
def test_philox_device_kat(gpu):
    import antpolarcodes_b200 as ap
    for ctr, key, want in PHILOX_KAT:
        assert [int(x) for x in ap.test_philox(ctr, key)] == want
    rng = np.random.default_rng(5)
    for _ in range(20):
        ctr, key = rng.integers(0, 1 << 32, 4, dtype=np.uint64), rng.integers(0, 1 << 32, 2, dtype=np.uint64)
        assert [int(x) for x in ap.test_philox(ctr, key)] == philox4x32_10(ctr, key)


def _sim(N, K, L, crc, dtype="char", batch=4096, ctype="BB", adaptive=False):
    import antpolarcodes_b200 as ap
    frozen = ap.frozen_bits(N, K, 0.0, ctype)
    return ap.LinkSimulator(N, L, frozen, dtype, crc=crc, batch=batch, adaptive=adaptive), frozen


@pytest.mark.gpu
@pytest.mark.parametrize("N,K,L,crc,dtype", [(1024, 512, 8, 16, "char"), (256, 128, 1, 0, "char"), (512, 256, 4, 8, "float"),
                                             (128, 64, 1, 8, "float"), (2048, 1024, 2, 32, "char")])
def test_sim_chain_against_oracle(gpu, N, K, L, crc, dtype):
    """One batch at 1.5 dB: the payload is the Philox stream of the frame, the sent block = payload + the oracle's CRC, the
    LLRs are the sent codeword plus noise of the right variance, the decisions are the oracle's on those LLRs, and the
    counters are what a recount of (sent, decoded, verdict) gives."""
    import torch
    from polar_testlib import (KIND_FASTSSC_F32, KIND_FASTSSC_I8, KIND_SCL_F32, KIND_SCL_I8, oracle_decode, oracle_encode, unpack_bits)
    sim, frozen = _sim(N, K, L, crc, dtype, batch=2048)
    counters = torch.zeros(5, dtype=torch.int64, device="cuda")
    seed, point, first, eb = 0x1234ABCD5678, 3, 1000, 1.5
    sim.run(eb, seed, point, first, 2048, counters)
    torch.cuda.synchronize()
    llr, sent, dec, ok = (t.cpu().numpy() for t in sim.last_batch())
    kb = K // 8
    # payload = Philox(counter = (frame, chunk, stream 0), key = (seed lo, seed hi ^ point * golden ratio))
    key = [seed & 0xFFFFFFFF, ((seed >> 32) ^ (point * 0x9E3779B9)) & 0xFFFFFFFF]
    for f in (0, 1, 777, 2047):
        fr = first + f
        want = b"".join(int(w).to_bytes(4, "little") for j in range((kb + 15) // 16)
                        for w in philox4x32_10([fr & 0xFFFFFFFF, fr >> 32, j, 0], key))[:kb]
        pay = kb - crc // 8 if crc != 32 else 4 * (kb // 4 - 1)
        assert bytes(sent[f][:pay]) == want[:pay]
    # CRC trailer + codeword: the oracle's encoder on the same payload reproduces `sent`, and sign(noise-free LLR) its codeword
    code, info = oracle_encode(N, frozen, sent.copy(), 1, crc)
    assert np.array_equal(info, sent)
    x = 1.0 - 2.0 * unpack_bits(code, N).astype(np.float64)
    amp = sim.default_amplification(eb)
    sigma = 1.0 / np.sqrt(2.0 * (K / N) * 10.0 ** (eb / 10.0))
    if dtype == "float":
        noise = llr.astype(np.float64) / amp - x
        assert abs(noise.mean()) < 5 * sigma / np.sqrt(noise.size)
        assert abs(noise.var() / sigma ** 2 - 1.0) < 0.01
        assert abs((noise ** 4).mean() / sigma ** 4 - 3.0) < 0.1           # Gaussian kurtosis
    else:
        resid = llr.astype(np.float64) / amp - x
        keep = np.abs(llr.astype(np.int32)) < 127                           # away from the clamp
        assert abs(resid[keep].mean()) < 0.01 and abs(resid.var() / (sigma ** 2 + 1.0 / 1200.0) - 1.0) < 0.03
    kind = {("char", True): KIND_SCL_I8, ("char", False): KIND_FASTSSC_I8, ("float", True): KIND_SCL_F32,
            ("float", False): KIND_FASTSSC_F32}[(dtype, L > 1)]
    eo, ek = oracle_decode(kind, N, frozen, llr, L, 1, crc)
    assert np.array_equal(eo, dec) and np.array_equal(ek, ok)
    be = np.unpackbits(sent ^ dec, axis=1).sum(axis=1)
    want = [2048, int((be > 0).sum()), int(be.sum()), int((ok == 0).sum()), int(((be > 0) & (ok != 0)).sum())]
    assert counters.tolist() == want and want[1] > 0


@pytest.mark.gpu
def test_sim_streams_do_not_depend_on_batching(gpu):
    """frames [0, 6000) in one call with 4096-frame batches == three calls with other batch sizes over the same range"""
    import torch
    a, _ = _sim(512, 256, 4, 16, batch=4096)
    b, _ = _sim(512, 256, 4, 16, batch=1000)
    ca = torch.zeros(5, dtype=torch.int64, device="cuda")
    cb = torch.zeros(5, dtype=torch.int64, device="cuda")
    a.run(1.0, 7, 1, 0, 6000, ca)
    for lo, n in ((3000, 3000), (0, 1234), (1234, 1766)):
        b.run(1.0, 7, 1, lo, n, cb)
    torch.cuda.synchronize()
    assert ca.tolist() == cb.tolist() and ca[0].item() == 6000 and ca[1].item() > 0
    cc = torch.zeros(5, dtype=torch.int64, device="cuda")
    a.run(1.0, 7, 2, 0, 6000, cc)   # another point = another stream
    torch.cuda.synchronize()
    assert cc.tolist() != ca.tolist()


@pytest.mark.gpu
def test_sweep_resume_csv_and_monotone(gpu, tmp_path):
    from antpolarcodes_b200 import bler
    out, csv = str(tmp_path / "s.jsonl"), str(tmp_path / "s.csv")
    kw = dict(Ns=[256], rate=0.5, configs=[(4, 8), (1, 0)], ebn0s=[0.0, 2.0, 4.0], min_errors=50, max_frames=60000, batch=4096,
              round_frames=8192, out=out, csv=csv, log=lambda s: None)
    recs = bler.run_sweep(**kw)
    assert len(recs) == 6 and [r["ebn0_db"] for r in recs[:3]] == [0.0, 2.0, 4.0]
    for cfg in (recs[:3], recs[3:]):
        b = [r["bler"] for r in cfg]
        assert b[0] > 0.05 and b[0] > b[1] > b[2], b
    for r in recs:
        assert r["runs"] % 8192 == 0 and (r["errors"] >= 50 or r["runs"] >= 60000)
        assert r["undetected"] <= r["errors"] <= r["undetected"] + r["reported"]
        assert r["bler_wilson95"][0] <= r["bler"] <= r["bler_wilson95"][1]
    lines = open(csv).read().strip().splitlines()
    assert lines[0] == bler.CSV_HEADER and len(lines) == 7 and lines[1].startswith("256,128,0.0,8,4,0.0,1,")
    # resume: nothing is simulated again (the records come back from the file, byte for byte)
    before = open(out).read()
    again = bler.run_sweep(**kw)
    assert open(out).read() == before and [r["key"] for r in again] == [r["key"] for r in recs]
    # one more point is appended, the old ones stay
    kw["ebn0s"] = [0.0, 2.0, 3.0, 4.0]
    more = bler.run_sweep(**kw)
    assert len(more) == 8 and len(open(out).read().strip().splitlines()) == 8


@pytest.mark.gpu
def test_sweep_adaptive_matches_list_decoder(gpu):
    """Same frames through the list decoder and through the adaptive decoder (Fast-SSC first): the block errors can only differ
    where Fast-SSC produced an undetected error."""
    import torch
    a, _ = _sim(512, 256, 8, 16, batch=8192)
    b, _ = _sim(512, 256, 8, 16, batch=8192, adaptive=True)
    ca = torch.zeros(5, dtype=torch.int64, device="cuda")
    cb = torch.zeros(5, dtype=torch.int64, device="cuda")
    a.run(1.5, 11, 0, 0, 32768, ca)
    b.run(1.5, 11, 0, 0, 32768, cb)
    torch.cuda.synchronize()
    assert ca[0] == cb[0] == 32768 and ca[1] > 100
    assert abs(int(ca[1]) - int(cb[1])) <= int(cb[4]) + int(ca[4]) + 2
