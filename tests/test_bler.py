"""BLER sweep driver (antpolarcodes_b200/bler.py): the whole chain on the device -- payload, CRC, encoder, AWGN,
quantiser, decoder, counters -- plus the counter reduction (single rank here; world_size 2 on gloo is covered by
test_host.py::test_sharding_world2_gloo)."""
import json

import numpy as np
import pytest


def test_sweep_parser():
    from antpolarcodes_b200.bler import _parse_sweep
    assert _parse_sweep("0:4:1") == [0.0, 1.0, 2.0, 3.0, 4.0]
    assert _parse_sweep("1.5,2") == [1.5, 2.0]


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", ["char", "float"])
def test_bler_sweep(gpu, capsys, dtype):
    from antpolarcodes_b200 import bler
    bler.main(["--N", "256", "--K", "128", "--L", "4", "--crc", "8", "--construction", "BB", "--type", dtype,
               "--ebn0", "0,2,4", "--frames", "30000", "--batch", "8192"])
    pts = [json.loads(l) for l in capsys.readouterr().out.strip().splitlines()]
    assert [p["ebn0_db"] for p in pts] == [0.0, 2.0, 4.0]
    assert all(p["frames"] == 30000 for p in pts)
    b = [p["bler"] for p in pts]
    assert b[0] > 0.05 and b[0] > b[1] > b[2] and b[2] < 5e-3, b
    # a block error is either flagged by the detector or counted as undetected
    for p in pts:
        assert p["undetected"] <= p["bler"] * p["frames"] + 1e-9
        assert p["detected"] + p["undetected"] >= p["bler"] * p["frames"] - 1e-9


@pytest.mark.gpu
def test_bler_sweep_adaptive_equals_list_decoder(gpu, capsys):
    """Same seeds, same frames: wherever Fast-SSC passes the CRC the adaptive decoder keeps its answer, so its block
    error rate can only differ from the list decoder's through undetected Fast-SSC errors (rare with CRC16)."""
    from antpolarcodes_b200 import bler
    common = ["--N", "512", "--K", "256", "--L", "8", "--crc", "16", "--construction", "BB", "--ebn0", "1,3",
              "--frames", "20000", "--batch", "8192"]
    bler.main(common)
    plain = [json.loads(l) for l in capsys.readouterr().out.strip().splitlines()]
    bler.main(common + ["--adaptive"])
    ada = [json.loads(l) for l in capsys.readouterr().out.strip().splitlines()]
    for p, a in zip(plain, ada):
        assert a["adaptive"] and a["frames"] == p["frames"]
        assert abs(a["bler"] - p["bler"]) <= 3 * p["bler_ci95"] + 1e-3, (p, a)
