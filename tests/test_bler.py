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
