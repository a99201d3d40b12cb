#!/usr/bin/env python
"""TEST INFRASTRUCTURE: BASELINE configs[4] with its parity gate -- the BLER-vs-Eb/N0 sweep of antpolarcodes_b200.bler on
all visible GPUs, and per point a 2^16-frame subsample (the first batch of rank 0: LLRs, decisions and verdicts copied back)
decoded again by the UNMODIFIED reference on the host cores (oracle/_ref).  Per point the record gains

    subsample: {frames, mismatches (info bytes or verdict differ, must be 0), gpu_errors, cpu_errors (must be equal for the
                int8 decoders), wilson95 of the CPU subsample}
    ci95_ok:   the BLER of the full run lies inside that 95 % Wilson interval (1 point in 20 falls outside by construction)
    ci_ok:     ... inside the 99.9 % interval of the CPU subsample: the gate

(SURVEY.md 8d "parity checks run with every measurement").  The product never imports oracle/: this driver
lives in tests/ and hands the comparison to the sweep as a callback.

    python tests/bler_check.py --N 512,1024,4096 --configs 8:16,1:0 --ebn0 0:4:0.5 --out profiles/r2_bler.jsonl
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tests/bler_check.py ...
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    from antpolarcodes_b200 import bler
    import antpolarcodes_b200 as ap
    from polar_testlib import (KIND_FASTSSC_F32, KIND_FASTSSC_I8, KIND_SCL_F32, KIND_SCL_I8, have_ref, ref_decode_ex)
    p = argparse.ArgumentParser()
    p.add_argument("--N", default="512,1024,4096")
    p.add_argument("--rate", type=float, default=0.5)
    p.add_argument("--configs", default="8:16,1:0")
    p.add_argument("--construction", default="BB")
    p.add_argument("--type", default="char", choices=["char", "float"])
    p.add_argument("--ebn0", default="0:4:0.5")
    p.add_argument("--min-errors", type=int, default=200)
    p.add_argument("--max-frames", type=float, default=1e8)
    p.add_argument("--batch", type=int, default=1 << 16)
    p.add_argument("--round-frames", type=int, default=1 << 20)
    p.add_argument("--subsample", type=int, default=1 << 16, help="frames of the first batch re-decoded on the CPU (0 = off)")
    p.add_argument("--out", default=None)
    p.add_argument("--csv", default=None)
    a = p.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    threads = os.cpu_count() or 1
    state = {}

    def make_cb(N, K, L, crc, frozen):
        kind = {("char", True): KIND_SCL_I8, ("char", False): KIND_FASTSSC_I8, ("float", True): KIND_SCL_F32,
                ("float", False): KIND_FASTSSC_F32}[(a.type, L > 1)]

        def cb(llr, sent, dec, ok):
            n = min(a.subsample, llr.shape[0])
            if not n or not have_ref():
                return None
            t0 = time.perf_counter()
            h_llr = np.ascontiguousarray(llr[:n].cpu().numpy())
            h_sent, h_dec, h_ok = sent[:n].cpu().numpy(), dec[:n].cpu().numpy(), ok[:n].cpu().numpy()
            fresh = a.type == "float" and L > 1   # SclAvxFloat keeps history between frames
            ro, rk, _ = ref_decode_ex(kind, N, frozen, h_llr, L, 1, crc, threads, 1, fresh=fresh)
            mism = int(((ro != h_dec).any(axis=1) | (rk != h_ok)).sum())
            cpu_err = int((ro != h_sent).any(axis=1).sum())
            gpu_err = int((h_dec != h_sent).any(axis=1).sum())
            lo, hi = bler.wilson(cpu_err, n)
            lo3, hi3 = bler.wilson(cpu_err, n, z=3.2905)
            return {"frames": n, "mismatches": mism, "gpu_errors": gpu_err, "cpu_errors": cpu_err, "cpu_wilson95": [lo, hi],
                    "cpu_wilson999": [lo3, hi3],
                    "cpu_threads": threads, "cpu_seconds": time.perf_counter() - t0}
        return cb

    # the sweep calls one callback for every point; it needs the code of the point, which run_sweep builds itself: wrap it
    class PerCode:
        def __call__(self, llr, sent, dec, ok):
            N = llr.shape[1]
            K = int(round(N * a.rate / 8.0)) * 8
            key = (N, state["L"], state["crc"])
            if key not in state:
                state[key] = make_cb(N, K, state["L"], state["crc"], ap.frozen_bits(N, K, 0.0, a.construction))
            return state[key](llr, sent, dec, ok)

    if int(os.environ.get("WORLD_SIZE", "1")) > 1:   # one process group for all the sweeps below
        import torch
        import torch.distributed as dist
        local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(local)
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    records = []
    cfgs = [tuple(int(x) for x in c.split(":")) for c in a.configs.split(",")]
    for (L, crc) in cfgs:   # one sweep per decoder configuration so that the callback knows (L, crc)
        state["L"], state["crc"] = L, crc
        recs = bler.run_sweep([int(x) for x in a.N.split(",")], a.rate, [(L, crc)], bler._parse_sweep(a.ebn0), a.type, False,
                              a.construction, 0.0, 0xB1E5, a.min_errors, a.max_frames, a.batch, a.round_frames, a.out, None, True,
                              PerCode() if a.subsample else None, log=lambda s: None)
        records += recs
    if rank == 0:
        bad = 0
        for r in records:
            sub = r.get("subsample")
            if sub:
                ok = sub["mismatches"] == 0 and sub["gpu_errors"] == sub["cpu_errors"] and r["ci_ok"]
                bad += 0 if ok else 1
            print(json.dumps({k: r[k] for k in ("N", "K", "L", "C", "ebn0_db", "runs", "errors", "bler", "bler_wilson95", "n_gpus",
                                                "frames_per_second", "stop") if k in r} |
                             {"subsample": sub, "ci95_ok": r.get("ci95_ok"), "ci_ok": r.get("ci_ok")}), flush=True)
        if a.csv:
            bler.write_csv(a.csv, records)
        out95 = sum(1 for r in records if r.get("subsample") and not r.get("ci95_ok", True))
        # a 95 % interval misses the true value for 1 point in 20 by construction: the gate is the 99.9 % interval, the count
        # of points outside the 95 % interval is reported next to its expectation
        print(json.dumps({"points": len(records), "points_failing_parity_or_ci": bad, "points_outside_cpu_wilson95": out95,
                          "expected_outside_wilson95": round(0.05 * sum(1 for r in records if r.get("subsample")), 1)}))
        rc = 1 if bad else 0
    else:
        rc = 0
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        import torch.distributed as dist
        dist.destroy_process_group()
    return rc


if __name__ == "__main__":
    sys.exit(main())
