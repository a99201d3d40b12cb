"""BLER-vs-Eb/N0 sweep on the GPUs (BASELINE configs[4]): what pcsim does with one CPU thread per job
(src/simulation/simulator.cpp:607-985) as batched device work.  Every stage of the worker loop is a kernel of
libpolar_cuda (LinkSimulator: Philox payload -> CRC + encoder -> BPSK / AWGN / amplifier / quantiser -> decoder -> error
counters); the host only launches rounds and reads five counters per round.

    python -m antpolarcodes_b200.bler --N 512,1024,4096 --rate 0.5 --configs 8:16,1:0 --ebn0 0:4:0.5 \
           --min-errors 200 --max-frames 1e8 --out sweep.jsonl --csv sweep.csv
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 -m antpolarcodes_b200.bler ...

Multi-GPU: a point's frames are numbered globally; round r gives rank k the frames [r*W*R + k*R, r*W*R + (k+1)*R) (W ranks,
R frames per rank and round) and ends with ONE SUM all-reduce of the five 64-bit counters (NCCL over NVLink under torchrun),
which is also where every rank learns whether the stop rule (>= min-errors block errors or >= max-frames frames) holds.
The random streams are keyed by (seed, point) and counted by the global frame number, so the frames of a point do not depend
on the number of GPUs.  Resume: every finished point is appended to the JSONL output; points found there are skipped.

Channel model as in the reference's simulator: sigma = 1/sqrt(2 R 10^(EbN0/10)) with R = K/N (K includes the CRC bits,
simulator.cpp:821-838, awgn.cpp:30-43), BPSK 0 -> +1 (bpsk.cpp:68-80); 8-bit decoders see clamp(rne(10 y)) (amplification 10,
setup.cpp:58; quantiser bitcontainer.cpp:468-516), float decoders 4 * 10^(EbN0/10) * y (simulator.cpp:370-378)."""
import argparse
import json
import math
import os
import time

CSV_HEADER = ('"N","K","dSNR","C","L","Eb/N0","BPS","BLER","BER","RER","Runs","Errors","Time","Blockspeed","Coded Bitrate",'
              '"Payload Bitrate","Effective Payload Bitrate","Encoder Bitrate","Amplification","time min","time max","time mean",'
              '"time deviation"')  # simulator.cpp:519-527


def _parse_sweep(text):
    if ":" in text:
        lo, hi, step = (float(x) for x in text.split(":"))
        n = int(round((hi - lo) / step)) + 1
        return [lo + i * step for i in range(n)]
    return [float(x) for x in text.split(",")]


def wilson(k, n, z=1.959964):
    """95 % Wilson score interval of k errors in n trials."""
    if n <= 0:
        return 0.0, 1.0
    p = k / n
    d = 1.0 + z * z / n
    c = (p + z * z / (2 * n)) / d
    h = z * math.sqrt(p * (1 - p) / n + z * z / (4.0 * n * n)) / d
    return max(0.0, c - h), min(1.0, c + h)


def point_key(N, K, L, crc, dtype, adaptive, construction, dsnr, ebn0, seed):
    return "N%d_K%d_L%d_C%d_%s%s_%s%g_E%.3f_S%d" % (N, K, L, crc, dtype, "_ad" if adaptive else "", construction, dsnr, ebn0, seed)


def simulate_point(sim, ebn0_db, seed, point, rank, world, device, min_errors, max_frames, round_frames, on_subsample=None,
                   amplification=None):
    """One Eb/N0 point on all ranks.  Returns (totals dict, per-round seconds list, subsample result or None)."""
    import torch
    import torch.distributed as dist
    counters = torch.zeros(5, dtype=torch.int64, device=device)
    cuda = torch.device(device).type == "cuda"
    stream = torch.cuda.current_stream(device) if cuda else None
    sp = stream.cuda_stream if cuda else 0
    done = 0          # frames of the whole job so far
    rounds = []
    sub = None
    tot = [0] * 5
    while True:
        first = done + rank * round_frames   # ranks take consecutive slices of the round's global frame range
        if cuda:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
        t0 = time.perf_counter()
        if on_subsample is not None and done == 0 and rank == 0:
            # the first batch of the point is what the CPU reference re-decodes (tests/bler_check.py)
            n0 = min(sim.batch, round_frames)
            sim.run(ebn0_db, seed, point, first, n0, counters, amplification, sp)
            if cuda:
                torch.cuda.synchronize(device)
            sub = on_subsample(*sim.last_batch())
            if round_frames > n0:
                sim.run(ebn0_db, seed, point, first + n0, round_frames - n0, counters, amplification, sp)
        else:
            sim.run(ebn0_db, seed, point, first, round_frames, counters, amplification, sp)
        if cuda:
            e1.record(stream)
        done += world * round_frames
        total = counters.clone()
        if world > 1:
            dist.all_reduce(total, op=dist.ReduceOp.SUM)   # the one collective of the job
        tot = [int(v) for v in total.tolist()]             # also the synchronisation point of the round
        rounds.append(e0.elapsed_time(e1) * 1e-3 if cuda else time.perf_counter() - t0)
        if tot[1] >= min_errors or done >= max_frames:
            break
    return dict(zip(("runs", "errors", "biterrors", "reported", "undetected"), tot)), rounds, sub


def run_sweep(Ns, rate, configs, ebn0s, dtype="char", adaptive=False, construction="BB", design_snr=0.0, seed=0xB1E5,
              min_errors=200, max_frames=1e8, batch=1 << 16, round_frames=1 << 20, out=None, csv=None, resume=True,
              on_subsample=None, log=print):
    """The whole sweep.  configs = [(L, crc), ...].  Returns the list of point records (rank 0; [] elsewhere)."""
    import torch
    import torch.distributed as dist

    import antpolarcodes_b200 as ap

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("the BLER sweep needs a CUDA device (there is no CPU fallback)")
    device = torch.device("cuda", local)
    torch.cuda.set_device(device)
    own_pg = False
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=device)
        own_pg = True
    have = {}
    if out and resume and os.path.exists(out):
        for line in open(out):
            try:
                r = json.loads(line)
                have[r["key"]] = r
            except Exception:
                pass
    records = []
    point = 0
    for N in Ns:
        K = int(round(N * rate / 8.0)) * 8
        frozen = ap.frozen_bits(N, K, design_snr, construction)
        for (L, crc) in configs:
            sim = None
            for eb in ebn0s:
                point += 1
                key = point_key(N, K, L, crc, dtype, adaptive, construction, design_snr, eb, seed)
                if key in have:                 # resume: the point is already in the output file
                    records.append(have[key])
                    continue
                if sim is None:
                    sim = ap.LinkSimulator(N, L, frozen, dtype, crc=crc, adaptive=adaptive, device=local, batch=batch)
                t0 = time.perf_counter()
                # rounds are a multiple of the batch and, at high SNR, large enough that the counter read-back is rare
                rf = max(batch, (int(round_frames) // batch) * batch)
                tot, rounds, sub = simulate_point(sim, eb, seed, point, rank, world, device, min_errors, int(max_frames), rf,
                                                  on_subsample)
                wall = time.perf_counter() - t0
                if rank == 0:
                    runs = max(tot["runs"], 1)
                    bler = tot["errors"] / runs
                    lo, hi = wilson(tot["errors"], runs)
                    tsum = sum(rounds)
                    per_block = [r / rf for r in rounds]   # seconds per block, per round (this rank)
                    mean = sum(per_block) / len(per_block)
                    dev = math.sqrt(sum((x - mean) ** 2 for x in per_block) / len(per_block))
                    rec = {"key": key, "N": N, "K": K, "dSNR": design_snr, "C": crc, "L": L, "ebn0_db": eb, "BPS": 1,
                           "type": dtype, "adaptive": bool(adaptive), "construction": construction, "seed": seed, "n_gpus": world,
                           "runs": tot["runs"], "errors": tot["errors"], "biterrors": tot["biterrors"], "reported": tot["reported"],
                           "undetected": tot["undetected"], "bler": bler, "bler_wilson95": [lo, hi],
                           "ber": tot["biterrors"] / (runs * float(K)), "rer": tot["reported"] / runs,
                           "amplification": sim.default_amplification(eb), "gpu_seconds": tsum, "wall_seconds": wall,
                           "rounds": len(rounds), "frames_per_second": tot["runs"] / max(tsum, 1e-9),
                           "time_per_block_ns": {"min": min(per_block) * 1e9, "max": max(per_block) * 1e9, "mean": mean * 1e9, "dev": dev * 1e9},
                           "stop": "errors" if tot["errors"] >= min_errors else "max_frames"}
                    if sub is not None:   # a checker's verdict on the first batch (tests/bler_check.py: the CPU reference)
                        rec["subsample"] = sub
                        if "cpu_wilson95" in sub:
                            rec["ci95_ok"] = bool(sub["cpu_wilson95"][0] <= bler <= sub["cpu_wilson95"][1])
                            wide = sub.get("cpu_wilson999", sub["cpu_wilson95"])
                            rec["ci_ok"] = bool(wide[0] <= bler <= wide[1])
                    records.append(rec)
                    if out:
                        with open(out, "a") as f:
                            f.write(json.dumps(rec) + "\n")
                    log(json.dumps(rec))
            if sim is not None:
                sim.close()
    if rank == 0 and csv:
        write_csv(csv, records)
    if own_pg:
        dist.destroy_process_group()
    return records if rank == 0 else []


def write_csv(path, records):
    """pcsim's result table (Simulator::saveResults, simulator.cpp:513-563), same column names and order.  Time columns:
    `Time` is the GPU time of the point, the per-block time statistics are taken over the rounds of the point."""
    with open(path, "w") as f:
        f.write(CSV_HEADER + "\n")
        for r in records:
            runs, t = max(r["runs"], 1), max(r["gpu_seconds"], 1e-12)
            val = lambda x: ("%g" % x) if x > 0.0 else "1e-99"   # simulator.cpp:535-546
            tb = r["time_per_block_ns"]
            f.write(",".join(str(x) for x in (
                r["N"], r["K"], r["dSNR"], r["C"], r["L"], r["ebn0_db"], r["BPS"], val(r["bler"]), val(r["ber"]), val(r["rer"]),
                r["runs"], r["errors"], "%g" % t, "%g" % (runs / t), "%g" % (runs * r["N"] / t),
                "%g" % (runs * (r["K"] - r["C"]) / t), "%g" % ((runs - r["errors"]) * (r["K"] - r["C"]) / t),
                "%g" % (runs * r["N"] / t), "%g" % r["amplification"], int(tb["min"]), int(tb["max"]), int(tb["mean"]),
                int(tb["dev"]))) + "\n")


def main(argv=None):
    p = argparse.ArgumentParser()
    p.add_argument("--N", default="1024", help="block lengths, comma separated")
    p.add_argument("--rate", type=float, default=0.5)
    p.add_argument("--configs", default="8:16", help="listSize:crcBits pairs, comma separated (1:0 = Fast-SSC without a detector)")
    p.add_argument("--construction", default="BB")
    p.add_argument("--design-snr", type=float, default=0.0)
    p.add_argument("--type", default="char", choices=["char", "float", "mixed"])
    p.add_argument("--adaptive", action="store_true",
                   help="Fast-SSC first, the list decoder only on detector failures (the Adaptive* decoders pcsim uses by default)")
    p.add_argument("--ebn0", default="0:4:0.5")
    p.add_argument("--min-errors", type=int, default=200)
    p.add_argument("--max-frames", type=float, default=1e8)
    p.add_argument("--batch", type=int, default=1 << 16)
    p.add_argument("--round-frames", type=int, default=1 << 20, help="frames per rank between two counter all-reduces")
    p.add_argument("--seed", type=int, default=0xB1E5)
    p.add_argument("--out", default=None, help="JSONL output (appended; finished points are skipped on a re-run)")
    p.add_argument("--csv", default=None, help="pcsim-style CSV of all points")
    p.add_argument("--no-resume", action="store_true")
    a = p.parse_args(argv)
    cfgs = [tuple(int(x) for x in c.split(":")) for c in a.configs.split(",")]
    run_sweep([int(x) for x in a.N.split(",")], a.rate, cfgs, _parse_sweep(a.ebn0), a.type, a.adaptive or a.type == "mixed",
              a.construction, a.design_snr, a.seed, a.min_errors, a.max_frames, a.batch, a.round_frames, a.out, a.csv,
              not a.no_resume)


if __name__ == "__main__":
    main()
