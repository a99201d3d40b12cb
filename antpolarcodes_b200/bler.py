"""BLER-vs-Eb/N0 sweep on the GPUs (BASELINE configs[4]): what pcsim does with one CPU thread per
job (src/simulation/simulator.cpp:607-985) as batched device work -- payload -> CRC -> encode -> BPSK -> AWGN ->
LLR -> decode -> compare, every stage on the device, frames sharded by rank, one SUM all-reduce of the counters
per point (sharding.reduce_counters; NCCL under torchrun).

    python -m antpolarcodes_b200.bler --N 1024 --K 512 --L 8 --crc 16 --ebn0 0:4:0.5 --frames 1000000
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 -m antpolarcodes_b200.bler ...

Channel model as in the reference's simulator: sigma = 1/sqrt(2 R 10^(EbN0/10)) with R = K/N (K includes the CRC
bits, simulator.cpp:832-838), BPSK 0 -> +1; int8 LLR = clamp(rne(10 y)) (setup.cpp:58), fp32 LLR = 2 y / sigma^2."""
import argparse
import json
import math
import os

import numpy as np


def _parse_sweep(text):
    if ":" in text:
        lo, hi, step = (float(x) for x in text.split(":"))
        n = int(round((hi - lo) / step)) + 1
        return [lo + i * step for i in range(n)]
    return [float(x) for x in text.split(",")]


def run_point(torch, enc, dec, N, K, crc, ebn0_db, frames, batch, gen, device, fmt="i8"):
    """Simulate `frames` frames at one Eb/N0 on this rank; returns the local counters."""
    sigma = 1.0 / math.sqrt(2.0 * (K / N) * 10.0 ** (ebn0_db / 10.0))
    shifts = torch.arange(7, -1, -1, device=device, dtype=torch.uint8)
    kb = K // 8
    payload_bytes = kb - crc // 8
    cnt = {"frames": 0, "block_errors": 0, "bit_errors": 0, "detected": 0, "undetected": 0}
    done = 0
    while done < frames:
        n = min(batch, frames - done)
        info = torch.randint(0, 256, (n, kb), dtype=torch.uint8, device=device, generator=gen)
        code = enc.encode_batch(info)                      # writes the CRC bytes into `info`
        torch.cuda.synchronize(device)
        bits = ((code.unsqueeze(-1) >> shifts) & 1).reshape(n, N).to(torch.float32)
        y = (1.0 - 2.0 * bits) + sigma * torch.randn((n, N), device=device, generator=gen)
        if fmt == "f32":
            llr = (2.0 / (sigma * sigma)) * y
        else:
            llr = torch.clamp(torch.round(10.0 * y), -128, 127).to(torch.int8)
        out, ok = dec.decode_batch(llr)
        torch.cuda.synchronize(device)
        diff = out[:, :payload_bytes] ^ info[:, :payload_bytes]
        wrong = diff.any(dim=1)
        biterr = ((diff.unsqueeze(-1) >> shifts) & 1).sum()
        cnt["frames"] += n
        cnt["block_errors"] += int(wrong.sum().item())
        cnt["bit_errors"] += int(biterr.item())
        cnt["detected"] += int((ok == 0).sum().item())
        cnt["undetected"] += int((wrong & (ok != 0)).sum().item())
        done += n
    return cnt


def main(argv=None):
    import torch
    import torch.distributed as dist

    import antpolarcodes_b200 as ap
    from antpolarcodes_b200.sharding import frame_range, reduce_counters

    p = argparse.ArgumentParser()
    p.add_argument("--N", type=int, default=1024)
    p.add_argument("--K", type=int, default=512)
    p.add_argument("--L", type=int, default=8)
    p.add_argument("--crc", type=int, default=16)
    p.add_argument("--construction", default="5G")
    p.add_argument("--design-snr", type=float, default=0.0)
    p.add_argument("--type", default="char", choices=["char", "float", "mixed"])
    p.add_argument("--adaptive", action="store_true",
                   help="Fast-SSC first, the list decoder only on detector failures (the Adaptive* decoders pcsim uses by default)")
    p.add_argument("--ebn0", default="0:4:1")
    p.add_argument("--frames", type=float, default=1e5, help="frames per point (whole job)")
    p.add_argument("--batch", type=int, default=1 << 16)
    a = p.parse_args(argv)
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("the BLER sweep needs a CUDA device (there is no CPU fallback)")
    device = torch.device("cuda", local)
    torch.cuda.set_device(device)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=device)
    frozen = ap.frozen_bits(a.N, a.K, a.design_snr, a.construction)
    enc = ap.PolarEncoder(a.N, frozen, device=local)
    enc.setErrorDetection(a.crc)
    dec = ap.PolarDecoder(a.N, a.L, frozen, a.type, device=local, adaptive=a.adaptive or a.type == "mixed")
    dec.setErrorDetection(a.crc)
    gen = torch.Generator(device=device)
    lo, hi = frame_range(int(a.frames), rank, world)
    for i, eb in enumerate(_parse_sweep(a.ebn0)):
        gen.manual_seed(0xB1E5 + 1000 * i + rank)
        c = run_point(torch, enc, dec, a.N, a.K, a.crc, eb, hi - lo, a.batch, gen, device, "i8" if a.type == "char" else "f32")
        tot = reduce_counters(c, device)                   # the one collective of the job
        if rank == 0:
            f = max(tot["frames"], 1)
            bler = tot["block_errors"] / f
            ci = 1.96 * math.sqrt(max(bler * (1 - bler), 1e-12) / f)
            print(json.dumps({"ebn0_db": eb, "N": a.N, "K": a.K, "L": a.L, "crc": a.crc, "type": a.type, "n_gpus": world,
                              "frames": tot["frames"], "bler": bler, "bler_ci95": ci,
                              "ber": tot["bit_errors"] / (f * (a.K - a.crc)), "detected": tot["detected"],
                              "undetected": tot["undetected"], "adaptive": bool(a.adaptive or a.type == "mixed")}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
