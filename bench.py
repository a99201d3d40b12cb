#!/usr/bin/env python
"""bench.py -- info-bit Gbit/s of the batched polar decoders on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload scl8|fastssc|...]

A "step" is one decode pass over one batch of synthetic noisy codewords (AWGN/BPSK, Eb/N0 = 2 dB,
int8 LLRs).  Headline workload = BASELINE configs[2]: SCL L=8 + CRC16, N=1024 K=512,
5G reliability-table frozen set.  `value` is timed with the LLRs resident in HBM; `e2e` is the same
decode through the public API with pinned HOST buffers (H2D of the LLRs and D2H of the decoded bytes
inside the timed region).

Every measurement is gated on parity with the UNMODIFIED reference (oracle/_ref): the first `parity`
frames of the timed batch are decoded again by the reference's own decoder objects on the host cores and
info bytes + verdicts must be identical (`check.parity_mismatches` == 0, else exit code 3).

The default run (one GPU, headline workload) also measures the other configurations BASELINE.json names in
the same process and attaches them as `also.<name>` sub-records (Fast-SSC int8 C1, Fast-SSC fp32 + encoder
C2, SCL L=32 C4, the pcbench-style rows), each with its own value / e2e / roofline / cpu_baseline / check.

Rank 0 also times the unmodified reference on the host cores on a bounded sample (`cpu_baseline`);
`--impl reference` times only that, through nothing but oracle/_ref (no product code is imported).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

WORKLOADS = {
    # frames = frames per step per GPU; parity = frames of the timed batch re-decoded by the reference; cpu_s = CPU baseline seconds
    "scl8": dict(N=1024, K=512, ctype="5G", L=8, crc=16, frames=1 << 19, e2e_frames=1 << 19, parity=1 << 16, cpu_s=8.0,
                 label="SCL L=8 + CRC16, N=1024 K=512, 5G frozen set, int8 LLR, Eb/N0=2dB"),
    "fastssc": dict(N=1024, K=512, ctype="BB", L=1, crc=0, frames=1 << 20, e2e_frames=1 << 19, parity=1 << 18, cpu_s=3.0,
                    label="Fast-SSC int8, N=1024 K=512, BB frozen set, Eb/N0=2dB"),
    "fastssc_f32": dict(N=256, K=128, ctype="BB", L=1, crc=0, frames=1 << 20, e2e_frames=1 << 19, fmt="f32", parity=1 << 18, cpu_s=3.0,
                        label="Fast-SSC fp32, N=256 K=128, BB frozen set, Eb/N0=2dB"),
    "scl8_f32": dict(N=1024, K=512, ctype="5G", L=8, crc=16, frames=1 << 17, e2e_frames=1 << 17, fmt="f32", parity=1 << 13, cpu_s=4.0,
                     label="SCL fp32 L=8 + CRC16, N=1024 K=512, 5G frozen set, Eb/N0=2dB"),
    "encode": dict(N=256, K=128, ctype="BB", L=1, crc=0, frames=1 << 22, e2e_frames=1 << 21, encoder=True, parity=1 << 18, cpu_s=3.0,
                   label="packed butterfly encoder, systematic, N=256 K=128, BB frozen set"),
    "encode_n1024": dict(N=1024, K=512, ctype="5G", L=1, crc=16, frames=1 << 21, e2e_frames=1 << 20, encoder=True, parity=1 << 17, cpu_s=3.0,
                         label="packed butterfly encoder + CRC16 generate, systematic, N=1024 K=512, 5G frozen set"),
    "adaptive8": dict(N=1024, K=512, ctype="5G", L=8, crc=16, frames=1 << 20, e2e_frames=1 << 19, adaptive=True, parity=1 << 16, cpu_s=4.0,
                      label="AdaptiveChar (Fast-SSC int8, SCL L=8 on CRC16 failures), N=1024 K=512, 5G frozen set, Eb/N0=2dB"),
    "scl32": dict(N=2048, K=1024, ctype="BB", L=32, crc=32, frames=1 << 16, e2e_frames=1 << 16, parity=1 << 13, cpu_s=4.0,
                  label="SCL L=32 + CRC32, N=2048 K=1024, BB frozen set, int8 LLR, Eb/N0=2dB"),
    "scl8_n512": dict(N=512, K=256, ctype="BB", L=8, crc=16, frames=1 << 19, e2e_frames=1 << 17, parity=1 << 14, cpu_s=2.0,
                      label="SCL L=8 + CRC16, N=512 K=256, BB frozen set, int8 LLR, Eb/N0=2dB (occupancy experiments)"),
    "fastssc_soft": dict(N=1024, K=512, ctype="BB", L=1, crc=0, frames=1 << 18, e2e_frames=1 << 17, soft=True, parity=1 << 14, cpu_s=2.0,
                         label="Fast-SSC int8 soft outputs (getSoftCodeword + getSoftInformation), N=1024 K=512, BB frozen set, Eb/N0=2dB"),
    # pcbench-style comparability rows (benchmark/main_benchmark.cpp:202-241): iid N(0, 0.5) floats, not codewords, fed to the
    # int8 decoders through decode_vector(const float*), i.e. the quantiser runs inside the timed region
    "pcbench_scl8": dict(N=1024, K=512, ctype="5G", L=8, crc=16, frames=1 << 18, e2e_frames=1 << 17, pcbench=True, parity=1 << 14, cpu_s=3.0,
                         label="pcbench-style: SCL L=8 + CRC16 int8 decoder fed iid N(0,0.5) float inputs, N=1024 K=512, 5G frozen set"),
    "pcbench_fastssc": dict(N=1024, K=512, ctype="BB", L=1, crc=0, frames=1 << 19, e2e_frames=1 << 18, pcbench=True, parity=1 << 17, cpu_s=3.0,
                            label="pcbench-style: Fast-SSC int8 decoder fed iid N(0,0.5) float inputs, N=1024 K=512, BB frozen set"),
}
ALSO = ["fastssc", "fastssc_f32", "encode", "scl32", "scl8_f32", "fastssc_soft", "pcbench_scl8", "pcbench_fastssc"]
METRIC = "info-bit Gbps (N=1024 K=512, SCL L=8+CRC16)"
EBN0_DB = 2.0


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d.get("hbm_gbs", 6650.0)), float(d.get("sm_max_mhz", 1965.0)), "measured"
    return 6650.0, 1965.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 7 and r[3 + i].lower() == "active" for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def ref_threads():
    return os.cpu_count() or 1


def ref_kind(wl):
    from polar_testlib import KIND_ADAPTIVE_I8, KIND_FASTSSC_F32, KIND_FASTSSC_I8, KIND_SCL_F32, KIND_SCL_I8
    if wl.get("adaptive"):
        return KIND_ADAPTIVE_I8
    if wl.get("fmt") == "f32":
        return KIND_SCL_F32 if wl["L"] > 1 else KIND_FASTSSC_F32
    return KIND_SCL_I8 if wl["L"] > 1 else KIND_FASTSSC_I8


# ---------------------------------------------------------------- reference legs (oracle/_ref only)
def time_reference(wl, frozen, llr_host, seconds_target, threads):
    """The unmodified reference decoders on the host cores: one decoder object per thread, disjoint slices of the pool."""
    from polar_testlib import ref_decode_ex
    kind = ref_kind(wl)
    if wl.get("adaptive"):  # AdaptiveChar only takes float LLRs (adaptive_char.cpp:60-64); its quantiser maps them back
        llr_host = np.ascontiguousarray(llr_host, np.float32)
    n = llr_host.shape[0]
    _, _, t1 = ref_decode_ex(kind, wl["N"], frozen, llr_host, wl["L"], 1, wl["crc"], threads, 1)
    rep = max(1, int(seconds_target / max(t1, 1e-6)))
    _, _, t = ref_decode_ex(kind, wl["N"], frozen, llr_host, wl["L"], 1, wl["crc"], threads, rep)
    return n * rep * wl["K"] / t / 1e9, n * rep, t


def time_reference_encoder(wl, frozen, info_host, seconds_target, threads):
    """ButterflyFipPacked on the host cores: one encoder object per thread, each on its own slice."""
    from concurrent.futures import ThreadPoolExecutor
    from polar_testlib import ref_encode
    parts = np.array_split(info_host, threads)

    def work(a):
        ref_encode(wl["N"], frozen, a, 1, wl["crc"])
    with ThreadPoolExecutor(threads) as ex:
        t0 = time.perf_counter()
        list(ex.map(work, parts))
        t1 = time.perf_counter() - t0
        rep = max(1, int(seconds_target / max(t1, 1e-6)))
        t0 = time.perf_counter()
        for _ in range(rep):
            list(ex.map(work, parts))
        t = time.perf_counter() - t0
    n = info_host.shape[0]
    return n * rep * wl["K"] / t / 1e9, n * rep, t


def reference_pool(wl, frozen, n):
    """>= 4096 seeded frames of the workload, built with the reference's own encoder."""
    from polar_testlib import awgn_llrs, pcbench_llrs, ref_encode
    if wl.get("pcbench"):
        return pcbench_llrs(wl["N"], n, fmt="f32")
    llr, _ = awgn_llrs(wl["N"], frozen, n, EBN0_DB, wl["crc"], fmt=wl.get("fmt", "i8"), encode=ref_encode)
    return llr


def run_reference_arm(args, wl):
    """--impl reference: the reference's own CPU implementation, all host threads.  Nothing of the product is imported:
    frozen set, encoder, decoders all come from oracle/_ref (the unmodified reference library)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from polar_testlib import have_ref, ref_frozen_bits
    if not have_ref():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libpolar_ref.so not present"}))
        return
    frozen = ref_frozen_bits(wl["N"], wl["K"], 0.0, wl["ctype"])
    threads = ref_threads()
    sample = max(4 * threads, 4096)
    per_step = min(20.0 / max(args.steps + args.warmup, 1), 4.0)
    vals, ms = [], []
    if wl.get("encoder"):
        pool = np.random.default_rng(1).integers(0, 256, (1 << 18, wl["K"] // 8), dtype=np.uint8)
        sample = len(pool)
    else:
        pool = reference_pool(wl, frozen, sample)
    for s in range(args.warmup + args.steps):
        if wl.get("encoder"):
            v, nfr, t = time_reference_encoder(wl, frozen, pool, per_step, threads)
        else:
            v, nfr, t = time_reference(wl, frozen, pool, per_step, threads)
        if s >= args.warmup:
            vals.append(v)
            ms.append(t * 1e3)
    v = float(np.mean(vals))
    line = {"metric": METRIC if args.workload == "scl8" else wl["label"], "value": v, "unit": "Gbit/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": float(np.mean(ms)), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None,
            "dtype": "u8" if wl.get("encoder") else ("f32" if wl.get("fmt") == "f32" else "int8"), "data": "synthetic", "impl": "reference",
            "config": {"workload": wl["label"], "frames_per_step": "bounded sample (%d frames x repeats)" % sample},
            "cpu_baseline": {"value": v, "unit": "Gbit/s", "cores": threads, "kind": "reference",
                             "sample": "%d seeded frames cycled for ~%.1f s per step, one reference decoder / encoder object per thread" % (sample, per_step)},
            "e2e": {"value": v, "unit": "Gbit/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


# ---------------------------------------------------------------- our arm
class Ctx:
    def __init__(self):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback)")
        self.device = torch.device("cuda", self.local)
        torch.cuda.set_device(self.device)
        if self.world > 1:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("nccl", device_id=self.device)
        self.stream = torch.cuda.current_stream(self.device)
        self.sp = self.stream.cuda_stream

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize(self.device)

    def close(self):
        if self.world > 1:
            self.dist.destroy_process_group()


def make_llrs_on_device(ctx, ap, wl, frozen, frames, seed):
    """payload -> CRC -> (our) encoder -> BPSK -> AWGN -> int8 LLR = clamp(rne(10 y)); all on the device.
    pcbench rows: iid N(0, 0.5) floats instead (not codewords)."""
    torch, device = ctx.torch, ctx.device
    N, K, crc = wl["N"], wl["K"], wl["crc"]
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    if wl.get("pcbench"):
        return 0.5 * torch.randn((frames, N), device=device, generator=g, dtype=torch.float32), None
    enc = ap.PolarEncoder(N, frozen, device=device.index)
    enc.setSystematic(True)
    enc.setErrorDetection(crc)
    f32 = wl.get("fmt") == "f32"
    llr = torch.empty((frames, N), dtype=torch.float32 if f32 else torch.int8, device=device)
    info_all = torch.empty((frames, K // 8), dtype=torch.uint8, device=device)
    sigma = 1.0 / np.sqrt(2.0 * (K / N) * 10.0 ** (EBN0_DB / 10.0))
    shifts = torch.arange(7, -1, -1, device=device, dtype=torch.uint8)
    chunk = 1 << 16
    for lo in range(0, frames, chunk):
        n = min(chunk, frames - lo)
        info = torch.randint(0, 256, (n, K // 8), dtype=torch.uint8, device=device, generator=g)
        code = enc.encode_batch(info)
        torch.cuda.synchronize(device)
        bits = ((code.unsqueeze(-1) >> shifts) & 1).reshape(n, N).to(torch.float32)
        y = (1.0 - 2.0 * bits) + sigma * torch.randn((n, N), device=device, generator=g)
        if f32:
            llr[lo:lo + n] = (2.0 / (sigma * sigma)) * y
        else:
            llr[lo:lo + n] = torch.clamp(torch.round(10.0 * y), -128, 127).to(torch.int8)
        info_all[lo:lo + n] = info
    return llr, info_all


def parity_check(wl, frozen, llr_host, out_host, ok_host, threads):
    """info bytes + verdict of the timed batch against the compiled reference on the same LLRs (every
    measurement carries its parity check).  fp32 list decoder: one fresh reference object per frame."""
    from polar_testlib import have_ref, ref_decode_ex
    if not have_ref():
        return {"parity_frames": 0, "parity_mismatches": None, "parity_against": "oracle/_ref missing"}
    kind = ref_kind(wl)
    fresh = (wl.get("fmt") == "f32" and wl["L"] > 1)
    src = np.ascontiguousarray(llr_host, np.float32) if wl.get("adaptive") else llr_host
    t0 = time.perf_counter()
    ro, rk, _ = ref_decode_ex(kind, wl["N"], frozen, src, wl["L"], 1, wl["crc"], threads, 1, fresh=fresh)
    bad = np.flatnonzero((ro != out_host).any(axis=1) | (rk != ok_host))
    return {"parity_frames": int(len(llr_host)), "parity_mismatches": int(len(bad)),
            "parity_first_bad_frames": [int(b) for b in bad[:8]],
            "parity_against": "unmodified reference (oracle/_ref), %d host threads, %.1f s%s" % (threads, time.perf_counter() - t0,
                                                                                               ", fresh object per frame" if fresh else "")}


def measure_decoder(ctx, args, name, wl, steps, warmup, with_cpu=True):
    import antpolarcodes_b200 as ap
    torch, device, world, rank = ctx.torch, ctx.device, ctx.world, ctx.rank
    N, K, L, crc = wl["N"], wl["K"], wl["L"], wl["crc"]
    frames = args.frames or wl["frames"]
    e2e_frames = min(args.e2e_frames or wl["e2e_frames"], frames)
    frozen = ap.frozen_bits(N, K, 0.0, wl["ctype"])
    f32 = wl.get("fmt") == "f32"
    in_f32 = f32 or bool(wl.get("pcbench"))
    adaptive = bool(wl.get("adaptive"))
    dec = ap.PolarDecoder(N, L, frozen, "float" if f32 else "char", device=ctx.local, adaptive=adaptive)
    dec.setErrorDetection(crc)
    dec.setSystematic(True)
    info = dec.info()
    lane_ops_list = 0.0
    if adaptive:  # algorithmic work of the list pass, charged per frame that falls through
        lane_ops_list = ap._Code(ap.KIND_SCL_I8, N, frozen, L, True, crc, ctx.local).info().laneOpsPerFrame

    # frames are sharded by rank: every rank decodes its own `frames` frames (weak scaling), no
    # data-path collective
    llr, sent = make_llrs_on_device(ctx, ap, wl, frozen, frames, 0xB200 + rank)
    out = torch.empty((frames, K // 8), dtype=torch.uint8, device=device)
    ok = torch.empty((frames,), dtype=torch.uint8, device=device)
    stream, sp = ctx.stream, ctx.sp

    def step():
        dec.decode_batch(llr, out, ok, stream=sp)

    for _ in range(warmup):
        step()
    ctx.barrier()
    sampler = ClockSampler(ctx.local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    launches0 = dec._get().launches()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    ctx.barrier()
    ev[0].record(stream)
    for s in range(steps):
        step()
        ev[s + 1].record(stream)
    ctx.barrier()
    total_ms = ev[0].elapsed_time(ev[-1])
    kernel_ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(steps)]
    launches = dec._get().launches() - launches0
    clocks = sampler.stop() if rank == 0 else None
    redo_frac = 0.0
    if adaptive:
        fb0 = dec.fallbacks()
        step()
        torch.cuda.synchronize(device)
        redo_frac = (dec.fallbacks() - fb0) / frames

    # correctness of what was timed: frame errors vs the transmitted payload + verdicts
    frame_err = int((out != sent).any(dim=1).sum().item()) if sent is not None else -1
    ok_count = int(ok.sum().item())

    # e2e: pinned host LLRs -> C ABI with HOST pointers (H2D + kernel + D2H inside the timed region)
    h_llr = torch.empty((e2e_frames, N), dtype=llr.dtype).pin_memory()
    h_llr.copy_(llr[:e2e_frames])
    h_out = torch.empty((e2e_frames, K // 8), dtype=torch.uint8).pin_memory()
    h_ok = torch.empty((e2e_frames,), dtype=torch.uint8).pin_memory()
    for _ in range(max(1, min(warmup, 3))):
        dec.decode_batch(h_llr, h_out, h_ok, stream=sp)
    ctx.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(steps):
        dec.decode_batch(h_llr, h_out, h_ok, stream=sp)
    e1.record(stream)
    ctx.barrier()
    e2e_ms = e0.elapsed_time(e1)
    e2e_match = bool((h_out == out[:e2e_frames].cpu()).all().item())

    t = torch.tensor([total_ms, e2e_ms], dtype=torch.float64, device=device)
    errs = torch.tensor([max(frame_err, 0), ok_count, frames], dtype=torch.int64, device=device)
    if world > 1:
        ctx.dist.all_reduce(t, op=ctx.dist.ReduceOp.MAX)
        ctx.dist.all_reduce(errs, op=ctx.dist.ReduceOp.SUM)  # the BLER counter allreduce
    total_ms, e2e_ms = float(t[0].item()), float(t[1].item())
    if rank != 0:
        return None

    hbm_peak, sm_max, which = peaks()
    value = world * frames * steps * K / (total_ms * 1e-3) / 1e9
    e2e_val = world * e2e_frames * steps * K / (e2e_ms * 1e-3) / 1e9
    k_ms = float(np.mean(kernel_ms))
    fps = frames / (k_ms * 1e-3)
    sm_mhz = (clocks or {}).get("sm_mhz") or sm_max
    alu_peak_max = 148 * 128 * sm_max * 1e6 / 1e12     # T lane-ops/s at max clock
    alu_peak_run = 148 * 128 * sm_mhz * 1e6 / 1e12
    lane_ops = info.laneOpsPerFrame + redo_frac * lane_ops_list
    bytes_per_frame = info.bytesPerFrame + (3 * N if (in_f32 and not f32) else 0)   # float LLRs into an int8 decoder: 4N in
    alu = fps * lane_ops / 1e12
    hbm = fps * bytes_per_frame / 1e9
    alu_bound = (lane_ops / (alu_peak_max * 1e12)) > (bytes_per_frame / (hbm_peak * 1e9))
    traffic, traffic_src = None, None
    try:  # DRAM bytes per launch from the committed ncu capture of this kernel (profiles/traffic.json)
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(name)
        if tj:
            traffic, traffic_src = tj["dram_bytes"] / tj["frames"] * frames, tj["source"]
    except Exception:
        pass
    roof = {"bound": "alu_issue" if alu_bound else "hbm",
            "achieved": alu if alu_bound else hbm, "peak": alu_peak_max if alu_bound else hbm_peak,
            "unit": "Tlane-op/s" if alu_bound else "GB/s",
            "frac": (alu / alu_peak_max) if alu_bound else hbm / hbm_peak, "traffic": traffic,
            "peak_source": which + (" sm_max_mhz x 148 SM x 128 lanes" if alu_bound else " hbm copy"),
            "frac_at_run_clock": alu / alu_peak_run,
            "alu": {"lane_ops_per_frame": lane_ops, "achieved_T": alu, "peak_T": alu_peak_max},
            "hbm": {"bytes_per_frame": bytes_per_frame, "achieved_GBs": hbm, "peak_GBs": hbm_peak,
                    "frac": hbm / hbm_peak},
            "traffic_source": traffic_src, "algorithmic_bytes": bytes_per_frame * frames,
            "kernel_ms": k_ms, "frames_per_launch": frames}
    th = ref_threads()
    check = {"frames": int(errs[2].item()), "frame_errors_vs_sent": int(errs[0].item()) if sent is not None else None,
             "detector_pass": int(errs[1].item())}
    npar = min(args.parity if args.parity >= 0 else wl["parity"], frames)
    if npar:
        check.update(parity_check(wl, frozen, np.ascontiguousarray(llr[:npar].cpu().numpy()), out[:npar].cpu().numpy(),
                                  ok[:npar].cpu().numpy(), th))
    cpu = None
    if with_cpu:
        try:
            from polar_testlib import have_ref
            if have_ref():
                sample = min(max(4 * th, 4096), frames)
                v, nfr, tt = time_reference(wl, frozen, np.ascontiguousarray(llr[:sample].cpu().numpy()), wl["cpu_s"], th)
                cpu = {"value": v, "unit": "Gbit/s", "cores": th, "kind": "reference",
                       "sample": "%d frames of this workload (%.1f MB, larger than the host L2), decoded %d times in %.1f s by %d threads (one reference decoder object each)"
                                 % (sample, sample * N * (4 if in_f32 else 1) / 1e6, nfr // sample, tt, th)}
        except Exception as e:  # the baseline is a reported number, never a reason to lose the bench line
            cpu = {"value": None, "unit": "Gbit/s", "cores": 0, "kind": "reference", "sample": "failed: %r" % (e,)}
    return {"metric": METRIC if name == "scl8" else wl["label"], "value": value, "unit": "Gbit/s",
            "n_gpus": world, "steps": steps, "warmup": warmup, "ms_per_step": total_ms / steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32" if f32 else "int8", "data": "synthetic",
            "config": {"workload": wl["label"], "frames_per_step_per_gpu": frames, "e2e_frames_per_step_per_gpu": e2e_frames,
                       "list_size": L, "crc_bits": crc, "cache": "inputs larger than L2 (%d MiB of LLRs per step)" % (frames * N * (4 if in_f32 else 1) >> 20),
                       "warps_per_block": info.warpsPerBlock, "blocks_per_sm": info.blocksPerSM,
                       "smem_per_warp": info.smemPerWarp, "schedule_ops": info.nOps, "parallelism": "frames sharded x%d" % world,
                       **({"list_decoder_frame_fraction": redo_frac} if adaptive else {})},
            "e2e": {"value": e2e_val, "unit": "Gbit/s", "h2d_bytes_per_step": e2e_frames * N * (4 if in_f32 else 1),
                    "d2h_bytes_per_step": e2e_frames * (K // 8 + 1), "matches_device_run": e2e_match},
            "gpu_launches": int(launches), "clocks": clocks, "roofline": roof, "cpu_baseline": cpu,
            "speedup_e2e_vs_cpu_baseline": (e2e_val / cpu["value"]) if cpu and cpu.get("value") else None,
            "check": check}


def measure_encoder(ctx, args, name, wl, steps, warmup, with_cpu=True):
    """Encoder::encode_vector over a batch (configs[1]): info bytes -> codeword bytes, CRC fused when configured."""
    import antpolarcodes_b200 as ap
    torch, device, world, rank = ctx.torch, ctx.device, ctx.world, ctx.rank
    N, K, crc = wl["N"], wl["K"], wl["crc"]
    frames = args.frames or wl["frames"]
    e2e_frames = min(args.e2e_frames or wl["e2e_frames"], frames)
    frozen = ap.frozen_bits(N, K, 0.0, wl["ctype"])
    enc = ap.PolarEncoder(N, frozen, device=ctx.local)
    enc.setSystematic(True)
    enc.setErrorDetection(crc)
    g = torch.Generator(device=device)
    g.manual_seed(0xE0C + rank)
    info = torch.randint(0, 256, (frames, K // 8), dtype=torch.uint8, device=device, generator=g)
    code = torch.empty((frames, N // 8), dtype=torch.uint8, device=device)
    stream, sp = ctx.stream, ctx.sp
    for _ in range(warmup):
        enc.encode_batch(info, code, stream=sp)
    ctx.barrier()
    sampler = ClockSampler(ctx.local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    l0 = enc._get().launches()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    ctx.barrier()
    ev[0].record(stream)
    for s_ in range(steps):
        enc.encode_batch(info, code, stream=sp)
        ev[s_ + 1].record(stream)
    ctx.barrier()
    total_ms = ev[0].elapsed_time(ev[-1])
    k_ms = float(np.mean([ev[i].elapsed_time(ev[i + 1]) for i in range(steps)]))
    launches = enc._get().launches() - l0
    clocks = sampler.stop() if rank == 0 else None
    # e2e through host buffers
    h_info = torch.empty((e2e_frames, K // 8), dtype=torch.uint8).pin_memory()
    h_info.copy_(info[:e2e_frames])
    h_code = torch.empty((e2e_frames, N // 8), dtype=torch.uint8).pin_memory()
    for _ in range(2):
        enc.encode_batch(h_info, h_code, stream=sp)
    ctx.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(steps):
        enc.encode_batch(h_info, h_code, stream=sp)
    e1.record(stream)
    ctx.barrier()
    e2e_ms = e0.elapsed_time(e1)
    e2e_match = bool((h_code == code[:e2e_frames].cpu()).all().item())
    t = torch.tensor([total_ms, e2e_ms], dtype=torch.float64, device=device)
    if world > 1:
        ctx.dist.all_reduce(t, op=ctx.dist.ReduceOp.MAX)
    total_ms, e2e_ms = float(t[0].item()), float(t[1].item())
    if rank != 0:
        return None
    hbm_peak, sm_max, which = peaks()
    value = world * frames * steps * K / (total_ms * 1e-3) / 1e9
    e2e_val = world * e2e_frames * steps * K / (e2e_ms * 1e-3) / 1e9
    bpf = K // 8 + N // 8
    hbm = frames * bpf / (k_ms * 1e-3) / 1e9
    traffic = None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(name)
        if tj:
            traffic = tj["dram_bytes"] / tj["frames"] * frames
    except Exception:
        pass
    roof = {"bound": "hbm", "achieved": hbm, "peak": hbm_peak, "unit": "GB/s", "frac": hbm / hbm_peak, "traffic": traffic,
            "peak_source": which + " hbm copy", "algorithmic_bytes": bpf * frames, "kernel_ms": k_ms,
            "frames_per_launch": frames}
    th = ref_threads()
    check = {"frames": frames}
    cpu = None
    try:
        from polar_testlib import have_ref, ref_encode
        if have_ref():
            npar = min(args.parity if args.parity >= 0 else wl["parity"], frames)
            if npar:  # codewords of the timed batch against the reference's ButterflyFipPacked
                rc, ri = ref_encode(N, frozen, np.ascontiguousarray(info[:npar].cpu().numpy()), 1, crc)
                bad = int(((rc != code[:npar].cpu().numpy()).any(axis=1)).sum())
                check.update({"parity_frames": npar, "parity_mismatches": bad, "parity_against": "unmodified reference (oracle/_ref)"})
            if with_cpu:
                sample = np.ascontiguousarray(h_info[:1 << 18].numpy())
                v, nfr, tt = time_reference_encoder(wl, frozen, sample, wl["cpu_s"], th)
                cpu = {"value": v, "unit": "Gbit/s", "cores": th, "kind": "reference",
                       "sample": "%d frames encoded %d times in %.1f s by %d threads (one ButterflyFipPacked each)" % (len(sample), nfr // len(sample), tt, th)}
    except Exception as e:
        cpu = {"value": None, "unit": "Gbit/s", "cores": 0, "kind": "reference", "sample": "failed: %r" % (e,)}
    return {"metric": wl["label"], "value": value, "unit": "Gbit/s", "n_gpus": world, "steps": steps, "warmup": warmup,
            "ms_per_step": total_ms / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic",
            "config": {"workload": wl["label"], "frames_per_step_per_gpu": frames, "e2e_frames_per_step_per_gpu": e2e_frames,
                       "crc_bits": crc, "cache": "inputs larger than L2 (%d MiB in+out per step)" % (frames * bpf >> 20),
                       "parallelism": "frames sharded x%d" % world},
            "e2e": {"value": e2e_val, "unit": "Gbit/s", "h2d_bytes_per_step": e2e_frames * (K // 8),
                    "d2h_bytes_per_step": e2e_frames * (N // 8), "matches_device_run": e2e_match},
            "gpu_launches": int(launches), "clocks": clocks, "roofline": roof, "cpu_baseline": cpu,
            "speedup_e2e_vs_cpu_baseline": (e2e_val / cpu["value"]) if cpu and cpu.get("value") else None,
            "check": check}


def measure_soft(ctx, args, name, wl, steps, warmup, with_cpu=True):
    """Decoder::getSoftCodeword / getSoftInformation over a batch: N int8 LLRs -> N soft bytes + K soft info bytes."""
    import antpolarcodes_b200 as ap
    torch, device, world, rank = ctx.torch, ctx.device, ctx.world, ctx.rank
    N, K = wl["N"], wl["K"]
    frames = args.frames or wl["frames"]
    e2e_frames = min(args.e2e_frames or wl["e2e_frames"], frames)
    frozen = ap.frozen_bits(N, K, 0.0, wl["ctype"])
    dec = ap.PolarDecoder(N, 1, frozen, "char", device=ctx.local)
    dec.setErrorDetection(0)
    llr, _ = make_llrs_on_device(ctx, ap, wl, frozen, frames, 0x50F7 + rank)
    soft = torch.empty((frames, N), dtype=torch.int8, device=device)
    info = torch.empty((frames, K), dtype=torch.int8, device=device)
    h = dec._get()._h
    L_ = ap.lib()

    def step(l=llr, s_=soft, i_=info, mem=ap.MEM_DEVICE):
        rc = L_.polar_cuda_decode_soft_batch(h, l.data_ptr(), mem, l.shape[0], s_.data_ptr(), i_.data_ptr(), ctx.sp)
        assert rc == 0, L_.polar_cuda_last_error()
    for _ in range(warmup):
        step()
    ctx.barrier()
    sampler = ClockSampler(ctx.local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    l0 = dec._get().launches()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    ctx.barrier()
    ev[0].record(ctx.stream)
    for s_i in range(steps):
        step()
        ev[s_i + 1].record(ctx.stream)
    ctx.barrier()
    total_ms = ev[0].elapsed_time(ev[-1])
    k_ms = float(np.mean([ev[i].elapsed_time(ev[i + 1]) for i in range(steps)]))
    launches = dec._get().launches() - l0
    clocks = sampler.stop() if rank == 0 else None
    h_llr = torch.empty((e2e_frames, N), dtype=torch.int8).pin_memory()
    h_llr.copy_(llr[:e2e_frames])
    h_soft = torch.empty((e2e_frames, N), dtype=torch.int8).pin_memory()
    h_info = torch.empty((e2e_frames, K), dtype=torch.int8).pin_memory()
    step(h_llr, h_soft, h_info, ap.MEM_HOST)
    ctx.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ctx.stream)
    for _ in range(steps):
        step(h_llr, h_soft, h_info, ap.MEM_HOST)
    e1.record(ctx.stream)
    ctx.barrier()
    e2e_ms = e0.elapsed_time(e1)
    e2e_match = bool((h_soft == soft[:e2e_frames].cpu()).all().item())
    t = torch.tensor([total_ms, e2e_ms], dtype=torch.float64, device=device)
    if world > 1:
        ctx.dist.all_reduce(t, op=ctx.dist.ReduceOp.MAX)
    total_ms, e2e_ms = float(t[0].item()), float(t[1].item())
    if rank != 0:
        return None
    hbm_peak, _, which = peaks()
    value = world * frames * steps * K / (total_ms * 1e-3) / 1e9
    e2e_val = world * e2e_frames * steps * K / (e2e_ms * 1e-3) / 1e9
    bpf = 2 * N + K
    hbm = frames * bpf / (k_ms * 1e-3) / 1e9
    roof = {"bound": "hbm", "achieved": hbm, "peak": hbm_peak, "unit": "GB/s", "frac": hbm / hbm_peak, "traffic": None,
            "peak_source": which + " hbm copy", "algorithmic_bytes": bpf * frames, "kernel_ms": k_ms, "frames_per_launch": frames}
    check = {"frames": frames}
    cpu = None
    try:
        from polar_testlib import KIND_FASTSSC_I8, have_ref, ref_soft
        if have_ref():
            npar = min(args.parity if args.parity >= 0 else wl["parity"], frames)
            hl = np.ascontiguousarray(llr[:npar].cpu().numpy())
            t0 = time.perf_counter()
            rs, ri = ref_soft(KIND_FASTSSC_I8, N, frozen, hl)
            tt = time.perf_counter() - t0
            bad = int(((rs != soft[:npar].cpu().numpy()).any(axis=1) | (ri != info[:npar].cpu().numpy()).any(axis=1)).sum())
            check.update({"parity_frames": npar, "parity_mismatches": bad, "parity_against": "unmodified reference (oracle/_ref): getSoftCodeword + getSoftInformation"})
            if with_cpu:
                cpu = {"value": npar * K / tt / 1e9, "unit": "Gbit/s", "cores": 1, "kind": "reference",
                       "sample": "%d frames: FastSscFipChar decode_vector + getSoftCodeword + getSoftInformation, one thread, %.2f s" % (npar, tt)}
    except Exception as e:
        cpu = {"value": None, "unit": "Gbit/s", "cores": 0, "kind": "reference", "sample": "failed: %r" % (e,)}
    return {"metric": wl["label"], "value": value, "unit": "Gbit/s", "n_gpus": world, "steps": steps, "warmup": warmup,
            "ms_per_step": total_ms / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int8", "data": "synthetic",
            "config": {"workload": wl["label"], "frames_per_step_per_gpu": frames, "e2e_frames_per_step_per_gpu": e2e_frames,
                       "cache": "inputs larger than L2 (%d MiB in+out per step)" % (frames * bpf >> 20), "parallelism": "frames sharded x%d" % world},
            "e2e": {"value": e2e_val, "unit": "Gbit/s", "h2d_bytes_per_step": e2e_frames * N, "d2h_bytes_per_step": e2e_frames * (N + K),
                    "matches_device_run": e2e_match},
            "gpu_launches": int(launches), "clocks": clocks, "roofline": roof, "cpu_baseline": cpu,
            "speedup_e2e_vs_cpu_baseline": (e2e_val / cpu["value"]) if cpu and cpu.get("value") else None, "check": check}


def measure(ctx, args, name, steps, warmup, with_cpu=True):
    wl = WORKLOADS[name]
    fn = measure_encoder if wl.get("encoder") else measure_soft if wl.get("soft") else measure_decoder
    rec = fn(ctx, args, name, wl, steps, warmup, with_cpu)
    ctx.torch.cuda.empty_cache()
    return rec


def main():
    ap_ = argparse.ArgumentParser()
    ap_.add_argument("--gpus", type=int, default=1)
    ap_.add_argument("--steps", type=int, default=10)
    ap_.add_argument("--warmup", type=int, default=3)
    ap_.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap_.add_argument("--workload", default="scl8", choices=sorted(WORKLOADS))
    ap_.add_argument("--frames", type=int, default=0)
    ap_.add_argument("--e2e-frames", type=int, default=0, dest="e2e_frames")
    ap_.add_argument("--parity", type=int, default=-1, help="frames re-decoded by the reference (default per workload, 0 = off)")
    ap_.add_argument("--also", default="auto", help="'auto' (the other BASELINE configs on a 1-GPU headline run), 'none', or a comma list")
    ap_.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap_.parse_args()
    if args.impl == "reference":
        run_reference_arm(args, WORKLOADS[args.workload])
        return 0
    ctx = Ctx()
    line = measure(ctx, args, args.workload, args.steps, args.warmup, with_cpu=not args.no_cpu)
    if args.also == "auto":
        also = ALSO if (ctx.world == 1 and args.workload == "scl8" and not args.frames) else []
    elif args.also == "none":
        also = []
    else:
        also = [a for a in args.also.split(",") if a]
    recs = {}
    for name in also:
        try:
            recs[name] = measure(ctx, args, name, max(3, min(args.steps, 10)), 3, with_cpu=not args.no_cpu)
        except Exception as e:  # a failing side record must not lose the headline line -- but it is reported and fails the run
            recs[name] = {"error": repr(e)}
    rc = 0
    if ctx.rank == 0:
        if recs:
            line["also"] = recs
        bad = [n for n, r in [(args.workload, line)] + list(recs.items())
               if r.get("error") or (r.get("check", {}).get("parity_mismatches") or 0) != 0 or r.get("e2e", {}).get("matches_device_run") is False]
        line["parity_gate"] = "green" if not bad else "FAILED: " + ",".join(bad)
        print(json.dumps(line))
        rc = 3 if bad else 0
    ctx.close()
    return rc


if __name__ == "__main__":
    sys.exit(main())
