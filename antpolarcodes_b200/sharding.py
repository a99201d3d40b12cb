"""Frame sharding and counter reduction for multi-GPU runs (one process per GPU).

Frames are independent (the reference decodes one frame per call with no cross-frame state,
src/polarcode/decoding/decoder.cpp:154-181), so ranks take contiguous frame ranges and never exchange data
on the decode path.  The only collective is one small SUM all-reduce of the error / frame counters per
BLER point (pcsim keeps the same counters per job: src/simulation/simulator.cpp:967-985)."""
import torch
import torch.distributed as dist


def frame_range(n_frames, rank, world):
    """Contiguous slice [lo, hi) of rank `rank` out of `world`; sizes differ by at most one frame."""
    lo = (n_frames * rank) // world
    hi = (n_frames * (rank + 1)) // world
    return lo, hi


def reduce_counters(counters, device=None):
    """SUM all-reduce of a dict of integer counters {frames, block_errors, bit_errors, detected_errors...}.
    NCCL over NVLink when the process group is NCCL (device must then be the rank's GPU), gloo on CPU."""
    keys = sorted(counters)
    t = torch.tensor([int(counters[k]) for k in keys], dtype=torch.int64, device=device or "cpu")
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return {k: int(v) for k, v in zip(keys, t.tolist())}


def max_over_ranks(value, device=None):
    t = torch.tensor([float(value)], dtype=torch.float64, device=device or "cpu")
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
