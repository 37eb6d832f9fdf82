"""Multi-GPU = replicas only: one frame does not shard (integral images and keyline ids are global
serial orders, the tracker reduces over all keylines every LM iteration), so scaling comes from partitioning independent
sequences over ranks, one process per GPU, with NO data-path collective. This module holds the host logic of that
partition and the only cross-rank operations the benchmark needs (a barrier and max/sum of scalars for timing), written
against torch.distributed so the same code runs over NCCL on the GPU box and over gloo in the CPU tests."""
from __future__ import annotations

LANES_PER_GPU = 64
SEED_BASE = 1000


def sequence_seeds(rank: int, lanes: int = LANES_PER_GPU):
    """seeds of the sequences rank `rank` replays (C4 of SURVEY.md 8(d): 1000 + 64 g + i)"""
    return [SEED_BASE + LANES_PER_GPU * rank + i for i in range(lanes)]


def owner_of(sequence_index: int, world: int, lanes: int = LANES_PER_GPU):
    """(rank, lane) that replays global sequence `sequence_index` when `world` ranks hold `lanes` lanes each"""
    if not 0 <= sequence_index < world * lanes:
        raise ValueError("sequence index out of range")
    return sequence_index // lanes, sequence_index % lanes


def partition(n_sequences: int, world: int):
    """contiguous block partition of n_sequences over ranks (sizes differ by at most one) -> list of ranges"""
    base, extra = divmod(n_sequences, world)
    out, start = [], 0
    for r in range(world):
        n = base + (1 if r < extra else 0)
        out.append(range(start, start + n))
        start += n
    return out


def reduce_scalar(x: float, op: str, device=None) -> float:
    """max / sum of a scalar over ranks (timing plumbing only). No-op without an initialised process group."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(x)
    t = torch.tensor([x], dtype=torch.float64, device=device if device is not None else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX if op == "max" else dist.ReduceOp.SUM)
    return float(t.item())


def whole_job_rate(units_this_rank: float, seconds_this_rank: float, device=None) -> float:
    """units all ranks processed / max over ranks of the elapsed time (bench.py's `value`)"""
    total = reduce_scalar(units_this_rank, "sum", device)
    tmax = reduce_scalar(seconds_this_rank, "max", device)
    return total / tmax
