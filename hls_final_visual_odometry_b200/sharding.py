"""Multi-GPU host logic: the hot path shards by INDEPENDENT sequences, one process per GPU, no
data-path collective.  The only cross-rank traffic is the timing
reduction (max over ranks) and an optional gather of per-rank counters -- done with
torch.distributed (NCCL on GPUs, gloo in the CPU tests)."""
from __future__ import annotations


def assign_sequences(n_sequences: int, world: int, rank: int) -> list[int]:
    """Round-robin ownership: sequence s belongs to rank s % world.  Every sequence has exactly one
    owner; ranks differ by at most one sequence."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad rank {rank} of {world}")
    return [s for s in range(n_sequences) if s % world == rank]


def split_frame_ranges(n_frames: int, world: int, rank: int) -> tuple[int, int]:
    """Second way to shard: ONE sequence split into contiguous frame ranges.  Matching pair p
    needs only frames (p-1, p), so rank r takes the pairs [a, b) of a balanced partition of 1..n_frames-1 and extracts
    the frames [a-1, b): the boundary frame a-1 is extracted twice (by this rank and by the one before it), nothing is
    exchanged.  Returns (first_frame, end_frame) with end exclusive; its pairs are first_frame+1 .. end_frame-1.  A rank
    without pairs gets an empty range (first == end)."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad rank {rank} of {world}")
    n_pairs = max(n_frames - 1, 0)
    a = 1 + n_pairs * rank // world
    b = 1 + n_pairs * (rank + 1) // world
    return (a - 1, b) if b > a else (a - 1, a - 1)


def sequence_seed(sequence: int, base: int = 1) -> int:
    """Synthetic sequence s uses texture seed base + s (seeds 1..8 for the 8 sharded sequences)."""
    return base + sequence


def reduce_max(values, dist=None, device=None):
    """Element-wise max of a list of floats over all ranks (identity when not distributed)."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return list(values)
    import torch
    t = torch.tensor(list(values), dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [float(x) for x in t]


def reduce_sum(values, dist=None, device=None):
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return list(values)
    import torch
    t = torch.tensor(list(values), dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return [float(x) for x in t]
