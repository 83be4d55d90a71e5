"""Stream partitioning over the GPUs of one box.

The path shards by independent IQ streams: every piece of state the reference keeps is per object
(libdsp/blkconv.h:51-61, decimate.h:48-59, resample.h:49-59), so rank r of W simply owns the contiguous
stream range [r*S/W, (r+1)*S/W) and runs its own sfe_chain on its own GPU.  There is no data-path
collective; torch.distributed is used only to agree on timings (max over ranks) and, when the caller wants
the whole result in one place, to gather the per-rank output rows.
"""
from __future__ import annotations


def shard_range(n_streams: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous, balanced: the first n_streams % world ranks carry one extra stream."""
    if world < 1 or not (0 <= rank < world) or n_streams < 0:
        raise ValueError("shard_range: need 0 <= rank < world and n_streams >= 0")
    base, extra = divmod(n_streams, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def max_over_ranks(value: float, device=None) -> float:
    """Timing rule for every multi-GPU number: the slowest rank defines the step."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def gather_rows(local_rows, n_streams: int):
    """All ranks' output rows (numpy, (s_local, n_out) complex64) in stream order on every rank.
    Convenience for callers that want one array; the hot path itself never needs it."""
    import numpy as np
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return local_rows
    world = dist.get_world_size()
    parts = [None] * world
    dist.all_gather_object(parts, (local_rows.shape, local_rows.tobytes()))
    rows = [np.frombuffer(b, dtype=np.complex64).reshape(shape) for shape, b in parts]
    n_out = {r.shape[1] for r in rows if r.shape[0] > 0}
    if len(n_out) > 1:
        raise RuntimeError(f"ranks disagree on the per-stream output count: {sorted(n_out)} (the schedule is data independent)")
    out = np.concatenate([r for r in rows if r.shape[0] > 0], axis=0) if n_out else np.zeros((0, 0), np.complex64)
    if out.shape[0] != n_streams:
        raise RuntimeError(f"gathered {out.shape[0]} streams, expected {n_streams}")
    return out
