"""Batch sharding of independent units (clips / sequences) over the ranks of one box -- SURVEY.md section 8e.

Every clip and every sequence is independent in every stage of the path (the reference itself loops over the batch
serially: layers/conv_1d.c:174-181, gru.c:259-283, lstm.c:438-465), so the multi-GPU plan has NO collective on the data
path: rank g of G owns the contiguous unit range `unit_range(n, g, G)` of the [B, ...] tensors, weights are replicated by
a plain per-device upload, and the only cross-rank traffic is (a) the optional final gather of the outputs to rank 0 and
(b) the max-over-ranks of the device timing.  Both go through `torch.distributed` (NCCL on the GPUs, gloo in the CPU
tests); results are bit-identical for any G because per-unit arithmetic never depends on the partition.
"""
from __future__ import annotations

from typing import Callable

import numpy as np


def unit_range(n_units: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous slice [first, last) of `n_units` owned by `rank`; the first `n_units % world` ranks get one extra."""
    if world < 1 or not 0 <= rank < world or n_units < 0:
        raise ValueError(f"bad partition request: n_units={n_units} rank={rank} world={world}")
    base, extra = divmod(n_units, world)
    first = rank * base + min(rank, extra)
    return first, first + base + (1 if rank < extra else 0)


def shard_apply(apply_batch: Callable[[np.ndarray], np.ndarray], units: np.ndarray, rank: int, world: int) -> np.ndarray:
    """Runs `apply_batch` (e.g. `Spectrogram.apply_batch`) on this rank's slice of `units` ([B, ...]) only."""
    first, last = unit_range(units.shape[0], rank, world)
    return apply_batch(units[first:last])


def gather_to_root(local_out, n_units: int, dist=None, root: int = 0):
    """Final host gather: rank `root` returns the [n_units, ...] concatenation in unit order, the others None.

    `local_out` is this rank's [n_local, ...] torch tensor (CPU for gloo, CUDA for NCCL).  Slices are ragged when
    `n_units % world != 0`, so every rank pads to the largest slice before `gather` and the root trims.
    """
    import torch
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return local_out
    rank, world = dist.get_rank(), dist.get_world_size()
    sizes = [b - a for a, b in (unit_range(n_units, r, world) for r in range(world))]
    assert local_out.shape[0] == sizes[rank], f"rank {rank}: {local_out.shape[0]} units, expected {sizes[rank]}"
    pad = max(sizes)
    buf = local_out.new_zeros((pad,) + tuple(local_out.shape[1:]))
    buf[: sizes[rank]] = local_out
    bucket = [torch.empty_like(buf) for _ in range(world)] if rank == root else None
    dist.gather(buf, bucket, dst=root)
    if rank != root:
        return None
    return torch.cat([b[:s] for b, s in zip(bucket, sizes)], dim=0)


def max_over_ranks(value: float, dist=None, device=None) -> float:
    """Device time of the slowest rank -- what bench.py divides the whole-job unit count by."""
    import torch
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
