"""Clip-sharded multi-GPU driver of the hot path: one process per GPU, no collective in the hot
loop, one gather of the per-file summary rows to rank 0.

The reference's only parallel idiom is a file-level multiprocessing.Pool
(/root/reference/04_feature_extraction_experiment/feature_extraction_for_student.py:168-171); batch_process itself is
a sequential loop (audio_feature_extraction_toolkit/core/feature_extractor.py:228-231).  Here rank r of W takes the
contiguous clip range [floor(r N / W), floor((r+1) N / W)) of the enumerated list, so the concatenation of the
gathered rows on rank 0 is in enumeration order, exactly what the sequential loop would return.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Sequence, Tuple

import numpy as np

__all__ = ["shard_range", "gather_rows", "extract_sharded"]


def shard_range(n: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous block partition of n clips over `world` ranks."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad rank {rank} / world {world}")
    return (rank * n) // world, ((rank + 1) * n) // world


def gather_rows(local: "torch.Tensor", dst: int = 0, group=None) -> Optional["torch.Tensor"]:
    """Gather 2-D row blocks of different heights to `dst` (NCCL for CUDA tensors, gloo for CPU tensors).
    Returns the row-wise concatenation in rank order on `dst`, None elsewhere."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    if world == 1:
        return local
    counts = [torch.zeros(1, dtype=torch.int64, device=local.device) for _ in range(world)]
    dist.all_gather(counts, torch.tensor([local.shape[0]], dtype=torch.int64, device=local.device), group=group)
    counts = [int(c.item()) for c in counts]
    width = local.shape[1]
    cap = max(max(counts), 1)
    padded = torch.zeros((cap, width), dtype=local.dtype, device=local.device)
    padded[: local.shape[0]] = local
    bufs = [torch.empty_like(padded) for _ in range(world)] if rank == dst else None
    dist.gather(padded, bufs, dst=dst, group=group)
    if rank != dst:
        return None
    return torch.cat([b[:c] for b, c in zip(bufs, counts)], dim=0)


def extract_sharded(extract_rows: Callable[[Sequence], Tuple["torch.Tensor", "torch.Tensor"]], items: Sequence,
                    dst: int = 0, group=None):
    """Run `extract_rows(shard) -> (summaries [m, SS] f32, meta [m, 4] i32)` on this rank's shard of `items` and gather
    both row blocks to `dst`.  Returns (summaries, meta) for all items on `dst`, (None, None) elsewhere."""
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized():
        rank, world = dist.get_rank(group), dist.get_world_size(group)
    else:
        rank, world = 0, 1
    lo, hi = shard_range(len(items), rank, world)
    summ, meta = extract_rows(items[lo:hi])
    if world == 1:
        return summ, meta
    return gather_rows(summ, dst, group), gather_rows(meta, dst, group)
