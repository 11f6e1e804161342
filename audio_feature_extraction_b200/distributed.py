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

__all__ = ["shard_range", "gather_rows", "extract_sharded", "bind_host_to_gpu_node", "parse_cpulist"]


def shard_range(n: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous block partition of n clips over `world` ranks."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad rank {rank} / world {world}")
    return (rank * n) // world, ((rank + 1) * n) // world


def gather_rows(local: "torch.Tensor", dst: int = 0, group=None, counts: Optional[Sequence[int]] = None) -> Optional["torch.Tensor"]:
    """Gather 2-D row blocks of different heights to `dst` (NCCL for CUDA tensors, gloo for CPU tensors).
    Returns the row-wise concatenation in rank order on `dst`, None elsewhere.  `counts` (rows per rank) skips the
    exchange of the heights -- and with it the host synchronisation -- when every rank already knows them."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    if world == 1:
        return local
    if counts is None:
        cs = [torch.zeros(1, dtype=torch.int64, device=local.device) for _ in range(world)]
        dist.all_gather(cs, torch.tensor([local.shape[0]], dtype=torch.int64, device=local.device), group=group)
        counts = [int(c.item()) for c in cs]
    counts = [int(c) for c in counts]
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
    # block partition: every rank can compute every shard's height, no exchange of counts (and no host sync) needed
    counts = [shard_range(len(items), r, world)[1] - shard_range(len(items), r, world)[0] for r in range(world)]
    return gather_rows(summ, dst, group, counts), gather_rows(meta, dst, group, counts)


def parse_cpulist(text: str) -> List[int]:
    """'0-3,8,10-11' (the sysfs cpulist format) -> [0, 1, 2, 3, 8, 10, 11]"""
    out: List[int] = []
    for part in text.strip().split(","):
        part = part.strip()
        if not part:
            continue
        if "-" in part:
            a, b = part.split("-", 1)
            out.extend(range(int(a), int(b) + 1))
        else:
            out.append(int(part))
    return out


def bind_host_to_gpu_node(device_index: int, sysfs: str = "/sys") -> dict:
    """Place this rank's host side next to its GPU: with one process per GPU on a multi-socket box, the pinned staging
    buffers of the host path (afx_extract_host_packed) must live on the NUMA node the GPU's PCIe root hangs off, or half
    the ranks pull their waves across the socket interconnect.  Sets the CPU affinity to that node's CPUs (when the
    process is allowed any of them) and the memory policy to MPOL_PREFERRED(node) so that pinned allocations made
    afterwards are node-local.  Call it BEFORE allocating pinned memory.  Returns what it found and did (never raises:
    a box without NUMA information is left as it is)."""
    import ctypes
    import os
    info = {"device": device_index, "node": None, "cpus_bound": None, "mempolicy": False}
    try:
        import torch
        pr = torch.cuda.get_device_properties(device_index)
        bdf = "%04x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        info["pci"] = bdf
        with open(f"{sysfs}/bus/pci/devices/{bdf}/numa_node") as f:
            node = int(f.read().strip())
        info["node"] = node
        if node < 0:
            return info
        with open(f"{sysfs}/devices/system/node/node{node}/cpulist") as f:
            cpus = set(parse_cpulist(f.read()))
        allowed = os.sched_getaffinity(0)
        mine = sorted(cpus & allowed)
        if mine:
            os.sched_setaffinity(0, mine)
            info["cpus_bound"] = len(mine)
        libc = ctypes.CDLL(None, use_errno=True)
        maxnode = 1024
        mask = (ctypes.c_ulong * (maxnode // (8 * ctypes.sizeof(ctypes.c_ulong))))()
        mask[node // (8 * ctypes.sizeof(ctypes.c_ulong))] = 1 << (node % (8 * ctypes.sizeof(ctypes.c_ulong)))
        MPOL_PREFERRED, SYS_set_mempolicy = 1, 238   # x86_64
        if os.uname().machine == "x86_64":
            rc = libc.syscall(SYS_set_mempolicy, MPOL_PREFERRED, ctypes.byref(mask), maxnode)
            info["mempolicy"] = rc == 0
    except Exception as e:   # no sysfs entry, no NUMA, restricted container: stay as we are
        info["error"] = repr(e)
    return info
