"""Data-parallel sharding of a head batch across ranks.

Every image is independent (its features, its ragged objects, its bin widths), so the batch dimension is split
contiguously across ranks, weights are replicated, and inference needs NO collective.  The helpers below are host logic
only (index arithmetic on the python lists the reference API uses); they run unchanged under ``gloo`` on CPU.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch


def shard_bounds(batch: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous [lo, hi) of rank ``rank``: the first ``batch % world`` ranks get one extra image."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    base, extra = divmod(batch, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_batch(features: torch.Tensor, object_features: Optional[Sequence[torch.Tensor]],
                object_xywh_list: Optional[Sequence[Optional[torch.Tensor]]], rank: int, world: int):
    """-> (features[lo:hi], object_features[lo:hi] as a NEW list, object_xywh_list[lo:hi]).  The object lists are split at
    image boundaries; the new list matters because the head overwrites the list it is given (reference quirk F11)."""
    lo, hi = shard_bounds(features.shape[0], rank, world)
    objs = None if object_features is None else list(object_features[lo:hi])
    boxes = None if object_xywh_list is None else list(object_xywh_list[lo:hi])
    return features[lo:hi], objs, boxes


def csr_offsets(counts: Sequence[int]) -> List[int]:
    out = [0]
    for c in counts:
        out.append(out[-1] + int(c))
    return out


def max_over_ranks_ms(ms: float, group=None) -> float:
    """Step time of the job = the slowest rank's device time (one scalar all-reduce, outside any timed region)."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return ms
    dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else torch.device("cpu")
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())


def gather_depth(depth: torch.Tensor, batch: int, group=None) -> Optional[torch.Tensor]:
    """Optional (not timed): collect the per-rank depth shards on rank 0 in batch order.  Shards may differ by one image."""
    import torch.distributed as dist
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    sizes = [shard_bounds(batch, r, world) for r in range(world)]
    biggest = max(hi - lo for lo, hi in sizes)
    pad = torch.zeros(biggest, *depth.shape[1:], dtype=depth.dtype, device=depth.device)
    pad[:depth.shape[0]] = depth
    bufs = [torch.empty_like(pad) for _ in range(world)] if rank == 0 else None
    dist.gather(pad, bufs, dst=0, group=group)
    if rank != 0:
        return None
    return torch.cat([b[:hi - lo] for b, (lo, hi) in zip(bufs, sizes)], 0)


def bind_host_to_gpu(local_rank: int) -> Optional[list]:
    """Pin this process to the CPU cores NVML reports as local to GPU ``local_rank`` (same NUMA node / PCIe root), intersected with
    the cores the process may use.  With one process per GPU the pinned host buffers of the host->device pipeline are then
    first-touched on the GPU's own NUMA node, so 8 ranks do not pull their 50 GB/s each through one socket's memory controllers
    and the inter-socket link.  Returns the core list, or None when nothing was changed (no NVML, empty intersection)."""
    import os
    try:
        import pynvml
        pynvml.nvmlInit()
        handle = pynvml.nvmlDeviceGetHandleByIndex(int(local_rank))
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(handle, words)
        cores = [64 * w + b for w, m in enumerate(mask) for b in range(64) if (int(m) >> b) & 1]
        allowed = os.sched_getaffinity(0)
        cores = sorted(set(cores) & set(allowed))
        if not cores or len(cores) == len(allowed):
            return None
        os.sched_setaffinity(0, cores)
        return cores
    except Exception:
        return None
