"""Head-gradient synchronisation for the data-parallel training step.

The reference trains under Lightning ``DDPStrategy(find_unused_parameters=True)`` (``main.py:66,104,129``): bucketed
all-reduces over every trainable parameter plus unused-parameter discovery, because in the single-SACA head several
modules never receive a gradient.  Here the head's gradients live in ONE flat, statically laid out fp32
buffer (32-46 MB): every parameter's ``.grad`` is a view into it, parameters that received no gradient simply leave their
slice at zero, and a step needs exactly one ``all_reduce`` (NCCL over NVLink on the GPUs; gloo in the CPU tests) followed by
the 1/world scaling — no bucketing, no unused-parameter search, no pack/unpack copies.
"""
from __future__ import annotations

from typing import Iterable, List, Optional

import torch
import torch.distributed as dist


class FlatGradBuffer:
    def __init__(self, params: Iterable[torch.nn.Parameter], align: int = 64):
        self.params: List[torch.nn.Parameter] = [p for p in params if p.requires_grad]
        if not self.params:
            raise ValueError("no trainable parameters")
        dev, dt = self.params[0].device, self.params[0].dtype
        self.offsets = []
        off = 0
        for p in self.params:
            if p.device != dev or p.dtype != dt:
                raise ValueError("all head parameters must share device and dtype")
            self.offsets.append(off)
            off += (p.numel() + align - 1) // align * align      # slices start on 256-byte boundaries (fp32)
        self.flat = torch.zeros(off, device=dev, dtype=dt)
        self.attach()

    def attach(self):
        """Make every ``.grad`` a view of the flat buffer (kernels / autograd then accumulate straight into it)."""
        for p, off in zip(self.params, self.offsets):
            p.grad = self.flat[off:off + p.numel()].view_as(p)

    def zero(self):
        self.flat.zero_()

    def numel(self) -> int:
        return sum(p.numel() for p in self.params)

    def all_reduce_mean(self, group=None, async_op: bool = False):
        """One collective for the whole head; returns the work handle when ``async_op`` (overlap with the loss tail)."""
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
            return None
        world = dist.get_world_size(group)
        work = dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group, async_op=async_op)
        if async_op:
            return _Scaled(work, self.flat, 1.0 / world)
        self.flat.mul_(1.0 / world)
        return None

    def clip_grad_norm_(self, max_norm: float) -> torch.Tensor:
        """``gradient_clip_val`` of the reference trainer (``main.py:122-137``; 0.1 in ``params/basicParams.yaml``): the
        norm of the flat buffer IS the global gradient norm (padding is zero)."""
        total = torch.linalg.vector_norm(self.flat)
        coef = torch.clamp(max_norm / (total + 1e-6), max=1.0)
        self.flat.mul_(coef)
        return total


class _Scaled:
    def __init__(self, work, flat, scale):
        self.work, self.flat, self.scale = work, flat, scale

    def wait(self):
        self.work.wait()
        self.flat.mul_(self.scale)
