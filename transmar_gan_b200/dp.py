"""Data-parallel training support for the drop-in generator: one process per GPU, batch
sharded over ranks, ONE flat fp32 gradient bucket all-reduced per step.

The reference has no distributed code (only `nn.DataParallel` behind `ngpu = 1`,
training/train_combined.py:95,234-235), so this is new: the generator's 504 gradient tensors
(990 811 floats = 3.96 MB) live as views into a single contiguous buffer, so the whole
exchange is one `all_reduce` (NCCL over NVLink on the GPU box, gloo in the CPU tests) —
latency-bound, no per-tensor launches.  NGswin has no BatchNorm and windows never cross
samples, so batch-sharded replicas are exact.
"""
from __future__ import annotations

from typing import Iterable, List

import torch
import torch.distributed as dist


class FlatGradBucket:
    """Owns one flat gradient buffer; every parameter's `.grad` is a view into it.  Views start on multiples of
    `align` elements (4 fp32 = 16 bytes: the backward kernels then flush weight gradients with 16-byte vector
    reductions); the padding stays zero and rides along in the all-reduce.  `numel` is the buffer's length
    (what is exchanged), `param_numel` the number of gradient elements, `offsets[i]` the start of parameter i."""

    def __init__(self, params: Iterable[torch.nn.Parameter], align: int = 4):
        self.params: List[torch.nn.Parameter] = [p for p in params if p.requires_grad]
        if not self.params:
            raise ValueError("FlatGradBucket needs at least one trainable parameter")
        dev, dt = self.params[0].device, self.params[0].dtype
        if any(p.device != dev or p.dtype != dt for p in self.params):
            raise ValueError("all parameters of a bucket must share device and dtype")
        align = max(int(align), 1)
        self.param_numel = sum(p.numel() for p in self.params)
        self.offsets, off = [], 0
        for p in self.params:
            self.offsets.append(off)
            off += -(-p.numel() // align) * align
        self.numel = off
        self.flat = torch.zeros(self.numel, device=dev, dtype=dt)
        for p, o in zip(self.params, self.offsets):
            p.grad = self.flat[o:o + p.numel()].view_as(p)

    def zero(self):
        """Zero all gradients with one memset (keeps the views attached)."""
        self.flat.zero_()
        for p in self.params:       # optimizer.zero_grad(set_to_none=True) would detach the views
            if p.grad is None:
                raise RuntimeError("a parameter's .grad was detached from the flat bucket; "
                                   "use bucket.zero() instead of zero_grad(set_to_none=True)")

    def all_reduce(self, group=None, async_op: bool = False):
        """Average the bucket over the ranks of `group` (no-op for a single process)."""
        if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
            return None
        world = dist.get_world_size(group)
        if dist.get_backend(group) == "nccl":
            return dist.all_reduce(self.flat, op=dist.ReduceOp.AVG, group=group, async_op=async_op)
        work = dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group, async_op=False)
        self.flat.div_(world)
        return work


def broadcast_parameters(module: torch.nn.Module, src: int = 0, group=None):
    """Make every rank start from rank `src`'s weights and buffers."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return
    for t in list(module.parameters()) + list(module.buffers()):
        dist.broadcast(t.data, src=src, group=group)


def shard_batch(x: torch.Tensor, rank: int, world: int) -> torch.Tensor:
    """Contiguous batch shard of rank `rank` (global batch must divide evenly)."""
    if x.shape[0] % world:
        raise ValueError(f"global batch {x.shape[0]} is not divisible by world size {world}")
    per = x.shape[0] // world
    return x[rank * per:(rank + 1) * per]
