"""The generator's training step as one replayable unit.

`L1TrainStep` is the baseline-variant recipe of the reference (`y = netG(x); F.l1_loss(y, t).backward();
optG.step()`, scripts/run_ablation_studies.py:1525-1535, Adam betas (.5, .999), lr 1e-4) on the drop-in
NGswin, with the data-parallel plumbing of `dp.py`:

    zero the flat gradient bucket -> forward -> L1 -> backward -> [all-reduce the bucket] -> Adam

A step at the reference's training crop (batch 16 x 128x128) is ~800 small kernel launches, so the host's
launch rate, not the GPU, bounds the eager step.  With `graph=True` the step is captured once into CUDA
graphs (inputs are copied into static buffers) and every later call is a replay: part A = zero + forward
+ loss + backward, part B = Adam.  On one process A and B are a single graph; with several ranks the
NCCL all-reduce of the flat bucket runs between the two graphs on the same stream.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.distributed as dist
import torch.nn.functional as F

from . import dp


class L1TrainStep:
    """Callable `step(x, target) -> loss` (a 0-d CUDA tensor; reading it synchronises)."""

    def __init__(self, model: torch.nn.Module, lr: float = 1e-4, betas: Tuple[float, float] = (0.5, 0.999),
                 graph: bool = True, group=None, warmup: int = 3):
        self.model, self.group, self.use_graph = model, group, graph
        self.bucket = dp.FlatGradBucket(model.parameters())
        # capturable: the step counter lives on the device, so the update can be replayed from a graph
        self.opt = torch.optim.Adam(model.parameters(), lr=lr, betas=betas, capturable=True)
        self.world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
        self._warmup = max(int(warmup), 1)
        self._x = self._t = self._loss = None
        self._gA: Optional[torch.cuda.CUDAGraph] = None
        self._gB: Optional[torch.cuda.CUDAGraph] = None

    # -- the recipe -----------------------------------------------------------------------
    def _fwd_bwd(self, x, t):
        self.bucket.zero()
        loss = F.l1_loss(self.model(x), t)
        loss.backward()
        return loss.detach()

    def _eager(self, x, t):
        loss = self._fwd_bwd(x, t)
        self.bucket.all_reduce(self.group)
        self.opt.step()
        return loss

    # -- capture --------------------------------------------------------------------------
    def _capture(self, x, t):
        if not x.is_cuda:
            raise RuntimeError("L1TrainStep: inputs must be CUDA tensors (there is no CPU path)")
        self._x, self._t = x.clone(), t.clone()
        # the warm-up steps must not count: parameters and optimizer state are restored in place afterwards
        params = [p for g in self.opt.param_groups for p in g["params"]]
        snap_p = [p.detach().clone() for p in params]
        snap_s = {p: {k: v.clone() for k, v in self.opt.state[p].items() if torch.is_tensor(v)}
                  for p in params if p in self.opt.state and self.opt.state[p]}
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):                 # warm-up: lazy kernel attributes, Adam state, cuDNN plans
            for _ in range(self._warmup):
                self._eager(self._x, self._t)
        torch.cuda.current_stream().wait_stream(side)
        with torch.no_grad():
            for p, q in zip(params, snap_p):
                p.copy_(q)
                for k, v in self.opt.state[p].items():
                    if torch.is_tensor(v):
                        v.copy_(snap_s[p][k]) if p in snap_s else v.zero_()
        torch.cuda.synchronize()
        self._gA = torch.cuda.CUDAGraph()
        if self.world == 1:
            with torch.cuda.graph(self._gA):
                self._loss = self._fwd_bwd(self._x, self._t)
                self.opt.step()
        else:
            with torch.cuda.graph(self._gA):
                self._loss = self._fwd_bwd(self._x, self._t)
            self._gB = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self._gB, pool=self._gA.pool()):
                self.opt.step()

    def __call__(self, x: torch.Tensor, t: torch.Tensor) -> torch.Tensor:
        if not self.use_graph:
            return self._eager(x, t)
        if self._gA is None:
            self._capture(x, t)
        if x.shape != self._x.shape or t.shape != self._t.shape:
            raise RuntimeError(f"L1TrainStep was captured for input {tuple(self._x.shape)}; got {tuple(x.shape)} "
                               "(build another step object for another shape)")
        if x.data_ptr() != self._x.data_ptr():
            self._x.copy_(x, non_blocking=True)
        if t.data_ptr() != self._t.data_ptr():
            self._t.copy_(t, non_blocking=True)
        self._gA.replay()
        if self._gB is not None:
            self.bucket.all_reduce(self.group)
            self._gB.replay()
        return self._loss
