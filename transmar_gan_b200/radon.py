"""Parallel-beam Radon projector for the physics-consistency loss, on radon.cu.

Interface mirror of the one class the reference uses from torch-radon:
    projector = Radon(PATCH_SIZE, angles)              # training/train_combined.py:269-272
    sino = projector.forward(x)                        # losses/gan_losses.py:241-246: [B, H, W] -> [B, n_angles, det]
with torch-radon's constructor arguments (`resolution, angles, det_count=-1, det_spacing=1.0`; det_count -1 =
resolution) and its autograd contract: the backward of `forward` is `backprojection` and vice versa.  torch-radon
is an empty submodule of the reference, so the discretisation is restated from its published algorithm and its
parity is unpinned; the tests pin the CUDA kernels to a float64 restatement and that restatement
to the analytic Radon transform of a disc.  CUDA only: there is no CPU path."""
from __future__ import annotations

from typing import Sequence, Union

import torch
from torch import Tensor

from . import ops


def _launch(name: str, src: Tensor, angles: Tensor, dst: Tensor, nimg: int, H: int, W: int, det: int,
            spacing: float) -> None:
    ops._chk(src, angles, dst)
    ops._call(name, ops._p(src), ops._p(angles), ops._p(dst), nimg, H, W, angles.numel(), det, float(spacing),
              ops._stream())


class _Forward(torch.autograd.Function):
    @staticmethod
    def forward(ctx_, x, angles, det, spacing):
        H, W = x.shape[-2:]
        x = x.contiguous()
        sino = x.new_empty(x.shape[:-2] + (angles.numel(), det))
        _launch("radon_fwd", x.detach(), angles, sino, x.numel() // (H * W), H, W, det, spacing)
        ctx_.save_for_backward(angles)
        ctx_.geom = (H, W, det, spacing)
        return sino

    @staticmethod
    def backward(ctx_, g):
        (angles,) = ctx_.saved_tensors
        H, W, det, spacing = ctx_.geom
        return _Backprojection.apply(g, angles, H, W, spacing), None, None, None


class _Backprojection(torch.autograd.Function):
    @staticmethod
    def forward(ctx_, sino, angles, H, W, spacing):
        sino = sino.contiguous()
        A, det = sino.shape[-2:]
        if A != angles.numel():
            raise RuntimeError(f"Radon.backprojection: sinogram has {A} angles, the projector {angles.numel()}")
        img = sino.new_empty(sino.shape[:-2] + (H, W))
        _launch("radon_bwd", sino.detach(), angles, img, sino.numel() // (A * det), H, W, det, spacing)
        ctx_.save_for_backward(angles)
        ctx_.geom = (det, spacing)
        return img

    @staticmethod
    def backward(ctx_, g):
        (angles,) = ctx_.saved_tensors
        det, spacing = ctx_.geom
        return _Forward.apply(g, angles, det, spacing), None, None, None, None


class Radon:
    """`Radon(resolution, angles, det_count=-1, det_spacing=1.0)`; `forward(x)` / `backprojection(sino)` (alias
    `backward`) on fp32 CUDA tensors [..., H, W] / [..., n_angles, det_count]."""

    def __init__(self, resolution: int, angles: Union[Sequence[float], Tensor], det_count: int = -1,
                 det_spacing: float = 1.0, clip_to_circle: bool = False):
        if clip_to_circle:
            raise NotImplementedError("clip_to_circle is never set by the reference (train_combined.py:271)")
        self.resolution = int(resolution)
        self.det_count = self.resolution if det_count <= 0 else int(det_count)
        self.det_spacing = float(det_spacing)
        self._angles_host = torch.as_tensor(angles, dtype=torch.float32).reshape(-1).clone()
        if not 1 <= self._angles_host.numel() <= 1024:
            raise ValueError("Radon: between 1 and 1024 angles")
        self._angles = {}

    def angles_on(self, device) -> Tensor:
        key = str(device)
        if key not in self._angles:
            self._angles[key] = self._angles_host.to(device)
        return self._angles[key]

    def forward(self, x: Tensor) -> Tensor:
        if x.dim() < 2 or x.shape[-1] != self.resolution or x.shape[-2] != self.resolution:
            raise RuntimeError(f"Radon.forward: expected [..., {self.resolution}, {self.resolution}], got {tuple(x.shape)}")
        return _Forward.apply(x, self.angles_on(x.device), self.det_count, self.det_spacing)

    def backprojection(self, sinogram: Tensor) -> Tensor:
        if sinogram.dim() < 2 or sinogram.shape[-1] != self.det_count:
            raise RuntimeError(f"Radon.backprojection: expected [..., n_angles, {self.det_count}]")
        return _Backprojection.apply(sinogram, self.angles_on(sinogram.device), self.resolution, self.resolution,
                                     self.det_spacing)

    backward = backprojection
