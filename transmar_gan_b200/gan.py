"""The parts of the TransMAR-GAN training step that stay on stock PyTorch (BASELINE.json north_star:
"The PatchGAN discriminator, the losses and torch-radon stay on stock PyTorch"), restated here only so that
the full step (BASELINE config 4) can be built and timed around the B200-native generator
on a machine that does not have the reference checkout.  Nothing in this file is a kernel of this repo.

Interface mirror (same names, arguments and state_dict keys as the reference):
  * `MultiScaleDiscriminator(in_channels=2, base_channels=64, num_layers=5, num_scales=3, use_sn=True)`
    - models/discriminator/ms_patchgan.py:9-152: per scale a stack of 4x4 convolutions (stride 2, the last
    one stride 1; widths base, 2b, 4b, 8b, 8b), LeakyReLU(0.2) after all but the last, a 1x1 convolution
    to one logit channel, spectral normalisation on every convolution; scale s sees the input average-
    pooled s times.  `forward(x, return_features)` -> (list of logits, list of lists of features | None).
  * loss functions of losses/gan_losses.py:6-219 (hinge, feature matching, metal-aware L1 / edge,
    metal consistency) and `physics_loss_syn` (:222-253), which takes the projector as an argument:
    torch-radon is an absent submodule of the reference, `transmar_gan_b200.radon.Radon` is
    this repo's projector with the same interface.
  * `weights_init` - training/train_combined.py:207-217.
Pinned against the reference by tests/test_gan_stock.py (golden fixtures generated from the reference files).
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F
from torch.nn.utils import spectral_norm


def weights_init(m: nn.Module) -> None:
    """N(0, 0.02) weights / zero biases for convolutions and linears, (1, 0) for normalisation layers."""
    if isinstance(m, (nn.Conv2d, nn.ConvTranspose2d, nn.Linear)):
        nn.init.normal_(m.weight, 0.0, 0.02)
        if m.bias is not None:
            nn.init.zeros_(m.bias)
    elif isinstance(m, (nn.BatchNorm2d, nn.InstanceNorm2d, nn.LayerNorm)):
        if getattr(m, "weight", None) is not None:
            nn.init.ones_(m.weight)
        if getattr(m, "bias", None) is not None:
            nn.init.zeros_(m.bias)


def _layer_widths(base: int, num_layers: int) -> List[int]:
    """Output width of each 4x4 convolution: doubling up to 8*base, frozen for the last two layers."""
    widths, w = [], base
    for i in range(num_layers):
        widths.append(w)
        if i < num_layers - 2:
            w = min(2 * w, 8 * base)
    return widths


class SingleScaleDiscriminator(nn.Module):
    def __init__(self, in_channels: int = 2, base_channels: int = 64, num_layers: int = 5,
                 use_sn: bool = True):
        super().__init__()
        self.num_layers = num_layers
        wrap = spectral_norm if use_sn else (lambda c: c)
        widths = _layer_widths(base_channels, num_layers)
        ins = [in_channels] + widths[:-1]
        self.convs = nn.ModuleList(
            wrap(nn.Conv2d(ci, co, kernel_size=4, stride=1 if i == num_layers - 1 else 2, padding=1))
            for i, (ci, co) in enumerate(zip(ins, widths)))
        self.final_conv = wrap(nn.Conv2d(widths[-1], 1, kernel_size=1))

    def forward(self, x, return_features: bool = True):
        feats, h = [], x
        for i, conv in enumerate(self.convs):
            h = conv(h)
            if i < self.num_layers - 1:
                h = F.leaky_relu(h, 0.2)
                if return_features:
                    feats.append(h)
        return self.final_conv(h), (feats if return_features else None)


class MultiScaleDiscriminator(nn.Module):
    def __init__(self, in_channels: int = 2, base_channels: int = 64, num_layers: int = 5,
                 num_scales: int = 3, use_sn: bool = True):
        super().__init__()
        self.num_scales = num_scales
        self.discriminators = nn.ModuleList(
            SingleScaleDiscriminator(in_channels, base_channels, num_layers, use_sn)
            for _ in range(num_scales))

    def forward(self, x, return_features: bool = True):
        logits, feats = [], []
        for d in self.discriminators:
            lg, ft = d(x, return_features=return_features)
            logits.append(lg)
            feats.append(ft)
            x = F.avg_pool2d(x, kernel_size=2, stride=2)
        return logits, (feats if return_features else None)


# ------------------------------------------------------------------------------ losses
def hinge_d_loss(real_logits: Sequence[torch.Tensor], fake_logits: Sequence[torch.Tensor]):
    return sum(F.relu(1.0 - r).mean() for r in real_logits) + sum(F.relu(1.0 + f).mean() for f in fake_logits)


def hinge_g_loss(fake_logits: Sequence[torch.Tensor]):
    return sum(-f.mean() for f in fake_logits)


def feature_matching_loss(real_features, fake_features):
    return sum((fr - ff).abs().mean()
               for rs, fs in zip(real_features, fake_features) for fr, ff in zip(rs, fs))


def extract_metal_mask(ct: torch.Tensor, threshold: float = 0.6) -> torch.Tensor:
    return (ct > threshold).float()


def dilate_mask(mask: torch.Tensor, radius: int = 5) -> torch.Tensor:
    return F.max_pool2d(mask, kernel_size=2 * radius + 1, stride=1, padding=radius)


def compute_weight_map(ct, beta: float = 1.0, radius: int = 5, w_max: float = 3.0, threshold: float = 0.6):
    band = dilate_mask(extract_metal_mask(ct, threshold), radius)
    return torch.clamp(1.0 + beta * band, max=w_max)


def compute_metal_aware_loss(fake, real_cpu, ct, beta: float = 1.0, radius: int = 5, w_max: float = 3.0,
                             threshold: float = 0.6):
    return (compute_weight_map(ct, beta, radius, w_max, threshold) * (fake - real_cpu)).abs().mean()


def compute_image_gradients(x) -> Tuple[torch.Tensor, torch.Tensor]:
    """Forward differences, last column / row replicated back to the input size."""
    gx = F.pad(x[..., :, 1:] - x[..., :, :-1], (0, 1, 0, 0), mode="replicate")
    gy = F.pad(x[..., 1:, :] - x[..., :-1, :], (0, 0, 0, 1), mode="replicate")
    return gx, gy


def compute_metal_aware_edge_loss(fake, real_cpu, w):
    fx, fy = compute_image_gradients(fake)
    rx, ry = compute_image_gradients(real_cpu)
    return (w * ((fx - rx).abs() + (fy - ry).abs())).mean()


def metal_consistency_loss(fake, real, M):
    return (M * (fake - real)).abs().mean()


def physics_loss_syn(fake, real_cpu, M, projector):
    """mean((1 - [P(M) > 0]) * |P(fake) - P(real)|): rays that cross metal do not count."""
    sq = lambda t: t.squeeze(1) if t.shape[1] == 1 else t
    proj_fake, proj_clean = projector.forward(sq(fake)), projector.forward(sq(real_cpu))
    Mp = (projector.forward(sq(M)) > 0).float()
    return ((1.0 - Mp) * (proj_fake - proj_clean).abs()).mean()
