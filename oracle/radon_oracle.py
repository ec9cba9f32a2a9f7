"""TEST INFRASTRUCTURE ONLY (tests/, smoke(), bench cpu_baseline) - never imported by the product path.

CPU restatement (numpy, float64) of the parallel-beam Radon projector behind the reference's physics-consistency
loss: `projector = Radon(PATCH_SIZE, angles)` (training/train_combined.py:269-272), `projector.forward([B,H,W])
-> [B, n_angles, det]` (losses/gan_losses.py:222-253).

PARITY UNPINNED: torch-radon lives in `external/torch-radon`, an EMPTY submodule of the reference checkout
(.SUBMODULES.json), is not installed here and the reference holds no golden sinogram.  The algorithm below is
torch-radon's published one (ray-driven forward with <= 1-pixel steps and bilinear interpolation, pixel-driven
backprojection with linear detector interpolation as the autograd backward); sub-pixel conventions that the absent
source would fix (midpoint sampling, step count) are this restatement's choice and are stated in
transmar_gan_b200/csrc/radon.cu.  What IS pinned, independently of torch-radon, is that this is a Radon transform:
tests/test_radon_oracle.py checks it against the analytic projection of a disc, mass conservation per angle,
linearity and approximate adjointness of the two operators."""
from __future__ import annotations

import numpy as np


def _bilinear0(img: np.ndarray, u: np.ndarray, v: np.ndarray) -> np.ndarray:
    """Texture-style sampling: pixel (i, j) centre at (u, v) = (j + .5, i + .5); zero outside."""
    H, W = img.shape
    xb, yb = u - 0.5, v - 0.5
    x0, y0 = np.floor(xb).astype(np.int64), np.floor(yb).astype(np.int64)
    fx, fy = xb - x0, yb - y0

    def at(y, x):
        ok = (y >= 0) & (y < H) & (x >= 0) & (x < W)
        return np.where(ok, img[np.clip(y, 0, H - 1), np.clip(x, 0, W - 1)], 0.0)

    top = at(y0, x0) + fx * (at(y0, x0 + 1) - at(y0, x0))
    bot = at(y0 + 1, x0) + fx * (at(y0 + 1, x0 + 1) - at(y0 + 1, x0))
    return top + fy * (bot - top)


def radon_forward(img: np.ndarray, angles: np.ndarray, det_count: int = -1, det_spacing: float = 1.0) -> np.ndarray:
    """img [H, W] -> sinogram [n_angles, det_count] (see radon.cu for the discretisation)."""
    img = np.asarray(img, dtype=np.float64)
    H, W = img.shape
    D = max(H, W) if det_count <= 0 else det_count
    hx, hy, T = 0.5 * W, 0.5 * H, 0.71 * max(H, W)
    out = np.zeros((len(angles), D))
    k = np.arange(D)
    s = (k - 0.5 * D + 0.5) * det_spacing
    for a, th in enumerate(np.asarray(angles, dtype=np.float64)):
        cs, sn = np.cos(th), np.sin(th)
        sx, sy = s * cs - T * sn, s * sn + T * cs
        dx, dy = 2 * T * sn, -2 * T * cs
        lo, hi = np.zeros(D), np.ones(D)
        for p0, d, h in ((sx, dx, hx), (sy, dy, hy)):
            if abs(d) > 1e-12:
                t0, t1 = (-h - p0) / d, (h - p0) / d
                lo, hi = np.maximum(lo, np.minimum(t0, t1)), np.minimum(hi, np.maximum(t0, t1))
            else:
                hi = np.where((p0 < -h) | (p0 > h), -1.0, hi)
        for r in np.nonzero(hi > lo)[0]:
            length = 2 * T * (hi[r] - lo[r])
            n = max(int(np.ceil(length - 1e-3)), 1)
            j = np.arange(n) + 0.5
            px = sx[r] + dx * lo[r] + hx + j * (dx * (hi[r] - lo[r]) / n)
            py = sy[r] + dy * lo[r] + hy + j * (dy * (hi[r] - lo[r]) / n)
            out[a, r] = _bilinear0(img, px, py).sum() * (length / n)
    return out


def radon_backprojection(sino: np.ndarray, angles: np.ndarray, H: int, W: int, det_spacing: float = 1.0) -> np.ndarray:
    """sinogram [n_angles, D] -> image [H, W]: pixel-driven, linear detector interpolation, zero outside."""
    sino = np.asarray(sino, dtype=np.float64)
    A, D = sino.shape
    y, x = np.meshgrid(np.arange(H) + 0.5 - 0.5 * H, np.arange(W) + 0.5 - 0.5 * W, indexing="ij")
    out = np.zeros((H, W))
    for a, th in enumerate(np.asarray(angles, dtype=np.float64)):
        kb = (x * np.cos(th) + y * np.sin(th)) / det_spacing + 0.5 * D - 0.5
        k0 = np.floor(kb).astype(np.int64)
        f = kb - k0
        v0 = np.where((k0 >= 0) & (k0 < D), sino[a, np.clip(k0, 0, D - 1)], 0.0)
        v1 = np.where((k0 + 1 >= 0) & (k0 + 1 < D), sino[a, np.clip(k0 + 1, 0, D - 1)], 0.0)
        out += v0 + f * (v1 - v0)
    return out


def physics_loss_syn(fake: np.ndarray, real: np.ndarray, M: np.ndarray, angles: np.ndarray) -> float:
    """losses/gan_losses.py:222-253 on [B, H, W] arrays: mean((1 - [P(M) > 0]) * |P(fake) - P(real)|)."""
    tot, cnt = 0.0, 0
    for f, r, m in zip(fake, real, M):
        pf, pr, pm = radon_forward(f, angles), radon_forward(r, angles), radon_forward(m, angles)
        tot += ((1.0 - (pm > 0)) * np.abs(pf - pr)).sum()
        cnt += pf.size
    return tot / cnt
