"""TEST INFRASTRUCTURE — structured synthetic CT-like slices (no dataset ships with the reference, there is
no network): a phantom of random ellipses (the clean target) corrupted by radial streaks through small
saturated "metal" discs plus noise (the input), both in [-1, 1] like data/datasets.py:107-109.
Used by oracle/make_w3.py (trained-like weights) and tests/test_gpu_convergence.py; never by the product path."""
from __future__ import annotations

import math

import torch


def phantom_batch(B: int, S: int, g: torch.Generator):
    """(corrupted, clean) pairs [B,1,S,S] in [-1,1]: ellipses + metal discs + streaks + noise."""
    yy, xx = torch.meshgrid(torch.linspace(-1, 1, S), torch.linspace(-1, 1, S), indexing="ij")
    clean = torch.full((B, 1, S, S), -1.0)
    corrupt = torch.empty_like(clean)
    for b in range(B):
        img = torch.full((S, S), -1.0)
        for _ in range(int(torch.randint(3, 7, (1,), generator=g))):
            cx, cy = (torch.rand(2, generator=g) * 1.2 - 0.6).tolist()
            ax, ay = (torch.rand(2, generator=g) * 0.5 + 0.15).tolist()
            th = float(torch.rand(1, generator=g)) * math.pi
            val = float(torch.rand(1, generator=g)) * 1.2 - 0.7
            xr = (xx - cx) * math.cos(th) + (yy - cy) * math.sin(th)
            yr = -(xx - cx) * math.sin(th) + (yy - cy) * math.cos(th)
            d = (xr / ax) ** 2 + (yr / ay) ** 2
            img = torch.where(d < 1.0, torch.full_like(img, val) + 0.1 * (1 - d), img)
        art = torch.zeros(S, S)
        metal = torch.zeros(S, S, dtype=torch.bool)
        for _ in range(int(torch.randint(1, 3, (1,), generator=g))):
            mx, my = (torch.rand(2, generator=g) * 1.0 - 0.5).tolist()
            metal |= ((xx - mx) ** 2 + (yy - my) ** 2) < (3.5 / S * 2) ** 2
            ang = torch.atan2(yy - my, xx - mx)
            r = ((xx - mx) ** 2 + (yy - my) ** 2).sqrt()
            k = int(torch.randint(6, 14, (1,), generator=g))
            ph = float(torch.rand(1, generator=g)) * 2 * math.pi
            art += 0.35 * torch.cos(k * ang + ph) * torch.exp(-1.5 * r)
        img_c = img.clone()
        img_c[metal] = 1.0
        noisy = (img_c + art + 0.03 * torch.randn(S, S, generator=g)).clamp(-1, 1)
        noisy[metal] = 1.0
        clean[b, 0], corrupt[b, 0] = img_c.clamp(-1, 1), noisy
    return corrupt, clean
