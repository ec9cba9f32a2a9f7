"""TEST INFRASTRUCTURE — writes tests/golden/w3_trained.pt: a TRAINED-LIKE weight set (W3) and the
UNMODIFIED reference's outputs on it.

Run in the build container only (needs /root/reference):  python oracle/make_w3.py [steps]

The reference's checkpoint (best_model_syndeeplesion.pth) is a missing large blob, so nobody knows where
trained weights sit relative to the init recipe (W1) and the stress set (W2).  W3 is produced by the
reference itself: the unmodified `NGswin` (through oracle/ref_shim.py) starts from W1 (seed 1) and is
trained on CPU with the reference's L1 recipe (scripts/run_ablation_studies.py:1525-1535: Adam betas
(0.5, 0.999); lr 2e-4 so that a few hundred steps move the weights by several init standard deviations) on
STRUCTURED synthetic slices: a phantom of random ellipses (the clean target) corrupted by radial streaks
through small saturated "metal" discs plus noise (the input), both in [-1, 1] like data/datasets.py:107-109.
Stored: the trained state_dict (fp32), the loss trajectory, the reference's outputs on the shapes the W1 / W2
fixtures use, and the L1-loss gradients of blocks 0 / 1 of every stage + all non-block parameters.
"""
from __future__ import annotations

import math
import os
import sys
import time

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ngswin_oracle as O  # noqa: E402
from oracle import ref_shim  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "w3_trained.pt")


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


def main():
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 400
    torch.manual_seed(0)
    torch.set_num_threads(max(1, (os.cpu_count() or 2) - 2))
    R = ref_shim.load_reference()
    model = R["NGswin"]()
    model.load_state_dict(O.synthetic_state_dict("W1", seed=1), strict=True)
    model.train()
    opt = torch.optim.Adam(model.parameters(), lr=2e-4, betas=(0.5, 0.999))
    g = torch.Generator().manual_seed(31337)
    losses, t0 = [], time.time()
    for it in range(steps):
        x, t = phantom_batch(4, 64, g)
        loss = F.l1_loss(model(x), t)
        opt.zero_grad(set_to_none=True)
        loss.backward()
        opt.step()
        losses.append(float(loss))
        if it % 20 == 0:
            print(f"step {it} loss {losses[-1]:.4f} ({time.time() - t0:.0f} s)", flush=True)
    model.eval()
    sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    gold = {"state_dict": sd, "losses": torch.tensor(losses), "steps": steps}
    for (B, H, W) in ((1, 64, 64), (2, 96, 160), (1, 100, 70), (1, 128, 128)):
        with torch.no_grad():
            gold[("out", B, H, W)] = model(O.synthetic_input(B, H, W, seed=999)).clone()
    # a structured (phantom) input too: what the trained weights were fitted to
    xs, ts = phantom_batch(2, 96, torch.Generator().manual_seed(77))
    with torch.no_grad():
        gold["phantom_in"], gold["phantom_target"], gold["phantom_out"] = xs, ts, model(xs).clone()
    model.zero_grad(set_to_none=True)
    x = O.synthetic_input(2, 64, 64, seed=999).requires_grad_(True)
    t = O.synthetic_input(2, 64, 64, seed=1000)
    loss = F.l1_loss(model(x), t)
    loss.backward()
    gold["l1_loss"], gold["grad_x"] = loss.detach().clone(), x.grad.clone()
    gold["grads"] = {n: p.grad.clone() for n, p in model.named_parameters()
                     if ".blocks." not in n or ".blocks.0." in n or ".blocks.1." in n}
    # how far the weights moved, for the record
    w1 = O.synthetic_state_dict("W1", seed=1)
    mv = {k: float((sd[k] - w1[k]).abs().max()) for k in sd if sd[k].is_floating_point() and not k.endswith("attn_mask")}
    gold["max_move"] = max(mv.values())
    ls = [float(v.max()) for k, v in sd.items() if k.endswith("logit_scale")]
    gold["logit_scale_range"] = (min(float(v.min()) for k, v in sd.items() if k.endswith("logit_scale")), max(ls))
    torch.save(gold, OUT)
    print(f"wrote {OUT} ({os.path.getsize(OUT) / 1e6:.2f} MB); loss {losses[0]:.4f} -> {losses[-1]:.4f}; "
          f"max weight move {gold['max_move']:.3f}; logit_scale range {gold['logit_scale_range']}")


if __name__ == "__main__":
    main()
