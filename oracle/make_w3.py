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
from oracle.phantoms import phantom_batch  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "w3_trained.pt")


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
        losses.append(float(loss.detach()))
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
