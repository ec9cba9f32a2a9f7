"""TEST INFRASTRUCTURE.  Pins transmar_gan_b200/gan.py (the stock-PyTorch discriminator and losses of the
full training step) to the UNMODIFIED reference files, which are imported here from /root/reference (build
container only) and used to write tests/golden/gan_stock.pt:
  * a small MultiScaleDiscriminator(2, 8, 5, 3, True) - state_dict before the forward, input, train-mode
    logits and features (one power iteration of the spectral norm);
  * key names and shapes of the default MultiScaleDiscriminator(2, 64, 5, 3, True);
  * every loss of losses/gan_losses.py on seeded inputs with a metal disc.
usage: python oracle/make_golden_gan.py
"""
import importlib.util
import os
import sys

import torch

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "gan_stock.pt")


def load(name, rel):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF, rel))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def main():
    disc = load("ref_ms_patchgan", "models/discriminator/ms_patchgan.py")
    losses = load("ref_gan_losses", "losses/gan_losses.py")
    torch.manual_seed(7)
    d = disc.MultiScaleDiscriminator(in_channels=2, base_channels=8, num_layers=5, num_scales=3, use_sn=True)
    for m in d.modules():
        if isinstance(m, torch.nn.Conv2d):
            torch.nn.init.normal_(m.weight_orig, 0.0, 0.2)
            torch.nn.init.normal_(m.bias, 0.0, 0.1)
    d.train()
    sd = {k: v.clone() for k, v in d.state_dict().items()}
    x = torch.randn(1, 2, 128, 128)
    logits, feats = d(x, return_features=True)
    big = disc.MultiScaleDiscriminator(2, 64, 5, 3, True)
    keys = [(k, tuple(v.shape)) for k, v in big.state_dict().items()]
    g = torch.Generator().manual_seed(11)
    fake = torch.rand(2, 1, 32, 32, generator=g) * 2 - 1
    real = torch.rand(2, 1, 32, 32, generator=g) * 2 - 1
    ct = torch.rand(2, 1, 32, 32, generator=g) * 1.2 - 1          # mostly below the 0.6 threshold ...
    yy, xx = torch.meshgrid(torch.arange(32), torch.arange(32), indexing="ij")
    ct[:, :, ((yy - 12) ** 2 + (xx - 20) ** 2) <= 9] = 1.0        # ... plus a metal disc
    M = losses.extract_metal_mask(ct, threshold=0.6)
    w = losses.compute_weight_map(ct, beta=1.0, radius=5, w_max=3.0, threshold=0.6)
    rl = [torch.randn(2, 1, 7, 7, generator=g), torch.randn(2, 1, 3, 3, generator=g)]
    fl = [torch.randn(2, 1, 7, 7, generator=g), torch.randn(2, 1, 3, 3, generator=g)]
    gold = {
        "small_state": sd, "x": x, "logits": [t.detach() for t in logits],
        "feats": [[t.detach() for t in fs] for fs in feats],
        "small_state_after": {k: v.clone() for k, v in d.state_dict().items() if k.endswith("_u")},
        "default_keys": keys,
        "fake": fake, "real": real, "ct": ct, "M": M, "w": w, "dilated": losses.dilate_mask(M, radius=5),
        "rl": rl, "fl": fl,
        "hinge_d": losses.hinge_d_loss(rl, fl), "hinge_g": losses.hinge_g_loss(fl),
        "fm": losses.feature_matching_loss([[t.detach() for t in fs] for fs in feats],
                                           [[t.detach() * 0.5 for t in fs] for fs in feats]),
        "rec_mw": losses.compute_metal_aware_loss(fake, real, ct, beta=1.0, radius=5, w_max=3.0, threshold=0.6),
        "edge_mw": losses.compute_metal_aware_edge_loss(fake, real, w),
        "metal": losses.metal_consistency_loss(fake, real, M),
        "grads": losses.compute_image_gradients(fake),
    }
    torch.save(gold, OUT)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    sys.exit(main())
