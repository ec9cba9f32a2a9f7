"""transmar_gan_b200/gan.py (the stock-PyTorch discriminator and losses around the generator, BASELINE
config 4) against the reference: fixtures written by oracle/make_golden_gan.py from the unmodified
models/discriminator/ms_patchgan.py and losses/gan_losses.py."""
import os

import pytest
import torch

from transmar_gan_b200 import gan

GOLD = os.path.join(os.path.dirname(__file__), "golden", "gan_stock.pt")


@pytest.fixture(scope="module")
def gold():
    return torch.load(GOLD, weights_only=False)


def test_discriminator_state_dict_keys_match_reference(gold):
    d = gan.MultiScaleDiscriminator(2, 64, 5, 3, True)
    assert [(k, tuple(v.shape)) for k, v in d.state_dict().items()] == [tuple(e) for e in gold["default_keys"]]
    assert sum(p.numel() for p in d.parameters()) == 20852547


def test_discriminator_forward_matches_reference(gold):
    d = gan.MultiScaleDiscriminator(2, 8, 5, 3, True)
    d.load_state_dict(gold["small_state"], strict=True)
    d.train()                                                           # one spectral-norm power iteration
    logits, feats = d(gold["x"], return_features=True)
    for a, b in zip(logits, gold["logits"]):
        assert a.shape == b.shape and torch.allclose(a, b, rtol=1e-5, atol=1e-6)
    for fa, fb in zip(feats, gold["feats"]):
        assert len(fa) == len(fb) == 4
        for a, b in zip(fa, fb):
            assert torch.allclose(a, b, rtol=1e-5, atol=1e-6)
    for k, v in gold["small_state_after"].items():                      # the power iteration updated u alike
        assert torch.allclose(d.state_dict()[k], v, rtol=1e-5, atol=1e-6)
    lg, none = d(gold["x"], return_features=False)
    assert none is None and len(lg) == 3


def test_losses_match_reference(gold):
    g = gold
    assert torch.equal(gan.extract_metal_mask(g["ct"], 0.6), g["M"])
    assert torch.equal(gan.dilate_mask(g["M"], 5), g["dilated"])
    assert torch.equal(gan.compute_weight_map(g["ct"], 1.0, 5, 3.0, 0.6), g["w"])
    for a, b in zip(gan.compute_image_gradients(g["fake"]), g["grads"]):
        assert torch.equal(a, b)
    close = lambda a, b: torch.allclose(a, b, rtol=1e-6, atol=1e-7)
    assert close(gan.hinge_d_loss(g["rl"], g["fl"]), g["hinge_d"])
    assert close(gan.hinge_g_loss(g["fl"]), g["hinge_g"])
    assert close(gan.feature_matching_loss(g["feats"], [[t * 0.5 for t in fs] for fs in g["feats"]]), g["fm"])
    assert close(gan.compute_metal_aware_loss(g["fake"], g["real"], g["ct"], 1.0, 5, 3.0, 0.6), g["rec_mw"])
    assert close(gan.compute_metal_aware_edge_loss(g["fake"], g["real"], g["w"]), g["edge_mw"])
    assert close(gan.metal_consistency_loss(g["fake"], g["real"], g["M"]), g["metal"])


def test_weights_init_recipe():
    m = torch.nn.Sequential(torch.nn.Conv2d(3, 4, 3), torch.nn.LayerNorm(4), torch.nn.Linear(4, 2))
    torch.manual_seed(0)
    m.apply(gan.weights_init)
    assert m[0].bias.abs().max() == 0 and m[2].bias.abs().max() == 0
    assert (m[1].weight == 1).all() and (m[1].bias == 0).all()
    assert 0.005 < m[0].weight.std() < 0.04
