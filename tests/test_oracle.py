"""CPU tests of the oracle (oracle/ngswin_oracle.py): pinned against the golden vectors that
oracle/make_golden.py wrote from the UNMODIFIED reference, and — when /root/reference is
mounted (build container only) — against the live reference modules."""
import math

import pytest
import torch

from oracle import ngswin_oracle as O
from oracle import ref_shim


def _sd(kind):
    return O.synthetic_state_dict(kind, seed=1)


def test_contract_matches_reference_dump(golden_contract):
    mine = [(k, s, str(dt)) for k, s, dt in O.reference_key_shapes()]
    assert mine == golden_contract["contract"]
    assert len(mine) == 554
    assert golden_contract["n_params"] == 990811


def test_relative_position_index_golden(golden_contract):
    assert torch.equal(O.relative_position_index(8, 8), golden_contract["rel_index_8"])
    assert torch.equal(O.relative_position_index(2, 2), golden_contract["rel_index_2"])
    assert O.relative_position_index(2, 2).tolist() == [[4, 3, 1, 0], [5, 4, 2, 1], [7, 6, 4, 3], [8, 7, 5, 4]]


def test_partition_index_maps_bit_exact(golden_maps):
    for (B, ph, pw, s), ref in golden_maps["partition_maps"].items():
        tok, _ = O.partition_source_index(B, ph, pw, s)
        assert torch.equal(tok, ref), (B, ph, pw, s)


def test_shift_masks_bit_exact(golden_maps):
    for (ph, pw), ref in golden_maps["masks_i8"].items():
        assert torch.equal(O.make_mask(ph, pw).to(torch.int8), ref), (ph, pw)


@pytest.mark.parametrize("kind", ["W1", "W2"])
@pytest.mark.parametrize("shape", [(1, 64, 64), (1, 100, 70)])
def test_whole_model_vs_golden(golden_outputs, kind, shape):
    B, H, W = shape
    with torch.no_grad():
        y = O.ngswin_forward(O.synthetic_input(B, H, W), _sd(kind))
    ref = golden_outputs[(kind, B, H, W)]
    assert y.shape == ref.shape == (B, 1, H, W)
    assert (y - ref).abs().max().item() < 2e-5


@pytest.mark.parametrize("kind", ["W1", "W2"])
def test_modules_vs_golden(golden_outputs, kind):
    sd = _sd(kind)
    g = torch.Generator().manual_seed(golden_outputs[(kind, "z_seed")])
    z = torch.randn(1, 24 * 40, 64, generator=g)
    with torch.no_grad():
        for bi in (0, 1):
            shift = 0 if bi == 0 else 4
            out = O.nstb(z, 24, 40, sd, f"encoder_layer1.blocks.{bi}.", 6, shift)
            assert (out - golden_outputs[(kind, "nstb6", bi)]).abs().max() < 5e-5
            win = O.ngram_window_partition(z.view(1, 24, 40, 64), sd,
                                           f"encoder_layer1.blocks.{bi}.ngram_window_partition.", 6, shift)
            assert (win - golden_outputs[(kind, "nwp6", bi)]).abs().max() < 2e-5
            ctx = O.ngram_context(z.view(1, 24, 40, 64), sd,
                                  f"encoder_layer1.blocks.{bi}.ngram_window_partition.ngram_context.", 6)
            assert (ctx - golden_outputs[(kind, "ctx6", bi)]).abs().max() < 2e-5
            out4 = O.nstb(z, 24, 40, sd, f"encoder_layer2.blocks.{bi}.", 4, shift)
            assert (out4 - golden_outputs[(kind, "nstb4", bi)]).abs().max() < 5e-5
        pm = O.patch_merging(z, 24, 40, sd, "encoder_layer1.downsample.")
        assert (pm - golden_outputs[(kind, "patch_merge")]).abs().max() < 2e-5


def test_pixel_shuffle_golden(golden_outputs):
    x2 = torch.arange(1 * 6 * 10 * 64, dtype=torch.float32).reshape(1, 60, 64)
    assert torch.equal(O.pixel_shuffle_permute(x2, 6, 10, 2), golden_outputs[("pixel_shuffle", 2)])
    x4 = torch.arange(1 * 3 * 5 * 64, dtype=torch.float32).reshape(1, 15, 64)
    assert torch.equal(O.pixel_shuffle_permute(x4, 3, 5, 4), golden_outputs[("pixel_shuffle", 4)])


def test_l1_gradients_vs_golden(golden_outputs):
    """autograd through the oracle reproduces the reference's gradients (config-3 recipe)."""
    kind = "W2"
    sd = {k: (v.clone().requires_grad_(True) if v.is_floating_point() and not k.endswith("attn_mask")
              else v) for k, v in _sd(kind).items()}
    x = O.synthetic_input(2, 64, 64, seed=999).requires_grad_(True)
    t = O.synthetic_input(2, 64, 64, seed=1000)
    loss = torch.nn.functional.l1_loss(O.ngswin_forward(x, sd), t)
    loss.backward()
    assert abs(loss.item() - golden_outputs[(kind, "l1_loss")].item()) < 1e-5
    assert (x.grad - golden_outputs[(kind, "grad_x")]).abs().max() < 1e-6 + 1e-3 * golden_outputs[(kind, "grad_x")].abs().max()
    grads = golden_outputs[(kind, "grads")]
    assert len(grads) == 504
    worst = 0.0
    for n, gref in grads.items():
        rel = (sd[n].grad - gref).norm() / gref.norm().clamp_min(1e-12)
        worst = max(worst, rel.item())
    assert worst < 1e-3, worst


def test_small_inputs_raise_like_reference():
    with pytest.raises(RuntimeError):
        O.ngswin_forward(torch.zeros(1, 1, 32, 32), _sd("W1"))


def test_psnr_ssim_sanity():
    a = O.synthetic_input(1, 64, 64)[0, 0]
    assert O.psnr(a, a) == float("inf")
    assert abs(O.ssim(a, a) - 1.0) < 1e-12
    b = a + 0.01
    assert abs(O.psnr(a, b) - 10 * math.log10(4 / 1e-4)) < 1e-3   # fp32 rounding of a+0.01


@pytest.mark.skipif(not ref_shim.reference_available(), reason="reference tree not mounted")
def test_oracle_pinned_to_live_reference():
    R = ref_shim.load_reference()
    model = R["NGswin"]().eval()
    for kind in ("W1", "W2"):
        sd = O.synthetic_state_dict(kind, seed=7)
        model.load_state_dict(sd, strict=True)
        x = O.synthetic_input(1, 96, 64, seed=3)
        with torch.no_grad():
            assert (model(x) - O.ngswin_forward(x, sd)).abs().max() < 2e-5
