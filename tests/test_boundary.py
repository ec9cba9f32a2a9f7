"""CPU tests of the drop-in boundary: state_dict contract, constructor, C-ABI export table.
No compute entry point is called (there is no GPU here)."""
import os
import re

import pytest
import torch

import transmar_gan_b200 as T
from transmar_gan_b200 import _cabi
from oracle import ngswin_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def model():
    torch.manual_seed(0)
    return T.NGswin()


def test_state_dict_contract(model, golden_contract):
    mine = [(k, tuple(v.shape), str(v.dtype)) for k, v in model.state_dict().items()]
    assert mine == golden_contract["contract"]
    assert sum(p.numel() for p in model.parameters()) == golden_contract["n_params"]
    assert len(list(model.parameters())) == 504


def test_buffers_match_reference_values(model, golden_contract, golden_maps):
    sd = model.state_dict()
    assert torch.equal(sd["encoder_layer1.blocks.0.attn.relative_position_index"],
                       golden_contract["rel_index_8"])
    k = "encoder_layer1.blocks.0.ngram_window_partition.ngram_context.ngram_attn.relative_position_index"
    assert torch.equal(sd[k], golden_contract["rel_index_2"])
    # encoder_layer1 masks are built for training_img_size 64 -> 64 windows
    assert torch.equal(sd["encoder_layer1.blocks.1.attn_mask"].to(torch.int8),
                       golden_maps["masks_i8"][(64, 64)])
    assert sd["decoder_layer1.blocks.1.attn_mask"].shape == (16, 64, 64)


def test_cross_load_with_reference_layout(model):
    for kind in ("W1", "W2"):
        sd = O.synthetic_state_dict(kind, seed=3)
        model.load_state_dict(sd, strict=True)
        back = model.state_dict()
        assert list(back.keys()) == list(sd.keys())
        for k in sd:
            assert torch.equal(back[k], sd[k]), k


def test_generator_wrapper_prefix_and_apply(model):
    """Callers wrap NGswin as Generator.main and run netG.apply(weights_init)
    (training/train_combined.py:222-237); checkpoints carry the `main.` prefix."""
    class Generator(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.main = T.NGswin()

    g = Generator()
    seen = {"lin": 0, "conv": 0, "ln": 0}

    def weights_init(m):
        if isinstance(m, torch.nn.Linear):
            seen["lin"] += 1
        elif isinstance(m, torch.nn.Conv2d):
            seen["conv"] += 1
        elif isinstance(m, torch.nn.LayerNorm):
            seen["ln"] += 1
    g.apply(weights_init)
    # per NSTB: qkv, proj (x2 attentions), fc1, fc2 = 6 Linear; 2 Conv; 2 LN
    assert seen["lin"] == 20 * 6 + 2 + 2 + 1
    assert seen["conv"] == 20 * 2 + 1 + 1 + 2
    assert seen["ln"] == 20 * 2 + 2 + 1 + 1
    assert all(k.startswith("main.") for k in g.state_dict())
    g.load_state_dict({"main." + k: v for k, v in O.synthetic_state_dict("W1").items()}, strict=True)


def test_reference_import_path():
    from models.generator.ngswin import NGswin
    assert NGswin is T.NGswin


def test_constructor_signature_and_refusals():
    import inspect
    names = list(inspect.signature(T.NGswin.__init__).parameters)[1:]
    assert names == ["training_img_size", "ngrams", "in_chans", "embed_dim", "depths", "num_heads",
                     "head_dim", "dec_dim", "dec_depths", "dec_num_heads", "dec_head_dim",
                     "target_mode", "window_size", "mlp_ratio", "qkv_bias", "img_norm", "drop_rate",
                     "attn_drop_rate", "drop_path_rate", "act_layer", "norm_layer", "weight_init",
                     "kwargs"]
    with pytest.raises(AssertionError):
        T.NGswin(target_mode="heavy")
    with pytest.raises(NotImplementedError):
        T.NGswin(embed_dim=96)
    with pytest.raises(NotImplementedError):
        T.NGswin(drop_path_rate=0.1)
    m = T.NGswin(training_img_size=128)          # mask buffers follow training_img_size
    assert m.state_dict()["encoder_layer1.blocks.1.attn_mask"].shape == (256, 64, 64)


def test_no_cpu_forward(model):
    with pytest.raises(RuntimeError, match="no CPU path"):
        model(torch.zeros(1, 1, 64, 64))
    with pytest.raises(RuntimeError, match="no CPU path"):
        T.ops.unigram_fwd(torch.zeros(1, 64, 64), torch.zeros(32, 2, 8, 8), torch.zeros(32), 8, 8)


def test_no_weight_decay_and_flops(model):
    nwd = model.no_weight_decay()
    assert len(nwd) == 40 and all("relative_position_bias_table" in n for n in nwd)
    # flops() reproduces the reference's analytic count (values printed by the unmodified reference's
    # NGswin().flops(), models/generator/ngswin.py:212-238); flops_exact() is the exact MAC count
    assert model.flops((416, 416)) == 97516238144
    assert model.flops((64, 64)) == 2308080704
    assert model.flops((100, 70)) == 6924233792
    assert model.flops((1024, 1024)) == 590867599424
    assert model.flops_exact((416, 416)) == 416 * 416 * 548013
    assert model.flops_exact((100, 70)) == 128 * 96 * 548013


def test_cabi_exports_every_declared_symbol():
    """The shared library loads and exports exactly what include/b200mar.h declares."""
    header = open(os.path.join(ROOT, "include", "b200mar.h")).read()
    declared = set(re.findall(r"\b(b200mar_[a-z0-9_]+)\s*\(", header))
    assert declared == set(_cabi.SIGNATURES), declared ^ set(_cabi.SIGNATURES)
    _cabi.build()                       # no-op when up to date; nvcc cross-compiles without a GPU
    lib = _cabi.lib()
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.b200mar_abi_version() == 7
    assert lib.b200mar_last_error() == b"no error"


def test_product_path_never_imports_oracle():
    pkg = os.path.join(ROOT, "transmar_gan_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert "oracle" not in src, fn


def test_flops_matches_live_reference(model):
    """flops() against the unmodified reference's own method, when the reference tree is mounted."""
    from oracle import ref_shim
    if not ref_shim.reference_available():
        pytest.skip("reference tree not mounted")
    ref = ref_shim.load_reference()["NGswin"]()
    for res in [(416, 416), (128, 128), (96, 160), (100, 70), (1024, 1024)]:
        assert model.flops(res) == ref.flops(res)
