"""TEST INFRASTRUCTURE — loader for the UNMODIFIED reference NGswin (container only).

`/root/reference` is not importable as shipped: its files were
renamed but their relative imports were not, and `timm` is absent.  This module builds
the import environment the reference files expect and executes them *where they lie*:

* a stand-in for the handful of `timm` names the path imports (`to_2tuple`, `to_ntuple`,
  `_assert`, `trunc_normal_`, `DropPath`, `Mlp`, and four imported-but-unused helpers);
  `Mlp` follows timm's published layout `fc1 → act → drop1 → fc2 → drop2`
  (reference call site: models/generator/components/transformer_blocks.py:60) — timm is
  un-vendored and unpinned (requirements.txt:6), so this boundary is "parity unpinned";
* the reference files loaded under a PRIVATE package name (`_tmg_ref…`) with module
  aliases for the stale relative-import names (`.pool`, `.main_branch`, `.nstb`,
  `.win_attention`, `.win_partition`).

Nothing here is copied from the reference, nothing is shipped in the product path, and
nothing in `-m gpu` tests, `smoke()` or `bench.py` imports this file: the reference tree
does not exist on the GPU box.  It is used by `oracle/make_golden.py` and by the
container-only oracle-pinning tests (skipped when `/root/reference` is absent).
"""
from __future__ import annotations

import collections.abc
import importlib.util
import os
import sys
import types
from itertools import repeat

import torch
import torch.nn as nn

REFERENCE_ROOT = os.environ.get("TMG_REFERENCE_ROOT", "/root/reference")
_PKG = "_tmg_ref"


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "models/generator/ngswin.py"))


# --------------------------------------------------------------------------- timm stand-in
def _ntuple(n):
    def parse(x):
        if isinstance(x, collections.abc.Iterable) and not isinstance(x, str):
            return tuple(x)
        return tuple(repeat(x, n))
    return parse


class _DropPath(nn.Module):
    def __init__(self, drop_prob: float = 0.0):
        super().__init__()
        self.drop_prob = drop_prob

    def forward(self, x):
        if self.drop_prob == 0.0 or not self.training:
            return x
        keep = 1 - self.drop_prob
        mask = x.new_empty((x.shape[0],) + (1,) * (x.ndim - 1)).bernoulli_(keep)
        return x * mask / keep


class _Mlp(nn.Module):
    """timm.layers.Mlp as published: fc1 -> act -> drop1 -> fc2 -> drop2."""

    def __init__(self, in_features, hidden_features=None, out_features=None,
                 act_layer=nn.GELU, drop=0.0):
        super().__init__()
        out_features = out_features or in_features
        hidden_features = hidden_features or in_features
        self.fc1 = nn.Linear(in_features, hidden_features)
        self.act = act_layer()
        self.drop1 = nn.Dropout(drop)
        self.fc2 = nn.Linear(hidden_features, out_features)
        self.drop2 = nn.Dropout(drop)

    def forward(self, x):
        return self.drop2(self.fc2(self.drop1(self.act(self.fc1(x)))))


def _install_timm_standin():
    if "timm" in sys.modules and not getattr(sys.modules["timm"], "_tmg_standin", False):
        return  # a real timm is importable; use it
    def mod(name):
        m = types.ModuleType(name)
        m._tmg_standin = True
        sys.modules[name] = m
        return m
    timm = mod("timm")
    models = mod("timm.models")
    layers = mod("timm.models.layers")
    helpers = mod("timm.models.helpers")
    vit = mod("timm.models.vision_transformer")
    eff = mod("timm.models._efficientnet_builder")
    fx = mod("timm.models.fx_features")
    timm.models = models
    models.layers, models.helpers, models.vision_transformer = layers, helpers, vit
    models._efficientnet_builder, models.fx_features = eff, fx
    layers.to_2tuple = _ntuple(2)
    layers.to_ntuple = _ntuple
    def _assert(cond, msg=""):
        assert cond, msg
    layers._assert = _assert
    layers.trunc_normal_ = nn.init.trunc_normal_
    layers.DropPath = _DropPath
    layers.Mlp = _Mlp
    helpers.named_apply = lambda *a, **k: None           # imported, never called
    vit.get_init_weights_vit = lambda *a, **k: None      # imported, never called
    eff._init_weight_goog = lambda *a, **k: None         # imported, never called
    fx.register_notrace_function = lambda f: f


# --------------------------------------------------------------------------- file loader
def _load(name: str, relpath: str, is_pkg: bool = False):
    full = f"{_PKG}.{name}" if name else _PKG
    if full in sys.modules:
        return sys.modules[full]
    path = os.path.join(REFERENCE_ROOT, relpath)
    if is_pkg:
        m = types.ModuleType(full)
        m.__path__ = [path]
        m.__package__ = full
        sys.modules[full] = m
        return m
    spec = importlib.util.spec_from_file_location(full, path)
    m = importlib.util.module_from_spec(spec)
    sys.modules[full] = m
    spec.loader.exec_module(m)
    return m


def load_reference():
    """Return the reference's module objects: dict with NGswin, NSTB, WindowAttention, …"""
    if not reference_available():
        raise FileNotFoundError(f"reference tree not found at {REFERENCE_ROOT}")
    _install_timm_standin()
    for pkg, rel in (("", ""), ("models", "models"), ("models.generator", "models/generator"),
                     ("models.generator.components", "models/generator/components")):
        _load(pkg, rel, is_pkg=True)
    # `from utils.etc_utils import denormalize` (ngswin.py:16) is an absolute import.
    saved_utils = {k: sys.modules.get(k) for k in ("utils", "utils.etc_utils")}
    utils_pkg = types.ModuleType("utils")
    utils_pkg.__path__ = [os.path.join(REFERENCE_ROOT, "utils")]
    sys.modules["utils"] = utils_pkg
    spec = importlib.util.spec_from_file_location(
        "utils.etc_utils", os.path.join(REFERENCE_ROOT, "utils/etc_utils.py"))
    etc = importlib.util.module_from_spec(spec)
    sys.modules["utils.etc_utils"] = etc
    spec.loader.exec_module(etc)
    try:
        comp = "models.generator.components"
        cdir = "models/generator/components"
        att = _load(f"{comp}.attention", f"{cdir}/attention.py")
        sys.modules[f"{_PKG}.{comp}.win_attention"] = att
        wu = _load(f"{comp}.window_utils", f"{cdir}/window_utils.py")
        sys.modules[f"{_PKG}.{comp}.win_partition"] = wu
        pool = _load(f"{comp}.pooling", f"{cdir}/pooling.py")
        sys.modules[f"{_PKG}.{comp}.pool"] = pool
        tb = _load(f"{comp}.transformer_blocks", f"{cdir}/transformer_blocks.py")
        sys.modules[f"{_PKG}.{comp}.nstb"] = tb
        ed = _load(f"{comp}.encoder_decoder", f"{cdir}/encoder_decoder.py")
        rec = _load(f"{comp}.reconstruction", f"{cdir}/reconstruction.py")
        # ngswin.py does `from .pool`, `.main_branch`, `.reconstruction` relative to models.generator
        sys.modules[f"{_PKG}.models.generator.pool"] = pool
        sys.modules[f"{_PKG}.models.generator.main_branch"] = ed
        sys.modules[f"{_PKG}.models.generator.reconstruction"] = rec
        ng = _load("models.generator.ngswin", "models/generator/ngswin.py")
    finally:
        for k, v in saved_utils.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    return {
        "NGswin": ng.NGswin, "NSTB": tb.NSTB, "WindowAttention": att.WindowAttention,
        "get_relative_position_index": att.get_relative_position_index,
        "NGramContext": wu.NGramContext, "NGramWindowPartition": wu.NGramWindowPartition,
        "window_partition": wu.window_partition, "window_unpartition": wu.window_unpartition,
        "PatchMerging": ed.PatchMerging, "pixel_shuffle_permute": ed.pixel_shuffle_permute,
        "SCDPBottleneck": ed.SCDPBottleneck, "EncoderLayer": ed.EncoderLayer,
        "DecoderLayer": ed.DecoderLayer, "InterPool": pool.InterPool,
        "BottleneckPool": pool.BottleneckPool, "Reconstruction": rec.Reconstruction,
    }


def weights_init_w1(model: nn.Module, seed: int = 0):
    """W1: the callers' `weights_init` (training/train_combined.py:207-217): N(0,.02)
    on Conv/Linear weights, zero biases, LayerNorm ones/zeros; bare Parameters untouched."""
    g = torch.Generator().manual_seed(seed)
    for m in model.modules():
        if isinstance(m, (nn.Conv2d, nn.Linear)):
            with torch.no_grad():
                m.weight.copy_(torch.randn(m.weight.shape, generator=g) * 0.02)
                if m.bias is not None:
                    m.bias.zero_()
        elif isinstance(m, nn.LayerNorm):
            with torch.no_grad():
                m.weight.fill_(1.0)
                m.bias.zero_()
    return model
