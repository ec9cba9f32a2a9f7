"""TEST INFRASTRUCTURE — CPU restatement (oracle) of the reference NGswin hot path.

This file is a from-scratch, functional restatement in plain PyTorch (no nn.Module tree,
no einops) of what `/root/reference/models/generator/ngswin.py::NGswin.forward` computes,
driven by a reference-layout ``state_dict``.  It is the checker that ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs
use; the product path (``transmar_gan_b200``) never imports it.

Parity status: **pinned** against the unmodified reference imported through
``oracle/ref_shim.py`` in the build container (tests/test_oracle_pinned.py, and the golden
vectors under tests/golden/ written by oracle/make_golden.py).  The reference itself ships
no golden vectors or tests, and its one third-party arithmetic dependency,
timm's ``Mlp`` (un-vendored, unpinned), is restated from its published definition — that
single boundary is "parity unpinned" by anything the reference holds.

Every function cites the reference file:line it follows (paths relative to /root/reference).
All functions are dtype-generic: run them in float64 for a tighter checker.
"""
from __future__ import annotations

import math
from typing import Dict, Optional, Tuple

import torch
import torch.nn.functional as F

Tensor = torch.Tensor
WINDOW = 8
LOGIT_SCALE_MAX = math.log(1.0 / 0.01)  # attention.py:84


# ----------------------------------------------------------------------------- index maps
def relative_position_index(win_h: int, win_w: int) -> Tensor:
    """attention.py:9-18 — idx[p,q] = (p_i-q_i+win_h-1)*(2*win_w-1) + (p_j-q_j+win_w-1)."""
    pi = torch.arange(win_h).repeat_interleave(win_w)
    pj = torch.arange(win_w).repeat(win_h)
    di = pi[:, None] - pi[None, :] + (win_h - 1)
    dj = pj[:, None] - pj[None, :] + (win_w - 1)
    return (di * (2 * win_w - 1) + dj).to(torch.int64)


def shift_region_ids(ph: int, pw: int, window: int = WINDOW, shift: int = WINDOW // 2) -> Tensor:
    """transformer_blocks.py:63-76 — 9-region id map of the SHIFTED frame, [ph, pw] int64.

    Python slices ``slice(0,-window)``, ``slice(-window,-shift)``, ``slice(-shift,None)``
    are applied in order; later writes overwrite earlier ones (matters only for ph<window,
    which the reference rejects)."""
    def ids(n):
        r = torch.zeros(n, dtype=torch.int64)
        for k, s in enumerate((slice(0, -window), slice(-window, -shift), slice(-shift, None))):
            r[s] = k
        return r
    return ids(ph)[:, None] * 3 + ids(pw)[None, :]


def make_mask(ph: int, pw: int, window: int = WINDOW, shift: int = WINDOW // 2,
              dtype=torch.float32) -> Tensor:
    """transformer_blocks.py:63-81 — [nW, N, N] with -100 where region ids differ, else 0."""
    rid = shift_region_ids(ph, pw, window, shift)
    w = window_partition(rid[None, :, :, None].to(dtype), window).reshape(-1, window * window)
    diff = w[:, None, :] - w[:, :, None]
    return torch.where(diff != 0, torch.full_like(diff, -100.0), torch.zeros_like(diff))


def window_partition(x: Tensor, window: int = WINDOW) -> Tensor:
    """window_utils.py:155-169 — [B,ph,pw,D] -> [B*wh*ww, window, window, D], order (b,h,w)."""
    B, ph, pw, D = x.shape
    wh, ww = ph // window, pw // window
    x = x.reshape(B, wh, window, ww, window, D).permute(0, 1, 3, 2, 4, 5)
    return x.reshape(B * wh * ww, window, window, D)


def window_unpartition(win: Tensor, wh: int, ww: int) -> Tensor:
    """window_utils.py:171-181 — inverse of window_partition."""
    Bn, window, _, D = win.shape
    B = Bn // (wh * ww)
    x = win.reshape(B, wh, ww, window, window, D).permute(0, 1, 3, 2, 4, 5)
    return x.reshape(B, wh * window, ww * window, D)


def partition_source_index(B: int, ph: int, pw: int, shift: int, window: int = WINDOW):
    """Integer map of NGramWindowPartition + roll (window_utils.py:134-145): for every
    output element (window n, slot t) the flat source token index b*ph*pw + r*pw + c and the
    flat context index b*wh*ww + (r//window)*ww + c//window, with r=(8h+i+s)%ph, c=(8w+j+s)%pw.
    Returns (token_idx[B*wh*ww, 64], ctx_idx[B*wh*ww, 64]) as int64 tensors."""
    wh, ww = ph // window, pw // window
    b = torch.arange(B).view(B, 1, 1, 1, 1)
    h = torch.arange(wh).view(1, wh, 1, 1, 1)
    w = torch.arange(ww).view(1, 1, ww, 1, 1)
    i = torch.arange(window).view(1, 1, 1, window, 1)
    j = torch.arange(window).view(1, 1, 1, 1, window)
    r = (h * window + i + shift) % ph
    c = (w * window + j + shift) % pw
    tok = (b * ph * pw + r * pw + c).expand(B, wh, ww, window, window)
    ctx = (b * wh * ww + (r // window) * ww + (c // window)).expand(B, wh, ww, window, window)
    return tok.reshape(B * wh * ww, window * window), ctx.reshape(B * wh * ww, window * window)


# ----------------------------------------------------------------------------- attention
def window_attention(x: Tensor, p: Dict[str, Tensor], prefix: str, num_heads: int,
                     win: int, mask: Optional[Tensor] = None) -> Tensor:
    """attention.py:70-101 — Swin-V2 scaled-cosine window attention.

    x [B_, N, C]; parameters `{prefix}qkv.weight` … in reference layout; mask [nW,N,N]|None."""
    B_, N, C = x.shape
    qkv = F.linear(x, p[prefix + "qkv.weight"], p[prefix + "qkv.bias"])
    qkv = qkv.reshape(B_, N, 3, num_heads, -1).permute(2, 0, 3, 1, 4)
    q, k, v = qkv[0], qkv[1], qkv[2]
    qn = q / q.norm(dim=-1, keepdim=True).clamp_min(1e-12)          # F.normalize, :83
    kn = k / k.norm(dim=-1, keepdim=True).clamp_min(1e-12)
    attn = qn @ kn.transpose(-2, -1)
    scale = p[prefix + "logit_scale"].clamp(max=LOGIT_SCALE_MAX).exp()  # :84
    attn = attn * scale
    table = p[prefix + "relative_position_bias_table"]
    idx = relative_position_index(win, win).to(x.device)
    bias = table[idx.reshape(-1)].reshape(N, N, num_heads).permute(2, 0, 1)  # :64-68
    attn = attn + bias.unsqueeze(0)
    if mask is not None:                                             # :88-91
        nW = mask.shape[0]
        attn = attn.reshape(B_ // nW, nW, num_heads, N, N) + mask[None, :, None]
        attn = attn.reshape(B_, num_heads, N, N)
    attn = attn.softmax(dim=-1)
    out = (attn @ v).transpose(1, 2).reshape(B_, N, -1)
    return F.linear(out, p[prefix + "proj.weight"], p[prefix + "proj.bias"])


def unigram_embed(x: Tensor, p: Dict[str, Tensor], prefix: str) -> Tensor:
    """window_utils.py:36-38,78 — grouped 8x8/stride-8 conv, x [B,ph,pw,D] -> [B,D/2,wh,ww]."""
    D = x.shape[-1]
    return F.conv2d(x.permute(0, 3, 1, 2), p[prefix + "unigram_embed.weight"],
                    p[prefix + "unigram_embed.bias"], stride=WINDOW, groups=D // 2)


def _sliding_ngram_attention(upad: Tensor, p, prefix, num_heads) -> Tensor:
    """window_utils.py:63-73 (ngram=2): upad [B,C,wh+1,ww+1] -> [B,C,wh,ww].
    Tokens of window (h,w) are upad[h+i, w+j], order (i,j) row-offset major; the attention
    output is averaged over the 4 tokens (AvgPool2d(2))."""
    B, C, hp, wp = upad.shape
    wh, ww = hp - 1, wp - 1
    toks = torch.stack([upad[:, :, i:i + wh, j:j + ww] for i in (0, 1) for j in (0, 1)], dim=-1)
    toks = toks.permute(0, 2, 3, 4, 1).reshape(B * wh * ww, 4, C)       # [B*wh*ww, 4, C]
    ctx = window_attention(toks, p, prefix + "ngram_attn.", num_heads, 2, None)
    return ctx.mean(dim=1).reshape(B, wh, ww, C).permute(0, 3, 1, 2)


def ngram_context(x: Tensor, p: Dict[str, Tensor], prefix: str, num_heads: int) -> Tensor:
    """window_utils.py:75-90 — x [B,ph,pw,D] -> context [B,wh,ww,D]."""
    u = unigram_embed(x, p, prefix)                                     # [B,C,wh,ww]
    B, C, wh, ww = u.shape
    if wh < 2 or ww < 2:
        # window_utils.py:51-54: the reference broadcasts an empty slice and raises.
        raise RuntimeError("NGramContext needs at least 2x2 windows (reference raises here)")
    # forward pad (:49-54): append row wh := row wh-2, then column ww := column ww-2
    uf = torch.cat([u, u[:, :, wh - 2:wh - 1, :]], dim=2)
    uf = torch.cat([uf, uf[:, :, :, ww - 2:ww - 1]], dim=3)
    # backward pad (:55-60): prepend row := row 1, then column := column 1 (of the padded map)
    ub = torch.cat([u[:, :, 1:2, :], u], dim=2)
    ub = torch.cat([ub[:, :, :, 1:2], ub], dim=3)
    cf = _sliding_ngram_attention(uf, p, prefix, num_heads)
    cb = _sliding_ngram_attention(ub, p, prefix, num_heads)
    cat = torch.cat([cf, cb], dim=1)                                    # [B,D,wh,ww]
    ctx = F.conv2d(cat, p[prefix + "merge.weight"], p[prefix + "merge.bias"])
    return ctx.permute(0, 2, 3, 1)


def ngram_window_partition(x: Tensor, p, prefix: str, num_heads: int, shift: int) -> Tensor:
    """window_utils.py:127-147 — add context per UNSHIFTED window, cyclic roll, partition."""
    ctx = ngram_context(x, p, prefix + "ngram_context.", num_heads)     # [B,wh,ww,D]
    xc = x + ctx.repeat_interleave(WINDOW, dim=1).repeat_interleave(WINDOW, dim=2)
    if shift > 0:
        xc = torch.roll(xc, shifts=(-shift, -shift), dims=(1, 2))
    return window_partition(xc, WINDOW)


# ----------------------------------------------------------------------------- NSTB
def nstb(z: Tensor, ph: int, pw: int, p, prefix: str, num_heads: int, shift: int,
         materialise_mask_like_reference: bool = True) -> Tensor:
    """transformer_blocks.py:83-120 — z [B, ph*pw, D] -> block output (same shape).
    a = z + LN1(attn(z)); out = a + LN2(fc2(GELU(fc1(a))))."""
    B, T, D = z.shape
    wh, ww = ph // WINDOW, pw // WINDOW
    win = ngram_window_partition(z.reshape(B, ph, pw, D), p, prefix + "ngram_window_partition.",
                                 num_heads, shift).reshape(-1, WINDOW * WINDOW, D)
    mask = make_mask(ph, pw, WINDOW, shift, z.dtype).to(z.device) if shift > 0 else None
    aw = window_attention(win, p, prefix + "attn.", num_heads, WINDOW, mask)
    y = window_unpartition(aw.reshape(-1, WINDOW, WINDOW, D), wh, ww)
    if shift > 0:
        y = torch.roll(y, shifts=(shift, shift), dims=(1, 2))
    y = y.reshape(B, T, D)
    a = z + F.layer_norm(y, (D,), p[prefix + "norm1.weight"], p[prefix + "norm1.bias"], 1e-5)
    h = F.gelu(F.linear(a, p[prefix + "ffn.fc1.weight"], p[prefix + "ffn.fc1.bias"]))
    m = F.linear(h, p[prefix + "ffn.fc2.weight"], p[prefix + "ffn.fc2.bias"])
    return a + F.layer_norm(m, (D,), p[prefix + "norm2.weight"], p[prefix + "norm2.bias"], 1e-5)


def stage(x: Tensor, ph: int, pw: int, p, prefix: str, depth: int, num_heads: int):
    """encoder_decoder.py:135-145 / :252-258 — block input = previous output + previous
    input.  Returns (last block output, last block output + last block input)."""
    z_prev = None
    for i in range(depth):
        z = x if z_prev is None else x + z_prev
        x = nstb(z, ph, pw, p, f"{prefix}blocks.{i}.", num_heads, 0 if i % 2 == 0 else WINDOW // 2)
        z_prev = z
    return x, x + z_prev


def patch_merging(x: Tensor, ph: int, pw: int, p, prefix: str) -> Tensor:
    """encoder_decoder.py:51-74 — 2x2 concat in order (even,even),(even,odd),(odd,even),
    (odd,odd) -> LN(4D) -> Linear(4D->D, no bias)."""
    B, T, D = x.shape
    g = x.reshape(B, ph // 2, 2, pw // 2, 2, D)
    cat = torch.cat([g[:, :, 0, :, 0], g[:, :, 0, :, 1], g[:, :, 1, :, 0], g[:, :, 1, :, 1]], -1)
    cat = cat.reshape(B, T // 4, 4 * D)
    cat = F.layer_norm(cat, (4 * D,), p[prefix + "norm.weight"], p[prefix + "norm.bias"], 1e-5)
    return F.linear(cat, p[prefix + "reduction.weight"])


def inter_pool(x: Tensor, ph: int, pw: int) -> Tensor:
    """pooling.py:13-17 — 2x2 max-pool on the token grid."""
    B, T, C = x.shape
    g = x.reshape(B, ph // 2, 2, pw // 2, 2, C)
    return g.amax(dim=(2, 4)).reshape(B, T // 4, C)


def pixel_shuffle_permute(x: Tensor, ph: int, pw: int, s: int) -> Tensor:
    """encoder_decoder.py:159-164 — 'b (h w) (c ch cw) -> b (h ch w cw) c'."""
    B, T, K = x.shape
    c = K // (s * s)
    g = x.reshape(B, ph, pw, c, s, s).permute(0, 1, 4, 2, 5, 3)      # b h ch w cw c
    return g.reshape(B, ph * s * pw * s, c)


def bottleneck(shallow: Tensor, feats, sizes, p, prefix: str = "bottleneck.") -> Tensor:
    """encoder_decoder.py:188-200 + pooling.py:32-37."""
    H, W = sizes[0]
    outs = []
    for i, (f, (ph, pw)) in enumerate(zip(feats, sizes)):
        s = shallow
        for _ in range(i):
            s = F.max_pool2d(s, 2)
        s = F.leaky_relu(s, 0.01)                                     # applied for i = 0 too
        s = s.flatten(2).transpose(1, 2)
        outs.append(pixel_shuffle_permute(f + s, ph, pw, H // ph))
    x = torch.cat(outs, dim=-1)                                       # [B, T, 84]
    B, T, C = x.shape
    x = x.transpose(1, 2).reshape(B, C, H, W)
    x = F.gelu(F.conv2d(x, p[prefix + "depthwise.weight"], p[prefix + "depthwise.bias"],
                        padding=1, groups=C))
    x = x.flatten(2).transpose(1, 2)
    x = F.linear(x, p[prefix + "pointwise.weight"], p[prefix + "pointwise.bias"])
    D = x.shape[-1]
    return F.layer_norm(x, (D,), p[prefix + "norm.weight"], p[prefix + "norm.bias"], 1e-5)


# ----------------------------------------------------------------------------- whole model
def ngswin_forward(x: Tensor, p: Dict[str, Tensor], depths=(6, 4, 4), num_heads=(6, 4, 4),
                   dec_depth: int = 6, dec_heads: int = 6) -> Tensor:
    """ngswin.py:158-210 — x [B,1,H,W] -> [B,1,H,W].  `p` is a reference-layout state_dict
    (keys without any `main.` prefix)."""
    _, _, H0, W0 = x.shape
    unit = 4 * WINDOW
    padh = (unit - H0 % unit) % unit
    padw = (unit - W0 % unit) % unit
    x = F.pad(x, (0, padw, 0, padh))                                   # zeros right/bottom
    B, _, H, W = x.shape
    if H < 2 * unit or W < 2 * unit:
        raise RuntimeError("padded input must be >= 64 per side (reference domain, SURVEY A.4)")
    shallow = F.conv2d(x, p["shallow_extract.conv1.weight"], p["shallow_extract.conv1.bias"],
                       padding=1)
    c0 = shallow.flatten(2).transpose(1, 2)                            # [B, HW, 64]
    sizes = [(H, W), (H // 2, W // 2), (H // 4, W // 4)]
    e1_, s1 = stage(c0, H, W, p, "encoder_layer1.", depths[0], num_heads[0])
    e1 = patch_merging(s1, H, W, p, "encoder_layer1.downsample.")
    c1 = torch.cat([inter_pool(c0, H, W), e1], dim=-1)
    x2 = F.linear(c1, p["encoder_layer2.across_cascade_proj.weight"],
                  p["encoder_layer2.across_cascade_proj.bias"])
    e2_, s2 = stage(x2, *sizes[1], p, "encoder_layer2.", depths[1], num_heads[1])
    e2 = patch_merging(s2, *sizes[1], p, "encoder_layer2.downsample.")
    c2 = torch.cat([inter_pool(c1, *sizes[1]), e2], dim=-1)
    x3 = F.linear(c2, p["encoder_layer3.across_cascade_proj.weight"],
                  p["encoder_layer3.across_cascade_proj.bias"])
    e3_, _ = stage(x3, *sizes[2], p, "encoder_layer3.", depths[2], num_heads[2])
    bt = bottleneck(shallow, [e1_, e2_, e3_], sizes, p)
    d, _ = stage(bt + e1_, H, W, p, "decoder_layer1.", dec_depth, dec_heads)
    d = F.layer_norm(d, (d.shape[-1],), p["norm.weight"], p["norm.bias"], 1e-5)
    d = d + c0                                                          # global skip, :204
    img = d.transpose(1, 2).reshape(B, -1, H, W)
    img = F.conv2d(img, p["to_target.before_shuffle.weight"], p["to_target.before_shuffle.bias"],
                   padding=1)
    img = F.conv2d(img, p["to_target.to_origin.weight"], p["to_target.to_origin.bias"], padding=1)
    return torch.tanh(img)[:, :, :H0, :W0]


# ----------------------------------------------------------------------------- metrics
def psnr(a: Tensor, b: Tensor) -> float:
    """scripts/run_ablation_studies.py:414-420 — 10*log10(4/mse) on [-1,1] data."""
    mse = torch.mean((a.double() - b.double()) ** 2).item()
    return float("inf") if mse == 0 else 10.0 * math.log10(4.0 / mse)


def ssim(a: Tensor, b: Tensor, data_range: float = 2.0, win: int = 7) -> float:
    """skimage.metrics.structural_similarity semantics used at
    scripts/run_ablation_studies.py:422-439: uniform win x win filter, K1=.01, K2=.03,
    sample covariance (N/(N-1)), mean over the valid interior; a,b are [H,W]."""
    a = a.double()[None, None]
    b = b.double()[None, None]
    k = torch.ones(1, 1, win, win, dtype=torch.float64) / (win * win)
    mu_a, mu_b = F.conv2d(a, k), F.conv2d(b, k)
    n = win * win
    cov = n / (n - 1.0)
    va = cov * (F.conv2d(a * a, k) - mu_a * mu_a)
    vb = cov * (F.conv2d(b * b, k) - mu_b * mu_b)
    vab = cov * (F.conv2d(a * b, k) - mu_a * mu_b)
    c1, c2 = (0.01 * data_range) ** 2, (0.03 * data_range) ** 2
    s = ((2 * mu_a * mu_b + c1) * (2 * vab + c2)) / ((mu_a ** 2 + mu_b ** 2 + c1) * (va + vb + c2))
    return s.mean().item()


# ----------------------------------------------------------------------------- weights
def reference_key_shapes(depths=(6, 4, 4), num_heads=(6, 4, 4), dec_depth=6, dec_heads=6,
                         training_img_size=64):
    """The 554-entry state_dict contract: ordered (name, shape, dtype)."""
    out = []
    def add(n, s, dt=torch.float32):
        out.append((n, tuple(s), dt))
    add("shallow_extract.conv1.weight", (64, 1, 3, 3)); add("shallow_extract.conv1.bias", (64,))
    def blocks(prefix, depth, nh, tsize):
        hd = 64 // nh
        a = hd * nh
        nhd = 32 // nh
        na = nhd * nh
        for i in range(depth):
            b = f"{prefix}blocks.{i}."
            if i % 2 == 1:
                add(b + "attn_mask", ((tsize // 8) ** 2, 64, 64))
            c = b + "ngram_window_partition.ngram_context."
            add(c + "unigram_embed.weight", (32, 2, 8, 8)); add(c + "unigram_embed.bias", (32,))
            add(c + "ngram_attn.logit_scale", (nh, 1, 1))
            add(c + "ngram_attn.relative_position_bias_table", (9, nh))
            add(c + "ngram_attn.relative_position_index", (4, 4), torch.int64)
            add(c + "ngram_attn.qkv.weight", (3 * na, 32)); add(c + "ngram_attn.qkv.bias", (3 * na,))
            add(c + "ngram_attn.proj.weight", (32, na)); add(c + "ngram_attn.proj.bias", (32,))
            add(c + "merge.weight", (64, 64, 1, 1)); add(c + "merge.bias", (64,))
            add(b + "attn.logit_scale", (nh, 1, 1))
            add(b + "attn.relative_position_bias_table", (225, nh))
            add(b + "attn.relative_position_index", (64, 64), torch.int64)
            add(b + "attn.qkv.weight", (3 * a, 64)); add(b + "attn.qkv.bias", (3 * a,))
            add(b + "attn.proj.weight", (64, a)); add(b + "attn.proj.bias", (64,))
            add(b + "norm1.weight", (64,)); add(b + "norm1.bias", (64,))
            add(b + "ffn.fc1.weight", (128, 64)); add(b + "ffn.fc1.bias", (128,))
            add(b + "ffn.fc2.weight", (64, 128)); add(b + "ffn.fc2.bias", (64,))
            add(b + "norm2.weight", (64,)); add(b + "norm2.bias", (64,))
    for s in range(3):
        pre = f"encoder_layer{s + 1}."
        if s > 0:
            add(pre + "across_cascade_proj.weight", (64, 64 * (s + 1)))
            add(pre + "across_cascade_proj.bias", (64,))
        blocks(pre, depths[s], num_heads[s], training_img_size // 2 ** s)
        if s < 2:
            add(pre + "downsample.norm.weight", (256,)); add(pre + "downsample.norm.bias", (256,))
            add(pre + "downsample.reduction.weight", (64, 256))
    add("bottleneck.depthwise.weight", (84, 1, 3, 3)); add("bottleneck.depthwise.bias", (84,))
    add("bottleneck.pointwise.weight", (64, 84)); add("bottleneck.pointwise.bias", (64,))
    add("bottleneck.norm.weight", (64,)); add("bottleneck.norm.bias", (64,))
    blocks("decoder_layer1.", dec_depth, dec_heads, training_img_size // 2)
    add("norm.weight", (64,)); add("norm.bias", (64,))
    add("to_target.before_shuffle.weight", (1, 64, 3, 3)); add("to_target.before_shuffle.bias", (1,))
    add("to_target.to_origin.weight", (1, 1, 3, 3)); add("to_target.to_origin.bias", (1,))
    return out


def synthetic_state_dict(kind: str = "W1", seed: int = 0, training_img_size: int = 64):
    """Seeded synthetic weights in the reference layout (SURVEY.md §8c 'Weights').

    W1: what a freshly built + `weights_init`-ed generator holds
        (training/train_combined.py:207-217): Conv/Linear weights N(0,.02), zero biases,
        LayerNorm (1,0); logit_scale = ln 10 (attention.py:48); bias tables ~ N(0,.02)
        (trunc-normal in the reference; the truncation at ±2 is >±100σ and never bites).
    W2: stress set — matrices x5, biases N(0,.05), LayerNorm weight 1+N(0,.1), bias N(0,.05),
        logit_scale ~ ln U(5,100)·(some >ln 100, exercising the clamp), tables N(0,.5)."""
    g = torch.Generator().manual_seed(seed)
    sd = {}
    for name, shape, dt in reference_key_shapes(training_img_size=training_img_size):
        if name.endswith("relative_position_index"):
            n = int(math.isqrt(shape[0]))
            sd[name] = relative_position_index(n, n)
        elif name.endswith("attn_mask"):
            side = int(math.isqrt(shape[0])) * 8
            sd[name] = make_mask(side, side)
        elif name.endswith("logit_scale"):
            if kind == "W1":
                sd[name] = torch.full(shape, math.log(10.0))
            else:
                u = torch.rand(shape, generator=g) * 115.0 + 5.0     # U(5,120): some > 100
                sd[name] = u.log()
        elif name.endswith("relative_position_bias_table"):
            sd[name] = torch.randn(shape, generator=g) * (0.02 if kind == "W1" else 0.5)
        elif ".norm" in name or name.startswith("norm."):
            is_w = name.endswith("weight")
            if kind == "W1":
                sd[name] = torch.ones(shape) if is_w else torch.zeros(shape)
            else:
                r = torch.randn(shape, generator=g)
                sd[name] = 1.0 + 0.1 * r if is_w else 0.05 * r
        elif name.endswith("bias"):
            sd[name] = torch.zeros(shape) if kind == "W1" else torch.randn(shape, generator=g) * 0.05
        else:
            sd[name] = torch.randn(shape, generator=g) * (0.02 if kind == "W1" else 0.1)
    return sd


def synthetic_input(B: int, H: int, W: int, seed: int = 999) -> Tensor:
    """SURVEY.md §8d — uniform [-1,1) single-channel slices (data range of data/datasets.py:107-109)."""
    g = torch.Generator().manual_seed(seed)
    return torch.rand(B, 1, H, W, generator=g) * 2.0 - 1.0
