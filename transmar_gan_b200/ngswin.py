"""Drop-in `NGswin` generator whose N-gram window-attention blocks run on hand-written
sm_100a CUDA kernels (libb200mar.so).

Boundary contract: same constructor signature as the reference
`models/generator/ngswin.py:43-49`, the same 554 `state_dict` entries (names, shapes, dtypes,
order — parameters live in real nn.Linear / nn.Conv2d / nn.LayerNorm submodules carrying the
reference attribute names so `netG.apply(weights_init)` and `load_state_dict(strict=True)`
behave identically), and the same `forward(x: [B,1,H,W]) -> [B,1,H,W]`.

What runs where
  * CUDA kernels (this repo): unigram conv, n-gram context attention, fused shifted-window
    partition + cosine attention + LN1 + residual, fused MLP + LN2 + residual(+within-stage
    sum), PatchMerging, pixel-shuffle + concat — forward and backward.
  * the glue between the stages (shallow conv, InterPool + cascade projections, SCDPBottleneck, final LayerNorm +
    skip + Reconstruction): fused channels-last kernels too, forward (glue_fwd.cu, *_tc.cu) and backward
    (glue_bwd.cu); the stock nn.Conv2d / nn.Linear / nn.LayerNorm submodules only hold the parameters.
There is NO CPU forward: calling the module on CPU tensors raises.
"""
from __future__ import annotations

import math
from itertools import repeat
from typing import Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import ops

WINDOW = 8


def _pair(v) -> Tuple[int, int]:
    return tuple(v) if isinstance(v, (tuple, list)) else (v, v)


def _ntuple(v, n):
    return tuple(v) if isinstance(v, (tuple, list)) else tuple(repeat(v, n))


def relative_position_index(win: int) -> torch.Tensor:
    """[win^2, win^2] int64 buffer, idx[p,q] = (p_i-q_i+win-1)*(2win-1) + (p_j-q_j+win-1)
    (reference: components/attention.py:9-18)."""
    t = torch.arange(win * win)
    i, j = t // win, t % win
    return ((i[:, None] - i[None, :] + win - 1) * (2 * win - 1)
            + (j[:, None] - j[None, :] + win - 1)).to(torch.int64)


def shift_attn_mask(ph: int, pw: int, window: int = WINDOW, shift: int = WINDOW // 2) -> torch.Tensor:
    """[nW, 64, 64] fp32 buffer (0 / -100) kept only for state_dict parity — the kernels
    evaluate the same 9-region rule analytically (reference: transformer_blocks.py:63-81)."""
    def ids(n):
        a = torch.arange(n)
        return (a >= n - window).to(torch.int64) + (a >= n - shift).to(torch.int64)
    rid = ids(ph)[:, None] * 3 + ids(pw)[None, :]
    wh, ww = ph // window, pw // window
    w = rid.reshape(wh, window, ww, window).permute(0, 2, 1, 3).reshape(wh * ww, window * window)
    return ((w[:, None, :] != w[:, :, None]).to(torch.float32)) * -100.0


# ------------------------------------------------------------------------------ containers
class WindowAttention(nn.Module):
    """Parameter container of a scaled-cosine window attention (attention.py:40-62)."""

    def __init__(self, dim, num_heads, window_size, head_dim=None, qkv_bias=True):
        super().__init__()
        self.dim, self.num_heads, self.window = dim, num_heads, window_size
        head_dim = head_dim or dim // num_heads
        self.attn_dim = head_dim * num_heads
        self.logit_scale = nn.Parameter(torch.log(10 * torch.ones((num_heads, 1, 1))))
        self.relative_position_bias_table = nn.Parameter(
            torch.zeros((2 * window_size - 1) ** 2, num_heads))
        self.register_buffer("relative_position_index", relative_position_index(window_size))
        self.qkv = nn.Linear(dim, self.attn_dim * 3, bias=qkv_bias)
        self.proj = nn.Linear(self.attn_dim, dim)
        nn.init.trunc_normal_(self.relative_position_bias_table, std=.02)


class NGramContext(nn.Module):
    def __init__(self, dim, window_size, ngram, num_heads):
        super().__init__()
        self.unigram_embed = nn.Conv2d(dim, dim // 2, kernel_size=window_size, stride=window_size,
                                       padding=0, groups=dim // 2)
        self.ngram_attn = WindowAttention(dim // 2, num_heads, ngram)
        self.merge = nn.Conv2d(dim, dim, 1, 1, 0)


class NGramWindowPartition(nn.Module):
    def __init__(self, dim, window_size, ngram, num_heads, shift_size):
        super().__init__()
        self.shift_size = shift_size
        self.ngram_context = NGramContext(dim, window_size, ngram, num_heads)


class Mlp(nn.Module):
    def __init__(self, dim, hidden):
        super().__init__()
        self.fc1 = nn.Linear(dim, hidden)
        self.fc2 = nn.Linear(hidden, dim)


class NSTB(nn.Module):
    """N-Gram Swin Transformer Block; forward is two+two CUDA kernels (ops.NSTBFunction)."""

    def __init__(self, dim, training_num_patches, ngram, num_heads, window_size, shift_size,
                 head_dim=None, mlp_ratio=2., qkv_bias=True):
        super().__init__()
        self.num_heads, self.shift_size = num_heads, shift_size
        # "tf32" / "fp32" per half, assigned by NGswin's precision plan; None = the selected mode
        self._mode_attn = self._mode_mlp = None
        self.ngram_window_partition = NGramWindowPartition(dim, window_size, ngram, num_heads,
                                                           shift_size)
        self.attn = WindowAttention(dim, num_heads, window_size, head_dim, qkv_bias)
        mask = shift_attn_mask(*training_num_patches, window_size, shift_size) if shift_size > 0 else None
        self.register_buffer("attn_mask", mask)
        self.norm1 = nn.LayerNorm(dim)
        self.ffn = Mlp(dim, int(dim * mlp_ratio))
        self.norm2 = nn.LayerNorm(dim)

    def kernel_params(self):
        """The block's 24 parameter tensors in ops.NSTB_PARAM_NAMES order."""
        c, a, f = self.ngram_window_partition.ngram_context, self.attn, self.ffn
        g = c.ngram_attn
        return (c.unigram_embed.weight, c.unigram_embed.bias, g.qkv.weight, g.qkv.bias,
                g.proj.weight, g.proj.bias, g.logit_scale, g.relative_position_bias_table,
                c.merge.weight, c.merge.bias,
                a.qkv.weight, a.qkv.bias, a.proj.weight, a.proj.bias, a.logit_scale,
                a.relative_position_bias_table, self.norm1.weight, self.norm1.bias,
                f.fc1.weight, f.fc1.bias, f.fc2.weight, f.fc2.bias,
                self.norm2.weight, self.norm2.bias)

    def forward(self, z, num_patches, want_out=True, want_zsum=True):
        ph, pw = num_patches
        out, zsum = ops.NSTBFunction.apply(z, ph, pw, self.num_heads, self.shift_size, want_out,
                                           want_zsum, torch.is_grad_enabled(), *self.kernel_params())
        return out, zsum


class PatchMerging(nn.Module):
    def __init__(self, dim, downsample_dim=None):
        super().__init__()
        self.norm = nn.LayerNorm(4 * dim)
        self.reduction = nn.Linear(4 * dim, downsample_dim or dim, bias=False)

    def forward(self, x, num_patches):
        ph, pw = num_patches
        if not torch.is_grad_enabled() and ops._tc():
            y = ops.patch_merge_tc_fwd(x, self.norm.weight.detach(), self.norm.bias.detach(),
                                       self.reduction.weight.detach(), ph, pw)
            return y, (ph // 2, pw // 2)
        y = ops.PatchMergeFunction.apply(x, ph, pw, self.norm.weight, self.norm.bias,
                                         self.reduction.weight)
        return y, (ph // 2, pw // 2)


class _Stage(nn.Module):
    """Shared body of EncoderLayer / DecoderLayer (encoder_decoder.py:135-145, :252-258)."""

    def _build_blocks(self, dim, training_num_patches, ngram, depth, num_heads, window_size,
                      head_dim, mlp_ratio, qkv_bias):
        self.blocks = nn.Sequential(*[
            NSTB(dim, training_num_patches, ngram, num_heads, window_size,
                 0 if i % 2 == 0 else window_size // 2, head_dim, mlp_ratio, qkv_bias)
            for i in range(depth)])

    def _run_blocks(self, x, num_patches, need_sum_last, owned=False):
        """Block i consumes z_i = out_{i-1} + z_{i-1}; each block's MLP kernel emits that sum
        for its successor, so intermediate `out` tensors are never written."""
        z, n = x, len(self.blocks)
        out = zsum = None
        private = owned                  # may z be updated in place?
        for i, blk in enumerate(self.blocks):
            last = i == n - 1
            want_zsum = (not last) or need_sum_last
            with ops.unit_precision(blk._mode_attn, blk._mode_mlp):
                ea, em = ops.unit_modes()
                if not torch.is_grad_enabled() and ea == "bf16" and em == "bf16":
                    # precision mode "bf16": the stream lives in BF16 inside the stage (cast once at stage entry,
                    # which also makes it private; cast back once at stage exit)
                    if z.dtype != torch.bfloat16:
                        z, private = ops.cast_stream(z.contiguous(), torch.bfloat16), True
                    elif want_zsum and not private:
                        z = z.clone()
                    out, zsum = ops.nstb_infer_s16_(z, *num_patches, blk.num_heads, blk.shift_size, last,
                                                    want_zsum, blk.kernel_params())
                else:
                    if z.dtype != torch.float32:
                        z, private = ops.cast_stream(z, torch.float32), True
                    if torch.is_grad_enabled() or (ea == "fp32" and em == "fp32"):
                        out, zsum = blk(z, num_patches, want_out=last, want_zsum=want_zsum)
                    elif ea != "tf32" or em != "tf32":
                        out, zsum = ops.nstb_infer_mixed_(z, *num_patches, blk.num_heads, blk.shift_size, last,
                                                          want_zsum, blk.kernel_params(), ea, em, private)
                    else:
                        # the block adds its output into z (TMA reduce-add), so z must be private to this
                        # loop: the stage input is copied once, later blocks own the previous block's sum
                        if want_zsum and not private:
                            z = z.clone()
                        out, zsum = ops.nstb_infer_(z, *num_patches, blk.num_heads, blk.shift_size, last,
                                                    want_zsum, blk.kernel_params())
            private = True               # zsum is this loop's own tensor (fresh, or the private z updated)
            z = zsum
        if out is not None and out.dtype != torch.float32:
            out = ops.cast_stream(out, torch.float32)
        if zsum is not None and zsum.dtype != torch.float32:
            zsum = ops.cast_stream(zsum, torch.float32) if need_sum_last else zsum.new_empty(0, dtype=torch.float32)
        return out, zsum


class EncoderLayer(_Stage):
    def __init__(self, dim, training_num_patches, ngram, depth, num_heads, window_size,
                 head_dim=None, mlp_ratio=2., qkv_bias=True, downsample=False, num_cas=1):
        super().__init__()
        self.across_cascade_proj = nn.Linear(dim * num_cas, dim) if num_cas != 1 else nn.Identity()
        self._build_blocks(dim, training_num_patches, ngram, depth, num_heads, window_size,
                           head_dim, mlp_ratio, qkv_bias)
        self.downsample = PatchMerging(dim, dim) if downsample else None

    def forward(self, x, num_patches, projected=False):
        if not projected:                                       # fused path projects beforehand
            x = self.across_cascade_proj(x).contiguous()
        # a stage input this layer projected itself (or that the fused path produced for it alone) may be
        # updated in place; encoder_layer1's input is the shallow feature map, which is reused later
        out, zsum = self._run_blocks(x, num_patches, need_sum_last=self.downsample is not None,
                                     owned=projected or not isinstance(self.across_cascade_proj, nn.Identity))
        if self.downsample is None:
            return out, out, num_patches
        down, np2 = self.downsample(zsum, num_patches)
        return out, down, np2


class DecoderLayer(_Stage):
    def __init__(self, dim, training_num_patches, ngram, depth, num_heads, window_size,
                 head_dim=None, mlp_ratio=2., qkv_bias=True):
        super().__init__()
        self._build_blocks(dim, training_num_patches, ngram, depth, num_heads, window_size,
                           head_dim, mlp_ratio, qkv_bias)

    def forward(self, x, num_patches, owned=False):
        out, _ = self._run_blocks(x.contiguous(), num_patches, need_sum_last=False, owned=owned)
        return out


class ShallowExtractor(nn.Module):
    def __init__(self, in_chans, out_chans):
        super().__init__()
        self.conv1 = nn.Conv2d(in_chans, out_chans, 3, 1, 1)

    def forward(self, x):
        return self.conv1(x)


class SCDPBottleneck(nn.Module):
    """pixel-Shuffle, Concat, Depth-wise conv, Point-wise projection (encoder_decoder.py:166-200)."""

    def __init__(self, num_encoder_stages, enc_dim, dec_dim):
        super().__init__()
        self.concat_dim = sum(4 ** i for i in range(num_encoder_stages)) * (enc_dim // 16)
        self.depthwise = nn.Conv2d(self.concat_dim, self.concat_dim, 3, 1, 1, groups=self.concat_dim)
        self.pointwise = nn.Linear(self.concat_dim, dec_dim)
        self.norm = nn.LayerNorm(dec_dim)

    def forward(self, pooled, feats, H, W):
        """pooled = (c0, maxpool2(c0), maxpool4(c0)) TOKEN-MAJOR [B, T_i, 64] (the cascade path computes them anyway;
        BottleneckPool of the reference pools the NCHW shallow map, pooling.py:32-37 - same values), feats =
        (e1_, e2_, e3_).  Everything stays channels-last: the 84-channel tensor is handed to the depthwise
        convolution as a channels_last view (no NCHW copies, no layout-conversion kernels)."""
        xs = [(f + F.leaky_relu(s, 0.01)).contiguous() for s, f in zip(pooled, feats)]
        cat = ops.PixelShuffleCatFunction.apply(xs[0], xs[1], xs[2], H, W)   # [B, HW, 84]
        B = cat.shape[0]
        x = cat.view(B, H, W, self.concat_dim).permute(0, 3, 1, 2)           # NCHW view of NHWC memory
        x = F.gelu(self.depthwise(x))
        x = x.permute(0, 2, 3, 1).reshape(B, H * W, self.concat_dim)         # a view when the conv kept NHWC
        return ops.LayerNorm64Function.apply(self.pointwise(x), self.norm.weight, self.norm.bias)


class Reconstruction(nn.Sequential):
    def __init__(self, dim, out_chans, upscale):
        super().__init__()
        self.add_module("before_shuffle", nn.Conv2d(dim, out_chans * upscale ** 2, 3, 1, 1))
        self.add_module("to_origin", nn.Conv2d(out_chans, out_chans, 3, 1, 1))
        self.add_module("tanh", nn.Tanh())


# ------------------------------------------------------------------------------ the model
class NGswin(nn.Module):
    """B200-native drop-in for `models.generator.ngswin.NGswin` (reference ngswin.py:18-210)."""

    def __init__(self, training_img_size=64, ngrams=(2, 2, 2, 2), in_chans=1, embed_dim=64,
                 depths=(6, 4, 4), num_heads=(6, 4, 4), head_dim=None, dec_dim=64, dec_depths=6,
                 dec_num_heads=6, dec_head_dim=None, target_mode='light_x1', window_size=8,
                 mlp_ratio=2., qkv_bias=True, img_norm=True, drop_rate=0.0, attn_drop_rate=0.0,
                 drop_path_rate=0.0, act_layer=nn.GELU, norm_layer=nn.LayerNorm, weight_init='',
                 **kwargs):
        assert target_mode in ['light_x1', 'light_x3', 'light_x4'], \
            "'target mode' should be one of ['light_x1', 'light_x3', 'light_x4']"
        super().__init__()
        # The fused kernels implement the one configuration every caller of the reference
        # instantiates (`NGswin()`, training/train_combined.py:226); anything else is refused
        # loudly instead of silently computing something different.
        unsupported = []
        if embed_dim != 64 or dec_dim != 64: unsupported.append("embed_dim/dec_dim != 64")
        if window_size != 8: unsupported.append("window_size != 8")
        if tuple(ngrams) != (2, 2, 2, 2): unsupported.append("ngrams != (2,2,2,2)")
        if len(depths) != 3 or len(num_heads) != 3: unsupported.append("other than 3 encoder stages")
        if any(h not in (4, 6) for h in tuple(num_heads) + (dec_num_heads,)):
            unsupported.append("head counts other than 4 or 6")
        if head_dim is not None or dec_head_dim is not None: unsupported.append("explicit head_dim")
        if float(mlp_ratio) != 2.0: unsupported.append("mlp_ratio != 2")
        if not qkv_bias: unsupported.append("qkv_bias=False")
        if drop_rate or attn_drop_rate or drop_path_rate: unsupported.append("non-zero drop rates")
        if act_layer is not nn.GELU or norm_layer is not nn.LayerNorm:
            unsupported.append("act_layer/norm_layer other than GELU/LayerNorm")
        if target_mode != 'light_x1' or in_chans != 1: unsupported.append("target_mode/in_chans")
        if unsupported:
            raise NotImplementedError("transmar_gan_b200.NGswin: unsupported configuration: "
                                      + "; ".join(unsupported))
        self.training_img_size = _pair(training_img_size)
        self.ngrams, self.embed_dim, self.depths, self.num_heads = ngrams, embed_dim, depths, num_heads
        self.dec_dim, self.dec_depths, self.dec_num_heads = dec_dim, (dec_depths,), (dec_num_heads,)
        self.window_size = window_size
        self.num_encoder_stages, self.num_decoder_stages = len(depths), 1
        self.scale = int(target_mode[-1])
        self.img_norm = False                                   # reference ngswin.py:79
        ts = self.training_img_size
        self.shallow_extract = ShallowExtractor(in_chans, embed_dim)
        for i in range(3):
            self.add_module(f'encoder_layer{i + 1}', EncoderLayer(
                embed_dim, (ts[0] // 2 ** i, ts[1] // 2 ** i), ngrams[i], depths[i], num_heads[i],
                window_size, None, 2., qkv_bias, downsample=(i != 2), num_cas=i + 1))
        self.bottleneck = SCDPBottleneck(3, embed_dim, dec_dim)
        self.add_module('decoder_layer1', DecoderLayer(
            dec_dim, (ts[0] // 2, ts[1] // 2), ngrams[3], dec_depths, dec_num_heads, window_size,
            None, 2., qkv_bias))
        self.norm = nn.LayerNorm(dec_dim)
        self.to_target = Reconstruction(dec_dim, in_chans, self.scale)
        self.apply(self._init_weights)

    @staticmethod
    def _init_weights(m):
        # Swin-V2 manner (reference ngswin.py:140-148)
        if isinstance(m, nn.Linear):
            nn.init.trunc_normal_(m.weight, std=.02)
            if m.bias is not None:
                nn.init.constant_(m.bias, 0)
        elif isinstance(m, nn.LayerNorm):
            nn.init.constant_(m.bias, 0)
            nn.init.constant_(m.weight, 1.0)

    @torch.jit.ignore
    def no_weight_decay(self):
        return {n for n, _ in self.named_parameters() if 'relative_position_bias_table' in n}

    # ---------------------------------------------------------------- precision plan ("auto" mode)
    def _units(self):
        """The 20 NSTBs in execution order.  A precision plan covers 41 units: (attention half, MLP half) of
        every block, then the stage glue (PatchMerging, cascade projections, bottleneck) as one unit."""
        return [b for st in (self.encoder_layer1, self.encoder_layer2, self.encoder_layer3, self.decoder_layer1)
                for b in st.blocks]

    def _weights_key(self):
        ps = list(self.parameters())
        return (sum(p._version for p in ps), ps[0].data_ptr(), str(ps[0].device))

    def _set_plan(self, plan):
        blocks = self._units()
        for i, b in enumerate(blocks):
            b._mode_attn = plan[2 * i] if plan else None
            b._mode_mlp = plan[2 * i + 1] if plan else None
        self._glue_mode = plan[2 * len(blocks)] if plan else None

    def _get_plan(self):
        return [m for b in self._units() for m in (b._mode_attn, b._mode_mlp)] + [getattr(self, "_glue_mode", None)]

    @torch.no_grad()
    def calibrate_precision(self, tol: float = 4e-4, size: int = 96, seed: int = 4242, x: torch.Tensor = None):
        """Measure, on four seeded probe slices (one uniform noise, three structured: flat regions, soft ellipses,
        saturated discs, radial streaks - piecewise-constant content is what 11-bit operands are worst at, since
        LayerNorm / F.normalize renormalise rows of almost identical values) and, if given, on the caller's own
        representative slices `x` [B,1,H,W] (recommended for real data), which units may run on the tensor cores (FP16 / TF32 operands)
        without moving the model output by more than `tol` (default 0.4 of the 1e-3 parity gate, the margin
        covering other inputs and larger images) relative to the all-fp32 kernels, and store that plan:
        it is what precision "auto" executes.  A block half that fails at 11-bit operands is tried at the
        compensated level "tf32c" (every GEMM of the half on operand pairs, three MMA passes, still tcgen05,
        fp32-level error) before it falls back to the fp32 FMA-pipe kernel.  All units pass on init-scale weights;
        trained-like weights need "tf32c" for the first block's attention; the W2 stress set (x100 cosine logits)
        needs it for its leading blocks.  Returns the report dict (also `.precision_report`)."""
        p0 = next(self.parameters())
        if not p0.is_cuda:
            raise RuntimeError("calibrate_precision: move the module to a B200 first (there is no CPU path)")
        probe = f"{'caller data ' + str(tuple(x.shape)) + ' + ' if x is not None else ''}4x1x{size}x{size} " \
                f"(uniform[-1,1) + 3 structured phantoms), seed {seed}"
        xs = [_probe_slices(size, seed).to(p0.device)]
        if x is not None:                                  # representative slices of the caller's own data
            xs.append(x.to(p0.device, torch.float32))
        n = 2 * len(self._units()) + 1
        saved = self._get_plan()

        def run(plan):
            self._set_plan(plan)
            return [self._forward_planned(xi) for xi in xs]

        def dist(ys, refs):
            return max((y - r).abs().max().item() for y, r in zip(ys, refs))

        n_blk = 2 * len(self._units())                      # units below this are block halves (the last is the glue)

        try:
            ref = run(["fp32"] * n)
            err_all = dist(run(["tf32"] * n), ref)
            marginal = marginal_c = None
            if err_all <= tol:
                plan, err = ["tf32"] * n, err_all
            else:
                def alone(u, level):
                    trial = ["fp32"] * n
                    trial[u] = level
                    return dist(run(trial), ref)

                marginal = [alone(u, "tf32") for u in range(n)]
                # block halves that fail at 11-bit operands get a second chance at the compensated level
                marginal_c = [alone(u, "tf32c") if (u < n_blk and marginal[u] > tol) else None for u in range(n)]
                level = ["tf32" if marginal[u] <= tol else ("tf32c" if marginal_c[u] is not None else None)
                         for u in range(n)]
                cost = [marginal[u] if level[u] == "tf32" else marginal_c[u] for u in range(n)]
                plan, err = ["fp32"] * n, 0.0
                for u in sorted((u for u in range(n) if level[u]), key=lambda i: cost[i]):
                    if cost[u] > tol:
                        break
                    for lv in ([level[u]] if level[u] == "tf32c" or u >= n_blk else ["tf32", "tf32c"]):
                        trial = list(plan)
                        trial[u] = lv
                        e = dist(run(trial), ref)
                        if e <= tol:
                            plan, err = trial, e
                            break
        finally:
            self._set_plan(saved)
        self._auto_plan, self._auto_key, self._auto_age = plan, self._weights_key(), 0
        self.precision_report = {"tol": tol, "probe": probe,
                                 "err_all_tensor_core": err_all, "err_plan": err,
                                 "units_on_tensor_cores": sum(m != "fp32" for m in plan),
                                 "units_compensated": sum(m == "tf32c" for m in plan), "units": n,
                                 "plan": list(plan), "marginal_err": marginal, "marginal_err_compensated": marginal_c}
        return self.precision_report

    #: in train() mode the weights change every step: the plan is re-measured every this many forwards
    AUTO_RECALIBRATE_EVERY = 500

    def _auto_refresh(self):
        capturing = torch.cuda.is_current_stream_capturing()
        plan = getattr(self, "_auto_plan", None)
        if plan is None:
            if capturing:
                raise RuntimeError("precision 'auto': call model.calibrate_precision() before CUDA-graph capture")
            self.calibrate_precision()
        elif not capturing:
            self._auto_age += 1
            if self.training:        # weights move a little every step: re-measure periodically
                stale = self._auto_age >= self.AUTO_RECALIBRATE_EVERY
            else:                    # eval: re-measure whenever the weights were modified or moved
                stale = self._weights_key() != self._auto_key
            if stale:
                self.calibrate_precision()
        return self._auto_plan

    def forward(self, x):
        if not x.is_cuda:
            raise RuntimeError("transmar_gan_b200.NGswin has no CPU path; move the module and the "
                               "input to a B200 (`.cuda()`)")
        if ops.get_precision() == "auto" and not ops._SCOPE:
            self._set_plan(self._auto_refresh())
        else:
            self._set_plan(None)
        return self._forward_planned(x)

    def _forward_planned(self, x):
        with ops.precision_scope(getattr(self, "_glue_mode", None)):
            return self._forward_impl(x)

    def _forward_impl(self, x):
        _, _, H0, W0 = x.shape
        unit = 4 * self.window_size
        padh, padw = (unit - H0 % unit) % unit, (unit - W0 % unit) % unit
        if padh or padw:
            x = F.pad(x, (0, padw, 0, padh))                     # zeros right / bottom
        B, _, H, W = x.shape
        if H < 2 * unit or W < 2 * unit:
            raise RuntimeError(f"input {H0}x{W0} pads to {H}x{W}; NGswin needs >= 64 per side "
                               "(stage 3 must have 2x2 windows; the reference raises too)")
        if not torch.is_grad_enabled():
            return self._forward_fused(x.float().contiguous(), H0, W0)
        return self._forward_train(x.float().contiguous(), H0, W0)

    def _forward_train(self, x, H0, W0):
        """Training path (autograd on): the same fused channels-last glue kernels as the inference path, each wrapped
        in an autograd Function with a hand-written backward (glue_bwd.cu) - no stock PyTorch / cuDNN op is left
        between the input image and the output image."""
        B, _, H, W = x.shape
        se = self.shallow_extract.conv1
        l2, l3, bn, tt = self.encoder_layer2, self.encoder_layer3, self.bottleneck, self.to_target
        c0 = ops.ShallowTokFunction.apply(x, se.weight, se.bias)
        e1_, e1, np1 = self.encoder_layer1(c0, (H, W))
        x2, p0 = ops.CascadeProjFunction.apply(c0, None, e1, l2.across_cascade_proj.weight,
                                               l2.across_cascade_proj.bias, H // 2, W // 2)
        e2_, e2, np2 = l2(x2, np1, projected=True)
        x3, p00 = ops.CascadeProjFunction.apply(p0, e1, e2, l3.across_cascade_proj.weight,
                                                l3.across_cascade_proj.bias, H // 4, W // 4)
        e3_, _, _ = l3(x3, np2, projected=True)
        dec_in = ops.BottleneckFunction.apply(c0, p0, p00, e1_, e2_, e3_, bn.depthwise.weight, bn.depthwise.bias,
                                              bn.pointwise.weight, bn.pointwise.bias, bn.norm.weight, bn.norm.bias,
                                              H, W)
        dec = self.decoder_layer1(dec_in, (H, W))
        return ops.FinalReconFunction.apply(dec, c0, self.norm.weight, self.norm.bias, tt.before_shuffle.weight,
                                            tt.before_shuffle.bias, tt.to_origin.weight, tt.to_origin.bias,
                                            H, W, H0, W0)

    def _forward_fused(self, x, H0, W0):
        """Inference path (no autograd): the glue between the stages runs as fused channels-last
        CUDA kernels (glue_fwd.cu) - same arithmetic as the composite path above, which training
        keeps so that autograd differentiates it."""
        B, _, H, W = x.shape
        se = self.shallow_extract.conv1
        c0 = ops.shallow_tok_fwd(x, se.weight, se.bias)
        e1_, e1, np1 = self.encoder_layer1(c0, (H, W))
        l2, l3, bn, tt = self.encoder_layer2, self.encoder_layer3, self.bottleneck, self.to_target
        x2, p0 = ops.cascade_proj_fwd(c0, None, e1, l2.across_cascade_proj.weight,
                                      l2.across_cascade_proj.bias, H // 2, W // 2)
        e2_, e2, np2 = l2(x2, np1, projected=True)
        x3, p00 = ops.cascade_proj_fwd(p0, e1, e2, l3.across_cascade_proj.weight,
                                       l3.across_cascade_proj.bias, H // 4, W // 4)
        e3_, _, _ = l3(x3, np2, projected=True)
        dec_in = ops.bottleneck_fwd(c0, p0, p00, e1_, e2_, e3_, bn.depthwise.weight, bn.depthwise.bias,
                                    bn.pointwise.weight, bn.pointwise.bias, bn.norm.weight,
                                    bn.norm.bias, H, W)
        dec = self.decoder_layer1(dec_in, (H, W), owned=True)      # dec_in is not used again
        return ops.final_recon_fwd(dec, c0, self.norm.weight, self.norm.bias,
                                   tt.before_shuffle.weight, tt.before_shuffle.bias,
                                   tt.to_origin.weight, tt.to_origin.bias, H, W, H0, W0)

    def flops(self, resolutions):
        """The reference's analytic count, reproduced term by term (models/generator/ngswin.py:212-238 and the
        `flops()` methods it sums: attention.py:103-115, window_utils.py:92-104, transformer_blocks.py:122-138,
        encoder_decoder.py:29-33,76-83,147-157,202-213,260-266, pooling.py:19-23,39-49, reconstruction.py:21-29).
        It counts MACs, adds LayerNorm / bias / activation terms, and over-counts the 6-head attention by taking
        A = dim; `flops_exact()` is the exact MAC count the roofline uses."""
        unit = 4 * self.window_size
        D = self.embed_dim
        H, W = resolutions
        H += (unit - H % unit) % unit
        W += (unit - W % unit) % unit

        def attn(dim, nh, area, nwin):
            f = area * dim * 3 * dim + 3 * dim
            f += 2 * nh * area * (dim // nh) * area
            f += area * dim * dim + dim
            return f * nwin

        def nstb(h, w, nh):
            wh, ww = h // 8, w // 8
            f = wh * ww * 64 * D + wh * ww * D                      # unigram embed
            f += 2 * attn(D // 2, nh, 4, wh * ww) + wh * ww * 4 * D + wh * ww * D * D
            f += attn(D, nh, 64, h // 8 * w // 8)                   # (S)W-MSA
            f += h * w * D                                          # norm1
            f = int(f + h * w * D * 2.0 * D + 2.0 * D + h * w * 2.0 * D * D + D)
            return f + h * w * D                                    # norm2

        def merge(h, w):
            return h * w * 4 * D + (h // 2) * (w // 2) * 4 * D * D + D

        f = H * W * 9 * D + H * W * D                                # shallow
        f += sum(nstb(H, W, self.num_heads[0]) for _ in range(self.depths[0])) + merge(H, W)
        f += (H // 2) * (W // 2) * 4 * D                             # inter_pool
        h2, w2 = H // 2, W // 2
        f += h2 * w2 * 2 * D * D + D + sum(nstb(h2, w2, self.num_heads[1]) for _ in range(self.depths[1]))
        f += merge(h2, w2)
        f += (h2 // 2) * (w2 // 2) * 4 * 2 * D
        h3, w3 = H // 4, W // 4
        f += h3 * w3 * 3 * D * D + D + sum(nstb(h3, w3, self.num_heads[2]) for _ in range(self.depths[2]))
        for i in range(3):                                           # bottleneck pools (pooling.py:39-49)
            hr, wr = H // 2, W // 2
            for _ in range(i):
                f += hr * wr * 4 * D
                hr, wr = hr // 2, wr // 2
            f += (H // 2 ** i) * (W // 2 ** i) * D
        cd = self.bottleneck.concat_dim
        f += H * W * 9 * cd + 2 * H * W * cd + H * W * cd * self.dec_dim + self.dec_dim
        f += sum(nstb(H, W, self.dec_num_heads[0]) for _ in range(self.dec_depths[0]))
        f += H * W * D                                               # norm
        f += H * W * 9 * self.dec_dim + H * W + H * W * 9 + H * W    # reconstruction
        return f

    def flops_exact(self, resolutions):
        """Exact multiply-accumulate count of one forward, unpadded head dims."""
        unit = 4 * self.window_size
        H, W = resolutions
        H += (unit - H % unit) % unit
        W += (unit - W % unit) % unit
        return H * W * 548013


def _probe_slices(size: int, seed: int, n_structured: int = 3) -> torch.Tensor:
    """[1 + n_structured, 1, size, size] calibration input in [-1,1]: slice 0 uniform noise, the others CT-like
    phantoms (air background, randomly placed constant / softly shaded ellipses, one or two saturated "metal" discs
    with radial streaks, weak noise): piecewise-constant content is what 11-bit operands are worst at.
    Deterministic (seeded CPU generator)."""
    g = torch.Generator().manual_seed(seed)
    out = [torch.rand(1, size, size, generator=g) * 2 - 1]
    yy, xx = torch.meshgrid(torch.linspace(-1, 1, size), torch.linspace(-1, 1, size), indexing="ij")
    for _ in range(n_structured):
        img = torch.full((size, size), -1.0)
        for _ in range(int(torch.randint(3, 7, (1,), generator=g))):
            cx, cy = (torch.rand(2, generator=g) * 1.2 - 0.6).tolist()
            ax, ay = (torch.rand(2, generator=g) * 0.5 + 0.15).tolist()
            th, val, shade = (torch.rand(3, generator=g) * torch.tensor([math.pi, 1.2, 0.15])).tolist()
            xr = (xx - cx) * math.cos(th) + (yy - cy) * math.sin(th)
            yr = -(xx - cx) * math.sin(th) + (yy - cy) * math.cos(th)
            d = (xr / ax) ** 2 + (yr / ay) ** 2
            img = torch.where(d < 1.0, (val - 0.7) + shade * (1 - d), img)
        art = torch.zeros(size, size)
        metal = torch.zeros(size, size, dtype=torch.bool)
        for _ in range(int(torch.randint(1, 3, (1,), generator=g))):
            mx, my = (torch.rand(2, generator=g) - 0.5).tolist()
            r = ((xx - mx) ** 2 + (yy - my) ** 2).sqrt()
            metal |= r < 7.0 / size
            k, ph = int(torch.randint(6, 14, (1,), generator=g)), float(torch.rand(1, generator=g)) * 2 * math.pi
            art += 0.35 * torch.cos(k * torch.atan2(yy - my, xx - mx) + ph) * torch.exp(-1.5 * r)
        img = torch.where(metal, torch.ones_like(img), img)
        img = (img + art + 0.03 * torch.randn(size, size, generator=g)).clamp(-1, 1)
        img = torch.where(metal, torch.ones_like(img), img)
        out.append(img[None])
    return torch.stack(out, 0)


def _token_maxpool(x, ph, pw):
    """InterPool (pooling.py:13-17): 2x2 max-pool on the token grid, channels last."""
    B, T, C = x.shape
    return x.reshape(B, ph // 2, 2, pw // 2, 2, C).amax(dim=(2, 4)).reshape(B, T // 4, C)
