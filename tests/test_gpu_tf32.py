"""GPU parity tests of the tensor-core path (tcgen05.mma kind::tf32, fp32 accumulate in TMEM),
called through the C ABI, against the oracle and the golden fixtures written from the reference.

Tolerances (written here, measured on B200):
  * north_star gate: whole-model max-abs <= 1e-3 on the output image with the reference's own
    init recipe (W1), PSNR / SSIM against a synthetic target within 0.01 dB / 1e-4;
  * W2 is a deliberately adversarial stress set (weights x5, logit scales up to the ln(100) clamp):
    TF32 operand rounding (2^-11 relative) is amplified by x100 cosine logits and 20 LayerNorms;
    with EVERY unit FORCED onto the tensor cores (set_precision("tf32"), which this file uses to test
    the tcgen05 kernels themselves) the whole-model error is ~2e-2; the W2 rows below only bound that
    forced mode (a characterisation, not the parity gate).  The parity gate on W2 - 1e-3, un-relaxed -
    is tests/test_gpu_auto_precision.py: the default mode measures a per-unit plan and runs the
    offending units on the fp32 kernels, which hold 3e-4 on the same set (tests/test_gpu_parity.py);
  * per-kernel: LayerNorm output has unit scale, so max-abs is relative to ~4.
"""
import pytest
import torch
import torch.nn.functional as F

from oracle import ngswin_oracle as O

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False


@pytest.fixture(scope="module")
def T():
    import transmar_gan_b200 as pkg
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    assert pkg._cabi.lib().b200mar_device_ok(0) == 1, "device 0 is not sm_100"
    return pkg


@pytest.fixture(autouse=True)
def tf32_kernels(T):
    old = T.ops.set_precision("tf32")
    yield
    assert T.ops.tc_error() == 0, "a tensor-core kernel timed out on an mbarrier"
    T.ops.set_precision(old)


def dev(sd):
    return {k: v.cuda() for k, v in sd.items()}


def mx_tol(kind):
    return 2e-2 if kind == "W1" else 1.5e-1


CASES = [("encoder_layer1.blocks.0.", 6, 0), ("encoder_layer1.blocks.1.", 6, 4),
         ("encoder_layer2.blocks.0.", 4, 0), ("encoder_layer3.blocks.1.", 4, 4)]


@pytest.mark.parametrize("kind", ["W1", "W2"])
@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("geom", [(1, 24, 40), (2, 16, 16), (1, 64, 48), (3, 8, 24)])
def test_attn_and_mlp_tc_forward(T, kind, case, geom):
    """attn_tc_fwd / mlp_tc_fwd against the oracle's window attention / MLP (odd window counts
    exercise the half-empty last tile; shift 4 the analytic mask and the cyclic gather)."""
    prefix, nh, shift = case
    B, ph, pw = geom
    sd = O.synthetic_state_dict(kind, seed=1)
    g = torch.Generator().manual_seed(11)
    z = torch.randn(B, ph * pw, 64, generator=g)
    ctx = torch.randn(B, (ph // 8) * (pw // 8), 64, generator=g) * 0.5
    P = [sd[prefix + n].cuda().contiguous() for n in T.ops.NSTB_PARAM_NAMES]
    with torch.no_grad():
        x = z.view(B, ph, pw, 64) + ctx.view(B, ph // 8, pw // 8, 64).repeat_interleave(8, 1).repeat_interleave(8, 2)
        if shift:
            x = torch.roll(x, (-shift, -shift), (1, 2))
        win = O.window_partition(x).reshape(-1, 64, 64)
        aw = O.window_attention(win, sd, prefix + "attn.", nh, 8, O.make_mask(ph, pw) if shift else None)
        y = O.window_unpartition(aw.reshape(-1, 8, 8, 64), ph // 8, pw // 8)
        if shift:
            y = torch.roll(y, (shift, shift), (1, 2))
        a_ref = z + F.layer_norm(y.reshape(B, -1, 64), (64,), sd[prefix + "norm1.weight"],
                                 sd[prefix + "norm1.bias"], 1e-5)
        h = F.gelu(F.linear(a_ref, sd[prefix + "ffn.fc1.weight"], sd[prefix + "ffn.fc1.bias"]))
        m = F.linear(h, sd[prefix + "ffn.fc2.weight"], sd[prefix + "ffn.fc2.bias"])
        out_ref = a_ref + F.layer_norm(m, (64,), sd[prefix + "norm2.weight"], sd[prefix + "norm2.bias"], 1e-5)
    a = T.ops.attn_tc_fwd(z.cuda(), ctx.cuda(), *P[10:18], ph, pw, nh, shift)
    out, zsum, _ = T.ops.mlp_tc_fwd(a_ref.cuda(), z.cuda(), *P[18:24], True, True)
    # the fused pair as the model runs it: attention writes LN1(attn) only, the MLP adds z
    t = T.ops.attn_tc_fwd(z.cuda(), ctx.cuda(), *P[10:18], ph, pw, nh, shift, False)
    assert torch.allclose(t + z.cuda(), a, atol=1e-5)
    out2, zsum2, a2 = T.ops.mlp_tc_fwd(t, z.cuda(), *P[18:24], True, True, True, True)
    assert torch.allclose(a2, a, atol=1e-5)
    assert (out2.cpu() - out_ref).abs().max() < mx_tol(kind) and torch.allclose(zsum2, out2 + z.cuda(), atol=1e-5)
    da = (a.cpu() - a_ref).abs()
    do = (out.cpu() - out_ref).abs()
    mx, mean = (2e-2, 2e-3) if kind == "W1" else (1.5e-1, 5e-3)
    assert da.max() < mx and da.mean() < mean, (da.max().item(), da.mean().item())
    assert do.max() < mx and do.mean() < mean, (do.max().item(), do.mean().item())
    assert torch.allclose(zsum, out + z.cuda(), atol=1e-6)


@pytest.mark.parametrize("kind,gate", [("W1", 1e-3), ("W2", 5e-2)])
@pytest.mark.parametrize("shape", [(1, 64, 64), (2, 96, 160), (1, 100, 70), (1, 128, 128)])
def test_whole_model_tf32_vs_golden(T, golden_outputs, kind, gate, shape):
    """Every unit FORCED onto tcgen05.  W1: the parity gate.  W2: bound of the forced mode only - the gate
    for W2 (1e-3) is met by the default per-unit plan, tests/test_gpu_auto_precision.py."""
    B, H, W = shape
    m = T.NGswin()
    m.load_state_dict(O.synthetic_state_dict(kind, seed=1), strict=True)
    m = m.cuda().eval()
    x = O.synthetic_input(B, H, W, seed=999)
    with torch.no_grad():
        y = m(x.cuda()).cpu()
    ref = golden_outputs[(kind, B, H, W)]
    err = (y - ref).abs().max().item()
    assert err < gate, err
    if kind == "W1":
        t = O.synthetic_input(B, H, W, seed=1000)
        assert abs(O.psnr(y, t) - O.psnr(ref, t)) < 0.01
        assert abs(O.ssim(y[0, 0], t[0, 0]) - O.ssim(ref[0, 0], t[0, 0])) < 1e-4


def test_whole_model_tf32_416(T):
    """BASELINE config-1/2 slice size; the oracle runs on the GPU in float64 as the checker."""
    sd = O.synthetic_state_dict("W1", seed=2)
    m = T.NGswin()
    m.load_state_dict(sd, strict=True)
    m = m.cuda().eval()
    x = O.synthetic_input(2, 416, 416, seed=999).cuda()
    with torch.no_grad():
        y = m(x)
        ref = O.ngswin_forward(x.double(), {k: (v.cuda().double() if v.is_floating_point() else v.cuda())
                                            for k, v in sd.items()})
    assert (y.double() - ref).abs().max().item() < 1e-3


def test_whole_model_tf32_1024_highres(T):
    """BASELINE config 5: 1024 x 1024 slice = 16 384 stage-1 windows, 127 + 127 masked edge windows per
    shifted block (the reference would materialise 6 x 268 MB masks per forward; the kernels evaluate the
    9-region rule analytically).  Checker: the oracle in float64 on the GPU."""
    sd = O.synthetic_state_dict("W1", seed=3)
    m = T.NGswin()
    m.load_state_dict(sd, strict=True)
    m = m.cuda().eval()
    x = O.synthetic_input(1, 1024, 1024, seed=999).cuda()
    with torch.no_grad():
        y = m(x)
        ref = O.ngswin_forward(x.double(), {k: (v.cuda().double() if v.is_floating_point() else v.cuda())
                                            for k, v in sd.items()})
    assert (y.double() - ref).abs().max().item() < 1e-3
    del ref
    torch.cuda.empty_cache()
    # non-multiple-of-32 high-res input: zero padding on the right / bottom, cropped back
    x2 = O.synthetic_input(1, 1000, 1010, seed=5).cuda()
    with torch.no_grad():
        y2 = m(x2)
        old = T.ops.set_precision("fp32")
        try:
            y2_32 = m(x2)
        finally:
            T.ops.set_precision(old)
    assert y2.shape == x2.shape and (y2 - y2_32).abs().max().item() < 1e-3


def test_tf32_determinism_and_batch_independence(T):
    m = T.NGswin()
    m.load_state_dict(O.synthetic_state_dict("W1", seed=1), strict=True)
    m = m.cuda().eval()
    x = O.synthetic_input(3, 64, 96, seed=5).cuda()
    with torch.no_grad():
        y, y2 = m(x), m(x)
        ys = torch.cat([m(x[i:i + 1]) for i in range(3)])
    assert torch.equal(y, y2)
    assert (y - ys).abs().max().item() < 2e-4


def test_tc_forward_tc_backward_gradients(T, golden_outputs):
    """Training in the default mode: tensor-core forward (fp16 / tf32 operands) and tensor-core backward
    (bf16 operands, tests/test_gpu_bwd_tc.py).  The L1 loss gradient is a sign function, so the
    per-tensor error of the whole-model gradients is mixed-precision-sized (median < 5e-2, worst < 0.25 on a
    1e-11-norm logit_scale gradient) rather than the 1e-3 the fp32 mode holds (tests/test_gpu_parity.py)."""
    m = T.NGswin()
    m.load_state_dict(O.synthetic_state_dict("W1", seed=1), strict=True)
    m = m.cuda().train()
    x = O.synthetic_input(2, 64, 64, seed=999).cuda()
    t = O.synthetic_input(2, 64, 64, seed=1000).cuda()
    loss = F.l1_loss(m(x), t)
    loss.backward()
    assert abs(loss.item() - golden_outputs[("W1", "l1_loss")].item()) < 1e-4
    named = dict(m.named_parameters())
    errs = []
    for n, gref in golden_outputs[("W1", "grads")].items():
        if gref.norm().item() < 1e-12:
            continue
        errs.append(((named[n].grad.cpu().double() - gref.double()).norm() / gref.double().norm()).item())
    errs.sort()
    assert errs[len(errs) // 2] < 5e-2 and errs[-1] < 0.25, (errs[len(errs) // 2], errs[-1])


def test_attn_tc_exact_softmax_fallback(T):
    """The tensor-core attention skips the row maximum for heads whose logits are bounded a priori
    ((scale + max|bias|) log2 e <= 60) and subtracts the exact row maximum for every other head.  Force the
    exact path: one huge relative-position-bias entry (200, i.e. 288 octaves; it also exercises the fp16
    bias operand of the one-hot MMA at a large magnitude)."""
    prefix, nh = "encoder_layer1.blocks.1.", 6
    sd = O.synthetic_state_dict("W1", seed=1)
    sd[prefix + "attn.relative_position_bias_table"] = sd[prefix + "attn.relative_position_bias_table"].clone()
    sd[prefix + "attn.relative_position_bias_table"][0, :] = 200.0       # (dy,dx) = (-7,-7): one key pair per window
    sd[prefix + "attn.logit_scale"] = torch.full_like(sd[prefix + "attn.logit_scale"], 4.0)
    B, ph, pw, shift = 1, 16, 24, 4
    g = torch.Generator().manual_seed(3)
    z = torch.randn(B, ph * pw, 64, generator=g)
    ctx = torch.randn(B, (ph // 8) * (pw // 8), 64, generator=g) * 0.5
    P = [sd[prefix + n].cuda().contiguous() for n in T.ops.NSTB_PARAM_NAMES]
    a_tc = T.ops.attn_tc_fwd(z.cuda(), ctx.cuda(), *P[10:18], ph, pw, nh, shift)
    a_32 = T.ops.attn_fwd(z.cuda(), ctx.cuda(), *P[10:18], ph, pw, nh, shift)
    assert torch.isfinite(a_tc).all()
    d = (a_tc - a_32).abs()
    assert d.max() < 5e-2 and d.mean() < 3e-3, (d.max().item(), d.mean().item())


@pytest.mark.gpu
@pytest.mark.parametrize("ntok", [64, 128, 192, 4160, 148 * 3 * 128 + 64])
def test_mlp_tma_inplace_matches_mlp_tc(ntok):
    """The TMA-fed in-place block tail (z += out, bulk reduce-add) is the same arithmetic as
    mlp_tc_fwd: bit-identical out and zsum, including ragged last tiles and extreme rows."""
    from transmar_gan_b200 import ops
    sd = {k: v.cuda() for k, v in O.synthetic_state_dict("W2", seed=3).items()}
    W = [sd["decoder_layer1.blocks.2." + n].contiguous() for n in ops.NSTB_PARAM_NAMES[18:24]]
    g = torch.Generator().manual_seed(ntok)
    a = (torch.randn(1, ntok, 64, generator=g) * 3).cuda()
    z = torch.randn(1, ntok, 64, generator=g).cuda()
    a[0, 0] = 0
    a[0, 1] *= 1e-6
    a[0, 2] *= 1e5                                   # would overflow an unscaled fp16 hidden row
    out_ref, zsum_ref, _ = ops.mlp_tc_fwd(a, z, *W, True, True)
    assert torch.isfinite(out_ref).all()
    z1 = z.clone()
    out = ops.mlp_tma_fwd_(a, z1, *W, True)
    assert torch.equal(out, out_ref) and torch.equal(z1, zsum_ref)
    z2 = z.clone()
    assert ops.mlp_tma_fwd_(a, z2, *W, False).numel() == 0
    assert torch.equal(z2, zsum_ref)
    assert torch.equal(ops.mlp_tma_fwd_(a, None, *W, True), out_ref)
    assert ops.tc_error() == 0


@pytest.mark.gpu
@pytest.mark.parametrize("B,ph,pw", [(1, 16, 32), (2, 24, 40), (1, 64, 64), (2, 104, 104), (1, 8, 8)])
def test_patch_merge_tc_matches_fp32_kernel(B, ph, pw):
    """TMA + tcgen05 PatchMerging (tf32 operands, LayerNorm folded around the GEMM) vs the fp32
    kernel: tf32-level agreement, ragged tiles (output grids that are not multiples of 16 x 8)."""
    from transmar_gan_b200 import ops
    sd = {k: v.cuda() for k, v in O.synthetic_state_dict("W2", seed=2).items()}
    lw, lb, rw = (sd["encoder_layer2.downsample." + n].contiguous()
                  for n in ("norm.weight", "norm.bias", "reduction.weight"))
    g = torch.Generator().manual_seed(ph * 1000 + pw)
    x = (torch.randn(B, ph * pw, 64, generator=g) * 2 + 0.7).cuda()     # non-zero mean rows
    ref = ops.patch_merge_fwd(x, lw, lb, rw, ph, pw)
    got = ops.patch_merge_tc_fwd(x, lw, lb, rw, ph, pw)
    d = (ref - got).abs()
    scale = ref.abs().max().item()
    assert d.max().item() <= 1.5e-3 * scale and d.mean().item() <= 2e-4 * scale
    assert ops.tc_error() == 0


@pytest.mark.gpu
@pytest.mark.parametrize("B,H,W", [(1, 16, 16), (2, 24, 40), (1, 64, 64), (1, 104, 72)])
def test_bottleneck_tc_matches_fp32_kernel(B, H, W):
    """TMA-halo + tcgen05 SCDP bottleneck tail vs the fp32 kernel (conv zero padding comes from the
    tensor map's out-of-bounds fill; ragged tiles when W is not a multiple of 16)."""
    from transmar_gan_b200 import ops
    sd = {k: v.cuda() for k, v in O.synthetic_state_dict("W2", seed=4).items()}
    Wt = [sd["bottleneck." + n].contiguous() for n in
          ("depthwise.weight", "depthwise.bias", "pointwise.weight", "pointwise.bias", "norm.weight", "norm.bias")]
    g = torch.Generator().manual_seed(H * 1000 + W)
    mk = lambda h, w: torch.randn(B, h * w, 64, generator=g).cuda()
    args = [mk(H, W), mk(H // 2, W // 2), mk(H // 4, W // 4), mk(H, W), mk(H // 2, W // 2),
            mk(H // 4, W // 4), *Wt, H, W]
    old = ops.set_precision("fp32")
    try:
        ref = ops.bottleneck_fwd(*args)
        ops.set_precision("tf32")
        got = ops.bottleneck_fwd(*args)
    finally:
        ops.set_precision(old)
    d = (ref - got).abs()
    assert d.max().item() <= 6e-3 and d.mean().item() <= 6e-4
    assert ops.tc_error() == 0


@pytest.mark.gpu
@pytest.mark.parametrize("two", [False, True])
@pytest.mark.parametrize("B,h,w", [(1, 8, 16), (2, 12, 20), (1, 32, 32), (2, 52, 36)])
def test_cascade_proj_tc_matches_fp32_kernel(two, B, h, w):
    """TMA + tcgen05 InterPool/concat/Linear vs the fp32 kernel; pooled features are bit-exact."""
    from transmar_gan_b200 import ops
    sd = {k: v.cuda() for k, v in O.synthetic_state_dict("W2", seed=6).items()}
    lay = "encoder_layer3" if two else "encoder_layer2"
    wt, bias = (sd[lay + ".across_cascade_proj." + n].contiguous() for n in ("weight", "bias"))
    g = torch.Generator().manual_seed(h * 100 + w + int(two))
    A = torch.randn(B, 4 * h * w, 64, generator=g).cuda()
    Bs = torch.randn(B, 4 * h * w, 64, generator=g).cuda() if two else None
    D = torch.randn(B, h * w, 64, generator=g).cuda()
    old = ops.set_precision("fp32")
    try:
        ro, rp = ops.cascade_proj_fwd(A, Bs, D, wt, bias, h, w)
        ops.set_precision("tf32")
        go, gp = ops.cascade_proj_fwd(A, Bs, D, wt, bias, h, w)
    finally:
        ops.set_precision(old)
    assert torch.equal(rp, gp)
    d = (ro - go).abs()
    scale = ro.abs().max().item()
    assert d.max().item() <= 2e-3 * scale and d.mean().item() <= 2e-4 * scale
    assert ops.tc_error() == 0
