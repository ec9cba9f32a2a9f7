"""GPU parity tests: the CUDA kernels, called through the C ABI (torch.ops.b200mar.* →
ctypes → libb200mar.so), against the oracle and the golden fixtures.

Tolerances (fp32 kernels, fp32 accumulate):
  * integer index maps / masks: bit-exact;
  * forward activations: max-abs <= 1e-3 on the output image is the north_star gate; the
    fp32 path is held to 2e-4 (W2 stress weights) / 2e-5 (W1);
  * gradients: relative L2 per tensor <= 1e-3.
"""
import math

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
    lib = pkg._cabi.lib()                       # fails loudly if libb200mar.so is missing
    assert lib.b200mar_device_ok(0) == 1, "device 0 is not sm_100"
    return pkg


@pytest.fixture(autouse=True)
def fp32_kernels(T):
    """This module pins the fp32 (FMA-pipe) kernels to their tight tolerances; the tcgen05 TF32
    kernels have their own parity tests in tests/test_gpu_tf32.py."""
    old = T.ops.set_precision("fp32")
    yield
    T.ops.set_precision(old)


def dev(sd, dtype=torch.float32):
    return {k: (v.to("cuda", dtype) if v.is_floating_point() else v.cuda()) for k, v in sd.items()}


def relerr(a, b):
    return ((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30)).item()


# ------------------------------------------------------------------------------ index maps
@pytest.mark.parametrize("geom", [(1, 16, 16), (2, 24, 40), (1, 64, 64), (3, 8, 16), (1, 416, 416)])
@pytest.mark.parametrize("shift", [0, 4])
def test_partition_index_bit_exact(T, golden_maps, geom, shift):
    B, ph, pw = geom
    tok, cwin = T.ops.partition_index(B, ph, pw, shift, "cuda")
    otok, octx = O.partition_source_index(B, ph, pw, shift)
    assert torch.equal(tok.cpu(), otok)
    assert torch.equal(cwin.cpu(), octx)
    key = (B, ph, pw, shift)
    if key in golden_maps["partition_maps"]:
        assert torch.equal(tok.cpu(), golden_maps["partition_maps"][key])
    # round trip: scatter(gather(x)) == x, every token exactly once
    assert torch.equal(torch.sort(tok.flatten()).values.cpu(), torch.arange(B * ph * pw))


@pytest.mark.parametrize("geom", [(16, 16), (24, 40), (8, 16), (64, 64), (128, 128)])
def test_shift_mask_and_rel_index_bit_exact(T, golden_maps, golden_contract, geom):
    ph, pw = geom
    m = T.ops.shift_mask(ph, pw, 4, "cuda").cpu()
    assert torch.equal(m, O.make_mask(ph, pw))
    if geom in golden_maps["masks_i8"]:
        assert torch.equal(m.to(torch.int8), golden_maps["masks_i8"][geom])
    assert torch.equal(T.ops.shift_mask(ph, pw, 0, "cuda").cpu(), torch.zeros_like(m))
    assert torch.equal(T.ops.rel_index("cuda").cpu(), golden_contract["rel_index_8"])


# ------------------------------------------------------------------------------ kernels, forward
def block_params(T, sd, prefix):
    return [sd[prefix + n].contiguous() for n in T.ops.NSTB_PARAM_NAMES]


CASES = [("encoder_layer1.blocks.0.", 6, 0), ("encoder_layer1.blocks.1.", 6, 4),
         ("encoder_layer2.blocks.0.", 4, 0), ("encoder_layer3.blocks.1.", 4, 4)]


@pytest.mark.parametrize("kind", ["W1", "W2"])
@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("geom", [(1, 24, 40), (2, 16, 16), (1, 64, 48)])
def test_nstb_pieces_forward(T, kind, case, geom):
    prefix, nh, shift = case
    B, ph, pw = geom
    sd = O.synthetic_state_dict(kind, seed=1)
    g = torch.Generator().manual_seed(11)
    z = torch.randn(B, ph * pw, 64, generator=g)
    gsd, P = dev(sd), None
    P = block_params(T, gsd, prefix)
    zc = z.cuda()
    # oracle pieces (CPU fp32)
    cpre = prefix + "ngram_window_partition.ngram_context."
    with torch.no_grad():
        u_ref = O.unigram_embed(z.view(B, ph, pw, 64), sd, cpre).permute(0, 2, 3, 1)
        ctx_ref = O.ngram_context(z.view(B, ph, pw, 64), sd, cpre, nh)
        out_ref = O.nstb(z, ph, pw, sd, prefix, nh, shift)
        # a = z + LN1(attn): recover from the oracle by undoing the MLP branch is awkward;
        # restate with the oracle's own functions instead
        win = O.ngram_window_partition(z.view(B, ph, pw, 64), sd, prefix + "ngram_window_partition.",
                                       nh, shift).reshape(-1, 64, 64)
        mask = O.make_mask(ph, pw) if shift else None
        aw = O.window_attention(win, sd, prefix + "attn.", nh, 8, mask)
        y = O.window_unpartition(aw.reshape(-1, 8, 8, 64), ph // 8, pw // 8)
        if shift:
            y = torch.roll(y, (shift, shift), (1, 2))
        a_ref = z + F.layer_norm(y.reshape(B, -1, 64), (64,), sd[prefix + "norm1.weight"],
                                 sd[prefix + "norm1.bias"], 1e-5)
    u = T.ops.unigram_fwd(zc, P[0], P[1], ph, pw)
    ctx = T.ops.ngram_ctx_fwd(u, *P[2:10], nh)
    a = T.ops.attn_fwd(zc, ctx, *P[10:18], ph, pw, nh, shift)
    out, zsum = T.ops.mlp_fwd(a, zc, *P[18:24], True, True)
    tol = 2e-5 if kind == "W1" else 3e-4
    assert (u.cpu() - u_ref).abs().max() < tol
    assert (ctx.cpu() - ctx_ref).abs().max() < tol
    assert (a.cpu() - a_ref).abs().max() < tol
    assert (out.cpu() - out_ref).abs().max() < tol
    assert (zsum.cpu() - (out_ref + z)).abs().max() < tol
    only_out, empty = T.ops.mlp_fwd(a, None, *P[18:24], True, False)
    assert empty.numel() == 0 and torch.equal(only_out, out)


@pytest.mark.parametrize("kind", ["W1", "W2"])
def test_nstb_vs_golden(T, golden_outputs, kind):
    """Same inputs as the fixtures written from the unmodified reference."""
    sd = dev(O.synthetic_state_dict(kind, seed=1))
    g = torch.Generator().manual_seed(golden_outputs[(kind, "z_seed")])
    z = torch.randn(1, 24 * 40, 64, generator=g).cuda()
    tol = 2e-5 if kind == "W1" else 3e-4
    for stage, nh, key in (("encoder_layer1", 6, "nstb6"), ("encoder_layer2", 4, "nstb4")):
        for bi in (0, 1):
            P = block_params(T, sd, f"{stage}.blocks.{bi}.")
            out, _ = T.ops.NSTBFunction.apply(z, 24, 40, nh, 0 if bi == 0 else 4, True, False, False, *P)
            assert (out.cpu() - golden_outputs[(kind, key, bi)]).abs().max() < tol, (stage, bi)
    pm = T.ops.patch_merge_fwd(z, sd["encoder_layer1.downsample.norm.weight"],
                               sd["encoder_layer1.downsample.norm.bias"],
                               sd["encoder_layer1.downsample.reduction.weight"], 24, 40)
    assert (pm.cpu() - golden_outputs[(kind, "patch_merge")]).abs().max() < tol


def test_pixel_shuffle_cat_bit_exact(T, golden_outputs):
    B, H, W = 2, 12, 20
    g = torch.Generator().manual_seed(3)
    x0 = torch.randn(B, H * W, 64, generator=g)
    x1 = torch.randn(B, H * W // 4, 64, generator=g)
    x2 = torch.randn(B, H * W // 16, 64, generator=g)
    ref = torch.cat([x0, O.pixel_shuffle_permute(x1, H // 2, W // 2, 2),
                     O.pixel_shuffle_permute(x2, H // 4, W // 4, 4)], dim=-1)
    cat = T.ops.pixel_shuffle_cat(x0.cuda(), x1.cuda(), x2.cuda(), H, W)
    assert torch.equal(cat.cpu(), ref)
    # golden (arange) vectors from the reference's pixel_shuffle_permute
    a1 = torch.arange(60 * 64, dtype=torch.float32).reshape(1, 60, 64)
    a2 = torch.arange(15 * 64, dtype=torch.float32).reshape(1, 15, 64)
    c = T.ops.pixel_shuffle_cat(torch.zeros(1, 240, 64).cuda(), a1.cuda(), a2.cuda(), 12, 20).cpu()
    assert torch.equal(c[..., 64:80], golden_outputs[("pixel_shuffle", 2)])
    assert torch.equal(c[..., 80:84], golden_outputs[("pixel_shuffle", 4)])
    # the inverse map is the exact transpose (round trip)
    d0, d1, d2 = T.ops.pixel_shuffle_cat_bwd(cat, H, W)
    assert torch.equal(d0.cpu(), x0) and torch.equal(d1.cpu(), x1) and torch.equal(d2.cpu(), x2)


# ------------------------------------------------------------------------------ whole model, forward
def load_model(T, kind, seed=1):
    m = T.NGswin()
    m.load_state_dict(O.synthetic_state_dict(kind, seed=seed), strict=True)
    return m.cuda().eval()


@pytest.mark.parametrize("kind", ["W1", "W2"])
@pytest.mark.parametrize("shape", [(1, 64, 64), (2, 96, 160), (1, 100, 70), (1, 128, 128)])
def test_whole_model_vs_golden(T, golden_outputs, kind, shape):
    B, H, W = shape
    m = load_model(T, kind)
    x = O.synthetic_input(B, H, W, seed=999)
    with torch.no_grad():
        y = m(x.cuda()).cpu()
    ref = golden_outputs[(kind, B, H, W)]
    assert y.shape == ref.shape
    err = (y - ref).abs().max().item()
    assert err < 1e-3                      # north_star gate
    assert err < (2e-5 if kind == "W1" else 3e-4), err   # what the fp32 path actually holds
    # PSNR / SSIM of {ours, reference} against a synthetic target within 0.01 dB / 1e-4
    t = O.synthetic_input(B, H, W, seed=1000)
    assert abs(O.psnr(y, t) - O.psnr(ref, t)) < 0.01
    assert abs(O.ssim(y[0, 0], t[0, 0]) - O.ssim(ref[0, 0], t[0, 0])) < 1e-4


def test_whole_model_416_vs_oracle_on_gpu(T):
    """BASELINE config-1 size: the oracle itself is run on the GPU in float64 as the checker."""
    sd = O.synthetic_state_dict("W2", seed=2)
    m = T.NGswin()
    m.load_state_dict(sd, strict=True)
    m = m.cuda().eval()
    x = O.synthetic_input(1, 416, 416, seed=999).cuda()
    with torch.no_grad():
        y = m(x)
        ref = O.ngswin_forward(x.double(), dev(sd, torch.float64))
    assert (y.double() - ref).abs().max().item() < 3e-4


def test_batch_independence_and_determinism(T):
    """Windows never cross samples: a batch equals its slices run one by one (bit-exact,
    same kernels, same per-window arithmetic), and reruns are deterministic."""
    m = load_model(T, "W2")
    x = O.synthetic_input(3, 64, 96, seed=5).cuda()
    with torch.no_grad():
        y = m(x)
        y2 = m(x)
        ys = torch.cat([m(x[i:i + 1]) for i in range(3)])
    assert torch.equal(y, y2)
    assert (y - ys).abs().max().item() < 5e-5     # stock cuDNN/cuBLAS pieces pick per-batch algorithms


def test_small_input_raises(T):
    m = load_model(T, "W1")
    with pytest.raises(RuntimeError):
        m(torch.zeros(1, 1, 32, 32, device="cuda"))


# ------------------------------------------------------------------------------ backward
@pytest.mark.parametrize("kind", ["W1", "W2"])
@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("geom", [(2, 16, 24), (1, 24, 40)])
def test_nstb_backward_vs_oracle_autograd(T, kind, case, geom):
    prefix, nh, shift = case
    B, ph, pw = geom
    sd = O.synthetic_state_dict(kind, seed=1)
    g = torch.Generator().manual_seed(21)
    z = torch.randn(B, ph * pw, 64, generator=g)
    w_out = torch.randn(B, ph * pw, 64, generator=g)
    w_zs = torch.randn(B, ph * pw, 64, generator=g)
    # ---- oracle: float64 autograd on the GPU
    sd64 = {k: (v.cuda().double().requires_grad_(True) if k.startswith(prefix) and v.is_floating_point()
                and not k.endswith("attn_mask") else v.cuda()) for k, v in sd.items() if k.startswith(prefix)}
    z64 = z.cuda().double().requires_grad_(True)
    out64 = O.nstb(z64, ph, pw, sd64, prefix, nh, shift)
    loss64 = (out64 * w_out.cuda().double()).sum() + ((out64 + z64) * w_zs.cuda().double()).sum()
    loss64.backward()
    # ---- CUDA kernels
    P = [sd[prefix + n].cuda().contiguous().requires_grad_(True) for n in T.ops.NSTB_PARAM_NAMES]
    zc = z.cuda().requires_grad_(True)
    out, zsum = T.ops.NSTBFunction.apply(zc, ph, pw, nh, shift, True, True, True, *P)
    loss = (out * w_out.cuda()).sum() + (zsum * w_zs.cuda()).sum()
    loss.backward()
    assert abs(loss.item() - loss64.item()) < 1e-3 * max(1.0, abs(loss64.item()))
    tol = 2e-4
    assert relerr(zc.grad, z64.grad) < tol
    for n, p in zip(T.ops.NSTB_PARAM_NAMES, P):
        gref = sd64[prefix + n].grad
        assert p.grad is not None, n
        denom = gref.norm().item()
        if denom < 1e-9:                       # e.g. clamped logit_scale entries: exact zeros
            assert p.grad.abs().max().item() < 1e-6, n
        else:
            assert relerr(p.grad, gref) < tol, (n, relerr(p.grad, gref))


@pytest.mark.parametrize("geom", [(1, 24, 40), (2, 16, 16)])
def test_patch_merge_backward(T, geom):
    B, ph, pw = geom
    sd = O.synthetic_state_dict("W2", seed=1)
    pre = "encoder_layer1.downsample."
    g = torch.Generator().manual_seed(4)
    x = torch.randn(B, ph * pw, 64, generator=g)
    w = torch.randn(B, ph * pw // 4, 64, generator=g)
    names = ["norm.weight", "norm.bias", "reduction.weight"]
    p64 = {pre + n: sd[pre + n].cuda().double().requires_grad_(True) for n in names}
    x64 = x.cuda().double().requires_grad_(True)
    (O.patch_merging(x64, ph, pw, p64, pre) * w.cuda().double()).sum().backward()
    P = [sd[pre + n].cuda().requires_grad_(True) for n in names]
    xc = x.cuda().requires_grad_(True)
    y = T.ops.PatchMergeFunction.apply(xc, ph, pw, *P)
    (y * w.cuda()).sum().backward()
    assert relerr(xc.grad, x64.grad) < 2e-4
    for n, p in zip(names, P):
        assert relerr(p.grad, p64[pre + n].grad) < 2e-4, n


@pytest.mark.parametrize("kind", ["W2", "W1"])
def test_whole_model_l1_gradients_vs_golden(T, golden_outputs, kind):
    """Config-3 recipe: F.l1_loss(netG(x), t).backward(); every parameter tensor the reference
    gives a gradient to gets the same gradient (relative L2 <= 1e-3 per tensor)."""
    m = T.NGswin()
    m.load_state_dict(O.synthetic_state_dict(kind, seed=1), strict=True)
    m = m.cuda().train()
    x = O.synthetic_input(2, 64, 64, seed=999).cuda().requires_grad_(True)
    t = O.synthetic_input(2, 64, 64, seed=1000).cuda()
    loss = F.l1_loss(m(x), t)
    loss.backward()
    assert abs(loss.item() - golden_outputs[(kind, "l1_loss")].item()) < 1e-5
    gx = golden_outputs[(kind, "grad_x")]
    assert relerr(x.grad.cpu(), gx) < 1e-3
    grads = golden_outputs[(kind, "grads")]
    named = dict(m.named_parameters())
    assert kind != "W2" or len(grads) == 504
    worst = ("", 0.0)
    for n, gref in grads.items():
        assert named[n].grad is not None, n
        if gref.norm().item() < 1e-12:
            assert named[n].grad.abs().max().item() < 1e-8, n
            continue
        r = relerr(named[n].grad.cpu(), gref)
        if r > worst[1]:
            worst = (n, r)
    assert worst[1] < 1e-3, worst


def test_training_step_reduces_loss(T):
    """A few Adam steps (betas .5/.999, lr 1e-4: training/train_combined.py:99-104) through the
    CUDA forward/backward lower the L1 loss — end-to-end sanity of the gradient sign/scale."""
    torch.manual_seed(0)
    m = load_model(T, "W1").train()
    opt = torch.optim.Adam(m.parameters(), lr=1e-4, betas=(0.5, 0.999))
    x = O.synthetic_input(2, 64, 64, seed=999).cuda()
    t = (x * 0.5).detach()
    losses = []
    for _ in range(8):
        opt.zero_grad(set_to_none=True)
        loss = F.l1_loss(m(x), t)
        loss.backward()
        opt.step()
        losses.append(loss.item())
    assert losses[-1] < losses[0]


def test_fused_glue_matches_composite_path(T):
    """The fused inference glue (glue_fwd.cu: shallow conv, pool+concat+cascade projection,
    bottleneck, LN + skip + reconstruction) against the stock-PyTorch composition that training
    uses, same fp32 block kernels, incl. a size that needs padding + cropping."""
    m = load_model(T, "W2")
    for shape in ((2, 64, 96), (1, 100, 70)):
        x = O.synthetic_input(*shape, seed=7).cuda()
        with torch.no_grad():
            y_fused = m(x)
        with torch.enable_grad():
            y_comp = m(x).detach()
        assert y_fused.shape == y_comp.shape
        assert (y_fused - y_comp).abs().max().item() < 2e-5


@pytest.mark.gpu
@pytest.mark.parametrize("nh,B,ph,pw", [(6, 2, 16, 16), (6, 1, 24, 40), (4, 3, 32, 16), (6, 1, 416, 416)])
def test_ngram_ctx_infer_matches_two_step(nh, B, ph, pw):
    """Inference pair (unigram + per-window qkv, then gather) == unigram_fwd + ngram_ctx_fwd."""
    from transmar_gan_b200 import ops
    sd = {k: v.cuda() for k, v in O.synthetic_state_dict("W2", seed=5).items()}
    pre = "encoder_layer1.blocks.3." if nh == 6 else "encoder_layer2.blocks.1."
    P = [sd[pre + n].contiguous() for n in ops.NSTB_PARAM_NAMES[:10]]
    g = torch.Generator().manual_seed(ph * pw + nh)
    z = torch.randn(B, ph * pw, 64, generator=g).cuda()
    ref = ops.ngram_ctx_fwd(ops.unigram_fwd(z, P[0], P[1], ph, pw), *P[2:10], nh)
    got = ops.ngram_ctx_infer(z, *P, ph, pw, nh)
    assert got.shape == ref.shape
    assert (got - ref).abs().max().item() <= 2e-5 * max(1.0, ref.abs().max().item())


@pytest.mark.gpu
@pytest.mark.parametrize("nh,B,ph,pw", [(6, 2, 16, 16), (6, 1, 24, 40), (4, 3, 32, 16), (6, 16, 128, 128)])
def test_ngram_ctx_train_fwd_matches_two_step(nh, B, ph, pw):
    """Training form of the inference pair: (u, ctx) == (unigram_fwd, ngram_ctx_fwd); u is what the backward reads."""
    from transmar_gan_b200 import ops
    sd = {k: v.cuda() for k, v in O.synthetic_state_dict("W2", seed=5).items()}
    pre = "encoder_layer1.blocks.3." if nh == 6 else "encoder_layer2.blocks.1."
    P = [sd[pre + n].contiguous() for n in ops.NSTB_PARAM_NAMES[:10]]
    g = torch.Generator().manual_seed(ph * pw + nh + 1)
    z = torch.randn(B, ph * pw, 64, generator=g).cuda()
    u_ref = ops.unigram_fwd(z, P[0], P[1], ph, pw)
    c_ref = ops.ngram_ctx_fwd(u_ref, *P[2:10], nh)
    u, c = ops.ngram_ctx_train_fwd(z, *P, ph, pw, nh)
    assert u.shape == u_ref.shape and c.shape == c_ref.shape
    assert (u - u_ref).abs().max().item() <= 2e-5 * max(1.0, u_ref.abs().max().item())
    assert (c - c_ref).abs().max().item() <= 2e-5 * max(1.0, c_ref.abs().max().item())


@pytest.mark.gpu
@pytest.mark.parametrize("n_tok", [1, 2, 33, 4096, 148 * 8 * 16 * 3 + 5])
def test_ln64_fwd_bwd_vs_torch(n_tok):
    """ln64.cu (LayerNorm(64) of the training-path stage glue: bottleneck.norm, NGswin.norm - ngswin.py:204)
    against torch.nn.functional.layer_norm in float64, forward and backward, odd row counts included."""
    import transmar_gan_b200 as T
    g = torch.Generator().manual_seed(n_tok)
    x = (torch.randn(n_tok, 64, generator=g) * 3 + 0.5).cuda().requires_grad_(True)
    w = torch.randn(64, generator=g).cuda().requires_grad_(True)
    b = torch.randn(64, generator=g).cuda().requires_grad_(True)
    dy = torch.randn(n_tok, 64, generator=g).cuda()
    y = T.ops.LayerNorm64Function.apply(x, w, b)
    y.backward(dy)
    x64, w64, b64 = (t.detach().double().requires_grad_(True) for t in (x, w, b))
    y64 = torch.nn.functional.layer_norm(x64, (64,), w64, b64, 1e-5)
    y64.backward(dy.double())
    rel = lambda a, r: ((a.double() - r).norm() / r.norm().clamp_min(1e-30)).item()
    assert rel(y, y64) < 1e-6
    assert rel(x.grad, x64.grad) < 1e-5 and rel(w.grad, w64.grad) < 1e-5 and rel(b.grad, b64.grad) < 1e-5
