"""Out-of-bounds WRITE check for the TMA / tcgen05 kernels (compute-sanitizer is closed on this GPU pool, so the
bounds evidence is our own): every output tensor of a launch is carved out of one arena with 64 KB sentinel bands
on both sides (and the inputs sit between bands too); after the launch all bands must still hold the sentinel and
the output must be fully written (no sentinel left) and finite.  Shapes are ragged on purpose: odd window counts
(half-empty last tile), shifted blocks (cyclically wrapped windows are written by direct stores, the others by TMA
boxes), token counts that are not multiples of the 128-row tile, the bf16-stream variants, the no-alloc entry."""
import pytest
import torch

from oracle import ngswin_oracle as O

pytestmark = pytest.mark.gpu
BAND = 16384                      # floats = 64 KB
SENT = -7.0e30


class Arena:
    def __init__(self, n_floats):
        self.buf = torch.full((n_floats,), SENT, device="cuda")
        self.off, self.bands = BAND, [(0, BAND)]

    def take(self, shape, dtype=torch.float32, fill=None):
        n = 1
        for s in shape:
            n *= s
        nf = (n * (2 if dtype == torch.bfloat16 else 4) + 3) // 4
        nf = (nf + 63) // 64 * 64                                   # keep 256-byte alignment
        t = self.buf[self.off:self.off + nf].view(dtype)[:n].view(*shape)
        if fill is not None:
            t.copy_(fill)
        self.bands.append((self.off + nf, self.off + nf + BAND))
        self.off += nf + BAND
        return t

    def check(self):
        torch.cuda.synchronize()
        for a, b in self.bands:
            assert bool((self.buf[a:b] == SENT).all()), f"guard band [{a},{b}) was written"


def written(t):
    f = t.float()
    return bool(torch.isfinite(f).all()) and not bool((f.abs() > 1e29).any())


@pytest.mark.parametrize("case", [("encoder_layer1.blocks.0.", 6, 0), ("encoder_layer1.blocks.1.", 6, 4),
                                  ("encoder_layer2.blocks.1.", 4, 4)])
@pytest.mark.parametrize("geom", [(1, 24, 40), (3, 8, 24), (1, 16, 24), (2, 16, 16)])
def test_block_kernels_stay_inside_their_outputs(case, geom):
    import transmar_gan_b200 as T
    ops, lib = T.ops, T._cabi.lib()
    prefix, nh, shift = case
    B, ph, pw = geom
    wh, ww = ph // 8, pw // 8
    sd = O.synthetic_state_dict("W2", seed=1)
    P = [sd[prefix + n].cuda().contiguous() for n in ops.NSTB_PARAM_NAMES]
    p = [t.data_ptr() for t in P]
    g = torch.Generator().manual_seed(3)
    n = B * ph * pw
    ar = Arena(40 * BAND + 12 * n * 64)
    z = ar.take((B, n // B, 64), fill=torch.randn(B, n // B, 64, generator=g).cuda())
    ctx = ar.take((B, wh * ww, 64), fill=(torch.randn(B, wh * ww, 64, generator=g) * 0.5).cuda())
    a = ar.take((B, n // B, 64))
    out = ar.take((B, n // B, 64))
    zs = ar.take((B, n // B, 64), fill=z)
    z16 = ar.take((B, n // B, 64), torch.bfloat16, fill=z.bfloat16())
    a16 = ar.take((B, n // B, 64), torch.bfloat16)
    out16 = ar.take((B, n // B, 64), torch.bfloat16)
    zs16 = ar.take((B, n // B, 64), torch.bfloat16, fill=z.bfloat16())
    nq = ar.take((B, wh, ww, 96))
    cx = ar.take((B, wh, ww, 64))
    cx2 = ar.take((B, wh, ww, 64))
    st = torch.cuda.current_stream().cuda_stream
    assert lib.b200mar_attn_tc_fwd(z.data_ptr(), ctx.data_ptr(), *p[10:18], a.data_ptr(), B, ph, pw, nh, shift, 1, st) == 0
    assert lib.b200mar_mlp_tma_fwd(a.data_ptr(), zs.data_ptr(), *p[18:24], out.data_ptr(), n, st) == 0
    assert lib.b200mar_attn_tc_fwd_s16(z16.data_ptr(), ctx.data_ptr(), *p[10:18], a16.data_ptr(), B, ph, pw, nh, shift, st) == 0
    assert lib.b200mar_mlp_tma_fwd_s16(a16.data_ptr(), zs16.data_ptr(), *p[18:24], out16.data_ptr(), n, st) == 0
    if wh >= 2 and ww >= 2:
        assert lib.b200mar_unigram_qkv_fwd_s16(z16.data_ptr(), *p[0:4], nq.data_ptr(), B, ph, pw, nh, st) == 0
        assert lib.b200mar_ngram_ctx_pre_fwd(nq.data_ptr(), *p[4:10], cx.data_ptr(), B, wh, ww, nh, st) == 0
        assert lib.b200mar_ngram_ctx_infer_noalloc(z.data_ptr(), *p[0:10], cx2.data_ptr(), B, ph, pw, nh, st) == 0
    ar.check()
    assert ops.tc_error() == 0
    for t in (a, out, zs, a16, out16, zs16) + ((nq, cx, cx2) if wh >= 2 and ww >= 2 else ()):
        assert written(t)


@pytest.mark.parametrize("geom", [(1, 16, 24), (2, 32, 16), (1, 48, 80)])
def test_glue_tc_kernels_stay_inside_their_outputs(geom):
    import transmar_gan_b200 as T
    ops = T.ops
    B, H, W = geom                                   # full-resolution token grid; H, W multiples of 16
    sd = {k: v.cuda() for k, v in O.synthetic_state_dict("W2", seed=1).items()}
    g = torch.Generator().manual_seed(5)
    ar = Arena(24 * BAND + 8 * B * H * W * 64)
    x = ar.take((B, H * W, 64), fill=torch.randn(B, H * W, 64, generator=g).cuda())
    old = ops.set_precision("tf32")
    try:
        y = ops.patch_merge_tc_fwd(x, sd["encoder_layer1.downsample.norm.weight"], sd["encoder_layer1.downsample.norm.bias"],
                                   sd["encoder_layer1.downsample.reduction.weight"], H, W)
        x2, p0 = ops.cascade_proj_fwd(x, None, y, sd["encoder_layer2.across_cascade_proj.weight"],
                                      sd["encoder_layer2.across_cascade_proj.bias"], H // 2, W // 2)
    finally:
        ops.set_precision(old)
    ar.check()                                        # the inputs' bands: nothing was written around x
    assert written(y) and written(x2) and written(p0)
    assert ops.tc_error() == 0


@pytest.mark.parametrize("case", [("encoder_layer1.blocks.1.", 6, 4), ("encoder_layer2.blocks.0.", 4, 0)])
@pytest.mark.parametrize("geom", [(1, 24, 40), (3, 8, 24), (2, 16, 16)])
def test_compensated_and_backward_kernels_stay_inside_their_outputs(case, geom):
    """The precision-level "tf32c" kernels (attention incl. the saved-O form, MLP) and the TMA-fed MLP backward, whose
    bf16 exports go to a caller-owned scratch of n_tok * 384 elements: bands around every output and the scratch."""
    import transmar_gan_b200 as T
    ops, lib = T.ops, T._cabi.lib()
    prefix, nh, shift = case
    B, ph, pw = geom
    wh, ww = ph // 8, pw // 8
    sd = O.synthetic_state_dict("W1", seed=1)
    P = [sd[prefix + n].cuda().contiguous() for n in ops.NSTB_PARAM_NAMES]
    p = [t.data_ptr() for t in P]
    g = torch.Generator().manual_seed(5)
    n = B * ph * pw                                   # a multiple of 64, not always of 128
    ntile = (B * wh * ww + 1) // 2
    ar = Arena(40 * BAND + 16 * n * 64 + 2 * ntile * 128 * 64)
    z = ar.take((B, n // B, 64), fill=torch.randn(B, n // B, 64, generator=g).cuda())
    ctx = ar.take((B, wh * ww, 64), fill=(torch.randn(B, wh * ww, 64, generator=g) * 0.5).cuda())
    dout = ar.take((B, n // B, 64), fill=torch.randn(B, n // B, 64, generator=g).cuda())
    a = ar.take((B, n // B, 64))
    a2 = ar.take((B, n // B, 64))
    o16 = ar.take((ntile * 128, 64), torch.bfloat16)
    out = ar.take((B, n // B, 64))
    zs = ar.take((B, n // B, 64), fill=z)
    da = ar.take((B, n // B, 64))
    scratch = ar.take((n * 384,), torch.bfloat16)
    G = [ar.take(tuple(t.shape), fill=torch.zeros_like(t)) for t in P[18:24]]
    st = torch.cuda.current_stream().cuda_stream
    assert lib.b200mar_attn_tc_fwd_comp(z.data_ptr(), ctx.data_ptr(), *p[10:18], a.data_ptr(), 0, B, ph, pw, nh, shift, st) == 0
    assert lib.b200mar_attn_tc_fwd_comp(z.data_ptr(), ctx.data_ptr(), *p[10:18], a2.data_ptr(), o16.data_ptr(), B, ph, pw, nh,
                                        shift, st) == 0
    assert lib.b200mar_mlp_tma_fwd_comp(a.data_ptr(), zs.data_ptr(), *p[18:24], out.data_ptr(), n, st) == 0
    assert lib.b200mar_mlp_bwd_tc(a.data_ptr(), dout.data_ptr(), *p[18:24], da.data_ptr(), *[t.data_ptr() for t in G],
                                  scratch.data_ptr(), n, st) == 0
    ar.check()
    assert ops.tc_error() == 0
    assert torch.equal(a, a2)
    for t in (a, out, zs, da, scratch):
        assert written(t)
    valid = B * wh * ww * 64                           # the rows of a half-empty last tile are never written
    assert written(o16[:valid])
