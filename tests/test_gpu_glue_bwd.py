"""The fused glue of the TRAINING path: every autograd Function of ops.py that wraps a fused
glue kernel (forward) and its hand-written backward, against stock PyTorch autograd of the same arithmetic
(float64 on the GPU as the checker; the stock composite restates the reference's modules: reconstruction.py:14-19,
encoder_decoder.py:23-27,117,188-200, pooling.py:13-17,32-37).  Tolerance: 1e-5 relative L2 per tensor (fp32
accumulation order) unless stated."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def rel(a, b):
    return ((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30)).item()


@pytest.fixture(autouse=True)
def fp32_kernels():
    import transmar_gan_b200 as T
    old = T.ops.set_precision("fp32")
    old_tf32 = torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = torch.backends.cudnn.allow_tf32 = False
    yield
    torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32 = old_tf32
    T.ops.set_precision(old)


@pytest.mark.parametrize("geom", [(2, 64, 96, 64, 96), (1, 128, 96, 100, 70), (3, 32, 32, 32, 17)])
@pytest.mark.parametrize("fused_acc", [False, True])
def test_final_recon_function(geom, fused_acc):
    from transmar_gan_b200 import ops
    B, H, W, H0, W0 = geom
    g = torch.Generator().manual_seed(1)
    dec = torch.randn(B, H * W, 64, generator=g).cuda().requires_grad_(True)
    c0 = torch.randn(B, H * W, 64, generator=g).cuda().requires_grad_(True)
    P = [(1 + 0.1 * torch.randn(64, generator=g)), 0.1 * torch.randn(64, generator=g),
         0.05 * torch.randn(1, 64, 3, 3, generator=g), 0.1 * torch.randn(1, generator=g),
         0.3 * torch.randn(1, 1, 3, 3, generator=g), 0.1 * torch.randn(1, generator=g)]
    P = [p.cuda().requires_grad_(True) for p in P]
    dy = torch.randn(B, 1, H0, W0, generator=g).cuda()
    # stock composite in float64
    D = [t.detach().double().requires_grad_(True) for t in (dec, c0, *P)]
    v = F.layer_norm(D[0], (64,), D[2], D[3], 1e-5) + D[1]
    img = v.transpose(1, 2).reshape(B, 64, H, W)
    ref = torch.tanh(F.conv2d(F.conv2d(img, D[4], D[5], padding=1), D[6], D[7], padding=1))[:, :, :H0, :W0]
    ref.backward(dy.double())
    if fused_acc:
        for p in P:
            p.grad = torch.zeros_like(p)
    with ops.fused_grad_accumulation(fused_acc):
        out = ops.FinalReconFunction.apply(dec, c0, *P, H, W, H0, W0)
        assert rel(out, ref) < 1e-5
        out.backward(dy)
    for got, want, name in zip((dec, c0, *P), D, ("ddec", "dc0", "dln_w", "dln_b", "dw1", "db1", "dw2", "db2")):
        assert rel(got.grad, want.grad) < 2e-5, (name, rel(got.grad, want.grad))


@pytest.mark.parametrize("geom", [(2, 64, 96), (1, 40, 24), (3, 16, 16)])
@pytest.mark.parametrize("need_dx", [True, False])
def test_shallow_tok_function(geom, need_dx):
    from transmar_gan_b200 import ops
    B, H, W = geom
    g = torch.Generator().manual_seed(2)
    x = torch.randn(B, 1, H, W, generator=g).cuda().requires_grad_(need_dx)
    w = (0.2 * torch.randn(64, 1, 3, 3, generator=g)).cuda().requires_grad_(True)
    b = (0.1 * torch.randn(64, generator=g)).cuda().requires_grad_(True)
    dc0 = torch.randn(B, H * W, 64, generator=g).cuda()
    X, Wd, Bd = (t.detach().double().requires_grad_(True) for t in (x, w, b))
    ref = F.conv2d(X, Wd, Bd, padding=1).flatten(2).transpose(1, 2)
    ref.backward(dc0.double())
    c0 = ops.ShallowTokFunction.apply(x, w, b)
    assert rel(c0, ref) < 1e-6
    c0.backward(dc0)
    assert rel(w.grad, Wd.grad) < 2e-5 and rel(b.grad, Bd.grad) < 2e-5
    if need_dx:
        assert rel(x.grad, X.grad) < 2e-5
    else:
        assert x.grad is None


def _pool(x, B, h, wd):
    """nn.MaxPool2d(2) on a token-major [B, 4*h*wd, C] map (first-maximum routing in the backward)."""
    C = x.shape[-1]
    y = F.max_pool2d(x.transpose(1, 2).reshape(B, C, 2 * h, 2 * wd), 2)
    return y.flatten(2).transpose(1, 2)


@pytest.mark.parametrize("geom", [(2, 16, 24), (1, 8, 8), (3, 12, 20)])
@pytest.mark.parametrize("with_b", [False, True])
@pytest.mark.parametrize("use_pool_grad", [True, False])
def test_cascade_proj_function(geom, with_b, use_pool_grad):
    from transmar_gan_b200 import ops
    B, h, wd = geom
    K = 192 if with_b else 128
    g = torch.Generator().manual_seed(3)
    srcA = torch.randn(B, 4 * h * wd, 64, generator=g).cuda().requires_grad_(True)
    srcB = torch.randn(B, 4 * h * wd, 64, generator=g).cuda().requires_grad_(True) if with_b else None
    direct = torch.randn(B, h * wd, 64, generator=g).cuda().requires_grad_(True)
    w = (0.1 * torch.randn(64, K, generator=g)).cuda().requires_grad_(True)
    b = (0.1 * torch.randn(64, generator=g)).cuda().requires_grad_(True)
    dout = torch.randn(B, h * wd, 64, generator=g).cuda()
    dpool = torch.randn(B, h * wd, 64, generator=g).cuda()
    ins = [t for t in (srcA, srcB, direct, w, b) if t is not None]
    D = [t.detach().double().requires_grad_(True) for t in ins]
    dA, dB_, dD, dW, dBi = (D[0], D[1], D[2], D[3], D[4]) if with_b else (D[0], None, D[1], D[2], D[3])
    pa = _pool(dA, B, h, wd)
    cat = torch.cat([pa] + ([_pool(dB_, B, h, wd)] if with_b else []) + [dD], dim=-1)
    ref = F.linear(cat, dW, dBi)
    loss = (ref * dout.double()).sum() + ((pa * dpool.double()).sum() if use_pool_grad else 0.0)
    loss.backward()
    out, pool = ops.CascadeProjFunction.apply(srcA, srcB, direct, w, b, h, wd)
    assert rel(out, ref) < 1e-5 and rel(pool, pa) < 1e-7
    ((out * dout).sum() + ((pool * dpool).sum() if use_pool_grad else 0.0)).backward()
    for got, want in zip(ins, D):
        assert rel(got.grad, want.grad) < 2e-5, rel(got.grad, want.grad)


@pytest.mark.parametrize("geom", [(2, 32, 48), (1, 16, 16), (1, 64, 40)])
def test_bottleneck_function(geom):
    from transmar_gan_b200 import ops
    B, H, W = geom
    g = torch.Generator().manual_seed(4)
    mk = lambda *s, sc=1.0: (sc * torch.randn(*s, generator=g)).cuda().requires_grad_(True)
    c0, e1 = mk(B, H * W, 64), mk(B, H * W, 64)
    p0, e2 = mk(B, H * W // 4, 64), mk(B, H * W // 4, 64)
    p00, e3 = mk(B, H * W // 16, 64), mk(B, H * W // 16, 64)
    dw_w, dw_b = mk(84, 1, 3, 3, sc=0.3), mk(84, sc=0.1)
    pw_w, pw_b = mk(64, 84, sc=0.1), mk(64, sc=0.1)
    ln_w = (1 + 0.1 * torch.randn(64, generator=g)).cuda().requires_grad_(True)
    ln_b = mk(64, sc=0.1)
    dout = torch.randn(B, H * W, 64, generator=g).cuda()
    ins = [c0, p0, p00, e1, e2, e3, dw_w, dw_b, pw_w, pw_b, ln_w, ln_b]
    D = [t.detach().double().requires_grad_(True) for t in ins]
    from einops import rearrange
    xs = []
    for i, (s, f) in enumerate(((D[0], D[3]), (D[1], D[4]), (D[2], D[5]))):
        x = f + F.leaky_relu(s, 0.01)
        h, w = H // 2 ** i, W // 2 ** i
        xs.append(rearrange(x, "b (h w) (c ch cw) -> b (h ch w cw) c", h=h, ch=2 ** i, cw=2 ** i))
    cat = torch.cat(xs, dim=-1)
    x = cat.transpose(1, 2).reshape(B, 84, H, W)
    x = F.gelu(F.conv2d(x, D[6], D[7], padding=1, groups=84)).flatten(2).transpose(1, 2)
    ref = F.layer_norm(F.linear(x, D[8], D[9]), (64,), D[10], D[11], 1e-5) + D[3]
    ref.backward(dout.double())
    out = ops.BottleneckFunction.apply(*ins, H, W)
    assert rel(out, ref) < 1e-5
    out.backward(dout)
    names = "dc0 dp0 dp00 de1 de2 de3 d_dw_w d_dw_b d_pw_w d_pw_b d_ln_w d_ln_b".split()
    for got, want, name in zip(ins, D, names):
        assert rel(got.grad, want.grad) < 3e-5, (name, rel(got.grad, want.grad))
