"""GPU parity tests of the tensor-core BACKWARD kernels (tcgen05, bf16 operands, fp32 accumulate),
called through the C ABI: wgrad_tc, mlp_bwd_tc, attn_bwd_tc (with and without the O tensor saved by
attn_tc_fwd_save), against float64 autograd of the oracle (oracle/ngswin_oracle.py) and against plain
matmuls.

Tolerances (written here, measured on B200): bf16 operands carry an 8-bit significand, so one operand
rounding is 2^-9 relative and a gradient tensor that is a sum over thousands of tokens lands at
3e-3 .. 1e-2 relative L2 (W2, logit scales up to the ln 100 clamp: 1.2e-2); the gate is 3e-2 per tensor.
logit_scale is a sum of strongly cancelling terms (rows of dS sum to zero) and is gated at 0.3.
The fp32 FMA-pipe backward (tests/test_gpu_parity.py) holds 2e-4 on the same cases.
"""
import pytest
import torch

from oracle import ngswin_oracle as O

pytestmark = pytest.mark.gpu


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


def relerr(a, b):
    return ((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30)).item()


@pytest.mark.parametrize("n_tok,cin,cout", [(64, 64, 64), (128, 64, 128), (192, 64, 192), (4160, 128, 64),
                                           (148 * 3 * 128 + 64, 64, 192), (1000, 128, 256)])
def test_wgrad_tc_vs_matmul(T, n_tok, cin, cout):
    """dw += g^T x with the tokens as the MMA K dimension (ragged last tile zero-filled by TMA); the
    bias gradient (column sums of g) comes out of the same MMAs through the ones block (lane 64)."""
    g = torch.Generator().manual_seed(n_tok)
    x = torch.randn(n_tok, cin, generator=g).cuda().bfloat16()
    gy = torch.randn(n_tok, cout, generator=g).cuda().bfloat16()
    dw = torch.ones(cout, cin, device="cuda")              # accumulates into what is there
    db = torch.ones(cout, device="cuda") if cin == 64 else None
    T.ops.wgrad_tc(x, gy, dw, db)
    ref = gy.double().T @ x.double() + 1.0
    assert relerr(dw, ref) < 1e-5
    if db is not None:
        assert relerr(db, gy.double().sum(0) + 1.0) < 1e-5


CASES = [("encoder_layer1.blocks.0.", 6, 0), ("encoder_layer1.blocks.1.", 6, 4),
         ("encoder_layer2.blocks.0.", 4, 0), ("encoder_layer3.blocks.1.", 4, 4)]


@pytest.mark.parametrize("kind", ["W1", "W2"])
@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("geom", [(2, 16, 24), (1, 24, 40), (3, 24, 24), (1, 64, 64)])
def test_nstb_tc_backward_vs_oracle_autograd(T, kind, case, geom):
    """One NSTB through NSTBFunction in the tensor-core mode (tcgen05 forward that saves O, mlp_bwd_tc,
    attn_bwd_tc core + dx kernels, wgrad_tc) against float64 autograd of the oracle.  (3, 24, 24) has an odd
    window count (half-empty last tile); shift 4 exercises the analytic mask and the wrapped gather."""
    prefix, nh, shift = case
    B, ph, pw = geom
    sd = O.synthetic_state_dict(kind, seed=1)
    g = torch.Generator().manual_seed(21)
    z = torch.randn(B, ph * pw, 64, generator=g)
    w_out = torch.randn(B, ph * pw, 64, generator=g)
    w_zs = torch.randn(B, ph * pw, 64, generator=g)
    sd64 = {k: (v.cuda().double().requires_grad_(True) if k.startswith(prefix) and v.is_floating_point()
                and not k.endswith("attn_mask") else v.cuda()) for k, v in sd.items() if k.startswith(prefix)}
    z64 = z.cuda().double().requires_grad_(True)
    out64 = O.nstb(z64, ph, pw, sd64, prefix, nh, shift)
    loss64 = (out64 * w_out.cuda().double()).sum() + ((out64 + z64) * w_zs.cuda().double()).sum()
    loss64.backward()
    P = [sd[prefix + n].cuda().contiguous().requires_grad_(True) for n in T.ops.NSTB_PARAM_NAMES]
    zc = z.cuda().requires_grad_(True)
    out, zsum = T.ops.NSTBFunction.apply(zc, ph, pw, nh, shift, True, True, True, *P)
    loss = (out * w_out.cuda()).sum() + (zsum * w_zs.cuda()).sum()
    loss.backward()
    assert relerr(zc.grad, z64.grad) < 3e-2
    for n, p in zip(T.ops.NSTB_PARAM_NAMES, P):
        gref = sd64[prefix + n].grad
        assert p.grad is not None, n
        if gref.norm().item() < 1e-9:                      # clamped logit_scale entries: exact zeros
            assert p.grad.abs().max().item() < 1e-6, n
        else:
            tol = 0.3 if n.endswith("logit_scale") else 3e-2
            assert relerr(p.grad, gref) < tol, (n, relerr(p.grad, gref))


@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("geom", [(1, 24, 40), (3, 8, 24)])
def test_attn_bwd_tc_saved_o_matches_recompute(T, case, geom):
    """attn_bwd_tc with the O tensor written by attn_tc_fwd_save agrees with its own recomputation pass
    (the two differ only by the rounding of O: tf32 P V in the forward, bf16 P V in the recomputation),
    and attn_tc_fwd_save returns the same `a` as attn_tc_fwd bit for bit."""
    prefix, nh, shift = case
    B, ph, pw = geom
    sd = O.synthetic_state_dict("W1", seed=1)
    P = [sd[prefix + n].cuda().contiguous() for n in T.ops.NSTB_PARAM_NAMES][10:18]
    g = torch.Generator().manual_seed(5)
    z = torch.randn(B, ph * pw, 64, generator=g).cuda()
    c = (torch.randn(B, (ph // 8) * (pw // 8), 64, generator=g) * 0.5).cuda()
    da = torch.randn(B, ph * pw, 64, generator=g).cuda()
    a_save, o16 = T.ops.attn_tc_fwd_save(z, c, *P, ph, pw, nh, shift)
    assert torch.equal(a_save, T.ops.attn_tc_fwd(z, c, *P, ph, pw, nh, shift))
    res = []
    for o in (None, o16):
        dz, dctx = torch.empty_like(z), torch.zeros_like(c)
        G = [torch.zeros_like(p) for p in P]
        T.ops.attn_bwd_tc(z, c, da, P, dz, dctx, G, ph, pw, nh, shift, o)
        res.append([dz - da, dctx, *G])
    for x, y in zip(*res):
        assert relerr(x, y) < 2e-2


@pytest.mark.parametrize("n_tok", [64, 128, 192, 4160])
def test_mlp_bwd_tc_vs_fp32_kernel(T, n_tok):
    """mlp_bwd_tc (row kernel + two wgrad launches) against the fp32 FMA-pipe mlp_bwd on the same inputs;
    64 and 192 tokens leave the last 128-row tile half empty."""
    sd = O.synthetic_state_dict("W2", seed=1)
    P = [sd["encoder_layer1.blocks.1." + n].cuda().contiguous() for n in T.ops.NSTB_PARAM_NAMES][18:24]
    g = torch.Generator().manual_seed(n_tok)
    a = torch.randn(n_tok, 64, generator=g).cuda()
    dout = torch.randn(n_tok, 64, generator=g).cuda()
    res = []
    for fn in (T.ops.mlp_bwd, T.ops.mlp_bwd_tc):
        da = torch.empty_like(a)
        G = [torch.zeros_like(p) for p in P]
        fn(a, dout, P, da, G)
        res.append([da, *G])
    for x, y in zip(res[1], res[0]):
        assert relerr(x, y) < 1.5e-2
