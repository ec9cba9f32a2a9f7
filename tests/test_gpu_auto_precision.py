"""GPU parity tests of the DEFAULT precision ("auto"): the whole model against outputs of the unmodified
reference, gate max-abs 1e-3 (north_star) on ALL THREE weight sets, un-relaxed:

  W1  the reference's init recipe                      -> every unit runs on the tensor cores (the benchmarked mode)
  W2  the stress set (weights x5, logit scales to the ln 100 clamp): every block GEMM alone breaks the gate at an
      11-bit operand significand (tools/tf32_ablation.py), so the measured plan routes the
      offending units to the fp32 FMA-pipe kernels instead of computing them out of tolerance
  W3  trained-like weights: the unmodified reference trained from W1 for 400 Adam steps on structured synthetic
      slices (oracle/make_w3.py); goldens are the reference's outputs on the trained weights

Everything goes through the C ABI (NGswin -> ops -> libb200mar.so)."""
import pytest
import torch

from oracle import ngswin_oracle as O

pytestmark = pytest.mark.gpu

SHAPES = [(1, 64, 64), (2, 96, 160), (1, 100, 70), (1, 128, 128)]
GATE = 1e-3

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
def auto_mode(T):
    old = T.ops.set_precision("auto")
    yield
    assert T.ops.tc_error() == 0, "a tensor-core kernel timed out on an mbarrier"
    T.ops.set_precision(old)


def weights(kind, golden_w3):
    return golden_w3["state_dict"] if kind == "W3" else O.synthetic_state_dict(kind, seed=1)


def golden(kind, shape, golden_outputs, golden_w3):
    return golden_w3[("out",) + shape] if kind == "W3" else golden_outputs[(kind,) + shape]


@pytest.fixture(scope="module")
def models(T, golden_w3):
    out = {}
    for kind in ("W1", "W2", "W3"):
        m = T.NGswin()
        m.load_state_dict(weights(kind, golden_w3), strict=True)
        out[kind] = m.cuda().eval()
    return out


@pytest.mark.parametrize("kind", ["W1", "W2", "W3"])
@pytest.mark.parametrize("shape", SHAPES)
def test_default_precision_meets_gate(models, golden_outputs, golden_w3, kind, shape):
    m = models[kind]
    x = O.synthetic_input(*shape, seed=999)
    with torch.no_grad():
        y = m(x.cuda()).cpu()
    ref = golden(kind, shape, golden_outputs, golden_w3)
    err = (y - ref).abs().max().item()
    rep = m.precision_report
    print(f"[auto] {kind} {shape}: max-abs {err:.2e}; {rep['units_on_tensor_cores']}/{rep['units']} units on "
          f"tcgen05 (probe err all-tc {rep['err_all_tensor_core']:.2e}, plan {rep['err_plan']:.2e})")
    assert err < GATE, (kind, shape, err, rep)
    t = O.synthetic_input(*shape, seed=1000)
    assert abs(O.psnr(y, t) - O.psnr(ref, t)) < 0.01
    assert abs(O.ssim(y[0, 0], t[0, 0]) - O.ssim(ref[0, 0], t[0, 0])) < 1e-4


def test_plan_w1_all_tensor_cores_w2_not(models):
    """The benchmarked configuration (W1) runs every unit on tcgen05 at 11-bit operands; the stress set cannot (its
    all-11-bit probe error exceeds the calibration tolerance) and moves units to the compensated level / the fp32
    kernels; every plan's own probe error is within the tolerance."""
    x = O.synthetic_input(1, 64, 64, seed=999).cuda()
    for kind in ("W1", "W2", "W3"):
        with torch.no_grad():
            models[kind](x)
    r1, r2, r3 = (models[k].precision_report for k in ("W1", "W2", "W3"))
    assert r1["units_on_tensor_cores"] == r1["units"] == 41 and r1["units_compensated"] == 0, r1
    assert r2["err_all_tensor_core"] > r2["tol"], r2
    assert sum(p == "tf32" for p in r2["plan"]) < 41 and r2["units_compensated"] + 41 - r2["units_on_tensor_cores"] >= 3, r2
    for r in (r1, r2, r3):
        assert r["err_plan"] <= r["tol"]
    print(f"[auto] W3 plan: {r3['units_on_tensor_cores']}/41 units on tcgen05, probe err all-tc "
          f"{r3['err_all_tensor_core']:.2e}, plan {r3['err_plan']:.2e}")


def test_trained_like_weights_on_structured_input(T, models, golden_w3):
    """W3 on the structured phantom slices it was trained on (the reference's outputs are in the fixture).
    Piecewise-constant content is harder for 11-bit operands than uniform noise: with EVERY unit forced onto
    tcgen05 the error is ~1.3e-3 (characterised here, bounded at 3e-3; uniform inputs stay within the gate);
    the default mode - whose calibration probe contains a structured slice - must meet the 1e-3 gate."""
    x = golden_w3["phantom_in"].cuda()
    ref = golden_w3["phantom_out"]
    with torch.no_grad():
        y = models["W3"](x).cpu()
    err_auto = (y - ref).abs().max().item()
    old = T.ops.set_precision("tf32")
    try:
        for shape in SHAPES:
            with torch.no_grad():
                yu = models["W3"](O.synthetic_input(*shape, seed=999).cuda()).cpu()
            assert (yu - golden_w3[("out",) + shape]).abs().max().item() < GATE, shape
        with torch.no_grad():
            yf = models["W3"](x).cpu()
    finally:
        T.ops.set_precision(old)
    err_forced = (yf - ref).abs().max().item()
    rep = models["W3"].precision_report
    print(f"[auto] W3 phantom: default mode {err_auto:.2e} ({rep['units_on_tensor_cores']}/41 units on tcgen05), "
          f"all units forced onto tcgen05 {err_forced:.2e}")
    assert err_auto < GATE, (err_auto, rep)
    assert err_forced < 3e-3, err_forced


def test_mixed_plan_plumbing(models, golden_outputs):
    """A hand-made alternating plan (tcgen05 / fp32 units interleaved, glue on fp32) exercises every hand-over
    between the in-place tensor-core blocks and the out-of-place fp32 blocks."""
    m = models["W1"]
    # 41 units: (attention, MLP) of each of the 20 blocks, then the glue; the pattern of period 3 exercises all four
    # combinations per block (tc/tc, tc/fp32, fp32/tc, fp32/fp32) and their hand-overs
    plan = ["tf32" if i % 3 != 1 else "fp32" for i in range(40)] + ["fp32"]
    for shape in SHAPES[:3]:
        x = O.synthetic_input(*shape, seed=999).cuda()
        with torch.no_grad():
            m._set_plan(plan)
            y = m._forward_planned(x).cpu()
            m._set_plan([p == "tf32" and "fp32" or "tf32" for p in plan])
            y2 = m._forward_planned(x).cpu()
        ref = golden_outputs[("W1",) + shape]
        assert (y - ref).abs().max().item() < GATE and (y2 - ref).abs().max().item() < GATE
        with torch.no_grad():
            assert torch.equal(m._forward_planned(x).cpu(), y2)      # repeatable, input not clobbered
        # all three levels interleaved (period 3 against the (attention, MLP) period 2: every pairing occurs)
        with torch.no_grad():
            m._set_plan([("tf32", "tf32c", "fp32")[i % 3] for i in range(40)] + ["tf32"])
            y3 = m._forward_planned(x).cpu()
            assert torch.equal(m._forward_planned(x).cpu(), y3)
        assert (y3 - ref).abs().max().item() < GATE


def test_stress_set_at_slice_size(T):
    """W2 at the SynDeepLesion slice size through the default mode; the oracle in float64 on the GPU checks."""
    sd = O.synthetic_state_dict("W2", seed=3)
    m = T.NGswin()
    m.load_state_dict(sd, strict=True)
    m = m.cuda().eval()
    x = O.synthetic_input(1, 416, 416, seed=999).cuda()
    with torch.no_grad():
        y = m(x)
        ref = O.ngswin_forward(x.double(), {k: (v.cuda().double() if v.is_floating_point() else v.cuda())
                                            for k, v in sd.items()})
    assert (y.double() - ref).abs().max().item() < GATE


def test_recalibrates_when_weights_change(T):
    m = T.NGswin()
    m.load_state_dict(O.synthetic_state_dict("W1", seed=1), strict=True)
    m = m.cuda().eval()
    x = O.synthetic_input(1, 64, 64, seed=999).cuda()
    with torch.no_grad():
        m(x)
        assert m.precision_report["units_on_tensor_cores"] == 41
        m.load_state_dict(O.synthetic_state_dict("W2", seed=1), strict=True)
        m(x)
    assert sum(p == "tf32" for p in m.precision_report["plan"]) < 41


def test_training_in_default_precision_matches_reference_gradients(T, golden_w3):
    """fwd + L1 + bwd on W3 in the default mode against the unmodified reference's loss and gradients."""
    m = T.NGswin()
    m.load_state_dict(golden_w3["state_dict"], strict=True)
    m = m.cuda().train()
    x = O.synthetic_input(2, 64, 64, seed=999).cuda().requires_grad_(True)
    t = O.synthetic_input(2, 64, 64, seed=1000).cuda()
    loss = torch.nn.functional.l1_loss(m(x), t)
    loss.backward()
    assert abs(loss.item() - golden_w3["l1_loss"].item()) < 1e-4
    named = dict(m.named_parameters())
    rel = sorted(((named[n].grad.cpu() - g).norm() / g.norm().clamp_min(1e-12)).item()
                 for n, g in golden_w3["grads"].items() if g.norm() > 1e-10)
    gx = ((x.grad.cpu() - golden_w3["grad_x"]).norm() / golden_w3["grad_x"].norm()).item()
    print(f"[auto/train] W3 grads: dx rel {gx:.2e}, param median {rel[len(rel) // 2]:.2e}, worst {rel[-1]:.2e}")
    assert gx < 5e-2 and rel[len(rel) // 2] < 5e-2 and rel[-1] < 0.25     # bf16-operand backward
