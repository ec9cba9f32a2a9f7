"""GPU parity of radon.cu with oracle/radon_oracle.py (float64 restatement; itself checked
against the analytic Radon transform in tests/test_radon_oracle.py), plus the size-independent properties at the
training size, the autograd contract of radon.Radon and the physics term of the GAN step."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _close(got: torch.Tensor, want: np.ndarray):
    """fp32 kernel vs float64 oracle: everything agrees to the fp32 rounding of a ~180-term sum (the mean gate).  A
    ray whose clipped length sat on a step-count boundary would be cut into one step more or fewer and, on a noise
    image, move by a few percent; the step rule's 0.001 margin keeps the integer-length axis-aligned rays off that
    boundary, and the max gate would catch a regression there."""
    want = torch.from_numpy(want).float()
    err = (got.cpu() - want).abs()
    scale = float(want.abs().max()) + 1e-12
    assert float(err.max()) <= 3e-3 * scale, float(err.max()) / scale
    assert float(err.mean()) <= 2e-5 * scale, float(err.mean()) / scale


@pytest.mark.parametrize("H,A,det,spacing", [(32, 16, -1, 1.0), (48, 20, 64, 1.0), (33, 7, 40, 1.5), (64, 180, -1, 1.0)])
def test_forward_and_backprojection_match_oracle(H, A, det, spacing):
    from oracle import radon_oracle as R
    from transmar_gan_b200 import radon
    rng = np.random.default_rng(H)
    angles = np.linspace(0, np.pi, A, dtype=np.float32)
    proj = radon.Radon(H, angles, det_count=det, det_spacing=spacing)
    x = rng.random((2, H, H)).astype(np.float32) * 2 - 1
    sino = proj.forward(torch.from_numpy(x).cuda())
    assert sino.shape == (2, A, proj.det_count)
    _close(sino, np.stack([R.radon_forward(i, angles, proj.det_count, spacing) for i in x]))
    y = rng.random((2, A, proj.det_count)).astype(np.float32) * 2 - 1
    img = proj.backprojection(torch.from_numpy(y).cuda())
    _close(img, np.stack([R.radon_backprojection(s, angles, H, H, spacing) for s in y]))


def test_large_image_takes_the_global_memory_path():
    """416 x 416 x 4 B = 692 KB does not fit in shared memory: same arithmetic through L1/L2."""
    from oracle import radon_oracle as R
    from transmar_gan_b200 import radon
    rng = np.random.default_rng(5)
    angles = np.linspace(0, np.pi, 6, dtype=np.float32)
    x = rng.random((1, 416, 416)).astype(np.float32)
    _close(radon.Radon(416, angles).forward(torch.from_numpy(x).cuda()), R.radon_forward(x[0], angles)[None])


def test_properties_at_training_size():
    """128 x 128, 180 angles (train_combined.py:269-271): analytic disc, mass per angle, linearity, adjointness,
    bitwise repeatability."""
    from transmar_gan_b200 import radon
    S, A = 128, 180
    angles = np.linspace(0, np.pi, A, dtype=np.float32)
    proj = radon.Radon(S, angles)
    yy, xx = np.meshgrid(np.arange(S) + 0.5 - S / 2, np.arange(S) + 0.5 - S / 2, indexing="ij")
    cx, cy, r = 11.0, -20.0, 25.0
    disc = torch.from_numpy((((xx - cx) ** 2 + (yy - cy) ** 2) <= r * r).astype(np.float32)).cuda()
    sino = proj.forward(disc[None])[0].cpu().numpy()
    s = np.arange(S) - S / 2 + 0.5
    s0 = cx * np.cos(angles)[:, None] + cy * np.sin(angles)[:, None]
    chord = 2 * np.sqrt(np.maximum(r * r - (s[None] - s0) ** 2, 0))
    inside = np.abs(s[None] - s0) < r - 2
    assert np.abs(sino - chord)[inside].max() < 1.2
    assert np.abs(sino.sum(1) - float(disc.sum())).max() < 0.01 * float(disc.sum())
    g = torch.Generator().manual_seed(0)
    a, b = torch.rand(4, S, S, generator=g).cuda(), torch.rand(4, S, S, generator=g).cuda()
    pa, pb = proj.forward(a), proj.forward(b)
    assert float((proj.forward(2 * a - 3 * b) - (2 * pa - 3 * pb)).abs().max()) < 2e-3
    assert torch.equal(pa, proj.forward(a))
    y = torch.rand(4, A, S, generator=g).cuda()
    mask = disc * 0 + torch.from_numpy(((xx ** 2 + yy ** 2) <= 60 ** 2).astype(np.float32)).cuda()
    lhs, rhs = float((proj.forward(a * mask) * y).sum()), float((a * mask * proj.backprojection(y)).sum())
    assert abs(lhs - rhs) < 0.02 * abs(lhs)
    assert torch.equal(proj.backprojection(y), proj.backward(y))


def test_autograd_contract_and_physics_loss():
    from oracle import radon_oracle as R
    from transmar_gan_b200 import gan, radon
    S, A = 32, 12
    angles = np.linspace(0, np.pi, A, dtype=np.float32)
    proj = radon.Radon(S, angles)
    g = torch.Generator().manual_seed(3)
    fake = (torch.rand(2, 1, S, S, generator=g) * 2 - 1).cuda().requires_grad_(True)
    real = (torch.rand(2, 1, S, S, generator=g) * 2 - 1).cuda()
    ct = torch.rand(2, 1, S, S, generator=g) * 1.2 - 1
    ct[:, :, 10:14, 18:21] = 1.0
    M = gan.extract_metal_mask(ct.cuda(), 0.6)
    loss = gan.physics_loss_syn(fake, real, M, proj)
    want = R.physics_loss_syn(fake.detach().cpu().numpy()[:, 0], real.cpu().numpy()[:, 0], M.cpu().numpy()[:, 0], angles)
    assert abs(float(loss.detach()) - want) < 1e-4 * want
    loss.backward()
    # the gradient is the backprojection of the masked sign sinogram / N  (torch-radon: backward of forward)
    with torch.no_grad():
        d = proj.forward(fake[:, 0]) - proj.forward(real[:, 0])
        w = (1.0 - (proj.forward(M[:, 0]) > 0).float()) * torch.sign(d) / d.numel()
        expect = proj.backprojection(w)[:, None]
    assert float((fake.grad - expect).abs().max()) <= 1e-6 * float(expect.abs().max())
    # and the backward of the backprojection is the forward projection
    y = torch.rand(2, A, S, generator=g).cuda().requires_grad_(True)
    k = torch.rand(2, S, S, generator=g).cuda()
    (proj.backprojection(y) * k).sum().backward()
    assert torch.equal(y.grad, proj.forward(k))


def test_bad_arguments_raise():
    from transmar_gan_b200 import radon
    proj = radon.Radon(32, np.linspace(0, np.pi, 8, dtype=np.float32))
    with pytest.raises(RuntimeError):
        proj.forward(torch.zeros(1, 32, 32))                       # CPU tensor: no CPU path
    with pytest.raises(RuntimeError):
        proj.forward(torch.zeros(1, 16, 32).cuda())                # wrong size
    with pytest.raises(RuntimeError):
        proj.backprojection(torch.zeros(1, 7, 32).cuda())          # wrong number of angles
    with pytest.raises(ValueError):
        radon.Radon(32, np.zeros(2000, dtype=np.float32))
    with pytest.raises(NotImplementedError):
        radon.Radon(32, [0.0], clip_to_circle=True)


def test_gan_step_with_physics_term_graph_matches_eager():
    """GANTrainStep(projector=Radon(128, 180 angles)): the reference's full loss (lambda_phys = 0.02) replayed from a
    CUDA graph follows the eager step, and the physics term moves the generator loss."""
    import transmar_gan_b200 as T
    from transmar_gan_b200 import gan, radon, train
    from oracle import ngswin_oracle as O

    def make(graph, with_proj=True):
        g = T.NGswin()
        g.load_state_dict(O.synthetic_state_dict("W1", seed=1), strict=True)
        torch.manual_seed(5)
        d = gan.MultiScaleDiscriminator(2, 16, 5, 3, True)
        d.apply(gan.weights_init)
        proj = radon.Radon(128, np.linspace(0, np.pi, 180, dtype=np.float32)) if with_proj else None
        return train.GANTrainStep(g.cuda().train(), d.cuda().train(), graph=graph, projector=proj)

    ct = 0.5 * O.synthetic_input(2, 128, 128, seed=999)           # below the 0.6 threshold except for the blob:
    ct[:, :, 30:40, 50:60] = 1.0                                   # most rays miss the metal and count
    ct, real = ct.cuda(), O.synthetic_input(2, 128, 128, seed=1000).cuda()
    se, sg, s0 = make(False), make(True), make(False, with_proj=False)
    le = [tuple(float(v) for v in se(ct, real)) for _ in range(3)]
    lg = [tuple(float(v) for v in sg(ct, real)) for _ in range(3)]
    l0 = tuple(float(v) for v in s0(ct, real))
    assert abs(le[0][1] - lg[0][1]) < 1e-3 * abs(le[0][1]), (le, lg)
    for (a, b), (c, d) in zip(le, lg):
        assert abs(a - c) < 5e-3 * max(1.0, abs(a)) and abs(b - d) < 5e-3 * max(1.0, abs(b)), (le, lg)
    assert le[0][1] > l0[1] + 1e-3, (le[0], l0)                    # 0.02 * mean |P(fake) - P(real)| is not small
