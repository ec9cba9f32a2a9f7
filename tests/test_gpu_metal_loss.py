"""GPU parity of metal_loss.cu against the stock-PyTorch loss functions of gan.py
(themselves pinned to losses/gan_losses.py by tests/test_gan_stock.py), values and dL/dfake, and of the two
GANTrainStep options built on it (fused_losses, reuse_fake)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

LAM = dict(rec=1.0, edge=0.2, metal=0.5)          # training/train_combined.py:108-113


def _images(n, H, W, seed, blobs=True):
    g = torch.Generator().manual_seed(seed)
    fake = torch.rand(n, 1, H, W, generator=g) * 2 - 1
    real = torch.rand(n, 1, H, W, generator=g) * 2 - 1
    ct = torch.rand(n, 1, H, W, generator=g) * 1.5 - 1            # a few pixels above the 0.6 threshold on their own
    if blobs:                                                      # "metal": interior, touching a corner, last column
        ct[0, :, H // 3:H // 3 + 6, W // 2:W // 2 + 5] = 1.0
        ct[-1, :, :3, :3] = 1.0
        ct[-1, :, H - 2:, W - 1:] = 1.0
    real[0, :, 0, :2] = fake[0, :, 0, :2]                          # exact zeros of fake - real: sign(0) = 0
    return fake.cuda(), real.cuda(), ct.cuda()


def _stock(fake, real, ct, lam, threshold, beta, radius, w_max):
    from transmar_gan_b200 import gan
    M = gan.extract_metal_mask(ct, threshold)
    w = gan.compute_weight_map(ct, beta, radius, w_max, threshold)
    rec = gan.compute_metal_aware_loss(fake, real, ct, beta, radius, w_max, threshold)
    edge = gan.compute_metal_aware_edge_loss(fake, real, w)
    metal = gan.metal_consistency_loss(fake, real, M)
    return lam["rec"] * rec + lam["edge"] * edge + lam["metal"] * metal, torch.stack([rec, edge, metal])


@pytest.mark.parametrize("n,H,W,radius,beta,w_max,blobs", [
    (2, 128, 128, 5, 1.0, 3.0, True),       # the training configuration (train_combined.py:116-119)
    (3, 100, 70, 5, 1.0, 3.0, True),        # ragged tiles
    (1, 2, 2, 5, 1.0, 3.0, False),          # smallest legal image: every difference is the replicated one
    (2, 33, 65, 0, 2.5, 3.0, True),         # no dilation; 1 + beta > w_max exercises the clamp
    (2, 64, 96, 12, 0.5, 3.0, True),        # wide dilation across tile borders
    (1, 416, 416, 5, 1.0, 3.0, True),       # test-slice size
])
def test_metal_losses_match_stock(n, H, W, radius, beta, w_max, blobs):
    from transmar_gan_b200 import ops
    fake, real, ct = _images(n, H, W, seed=11 + H, blobs=blobs)
    f0 = fake.clone().requires_grad_(True)
    tot0, parts0 = _stock(f0, real, ct, LAM, 0.6, beta, radius, w_max)
    (tot0 * 1.7).backward()
    f1 = fake.clone().requires_grad_(True)
    tot1, parts1 = ops.MetalLossFunction.apply(f1, real, ct, LAM["rec"], LAM["edge"], LAM["metal"], 0.6, beta,
                                               radius, w_max)
    assert not parts1.requires_grad
    (tot1 * 1.7).backward()
    # fp32 sums in another order
    assert torch.allclose(parts1, parts0.detach(), rtol=2e-5, atol=1e-7), (parts1, parts0)
    assert abs(float(tot1.detach()) - float(tot0.detach())) <= 2e-5 * abs(float(tot0.detach()))
    # the gradient is +-weights / N: exact up to the rounding of a handful of products
    scale = float(f0.grad.abs().max())
    assert float((f1.grad - f0.grad).abs().max()) <= 1e-6 * scale + 1e-12, \
        float((f1.grad - f0.grad).abs().max()) / scale


def test_metal_losses_rejects_bad_arguments():
    from transmar_gan_b200 import ops
    fake, real, ct = _images(1, 16, 16, seed=3)
    with pytest.raises(RuntimeError):
        ops.MetalLossFunction.apply(fake, real, ct, 1.0, 0.2, 0.5, 0.6, 1.0, 33, 3.0)          # radius > 32
    with pytest.raises(RuntimeError):
        ops.MetalLossFunction.apply(fake, real[:, :, :8], ct, 1.0, 0.2, 0.5, 0.6, 1.0, 5, 3.0)  # shape mismatch
    with pytest.raises(RuntimeError):
        ops.MetalLossFunction.apply(fake.cpu(), real.cpu(), ct.cpu(), 1.0, 0.2, 0.5, 0.6, 1.0, 5, 3.0)  # no CPU path
    with pytest.raises(RuntimeError):
        f1 = fake[:, :, :1].contiguous()
        ops.MetalLossFunction.apply(f1, f1, f1, 1.0, 0.2, 0.5, 0.6, 1.0, 5, 3.0)              # H < 2


def _gan_step(**kw):
    import transmar_gan_b200 as T
    from transmar_gan_b200 import gan, train
    from oracle import ngswin_oracle as O
    g = T.NGswin()
    g.load_state_dict(O.synthetic_state_dict("W1", seed=1), strict=True)
    torch.manual_seed(5)
    d = gan.MultiScaleDiscriminator(2, 16, 5, 3, True)
    d.apply(gan.weights_init)
    return train.GANTrainStep(g.cuda().train(), d.cuda().train(), **kw)


def _gan_inputs():
    from oracle import ngswin_oracle as O
    ct = O.synthetic_input(2, 128, 128, seed=999)
    ct[:, :, 30:40, 50:60] = 1.0
    return ct.cuda(), O.synthetic_input(2, 128, 128, seed=1000).cuda()


def test_gan_step_fused_losses_match_stock_losses():
    """One eager step: same losses, and the generator gradient that reaches the flat bucket agrees to the
    run-to-run noise of the bf16 backward (the loss gradient itself is exact, see above)."""
    ct, real = _gan_inputs()
    a, b = _gan_step(graph=False, fused_losses=False), _gan_step(graph=False, fused_losses=True)
    la, lb = a(ct, real), b(ct, real)
    assert abs(float(la[0]) - float(lb[0])) < 1e-5
    assert abs(float(la[1]) - float(lb[1])) < 1e-4 * abs(float(la[1])), (la, lb)
    ga, gb = a.bucketG.flat, b.bucketG.flat
    assert float((ga - gb).norm()) <= 2e-2 * float(ga.norm())


@pytest.mark.parametrize("graph", [False, True])
def test_gan_step_reuse_fake_follows_two_forward_recipe(graph):
    ct, real = _gan_inputs()
    a, b = _gan_step(graph=graph, reuse_fake=False), _gan_step(graph=graph, reuse_fake=True)
    la = [tuple(float(v) for v in a(ct, real)) for _ in range(3)]
    lb = [tuple(float(v) for v in b(ct, real)) for _ in range(3)]
    # first step: same weights, the two forwards differ only in kernel path (inference vs training kernels)
    assert abs(la[0][0] - lb[0][0]) < 2e-3 * max(1.0, abs(la[0][0])), (la, lb)
    assert abs(la[0][1] - lb[0][1]) < 1e-3 * abs(la[0][1]), (la, lb)
    for (x, y), (u, v) in zip(la, lb):
        assert abs(x - u) < 5e-3 * max(1.0, abs(x)) and abs(y - v) < 5e-3 * max(1.0, abs(y)), (la, lb)
    assert b._fake is None                               # the retained forward is released after the G step
