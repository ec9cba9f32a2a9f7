"""CPU checks that oracle/radon_oracle.py is a Radon transform (torch-radon itself is an absent submodule of the
reference, so there is no golden sinogram to pin against: SURVEY.md §8f rank 4, parity unpinned)."""
import numpy as np

from oracle import radon_oracle as R


def _disc(H, W, cy, cx, r, value=1.0):
    y, x = np.meshgrid(np.arange(H) + 0.5 - 0.5 * H, np.arange(W) + 0.5 - 0.5 * W, indexing="ij")
    return value * (((x - cx) ** 2 + (y - cy) ** 2) <= r * r).astype(np.float64)


def test_disc_projection_is_the_analytic_chord_length():
    H = W = 64
    angles = np.linspace(0, np.pi, 12, dtype=np.float32)
    cx, cy, r = 5.0, -7.0, 14.0
    sino = R.radon_forward(_disc(H, W, cy, cx, r, 0.8), angles)
    s = np.arange(W) - 0.5 * W + 0.5
    for a, th in enumerate(angles):
        s0 = cx * np.cos(th) + cy * np.sin(th)                      # the centre's detector coordinate
        chord = 0.8 * 2 * np.sqrt(np.maximum(r * r - (s - s0) ** 2, 0.0))
        inside = np.abs(s - s0) < r - 2                              # away from the pixelated rim
        assert np.abs(sino[a] - chord)[inside].max() < 1.2           # a pixel of chord at most
        assert abs(sino[a].sum() - chord.sum()) < 0.01 * chord.sum()
        assert np.abs(sino[a][np.abs(s - s0) > r + 2]).max() == 0.0  # rays that miss the disc


def test_mass_is_conserved_at_every_angle_and_the_operator_is_linear():
    rng = np.random.default_rng(0)
    H, W = 32, 32
    a = rng.random((H, W)) * _disc(H, W, 0, 0, 13)                  # support inside every view
    b = rng.random((H, W)) * _disc(H, W, 0, 0, 13)
    angles = np.linspace(0, np.pi, 9, dtype=np.float32)
    pa, pb = R.radon_forward(a, angles), R.radon_forward(b, angles)
    assert np.abs(pa.sum(1) - a.sum()).max() < 0.01 * a.sum()
    assert np.abs(R.radon_forward(2 * a - 3 * b, angles) - (2 * pa - 3 * pb)).max() < 1e-10
    # axis-aligned views of a pixel image are exact column / row sums
    assert np.abs(R.radon_forward(a, np.array([0.0]))[0] - a.sum(0)).max() < 1e-9
    assert np.abs(R.radon_forward(a, np.array([np.pi / 2]))[0] - a.sum(1)).max() < 1e-9


def test_backprojection_is_the_adjoint_up_to_discretisation():
    rng = np.random.default_rng(1)
    H = W = 32
    x = rng.random((H, W)) * _disc(H, W, 0, 0, 13)
    angles = np.linspace(0, np.pi, 16, dtype=np.float32)
    y = rng.random((16, W))
    lhs = (R.radon_forward(x, angles) * y).sum()
    rhs = (x * R.radon_backprojection(y, angles, H, W)).sum()
    assert abs(lhs - rhs) < 0.02 * abs(lhs)


def test_ragged_shapes_and_detector_options():
    img = np.ones((20, 36))
    sino = R.radon_forward(img, np.array([0.0, np.pi / 2]), det_count=48, det_spacing=1.0)
    assert sino.shape == (2, 48)
    assert np.allclose(sino[0, 6:42], 20.0) and np.allclose(sino[0, :6], 0) and np.allclose(sino[0, 42:], 0)
    assert np.allclose(sino[1, 14:34], 36.0, rtol=2e-3) and np.allclose(sino[1, :14], 0)   # step < 1: rim samples
    wide = R.radon_forward(img, np.array([0.0]), det_count=18, det_spacing=2.0)   # detector lines between columns
    assert np.allclose(wide[0], 20.0)


def test_physics_loss_ignores_metal_rays():
    rng = np.random.default_rng(2)
    H = W = 32
    real = rng.random((1, H, W)) * _disc(H, W, 0, 0, 13)
    M = _disc(H, W, 2, -3, 2.5)[None]
    fake = real + 5.0 * M                                             # differs from real only inside the metal
    angles = np.linspace(0, np.pi, 8, dtype=np.float32)
    assert R.physics_loss_syn(fake, real, M, angles) < 1e-12          # every ray that sees the change is masked
    assert R.physics_loss_syn(real + 0.1 * _disc(H, W, -6, 6, 2)[None], real, M, angles) > 1e-4
