"""N3: the photometric loss.  The oracle's numpy restatement (oracle/gs_oracle.py: ssim_window,
photometric_loss) against tests/golden/loss.npz, which tests/golden/make_golden.py::loss_fixture produced by
calling the reference's own utils/loss_utils.py (l1_loss, ssim) and autograd on the CPU."""
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(HERE, "golden", "loss.npz"))


def close_grad(got, want, n_elems, rel=1e-4):
    """Relative L1 <= rel where the gradient is not pure rounding noise, and no single entry further than
    rel of the largest entry (or of the 1/N scale every entry of this loss's gradient lives on)."""
    got, want = np.asarray(got, np.float64), np.asarray(want, np.float64)
    scale = max(np.abs(want).max(), 1.0 / n_elems)
    assert np.abs(got - want).max() <= rel * scale, (np.abs(got - want).max(), scale)
    if np.abs(want).sum() > 1e-3 * scale * want.size:
        assert np.abs(got - want).sum() / np.abs(want).sum() <= rel


def test_window_is_the_reference_window(oracle):
    w, w2d = oracle.ssim_window()
    assert w.dtype == np.float32 and w.shape == (11,) and w2d.shape == (11, 11)
    assert abs(float(w.sum()) - 1.0) < 1e-6 and np.array_equal(w, w[::-1])
    np.testing.assert_array_equal(w2d, np.outer(w, w).astype(np.float32))
    # known answer: centre tap of a sigma-1.5, 11-tap normalised Gaussian
    assert abs(float(w[5]) - 0.26601171) < 1e-6


def test_oracle_loss_matches_the_reference(oracle, golden):
    lam = float(golden["lambda_dssim"])
    assert lam == 0.2                                    # arguments/__init__.py OptimizationParams default
    for n in range(int(golden["cases"])):
        x, y = golden[f"image{n}"], golden[f"gt{n}"]
        r = oracle.photometric_loss(x, y, lam)
        np.testing.assert_allclose(r["l1"], golden[f"l1_{n}"], rtol=2e-6, atol=1e-8)
        # binary32 sigma = E[x^2] - mu^2 cancels on smooth images: the reference's own value carries ~2e-5 noise
        np.testing.assert_allclose(r["ssim"], golden[f"ssim{n}"], rtol=1e-4)
        np.testing.assert_allclose(r["loss"], golden[f"loss{n}"], rtol=1e-4, atol=1e-7)
        close_grad(r["grad"], golden[f"grad{n}"], x.size)
        close_grad(r["grad_ssim"], golden[f"grad_ssim{n}"], x.size)


def test_oracle_gradient_is_the_derivative_of_the_oracle_loss(oracle):
    rng = np.random.default_rng(3)
    y = rng.random((2, 13, 17))
    x = np.clip(y + 0.05 * rng.normal(size=y.shape), 0.01, 0.99)
    r = oracle.photometric_loss(x, y, 0.2)
    for _ in range(8):
        i = tuple(rng.integers(0, s) for s in x.shape)
        h = 1e-6
        xp, xm = x.copy(), x.copy()
        xp[i] += h
        xm[i] -= h
        num = (oracle.photometric_loss(xp, y, 0.2)["loss"] - oracle.photometric_loss(xm, y, 0.2)["loss"]) / (2 * h)
        assert abs(num - r["grad"][i]) <= 1e-6 * max(abs(num), 1.0 / x.size), (i, num, r["grad"][i])


def test_identical_images(oracle):
    y = np.random.default_rng(1).random((3, 20, 20))
    r = oracle.photometric_loss(y, y, 0.2)
    assert r["l1"] == 0.0 and abs(r["ssim"] - 1.0) < 1e-12 and abs(r["loss"]) < 1e-12
    assert np.abs(r["grad"]).max() < 1e-12               # SSIM is maximal and sign(0) = 0


def test_product_has_no_cpu_path():
    import torch
    from mvs_gaussian_splatting_b200 import loss
    with pytest.raises(RuntimeError, match="no CPU path"):
        loss.l1_dssim_loss(torch.rand(3, 8, 8), torch.rand(3, 8, 8))
