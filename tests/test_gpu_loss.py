"""N3 on the GPU: the fused photometric loss (csrc/gsr_loss.cu through the C ABI) against
(1) tests/golden/loss.npz — the reference's own utils/loss_utils.py + autograd, (2) the oracle at sizes that
span tile borders, (3) at 1920x1080, torch's conv2d formulation of the same formula on the same device."""
import os
import sys

import numpy as np
import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from test_loss_cpu import close_grad  # noqa: E402

pytestmark = pytest.mark.gpu
TOL = 1e-4   # north_star tolerance for floating-point outputs


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(HERE, "golden", "loss.npz"))


def run(cuda, x, y, lam):
    from mvs_gaussian_splatting_b200 import loss as L
    xi = torch.from_numpy(np.ascontiguousarray(x)).to(cuda).requires_grad_(True)
    yi = torch.from_numpy(np.ascontiguousarray(y)).to(cuda)
    out = L.l1_dssim_loss(xi, yi, lam)
    out.backward()
    return float(out.detach()), xi.grad.cpu().numpy()


def test_matches_the_reference_vectors(cuda, native, golden):
    from mvs_gaussian_splatting_b200 import loss as L
    lam = float(golden["lambda_dssim"])
    for n in range(int(golden["cases"])):
        x, y = golden[f"image{n}"], golden[f"gt{n}"]
        before = native.lib().gsr_launch_count()
        loss, grad = run(cuda, x, y, lam)
        assert native.lib().gsr_launch_count() - before == 2            # one forward, one backward launch
        np.testing.assert_allclose(loss, golden[f"loss{n}"], rtol=TOL, atol=1e-7)
        close_grad(grad, golden[f"grad{n}"], x.size, TOL)
        xi = torch.from_numpy(x).to(cuda).requires_grad_(True)
        yi = torch.from_numpy(y).to(cuda)
        terms = L.photometric_terms(xi, yi, 0.0, 0.0, 1.0)
        np.testing.assert_allclose(terms[1].item(), golden[f"l1_{n}"], rtol=2e-6, atol=1e-8)
        np.testing.assert_allclose(terms[2].item(), golden[f"ssim{n}"], rtol=TOL)
        (3.0 * terms[0]).backward()                                       # upstream gradient != 1
        close_grad(xi.grad.cpu().numpy() / 3.0, golden[f"grad_ssim{n}"], x.size, TOL)
        np.testing.assert_allclose(L.l1_loss(xi, yi).item(), golden[f"l1_{n}"], rtol=2e-6, atol=1e-8)
        np.testing.assert_allclose(L.ssim(xi, yi).item(), golden[f"ssim{n}"], rtol=TOL)


@pytest.mark.parametrize("shape", [(3, 32, 32), (3, 33, 65), (1, 5, 100), (4, 70, 31), (2, 3, 64, 40)])
def test_matches_the_oracle_across_tile_borders(cuda, oracle, shape):
    rng = np.random.default_rng(sum(shape))
    y = rng.random(shape).astype(np.float32)
    x = np.clip(y + 0.1 * rng.normal(size=shape), 0, 1).astype(np.float32)
    loss, grad = run(cuda, x, y, 0.2)
    flat = (-1,) + shape[-2:]
    r = oracle.photometric_loss(x.reshape(flat), y.reshape(flat), 0.2)
    np.testing.assert_allclose(loss, r["loss"], rtol=TOL)
    close_grad(grad.reshape(flat), r["grad"], x.size, TOL)


def test_identical_images_and_no_grad(cuda):
    from mvs_gaussian_splatting_b200 import loss as L
    y = torch.rand(3, 50, 50, device=cuda)
    out = L.l1_dssim_loss(y.clone(), y, 0.2)                             # no input requires grad: forward only
    assert not out.requires_grad and abs(out.item()) < 1e-6
    x = y.clone().requires_grad_(True)
    L.l1_dssim_loss(x, y, 0.2).backward()
    assert x.grad.abs().max().item() < 1e-4 / y.numel()


def test_rejects_bad_arguments(cuda, native):
    from mvs_gaussian_splatting_b200 import loss as L
    with pytest.raises(ValueError):
        L.l1_dssim_loss(torch.rand(3, 8, 8, device=cuda), torch.rand(3, 8, 9, device=cuda))
    with pytest.raises(NotImplementedError):
        L.ssim(torch.rand(3, 8, 8, device=cuda), torch.rand(3, 8, 8, device=cuda), window_size=7)
    lib = native.lib()
    import ctypes
    n = ctypes.c_size_t()
    assert lib.gsr_loss_workspace_bytes(0, 8, 8, 1, ctypes.byref(n)) == 1       # GSR_ERR_INVALID_ARGUMENT
    assert lib.gsr_loss_forward(3, 8, 8, None, None, 0.0, 1.0, 0.0, None, 0, None, None) == 1


def torch_formulation(x, y, lam):
    """The same formula with torch ops (grouped conv2d, zero padding 5), autograd for the gradient."""
    g = torch.tensor([np.exp(-(i - 5) ** 2 / (2 * 1.5 ** 2)) for i in range(11)], dtype=torch.float32)
    g = g / g.sum()
    C = x.shape[0]
    win = (g[:, None] @ g[None, :]).expand(C, 1, 11, 11).contiguous().to(x.device, x.dtype)
    conv = lambda a: torch.nn.functional.conv2d(a, win, padding=5, groups=C)
    mu1, mu2 = conv(x), conv(y)
    s1, s2, s12 = conv(x * x) - mu1 * mu1, conv(y * y) - mu2 * mu2, conv(x * y) - mu1 * mu2
    m = ((2 * mu1 * mu2 + 1e-4) * (2 * s12 + 9e-4)) / ((mu1 * mu1 + mu2 * mu2 + 1e-4) * (s1 + s2 + 9e-4))
    return (1 - lam) * (x - y).abs().mean() + lam * (1 - m.mean())


def test_full_hd_against_the_torch_formulation(cuda):
    from mvs_gaussian_splatting_b200 import loss as L
    gen = torch.Generator(device="cpu").manual_seed(9)
    y = torch.rand(3, 1080, 1920, generator=gen).to(cuda)
    x = (y + 0.1 * torch.randn(3, 1080, 1920, generator=gen).to(cuda)).clamp(0, 1)
    a = x.clone().requires_grad_(True)
    la = L.l1_dssim_loss(a, y, 0.2)
    la.backward()
    b = x.clone().double().requires_grad_(True)
    lb = torch_formulation(b, y.double(), 0.2)                            # binary64: the arbiter
    lb.backward()
    assert abs(la.item() - lb.item()) <= TOL * abs(lb.item())
    rel = ((a.grad.double() - b.grad).abs().sum() / b.grad.abs().sum()).item()
    assert rel <= TOL, rel
    assert (a.grad.double() - b.grad).abs().max().item() <= TOL * b.grad.abs().max().item()


def test_l1_only_path(cuda, native):
    """a_ssim == 0 takes the streaming L1 kernels: same value and gradient as torch, odd sizes and an unaligned view."""
    from mvs_gaussian_splatting_b200 import loss as L
    gen = torch.Generator(device="cpu").manual_seed(4)
    for shape in ((3, 1080, 1920), (3, 37, 53), (1, 1, 1), (2, 5, 7)):
        y = torch.rand(*shape, generator=gen).to(cuda)
        x = (y + 0.1 * torch.randn(*shape, generator=gen).to(cuda)).clamp(0, 1)
        x[..., ::3] = y[..., ::3]
        a = x.clone().requires_grad_(True)
        n0 = native.lib().gsr_launch_count()
        la = L.l1_loss(a, y)
        (2.0 * la).backward()
        assert native.lib().gsr_launch_count() - n0 == 2
        b = x.clone().double().requires_grad_(True)
        lb = (b - y.double()).abs().mean()
        (2.0 * lb).backward()
        assert abs(la.item() - lb.item()) <= 2e-6 * abs(lb.item()) + 1e-9
        np.testing.assert_allclose(a.grad.cpu().numpy(), b.grad.float().cpu().numpy(), rtol=1e-6, atol=0)
    # a view whose storage is not 16-byte aligned takes the scalar loop
    base = torch.rand(3 * 9 * 11 + 1, generator=gen).to(cuda)
    xv, yv = base[1:].view(3, 9, 11), torch.rand(3, 9, 11, generator=gen).to(cuda)
    assert xv.data_ptr() % 16 != 0
    a = xv.requires_grad_(True)
    L.l1_loss(a, yv).backward()
    want = torch.sign(xv.detach() - yv) / xv.numel()
    np.testing.assert_allclose(a.grad.cpu().numpy(), want.cpu().numpy(), rtol=1e-6)
