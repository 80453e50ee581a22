"""CPU restatement of the reference's pure-PyTorch dense 2-D Gaussian splatting, for TIMING only.

TEST / BENCH INFRASTRUCTURE.  north_star asks for "the pure-PyTorch CPU rasterizer path
(2D-Gaussian-Splatting-main style dense splat compositing) timed on the box's host cores in the same run
with core count stated; that CPU path is a reported baseline only".  The function it means is
`generate_2D_gaussian_splatting` at /root/reference/2D-Gaussian-Splatting-main/2d_gaussian_splatting.py:44-123
(sizes: config.yml:1-11 — 1000 primary samples, 256x256x3 image, 101-tap kernel).  That module cannot be
imported (its body downloads an image and needs matplotlib / imageio), so the function is restated here with
the same tensor program, in the same order:
  :46-58   sigma_x, sigma_y, rho -> batch of 2x2 covariances, PSD check            (:56-58 raises ValueError)
  :60-76   inverse covariance; a [-5,5] grid of kernel_size^2 points; z = -1/2 x^T S^-1 x by einsum
  :77-82   kernel = exp(z) / (2 pi sqrt(det)), normalised by its own maximum
  :84-101  replicate to RGB, zero-pad the kernel to the image size
  :106-114 translate every padded kernel to its coordinate with affine_grid + grid_sample (align_corners=True)
  :116-121 multiply by the colours, SUM over the batch (additive, unsorted, un-occluded), clamp to [0,1], HWC
It is NOT the function the 3-D rasterizer computes (no depth order, no transmittance), so it is a throughput
baseline — ns per splat*pixel — never an accuracy oracle.
"""
from __future__ import annotations

import math
import time


def dense_splat_2d(kernel_size, sigma_x, sigma_y, rho, coords, colours, image_size=(256, 256, 3), device="cpu"):
    import torch
    import torch.nn.functional as F
    n = colours.shape[0]
    sx, sy, r = (t.view(n, 1, 1) for t in (sigma_x, sigma_y, rho))
    off = r * sx * sy
    cov = torch.stack([torch.stack([sx ** 2, off], dim=-1), torch.stack([off, sy ** 2], dim=-1)], dim=-2)
    if ((sx ** 2) * (sy ** 2) - off ** 2 <= 0).any():
        raise ValueError("Covariance matrix must be positive semi-definite")
    inv_cov = torch.inverse(cov)
    axis = -5.0 + 10.0 * torch.linspace(0, 1, steps=kernel_size, device=device).view(1, -1)
    xx = axis.unsqueeze(-1).expand(-1, -1, kernel_size)
    yy = axis.unsqueeze(1).expand(-1, kernel_size, -1)
    xy = torch.stack([xx, yy], dim=-1)
    z = torch.einsum("b...i,b...ij,b...j->b...", xy, -0.5 * inv_cov, xy)
    kernel = torch.exp(z) / (2 * math.pi * torch.sqrt(torch.det(cov)).view(n, 1, 1))
    kernel = kernel / kernel.amax(dim=(-2, -1), keepdim=True)
    rgb = kernel.repeat(1, 3, 1).view(n * 3, kernel_size, kernel_size).unsqueeze(0).reshape(n, 3, kernel_size, kernel_size)
    pad_h, pad_w = image_size[0] - kernel_size, image_size[1] - kernel_size
    if pad_h < 0 or pad_w < 0:
        raise ValueError("Kernel size should be smaller or equal to the image size.")
    rgb = F.pad(rgb, (pad_w // 2, pad_w // 2 + pad_w % 2, pad_h // 2, pad_h // 2 + pad_h % 2), "constant", 0)
    b, c, h, w = rgb.shape
    theta = torch.zeros(b, 2, 3, dtype=torch.float32, device=device)
    theta[:, 0, 0] = 1.0
    theta[:, 1, 1] = 1.0
    theta[:, :, 2] = coords
    grid = F.affine_grid(theta, size=(b, c, h, w), align_corners=True)
    moved = F.grid_sample(rgb, grid, align_corners=True)
    image = (colours.unsqueeze(-1).unsqueeze(-1) * moved).sum(dim=0)
    return torch.clamp(image, 0, 1).permute(1, 2, 0)


def time_dense_splat_2d(n, height, width, kernel_size=101, reps=2, threads=None, seed=0):
    """Forward and forward+backward wall time of one call on the host cores (torch CPU, `threads` intra-op
    threads).  Returns ms and ns per splat*pixel."""
    import torch
    if threads:
        torch.set_num_threads(int(threads))
    g = torch.Generator().manual_seed(seed)
    u = lambda *s: torch.rand(*s, generator=g)
    leaves = dict(sigma_x=(0.2 + u(n)), sigma_y=(0.2 + u(n)), rho=(1.6 * u(n) - 0.8), coords=(2 * u(n, 2) - 1),
                  colours=u(n, 3))
    leaves = {k: v.requires_grad_(True) for k, v in leaves.items()}
    target = u(height, width, 3)

    def call():
        return dense_splat_2d(kernel_size, leaves["sigma_x"], leaves["sigma_y"], leaves["rho"], leaves["coords"],
                              leaves["colours"], (height, width, 3))

    with torch.no_grad():
        call()                                   # warm-up (allocator, thread pool)
    tf = tfb = 1e30
    for _ in range(reps):
        t0 = time.perf_counter()
        with torch.no_grad():
            call()
        tf = min(tf, time.perf_counter() - t0)
        for v in leaves.values():
            v.grad = None
        t0 = time.perf_counter()
        loss = (call() - target).abs().mean()    # the toy's L1 term (2d_gaussian_splatting.py: combined_loss)
        loss.backward()
        tfb = min(tfb, time.perf_counter() - t0)
    sp = float(n) * height * width
    return {"gaussians": n, "image": [height, width], "kernel_size": kernel_size, "threads": torch.get_num_threads(),
            "fwd_ms": 1e3 * tf, "fwd_bwd_ms": 1e3 * tfb, "fwd_ns_per_splat_pixel": 1e9 * tf / sp,
            "fwd_bwd_ns_per_splat_pixel": 1e9 * tfb / sp}
