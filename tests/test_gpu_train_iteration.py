"""SURVEY 8f N1 + N2 + N3 together: three training iterations (train.py:90-142) run twice from the same raw
leaf parameters — once with PyTorch ops around the rasterizer the way the reference has them (getters,
conv2d SSIM, torch.optim.Adam, indexed densification statistics), once through the fused entries
(forward_raw, l1_dssim_loss, densification_stats, FusedAdam).  Losses agree within 1e-4 at every iteration;
parameters after the last step agree to a small fraction of one learning-rate step (Adam's first steps are
sign(g) * lr, so an element whose gradient is rounding noise may legitimately land a full step apart: those
are bounded in number, not excluded)."""
import math
import os
import sys

import numpy as np
import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
pytestmark = pytest.mark.gpu

LRS = dict(xyz=0.00016, f_dc=0.0025, f_rest=0.0025 / 20.0, opacity=0.05, scaling=0.005, rotation=0.001)


def torch_loss(image, gt, lam):
    F = torch.nn.functional
    g = torch.tensor([math.exp(-(i - 5) ** 2 / (2 * 1.5 ** 2)) for i in range(11)])
    g = (g / g.sum()).unsqueeze(1)
    win = g.mm(g.t()).float().expand(3, 1, 11, 11).contiguous().to(image.device)
    conv = lambda a: F.conv2d(a, win, padding=5, groups=3)
    mu1, mu2 = conv(image), conv(gt)
    s1, s2, s12 = conv(image * image) - mu1 * mu1, conv(gt * gt) - mu2 * mu2, conv(image * gt) - mu1 * mu2
    m = ((2 * mu1 * mu2 + 1e-4) * (2 * s12 + 9e-4)) / ((mu1 * mu1 + mu2 * mu2 + 1e-4) * (s1 + s2 + 9e-4))
    return (1.0 - lam) * torch.abs(image - gt).mean() + lam * (1.0 - m.mean())


def test_three_iterations_fused_vs_torch_formulation(cuda):
    from mvs_gaussian_splatting_b200 import GaussianRasterizationSettings, GaussianRasterizer
    from mvs_gaussian_splatting_b200 import synthetic as S
    from mvs_gaussian_splatting_b200.loss import l1_dssim_loss
    from mvs_gaussian_splatting_b200.optim import FusedAdam, densification_stats
    W, H, P, M, D, lam, steps = 256, 192, 20_000, 16, 3, 0.2, 3
    cam = S.front_camera(W, H)
    g = S.make_gaussians(P, cam, sh_coeffs=M, seed=3)
    t = {k: torch.from_numpy(v).to(cuda) for k, v in g.items()}
    kw = S.settings_kwargs(cam, [0.1, 0.2, 0.3], D)
    for k in ("bg", "viewmatrix", "projmatrix", "campos"):
        kw[k] = torch.from_numpy(np.ascontiguousarray(kw[k])).to(cuda)
    R = GaussianRasterizer(GaussianRasterizationSettings(**kw))
    raw = dict(xyz=t["means3D"], f_dc=t["shs"][:, :1].contiguous(), f_rest=t["shs"][:, 1:].contiguous(),
               opacity=torch.logit(t["opacities"].clamp(1e-6, 1 - 1e-6)), scaling=torch.log(t["scales"]),
               rotation=t["rotations"] * 1.3)
    gen = torch.Generator(device="cpu").manual_seed(8)
    gt = torch.rand(3, H, W, generator=gen).to(cuda)

    def run(fused):
        ps = {k: v.detach().clone().requires_grad_(True) for k, v in raw.items()}
        opt = (FusedAdam if fused else torch.optim.Adam)([dict(params=[ps[k]], lr=LRS[k], name=k) for k in ps],
                                                         lr=0.0, eps=1e-15)
        stats = [torch.zeros(P, device=cuda), torch.zeros(P, 1, device=cuda), torch.zeros(P, 1, device=cuda)]
        losses = []
        for _ in range(steps):
            m2 = torch.zeros_like(ps["xyz"], requires_grad=True)
            if fused:
                image, radii = R.forward_raw(ps["xyz"], m2, ps["f_dc"], ps["f_rest"], ps["opacity"], ps["scaling"],
                                             ps["rotation"])
                loss = l1_dssim_loss(image, gt, lam)
            else:
                image, radii = R(means3D=ps["xyz"], means2D=m2, shs=torch.cat((ps["f_dc"], ps["f_rest"]), dim=1),
                                 opacities=torch.sigmoid(ps["opacity"]), scales=torch.exp(ps["scaling"]),
                                 rotations=torch.nn.functional.normalize(ps["rotation"]))
                loss = torch_loss(image, gt, lam)
            loss.backward()
            losses.append(loss.item())
            with torch.no_grad():
                if fused:
                    densification_stats(radii, m2.grad, *stats)
                else:
                    vis = radii > 0
                    stats[0][vis] = torch.max(stats[0][vis], radii[vis])
                    stats[1][vis] += torch.norm(m2.grad[vis, :2], dim=-1, keepdim=True)
                    stats[2][vis] += 1
                opt.step()
                opt.zero_grad(set_to_none=True)
        return losses, {k: v.detach().clone() for k, v in ps.items()}, stats

    la, pa, sa = run(True)
    lb, pb, sb = run(False)
    for a, b in zip(la, lb):
        assert abs(a - b) <= 1e-4 * abs(b), (la, lb)
    assert la[-1] != la[0]                                            # the parameters did move
    for k in pa:
        moved = (pb[k] - raw[k]).abs()
        assert moved.max().item() > 0.5 * LRS[k], k                  # Adam's first steps are ~lr in size
        d = (pa[k] - pb[k]).abs()
        frac_far = (d > 0.05 * LRS[k] * steps).float().mean().item()
        assert frac_far <= 0.01, (k, frac_far)
        assert d.max().item() <= 2.001 * LRS[k] * steps, k            # never more than sign flips every step
        assert d.median().item() <= 1e-3 * LRS[k] * steps, (k, d.median().item())
    assert torch.equal(sa[0], sb[0]) and torch.equal(sa[2], sb[2])    # max radii, visibility counts: exact
    rel = ((sa[1] - sb[1]).abs().sum() / sb[1].abs().sum()).item()
    assert rel <= 1e-4, rel
