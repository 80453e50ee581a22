"""Scratch: host-side cost of one fwd+bwd step through the public autograd API (no synchronisation inside the
loop, so the figure is what the Python side needs to enqueue a step) + a cProfile of 60 steps."""
import cProfile
import os
import pstats
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from mvs_gaussian_splatting_b200 import GaussianRasterizationSettings, GaussianRasterizer  # noqa: E402
from mvs_gaussian_splatting_b200 import synthetic as S  # noqa: E402
from mvs_gaussian_splatting_b200.loss import l1_loss  # noqa: E402

dev = torch.device("cuda:0")
P, W, H, D, M = 3_000_000, 1920, 1080, 3, 16
cam = S.front_camera(W, H)
g = S.make_gaussians(P, cam, sh_coeffs=M, seed=0)
params = {k: torch.from_numpy(v).to(dev).requires_grad_(True) for k, v in g.items()}
kw = S.settings_kwargs(cam, [0.0, 0.0, 0.0], D)
for k in ("bg", "viewmatrix", "projmatrix", "campos"):
    kw[k] = torch.from_numpy(np.ascontiguousarray(kw[k])).to(dev)
rs = GaussianRasterizationSettings(**kw)
gt = torch.rand(3, H, W, device=dev)


def step():
    m2 = torch.zeros_like(params["means3D"], requires_grad=True)
    color, radii = GaussianRasterizer(rs)(means3D=params["means3D"], means2D=m2, shs=params["shs"],
                                          opacities=params["opacities"], scales=params["scales"],
                                          rotations=params["rotations"])
    loss = l1_loss(color, gt)
    loss.backward()
    for p in params.values():
        p.grad = None


for _ in range(5):
    step()
torch.cuda.synchronize()
ts = []
for _ in range(40):
    torch.cuda.synchronize()       # empty queue: the call never blocks on a full launch queue
    t0 = time.perf_counter()
    step()
    ts.append(1e3 * (time.perf_counter() - t0))
torch.cuda.synchronize()
print(f"host ms per step (queue empty): median {np.median(ts):.3f} min {min(ts):.3f} max {max(ts):.3f}")
t0 = time.perf_counter()
for _ in range(40):
    step()
torch.cuda.synchronize()
print(f"back-to-back wall ms per step: {1e3 * (time.perf_counter() - t0) / 40:.3f}")
pr = cProfile.Profile()
pr.enable()
for _ in range(60):
    step()
torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(18)
