"""Scratch: where the host time of one e2e step goes relative to the GPU (perf_counter stamps, medians over 60 steps):
forward enqueue incl. the capacity wait | loss forward | backward enqueue | the rest (zero_grad, copies)."""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mvs_gaussian_splatting_b200 import GaussianRasterizationSettings, GaussianRasterizer
from mvs_gaussian_splatting_b200 import synthetic as S
from mvs_gaussian_splatting_b200.loss import l1_loss
dev = torch.device("cuda:0")
P, W, H, D, M = 3_000_000, 1920, 1080, 3, 16
cam = S.front_camera(W, H)
g = S.make_gaussians(P, cam, sh_coeffs=M, seed=0)
params = {k: torch.from_numpy(v).to(dev).requires_grad_(True) for k, v in g.items()}
kw = S.settings_kwargs(cam, [0.0, 0.0, 0.0], D)
for k in ("bg", "viewmatrix", "projmatrix", "campos"):
    kw[k] = torch.from_numpy(np.ascontiguousarray(kw[k])).to(dev)
rs = GaussianRasterizationSettings(**kw)
gt_host = (torch.rand(3, H, W) * 255).to(torch.uint8).pin_memory()
gt_u8 = [torch.empty_like(gt_host, device=dev) for _ in range(2)]
gt = torch.empty(3, H, W, device=dev)
copy_stream = torch.cuda.Stream(device=dev)
main = torch.cuda.current_stream(dev)
ready = [torch.cuda.Event(), torch.cuda.Event()]
free = [torch.cuda.Event(), torch.cuda.Event()]
loss_host = [torch.zeros(1).pin_memory() for _ in range(2)]
loss_ready = [torch.cuda.Event(), torch.cuda.Event()]
for e in free: e.record(main)
MODE = os.environ.get("MODE", "full")
def prefetch(i):
    with torch.cuda.stream(copy_stream):
        copy_stream.wait_event(free[i]); gt_u8[i].copy_(gt_host, non_blocking=True); ready[i].record(copy_stream)
stamps = []
k = 0
def step():
    global k
    i = k & 1
    t0 = time.perf_counter()
    if MODE != "nocopy":
        prefetch((k + 1) & 1); main.wait_event(ready[i]); torch.mul(gt_u8[i], 1 / 255.0, out=gt)
    m2 = torch.zeros_like(params["means3D"], requires_grad=True)
    t1 = time.perf_counter()
    color, radii = GaussianRasterizer(rs)(means3D=params["means3D"], means2D=m2, shs=params["shs"], opacities=params["opacities"],
                                          scales=params["scales"], rotations=params["rotations"])
    t2 = time.perf_counter()
    loss = l1_loss(color, gt)
    t3 = time.perf_counter()
    loss.backward()
    t4 = time.perf_counter()
    if MODE != "nocopy":
        loss_host[i].copy_(loss.detach().reshape(1), non_blocking=True); loss_ready[i].record(main); free[i].record(main)
        if k > 0: loss_ready[i ^ 1].synchronize()
    for p in params.values(): p.grad = None
    t5 = time.perf_counter()
    stamps.append((t1 - t0, t2 - t1, t3 - t2, t4 - t3, t5 - t4))
    k += 1
prefetch(0)
for _ in range(10): step()
torch.cuda.synchronize(); stamps.clear()
t0 = time.perf_counter()
for _ in range(60): step()
torch.cuda.synchronize()
wall = 1e3 * (time.perf_counter() - t0) / 60
a = 1e3 * np.median(np.array(stamps), axis=0)
print(f"MODE={MODE} wall {wall:.3f} ms/step | host medians: pre {a[0]:.3f} forward(incl. capacity wait) {a[1]:.3f} loss {a[2]:.3f} backward {a[3]:.3f} post {a[4]:.3f} | sum {a.sum():.3f}")
