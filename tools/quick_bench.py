"""Scratch timing of fwd / bwd at a given size (not the contract bench; see bench.py)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from mvs_gaussian_splatting_b200 import synthetic as S, rasterize_forward, rasterize_backward, GaussianRasterizationSettings, last_forward_stats, _native

P = int(sys.argv[1]) if len(sys.argv) > 1 else 3_000_000
W = int(sys.argv[2]) if len(sys.argv) > 2 else 1920
H = int(sys.argv[3]) if len(sys.argv) > 3 else 1080
dev = torch.device("cuda:0")
cam = S.front_camera(W, H)
t0 = time.time(); g = S.make_gaussians(P, cam); print("gen", time.time() - t0, flush=True)
kw = S.settings_kwargs(cam, [0, 0, 0], 3)
for k in ("bg", "viewmatrix", "projmatrix", "campos"):
    kw[k] = torch.from_numpy(np.ascontiguousarray(kw[k])).to(dev)
rs = GaussianRasterizationSettings(**kw)
t = {k: torch.from_numpy(v).to(dev) for k, v in g.items()}
dL = torch.randn(3, H, W, device=dev)
def fwd():
    return rasterize_forward(t["means3D"], t["shs"], None, t["opacities"], t["scales"], t["rotations"], None, rs)
for it in range(3):
    c, r, ctx, inp = fwd(); rasterize_backward(ctx, inp, rs, dL)
torch.cuda.synchronize()
print(last_forward_stats(dev), flush=True)
e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
N = 10; tf = tb = 0
l0 = _native.launch_count()
for it in range(N):
    e[0].record(); c, r, ctx, inp = fwd(); e[1].record(); rasterize_backward(ctx, inp, rs, dL); e[2].record()
    torch.cuda.synchronize(); tf += e[0].elapsed_time(e[1]); tb += e[1].elapsed_time(e[2])
print(f"P={P} {W}x{H} fwd {tf/N:.3f} ms  bwd {tb/N:.3f} ms  launches/iter {(_native.launch_count()-l0)/N}")
print("n_contrib sum", int(ctx.array("n_contrib").long().sum()), "R", ctx.num_rendered)
