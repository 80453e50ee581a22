"""Diagnostic: pixels where a skip/stop decision of the CUDA path differs from the oracles' although the oracle
did not flag them fragile (tests/test_gpu_parity.py::check_backward's first assertion).  Prints what differs."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch
import gs_oracle as O
import helpers as Hh
from mvs_gaussian_splatting_b200 import rasterize_forward

O.build()
P, W, H, D, M, seed = 3_000_000, 1920, 1080, 3, 16, 0
camera = sys.argv[1] if len(sys.argv) > 1 else "rolled"
cuda = torch.device("cuda:0")
cam, g = Hh.make_case(P, W, H, seed=seed, sh_coeffs=M, camera=camera)
bg = [0.3, 0.2, 0.7]
thr = len(os.sched_getaffinity(0))
dL, ref, refre, fragile = Hh.gradient_reference(O, cam, g, bg, D, np.zeros((3, H, W)), threads=thr, zero_fragile=False,
                                                run_backward=False)
rs = Hh.settings_for(cam, bg, D, cuda)
t = Hh.to_dev(g, cuda)
color, radii, ctx, _ = rasterize_forward(t["means3D"], t["shs"], None, t["opacities"], t["scales"], t["rotations"], None, rs)
torch.cuda.synchronize()
n_c = ctx.array("n_contrib").cpu().numpy().view(np.uint32).reshape(H, W)
T_c = ctx.array("final_T").cpu().numpy().reshape(H, W)
T_r, n_r = ref["fwd"]["final_T"], ref["fwd"]["n_contrib"]
T_e, n_e = refre["fwd"]["final_T"]["v"], refre["fwd"]["n_contrib"]
flipped = Hh.flipped_pixels(n_c, T_c, ref, refre)
bad = flipped & ~fragile
print(camera, "fragile", fragile.mean(), "(f32", ref["fwd"]["fragile"].mean(), "re", refre["fwd"]["fragile"].mean(), ") flipped",
      flipped.mean(), "flipped&~fragile", int(bad.sum()))
tol = 1e-3 * T_r + 1e-7
for y, x in zip(*np.nonzero(bad)):
    print(f"  pixel ({x},{y}) tile ({x//16},{y//16}): n cuda {n_c[y,x]} f32 {n_r[y,x]} re {n_e[y,x]} | T cuda {T_c[y,x]:.9g} f32 {T_r[y,x]:.9g} "
          f"re {T_e[y,x]:.9g} tol {tol[y,x]:.3g} | range len {int(np.diff(ref['binned']['ranges'][(y//16)*((W+15)//16)+x//16])[0])}")
    if bad.sum() > 40:
        break
img = color.cpu().numpy()
d = np.abs(img - ref["color"]).max(0)
print("image: max |diff| on non-fragile", d[~fragile].max(), " on bad pixels", d[bad].max() if bad.any() else 0)
