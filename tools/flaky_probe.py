"""Scratch: run-to-run spread of the elementwise gradient error on the opaque-wall stress case
(tests/test_gpu_variants.py::test_screen_filling_and_opaque_wall) — the blend backward's vector reductions land
in a different order every launch.  Prints, per parameter, the worst element error over N launches."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)
import gs_oracle  # noqa: E402
import helpers as Hh  # noqa: E402
from test_gpu_variants import _run_gpu  # noqa: E402

gs_oracle.build()
cuda = torch.device("cuda:0")
cam, g = Hh.make_case(4000, 256, 144, seed=16)
g["opacities"][:] = 0.98
g["scales"][:] *= 3.0
g["means3D"][0] = [0.0, 0.0, 2.0]
g["scales"][0] = [3.0, 3.0, 3.0]
g["opacities"][0] = 0.5
bg, D = [0.0, 0.0, 0.0], 1
fwd = Hh.run_oracle(gs_oracle, cam, g, bg, D)
dL = np.random.default_rng(26).normal(size=(3, cam.height, cam.width)).astype(np.float32)
dL[:, fwd["fwd"]["fragile"]] = 0.0
ref = Hh.run_oracle(gs_oracle, cam, g, bg, D, dL_dpix=dL)
ref64 = Hh.run_oracle(gs_oracle, cam, g, bg, D, dL_dpix=dL, dtype=np.float64)
N = int(sys.argv[1]) if len(sys.argv) > 1 else 12
out = {}
for it in range(N):
    _, _, grads = _run_gpu(cuda, cam, g, bg, D, dL)
    for k, a in grads.items():
        good, rep = Hh.grad_report(a, ref64["grads"][k], ref["grads"][k])
        out.setdefault(k, []).append((rep["worst"], rep["worst_fp32_oracle"], rep["frac"], rep["frac_fp32_oracle"], good))
for k, v in out.items():
    w = [x[0] for x in v]
    print(f"{k:10s} worst over {N} launches: min {min(w):.2e} max {max(w):.2e}  fp32-oracle worst {v[0][1]:.2e}  "
          f"frac max {max(x[2] for x in v):.1e} (oracle {v[0][3]:.1e})  failures {sum(not x[4] for x in v)}")
