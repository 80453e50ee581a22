"""Helpers shared by the parity tests: seeded cases, oracle and CUDA runners, tolerances."""
import numpy as np

from mvs_gaussian_splatting_b200 import synthetic as S

REL_TOL = 1e-4  # north_star: image / gradient tensors within 1e-4 relative (fp32)


def make_case(P, W, H, seed=0, sh_coeffs=16, **kw):
    cam = S.front_camera(W, H)
    g = S.make_gaussians(P, cam, sh_coeffs=sh_coeffs, seed=seed, **kw)
    return cam, g


def run_oracle(O, cam, g, bg, sh_degree, dL_dpix=None, dtype=np.float32, threads=8, colors_precomp=None,
               cov3D_precomp=None, scale_modifier=1.0, margin=1e-4):
    use_sr = cov3D_precomp is None
    return O.rasterize(g["means3D"], g["opacities"], cam.world_view_transform, cam.full_proj_transform,
                       cam.camera_center, np.asarray(bg, np.float32), cam.width, cam.height, cam.tanfovx,
                       cam.tanfovy, sh_degree=sh_degree, shs=None if colors_precomp is not None else g["shs"],
                       colors_precomp=colors_precomp, scales=g["scales"] if use_sr else None,
                       rotations=g["rotations"] if use_sr else None, cov3D_precomp=cov3D_precomp,
                       scale_modifier=scale_modifier, dL_dpix=dL_dpix, dtype=dtype, threads=threads,
                       margin=margin)


def settings_for(cam, bg, sh_degree, device, scale_modifier=1.0, debug=False):
    import torch
    from mvs_gaussian_splatting_b200 import GaussianRasterizationSettings
    kw = S.settings_kwargs(cam, bg, sh_degree, scale_modifier, debug)
    for k in ("bg", "viewmatrix", "projmatrix", "campos"):
        kw[k] = torch.from_numpy(np.ascontiguousarray(kw[k])).to(device)
    return GaussianRasterizationSettings(**kw)


def to_dev(g, device, requires_grad=False):
    import torch
    out = {}
    for k, v in g.items():
        t = torch.from_numpy(v).to(device)
        if requires_grad:
            t.requires_grad_(True)
        out[k] = t
    return out


def close(a, b, scale=None, rel=REL_TOL):
    """|a-b| <= rel * max(|b|, scale) elementwise; returns (ok, worst ratio)."""
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    if scale is None:
        nz = np.abs(b[b != 0])
        scale = float(nz.mean()) if nz.size else 1.0
    err = np.abs(a - b) / np.maximum(np.abs(b), scale)
    worst = float(err.max()) if err.size else 0.0
    return worst <= rel, worst
