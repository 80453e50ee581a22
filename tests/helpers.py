"""Helpers shared by the parity tests: seeded cases, oracle and CUDA runners, tolerances."""
import numpy as np

from mvs_gaussian_splatting_b200 import synthetic as S

REL_TOL = 1e-4  # north_star: image / gradient tensors within 1e-4 relative (fp32)


def make_case(P, W, H, seed=0, sh_coeffs=16, **kw):
    cam = S.front_camera(W, H)
    g = S.make_gaussians(P, cam, sh_coeffs=sh_coeffs, seed=seed, **kw)
    return cam, g


def run_oracle(O, cam, g, bg, sh_degree, dL_dpix=None, dtype=np.float32, threads=8, colors_precomp=None,
               cov3D_precomp=None, scale_modifier=1.0, margin=1e-4, lists_from=None):
    use_sr = cov3D_precomp is None
    return O.rasterize(g["means3D"], g["opacities"], cam.world_view_transform, cam.full_proj_transform,
                       cam.camera_center, np.asarray(bg, np.float32), cam.width, cam.height, cam.tanfovx,
                       cam.tanfovy, sh_degree=sh_degree, shs=None if colors_precomp is not None else g["shs"],
                       colors_precomp=colors_precomp, scales=g["scales"] if use_sr else None,
                       rotations=g["rotations"] if use_sr else None, cov3D_precomp=cov3D_precomp,
                       scale_modifier=scale_modifier, dL_dpix=dL_dpix, dtype=dtype, threads=threads,
                       margin=margin, lists_from=lists_from)


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


GRAD_K = 1.0   # multiples of the running rounding-error bound u*e accepted per element (see grad_check)


def gradient_reference(O, cam, g, bg, sh_degree, dL, threads=8, zero_fragile=True, run_backward=True, **kw):
    """The references a gradient test compares with, sharing ONE set of integer decisions:
      ref32 — the binary32 oracle (forward + backward): radii / lists / image are compared with it;
      refre — the running-error build (binary64 values `v` + per-element bound `e` on what a binary32
              evaluation of the same expressions accumulates, in units of u = 2^-24), evaluated on the
              binary32 run's visible set, radii, rects and sorted lists.
    dL is zeroed on pixels either run flags `fragile` (a skip / stop decision within its own rounding
    error of the threshold: the binary32 reference itself may fall either way there).  Returns
    (dL, ref32, refre, fragile)."""
    ref32 = run_oracle(O, cam, g, bg, sh_degree, threads=threads, **kw)
    refre = run_oracle(O, cam, g, bg, sh_degree, dtype=O.RE, threads=threads, lists_from=ref32, **kw)
    fragile = ref32["fwd"]["fragile"] | refre["fwd"]["fragile"]
    dL = np.array(dL, np.float32, copy=True)
    if zero_fragile:
        dL[:, fragile] = 0.0
    if run_backward:
        O.backward(ref32, dL)
        O.backward(refre, dL)
    return dL, ref32, refre, fragile


def flipped_pixels(n_c, T_c, ref32, refre):
    """Pixels where a skip / stop decision actually DIFFERS between the three evaluations in play — the CUDA path
    (its n_contrib [H,W] uint32 and final_T [H,W]), the binary32 oracle and the binary64 running-error oracle the
    gradients are compared with (it decides on its own binary64 values).  Observable: a flipped stop changes
    n_contrib, a flipped skip changes final_T by at least the factor 1 - 1/255.  Always a subset of `fragile`
    (asserted by the callers); everywhere else the three differentiate the same function."""
    T_r, n_r = ref32["fwd"]["final_T"], ref32["fwd"]["n_contrib"]
    T_e, n_e = refre["fwd"]["final_T"]["v"], refre["fwd"]["n_contrib"]
    tol = 1e-3 * T_r + 1e-7
    return (n_c != n_r) | (np.abs(T_c - T_r) > tol) | (n_e != n_r) | (np.abs(T_e - T_r) > tol)


def grad_check(a, re, rel=REL_TOL, k=GRAD_K):
    """Conditioning-aware ELEMENTWISE gradient bound, every element, no outlier budget.

    `re` is a gradient tensor of the running-error oracle: re["v"] the binary64 value, re["e"] a
    first-order bound (units of u = 2^-24) on the absolute error that ANY binary32 evaluation of the
    reference's expressions accumulates for that element — it grows with the magnitude of the terms that
    are summed and cancelled on the way (hundreds of signed per-pixel terms, det = ac - b^2 of needle
    splats, 1 - alpha, the T chain), which is what makes some elements ill-conditioned in binary32.
    Rule:   |a - v| <= max(rel * |v|, k * u * e)   for EVERY element,
    i.e. 1e-4 relative (north_star), or within k times the rounding-error bound of the reference algorithm
    itself where that is larger.  Measured: the binary32 oracle stays below 0.35 u e on every
    element; typical elements sit at 0.002 u e.  Also asserted: tensor-level sum|a-v| / sum|v| <= rel and the
    99.9th percentile of |a-v| / (u e) <= k / 2.
    Returns (ok, dict of measured figures)."""
    a = np.asarray(a, np.float64)
    v, e = np.asarray(re["v"], np.float64), np.asarray(re["e"], np.float64)
    assert a.shape == v.shape, (a.shape, v.shape)
    err = np.abs(a - v)
    ue = (2.0 ** -24) * e
    bound = np.maximum(rel * np.abs(v), k * ue)
    with np.errstate(divide="ignore", invalid="ignore"):
        ratio = np.where(ue > 0, err / ue, np.where(err > 0, np.inf, 0.0))
    tot = float(np.abs(v).sum()) or 1.0
    bad = err > bound
    rep = dict(rel_l1=float(err.sum()) / tot, worst_over_ue=float(ratio.max()) if ratio.size else 0.0,
               p999_over_ue=float(np.quantile(ratio, 0.999)) if ratio.size else 0.0,
               median_over_ue=float(np.median(ratio)) if ratio.size else 0.0,
               n_outside=int(bad.sum()), n=int(a.size),
               frac_rel_gt_tol=float((err > rel * np.abs(v)).mean()) if a.size else 0.0)
    if bad.any():
        i = np.unravel_index(int(np.argmax(np.where(bad, err / np.maximum(bound, 1e-300), 0.0))), a.shape)
        rep["worst_element"] = dict(index=tuple(int(x) for x in i), got=float(a[i]), want=float(v[i]), ue=float(ue[i]))
    ok = (not bad.any()) and rep["rel_l1"] <= rel and rep["p999_over_ue"] <= 0.5 * k
    return ok, rep


def dump_grad_report(title, report):
    """Appends the measured per-tensor figures of grad_check to $GRAD_REPORT (default gpurun_out/grad_report.txt)."""
    import os
    path = os.environ.get("GRAD_REPORT", os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                      "gpurun_out", "grad_report.txt"))
    try:
        os.makedirs(os.path.dirname(path), exist_ok=True)
        with open(path, "a") as f:
            for k, r in report.items():
                f.write(f"{title} {k}: worst |a-v|/(u e) {r['worst_over_ue']:.3f}  p99.9 {r['p999_over_ue']:.3f}  median "
                        f"{r['median_over_ue']:.4f}  rel_l1 {r['rel_l1']:.2e}  outside {r['n_outside']}/{r['n']}  "
                        f"(elements beyond plain 1e-4 relative: {r['frac_rel_gt_tol']:.2e})\n")
    except OSError:
        pass


def depth_order_emission(pre, binned):
    """The oracle's unsorted (key, value) lists re-ordered block-wise into the emission order the
    CUDA path documents: Gaussians by (depth bits, index), each Gaussian's tiles row-major."""
    tt = pre["tiles_touched"].astype(np.int64)
    ends = np.cumsum(tt)
    starts = ends - tt
    vis = pre["radii"] > 0
    ids = np.flatnonzero(vis)
    ids = ids[np.argsort(pre["depths"].view(np.uint32)[vis], kind="stable")]
    n = tt[ids]
    total = int(n.sum())
    if total == 0:
        idx = np.zeros(0, np.int64)
    else:
        block_first = np.cumsum(n) - n                      # first output slot of each Gaussian's block
        idx = np.repeat(starts[ids] - block_first, n) + np.arange(total, dtype=np.int64)
    return binned["keys_unsorted"][idx], binned["values_unsorted"][idx]
