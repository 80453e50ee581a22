"""numpy/ctypes front-end of the CPU oracle (oracle/gs_oracle.c).

TEST INFRASTRUCTURE ONLY — *** parity unpinned *** (see the header of gs_oracle.c).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module; the product package (mvs_gaussian_splatting_b200) never does.

The functions mirror the stages of the rasterizer the reference calls at
/root/reference/gaussian_renderer/__init__.py:257-265:
  preprocess -> bin (emit keys, stable sort, tile ranges) -> blend forward
  blend backward -> preprocess backward
"""
from __future__ import annotations

import ctypes
import math
import os
import subprocess
from concurrent.futures import ThreadPoolExecutor

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "libgs_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    """Compile the oracle with its Makefile (gcc only; no reference sources involved)."""
    if force or not os.path.exists(_LIB_PATH) or (
        os.path.getmtime(_LIB_PATH) < os.path.getmtime(os.path.join(_HERE, "gs_oracle.c"))
    ):
        subprocess.run(["make", "-C", _HERE], check=True, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_LIB_PATH)
        _lib.orc_count_instances_f32.restype = ctypes.c_int64
        _lib.orc_count_instances_f64.restype = ctypes.c_int64
        _lib.orc_count_instances_re.restype = ctypes.c_int64
    return _lib


# Third build of the oracle (gs_oracle.c compiled as C++ with -DORC_RE): every `real` is a pair
# (v = binary64 value, e = first-order bound, in units of u = 2^-24, on the absolute error a binary32
# evaluation of the same expression accumulates).  Pass dtype=RE to any stage function / rasterize().
RE = np.dtype([("v", np.float64), ("e", np.float64)])
U32 = 2.0 ** -24   # binary32 unit roundoff


class _ReScalar(ctypes.Structure):
    _fields_ = [("v", ctypes.c_double), ("e", ctypes.c_double)]


def _re_scalar(x):
    x = float(x)
    return _ReScalar(x, 0.0 if float(np.float32(x)) == x else abs(x))   # the rule of real(double) in gs_oracle.c


def _is_re(dt):
    return np.dtype(dt) == RE


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def _c(a, dt):
    if a is None:
        return None
    if _is_re(dt):
        a = np.asarray(a)
        if a.dtype == RE:
            return np.ascontiguousarray(a)
        out = np.zeros(a.shape, RE)      # inputs are exact binary32 values: e = 0
        out["v"] = a
        return out
    return np.ascontiguousarray(a, dtype=dt)


def _fn(name, dt):
    suffix = "re" if _is_re(dt) else ("f64" if dt == np.float64 else "f32")
    return getattr(lib(), f"{name}_{suffix}")


def _real(dt):
    if _is_re(dt):
        return _re_scalar
    return ctypes.c_double if dt == np.float64 else ctypes.c_float


def grid(W, H):
    return (W + 15) // 16, (H + 15) // 16


def _ranges(n, threads):
    if n <= 0:
        return []
    threads = max(1, min(threads, n))
    step = (n + threads - 1) // threads
    return [(b, min(n, b + step)) for b in range(0, n, step)]


def preprocess(means3D, scales, rotations, opacities, shs, viewmatrix, projmatrix, campos, W, H,
               tanfovx, tanfovy, sh_degree, scale_modifier=1.0, cov3D_precomp=None,
               colors_precomp=None, dtype=np.float32, threads=1):
    """K1. viewmatrix/projmatrix are the reference's *transposed* 4x4 tensors
    (scene/cameras.py:54-56); flattened row-major they are the column-major arrays the
    kernels index."""
    dt = dtype
    R = _real(dt)
    means3D = _c(means3D, dt)
    P = means3D.shape[0]
    shs = _c(shs, dt)
    M = 0 if shs is None else shs.shape[1]
    scales, rotations = _c(scales, dt), _c(rotations, dt)
    opacities = _c(opacities, dt).reshape(-1)
    cov3D_precomp, colors_precomp = _c(cov3D_precomp, dt), _c(colors_precomp, dt)
    view, proj = _c(viewmatrix, dt).reshape(16), _c(projmatrix, dt).reshape(16)
    campos = _c(campos, dt).reshape(3)
    out = dict(
        radii=np.zeros(P, np.int32), means2D=np.zeros((P, 2), dt), depths=np.zeros(P, dt),
        cov3D=np.zeros((P, 6), dt), rgb=np.zeros((P, 3), dt), conic_opacity=np.zeros((P, 4), dt),
        tiles_touched=np.zeros(P, np.uint32), rects=np.zeros((P, 4), np.int32),
        clamped=np.zeros((P, 3), np.uint8),
    )
    f = _fn("orc_preprocess", dt)

    def run(rg):
        f(P, int(sh_degree), M, _p(means3D), _p(scales), R(scale_modifier), _p(rotations),
          _p(opacities), _p(shs), _p(cov3D_precomp), _p(colors_precomp), _p(view), _p(proj),
          _p(campos), W, H, R(tanfovx), R(tanfovy), _p(out["radii"]), _p(out["means2D"]),
          _p(out["depths"]), _p(out["cov3D"]), _p(out["rgb"]), _p(out["conic_opacity"]),
          _p(out["tiles_touched"]), _p(out["rects"]), _p(out["clamped"]), rg[0], rg[1])

    _run_ranges(run, _ranges(P, threads))
    return out


def sh_eval(xyz, campos, shs, sh_degree, dtype=np.float32):
    """SH colour of each point seen from campos: (rgb [P,3] after +0.5 and clamp, clamped [P,3])."""
    xyz, shs, campos = _c(xyz, dtype), _c(shs, dtype), _c(campos, dtype).reshape(3)
    P, M = xyz.shape[0], shs.shape[1]
    rgb, clamped = np.zeros((P, 3), dtype), np.zeros((P, 3), np.uint8)
    _fn("orc_sh_eval", dtype)(P, int(sh_degree), M, _p(xyz), _p(campos), _p(shs), _p(rgb), _p(clamped))
    return rgb, clamped.astype(bool)


def cov3d(scales, rotations, scale_modifier=1.0, dtype=np.float32):
    """6-float upper-triangular 3D covariance from scale and (un-normalised) quaternion."""
    scales, rotations = _c(scales, dtype), _c(rotations, dtype)
    P = scales.shape[0]
    out = np.zeros((P, 6), dtype)
    _fn("orc_cov3d", dtype)(P, _p(scales), _real(dtype)(scale_modifier), _p(rotations), _p(out))
    return out


def _run_ranges(run, ranges):
    if len(ranges) <= 1:
        for rg in ranges:
            run(rg)
    else:
        with ThreadPoolExecutor(len(ranges)) as ex:
            list(ex.map(run, ranges))


def mark_visible(means3D, viewmatrix, dtype=np.float32):
    means3D = _c(means3D, dtype)
    P = means3D.shape[0]
    view = _c(viewmatrix, dtype).reshape(16)
    present = np.zeros(P, np.uint8)
    _fn("orc_mark_visible", dtype)(P, _p(means3D), _p(view), _p(present))
    return present.astype(bool)


def bin_instances(pre, W, H, dtype=np.float32):
    """K2..K6: returns unsorted + sorted key/value lists and per-tile ranges."""
    P = pre["radii"].shape[0]
    gx, gy = grid(W, H)
    R = int(_fn("orc_count_instances", dtype)(P, _p(pre["tiles_touched"])))
    n = max(R, 1)
    ku, vu = np.zeros(n, np.uint64), np.zeros(n, np.uint32)
    ks, vs = np.zeros(n, np.uint64), np.zeros(n, np.uint32)
    ranges = np.zeros((gx * gy, 2), np.uint32)
    _fn("orc_bin", dtype)(P, _p(pre["radii"]), _p(pre["rects"]), _p(pre["depths"]), W, H,
                          ctypes.c_int64(R), _p(ku), _p(vu), _p(ks), _p(vs), _p(ranges))
    return dict(num_rendered=R, keys_unsorted=ku[:R], values_unsorted=vu[:R], keys=ks[:R],
                point_list=vs[:R], ranges=ranges)


def blend_forward(pre, binned, bg, W, H, dtype=np.float32, margin=1e-4, threads=1, tile_ranges=None):
    """K7.  tile_ranges: optional list of [t0,t1) tile-index ranges to process instead of all tiles
    (bench.py's bounded CPU sample); other pixels stay zero."""
    dt = dtype
    gx, gy = grid(W, H)
    out = np.zeros((3, H, W), dt)
    final_T = np.zeros((H, W), dt)
    n_contrib = np.zeros((H, W), np.uint32)
    fragile = np.zeros((H, W), np.uint8)
    bg = _c(bg, dt).reshape(3)
    pl = np.ascontiguousarray(binned["point_list"]) if binned["num_rendered"] else np.zeros(1, np.uint32)
    f = _fn("orc_blend_fwd", dt)

    def run(rg):
        f(W, H, _p(binned["ranges"]), _p(pl), _p(pre["means2D"]), _p(pre["rgb"]),
          _p(pre["conic_opacity"]), _p(bg), _p(out), _p(final_T), _p(n_contrib), _p(fragile),
          _real(dt)(margin), rg[0], rg[1])

    if tile_ranges is None:
        _run_ranges(run, _ranges(gx * gy, threads * 8 if threads > 1 else 1))
    else:
        with ThreadPoolExecutor(max(1, threads)) as ex:
            list(ex.map(run, tile_ranges))
    return dict(color=out, final_T=final_T, n_contrib=n_contrib, fragile=fragile.astype(bool))


def blend_backward(pre, binned, fwd, bg, dL_dpix, W, H, dtype=np.float32, threads=1, tile_ranges=None):
    """K8 (tile_ranges as in blend_forward). Returns dL/dmean2D [P,2] (NDC-scaled), dL/dconic [P,3] (true partials of A,B,C),
    dL/dopacity [P], dL/dcolor [P,3]."""
    dt = dtype
    P = pre["radii"].shape[0]
    gx, gy = grid(W, H)
    bg = _c(bg, dt).reshape(3)
    dL_dpix = _c(dL_dpix, dt)
    pl = np.ascontiguousarray(binned["point_list"]) if binned["num_rendered"] else np.zeros(1, np.uint32)
    f = _fn("orc_blend_bwd", dt)
    if tile_ranges is None:
        tile_ranges = _ranges(gx * gy, threads * 8 if threads > 1 else 1)
    # one accumulator per thread, ranges dealt round-robin (balances dense and sparse image regions)
    nt = max(1, min(threads, len(tile_ranges)))
    groups = [list(tile_ranges[i::nt]) for i in range(nt)]
    bufs = []

    def run(group):
        g = dict(dL_dmean2D=np.zeros((P, 2), dt), dL_dconic=np.zeros((P, 3), dt),
                 dL_dopacity=np.zeros(P, dt), dL_dcolor=np.zeros((P, 3), dt))
        for rg in group:
            f(W, H, _p(binned["ranges"]), _p(pl), _p(pre["means2D"]), _p(pre["rgb"]),
              _p(pre["conic_opacity"]), _p(bg), _p(fwd["final_T"]), _p(fwd["n_contrib"]), _p(dL_dpix),
              _p(g["dL_dmean2D"]), _p(g["dL_dconic"]), _p(g["dL_dopacity"]), _p(g["dL_dcolor"]),
              rg[0], rg[1])
        bufs.append(g)

    _run_ranges(run, groups)
    acc = bufs[0]
    for g in bufs[1:]:
        for k in acc:
            if _is_re(dt):   # one more addition: bounds add, plus its own rounding
                acc[k]["v"] += g[k]["v"]
                acc[k]["e"] += g[k]["e"] + np.abs(acc[k]["v"])
            else:
                acc[k] += g[k]
    return acc


def preprocess_backward(means3D, scales, rotations, shs, viewmatrix, projmatrix, campos, W, H,
                        tanfovx, tanfovy, sh_degree, pre, gblend, scale_modifier=1.0,
                        cov3D_precomp=None, colors_precomp=None, dtype=np.float32, threads=1):
    """K9 + K10. All outputs fully written (zeros for culled Gaussians)."""
    dt = dtype
    R = _real(dt)
    means3D = _c(means3D, dt)
    P = means3D.shape[0]
    shs = _c(shs, dt)
    M = 0 if shs is None else shs.shape[1]
    scales, rotations = _c(scales, dt), _c(rotations, dt)
    view, proj = _c(viewmatrix, dt).reshape(16), _c(projmatrix, dt).reshape(16)
    campos = _c(campos, dt).reshape(3)
    cov_in = pre["cov3D"] if cov3D_precomp is None else _c(cov3D_precomp, dt)
    out = dict(dL_dmeans3D=np.zeros((P, 3), dt), dL_dcov3D=np.zeros((P, 6), dt),
               dL_dsh=None if shs is None else np.zeros((P, M, 3), dt),
               dL_dscales=None if scales is None else np.zeros((P, 3), dt),
               dL_drots=None if rotations is None else np.zeros((P, 4), dt))
    f = _fn("orc_preprocess_bwd", dt)

    def run(rg):
        f(P, int(sh_degree), M, _p(means3D), _p(pre["radii"]), _p(shs), _p(pre["clamped"]),
          _p(scales), _p(rotations), R(scale_modifier), _p(cov_in),
          int(cov3D_precomp is not None), int(colors_precomp is not None), _p(view), _p(proj),
          _p(campos), W, H, R(tanfovx), R(tanfovy), _p(gblend["dL_dmean2D"]),
          _p(gblend["dL_dconic"]), _p(gblend["dL_dcolor"]), _p(out["dL_dmeans3D"]),
          _p(out["dL_dcov3D"]), _p(out["dL_dsh"]), _p(out["dL_dscales"]), _p(out["dL_drots"]),
          rg[0], rg[1])

    _run_ranges(run, _ranges(P, threads))
    return out


def rasterize(means3D, opacities, viewmatrix, projmatrix, campos, bg, W, H, tanfovx, tanfovy,
              sh_degree=0, shs=None, colors_precomp=None, scales=None, rotations=None,
              cov3D_precomp=None, scale_modifier=1.0, dL_dpix=None, dtype=np.float32,
              margin=1e-4, threads=1, lists_from=None):
    """Whole path. With dL_dpix, also runs the reverse pass and returns the gradients the
    reference's autograd function hands back (same order/meaning as SURVEY.md §8a a5).

    lists_from: result of a binary32 run of the same scene.  The binary64 / running-error evaluations
    then take every INTEGER decision of the geometry stage from it (visible set, radii, tile rects, the
    depth-sorted per-tile lists), so that they differentiate the function the binary32 path evaluates
    instead of a neighbouring one (a radius that rounds up in one precision and not in the other changes
    which tiles a splat reaches)."""
    pre = preprocess(means3D, scales, rotations, opacities, shs, viewmatrix, projmatrix, campos,
                     W, H, tanfovx, tanfovy, sh_degree, scale_modifier, cov3D_precomp,
                     colors_precomp, dtype, threads)
    if lists_from is not None:
        p32 = lists_from["pre"]
        flipped = np.flatnonzero((p32["radii"] > 0) != (pre["radii"] > 0))
        if flipped.size:
            raise RuntimeError(f"visible sets differ between precisions for Gaussians {flipped[:8]} — "
                               "a cull decision sits on its threshold; pick another seed")
        for k in ("radii", "rects", "tiles_touched"):
            pre[k] = p32[k]
        binned = lists_from["binned"]
    else:
        binned = bin_instances(pre, W, H, dtype)
    fwd = blend_forward(pre, binned, bg, W, H, dtype, margin, threads)
    res = dict(pre=pre, binned=binned, fwd=fwd, color=fwd["color"], radii=pre["radii"])
    res["call"] = dict(means3D=means3D, scales=scales, rotations=rotations, shs=shs, viewmatrix=viewmatrix,
                       projmatrix=projmatrix, campos=campos, bg=bg, W=W, H=H, tanfovx=tanfovx, tanfovy=tanfovy,
                       sh_degree=sh_degree, scale_modifier=scale_modifier, cov3D_precomp=cov3D_precomp,
                       colors_precomp=colors_precomp, dtype=dtype, threads=threads)
    if dL_dpix is not None:
        backward(res, dL_dpix)
    return res


def backward(res, dL_dpix):
    """Reverse pass of a finished `rasterize` result (its forward state is reused, not recomputed);
    adds res["grads"] (+ the stage outputs "gblend", "gpre") and returns res."""
    c = res["call"]
    pre, binned, fwd, dtype, threads = res["pre"], res["binned"], res["fwd"], c["dtype"], c["threads"]
    W, H = c["W"], c["H"]
    gb = blend_backward(pre, binned, fwd, c["bg"], dL_dpix, W, H, dtype, threads)
    gp = preprocess_backward(c["means3D"], c["scales"], c["rotations"], c["shs"], c["viewmatrix"], c["projmatrix"],
                             c["campos"], W, H, c["tanfovx"], c["tanfovy"], c["sh_degree"], pre, gb,
                             c["scale_modifier"], c["cov3D_precomp"], c["colors_precomp"], dtype, threads)
    P = pre["radii"].shape[0]
    g2 = np.zeros((P, 3), dtype)
    g2[:, :2] = gb["dL_dmean2D"]
    res["grads"] = dict(
        means3D=gp["dL_dmeans3D"], means2D=g2, shs=gp["dL_dsh"],
        colors_precomp=gb["dL_dcolor"] if c["colors_precomp"] is not None else None,
        opacities=gb["dL_dopacity"].reshape(P, 1), scales=gp["dL_dscales"],
        rotations=gp["dL_drots"],
        cov3D_precomp=gp["dL_dcov3D"] if c["cov3D_precomp"] is not None else None,
    )
    res["gblend"] = gb
    res["gpre"] = gp
    return res


# ---------------------------------------------------------------------------------------------
# Raw-parameter mode: the getters of scene/gaussian_model.py:152-183 with the
# activations installed at :28-45 (torch.exp, torch.nn.functional.normalize eps=1e-12,
# torch.sigmoid, torch.cat along dim 1), and their chain rules.  numpy restatement; pinned against
# tests/golden/activations.npz, which is produced by the reference's own GaussianModel getters.
# ---------------------------------------------------------------------------------------------
def activate(features_dc, features_rest, opacity_raw, scaling_raw, rotation_raw, dtype=np.float32):
    dt = dtype
    sc = np.exp(np.asarray(scaling_raw, dt))
    q = np.asarray(rotation_raw, dt)
    n = np.maximum(np.sqrt((q * q).sum(1, keepdims=True, dtype=dt)), dt(1e-12))
    rot = q / n
    op = dt(1) / (dt(1) + np.exp(-np.asarray(opacity_raw, dt)))
    dc = np.asarray(features_dc, dt)
    shs = dc if features_rest is None else np.concatenate([dc, np.asarray(features_rest, dt)], axis=1)
    return dict(shs=shs, opacities=op.reshape(-1, 1), scales=sc, rotations=rot)


def activate_backward(opacity_raw, scaling_raw, rotation_raw, g_shs, g_opacities, g_scales, g_rotations,
                      dtype=np.float64):
    """Gradients w.r.t. the raw tensors given gradients w.r.t. the activated ones."""
    dt = dtype
    act = activate(np.zeros((len(scaling_raw), 1, 3)), None, opacity_raw, scaling_raw, rotation_raw, dtype=dt)
    op, sc, qh = act["opacities"], act["scales"], act["rotations"]
    q = np.asarray(rotation_raw, dt)
    n = np.maximum(np.sqrt((q * q).sum(1, keepdims=True)), dt(1e-12))
    gq = np.asarray(g_rotations, dt)
    g_shs = np.asarray(g_shs, dt)
    return dict(features_dc=g_shs[:, :1], features_rest=g_shs[:, 1:],
                opacity_raw=np.asarray(g_opacities, dt).reshape(-1, 1) * op * (1 - op),
                scaling_raw=np.asarray(g_scales, dt) * sc,
                rotation_raw=(gq - qh * (qh * gq).sum(1, keepdims=True)) / n)


# ---------------------------------------------------------------------------------------------
# The step after backward: torch.optim.Adam's single-tensor update as the reference
# configures it (scene/gaussian_model.py:262: lr per group, betas (0.9, 0.999), eps 1e-15, no amsgrad, no
# weight decay) and the densification statistics of train.py:130-131 / scene/gaussian_model.py:775-777.
# numpy restatement in binary32, pinned against tests/golden/adam.npz and densification.npz.
# ---------------------------------------------------------------------------------------------
def adam_step(p, g, m, v, lr, step, beta1=0.9, beta2=0.999, eps=1e-15):
    """One step in place on float32 arrays (same operation order as the CUDA kernel)."""
    f = np.float32
    w1, w2 = f(1.0 - beta1), f(1.0 - beta2)
    bc1, bc2 = 1.0 - beta1 ** step, 1.0 - beta2 ** step
    step_size, bc2_sqrt = f(lr / bc1), f(np.sqrt(bc2))
    m += w1 * (g - m)
    v *= f(beta2)
    v += (w2 * g) * g
    denom = np.sqrt(v) / bc2_sqrt + f(eps)
    p -= step_size * (m / denom)


def densification_stats(radii, grad, max_radii2D, xyz_gradient_accum, denom):
    vis = radii > 0
    max_radii2D[vis] = np.maximum(max_radii2D[vis], radii[vis].astype(np.float32))
    xyz_gradient_accum[vis, 0] += np.sqrt(grad[vis, 0] ** 2 + grad[vis, 1] ** 2, dtype=np.float32)
    denom[vis] += 1


# ---- N3: photometric loss (utils/loss_utils.py:17-63, train.py:99-101) -------------------------------------------
def ssim_window(window_size=11, sigma=1.5):
    """utils/loss_utils.py:23-31: binary32 1-D Gaussian normalised by its binary32 sum, 2-D window = outer product."""
    g = np.array([math.exp(-(x - window_size // 2) ** 2 / float(2 * sigma ** 2)) for x in range(window_size)], np.float32)
    g = (g / g.sum(dtype=np.float32)).astype(np.float32)
    return g, np.outer(g, g).astype(np.float32)


def _conv_same(a, w2d):
    """Zero-padded 'same' correlation of every channel of a [C,H,W] with the 2-D window (F.conv2d, groups=C);
    accumulated in binary64."""
    C, H, W = a.shape
    k = w2d.shape[0] // 2
    pad = np.zeros((C, H + 2 * k, W + 2 * k), np.float64)
    pad[:, k:k + H, k:k + W] = a
    out = np.zeros((C, H, W), np.float64)
    for dy in range(2 * k + 1):
        for dx in range(2 * k + 1):
            out += float(w2d[dy, dx]) * pad[:, dy:dy + H, dx:dx + W]
    return out


def photometric_loss(image, gt, lambda_dssim=0.2):
    """Returns dict(l1, ssim, loss, grad, grad_ssim): l1_loss (loss_utils.py:17-18), ssim with size_average
    (:39-63), loss = (1-lambda) l1 + lambda (1-ssim) (train.py:100-101), and the analytic gradients of loss and
    of ssim w.r.t. `image` (what autograd computes through the five grouped convolutions)."""
    x, y = np.asarray(image, np.float64), np.asarray(gt, np.float64)
    _, w2d = ssim_window()
    N = x.size
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    mu1, mu2 = _conv_same(x, w2d), _conv_same(y, w2d)
    e11, e22, e12 = _conv_same(x * x, w2d), _conv_same(y * y, w2d), _conv_same(x * y, w2d)
    s1, s2, s12 = e11 - mu1 * mu1, e22 - mu2 * mu2, e12 - mu1 * mu2
    A1, A2 = 2 * mu1 * mu2 + C1, 2 * s12 + C2
    B1, B2 = mu1 * mu1 + mu2 * mu2 + C1, s1 + s2 + C2
    S = A1 * A2 / (B1 * B2)
    l1, ssim = np.abs(x - y).mean() if N else 0.0, S.mean() if N else 0.0
    # partial derivatives of the map w.r.t. the three convolution outputs that depend on `image`
    d_mu1 = 2 * mu2 * (A2 - A1) / (B1 * B2) - 2 * mu1 * S * (1 / B1 - 1 / B2)
    d_e11 = -S / B2
    d_e12 = 2 * A1 / (B1 * B2)
    # the window is symmetric: the adjoint of the correlation is the same correlation
    g_ssim = (_conv_same(d_mu1, w2d) + 2 * x * _conv_same(d_e11, w2d) + y * _conv_same(d_e12, w2d)) / max(N, 1)
    g_l1 = np.sign(x - y) / max(N, 1)
    lam = float(lambda_dssim)
    return dict(l1=l1, ssim=ssim, loss=(1.0 - lam) * l1 + lam * (1.0 - ssim),
                grad=(1.0 - lam) * g_l1 - lam * g_ssim, grad_ssim=g_ssim)


# ---- N4: Gaussian .ply layout (scene/gaussian_model.py:279-358) ----------------------------------------------------
def ply_attribute_names(n_dc=3, n_rest=45):
    """construct_list_of_attributes (:279-291)."""
    return (["x", "y", "z", "nx", "ny", "nz"] + [f"f_dc_{i}" for i in range(n_dc)]
            + [f"f_rest_{i}" for i in range(n_rest)] + ["opacity"] + [f"scale_{i}" for i in range(3)]
            + [f"rot_{i}" for i in range(4)])


def ply_rows_from_params(xyz, features_dc, features_rest, opacity, scaling, rotation):
    """save_ply (:293-309): the [P, n_props] binary32 matrix; features are stored transposed
    ([P,K,3] -> [P,3,K] flattened), normals as zeros."""
    P = xyz.shape[0]
    f_dc = np.transpose(features_dc, (0, 2, 1)).reshape(P, 3 * features_dc.shape[1])
    f_rest = np.transpose(features_rest, (0, 2, 1)).reshape(P, 3 * features_rest.shape[1])
    rows = np.concatenate((xyz, np.zeros_like(xyz), f_dc, f_rest, opacity, scaling, rotation), axis=1)
    return ply_attribute_names(f_dc.shape[1], f_rest.shape[1]), rows.astype(np.float32)


def ply_params_from_rows(names, rows, max_sh_degree=3):
    """load_ply (:316-358): columns found by property name; f_rest_* / scale_* / rot* sorted by numeric suffix;
    features reshaped (P, 3, K) and transposed back to (P, K, 3)."""
    names = [str(n) for n in names]
    col = {n: i for i, n in enumerate(names)}
    P = rows.shape[0]
    by_suffix = lambda prefix: sorted((n for n in names if n.startswith(prefix)), key=lambda n: int(n.split("_")[-1]))
    xyz = np.stack([rows[:, col[k]] for k in ("x", "y", "z")], axis=1)
    opacity = rows[:, col["opacity"]][:, None]
    f_dc = np.stack([rows[:, col[f"f_dc_{c}"]] for c in range(3)], axis=1).reshape(P, 3, 1)
    extra = by_suffix("f_rest_")
    if len(extra) != 3 * (max_sh_degree + 1) ** 2 - 3:
        raise AssertionError("f_rest_* count does not match max_sh_degree")
    f_rest = np.stack([rows[:, col[n]] for n in extra], axis=1).reshape(P, 3, (max_sh_degree + 1) ** 2 - 1) \
        if extra else np.zeros((P, 3, 0), rows.dtype)
    scaling = np.stack([rows[:, col[n]] for n in by_suffix("scale_")], axis=1)
    rotation = np.stack([rows[:, col[n]] for n in by_suffix("rot")], axis=1)
    f32 = lambda a: np.ascontiguousarray(a, np.float32)
    return dict(xyz=f32(xyz), features_dc=f32(np.transpose(f_dc, (0, 2, 1))),
                features_rest=f32(np.transpose(f_rest, (0, 2, 1))), opacity=f32(opacity), scaling=f32(scaling),
                rotation=f32(rotation))
