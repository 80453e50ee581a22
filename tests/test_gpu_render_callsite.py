"""SURVEY §8 row a1 — the drop-in test: the reference's `render()` call sequence, replayed literally on top of
this repo's `diff_gaussian_rasterization` import name, followed by what train.py does with the result.

Replayed statements (all paths relative to /root/reference):
  gaussian_renderer/__init__.py:15        from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
  gaussian_renderer/__init__.py:32-36     screenspace_points = zeros_like(xyz, requires_grad=True) + 0 (a NON-LEAF) ; retain_grad()
  gaussian_renderer/__init__.py:39-57     tanfov from FoV, settings by keyword with debug=pipe.debug, GaussianRasterizer(raster_settings=...)
  gaussian_renderer/__init__.py:59-88     getters; compute_cov3D_python -> pc.get_covariance; override_color -> colors_precomp
  gaussian_renderer/__init__.py:112-119   grown points: every input torch.cat((x, x[mask])) — fresh non-leaf tensors
  gaussian_renderer/__init__.py:257-269   the keyword call, `radii = radii[:origin_num]`
  gaussian_renderer/__init__.py:309-313   the returned dict, visibility_filter = radii > 0
  train.py:130-131                        max_radii2D update + add_densification_stats
  scene/gaussian_model.py:775-777         add_densification_stats on viewspace_point_tensor.grad
The model below carries the reference's getters (scene/gaussian_model.py:28-45,152-183) and its covariance
builder (utils/general_utils.py:64-110) as plain torch ops; results are checked against the CPU oracle.
"""
import math
import os

import numpy as np
import pytest

import helpers as Hh

pytestmark = pytest.mark.gpu


def _t():
    import torch   # lazily: `-m "not gpu"` collection stays light
    return torch


# ---- what the reference module does, statement for statement --------------------------------------------
def _build_rotation(r):   # utils/general_utils.py:78-99
    import torch
    norm = torch.sqrt(r[:, 0] * r[:, 0] + r[:, 1] * r[:, 1] + r[:, 2] * r[:, 2] + r[:, 3] * r[:, 3])
    q = r / norm[:, None]
    R = torch.zeros((q.size(0), 3, 3), device=r.device)
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R[:, 0, 0] = 1 - 2 * (y * y + z * z)
    R[:, 0, 1] = 2 * (x * y - r * z)
    R[:, 0, 2] = 2 * (x * z + r * y)
    R[:, 1, 0] = 2 * (x * y + r * z)
    R[:, 1, 1] = 1 - 2 * (x * x + z * z)
    R[:, 1, 2] = 2 * (y * z - r * x)
    R[:, 2, 0] = 2 * (x * z - r * y)
    R[:, 2, 1] = 2 * (y * z + r * x)
    R[:, 2, 2] = 1 - 2 * (x * x + y * y)
    return R


class _Model:
    """The slice of scene/gaussian_model.py:GaussianModel that render() and train.py:130-131 touch."""

    def __init__(self, g, device, max_sh_degree, active_sh_degree):
        import torch
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(device).requires_grad_(True)
        self._xyz = t(g["means3D"])
        self._features_dc = t(g["shs"][:, :1])
        self._features_rest = t(g["shs"][:, 1:])
        self._scaling = t(np.log(g["scales"]))
        self._rotation = t(g["rotations"] * 1.7)                         # un-normalised storage, as trained models have
        op = np.clip(g["opacities"], 1e-6, 1 - 1e-6)
        self._opacity = t(np.log(op / (1 - op)).astype(np.float32))
        self.max_sh_degree, self.active_sh_degree = max_sh_degree, active_sh_degree
        P = g["means3D"].shape[0]
        self.max_radii2D = torch.zeros(P, device=device)
        self.xyz_gradient_accum = torch.zeros((P, 1), device=device)
        self.denom = torch.zeros((P, 1), device=device)

    get_xyz = property(lambda s: s._xyz)
    get_scaling = property(lambda s: _t().exp(s._scaling))                                    # :152-154, :35
    get_rotation = property(lambda s: _t().nn.functional.normalize(s._rotation))              # :168-170, :44
    get_opacity = property(lambda s: _t().sigmoid(s._opacity))                                # :182-184, :41
    get_features = property(lambda s: _t().cat((s._features_dc, s._features_rest), dim=1))    # :176-180

    def get_covariance(self, scaling_modifier=1):   # scene/gaussian_model.py:28-32 + utils/general_utils.py:64-110
        import torch
        s = scaling_modifier * self.get_scaling
        L = torch.zeros((s.shape[0], 3, 3), dtype=torch.float, device=s.device)
        L[:, 0, 0], L[:, 1, 1], L[:, 2, 2] = s[:, 0], s[:, 1], s[:, 2]
        L = _build_rotation(self._rotation) @ L
        cov = L @ L.transpose(1, 2)
        return torch.stack((cov[:, 0, 0], cov[:, 0, 1], cov[:, 0, 2], cov[:, 1, 1], cov[:, 1, 2], cov[:, 2, 2]), dim=1)

    def add_densification_stats(self, viewspace_point_tensor, update_filter):   # scene/gaussian_model.py:775-777
        import torch
        self.xyz_gradient_accum[update_filter] += torch.norm(viewspace_point_tensor.grad[update_filter, :2], dim=-1,
                                                             keepdim=True)
        self.denom[update_filter] += 1


class _Pipe:
    def __init__(self, debug=False, compute_cov3D_python=False, convert_SHs_python=False):
        self.debug, self.compute_cov3D_python, self.convert_SHs_python = debug, compute_cov3D_python, convert_SHs_python


class _Camera:
    def __init__(self, cam, device):
        import torch
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(device)
        self.FoVx, self.FoVy = cam.fovx, cam.fovy
        self.image_height, self.image_width = cam.height, cam.width
        self.world_view_transform, self.full_proj_transform = t(cam.world_view_transform), t(cam.full_proj_transform)
        self.camera_center = t(cam.camera_center)


def render(viewpoint_camera, pc, pipe, bg_color, scaling_modifier=1.0, override_color=None, grow_mask=None):
    """gaussian_renderer/__init__.py:19-313 without the split-point bookkeeping: `grow_mask` stands for
    `selected_pts_mask` of :91-253 (every input is concatenated with its selected rows, :112-119)."""
    import torch
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer   # :15
    screenspace_points = torch.zeros_like(pc.get_xyz, dtype=pc.get_xyz.dtype, requires_grad=True, device="cuda") + 0
    try:
        screenspace_points.retain_grad()
    except Exception:
        pass
    tanfovx = math.tan(viewpoint_camera.FoVx * 0.5)
    tanfovy = math.tan(viewpoint_camera.FoVy * 0.5)
    raster_settings = GaussianRasterizationSettings(
        image_height=int(viewpoint_camera.image_height),
        image_width=int(viewpoint_camera.image_width),
        tanfovx=tanfovx,
        tanfovy=tanfovy,
        bg=bg_color,
        scale_modifier=scaling_modifier,
        viewmatrix=viewpoint_camera.world_view_transform,
        projmatrix=viewpoint_camera.full_proj_transform,
        sh_degree=pc.active_sh_degree,
        campos=viewpoint_camera.camera_center,
        prefiltered=False,
        debug=pipe.debug,
    )
    rasterizer = GaussianRasterizer(raster_settings=raster_settings)
    means3D = pc.get_xyz
    means2D = screenspace_points
    opacity = pc.get_opacity
    scales = rotations = cov3D_precomp = None
    if pipe.compute_cov3D_python:
        cov3D_precomp = pc.get_covariance(scaling_modifier)
    else:
        scales = pc.get_scaling
        rotations = pc.get_rotation
    shs = colors_precomp = None
    if override_color is None:
        shs = pc.get_features
    else:
        colors_precomp = override_color
    origin_num = means3D.shape[0]
    if grow_mask is not None:   # :112-119
        cat = lambda x: None if x is None else torch.cat((x, x[grow_mask]), 0)
        means3D, means2D, opacity, shs, colors_precomp = cat(means3D), cat(means2D), cat(opacity), cat(shs), cat(colors_precomp)
        scales, rotations, cov3D_precomp = cat(scales), cat(rotations), cat(cov3D_precomp)
    rendered_image, radii = rasterizer(
        means3D=means3D,
        means2D=means2D,
        shs=shs,
        colors_precomp=colors_precomp,
        opacities=opacity,
        scales=scales,
        rotations=rotations,
        cov3D_precomp=cov3D_precomp)
    if grow_mask is not None:
        radii = radii[:origin_num]   # :266-269
    return {"render": rendered_image, "viewspace_points": screenspace_points, "visibility_filter": radii > 0,
            "radii": radii, "selected_pts_mask": grow_mask}


# ---- the test ---------------------------------------------------------------------------------------------
def _activated(pc):
    """What the getters hand to the rasterizer, as float32 numpy (the oracle's inputs)."""
    f = lambda x: x.detach().cpu().numpy()
    return dict(means3D=f(pc.get_xyz), shs=f(pc.get_features), opacities=f(pc.get_opacity), scales=f(pc.get_scaling),
                rotations=f(pc.get_rotation))


@pytest.mark.parametrize("variant", ["default", "debug", "cov3D_python", "override_color", "grown"])
def test_render_call_sequence_and_densification_stats(cuda, oracle, variant):
    import torch
    P, W, H, D = 6000, 320, 200, 2          # active degree 2 of max degree 3: M = 16 coefficients, 9 in use
    cam, g = Hh.make_case(P, W, H, seed=31)
    pc = _Model(g, cuda, max_sh_degree=3, active_sh_degree=D)
    camera = _Camera(cam, cuda)
    bg_np = [0.1, 0.4, 0.7]
    bg = torch.tensor(bg_np, dtype=torch.float32, device=cuda)
    pipe = _Pipe(debug=(variant == "debug"), compute_cov3D_python=(variant == "cov3D_python"))
    override = None
    if variant == "override_color":
        override = torch.from_numpy(np.random.default_rng(2).uniform(0, 1, (P, 3)).astype(np.float32)).to(cuda)
    grow_mask = None
    if variant == "grown":
        grow_mask = torch.from_numpy(np.random.default_rng(3).uniform(size=P) < 0.1).to(cuda)

    # the oracle sees exactly the tensors the getters produce
    act = _activated(pc)
    kw = {}
    if variant == "cov3D_python":
        kw["cov3D_precomp"] = pc.get_covariance(1.0).detach().cpu().numpy()
    if override is not None:
        kw["colors_precomp"] = override.cpu().numpy()
    if grow_mask is not None:
        m = grow_mask.cpu().numpy()
        act = {k: np.concatenate([v, v[m]], 0) for k, v in act.items()}
    w = np.random.default_rng(4).normal(size=(3, H, W)).astype(np.float32)
    dL, ref, refre, fragile = Hh.gradient_reference(oracle, cam, act, bg_np, D, w, **kw)

    out = render(camera, pc, pipe, bg, override_color=override, grow_mask=grow_mask)
    image, viewspace_point_tensor = out["render"], out["viewspace_points"]
    visibility_filter, radii = out["visibility_filter"], out["radii"]
    assert not viewspace_point_tensor.is_leaf                       # `+ 0`: gradients arrive through retain_grad()
    loss = (image * torch.from_numpy(dL).to(cuda)).sum()
    loss.backward()
    # train.py:130-131
    pc.max_radii2D[visibility_filter] = torch.max(pc.max_radii2D[visibility_filter], radii[visibility_filter])
    pc.add_densification_stats(viewspace_point_tensor, visibility_filter)
    torch.cuda.synchronize()

    # ---- forward results
    assert tuple(image.shape) == (3, H, W) and radii.dtype == torch.int32 and radii.shape[0] == P
    np.testing.assert_array_equal(radii.cpu().numpy(), ref["radii"][:P])
    np.testing.assert_array_equal(visibility_filter.cpu().numpy(), ref["radii"][:P] > 0)
    ok = ~fragile
    good, worst = Hh.close(image.detach().cpu().numpy()[:, ok], ref["color"][:, ok], scale=1.0)
    assert good, f"image worst rel err {worst:.3e}"

    # ---- dL/dmeans2D arrives on the non-leaf screenspace_points, in the convention add_densification_stats reads
    g2 = viewspace_point_tensor.grad
    assert g2 is not None and tuple(g2.shape) == (P, 3)
    want2 = refre["grads"]["means2D"]
    if grow_mask is not None:
        # means2D = cat(screenspace_points, screenspace_points[mask]): the gradient of a duplicated row is the sum
        m = grow_mask.cpu().numpy()
        tot = {f: want2[f][:P].copy() for f in ("v", "e")}
        for f in ("v", "e"):
            tot[f][m] += want2[f][P:]
        want2 = tot
    good, rep = Hh.grad_check(g2.cpu().numpy(), want2)
    assert good, rep
    vis = ref["radii"][:P] > 0
    np.testing.assert_array_equal(pc.denom.cpu().numpy()[:, 0], vis.astype(np.float32))
    np.testing.assert_array_equal(pc.max_radii2D.cpu().numpy(), np.where(vis, ref["radii"][:P], 0).astype(np.float32))
    want_acc = np.where(vis, np.linalg.norm(g2.cpu().numpy()[:, :2].astype(np.float64), axis=1), 0.0)
    np.testing.assert_allclose(pc.xyz_gradient_accum.cpu().numpy()[:, 0], want_acc, rtol=1e-5, atol=1e-30)

    # ---- parameter gradients reach the RAW leaves through the getters' autograd
    if variant in ("default", "debug"):
        gsh = torch.cat((pc._features_dc.grad, pc._features_rest.grad), dim=1).cpu().numpy()
        good, rep = Hh.grad_check(gsh, refre["grads"]["shs"])
        assert good, rep
        good, rep = Hh.grad_check(pc._xyz.grad.cpu().numpy(), refre["grads"]["means3D"])
        assert good, rep
        # chain rule of sigmoid on the oracle's dL/dopacity
        o = act["opacities"].astype(np.float64)
        chain = (o * (1 - o))
        wo = dict(v=refre["grads"]["opacities"]["v"] * chain, e=refre["grads"]["opacities"]["e"] * chain + 4 * np.abs(
            refre["grads"]["opacities"]["v"] * chain))
        good, rep = Hh.grad_check(pc._opacity.grad.cpu().numpy(), wo)
        assert good, rep
    for p_ in (pc._xyz, pc._opacity):
        assert p_.grad is not None and bool(torch.isfinite(p_.grad).all())


def test_debug_mode_equals_normal_mode(cuda):
    """debug=True synchronises after every launch (the reference's `debug` flag); results are unchanged."""
    import torch
    cam, g = Hh.make_case(3000, 200, 120, seed=32)
    outs = []
    for debug in (False, True):
        pc = _Model(g, cuda, 3, 3)
        out = render(_Camera(cam, cuda), pc, _Pipe(debug=debug), torch.zeros(3, device=cuda))
        out["render"].sum().backward()
        outs.append((out["render"].detach().clone(), out["radii"].clone(), pc._xyz.grad.clone(),
                     out["viewspace_points"].grad.clone()))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])
    for a, b in zip(outs[0][2:], outs[1][2:]):      # atomics reorder sums: tolerance, not bits
        assert float((a - b).abs().sum()) <= 1e-5 * float(a.abs().sum())


def test_debug_mode_dumps_snapshots_on_failure(cuda, tmp_path, monkeypatch):
    """An error inside the forward (backward) with debug=True leaves snapshot_fw.dump (snapshot_bw.dump) in the
    working directory — a CPU deep copy of the arguments — and re-raises, as the reference's autograd bridge does
    (SURVEY §8a a5)."""
    import torch
    from mvs_gaussian_splatting_b200 import _native, rasterizer as Rz
    monkeypatch.chdir(tmp_path)
    cam, g = Hh.make_case(500, 96, 64, seed=33)
    pc = _Model(g, cuda, 3, 3)
    camera, bg = _Camera(cam, cuda), torch.zeros(3, device=cuda)

    # forward: the library rejects an SH degree above 3 -> GsrError out of rasterize_forward
    pc.active_sh_degree = 7
    with pytest.raises(_native.GsrError):
        render(camera, pc, _Pipe(debug=True), bg)
    assert os.path.exists("snapshot_fw.dump")
    snap = torch.load("snapshot_fw.dump", weights_only=False)
    assert isinstance(snap, tuple) and snap[0].device.type == "cpu" and tuple(snap[0].shape) == (500, 3)
    assert torch.equal(snap[0], pc._xyz.detach().cpu())
    # without debug nothing is written
    os.remove("snapshot_fw.dump")
    with pytest.raises(_native.GsrError):
        render(camera, pc, _Pipe(debug=False), bg)
    assert not os.path.exists("snapshot_fw.dump")

    # backward: make the native backward call fail once
    pc.active_sh_degree = 3
    out = render(camera, pc, _Pipe(debug=True), bg)
    real = Rz.rasterize_backward

    def boom(*a, **k):
        raise RuntimeError("injected backward failure")

    monkeypatch.setattr(Rz, "rasterize_backward", boom)
    with pytest.raises(RuntimeError, match="injected"):
        out["render"].sum().backward()
    assert os.path.exists("snapshot_bw.dump")
    snap = torch.load("snapshot_bw.dump", weights_only=False)
    assert isinstance(snap, tuple) and tuple(snap[-1].shape) == (3, 64, 96) and snap[-1].device.type == "cpu"
    monkeypatch.setattr(Rz, "rasterize_backward", real)


def test_second_backward_with_retain_graph(cuda):
    """The workspaces live as long as the graph: backward(retain_graph=True) twice accumulates twice the gradient."""
    import torch
    cam, g = Hh.make_case(2000, 160, 96, seed=34)
    pc = _Model(g, cuda, 3, 3)
    out = render(_Camera(cam, cuda), pc, _Pipe(), torch.zeros(3, device=cuda))
    loss = out["render"].sum()
    loss.backward(retain_graph=True)
    g1 = pc._xyz.grad.clone()
    loss.backward()
    assert float((pc._xyz.grad - 2 * g1).abs().sum()) <= 1e-5 * float(g1.abs().sum())
    with pytest.raises(RuntimeError):      # graph freed: autograd's own error, not an AttributeError on None
        loss.backward()
