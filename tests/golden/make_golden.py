"""Generates the golden fixtures in this directory by IMPORTING the reference's own Python code
from /root/reference (run in the build container only; the GPU box has no /root/reference, so the
.npz outputs are committed next to this script).

What the reference tree holds for the rasterizer path is the Python restatement of two preprocess
sub-steps plus the camera conventions:
  utils/sh_utils.py:57-112            eval_sh            -> sh_eval.npz
  gaussian_renderer/__init__.py:80-84 direction / +0.5 / clamp recipe around it
  utils/general_utils.py:64-110       build_rotation, build_scaling_rotation, strip_symmetric
  scene/gaussian_model.py:28-32       Sigma = L L^T                      -> cov3d.npz
  utils/graphics_utils.py:38-71       getWorld2View2, getProjectionMatrix
  scene/cameras.py:54-57              transposes, full_proj, camera_center -> cameras.npz
  scene/gaussian_model.py:28-45,152-183  the getters the rasterizer's inputs come from (exp, normalize,
                                         sigmoid, cat of _features_dc | _features_rest) AND their autograd
                                         gradients                        -> activations.npz
                                         (scene.gaussian_model imports simple_knn and plyfile, which are
                                         absent here and unused by the getters: both are stubbed)
  scene/gaussian_model.py:245-262 + arguments/__init__.py:83-95  the optimizer: torch.optim.Adam(l, lr=0.0,
                                         eps=1e-15) over the six named groups with the reference's learning
                                         rates, stepped on seeded gradients   -> adam.npz
  train.py:130-131 + scene/gaussian_model.py:775-777  max_radii2D / xyz_gradient_accum / denom updates
                                                                             -> densification.npz
  utils/loss_utils.py:17-63 + train.py:99-101  l1_loss, ssim and the training loss
                                         (1 - lambda) L1 + lambda (1 - SSIM) with its autograd gradient
                                         w.r.t. the rendered image             -> loss.npz
  scene/gaussian_model.py:279-358    save_ply / load_ply: the structured array save_ply hands to plyfile
                                         (property names in order + the row matrix) and the six tensors
                                         load_ply builds from it.  plyfile itself is absent here: a stand-in
                                         captures the array save_ply describes and serves it back to load_ply
                                         by property name, so what is pinned is the reference's column order,
                                         feature transposition and name sorting — not plyfile's container
                                         syntax, which is the PLY standard           -> ply.npz
Those helpers hard-code device="cuda"; torch.zeros is patched to ignore the device so they run on
the CPU unmodified.

    python tests/golden/make_golden.py
"""
import math
import os
import sys

import numpy as np
import torch

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REF)

_zeros = torch.zeros


def _cpu_zeros(*a, **k):
    k.pop("device", None)
    return _zeros(*a, **k)


torch.zeros = _cpu_zeros
from utils.sh_utils import eval_sh  # noqa: E402
from utils.general_utils import build_rotation, build_scaling_rotation, strip_symmetric  # noqa: E402
from utils.graphics_utils import getWorld2View2, getProjectionMatrix  # noqa: E402


def sh_fixture():
    rng = np.random.Generator(np.random.PCG64(11))
    P = 512
    xyz = rng.normal(0, 3, (P, 3))
    campos = np.array([0.3, -0.2, 0.1])
    shs = rng.normal(0, 0.6, (P, 16, 3))  # [P, coeff, rgb] as the rasterizer receives it
    out = {"xyz": xyz, "campos": campos, "shs": shs}
    for dt, tag in ((torch.float64, "f64"), (torch.float32, "f32")):
        feats = torch.from_numpy(shs).to(dt)
        x = torch.from_numpy(xyz).to(dt)
        c = torch.from_numpy(campos).to(dt)
        for deg in range(4):
            # gaussian_renderer/__init__.py:80-84
            shs_view = feats.transpose(1, 2).view(-1, 3, 16)
            dir_pp = x - c.repeat(feats.shape[0], 1)
            dir_pp_normalized = dir_pp / dir_pp.norm(dim=1, keepdim=True)
            sh2rgb = eval_sh(deg, shs_view, dir_pp_normalized)
            out[f"colors_deg{deg}_{tag}"] = torch.clamp_min(sh2rgb + 0.5, 0.0).numpy()
            out[f"raw_deg{deg}_{tag}"] = (sh2rgb + 0.5).numpy()
    np.savez_compressed(os.path.join(HERE, "sh_eval.npz"), **out)


def cov_fixture():
    rng = np.random.Generator(np.random.PCG64(12))
    P = 512
    scales = np.exp(rng.normal(-2, 1, (P, 3)))
    q = rng.normal(0, 1, (P, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)  # the rasterizer uses q as given; the helper normalises
    out = {"scales": scales, "rotations": q}
    for mod in (1.0, 0.37):
        for dt, tag in ((torch.float64, "f64"), (torch.float32, "f32")):
            s = torch.from_numpy(scales).to(dt)
            r = torch.from_numpy(q).to(dt)
            if dt == torch.float32:
                L = build_scaling_rotation(mod * s, r)           # the helper as shipped (float32 buffers)
            else:
                # the helper allocates float32 buffers; for a float64 truth run its rotation formula
                # with the allocation patched to float64 and compose L = R diag(s) as it does
                torch.zeros = lambda *a, **k: _cpu_zeros(*a, **{**k, "dtype": torch.float64})
                Rm = build_rotation(r)
                torch.zeros = _cpu_zeros
                L = Rm @ torch.diag_embed(mod * s)
            cov = L @ L.transpose(1, 2)
            if dt == torch.float32:
                sym = strip_symmetric(cov)
            else:
                sym = torch.stack([cov[:, 0, 0], cov[:, 0, 1], cov[:, 0, 2], cov[:, 1, 1], cov[:, 1, 2], cov[:, 2, 2]], 1)
            out[f"cov6_mod{mod}_{tag}"] = sym.numpy()
    np.savez_compressed(os.path.join(HERE, "cov3d.npz"), **out)


def camera_fixture():
    rng = np.random.Generator(np.random.PCG64(13))
    out = {}
    n = 6
    out["n"] = np.array(n)
    for i in range(n):
        A = rng.normal(size=(3, 3))
        Q, _ = np.linalg.qr(A)
        if np.linalg.det(Q) < 0:
            Q[:, 0] = -Q[:, 0]
        if i == 0:
            Q = np.eye(3)
        T = rng.normal(0, 2, 3) if i else np.zeros(3)
        fovx = math.radians(40 + 10 * i)
        W, H = [(1920, 1080), (800, 800), (100, 70), (3840, 2160), (1008, 756), (640, 360)][i]
        fovy = 2 * math.atan(H / (2 * (W / (2 * math.tan(fovx / 2)))))
        # scene/cameras.py:54-57 on the CPU
        wvt = torch.tensor(getWorld2View2(Q, T, np.array([0.0, 0.0, 0.0]), 1.0)).transpose(0, 1)
        proj = getProjectionMatrix(znear=0.01, zfar=100.0, fovX=fovx, fovY=fovy).transpose(0, 1)
        full = (wvt.unsqueeze(0).bmm(proj.unsqueeze(0))).squeeze(0)
        center = wvt.inverse()[3, :3]
        out[f"R{i}"], out[f"T{i}"] = Q, T
        out[f"fovx{i}"], out[f"fovy{i}"], out[f"WH{i}"] = np.array(fovx), np.array(fovy), np.array([W, H])
        out[f"view{i}"], out[f"proj{i}"] = wvt.contiguous().numpy(), proj.contiguous().numpy()
        out[f"full{i}"], out[f"center{i}"] = full.contiguous().numpy(), center.numpy()
    np.savez_compressed(os.path.join(HERE, "cameras.npz"), **out)


def activation_fixture():
    """Raw parameters -> GaussianModel.get_scaling / get_rotation / get_opacity / get_features of the
    reference, plus the gradients autograd sends back through them for seeded upstream gradients."""
    import types
    for name in ("simple_knn", "simple_knn._C", "plyfile"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["simple_knn._C"].distCUDA2 = None
    sys.modules["plyfile"].PlyData = sys.modules["plyfile"].PlyElement = None
    from scene.gaussian_model import GaussianModel
    gm = GaussianModel.__new__(GaussianModel)
    gm.setup_functions()
    gen = torch.Generator().manual_seed(11)
    P = 512
    gm._scaling = (torch.randn(P, 3, generator=gen) * 1.5 - 3.0).requires_grad_(True)
    rot = torch.randn(P, 4, generator=gen) * torch.exp(torch.randn(P, 1, generator=gen) * 2.0)
    rot[0] = torch.tensor([1.0, 0.0, 0.0, 0.0])
    gm._rotation = rot.requires_grad_(True)
    gm._opacity = (torch.randn(P, 1, generator=gen) * 4.0).requires_grad_(True)
    gm._features_dc = torch.randn(P, 1, 3, generator=gen).requires_grad_(True)
    gm._features_rest = (torch.randn(P, 15, 3, generator=gen) * 0.15).requires_grad_(True)
    scaling, rotation, opacity, feats = gm.get_scaling, gm.get_rotation, gm.get_opacity, gm.get_features
    ups = dict(scaling=torch.randn(P, 3, generator=gen), rotation=torch.randn(P, 4, generator=gen),
               opacity=torch.randn(P, 1, generator=gen), features=torch.randn(P, 16, 3, generator=gen))
    loss = (scaling * ups["scaling"]).sum() + (rotation * ups["rotation"]).sum() + \
           (opacity * ups["opacity"]).sum() + (feats * ups["features"]).sum()
    loss.backward()
    out = dict(raw_scaling=gm._scaling, raw_rotation=gm._rotation, raw_opacity=gm._opacity,
               raw_features_dc=gm._features_dc, raw_features_rest=gm._features_rest,
               scaling=scaling, rotation=rotation, opacity=opacity, features=feats,
               up_scaling=ups["scaling"], up_rotation=ups["rotation"], up_opacity=ups["opacity"],
               up_features=ups["features"], grad_scaling=gm._scaling.grad, grad_rotation=gm._rotation.grad,
               grad_opacity=gm._opacity.grad, grad_features_dc=gm._features_dc.grad,
               grad_features_rest=gm._features_rest.grad)
    np.savez_compressed(os.path.join(HERE, "activations.npz"),
                        **{k: v.detach().numpy() for k, v in out.items()})


def adam_fixture():
    """The optimizer exactly as GaussianModel.training_setup builds it, with OptimizationParams' defaults."""
    from argparse import ArgumentParser
    from arguments import OptimizationParams
    opt = OptimizationParams(ArgumentParser())
    gen = torch.Generator().manual_seed(21)
    P, spatial_lr_scale = 301, 4.2          # P*3, P*45 ... not multiples of 4: exercises the kernel's tail
    shapes = dict(xyz=(P, 3), f_dc=(P, 1, 3), f_rest=(P, 15, 3), opacity=(P, 1), scaling=(P, 3), rotation=(P, 4))
    lrs = dict(xyz=opt.position_lr_init * spatial_lr_scale, f_dc=opt.feature_lr, f_rest=opt.feature_lr / 20.0,
               opacity=opt.opacity_lr, scaling=opt.scaling_lr, rotation=opt.rotation_lr)
    params = {k: torch.nn.Parameter(torch.randn(*s, generator=gen)) for k, s in shapes.items()}
    l = [{'params': [params[k]], 'lr': lrs[k], "name": k} for k in shapes]
    optimizer = torch.optim.Adam(l, lr=0.0, eps=1e-15)          # scene/gaussian_model.py:262
    out = {f"p0_{k}": v.detach().clone().numpy() for k, v in params.items()}
    out["lrs"] = np.array([lrs[k] for k in shapes], np.float64)
    steps = 5
    for it in range(steps):
        for k, v in params.items():
            scale = 10.0 ** float(torch.randint(-6, 1, (1,), generator=gen))
            g = torch.randn(*shapes[k], generator=gen) * scale
            if it == 2:
                g[::3] = 0.0                                     # rows of culled Gaussians have zero gradient
            v.grad = g
            out[f"g{it}_{k}"] = g.numpy().copy()
        optimizer.step()
        for k, v in params.items():
            out[f"p{it + 1}_{k}"] = v.detach().clone().numpy()
    for k, v in params.items():
        st = optimizer.state[v]
        out[f"m_{k}"], out[f"v_{k}"] = st["exp_avg"].numpy().copy(), st["exp_avg_sq"].numpy().copy()
    out["steps"] = np.array(steps)
    np.savez_compressed(os.path.join(HERE, "adam.npz"), **out)


def densification_fixture():
    gen = torch.Generator().manual_seed(31)
    P = 1000
    max_radii2D = torch.rand(P, generator=gen) * 30                       # scene/gaussian_model.py:238 (float)
    xyz_gradient_accum, denom = torch.rand(P, 1, generator=gen), torch.randint(0, 5, (P, 1), generator=gen).float()
    radii = torch.randint(-1, 60, (P,), generator=gen).clamp_min(0).int()
    grad = torch.randn(P, 3, generator=gen)
    out = dict(max_radii2D0=max_radii2D.clone(), accum0=xyz_gradient_accum.clone(), denom0=denom.clone(),
               radii=radii, grad=grad)
    visibility_filter = radii > 0                                          # gaussian_renderer/__init__.py:311
    max_radii2D[visibility_filter] = torch.max(max_radii2D[visibility_filter], radii[visibility_filter])   # train.py:130
    xyz_gradient_accum[visibility_filter] += torch.norm(grad[visibility_filter, :2], dim=-1, keepdim=True)  # gaussian_model.py:776
    denom[visibility_filter] += 1                                          # gaussian_model.py:777
    out.update(max_radii2D1=max_radii2D, accum1=xyz_gradient_accum, denom1=denom)
    np.savez_compressed(os.path.join(HERE, "densification.npz"), **{k: v.numpy() for k, v in out.items()})


def ply_fixture():
    import types
    for name in ("simple_knn", "simple_knn._C", "plyfile"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["simple_knn._C"].distCUDA2 = None
    captured = {}

    class PlyElement:
        @staticmethod
        def describe(elements, name):
            captured["elements"], captured["name"] = elements.copy(), name
            class El:
                data = captured["elements"]
                properties = [types.SimpleNamespace(name=n) for n in elements.dtype.names]

                def __getitem__(self, k):
                    return self.data[k]
            return El()

    class PlyData:
        def __init__(self, elements):
            self.elements = elements

        def write(self, path):
            captured["path"] = path

        @staticmethod
        def read(path):
            return PlyData([PlyElement.describe(captured["elements"], "vertex")])

    sys.modules["plyfile"].PlyData, sys.modules["plyfile"].PlyElement = PlyData, PlyElement
    import scene.gaussian_model as gmod
    gmod.PlyData, gmod.PlyElement = PlyData, PlyElement
    gmod.mkdir_p = lambda d: None
    _tensor = torch.tensor
    gmod.torch.tensor = lambda *a, **k: _tensor(*a, **{kk: v for kk, v in k.items() if kk != "device"})
    try:
        gen = torch.Generator().manual_seed(51)
        P = 77
        gm = gmod.GaussianModel.__new__(gmod.GaussianModel)
        torch.nn.Module.__init__(gm) if isinstance(gm, torch.nn.Module) else None
        gm.max_sh_degree = 3
        gm._xyz = torch.randn(P, 3, generator=gen)
        gm._features_dc = torch.randn(P, 1, 3, generator=gen)
        gm._features_rest = torch.randn(P, 15, 3, generator=gen)
        gm._opacity = torch.randn(P, 1, generator=gen)
        gm._scaling = torch.randn(P, 3, generator=gen)
        gm._rotation = torch.randn(P, 4, generator=gen)
        gm.save_ply("/tmp/unused/point_cloud.ply")
        el = captured["elements"]
        names = list(el.dtype.names)
        rows = np.stack([el[n] for n in names], axis=1).astype(np.float32)
        g2 = gmod.GaussianModel.__new__(gmod.GaussianModel)
        g2.max_sh_degree = 3
        g2.load_ply("/tmp/unused/point_cloud.ply")
        out = dict(names=np.array(names), rows=rows)
        for k in ("_xyz", "_features_dc", "_features_rest", "_opacity", "_scaling", "_rotation"):
            out["saved" + k] = getattr(gm, k).detach().numpy()
            out["loaded" + k] = getattr(g2, k).detach().numpy()
        np.savez_compressed(os.path.join(HERE, "ply.npz"), **out)
    finally:
        gmod.torch.tensor = _tensor


def loss_fixture():
    """The photometric loss exactly as train.py:99-101 forms it from utils/loss_utils.py (CPU torch), with
    autograd's gradient w.r.t. the rendered image.  Sizes: ragged (not tile multiples), narrower than the
    11-tap window, one pixel, and a rendered image equal to the ground truth in places (|x-y| = 0)."""
    from argparse import ArgumentParser
    from arguments import OptimizationParams
    from utils.loss_utils import l1_loss, ssim
    lam = OptimizationParams(ArgumentParser()).lambda_dssim
    gen = torch.Generator().manual_seed(41)
    out = dict(lambda_dssim=np.array(lam, np.float64))
    for n, (H, W) in enumerate(((37, 53), (9, 7), (1, 1), (64, 96))):
        gt = torch.rand(3, H, W, generator=gen)
        if n == 3:   # smooth images: small local variance, the regime rendered images live in
            yy, xx = torch.meshgrid(torch.linspace(0, 1, H), torch.linspace(0, 1, W), indexing="ij")
            gt = torch.stack([0.5 + 0.4 * torch.sin(6 * xx + c) * torch.cos(4 * yy) for c in range(3)])
        image = (gt + 0.1 * torch.randn(3, H, W, generator=gen)).clamp(0, 1)
        image[:, ::4, ::5] = gt[:, ::4, ::5]
        image.requires_grad_(True)
        Ll1 = l1_loss(image, gt)
        s = ssim(image, gt)
        loss = (1.0 - lam) * Ll1 + lam * (1.0 - s)           # train.py:100-101
        loss.backward()
        g_loss = image.grad.clone()
        image.grad = None
        ssim(image, gt).backward()
        out.update({f"image{n}": image.detach().numpy(), f"gt{n}": gt.numpy(), f"l1_{n}": Ll1.detach().numpy(),
                    f"ssim{n}": s.detach().numpy(), f"loss{n}": loss.detach().numpy(), f"grad{n}": g_loss.numpy(),
                    f"grad_ssim{n}": image.grad.numpy().copy()})
    out["cases"] = np.array(4)
    np.savez_compressed(os.path.join(HERE, "loss.npz"), **out)


if __name__ == "__main__":
    ply_fixture()
    loss_fixture()
    adam_fixture()
    densification_fixture()
    activation_fixture()
    sh_fixture()
    cov_fixture()
    camera_fixture()
    print("wrote", [f for f in os.listdir(HERE) if f.endswith(".npz")])
