"""Raw-parameter entry point on the GPU.

1. `gsr_activate` (the library's activation kernel) against tests/golden/activations.npz — the outputs of the
   reference's own GaussianModel getters — within a few ulps, and against torch's ops on the same device.
2. Fused == unfused, BIT FOR BIT: `rasterize_forward_raw` on the raw tensors equals `rasterize_forward` on
   `activate(...)` of them: radii, sorted key list, point list, ranges and the image are identical.
3. Gradients w.r.t. the raw tensors equal what autograd produces through the reference's PyTorch
   activations wrapped around the standard (oracle-checked) rasterizer path, within 1e-4 relative.
"""
import os

import numpy as np
import pytest

import helpers as Hh

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _ulps(a, b):
    a, b = np.asarray(a, np.float32), np.asarray(b, np.float32)
    return int(np.abs(a.view(np.int32).astype(np.int64) - b.view(np.int32).astype(np.int64)).max())


def _raw_case(P, W, H, M, seed, device):
    """Raw parameters whose activations are a synthetic scene of the usual statistics."""
    import torch
    cam, g = Hh.make_case(P, W, H, seed=seed, sh_coeffs=M)
    rng = np.random.default_rng(seed + 50)
    raw = dict(means3D=g["means3D"], features_dc=g["shs"][:, :1].copy(),
               features_rest=g["shs"][:, 1:].copy() if M > 1 else None,
               scaling_raw=np.log(g["scales"]).astype(np.float32),
               rotation_raw=(g["rotations"] * np.exp(rng.normal(0, 1, (P, 1)))).astype(np.float32),
               opacity_raw=np.log(g["opacities"] / (1 - g["opacities"])).astype(np.float32))
    t = {k: (None if v is None else torch.from_numpy(np.ascontiguousarray(v)).to(device)) for k, v in raw.items()}
    return cam, t


def test_activation_kernel_matches_the_reference_getters(cuda):
    import torch
    from mvs_gaussian_splatting_b200 import activate
    g = np.load(os.path.join(HERE, "golden", "activations.npz"))
    t = {k: torch.from_numpy(g[k]).to(cuda) for k in g.files if k.startswith("raw_")}
    shs, op, sc, rot = activate(t["raw_features_dc"], t["raw_features_rest"], t["raw_opacity"], t["raw_scaling"],
                                t["raw_rotation"])
    np.testing.assert_array_equal(shs.cpu().numpy(), g["features"])
    assert _ulps(sc.cpu().numpy(), g["scaling"]) <= 2
    assert _ulps(op.cpu().numpy(), g["opacity"]) <= 4
    assert _ulps(rot.cpu().numpy(), g["rotation"]) <= 2
    # same device, torch's own kernels (what the reference would run on this GPU)
    assert _ulps(sc.cpu().numpy(), torch.exp(t["raw_scaling"]).cpu().numpy()) <= 1
    assert _ulps(op.cpu().numpy(), torch.sigmoid(t["raw_opacity"]).cpu().numpy()) <= 2
    assert _ulps(rot.cpu().numpy(), torch.nn.functional.normalize(t["raw_rotation"]).cpu().numpy()) <= 2


@pytest.mark.parametrize("P,W,H,D,M,seed", [(20000, 640, 360, 3, 16, 3), (3000, 200, 120, 0, 1, 1),
                                            (5000, 320, 200, 1, 4, 2), (100000, 800, 800, 3, 16, 0)])
def test_fused_equals_unfused_bit_for_bit(cuda, P, W, H, D, M, seed):
    import torch
    from mvs_gaussian_splatting_b200 import activate, rasterize_forward, rasterize_forward_raw
    cam, t = _raw_case(P, W, H, M, seed, cuda)
    rs = Hh.settings_for(cam, [0.1, 0.5, 0.9], D, cuda)
    shs, op, sc, rot = activate(t["features_dc"], t["features_rest"], t["opacity_raw"], t["scaling_raw"],
                                t["rotation_raw"])
    c0, r0, ctx0, _ = rasterize_forward(t["means3D"], shs, None, op, sc, rot, None, rs)
    c1, r1, ctx1, _ = rasterize_forward_raw(t["means3D"], t["features_dc"], t["features_rest"], t["opacity_raw"],
                                            t["scaling_raw"], t["rotation_raw"], rs)
    torch.cuda.synchronize()
    assert torch.equal(r0, r1)
    assert ctx0.num_rendered == ctx1.num_rendered and ctx0.num_visible == ctx1.num_visible
    R = ctx0.num_rendered
    for name in ("keys", "point_list"):
        assert torch.equal(ctx0.array(name)[:R], ctx1.array(name)[:R]), name
    assert torch.equal(ctx0.array("ranges"), ctx1.array("ranges"))
    vis = r0 > 0
    for name in ("conic_opacity", "rgb", "means2D"):
        assert torch.equal(ctx0.array(name)[vis].view(torch.int32), ctx1.array(name)[vis].view(torch.int32)), name
    assert torch.equal(c0.view(torch.int32), c1.view(torch.int32))


@pytest.mark.parametrize("P,W,H,D,M,seed", [(20000, 640, 360, 3, 16, 3), (3000, 200, 120, 0, 1, 1),
                                            (5000, 320, 200, 2, 9, 2)])
def test_raw_gradients_match_autograd_through_the_reference_activations(cuda, P, W, H, D, M, seed):
    import torch
    from mvs_gaussian_splatting_b200 import GaussianRasterizer
    cam, t = _raw_case(P, W, H, M, seed, cuda)
    rs = Hh.settings_for(cam, [0.3, 0.2, 0.7], D, cuda)
    dL = torch.from_numpy(np.random.default_rng(seed).normal(size=(3, H, W)).astype(np.float32)).to(cuda)
    names = ["means3D", "features_dc", "opacity_raw", "scaling_raw", "rotation_raw"] + (["features_rest"] if M > 1 else [])

    def leaves():
        return {k: t[k].clone().requires_grad_(True) for k in names}

    # reference composition: PyTorch activations (scene/gaussian_model.py:152-183) around the standard path
    a = leaves()
    shs = a["features_dc"] if M == 1 else torch.cat((a["features_dc"], a["features_rest"]), dim=1)
    m2a = torch.zeros_like(a["means3D"], requires_grad=True)
    color, radii = GaussianRasterizer(rs)(means3D=a["means3D"], means2D=m2a, shs=shs,
                                          opacities=torch.sigmoid(a["opacity_raw"]), scales=torch.exp(a["scaling_raw"]),
                                          rotations=torch.nn.functional.normalize(a["rotation_raw"]))
    color.backward(dL)
    # fused
    b = leaves()
    m2b = torch.zeros_like(b["means3D"], requires_grad=True)
    color2, radii2 = GaussianRasterizer(rs).forward_raw(b["means3D"], m2b, b["features_dc"], b.get("features_rest"),
                                                        b["opacity_raw"], b["scaling_raw"], b["rotation_raw"])
    color2.backward(dL)
    torch.cuda.synchronize()
    same = radii == radii2   # torch's exp/normalize may differ from the library's by an ulp: a handful of radii may flip
    assert float((~same).float().mean()) <= 1e-4
    good, worst = Hh.close(color2.detach().cpu().numpy(), color.detach().cpu().numpy(), scale=1.0)
    assert good, worst
    for k in names + ["means2D"]:
        ga = (m2a if k == "means2D" else a[k]).grad.double()
        gb = (m2b if k == "means2D" else b[k]).grad.double()
        assert torch.isfinite(gb).all(), k
        rel = float((ga - gb).abs().sum() / ga.abs().sum().clamp_min(1e-30))
        assert rel <= 1e-4, (k, rel)
        assert not gb[radii2 == 0].any(), k    # culled rows: exact zeros
