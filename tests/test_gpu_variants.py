"""GPU parity for the input modes and edge cases of the reference call site
(gaussian_renderer/__init__.py:65-88 selects SH vs precomputed colours and scale/rotation vs
precomputed covariance through PipelineParams flags, arguments/__init__.py:76-77) and for the
degenerate scenes a training run meets: nothing visible, one screen-filling splat, an opaque wall."""
import numpy as np
import pytest

import helpers as Hh

pytestmark = pytest.mark.gpu


def _run_gpu(cuda, cam, g, bg, D, dL, colors=None, cov=None, scale_modifier=1.0):
    import torch
    from mvs_gaussian_splatting_b200 import GaussianRasterizer
    rs = Hh.settings_for(cam, bg, D, cuda, scale_modifier=scale_modifier)
    t = Hh.to_dev(g, cuda, requires_grad=True)
    means2D = torch.zeros_like(t["means3D"], requires_grad=True)
    kw = dict(means3D=t["means3D"], means2D=means2D, opacities=t["opacities"])
    leaves = dict(means3D=t["means3D"], means2D=means2D, opacities=t["opacities"])
    if colors is None:
        kw["shs"] = leaves["shs"] = t["shs"]
    else:
        kw["colors_precomp"] = leaves["colors_precomp"] = torch.from_numpy(colors).to(cuda).requires_grad_(True)
    if cov is None:
        kw["scales"], kw["rotations"] = t["scales"], t["rotations"]
        leaves["scales"], leaves["rotations"] = t["scales"], t["rotations"]
    else:
        kw["cov3D_precomp"] = leaves["cov3D_precomp"] = torch.from_numpy(cov).to(cuda).requires_grad_(True)
    color, radii = GaussianRasterizer(rs)(**kw)
    color.backward(torch.from_numpy(dL).to(cuda))
    torch.cuda.synchronize()
    return color.detach().cpu().numpy(), radii.cpu().numpy(), {k: v.grad.cpu().numpy() for k, v in leaves.items()}


def _check(cuda, oracle, cam, g, bg, D, seed, colors=None, cov=None, scale_modifier=1.0):
    W, H = cam.width, cam.height
    kw = dict(colors_precomp=colors, cov3D_precomp=cov, scale_modifier=scale_modifier)
    dL, ref, refre, fragile = Hh.gradient_reference(oracle, cam, g, bg, D,
                                                    np.random.default_rng(seed).normal(size=(3, H, W)), **kw)
    img, radii, grads = _run_gpu(cuda, cam, g, bg, D, dL, colors, cov, scale_modifier)
    np.testing.assert_array_equal(radii, ref["radii"])
    ok = ~fragile
    good, worst = Hh.close(img[:, ok], ref["color"][:, ok], scale=1.0)
    assert good, f"image worst rel err {worst:.3e}"
    for k, a in grads.items():
        want = ref["grads"][k]
        assert want is not None and a.shape == want.shape, k
        good, rep = Hh.grad_check(a, refre["grads"][k])
        assert good, f"{k}: {rep}"
    return ref


def test_precomputed_colours(cuda, oracle):
    cam, g = Hh.make_case(3000, 200, 120, seed=11)
    colors = np.random.default_rng(5).uniform(0, 1, size=(3000, 3)).astype(np.float32)
    _check(cuda, oracle, cam, g, [0.1, 0.2, 0.3], 0, 21, colors=colors)


@pytest.mark.parametrize("mod", [1.0, 0.6])
def test_precomputed_covariance_and_scale_modifier(cuda, oracle, mod):
    cam, g = Hh.make_case(3000, 200, 120, seed=12)
    cov = oracle.cov3d(g["scales"], g["rotations"], mod).astype(np.float32)
    _check(cuda, oracle, cam, g, [1.0, 1.0, 1.0], 3, 22, cov=cov, scale_modifier=mod)


def test_scale_modifier_with_scales(cuda, oracle):
    cam, g = Hh.make_case(3000, 200, 120, seed=13)
    _check(cuda, oracle, cam, g, [0.0, 0.0, 0.0], 2, 23, scale_modifier=1.7)


def test_both_precomputed(cuda, oracle):
    cam, g = Hh.make_case(2000, 160, 96, seed=14)
    colors = np.random.default_rng(6).uniform(0, 1, size=(2000, 3)).astype(np.float32)
    cov = oracle.cov3d(g["scales"], g["rotations"], 1.0).astype(np.float32)
    _check(cuda, oracle, cam, g, [0.5, 0.5, 0.5], 0, 24, colors=colors, cov=cov)


def test_nothing_visible_gives_background_and_zero_gradients(cuda, oracle):
    cam, g = Hh.make_case(500, 96, 64, seed=15)
    g["means3D"][:, 2] = -np.abs(g["means3D"][:, 2]) - 1.0      # everything behind the camera
    bg = [0.25, 0.5, 0.75]
    dL = np.ones((3, 64, 96), np.float32)
    img, radii, grads = _run_gpu(cuda, cam, g, bg, 3, dL)
    assert not radii.any()
    np.testing.assert_array_equal(img, np.broadcast_to(np.asarray(bg, np.float32)[:, None, None], img.shape))
    for k, a in grads.items():
        assert not a.any(), k


def test_screen_filling_and_opaque_wall(cuda, oracle):
    """One huge splat covering every tile (cooperative key emission) in front of an opaque wall of
    small ones (pixels saturate: early termination, n_contrib < list length)."""
    cam, g = Hh.make_case(4000, 256, 144, seed=16)
    g["opacities"][:] = 0.98
    g["scales"][:] *= 3.0
    g["means3D"][0] = [0.0, 0.0, 2.0]
    g["scales"][0] = [3.0, 3.0, 3.0]
    g["opacities"][0] = 0.5
    ref = _check(cuda, oracle, cam, g, [0.0, 0.0, 0.0], 1, 26)
    tiles = ((256 + 15) // 16) * ((144 + 15) // 16)
    assert ref["pre"]["tiles_touched"][0] == tiles
    lens = ref["binned"]["ranges"][:, 1] - ref["binned"]["ranges"][:, 0]
    assert (ref["fwd"]["n_contrib"].reshape(-1) < lens.max()).any()


def test_mark_visible(cuda, oracle):
    import torch
    from mvs_gaussian_splatting_b200 import GaussianRasterizer
    cam, g = Hh.make_case(5000, 128, 128, seed=17)
    rs = Hh.settings_for(cam, [0, 0, 0], 0, cuda)
    got = GaussianRasterizer(rs).markVisible(torch.from_numpy(g["means3D"]).to(cuda)).cpu().numpy()
    want = oracle.mark_visible(g["means3D"], cam.world_view_transform)
    assert got.dtype == np.bool_
    np.testing.assert_array_equal(got, want)
    assert 0 < want.sum() < want.size


def test_empty_model(cuda):
    import torch
    from mvs_gaussian_splatting_b200 import GaussianRasterizer
    cam, _ = Hh.make_case(1, 64, 48, seed=0)
    rs = Hh.settings_for(cam, [0.3, 0.3, 0.3], 3, cuda)
    z = lambda *s: torch.zeros(*s, device=cuda, requires_grad=True)
    color, radii = GaussianRasterizer(rs)(means3D=z(0, 3), means2D=z(0, 3), shs=z(0, 16, 3), opacities=z(0, 1),
                                          scales=z(0, 3), rotations=z(0, 4))
    assert tuple(color.shape) == (3, 48, 64) and radii.numel() == 0
    assert not color.any()   # the reference's glue returns its zero-initialised image when P == 0


def test_non_contiguous_and_concatenated_inputs(cuda, oracle):
    """render() hands over torch.cat results and views (gaussian_renderer/__init__.py:112-119)."""
    import torch
    from mvs_gaussian_splatting_b200 import GaussianRasterizer
    cam, g = Hh.make_case(2000, 160, 96, seed=18)
    ref = Hh.run_oracle(oracle, cam, g, [0, 0, 0], 3)
    rs = Hh.settings_for(cam, [0, 0, 0], 3, cuda)
    t = Hh.to_dev(g, cuda)
    shs = torch.cat((t["shs"][:, :1, :], t["shs"][:, 1:, :]), dim=1)
    means = torch.stack((t["means3D"], t["means3D"]), dim=0)[1].t().contiguous().t()   # non-contiguous view
    assert not means.is_contiguous()
    color, radii = GaussianRasterizer(rs)(means3D=means, means2D=torch.zeros_like(t["means3D"]), shs=shs,
                                          opacities=t["opacities"], scales=t["scales"], rotations=t["rotations"])
    np.testing.assert_array_equal(radii.cpu().numpy(), ref["radii"])
    ok = ~ref["fwd"]["fragile"]
    good, worst = Hh.close(color.cpu().numpy()[:, ok], ref["color"][:, ok], scale=1.0)
    assert good, worst
