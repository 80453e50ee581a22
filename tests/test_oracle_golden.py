"""CPU tests: the oracle against the golden fixtures generated from the reference's own Python
(tests/golden/make_golden.py imports /root/reference/utils/sh_utils.py, utils/general_utils.py,
utils/graphics_utils.py, scene/cameras.py conventions) and against paper-derivable known answers
(SURVEY 8c).  The reference has no tests of its own for this path."""
import math
import os

import numpy as np
import pytest

import helpers as Hh
from mvs_gaussian_splatting_b200 import synthetic as S

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


# ------------------------------ golden fixtures ---------------------------------------------
@pytest.mark.parametrize("deg", [0, 1, 2, 3])
def test_sh_colour_matches_reference_eval_sh(oracle, deg):
    d = np.load(os.path.join(GOLD, "sh_eval.npz"))
    rgb64, cl64 = oracle.sh_eval(d["xyz"], d["campos"], d["shs"], deg, dtype=np.float64)
    np.testing.assert_allclose(rgb64, d[f"colors_deg{deg}_f64"], rtol=0, atol=1e-13)
    np.testing.assert_array_equal(cl64, d[f"raw_deg{deg}_f64"] < 0)
    rgb32, _ = oracle.sh_eval(d["xyz"], d["campos"], d["shs"], deg, dtype=np.float32)
    # binary32: the oracle's documented FMA order vs torch's op-by-op evaluation: a few ulp
    np.testing.assert_allclose(rgb32, d[f"colors_deg{deg}_f32"], rtol=0, atol=2e-6)
    np.testing.assert_allclose(rgb32, d[f"colors_deg{deg}_f64"], rtol=0, atol=4e-6)


@pytest.mark.parametrize("mod", [1.0, 0.37])
def test_cov3d_matches_reference_build_scaling_rotation(oracle, mod):
    c = np.load(os.path.join(GOLD, "cov3d.npz"))
    want64 = c[f"cov6_mod{mod}_f64"]
    got64 = oracle.cov3d(c["scales"], c["rotations"], mod, dtype=np.float64)
    ok, worst = Hh.close(got64, want64, rel=1e-12)
    assert ok, worst
    got32 = oracle.cov3d(c["scales"], c["rotations"], mod, dtype=np.float32)
    ok, worst = Hh.close(got32, want64, rel=1e-5)
    assert ok, worst


def test_camera_matrices_match_reference_conventions():
    c = np.load(os.path.join(GOLD, "cameras.npz"))
    for i in range(int(c["n"])):
        W, H = (int(v) for v in c[f"WH{i}"])
        cam = S.make_camera(c[f"R{i}"], c[f"T{i}"], float(c[f"fovx{i}"]), W, H)
        assert abs(cam.fovy - float(c[f"fovy{i}"])) < 1e-12
        np.testing.assert_allclose(cam.world_view_transform, c[f"view{i}"], rtol=0, atol=2e-6)
        np.testing.assert_allclose(cam.projection_matrix, c[f"proj{i}"], rtol=1e-6, atol=1e-7)
        np.testing.assert_allclose(cam.full_proj_transform, c[f"full{i}"], rtol=2e-6, atol=2e-6)
        np.testing.assert_allclose(cam.camera_center, c[f"center{i}"], rtol=0, atol=5e-6)


# ------------------------------ known answers ------------------------------------------------
def _one(mean, scale, opacity=0.9, rgb=(0.2, 0.5, 0.8), quat=(1, 0, 0, 0)):
    C0 = 0.28209479177387814
    shs = np.zeros((1, 1, 3), np.float32)
    shs[0, 0] = (np.asarray(rgb, np.float64) - 0.5) / C0
    return dict(means3D=np.asarray([mean], np.float32), scales=np.asarray([[scale] * 3], np.float32),
                rotations=np.asarray([quat], np.float32), opacities=np.asarray([[opacity]], np.float32), shs=shs)


def _cat(*gs):
    return {k: np.concatenate([g[k] for g in gs], 0) for k in gs[0]}


def test_isotropic_gaussian_on_axis_closed_form(oracle):
    W = H = 64
    cam = S.front_camera(W, H)
    z, s, o = 5.0, 0.2, 0.9
    rgb, bg = np.array([0.2, 0.5, 0.8]), np.array([0.1, 0.1, 0.3])
    g = _one((0, 0, z), s, o, rgb)
    r = Hh.run_oracle(oracle, cam, g, bg, 0, dtype=np.float64)
    f = W / (2 * cam.tanfovx)
    var = (f * s / z) ** 2 + 0.3                       # EWA footprint + 0.3 px^2 low-pass
    # lambda_max = mid + sqrt(max(0.1, mid^2 - det)); isotropic => mid^2 - det = 0 => + sqrt(0.1)
    assert r["radii"][0] == math.ceil(3 * math.sqrt(var + math.sqrt(0.1)))
    px = (0 + 1) * W * 0.5 - 0.5                        # ndc 0 -> pixel centre ((ndc+1)*S-1)/2
    np.testing.assert_allclose(r["pre"]["means2D"][0], [px, px], atol=1e-9)
    ys, xs = np.mgrid[0:H, 0:W]
    d2 = (xs - px) ** 2 + (ys - px) ** 2
    alpha = np.minimum(0.99, o * np.exp(-0.5 * d2 / var))
    alpha[alpha < 1 / 255] = 0.0                        # the cut-off ring
    want = rgb[:, None, None] * alpha + (1 - alpha) * bg[:, None, None]
    # pixels outside the splat's tile rect see only background; the rect covers radius 3 sigma,
    # where alpha is already below 1/255 * ... check everywhere
    np.testing.assert_allclose(r["color"], want, atol=1e-7)
    np.testing.assert_allclose(r["fwd"]["final_T"], 1 - alpha, atol=1e-7)   # inputs are binary32-rounded
    assert set(np.unique(r["fwd"]["n_contrib"])) <= {0, 1}


def test_depth_order_and_swap(oracle):
    W = H = 32
    cam = S.front_camera(W, H)
    near = _one((0, 0, 3.0), 0.3, 0.6, (1.0, 0.0, 0.0))
    far = _one((0, 0, 6.0), 0.6, 0.6, (0.0, 0.0, 1.0))
    bg = [0, 0, 0]
    a = Hh.run_oracle(oracle, cam, _cat(near, far), bg, 0, dtype=np.float64)
    b = Hh.run_oracle(oracle, cam, _cat(far, near), bg, 0, dtype=np.float64)   # storage order swapped
    np.testing.assert_allclose(a["color"], b["color"], atol=1e-12)              # depth decides, not index
    c = a["color"][:, 16, 16]
    a_near = min(0.99, 0.6 * math.exp(-0.5 * (0.5 ** 2 * 2) / ((W / (2 * cam.tanfovx) * 0.3 / 3.0) ** 2 + 0.3)))
    assert abs(c[0] - a_near) < 1e-6 and c[2] < (1 - a_near) * 0.6 + 1e-6       # red in front, blue attenuated
    np.testing.assert_array_equal(a["binned"]["point_list"][:1], [0])
    np.testing.assert_array_equal(b["binned"]["point_list"][:1], [1])


def test_near_plane_cull_threshold(oracle):
    cam = S.front_camera(64, 64)
    z0 = np.float32(0.2)
    for z, visible in ((z0, False), (np.nextafter(z0, np.float32(1)), True), (np.nextafter(z0, np.float32(0)), False)):
        g = _one((0, 0, float(z)), 0.01)
        r = Hh.run_oracle(oracle, cam, g, [0, 0, 0], 0)
        assert (r["radii"][0] > 0) == visible, z
        assert bool(oracle.mark_visible(g["means3D"], cam.world_view_transform)[0]) == visible


def test_opaque_stack_stops_and_n_contrib(oracle):
    W = H = 16
    cam = S.front_camera(W, H)
    gs = [_one((0, 0, 2.0 + 0.1 * k), 5.0, 0.99, (1.0, 1.0, 1.0)) for k in range(8)]
    r = Hh.run_oracle(oracle, cam, _cat(*gs), [0, 0, 0], 0, dtype=np.float64)
    # alpha is capped at 0.99: T = 0.01 after one, 1e-4 after two (not < 1e-4 in exact arithmetic:
    # 0.01*0.01 == 1.0000000000000002e-4 in binary64), so the third entry trips the stop
    n = r["fwd"]["n_contrib"][8, 8]
    T = r["fwd"]["final_T"][8, 8]
    assert n in (1, 2) and T >= 1e-4 * 0.999
    assert n < 8                                        # later splats are ignored
    assert r["fwd"]["n_contrib"].max() < 8


def test_sh_negative_colour_clamps_and_blocks_gradient(oracle):
    cam = S.front_camera(32, 32)
    g = _one((0, 0, 4.0), 0.3, 0.8, (-0.4, 0.5, 0.5))
    dL = np.ones((3, 32, 32), np.float32)
    r = Hh.run_oracle(oracle, cam, g, [0, 0, 0], 0, dL_dpix=dL, dtype=np.float64)
    assert r["pre"]["rgb"][0, 0] == 0.0 and r["pre"]["clamped"][0, 0] == 1
    assert r["grads"]["shs"][0, 0, 0] == 0.0 and r["grads"]["shs"][0, 0, 1] > 0.0


@pytest.mark.parametrize("W,H", [(100, 70), (64, 1080 // 8)])
def test_tile_rect_edges_and_partial_tiles(oracle, W, H):
    cam = S.front_camera(W, H)
    gx, gy = (W + 15) // 16, (H + 15) // 16
    f = W / (2 * cam.tanfovx)
    # centre far off-screen to the right, radius reaching in; and a splat covering the whole image
    off = _one((1.6 * cam.tanfovx * 4.0, 0, 4.0), 1.2, 0.9)
    full = _one((0, 0, 2.0), 50.0, 0.5)
    r = Hh.run_oracle(oracle, cam, _cat(off, full), [0, 0, 0], 0)
    rects, tt = r["pre"]["rects"], r["pre"]["tiles_touched"]
    assert r["radii"][0] > 0 and rects[0, 2] == gx and rects[0, 0] > 0
    assert tuple(rects[1]) == (0, 0, gx, gy) and tt[1] == gx * gy
    assert r["binned"]["num_rendered"] == int(tt.sum())
    rg = r["binned"]["ranges"]
    assert (rg[:, 1] >= rg[:, 0]).all() and int((rg[:, 1] - rg[:, 0]).sum()) == int(tt.sum())
    keys = r["binned"]["keys"]
    assert (np.diff(keys.astype(np.uint64)) >= 0).all() if keys.size > 1 else True


def test_keys_are_tile_shl32_or_depth_bits_and_sort_is_stable(oracle):
    cam, g = Hh.make_case(3000, 200, 120, seed=9, sh_coeffs=1)
    g["means3D"][1::2] = g["means3D"][0::2]            # exact depth ties between neighbours
    r = Hh.run_oracle(oracle, cam, g, [0, 0, 0], 0)
    b, pre = r["binned"], r["pre"]
    ku, vu = b["keys_unsorted"], b["values_unsorted"]
    np.testing.assert_array_equal((ku & np.uint64(0xFFFFFFFF)).astype(np.uint32),
                                  pre["depths"].view(np.uint32)[vu])
    order = np.argsort(ku, kind="stable")
    np.testing.assert_array_equal(b["keys"], ku[order])
    np.testing.assert_array_equal(b["point_list"], vu[order])        # ties keep Gaussian-index order


def test_empty_inputs(oracle):
    cam = S.front_camera(48, 32)
    g = dict(means3D=np.zeros((0, 3), np.float32), scales=np.zeros((0, 3), np.float32),
             rotations=np.zeros((0, 4), np.float32), opacities=np.zeros((0, 1), np.float32),
             shs=np.zeros((0, 16, 3), np.float32))
    r = Hh.run_oracle(oracle, cam, g, [0.2, 0.3, 0.4], 3)
    assert r["binned"]["num_rendered"] == 0
    np.testing.assert_allclose(r["color"], np.broadcast_to(np.array([0.2, 0.3, 0.4], np.float32)[:, None, None], (3, 32, 48)))


# ------------------------------ gradient check -----------------------------------------------
@pytest.mark.parametrize("camera,W,H", [("front", 32, 32), ("orbit", 40, 24), ("rolled", 32, 32), ("tele", 24, 40)])
def test_backward_matches_finite_differences_f64(oracle, camera, W, H):
    """Every input tensor, under the identity camera and three general ones (rotation about all three axes incl.
    roll, translated eye, fovx != fovy): the oracle's backward through Rv (dL/dmean3D via J W), the SH view
    direction with campos != 0 and the projective divide is checked against central differences of its own
    forward, so the thing the CUDA path is compared with is itself verified away from view = I."""
    cam = Hh.named_camera(camera, W, H)
    if camera != "front":
        assert np.abs(cam.world_view_transform[:3, :3] - np.eye(3)).max() > 0.1 and np.abs(cam.camera_center).min() > 0.1
    g = S.make_gaussians(24, cam, sh_coeffs=16, seed=4, sigma_px=2.5, huge_frac=0.0, needle_frac=0.0, behind_frac=0.0)
    g = {k: v.astype(np.float64) for k, v in g.items()}
    bg = [0.3, 0.1, 0.2]
    rng = np.random.default_rng(0)
    dL = rng.normal(size=(3, H, W))

    def loss(gg):
        return float((Hh.run_oracle(oracle, cam, gg, bg, 3, dtype=np.float64, threads=1)["color"] * dL).sum())

    base = Hh.run_oracle(oracle, cam, g, bg, 3, dL_dpix=dL, dtype=np.float64, threads=1)
    vis = np.nonzero(base["radii"] > 0)[0]
    assert vis.size >= 8
    checked = 0
    for name in ("means3D", "scales", "rotations", "opacities", "shs"):
        arr = g[name]
        for _ in range(12):
            i = int(rng.choice(vis))
            idx = (i,) + tuple(int(rng.integers(0, s)) for s in arr.shape[1:])
            h = 1e-6 * max(1.0, abs(arr[idx]))
            gp = {k: v.copy() for k, v in g.items()}
            gm = {k: v.copy() for k, v in g.items()}
            gp[name][idx] += h
            gm[name][idx] -= h
            fd = (loss(gp) - loss(gm)) / (2 * h)
            an = base["grads"][name][idx]
            # discrete skip/stop decisions make the loss piecewise smooth: allow a loose tolerance
            assert abs(fd - an) <= 2e-4 * max(1.0, abs(an), abs(fd)), (name, idx, fd, an)
            checked += 1
    assert checked == 60


# ---------------------------------------------------------------------------------------------------
# running-error build (oracle/gs_oracle.c compiled with -DORC_RE): the conditioning-aware scale of the
# elementwise gradient rule (helpers.grad_check)
# ---------------------------------------------------------------------------------------------------
def test_running_error_build_values_equal_f64_and_bound_the_fp32_oracle(oracle):
    P, W, H, D = 3000, 200, 120, 3
    cam, g = Hh.make_case(P, W, H, seed=21)
    bg = [0.3, 0.2, 0.7]
    dL0 = np.random.default_rng(5).normal(size=(3, H, W))
    dL, ref32, refre, fragile = Hh.gradient_reference(oracle, cam, g, bg, D, dL0, threads=4)
    assert fragile.mean() < 0.06 and not dL[:, fragile].any()
    # same integer decisions in both runs (lists_from), same values as the plain binary64 build on them
    ref64 = Hh.run_oracle(oracle, cam, g, bg, D, dtype=np.float64, threads=4, lists_from=ref32)
    oracle.backward(ref64, dL)
    np.testing.assert_array_equal(refre["fwd"]["n_contrib"], ref32["fwd"]["n_contrib"])
    culled = ref32["radii"] == 0
    for k in ("means3D", "means2D", "shs", "opacities", "scales", "rotations"):
        v, e = refre["grads"][k]["v"], refre["grads"][k]["e"]
        np.testing.assert_allclose(v, ref64["grads"][k], rtol=1e-9, atol=1e-300, err_msg=k)
        assert (e >= 0).all() and np.isfinite(e).all(), k
        assert not v[culled].any() and not e[culled].any(), k          # exact zeros carry no error
        assert (e[v != 0] >= np.abs(v[v != 0])).all(), k            # at least its own final rounding
        # the binary32 restatement of the same algorithm sits well inside the bound, on every element
        ok, rep = Hh.grad_check(ref32["grads"][k], refre["grads"][k])
        assert ok and rep["worst_over_ue"] < 0.6, (k, rep)
        # a 1 % error on every element is far outside it
        ok, rep = Hh.grad_check(ref32["grads"][k] * 1.01, refre["grads"][k])
        assert not ok, k
        # ... and so is one on 50 elements only (no outlier budget)
        bad = ref32["grads"][k].copy().reshape(-1)
        idx = np.argsort(-np.abs(bad))[:50]
        bad[idx] *= 1.05
        ok, rep = Hh.grad_check(bad.reshape(ref32["grads"][k].shape), refre["grads"][k])
        assert not ok and rep["n_outside"] >= 1, (k, rep)


def test_running_error_scalar_rules(oracle):
    """The arithmetic of the (v, e) pairs on a case small enough to follow by hand: cov3D of an axis-aligned
    Gaussian with unit quaternion is diag(s^2) — products of exact inputs, so e is one rounding of the result
    for the diagonal and exactly zero... for entries that are sums of exact zeros only their own rounding."""
    sc = np.array([[0.5, 2.0, 3.0]], np.float32)
    q = np.array([[1.0, 0.0, 0.0, 0.0]], np.float32)
    out = oracle.cov3d(sc, q, 1.0, dtype=oracle.RE)
    np.testing.assert_array_equal(out["v"][0], [0.25, 0.0, 0.0, 4.0, 0.0, 9.0])
    assert (out["e"][0][[1, 2, 4]] == 0).all()                    # sums of exact zeros
    assert (out["e"][0][[0, 3, 5]] > 0).all() and (out["e"][0][[0, 3, 5]] < 16 * out["v"][0][[0, 3, 5]]).all()


def test_flipped_pixel_rule_on_the_cpu(oracle):
    """The rule the GPU gradient tests use (helpers.flipped_pixels + grad_check), exercised without a GPU with
    the binary32 oracle standing in for the CUDA path: dL/dpixel is zeroed only where a skip / stop decision
    actually differs between the binary32 run and the binary64 running-error run — fragile pixels whose decisions
    agree carry gradient — and every gradient element then obeys the elementwise bound.  With dL left on the
    flipped pixels some elements must leave it (that is what makes them worth excluding)."""
    P, W, H, D = 20000, 640, 360, 3           # the case in which the rule was found necessary (tests/test_gpu_fragile.py)
    cam, g = Hh.make_case(P, W, H, seed=3)
    bg = [0.3, 0.2, 0.7]
    dL_all = np.random.default_rng(203).normal(size=(3, H, W)).astype(np.float32)
    _, ref32, refre, fragile = Hh.gradient_reference(oracle, cam, g, bg, D, dL_all, threads=8, zero_fragile=False,
                                                     run_backward=False)
    flipped = Hh.flipped_pixels(ref32["fwd"]["n_contrib"], ref32["fwd"]["final_T"], ref32, refre)
    assert flipped.any() and not Hh.unflagged_flips(flipped, fragile, ref32["fwd"]["n_contrib"], ref32["fwd"]["final_T"],
                                                      ref32, refre).any()
    assert flipped.mean() < 0.005 and fragile.sum() > 20 * flipped.sum()   # almost every fragile pixel keeps its gradient
    dL = dL_all.copy()
    dL[:, flipped] = 0.0
    oracle.backward(ref32, dL)
    oracle.backward(refre, dL)
    for k in ("means3D", "means2D", "shs", "opacities", "scales", "rotations"):
        ok, rep = Hh.grad_check(ref32["grads"][k], refre["grads"][k])
        assert ok, (k, rep)
    oracle.backward(ref32, dL_all)
    oracle.backward(refre, dL_all)
    outside = sum(Hh.grad_check(ref32["grads"][k], refre["grads"][k])[1]["n_outside"]
                  for k in ("means3D", "means2D", "opacities"))
    assert outside > 0
