"""Single-GPU parity of the multi-view slot entries (gsr_preprocess_views /
gsr_preprocess_backward_views) that the Gaussian-sharded step is built from: one multi-view call over a
row range must equal the per-view single calls of the ordinary path — forward rows bit-exact, parameter
gradients equal to the per-view gradients summed over the views (fp32 rounding of the sum only)."""
import ctypes

import numpy as np
import pytest

import helpers as Hh

pytestmark = pytest.mark.gpu


def _cams(W, H, n):
    from mvs_gaussian_splatting_b200 import synthetic as S
    cams = [S.front_camera(W, H)]
    for k in range(1, n):
        a = 2 * np.pi * k / n
        R, t = S.look_at([0.6 * np.cos(a), 0.6 * np.sin(a), 0.0], [0.0, 0.0, 10.0])
        cams.append(S.make_camera(R, t, np.radians(60.0), W, H))
    return cams


@pytest.mark.parametrize("P,row0,Ps,D,M,nv", [(6000, 1024, 3000, 3, 16, 3), (900, 0, 900, 1, 4, 2), (5000, 4096, 904, 2, 9, 5)])
def test_multiview_slot_calls_match_single_view_calls(cuda, P, row0, Ps, D, M, nv):
    import torch
    from mvs_gaussian_splatting_b200 import _native, rasterize_backward, rasterize_forward
    from mvs_gaussian_splatting_b200.rasterizer import _ptr
    from mvs_gaussian_splatting_b200.sharded import _ws_arrays
    W, H = 320, 200
    cams = _cams(W, H, nv)
    _, g = Hh.make_case(P, W, H, seed=3, sh_coeffs=M)
    t = Hh.to_dev(g, cuda)
    L = _native.lib()
    lay = _native.workspace_layout(P, W, H, 0)
    gb, _, _ = _native.workspace_bytes(P, W, H, 0)
    stream = ctypes.c_void_p(torch.cuda.current_stream(cuda).cuda_stream)
    rows = slice(row0, row0 + Ps)
    sl = {k: t[k][rows].contiguous() for k in ("means3D", "shs", "opacities", "scales", "rotations")}
    gen = torch.Generator(device="cpu").manual_seed(1)

    # ---- reference: the ordinary single-view path on all P Gaussians, per view ----
    ref = []
    for v in range(nv):
        rs = Hh.settings_for(cams[v], [0.0, 0.0, 0.0], D, cuda)
        color, radii, ctx, inp = rasterize_forward(t["means3D"], t["shs"], None, t["opacities"], t["scales"],
                                                   t["rotations"], None, rs)
        dL = torch.randn(3, H, W, generator=gen).to(cuda)
        grads = rasterize_backward(ctx, inp, rs, dL)
        torch.cuda.synchronize()
        ref.append(dict(rs=rs, ctx=ctx, radii=radii, grads=[None if x is None else x.clone() for x in grads],
                        geom={k: a.clone() for k, a in _ws_arrays(ctx.geom_ws, lay, P).items()}))

    # ---- forward: ONE multi-view call writes rows [row0, row0+Ps) of nv fresh workspaces ----
    ws = [torch.zeros(gb, dtype=torch.uint8, device=cuda) for _ in range(nv)]
    radii_mv = [torch.zeros(P, dtype=torch.int32, device=cuda) for _ in range(nv)]
    m2d = torch.zeros(nv, Ps, 3, device=cuda)
    slots = (_native.ViewSlot * nv)()
    keep = []
    for v in range(nv):
        rs = ref[v]["rs"]
        view, proj, campos = rs.viewmatrix.contiguous(), rs.projmatrix.contiguous(), rs.campos.contiguous()
        keep += [view, proj, campos]
        slots[v] = _native.ViewSlot(view.data_ptr(), proj.data_ptr(), campos.data_ptr(), float(rs.tanfovx),
                                    float(rs.tanfovy), ws[v].data_ptr(), radii_mv[v].data_ptr() + 4 * row0,
                                    m2d[v].data_ptr())
    _native.check(L.gsr_preprocess_views(P, row0, Ps, D, M, W, H, _ptr(sl["means3D"]), _ptr(sl["shs"]), None,
                                         _ptr(sl["opacities"]), _ptr(sl["scales"]), 1.0, _ptr(sl["rotations"]), None,
                                         nv, slots, stream, 0), "gsr_preprocess_views")
    torch.cuda.synchronize()
    for v in range(nv):
        assert torch.equal(radii_mv[v][rows], ref[v]["radii"][rows])
        vis = ref[v]["radii"][rows] > 0
        got = _ws_arrays(ws[v], lay, P)
        for name in ("depths", "tiles_touched"):
            assert torch.equal(got[name][rows].view(torch.int32), ref[v]["geom"][name][rows].view(torch.int32)), name
        for name in ("means2D", "conic_opacity", "rgb"):   # rows of culled Gaussians are not written
            assert torch.equal(got[name][rows][vis].view(torch.int32), ref[v]["geom"][name][rows][vis].view(torch.int32)), name

    # ---- backward: the views' accumulators come from the reference workspaces; ONE multi-view call ----
    for v in range(nv):
        slots[v].geom_ws = ref[v]["ctx"].geom_ws.data_ptr()
        slots[v].radii_slots = ref[v]["radii"].data_ptr() + 4 * row0
    f = lambda *s: torch.full(s, float("nan"), dtype=torch.float32, device=cuda)
    G = dict(means3D=f(Ps, 3), shs=f(Ps, M, 3), opacities=f(Ps, 1), scales=f(Ps, 3), rotations=f(Ps, 4))
    _native.check(L.gsr_preprocess_backward_views(
        P, row0, Ps, D, M, W, H, _ptr(sl["means3D"]), _ptr(sl["shs"]), None, _ptr(sl["scales"]), 1.0,
        _ptr(sl["rotations"]), None, nv, slots, None, _ptr(G["opacities"]), _ptr(G["means3D"]), None, _ptr(G["shs"]),
        _ptr(G["scales"]), _ptr(G["rotations"]), 0, stream, 0), "gsr_preprocess_backward_views")
    torch.cuda.synchronize()
    idx = dict(means3D=0, shs=2, opacities=4, scales=5, rotations=6)
    for k, j in idx.items():
        want = sum(ref[v]["grads"][j][rows].double() for v in range(nv))
        got = G[k].double()
        assert torch.isfinite(got).all(), k
        rel = float((got - want).abs().sum() / want.abs().sum().clamp_min(1e-30))
        assert rel <= 2e-6, (k, rel)
    for v in range(nv):
        assert torch.equal(m2d[v], ref[v]["grads"][1][rows]), f"means2D grads of view {v}"

    # accumulate = 1 adds on top (rows no view sees stay untouched)
    before = {k: x.clone() for k, x in G.items()}
    _native.check(L.gsr_preprocess_backward_views(
        P, row0, Ps, D, M, W, H, _ptr(sl["means3D"]), _ptr(sl["shs"]), None, _ptr(sl["scales"]), 1.0,
        _ptr(sl["rotations"]), None, nv, slots, None, _ptr(G["opacities"]), _ptr(G["means3D"]), None, _ptr(G["shs"]),
        _ptr(G["scales"]), _ptr(G["rotations"]), 1, stream, 0), "gsr_preprocess_backward_views")
    torch.cuda.synchronize()
    for k in G:
        assert torch.allclose(G[k], 2 * before[k], rtol=1e-6, atol=0), k
