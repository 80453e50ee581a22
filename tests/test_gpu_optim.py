"""SURVEY 8f N2 on the GPU: FusedAdam (one launch for all groups) and densification_stats against the
fixtures made with the reference's recipe, and against torch.optim.Adam itself on the same device."""
import os

import numpy as np
import pytest

from test_optim_cpu import within_ulps

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
GROUPS = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")


def test_fused_adam_matches_the_reference_optimizer(cuda):
    import torch
    from mvs_gaussian_splatting_b200 import _native
    from mvs_gaussian_splatting_b200.optim import FusedAdam
    g = np.load(os.path.join(HERE, "golden", "adam.npz"))
    params = {k: torch.nn.Parameter(torch.from_numpy(g[f"p0_{k}"]).to(cuda)) for k in GROUPS}
    l = [{'params': [params[k]], 'lr': float(g["lrs"][i]), "name": k} for i, k in enumerate(GROUPS)]
    opt = FusedAdam(l, lr=0.0, eps=1e-15)                       # as scene/gaussian_model.py:262
    for it in range(int(g["steps"])):
        for k in GROUPS:
            params[k].grad = torch.from_numpy(g[f"g{it}_{k}"]).to(cuda)
        n0 = _native.launch_count()
        opt.step()
        assert _native.launch_count() - n0 == 1                 # all six groups in ONE launch
        opt.zero_grad(set_to_none=True)
        for k in GROUPS:
            assert within_ulps(params[k].detach().cpu().numpy(), g[f"p{it + 1}_{k}"], g[f"p{it}_{k}"]), f"{k} step {it + 1}"
    for k in GROUPS:
        st = opt.state[params[k]]
        np.testing.assert_allclose(st["exp_avg"].cpu().numpy(), g[f"m_{k}"], rtol=2e-6,
                                   atol=2e-6 * np.abs(g[f"m_{k}"]).max())   # signed running mean
        np.testing.assert_allclose(st["exp_avg_sq"].cpu().numpy(), g[f"v_{k}"], rtol=2e-6, atol=1e-30)
        assert int(st["step"].item()) == int(g["steps"])
    assert [grp["name"] for grp in opt.param_groups] == list(GROUPS)


def test_fused_adam_tracks_torch_adam_on_device_at_scale(cuda):
    import torch
    from mvs_gaussian_splatting_b200.optim import FusedAdam
    gen = torch.Generator(device="cpu").manual_seed(3)
    P = 200_003
    shapes = dict(xyz=(P, 3), f_rest=(P, 15, 3), rotation=(P, 4))
    a = {k: torch.nn.Parameter(torch.randn(*s, generator=gen).to(cuda)) for k, s in shapes.items()}
    b = {k: torch.nn.Parameter(v.detach().clone()) for k, v in a.items()}
    lrs = dict(xyz=1.6e-4, f_rest=1.25e-4, rotation=1e-3)
    oa = FusedAdam([{'params': [a[k]], 'lr': lrs[k], "name": k} for k in shapes], lr=0.0, eps=1e-15)
    ob = torch.optim.Adam([{'params': [b[k]], 'lr': lrs[k], "name": k} for k in shapes], lr=0.0, eps=1e-15)
    for it in range(4):
        for k in shapes:
            gr = (torch.randn(*shapes[k], generator=gen) * 10.0 ** (-it)).to(cuda)
            a[k].grad, b[k].grad = gr, gr.clone()
        oa.step()
        ob.step()
    torch.cuda.synchronize()
    for k in shapes:
        rel = float((a[k] - b[k]).abs().max() / b[k].abs().max())
        assert rel <= 1e-6, (k, rel)


def test_densification_stats_match_the_reference_lines(cuda):
    import torch
    from mvs_gaussian_splatting_b200.optim import densification_stats
    g = np.load(os.path.join(HERE, "golden", "densification.npz"))
    t = {k: torch.from_numpy(g[k]).to(cuda) for k in g.files}
    mr, acc, den = t["max_radii2D0"].clone(), t["accum0"].clone(), t["denom0"].clone()
    densification_stats(t["radii"], t["grad"], mr, acc, den)
    np.testing.assert_array_equal(mr.cpu().numpy(), g["max_radii2D1"])
    np.testing.assert_allclose(acc.cpu().numpy(), g["accum1"], rtol=2e-7)
    np.testing.assert_array_equal(den.cpu().numpy(), g["denom1"])
