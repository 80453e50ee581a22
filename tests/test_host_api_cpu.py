"""CPU tests of the host-side mirror of the reference's rasterizer interface
(gaussian_renderer/__init__.py:15,42-57,257-265): names, argument checks, error behaviour."""
import numpy as np
import pytest
import torch

import helpers as Hh
from mvs_gaussian_splatting_b200 import synthetic as S


def _settings(W=32, H=16, D=0):
    from mvs_gaussian_splatting_b200 import GaussianRasterizationSettings
    cam = S.front_camera(W, H)
    kw = S.settings_kwargs(cam, [0, 0, 0], D)
    for k in ("bg", "viewmatrix", "projmatrix", "campos"):
        kw[k] = torch.from_numpy(np.ascontiguousarray(kw[k]))
    return GaussianRasterizationSettings(**kw)


def test_import_name_the_reference_uses():
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer  # noqa: F401
    assert GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
        "sh_degree", "campos", "prefiltered", "debug")


def test_one_of_argument_checks_raise_like_the_reference():
    from mvs_gaussian_splatting_b200 import GaussianRasterizer
    r = GaussianRasterizer(_settings())
    P = 4
    m, m2, o = torch.zeros(P, 3), torch.zeros(P, 3), torch.ones(P, 1)
    sh, col = torch.zeros(P, 1, 3), torch.zeros(P, 3)
    s, q, cov = torch.ones(P, 3), torch.zeros(P, 4), torch.zeros(P, 6)
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        r(means3D=m, means2D=m2, opacities=o, scales=s, rotations=q)
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        r(means3D=m, means2D=m2, opacities=o, shs=sh, colors_precomp=col, scales=s, rotations=q)
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(means3D=m, means2D=m2, opacities=o, shs=sh)
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(means3D=m, means2D=m2, opacities=o, shs=sh, scales=s, rotations=q, cov3D_precomp=cov)


def test_no_cpu_fallback():
    from mvs_gaussian_splatting_b200 import GaussianRasterizer, mark_visible
    r = GaussianRasterizer(_settings())
    P = 4
    with pytest.raises(RuntimeError, match="no CPU path"):
        r(means3D=torch.zeros(P, 3), means2D=torch.zeros(P, 3), opacities=torch.ones(P, 1), shs=torch.zeros(P, 1, 3),
          scales=torch.ones(P, 3), rotations=torch.zeros(P, 4))
    with pytest.raises(RuntimeError, match="no CPU path"):
        mark_visible(torch.zeros(P, 3), _settings())


def test_product_package_never_imports_the_oracle():
    import os
    import re
    root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "mvs_gaussian_splatting_b200")
    for dirpath, _, files in os.walk(root):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(import|from)\s+gs_oracle|oracle/|gs_oracle\.c", src, re.M) or f.endswith((".cu", ".cuh")), f
                assert "import gs_oracle" not in src and "from gs_oracle" not in src, f


def test_gradient_arena_layout():
    from mvs_gaussian_splatting_b200 import GradientArena
    P, M = 1001, 16
    a = GradientArena(P, M, "cpu")
    assert a.flat.numel() >= 59 * P and a.flat.numel() < 59 * P + 32
    assert a.means3D.shape == (P, 3) and a.shs.shape == (P, M, 3) and a.opacities.shape == (P, 1)
    assert a.scales.shape == (P, 3) and a.rotations.shape == (P, 4)
    for v in (a.means3D, a.shs, a.opacities, a.scales, a.rotations):
        assert v.data_ptr() % 16 == 0
        assert v.untyped_storage().data_ptr() == a.flat.untyped_storage().data_ptr()
    a.shs.fill_(2.0)
    assert float(a.flat.sum()) == 2.0 * 3 * M * P


def test_capacity_hints_are_kept_per_workload():
    """rasterizer._capacity_hint: one speculative instance capacity per (device, P, W, H); set_capacity_hint installs a
    device-wide value that a workload's own entry overrides, and forgets every entry of that device."""
    from mvs_gaussian_splatting_b200 import rasterizer as R
    R._capacity_hint.clear()
    assert R._hint_lookup(0, 1000, 64, 64) is None
    R._hint_store(0, 1000, 64, 64, 5000)
    R._hint_store(0, 1000, 128, 128, 20000)          # another resolution of the same model keeps its own hint
    assert R._hint_lookup(0, 1000, 64, 64) == 5000 and R._hint_lookup(0, 1000, 128, 128) == 20000
    assert R._hint_lookup(1, 1000, 64, 64) is None   # per device
    R.set_capacity_hint("cuda:0", 16)                # device-wide: drops the per-workload entries of cuda:0 ...
    assert R._hint_lookup(0, 1000, 64, 64) == 16 and R._hint_lookup(0, 7, 8, 8) == 16
    R._hint_store(0, 1000, 64, 64, 6000)             # ... until a forward of that workload stores its own again
    assert R._hint_lookup(0, 1000, 64, 64) == 6000 and R._hint_lookup(0, 1000, 128, 128) == 16
    R.set_capacity_hint("cuda:0", None)
    assert R._hint_lookup(0, 1000, 64, 64) is None
    for i in range(3 * R._HINT_ENTRIES):             # bounded
        R._hint_store(0, i, 8, 8, 1)
    assert len(R._capacity_hint) <= R._HINT_ENTRIES
    R._capacity_hint.clear()
