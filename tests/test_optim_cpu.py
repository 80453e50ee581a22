"""SURVEY 8f N2, CPU side: the oracle's restatement of the reference's optimizer step and densification
statistics against fixtures produced by the reference's own recipe (tests/golden/make_golden.py:
torch.optim.Adam(l, lr=0.0, eps=1e-15) over the six named groups with arguments/__init__.py's learning
rates; train.py:130-131 and scene/gaussian_model.py:775-777 verbatim)."""
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
GROUPS = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")


def within_ulps(got, want, before, n=8):
    """|got - want| <= n ulps of the larger of the parameter before and after the step (the rounding
    granularity of p - update) + 2e-6 of the update itself (exp_avg / exp_avg_sq carry a few ulps of
    binary32 rounding from step to step, and torch's vectorised CPU kernels contract differently)."""
    mag = np.maximum(np.abs(before), np.abs(want)).astype(np.float32)
    tol = n * np.spacing(mag) + 2e-6 * np.abs(want.astype(np.float64) - before)
    return bool((np.abs(got.astype(np.float64) - want) <= tol).all())


def test_adam_steps_match_torch_optim_adam(oracle):
    g = np.load(os.path.join(HERE, "golden", "adam.npz"))
    for gi, k in enumerate(GROUPS):
        p = g[f"p0_{k}"].copy()
        m, v = np.zeros_like(p), np.zeros_like(p)
        for it in range(int(g["steps"])):
            oracle.adam_step(p, g[f"g{it}_{k}"], m, v, float(g["lrs"][gi]), it + 1)
            assert within_ulps(p, g[f"p{it + 1}_{k}"], g[f"p{it}_{k}"]), f"{k} step {it + 1}"
        np.testing.assert_allclose(m, g[f"m_{k}"], rtol=2e-6, atol=2e-6 * np.abs(g[f"m_{k}"]).max())  # signed running mean
        np.testing.assert_allclose(v, g[f"v_{k}"], rtol=2e-6, atol=1e-30)


def test_densification_stats_match_the_reference_lines(oracle):
    g = np.load(os.path.join(HERE, "golden", "densification.npz"))
    mr, acc, den = g["max_radii2D0"].copy(), g["accum0"].copy(), g["denom0"].copy()
    oracle.densification_stats(g["radii"], g["grad"], mr, acc, den)
    np.testing.assert_array_equal(mr, g["max_radii2D1"])
    np.testing.assert_allclose(acc, g["accum1"], rtol=2e-7)
    np.testing.assert_array_equal(den, g["denom1"])
