"""Raw-parameter mode, CPU side: the oracle's numpy restatement of the reference's getters
and of the gradients autograd sends through them, against tests/golden/activations.npz — a fixture made
by CALLING the reference's own GaussianModel.get_scaling / get_rotation / get_opacity / get_features
(scene/gaussian_model.py:152-183) and autograd in the build container (tests/golden/make_golden.py)."""
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def _ulps(a, b):
    a, b = np.asarray(a, np.float32), np.asarray(b, np.float32)
    return np.abs(a.view(np.int32).astype(np.int64) - b.view(np.int32).astype(np.int64)).max()


def test_forward_activations_match_the_reference_getters(oracle):
    g = np.load(os.path.join(HERE, "golden", "activations.npz"))
    act = oracle.activate(g["raw_features_dc"], g["raw_features_rest"], g["raw_opacity"], g["raw_scaling"],
                          g["raw_rotation"])
    np.testing.assert_array_equal(act["shs"], g["features"])          # concatenation: exact
    # exp is not bit-defined by the reference (torch's vectorised CPU exp, numpy's and CUDA's libdevice exp
    # are each faithful to ~1 ulp and differ from one another by up to 2): a few ulps is the bar
    assert _ulps(act["scales"], g["scaling"]) <= 2
    assert _ulps(act["opacities"], g["opacity"]) <= 4
    assert _ulps(act["rotations"], g["rotation"]) <= 2
    assert np.array_equal(act["rotations"][0], [1, 0, 0, 0])


def test_backward_chain_rules_match_autograd(oracle):
    g = np.load(os.path.join(HERE, "golden", "activations.npz"))
    got = oracle.activate_backward(g["raw_opacity"], g["raw_scaling"], g["raw_rotation"], g["up_features"],
                                   g["up_opacity"], g["up_scaling"], g["up_rotation"])
    for k, ref in (("features_dc", "grad_features_dc"), ("features_rest", "grad_features_rest"),
                   ("opacity_raw", "grad_opacity"), ("scaling_raw", "grad_scaling"),
                   ("rotation_raw", "grad_rotation")):
        want = g[ref].astype(np.float64)
        scale = np.abs(want).mean()
        err = np.abs(got[k] - want) / np.maximum(np.abs(want), scale)
        assert err.max() <= 1e-5, (k, err.max())
