"""N4: Gaussian .ply layout.  The oracle's restatement of save_ply / load_ply
(oracle/gs_oracle.py: ply_rows_from_params, ply_params_from_rows) against tests/golden/ply.npz — the
structured array the reference's save_ply builds and the tensors its load_ply returns (make_golden.py::
ply_fixture) — and the host logic of mvs_gaussian_splatting_b200/ply.py (header, name -> column resolution)."""
import io
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
KEYS = ("xyz", "features_dc", "features_rest", "opacity", "scaling", "rotation")


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(HERE, "golden", "ply.npz"))


def test_oracle_rows_are_the_reference_rows(oracle, golden):
    names, rows = oracle.ply_rows_from_params(*[golden["saved_" + k] for k in KEYS])
    assert names == [str(n) for n in golden["names"]]
    np.testing.assert_array_equal(rows, golden["rows"])            # data movement: bit-exact
    assert names == oracle.ply_attribute_names()


def test_oracle_load_is_the_reference_load(oracle, golden):
    back = oracle.ply_params_from_rows(golden["names"], golden["rows"], 3)
    for k in KEYS:
        np.testing.assert_array_equal(back[k], golden["loaded_" + k], err_msg=k)
        np.testing.assert_array_equal(back[k], golden["saved_" + k], err_msg=k)   # save -> load is the identity
    with pytest.raises(AssertionError):
        oracle.ply_params_from_rows(golden["names"], golden["rows"], 2)


def test_header_round_trip_and_column_resolution(oracle, golden):
    from mvs_gaussian_splatting_b200 import ply
    names = [str(n) for n in golden["names"]]
    assert ply.attribute_names(45) == names
    hdr = ply.header_bytes(77, names)
    assert hdr.startswith(b"ply\nformat binary_little_endian 1.0\nelement vertex 77\nproperty float x\n")
    assert hdr.endswith(b"property float rot_3\nend_header\n")
    n, got, off = ply.read_header(io.BytesIO(hdr + b"\x00" * 16))
    assert (n, got, off) == (77, names, len(hdr))
    cols, M = ply.resolve_columns(names, 3)
    assert M == 16 and cols.n_props == 62 and cols.n_rest == 45
    # the map must reproduce load_ply on the golden rows
    rows = golden["rows"]
    np.testing.assert_array_equal(rows[:, list(cols.f_rest)].reshape(-1, 15, 3), golden["loaded_features_rest"])
    np.testing.assert_array_equal(rows[:, list(cols.f_dc)].reshape(-1, 1, 3), golden["loaded_features_dc"])
    np.testing.assert_array_equal(rows[:, list(cols.xyz)], golden["loaded_xyz"])
    np.testing.assert_array_equal(rows[:, [cols.opacity]], golden["loaded_opacity"])
    np.testing.assert_array_equal(rows[:, list(cols.scale)], golden["loaded_scaling"])
    np.testing.assert_array_equal(rows[:, list(cols.rot)], golden["loaded_rotation"])
    # property order in the file is free: resolution is by name
    perm = np.random.default_rng(0).permutation(len(names))
    cols2, _ = ply.resolve_columns([names[i] for i in perm], None)
    np.testing.assert_array_equal(rows[:, perm][:, list(cols2.f_rest)], rows[:, list(cols.f_rest)])


def test_header_errors():
    from mvs_gaussian_splatting_b200 import ply
    with pytest.raises(ValueError):
        ply.read_header(io.BytesIO(b"plx\n"))
    with pytest.raises(NotImplementedError):
        ply.read_header(io.BytesIO(b"ply\nformat ascii 1.0\nelement vertex 1\nproperty float x\nend_header\n"))
    with pytest.raises(NotImplementedError):
        ply.read_header(io.BytesIO(b"ply\nformat binary_little_endian 1.0\nelement vertex 1\nproperty uchar red\nend_header\n"))
    with pytest.raises(ValueError):
        ply.read_header(io.BytesIO(b"ply\nformat binary_little_endian 1.0\nelement vertex 1\nproperty float x\n"))
    names = ply.attribute_names(45)
    with pytest.raises(AssertionError):
        ply.resolve_columns(names, 2)                      # scene/gaussian_model.py:331
    with pytest.raises(ValueError):
        ply.resolve_columns([n for n in names if n != "opacity"], 3)
    with pytest.raises(RuntimeError, match="CUDA"):
        ply.load_ply("/nonexistent.ply", device="cpu")
