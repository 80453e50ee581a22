"""N4 on the GPU: save_ply / load_ply through gsr_ply_pack / gsr_ply_unpack.  Pure data movement:
everything is bit-exact — against the reference's own rows / tensors (tests/golden/ply.npz) and against the
oracle for ragged sizes, every SH degree, permuted and extra properties."""
import os

import numpy as np
import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
KEYS = ("xyz", "features_dc", "features_rest", "opacity", "scaling", "rotation")
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(HERE, "golden", "ply.npz"))


def params(P, M, seed):
    rng = np.random.default_rng(seed)
    shapes = dict(xyz=(P, 3), features_dc=(P, 1, 3), features_rest=(P, M - 1, 3), opacity=(P, 1), scaling=(P, 3),
                  rotation=(P, 4))
    return {k: rng.normal(size=s).astype(np.float32) for k, s in shapes.items()}


def test_save_writes_the_reference_rows_and_load_returns_the_reference_tensors(cuda, native, golden, tmp_path):
    from mvs_gaussian_splatting_b200 import ply
    path = str(tmp_path / "iteration_7" / "point_cloud.ply")
    n0 = native.lib().gsr_launch_count()
    ply.save_ply(path, *[torch.from_numpy(golden["saved_" + k]).to(cuda) for k in KEYS])
    assert native.lib().gsr_launch_count() - n0 == 1
    raw = open(path, "rb").read()
    hdr = ply.header_bytes(77, [str(n) for n in golden["names"]])
    assert raw[:len(hdr)] == hdr
    assert raw[len(hdr):] == golden["rows"].astype("<f4").tobytes()          # byte-identical body
    n0 = native.lib().gsr_launch_count()
    got = ply.load_ply(path, cuda, 3)
    assert native.lib().gsr_launch_count() - n0 == 1
    for k in KEYS:
        assert got[k].dtype == torch.float32 and got[k].is_contiguous()
        np.testing.assert_array_equal(got[k].cpu().numpy(), golden["loaded_" + k], err_msg=k)


@pytest.mark.parametrize("P,M", [(1, 16), (127, 16), (128, 16), (129, 16), (1000, 9), (333, 4), (260, 1), (0, 16)])
def test_round_trip_against_the_oracle(cuda, oracle, tmp_path, P, M):
    from mvs_gaussian_splatting_b200 import ply
    p = params(P, M, P + M)
    path = str(tmp_path / "a.ply")
    ply.save_ply(path, *[torch.from_numpy(p[k]).to(cuda) for k in KEYS])
    names, rows = oracle.ply_rows_from_params(*[p[k] for k in KEYS])
    raw = open(path, "rb").read()
    hdr = ply.header_bytes(P, names)
    assert raw == hdr + rows.astype("<f4").tobytes()
    deg = {1: 0, 4: 1, 9: 2, 16: 3}[M]
    got = ply.load_ply(path, cuda, deg)
    want = oracle.ply_params_from_rows(names, rows, deg)
    for k in KEYS:
        np.testing.assert_array_equal(got[k].cpu().numpy(), want[k].reshape(got[k].shape), err_msg=k)
    inferred = ply.load_ply(path, cuda, None)                                  # degree from the f_rest_* count
    assert inferred["features_rest"].shape == (P, M - 1, 3)


def test_permuted_and_extra_properties(cuda, oracle, tmp_path):
    from mvs_gaussian_splatting_b200 import ply
    P, M = 777, 16
    p = params(P, M, 5)
    names, rows = oracle.ply_rows_from_params(*[p[k] for k in KEYS])
    rng = np.random.default_rng(1)
    names2 = names + ["confidence", "label_0"]
    rows2 = np.concatenate([rows, rng.normal(size=(P, 2)).astype(np.float32)], axis=1)
    perm = rng.permutation(len(names2))
    names2, rows2 = [names2[i] for i in perm], np.ascontiguousarray(rows2[:, perm])
    path = str(tmp_path / "perm.ply")
    with open(path, "wb") as f:
        f.write(b"ply\nformat binary_little_endian 1.0\ncomment made by a test\nelement vertex %d\n" % P)
        f.write("".join(f"property float32 {n}\n" for n in names2).encode())
        f.write(b"element face 0\nproperty list uchar int vertex_indices\nend_header\n")
        f.write(rows2.astype("<f4").tobytes())
    got = ply.load_ply(path, cuda, 3)
    want = oracle.ply_params_from_rows(names2, rows2, 3)
    for k in KEYS:
        np.testing.assert_array_equal(got[k].cpu().numpy(), want[k], err_msg=k)
        np.testing.assert_array_equal(got[k].cpu().numpy(), p[k], err_msg=k)


def test_errors(cuda, native, tmp_path):
    from mvs_gaussian_splatting_b200 import ply
    path = str(tmp_path / "short.ply")
    names = ply.attribute_names(45)
    with open(path, "wb") as f:
        f.write(ply.header_bytes(10, names) + b"\x00" * 100)
    with pytest.raises(ValueError, match="truncated"):
        ply.load_ply(path, cuda, 3)
    with pytest.raises(ValueError):
        ply.save_ply(str(tmp_path / "b.ply"), torch.zeros(4, 3, device=cuda), torch.zeros(4, 1, 3, device=cuda),
                     torch.zeros(4, 15, 3, device=cuda), torch.zeros(3, 1, device=cuda), torch.zeros(4, 3, device=cuda),
                     torch.zeros(4, 4, device=cuda))
    cols, _ = ply.resolve_columns(names, 3)
    cols.opacity = cols.rot[0]                                                   # a column named twice
    import ctypes
    buf = torch.zeros(62 * 4, device=cuda)
    assert native.lib().gsr_ply_unpack(4, ctypes.byref(cols), *[ctypes.c_void_p(buf.data_ptr())] * 7, None) == 1


def test_million_gaussians_round_trip(cuda, tmp_path):
    from mvs_gaussian_splatting_b200 import ply
    P = 1_000_003
    gen = torch.Generator(device="cpu").manual_seed(2)
    t = [torch.randn(P, *s, generator=gen).to(cuda) for s in ((3,), (1, 3), (15, 3), (1,), (3,), (4,))]
    path = str(tmp_path / "big.ply")
    ply.save_ply(path, *t)
    assert os.path.getsize(path) == len(ply.header_bytes(P, ply.attribute_names(45))) + P * 62 * 4
    got = ply.load_ply(path, cuda, 3)
    for k, x in zip(KEYS, t):
        assert torch.equal(got[k], x), k
