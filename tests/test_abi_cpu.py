"""CPU tests of the drop-in boundary: the C-ABI library builds for sm_100a, loads, exports every
symbol include/gsrast_b200.h declares, and its host-only entry points behave (no compute calls)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "gsrast_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"GSR_API\s+[\w\s\*]+?\b(gsr_\w+)\s*\(", src)))


def test_header_and_binding_list_the_same_symbols(native):
    assert _declared_symbols() == sorted(native.EXPORTS)


def test_library_exports_every_declared_symbol(native):
    L = ctypes.CDLL(native.LIB_PATH)
    for name in _declared_symbols():
        assert hasattr(L, name), name
    assert native.lib().gsr_abi_version() == native.ABI_VERSION


def test_library_is_sm100a_only(native):
    import subprocess
    out = subprocess.run(["cuobjdump", "--list-elf", native.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


def test_status_strings(native):
    L = native.lib()
    assert L.gsr_status_string(0) == b"ok"
    for code in (1, 2, 3, 4, 99):
        assert L.gsr_status_string(code)
    assert L.gsr_last_cuda_error() == b""
    names = [L.gsr_stage_name(i).decode() for i in range(L.gsr_stage_count())]
    assert names == ["preprocess", "depth_sort", "emit_keys", "tile_ranges", "radix_sort", "blend_forward",
                     "blend_backward", "preprocess_backward"]


def test_workspace_sizes_and_layout(native):
    P, W, H, C = 100_000, 800, 800, 250_000
    g, b, i = native.workspace_bytes(P, W, H, C)
    L = native.workspace_layout(P, W, H, C)
    assert (g, b, i) == (L.geom_bytes, L.binning_bytes, L.image_bytes)
    # per Gaussian: depth 4 + mean2D 8 + conic 16 + rgb 16 + tiles 4 + accumulators 48
    # + depth-order sort (key, index) x 2 = 16, plus the sort's control words (1 KB per 6144 Gaussians x 4 passes)
    assert g >= 112 * P and g < 114 * P + (1 << 20)
    # per instance: three (u16 tile id, u32 Gaussian index) pairs (up to 65536 tiles; u32 tile ids beyond)
    assert L.b_key_bytes == 2 and b >= 18 * C and b < 19 * C + (1 << 20)
    assert native.workspace_layout(P, 8192, 8192, C).b_key_bytes == 4
    tiles = 50 * 50
    assert i >= 8 * tiles + 8 * W * H
    offs = [L.g_depths, L.g_means2D, L.g_conic_opacity, L.g_rgb, L.g_tiles_touched, L.g_info,
            L.g_grad_accum, L.g_order_keys, L.g_order, L.g_order_keys_alt, L.g_order_alt, L.g_sort_ctl]
    assert offs == sorted(offs) and all(o % 256 == 0 for o in offs)
    # the geometry and image layouts do not depend on the instance capacity
    L0 = native.workspace_layout(P, W, H, 0)
    assert (L0.geom_bytes, L0.image_bytes) == (L.geom_bytes, L.image_bytes)
    # 3 M Gaussians at 1080p with 10 M instances stays far below 180 GB
    g3, b3, i3 = native.workspace_bytes(3_000_000, 1920, 1080, 10_000_000)
    assert g3 + b3 + i3 < 2 << 30


def test_invalid_arguments_are_reported_not_crashed(native):
    L = native.lib()
    z = ctypes.c_size_t()
    assert L.gsr_workspace_bytes(-1, 10, 10, 0, ctypes.byref(z), ctypes.byref(z), ctypes.byref(z)) == native.GSR_ERR_INVALID_ARGUMENT
    assert L.gsr_workspace_bytes(10, 0, 10, 0, ctypes.byref(z), ctypes.byref(z), ctypes.byref(z)) == native.GSR_ERR_INVALID_ARGUMENT
    assert L.gsr_workspace_bytes(10, 10, 10, 1 << 33, ctypes.byref(z), ctypes.byref(z), ctypes.byref(z)) == native.GSR_ERR_UNSUPPORTED
    assert L.gsr_workspace_layout(10, 10, 10, 0, None) == native.GSR_ERR_INVALID_ARGUMENT
    # NULL workspace: rejected before anything touches the device
    st = L.gsr_forward_geometry(1, 0, 1, 16, 16, None, None, None, None, None, 1.0, None, None, None, None, None,
                                1.0, 1.0, 0, None, None, None, None, 0)
    assert st == native.GSR_ERR_WORKSPACE
    st = L.gsr_mark_visible(-1, None, None, None, None, None)
    assert st == native.GSR_ERR_INVALID_ARGUMENT
    with pytest.raises(native.GsrError) as e:
        native.check(native.GSR_ERR_WORKSPACE, "here")
    assert "workspace" in str(e.value)


def test_static_scratch_entry_point_refuses_oversized_input_on_the_host(native):
    """gsr_loss_value_and_grad_static checks its size limit before it touches the device, so the refusal can
    be seen without a GPU (the pointers are never dereferenced on the host)."""
    L = native.lib()
    fake = ctypes.c_void_p(0x1000)
    C, W = 3, 1920
    H = native.LOSS_STATIC_MAX_ELEMS // (C * W) + 1
    st = L.gsr_loss_value_and_grad_static(C, H, W, fake, fake, 0.2, 0.8, -0.2, None, fake, fake, None)
    assert st == native.GSR_ERR_SCRATCH_LIMIT
    assert b"scratch" in L.gsr_status_string(st)
    assert L.gsr_loss_value_and_grad_static(C, 0, W, fake, fake, 0.2, 0.8, -0.2, None, fake, fake, None) \
        == native.GSR_ERR_INVALID_ARGUMENT
    hdr = open(os.path.join(ROOT, "include", "gsrast_b200.h")).read()
    assert re.search(r"#define\s+GSR_LOSS_STATIC_MAX_ELEMS\s+\(3 \* 1080 \* 1920\)", hdr)
    assert native.LOSS_STATIC_MAX_ELEMS == 3 * 1080 * 1920


def test_missing_library_fails_loudly(native, monkeypatch, tmp_path):
    monkeypatch.setattr(native, "_lib", None)
    monkeypatch.setattr(native, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(native.NativeLibraryError):
        native.lib()


def test_every_stage_the_library_times_has_a_cost_model_in_the_bench(native):
    """bench.py divides each stage's measured time into an algorithmic byte / flop count: a stage the library
    reports without a model would silently vanish from the roofline table."""
    import importlib.util
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("bench_for_test", os.path.join(root, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    L = native.lib()
    names = [L.gsr_stage_name(i).decode() for i in range(L.gsr_stage_count())]
    counts = dict(P=1000, P_visible=800, R=3000, tiles=100, pairs_evaluated=10 ** 6, pairs_blended=10 ** 5,
                  sum_n_contrib=10 ** 6, views_rendered_by_this_rank=1, depth_counts_in_preprocess=True)
    roofs = bench.stage_rooflines({n: 0.1 for n in names}, counts, bench.load_peaks(), None)
    assert sorted(roofs) == sorted(names)
    for n, r in roofs.items():
        assert r["achieved"] > 0 and r["peak"] > 0 and r["bound"] in ("hbm", "fp32_simt_issue"), n
