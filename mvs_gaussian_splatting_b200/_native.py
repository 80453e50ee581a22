"""ctypes binding of libgsrast_b200.so — the C ABI declared in include/gsrast_b200.h.

This is the only place the package touches native code.  There is NO fallback: if the shared
library is missing or does not export the ABI this module raises, and every op built on it
fails loudly (north_star: "no CPU fallback").  The library replaces the pybind11 module `_C`
of the reference's rasterizer dependency (imported at
/root/reference/gaussian_renderer/__init__.py:15).
"""
from __future__ import annotations

import ctypes
import functools
import os
import subprocess
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
# GSRAST_B200_LIB: A/B builds of the same ABI during kernel tuning (never a different backend)
LIB_PATH = os.environ.get("GSRAST_B200_LIB") or os.path.join(_HERE, "lib", "libgsrast_b200.so")
CSRC_DIR = os.path.join(_HERE, "csrc")
ABI_VERSION = 9

GSR_OK = 0
GSR_ERR_INVALID_ARGUMENT = 1
GSR_ERR_CUDA = 2
GSR_ERR_WORKSPACE = 3
GSR_ERR_UNSUPPORTED = 4
GSR_ERR_SCRATCH_LIMIT = 5
LOSS_STATIC_MAX_ELEMS = 3 * 1080 * 1920   # GSR_LOSS_STATIC_MAX_ELEMS

# Every symbol include/gsrast_b200.h declares (tests check the .so exports all of them).
EXPORTS = (
    "gsr_abi_version", "gsr_status_string", "gsr_last_cuda_error", "gsr_workspace_bytes",
    "gsr_workspace_layout", "gsr_forward_geometry", "gsr_forward_render", "gsr_backward",
    "gsr_mark_visible", "gsr_launch_count", "gsr_stage_count", "gsr_stage_name",
    "gsr_stage_timing_enable", "gsr_stage_timing_read", "gsr_blend_pair_counts",
    "gsr_preprocess_slots", "gsr_depth_order", "gsr_blend_backward", "gsr_preprocess_backward_slots",
    "gsr_preprocess_views", "gsr_preprocess_backward_views",
    "gsr_forward_geometry_raw", "gsr_backward_raw", "gsr_activate",
    "gsr_adam_step", "gsr_densification_stats",
    "gsr_loss_workspace_bytes", "gsr_loss_forward", "gsr_loss_backward",
    "gsr_ply_unpack", "gsr_ply_pack", "gsr_materialize_keys",
    "gsr_loss_value_and_grad_static",
)
PLY_MAX_PROPS = 96
MAX_ADAM_GROUPS = 16
MAX_VIEWS = 16


class PlyColumns(ctypes.Structure):
    """gsr_ply_columns (include/gsrast_b200.h)."""
    _fields_ = [("n_props", ctypes.c_int32), ("n_rest", ctypes.c_int32), ("xyz", ctypes.c_int16 * 3),
                ("f_dc", ctypes.c_int16 * 3), ("f_rest", ctypes.c_int16 * 45), ("opacity", ctypes.c_int16),
                ("scale", ctypes.c_int16 * 3), ("rot", ctypes.c_int16 * 4)]


class ViewSlot(ctypes.Structure):
    """gsr_view_slot of include/gsrast_b200.h."""
    _fields_ = [("viewmatrix", ctypes.c_void_p), ("projmatrix", ctypes.c_void_p), ("campos", ctypes.c_void_p),
                ("tanfovx", ctypes.c_float), ("tanfovy", ctypes.c_float), ("geom_ws", ctypes.c_void_p),
                ("radii_slots", ctypes.c_void_p), ("dL_dmeans2D", ctypes.c_void_p),
                ("radii_local", ctypes.c_void_p), ("conic_local", ctypes.c_void_p)]


LOCAL_RADIUS_MASK = 0x0FFFFFFF   # GSR_LOCAL_RADIUS_MASK


class AdamGroup(ctypes.Structure):
    """gsr_adam_group of include/gsrast_b200.h."""
    _fields_ = [("param", ctypes.c_void_p), ("grad", ctypes.c_void_p), ("exp_avg", ctypes.c_void_p),
                ("exp_avg_sq", ctypes.c_void_p), ("n", ctypes.c_int64), ("lr", ctypes.c_double)]


class ForwardInfo(ctypes.Structure):
    _fields_ = [("num_rendered", ctypes.c_int32), ("overflow", ctypes.c_int32),
                ("num_visible", ctypes.c_int32), ("reserved", ctypes.c_int32)]


class Layout(ctypes.Structure):
    _fields_ = [(n, ctypes.c_size_t) for n in (
        "g_depths", "g_means2D", "g_conic_opacity", "g_rgb", "g_tiles_touched",
        "g_info", "g_grad_accum", "g_order_keys", "g_order", "g_order_keys_alt", "g_order_alt", "g_sort_ctl",
        "geom_bytes",
        "b_keys_unsorted", "b_vals_unsorted", "b_keys", "b_point_list", "b_keys_alt", "b_vals_alt",
        "b_sort_ctl", "binning_bytes", "b_key_bytes",
        "i_ranges", "i_final_T", "i_n_contrib", "image_bytes")]


class NativeLibraryError(RuntimeError):
    pass


class GsrError(RuntimeError):
    def __init__(self, status: int, where: str):
        L = lib()
        msg = L.gsr_status_string(status).decode()
        cuda = L.gsr_last_cuda_error().decode()
        super().__init__(f"{where}: {msg}" + (f" ({cuda})" if status == GSR_ERR_CUDA and cuda else ""))
        self.status = status


def build(verbose: bool = False, extra_nvflags: str = "") -> str:
    """Compile the CUDA sources for sm_100a with the in-tree Makefile (nvcc cross-compiles
    without a GPU).  Output stays in-tree so it travels to the GPU box."""
    cmd = ["make", "-C", CSRC_DIR, "-j", str(os.cpu_count() or 4)]
    if extra_nvflags:
        cmd.append(f"EXTRA_NVFLAGS={extra_nvflags}")
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if verbose or r.returncode != 0:
        print(r.stdout)
    if r.returncode != 0:
        raise NativeLibraryError("building libgsrast_b200.so failed:\n" + r.stdout[-4000:])
    return LIB_PATH


_lib = None
_lock = threading.Lock()
_vp, _i, _f, _i64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_int64


def lib() -> ctypes.CDLL:
    """Load (once) and type the shared library.  Raises NativeLibraryError if it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise NativeLibraryError(
                f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "or `make -C mvs_gaussian_splatting_b200/csrc`. There is no CPU or PyTorch fallback.")
        L = ctypes.CDLL(LIB_PATH)
        for name in EXPORTS:
            if not hasattr(L, name):
                raise NativeLibraryError(f"{LIB_PATH} does not export {name}")
        L.gsr_abi_version.restype = _i
        if L.gsr_abi_version() != ABI_VERSION:
            raise NativeLibraryError(f"ABI version mismatch: library {L.gsr_abi_version()}, binding {ABI_VERSION}")
        L.gsr_status_string.restype = ctypes.c_char_p
        L.gsr_status_string.argtypes = [_i]
        L.gsr_last_cuda_error.restype = ctypes.c_char_p
        L.gsr_launch_count.restype = _i64
        L.gsr_workspace_bytes.argtypes = [_i, _i, _i, _i64] + [ctypes.POINTER(ctypes.c_size_t)] * 3
        L.gsr_workspace_layout.argtypes = [_i, _i, _i, _i64, ctypes.POINTER(Layout)]
        L.gsr_forward_geometry.argtypes = [
            _i, _i, _i, _i, _i,          # P D M W H
            _vp, _vp, _vp, _vp,          # means3D shs colors_precomp opacities
            _vp, _f, _vp, _vp,           # scales scale_modifier rotations cov3D_precomp
            _vp, _vp, _vp, _f, _f,       # view proj campos tanfovx tanfovy
            _i, _vp, _vp, _vp, _vp, _i]  # prefiltered radii geom_ws h_info stream debug
        L.gsr_forward_render.argtypes = [
            _i, _i, _i, _vp, _vp,        # P W H bg radii
            _vp, _vp, _i64, _vp,         # geom_ws binning_ws capacity image_ws
            _vp, _vp, _vp, _i]           # out_color h_info stream debug
        L.gsr_backward.argtypes = [
            _i, _i, _i, _i, _i, _vp,     # P D M W H bg
            _vp, _vp, _vp,               # means3D shs colors_precomp
            _vp, _f, _vp, _vp,           # scales scale_modifier rotations cov3D_precomp
            _vp, _vp, _vp, _f, _f, _vp,  # view proj campos tanfovx tanfovy radii
            _vp, _vp, _i64, _vp,         # geom_ws binning_ws capacity image_ws
            _vp,                         # dL_dpix
            _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp,  # dL_d{means2D,colors,opacity,means3D,cov3D,sh,scales,rots}
            _vp, _i]                     # stream debug
        L.gsr_mark_visible.argtypes = [_i, _vp, _vp, _vp, _vp, _vp]
        L.gsr_preprocess_slots.argtypes = [
            _i, _i, _i, _i, _i, _i, _i,  # P_ws slot_offset Ps D M W H
            _vp, _vp, _vp, _vp,          # means3D shs colors_precomp opacities (slice rows)
            _vp, _f, _vp, _vp,           # scales scale_modifier rotations cov3D_precomp
            _vp, _vp, _vp, _f, _f,       # view proj campos tanfovx tanfovy
            _vp, _vp, _vp, _i]           # radii_slots geom_ws stream debug
        L.gsr_depth_order.argtypes = [_i, _i, _i, _vp, _vp, _vp, _i]  # P W H geom_ws h_info stream debug
        L.gsr_blend_backward.argtypes = [_i, _i, _i, _vp, _vp, _vp, _i64, _vp, _vp, _vp, _i]
        L.gsr_preprocess_backward_slots.argtypes = [
            _i, _i, _i, _i, _i, _i, _i,  # P_ws slot_offset Ps D M W H
            _vp, _vp, _vp,               # means3D shs colors_precomp (slice rows)
            _vp, _f, _vp, _vp,           # scales scale_modifier rotations cov3D_precomp
            _vp, _vp, _vp, _f, _f, _vp,  # view proj campos tanfovx tanfovy radii_slots
            _vp,                         # geom_ws
            _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp,  # dL_d{means2D,colors,opacity,means3D,cov3D,sh,scales,rots}
            _i, _vp, _i]                 # accumulate stream debug
        L.gsr_blend_pair_counts.argtypes = [_i, _i, _i, _vp, _vp, _i64, _vp, _vp, _vp]
        L.gsr_blend_pair_counts.restype = _i
        L.gsr_materialize_keys.argtypes = [_i, _i, _i, _vp, _vp, _i64, _i, _vp, _vp]
        L.gsr_materialize_keys.restype = _i
        L.gsr_stage_count.restype = _i
        L.gsr_stage_name.restype = ctypes.c_char_p
        L.gsr_stage_name.argtypes = [_i]
        L.gsr_stage_timing_enable.argtypes = [_i]
        L.gsr_stage_timing_enable.restype = _i
        L.gsr_stage_timing_read.argtypes = [ctypes.POINTER(ctypes.c_double), ctypes.POINTER(_i64), _i]
        L.gsr_stage_timing_read.restype = _i
        L.gsr_preprocess_views.argtypes = [
            _i, _i, _i, _i, _i, _i, _i,  # P_ws slot_offset Ps D M W H
            _vp, _vp, _vp, _vp,          # means3D shs colors_precomp opacities
            _vp, _f, _vp, _vp,           # scales scale_modifier rotations cov3D_precomp
            _i, ctypes.POINTER(ViewSlot), _vp, _i]  # n_views views stream debug
        L.gsr_preprocess_backward_views.argtypes = [
            _i, _i, _i, _i, _i, _i, _i,  # P_ws slot_offset Ps D M W H
            _vp, _vp, _vp,               # means3D shs colors_precomp
            _vp, _f, _vp, _vp,           # scales scale_modifier rotations cov3D_precomp
            _i, ctypes.POINTER(ViewSlot),
            _vp, _vp, _vp, _vp, _vp, _vp, _vp,  # dL_d{colors,opacity,means3D,cov3D,sh,scales,rots}
            _i, _vp, _i]                 # accumulate stream debug
        L.gsr_forward_geometry_raw.argtypes = [
            _i, _i, _i, _i, _i,          # P D M W H
            _vp, _vp, _vp, _vp,          # means3D features_dc features_rest opacity_raw
            _vp, _f, _vp,                # scaling_raw scale_modifier rotation_raw
            _vp, _vp, _vp, _f, _f,       # view proj campos tanfovx tanfovy
            _vp, _vp, _vp, _vp, _i]      # radii geom_ws h_info stream debug
        L.gsr_backward_raw.argtypes = [
            _i, _i, _i, _i, _i, _vp,     # P D M W H bg
            _vp, _vp, _vp,               # means3D features_dc features_rest
            _vp, _f, _vp,                # scaling_raw scale_modifier rotation_raw
            _vp, _vp, _vp, _f, _f, _vp,  # view proj campos tanfovx tanfovy radii
            _vp, _vp, _i64, _vp,         # geom_ws binning_ws capacity image_ws
            _vp,                         # dL_dpix
            _vp, _vp, _vp, _vp, _vp, _vp, _vp,  # dL_d{means2D,opacity_raw,means3D,features_dc,features_rest,scaling_raw,rotation_raw}
            _vp, _i]                     # stream debug
        L.gsr_activate.argtypes = [_i, _i] + [_vp] * 10
        for name in ("gsr_forward_geometry_raw", "gsr_backward_raw", "gsr_activate"):
            getattr(L, name).restype = _i
        _d = ctypes.c_double
        L.gsr_adam_step.argtypes = [_i, ctypes.POINTER(AdamGroup), _d, _d, _d, _i, _vp]
        L.gsr_densification_stats.argtypes = [_i, _vp, _vp, _vp, _vp, _vp, _vp]
        L.gsr_adam_step.restype = _i
        L.gsr_loss_workspace_bytes.argtypes = [_i, _i, _i, _i, ctypes.POINTER(ctypes.c_size_t)]
        L.gsr_loss_forward.argtypes = [_i, _i, _i, _vp, _vp, _d, _d, _d, _vp, _i, _vp, _vp]
        L.gsr_loss_backward.argtypes = [_i, _i, _i, _vp, _vp, _d, _d, _vp, _vp, _vp, _vp]
        for name in ("gsr_loss_workspace_bytes", "gsr_loss_forward", "gsr_loss_backward"):
            getattr(L, name).restype = _i
        L.gsr_loss_value_and_grad_static.argtypes = [_i, _i, _i, _vp, _vp, _d, _d, _d, _vp, _vp, _vp, _vp]
        L.gsr_loss_value_and_grad_static.restype = _i
        for name in ("gsr_ply_unpack", "gsr_ply_pack"):
            getattr(L, name).argtypes = [ctypes.c_int64, ctypes.POINTER(PlyColumns)] + [_vp] * 8
            getattr(L, name).restype = _i
        L.gsr_densification_stats.restype = _i
        L.gsr_preprocess_views.restype = _i
        L.gsr_preprocess_backward_views.restype = _i
        for name in ("gsr_workspace_bytes", "gsr_workspace_layout", "gsr_forward_geometry",
                     "gsr_forward_render", "gsr_backward", "gsr_mark_visible", "gsr_preprocess_slots",
                     "gsr_depth_order", "gsr_blend_backward", "gsr_preprocess_backward_slots"):
            getattr(L, name).restype = _i
        _lib = L
    return _lib


def check(status: int, where: str) -> None:
    if status != GSR_OK:
        raise GsrError(status, where)


@functools.lru_cache(maxsize=256)
def workspace_bytes(P: int, W: int, H: int, capacity: int):
    g, b, i = ctypes.c_size_t(), ctypes.c_size_t(), ctypes.c_size_t()
    check(lib().gsr_workspace_bytes(P, W, H, capacity, ctypes.byref(g), ctypes.byref(b), ctypes.byref(i)),
          "gsr_workspace_bytes")
    return g.value, b.value, i.value


@functools.lru_cache(maxsize=256)
def workspace_layout(P: int, W: int, H: int, capacity: int) -> Layout:
    out = Layout()
    check(lib().gsr_workspace_layout(P, W, H, capacity, ctypes.byref(out)), "gsr_workspace_layout")
    return out


def launch_count() -> int:
    return int(lib().gsr_launch_count())


def stage_timing_enable(on: bool) -> None:
    check(lib().gsr_stage_timing_enable(1 if on else 0), "gsr_stage_timing_enable")


def stage_timing_read() -> dict:
    """{stage name: (summed ms, number of brackets)} since the last read; waits for the events."""
    L = lib()
    n = L.gsr_stage_count()
    ms = (ctypes.c_double * n)()
    calls = (_i64 * n)()
    check(L.gsr_stage_timing_read(ms, calls, n), "gsr_stage_timing_read")
    return {L.gsr_stage_name(i).decode(): (float(ms[i]), int(calls[i])) for i in range(n)}
