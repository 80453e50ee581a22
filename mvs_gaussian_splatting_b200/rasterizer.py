"""Host-side mirror of the reference's rasterizer interface, on top of the C ABI.

The reference imports `GaussianRasterizationSettings` and `GaussianRasterizer` from the
third-party module `diff_gaussian_rasterization` (/root/reference/gaussian_renderer/__init__.py:15),
builds the settings with 12 keyword fields (:42-55), constructs the module (:57) and calls it with
`means3D, means2D, shs, colors_precomp, opacities, scales, rotations, cov3D_precomp` (:257-265),
unpacking `(rendered_image, radii)`.  This file keeps those names, argument meanings, return
values, gradient set and error behaviour; the work underneath is libgsrast_b200.so
(include/gsrast_b200.h) — hand-written sm_100a CUDA, no CPU / PyTorch fallback.

PyTorch is used for device memory, the current stream and autograd plumbing only.
"""
from __future__ import annotations

import ctypes
import threading
from typing import NamedTuple, Optional

import torch
from torch import nn

from . import _native

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians",
           "RasterContext", "rasterize_forward", "rasterize_backward", "mark_visible",
           "set_capacity_hint", "last_forward_stats", "GradientArena", "blend_pair_counts",
           "activate", "rasterize_forward_raw", "rasterize_backward_raw"]


class GaussianRasterizationSettings(NamedTuple):
    """Same 12 fields, same order, as the settings tuple the reference constructs at
    gaussian_renderer/__init__.py:42-55."""
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


# --------------------------------------------------------------------------------------------
# small helpers
# --------------------------------------------------------------------------------------------
def _dense(t: Optional[torch.Tensor], device) -> Optional[torch.Tensor]:
    """float32, contiguous, on `device`, 16-byte aligned (the ABI's pointer contract).
    Inputs at the call site can be freshly concatenated non-leaf tensors or views
    (gaussian_renderer/__init__.py:112-119), so this is enforced, not assumed."""
    if t is None or t.numel() == 0:
        return None
    if t.dtype != torch.float32 or t.device != device:
        t = t.to(device=device, dtype=torch.float32)
    t = t.contiguous()
    if t.data_ptr() % 16:
        t = t.clone(memory_format=torch.contiguous_format)
    return t


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


_tls = threading.local()
# Instance capacity to speculate with, per WORKLOAD: (device index, P, W, H) -> capacity.  A loop that alternates
# resolutions or models (the reference evaluates its test cameras every few thousand iterations, train.py:210-222)
# keeps one hint per shape instead of re-running binning + blend on every upward switch.  (device index,) alone is
# the device-wide value set_capacity_hint() installs; a workload's own entry wins over it.
_capacity_hint: dict = {}
_HINT_ENTRIES = 64


def _hint_lookup(dev_idx: int, P: int, W: int, H: int):
    h = _capacity_hint.get((dev_idx, P, W, H))
    return _capacity_hint.get((dev_idx,)) if h is None else h


def _hint_store(dev_idx: int, P: int, W: int, H: int, capacity: int) -> None:
    key = (dev_idx, P, W, H)
    if key not in _capacity_hint and len(_capacity_hint) >= _HINT_ENTRIES:
        _capacity_hint.pop(next(iter(_capacity_hint)))
    _capacity_hint[key] = int(capacity)
_last_stats: dict = {}


def _pinned_info(device_index: int) -> torch.Tensor:
    """One pinned 2x4-int32 block per (thread, device): row 0 = after geometry, row 1 = after render."""
    cache = getattr(_tls, "info", None)
    if cache is None:
        cache = _tls.info = {}
    buf = cache.get(device_index)
    if buf is None:
        buf = cache[device_index] = torch.zeros((2, 4), dtype=torch.int32).pin_memory()
    return buf


def set_capacity_hint(device, capacity: Optional[int]) -> None:
    """Instance capacity the next forward on `device` speculates with (None: forget; the next
    call then waits for the exact count before binning, which is what the reference always does)."""
    idx = torch.device(device).index or 0
    for key in [k for k in _capacity_hint if k[0] == idx]:   # the per-workload entries of this device too
        del _capacity_hint[key]
    if capacity is not None:
        _capacity_hint[(idx,)] = int(capacity)


def last_forward_stats(device=None) -> dict:
    """num_rendered / num_visible / capacity / reruns of the most recent forward (for reports)."""
    idx = 0 if device is None else (torch.device(device).index or 0)
    return dict(_last_stats.get(idx, {}))


class RasterContext:
    """Everything a backward call needs from its forward: the reference keeps the same things
    alive as three byte tensors + num_rendered in its autograd ctx."""

    __slots__ = ("P", "D", "M", "W", "H", "capacity", "num_rendered", "num_visible", "geom_ws",
                 "binning_ws", "image_ws", "radii", "reruns")

    def layout(self) -> _native.Layout:
        return _native.workspace_layout(self.P, self.W, self.H, self.capacity)

    def array(self, name: str) -> torch.Tensor:
        """Typed view of one workspace array (tests / tools only).  "keys" / "keys_unsorted" are the reference's
        64-bit (tile << 32 | depth bits) lists: the render path stores tile ids and indices only, so they are
        rebuilt on request by gsr_materialize_keys into a new tensor ([capacity] int64, first num_rendered valid)."""
        L = self.layout()
        P, C, W, H = max(self.P, 1), max(self.capacity, 1), self.W, self.H
        tiles = ((W + 15) // 16) * ((H + 15) // 16)
        if name in ("keys", "keys_unsorted"):
            dev = self.geom_ws.device
            out = torch.zeros(C, dtype=torch.int64, device=dev)
            with torch.cuda.device(dev):
                stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
                _native.check(_native.lib().gsr_materialize_keys(
                    self.P, W, H, _ptr(self.geom_ws), _ptr(self.binning_ws), self.capacity,
                    1 if name == "keys" else 0, _ptr(out), stream), "gsr_materialize_keys")
            return out
        tk = torch.int16 if L.b_key_bytes == 2 else torch.int32
        spec = {
            "depths": (self.geom_ws, L.g_depths, torch.float32, (P,)),
            "means2D": (self.geom_ws, L.g_means2D, torch.float32, (P, 2)),
            "conic_opacity": (self.geom_ws, L.g_conic_opacity, torch.float32, (P, 4)),
            "rgb": (self.geom_ws, L.g_rgb, torch.float32, (P, 4)),
            "tiles_touched": (self.geom_ws, L.g_tiles_touched, torch.int32, (P,)),
            "grad_accum": (self.geom_ws, L.g_grad_accum, torch.float32, (P, 12)),
            "order_keys": (self.geom_ws, L.g_order_keys, torch.int32, (P,)),
            "order": (self.geom_ws, L.g_order, torch.int32, (P,)),
            "tiles_unsorted": (self.binning_ws, L.b_keys_unsorted, tk, (C,)),
            "values_unsorted": (self.binning_ws, L.b_vals_unsorted, torch.int32, (C,)),
            "tiles_sorted": (self.binning_ws, L.b_keys, tk, (C,)),
            "point_list": (self.binning_ws, L.b_point_list, torch.int32, (C,)),
            "ranges": (self.image_ws, L.i_ranges, torch.int32, (tiles, 2)),
            "final_T": (self.image_ws, L.i_final_T, torch.float32, (H, W)),
            "n_contrib": (self.image_ws, L.i_n_contrib, torch.int32, (H, W)),
        }[name]
        ws, off, dt, shape = spec
        n = 1
        for s in shape:
            n *= s
        nbytes = n * torch.empty((), dtype=dt).element_size()
        return ws[off:off + nbytes].view(dt).view(shape)


_settings_cache: dict = {}    # id(settings) -> (settings, device, dense tensors); a training loop reuses its cameras


def _settings_tensors(rs: GaussianRasterizationSettings, device):
    """(view, proj, campos, bg) as dense float32 device tensors.  Cached per settings OBJECT (the entry keeps the
    object alive, so an id cannot be recycled): settings are immutable NamedTuples, and when their tensors are
    already dense the cached entries ARE those tensors — a caller that refreshes a camera block in place is seen."""
    hit = _settings_cache.get(id(rs))
    if hit is not None and hit[0] is rs and hit[1] == device:
        return hit[2]
    out = _settings_tensors_uncached(rs, device)
    if all(o is t for o, t in zip(out, (rs.viewmatrix, rs.projmatrix, rs.campos, rs.bg))):   # no copies were made
        if len(_settings_cache) >= 64:
            _settings_cache.pop(next(iter(_settings_cache)))
        _settings_cache[id(rs)] = (rs, device, out)
    return out


def _settings_tensors_uncached(rs: GaussianRasterizationSettings, device):
    view = _dense(rs.viewmatrix, device)
    proj = _dense(rs.projmatrix, device)
    campos = _dense(rs.campos, device)
    bg = _dense(rs.bg, device)
    if view is None or view.numel() != 16 or proj is None or proj.numel() != 16:
        raise ValueError("viewmatrix and projmatrix must be 4x4")
    if campos is None or campos.numel() != 3 or bg is None or bg.numel() != 3:
        raise ValueError("campos and bg must have 3 elements")
    return view, proj, campos, bg


def _render_stage(L, ctx, P, W, H, bg, radii, color, info, capacity, device, dev_idx, stream, debug):
    """Binning + blend for a geometry workspace whose depth order is queued on `stream` and whose
    info block is being copied to the pinned `info` row 0.  Speculates on the instance capacity
    (see rasterize_forward) and re-runs only this stage on overflow.  Must run inside
    torch.cuda.device(device)."""
    ev = torch.cuda.Event()
    ev.record(torch.cuda.current_stream(device))

    hint = _hint_lookup(dev_idx, P, W, H)
    speculative = capacity is None and hint is not None
    if capacity is None:
        if speculative:
            capacity = hint
        else:
            ev.synchronize()           # first call of this workload on this device: size exactly, like the reference
            capacity = int(info[0, 0])
    capacity = max(int(capacity), 1)

    def render(cap):
        _, bb, _ = _native.workspace_bytes(P, W, H, cap)
        ctx.binning_ws = torch.empty(bb, dtype=torch.uint8, device=device)
        ctx.capacity = cap
        _native.check(L.gsr_forward_render(
            P, W, H, _ptr(bg), _ptr(radii), _ptr(ctx.geom_ws), _ptr(ctx.binning_ws), cap,
            _ptr(ctx.image_ws), _ptr(color), None, stream, debug), "gsr_forward_render")

    render(capacity)
    ev.synchronize()                   # waits for the geometry stage only; blend keeps running
    R, nvis = int(info[0, 0]), int(info[0, 2])
    if R > capacity:
        ctx.reruns = 1
        render(R + R // 8 + 1024)
    ctx.num_rendered, ctx.num_visible, ctx.radii = R, nvis, radii
    # speculate next time with 25% head-room over the largest recent count
    prev = _capacity_hint.get((dev_idx, P, W, H), 0)
    _hint_store(dev_idx, P, W, H, max(int(prev * 0.97), R + R // 4 + 4096))
    _last_stats[dev_idx] = dict(num_rendered=R, num_visible=nvis, capacity=ctx.capacity,
                                reruns=ctx.reruns, speculative=bool(speculative))


def rasterize_forward(means3D, shs, colors_precomp, opacities, scales, rotations, cov3D_precomp,
                      raster_settings: GaussianRasterizationSettings, capacity: Optional[int] = None):
    """Forward through the C ABI.  Returns (color[3,H,W], radii[P] int32, RasterContext, inputs) —
    `inputs` is the tuple of dense tensors `rasterize_backward` takes as `ctx_inputs`.

    `capacity`: instance capacity of the binning workspace.  None = use the per-device hint from
    earlier calls and verify it against the device's count WITHOUT stalling the GPU (the count is
    copied to pinned memory right after the prefix sum while the binning/blend launches are
    already queued); on overflow only the binning+blend stage is re-run.  The reference instead
    blocks the host in the middle of every forward."""
    rs = raster_settings
    device = means3D.device
    if device.type != "cuda":
        raise RuntimeError("the B200 rasterizer has no CPU path: tensors must be on a CUDA device")
    L = _native.lib()
    if means3D.dim() != 2 or means3D.shape[1] != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    P = int(means3D.shape[0])
    H, W = int(rs.image_height), int(rs.image_width)
    debug = 1 if rs.debug else 0
    dev_idx = device.index if device.index is not None else torch.cuda.current_device()

    means3D = _dense(means3D, device)
    shs, colors_precomp = _dense(shs, device), _dense(colors_precomp, device)
    opacities = _dense(opacities, device)
    scales, rotations = _dense(scales, device), _dense(rotations, device)
    cov3D_precomp = _dense(cov3D_precomp, device)
    view, proj, campos, bg = _settings_tensors(rs, device)
    M = 0 if shs is None else int(shs.shape[1])
    D = int(rs.sh_degree)

    ctx = RasterContext()
    ctx.P, ctx.D, ctx.M, ctx.W, ctx.H = P, D, M, W, H
    ctx.reruns = 0
    with torch.cuda.device(device):
        stream = ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)
        gb, _, ib = _native.workspace_bytes(P, W, H, 0)
        ctx.geom_ws = torch.empty(gb, dtype=torch.uint8, device=device)
        ctx.image_ws = torch.empty(ib, dtype=torch.uint8, device=device)
        radii = torch.empty(P, dtype=torch.int32, device=device)
        color = torch.empty((3, H, W), dtype=torch.float32, device=device)
        info = _pinned_info(dev_idx)
        info_p0 = ctypes.c_void_p(info.data_ptr())
        _native.check(L.gsr_forward_geometry(
            P, D, M, W, H, _ptr(means3D), _ptr(shs), _ptr(colors_precomp), _ptr(opacities),
            _ptr(scales), float(rs.scale_modifier), _ptr(rotations), _ptr(cov3D_precomp), _ptr(view),
            _ptr(proj), _ptr(campos), float(rs.tanfovx), float(rs.tanfovy), 1 if rs.prefiltered else 0,
            _ptr(radii), _ptr(ctx.geom_ws), info_p0, stream, debug), "gsr_forward_geometry")
        _render_stage(L, ctx, P, W, H, bg, radii, color, info, capacity, device, dev_idx, stream, debug)
    ctx_inputs = (means3D, shs, colors_precomp, scales, rotations, cov3D_precomp, view, proj, campos, bg)
    return color, radii, ctx, ctx_inputs


# --------------------------------------------------------------------------------------------
# raw-parameter entry point: activations fused into preprocess / preprocess_backward
# --------------------------------------------------------------------------------------------
def activate(features_dc, features_rest, opacity_raw, scaling_raw, rotation_raw):
    """The reference's getters (scene/gaussian_model.py:152-183) as ONE call of the library's own
    activation kernel: returns (shs [P,M,3], opacities [P,1], scales [P,3], rotations [P,4]).  Feeding
    these to `rasterize_forward` reproduces `rasterize_forward_raw` bit for bit."""
    device = features_dc.device
    if device.type != "cuda":
        raise RuntimeError("the B200 rasterizer has no CPU path: tensors must be on a CUDA device")
    dc, rest = _dense(features_dc, device), _dense(features_rest, device)
    op, sc, rot = _dense(opacity_raw, device), _dense(scaling_raw, device), _dense(rotation_raw, device)
    P = int(features_dc.shape[0])
    M = 1 + (0 if rest is None else int(rest.shape[1]))
    e = lambda *s: torch.empty(s, dtype=torch.float32, device=device)
    shs, opac, scales, rots = e(P, M, 3), e(P, 1), e(P, 3), e(P, 4)
    if P:
        with torch.cuda.device(device):
            stream = ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)
            _native.check(_native.lib().gsr_activate(P, M, _ptr(dc), _ptr(rest), _ptr(op), _ptr(sc), _ptr(rot),
                                                     _ptr(shs), _ptr(opac), _ptr(scales), _ptr(rots), stream),
                          "gsr_activate")
    return shs, opac, scales, rots


def rasterize_forward_raw(means3D, features_dc, features_rest, opacity_raw, scaling_raw, rotation_raw,
                          raster_settings: GaussianRasterizationSettings, capacity: Optional[int] = None):
    """`rasterize_forward` on the reference's PRE-activation parameters (`_xyz, _features_dc [P,1,3],
    _features_rest [P,M-1,3], _opacity, _scaling, _rotation` of scene/gaussian_model.py:56-61): exp /
    sigmoid / normalize and the dc|rest concatenation happen inside the preprocess kernel instead of as
    PyTorch ops around it.  Returns (color, radii, RasterContext, inputs)."""
    rs = raster_settings
    device = means3D.device
    if device.type != "cuda":
        raise RuntimeError("the B200 rasterizer has no CPU path: tensors must be on a CUDA device")
    L = _native.lib()
    if means3D.dim() != 2 or means3D.shape[1] != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    P = int(means3D.shape[0])
    H, W = int(rs.image_height), int(rs.image_width)
    debug = 1 if rs.debug else 0
    dev_idx = device.index if device.index is not None else torch.cuda.current_device()
    means3D = _dense(means3D, device)
    dc, rest = _dense(features_dc, device), _dense(features_rest, device)
    op, sc, rot = _dense(opacity_raw, device), _dense(scaling_raw, device), _dense(rotation_raw, device)
    view, proj, campos, bg = _settings_tensors(rs, device)
    M = 1 + (0 if rest is None else int(rest.shape[1]))
    D = int(rs.sh_degree)
    ctx = RasterContext()
    ctx.P, ctx.D, ctx.M, ctx.W, ctx.H = P, D, M, W, H
    ctx.reruns = 0
    with torch.cuda.device(device):
        stream = ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)
        gb, _, ib = _native.workspace_bytes(P, W, H, 0)
        ctx.geom_ws = torch.empty(gb, dtype=torch.uint8, device=device)
        ctx.image_ws = torch.empty(ib, dtype=torch.uint8, device=device)
        radii = torch.empty(P, dtype=torch.int32, device=device)
        color = torch.empty((3, H, W), dtype=torch.float32, device=device)
        info = _pinned_info(dev_idx)
        _native.check(L.gsr_forward_geometry_raw(
            P, D, M, W, H, _ptr(means3D), _ptr(dc), _ptr(rest), _ptr(op), _ptr(sc), float(rs.scale_modifier),
            _ptr(rot), _ptr(view), _ptr(proj), _ptr(campos), float(rs.tanfovx), float(rs.tanfovy), _ptr(radii),
            _ptr(ctx.geom_ws), ctypes.c_void_p(info.data_ptr()), stream, debug), "gsr_forward_geometry_raw")
        _render_stage(L, ctx, P, W, H, bg, radii, color, info, capacity, device, dev_idx, stream, debug)
    return color, radii, ctx, (means3D, dc, rest, sc, rot, view, proj, campos, bg)


def rasterize_backward_raw(ctx: RasterContext, ctx_inputs, raster_settings: GaussianRasterizationSettings,
                           grad_out_color: torch.Tensor):
    """Backward of `rasterize_forward_raw`: gradients w.r.t. the raw tensors, in the order
    (means3D, means2D, features_dc, features_rest, opacity_raw, scaling_raw, rotation_raw)."""
    rs = raster_settings
    means3D, dc, rest, sc, rot, view, proj, campos, bg = ctx_inputs
    device = means3D.device
    P, M = ctx.P, ctx.M
    g = _dense(grad_out_color, device)
    if g is None or tuple(g.shape) != (3, ctx.H, ctx.W):
        raise RuntimeError("grad_out_color must be [3,H,W]")
    e = lambda *s: torch.empty(s, dtype=torch.float32, device=device)
    d_means2D, d_means3D, d_op = e(P, 3), e(P, 3), e(P, 1)
    d_dc, d_rest = e(P, 1, 3), (e(P, M - 1, 3) if M > 1 else None)
    d_sc, d_rot = e(P, 3), e(P, 4)
    with torch.cuda.device(device):
        stream = ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)
        _native.check(_native.lib().gsr_backward_raw(
            P, ctx.D, M, ctx.W, ctx.H, _ptr(bg), _ptr(means3D), _ptr(dc), _ptr(rest), _ptr(sc),
            float(rs.scale_modifier), _ptr(rot), _ptr(view), _ptr(proj), _ptr(campos), float(rs.tanfovx),
            float(rs.tanfovy), _ptr(ctx.radii), _ptr(ctx.geom_ws), _ptr(ctx.binning_ws), ctx.capacity,
            _ptr(ctx.image_ws), _ptr(g), _ptr(d_means2D), _ptr(d_op), _ptr(d_means3D), _ptr(d_dc), _ptr(d_rest),
            _ptr(d_sc), _ptr(d_rot), stream, 1 if rs.debug else 0), "gsr_backward_raw")
    return d_means3D, d_means2D, d_dc, d_rest, d_op, d_sc, d_rot


class _RasterizeGaussiansRaw(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, features_dc, features_rest, opacity_raw, scaling_raw, rotation_raw,
                raster_settings):
        color, radii, rctx, inputs = rasterize_forward_raw(means3D, features_dc, features_rest, opacity_raw,
                                                           scaling_raw, rotation_raw, raster_settings)
        ctx.raster_settings, ctx.rctx = raster_settings, rctx
        ctx.save_for_backward(*inputs)   # in-place edits before backward are detected; freed with the graph
        ctx.has_rest = features_rest is not None and features_rest.numel() > 0
        ctx.mark_non_differentiable(radii)
        return color, radii

    @staticmethod
    def backward(ctx, grad_out_color, _grad_radii):
        # the workspaces live as long as the graph does, so backward(retain_graph=True) can run again
        d_means3D, d_means2D, d_dc, d_rest, d_op, d_sc, d_rot = rasterize_backward_raw(
            ctx.rctx, ctx.saved_tensors, ctx.raster_settings, grad_out_color)
        return d_means3D, d_means2D, d_dc, (d_rest if ctx.has_rest else None), d_op, d_sc, d_rot, None


class GradientArena:
    """One flat float32 buffer holding the parameter gradients of a view back to back
    (means3D [P,3] | shs [P,M,3] | opacities [P,1] | scales [P,3] | rotations [P,4], each segment
    16-byte aligned), so that the multi-GPU step sums them with ONE all-reduce and no packing copy:
    the backward kernel writes straight into the segments.  59 floats per Gaussian at M = 16
    (the six Adam groups of scene/gaussian_model.py:245-252)."""

    def __init__(self, P: int, M: int, device):
        sizes = dict(means3D=3 * P, shs=3 * M * P, opacities=P, scales=3 * P, rotations=4 * P)
        off, self.offsets = 0, {}
        for k, n in sizes.items():
            self.offsets[k] = (off, n)
            off += (n + 3) // 4 * 4
        self.flat = torch.zeros(off, dtype=torch.float32, device=device)
        self.P, self.M = P, M
        v = lambda k, *shape: self.flat[self.offsets[k][0]:self.offsets[k][0] + self.offsets[k][1]].view(*shape)
        self.means3D, self.shs = v("means3D", P, 3), v("shs", P, M, 3)
        self.opacities, self.scales, self.rotations = v("opacities", P, 1), v("scales", P, 3), v("rotations", P, 4)


def rasterize_backward(ctx: RasterContext, ctx_inputs, raster_settings: GaussianRasterizationSettings,
                       grad_out_color: torch.Tensor, arena: Optional[GradientArena] = None):
    """Backward through the C ABI.  Returns the 8 gradients in the reference's order
    (means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3D_precomp);
    entries for inputs that were not given are None.  With `arena`, the parameter gradients are
    written into its segments (overwritten, not accumulated)."""
    rs = raster_settings
    means3D, shs, colors_precomp, scales, rotations, cov3D_precomp, view, proj, campos, bg = ctx_inputs
    device = means3D.device
    L = _native.lib()
    P, M = ctx.P, ctx.M
    g = _dense(grad_out_color, device)
    if g is None or tuple(g.shape) != (3, ctx.H, ctx.W):
        raise RuntimeError("grad_out_color must be [3,H,W]")
    e = lambda *s: torch.empty(s, dtype=torch.float32, device=device)
    if arena is not None and (arena.P != P or arena.M != M or shs is None or scales is None):
        raise RuntimeError("gradient arena does not match this call (P, M, SH + scale/rotation inputs)")
    d_means2D = e(P, 3)
    d_opacity = arena.opacities if arena else e(P, 1)
    d_means3D = arena.means3D if arena else e(P, 3)
    d_colors = e(P, 3) if colors_precomp is not None else None
    d_cov3D = e(P, 6) if cov3D_precomp is not None else None
    d_sh = (arena.shs if arena else e(P, M, 3)) if shs is not None else None
    d_scales = (arena.scales if arena else e(P, 3)) if scales is not None else None
    d_rots = (arena.rotations if arena else e(P, 4)) if rotations is not None else None
    with torch.cuda.device(device):
        stream = ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)
        _native.check(L.gsr_backward(
            P, ctx.D, M, ctx.W, ctx.H, _ptr(bg), _ptr(means3D), _ptr(shs), _ptr(colors_precomp),
            _ptr(scales), float(rs.scale_modifier), _ptr(rotations), _ptr(cov3D_precomp), _ptr(view),
            _ptr(proj), _ptr(campos), float(rs.tanfovx), float(rs.tanfovy), _ptr(ctx.radii),
            _ptr(ctx.geom_ws), _ptr(ctx.binning_ws), ctx.capacity, _ptr(ctx.image_ws), _ptr(g),
            _ptr(d_means2D), _ptr(d_colors), _ptr(d_opacity), _ptr(d_means3D), _ptr(d_cov3D), _ptr(d_sh),
            _ptr(d_scales), _ptr(d_rots), stream, 1 if rs.debug else 0), "gsr_backward")
    return d_means3D, d_means2D, d_sh, d_colors, d_opacity, d_scales, d_rots, d_cov3D


def blend_pair_counts(ctx: RasterContext):
    """(evaluated, blended) (pixel, splat) pair counts of the reference algorithm for a finished
    forward, then the pairs this library evaluates after culling and its (warp, splat) iterations
    (instrumented kernel; benchmarks only).  Synchronises."""
    device = ctx.geom_ws.device
    out = torch.zeros(4, dtype=torch.int64, device=device)
    with torch.cuda.device(device):
        stream = ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)
        _native.check(_native.lib().gsr_blend_pair_counts(
            ctx.P, ctx.W, ctx.H, _ptr(ctx.geom_ws), _ptr(ctx.binning_ws), ctx.capacity, _ptr(ctx.image_ws),
            _ptr(out), stream), "gsr_blend_pair_counts")
    return tuple(int(v) for v in out.cpu().tolist())


def mark_visible(positions: torch.Tensor, raster_settings: GaussianRasterizationSettings) -> torch.Tensor:
    device = positions.device
    if device.type != "cuda":
        raise RuntimeError("the B200 rasterizer has no CPU path: tensors must be on a CUDA device")
    pos = _dense(positions, device)
    P = int(positions.shape[0])
    present = torch.zeros(P, dtype=torch.bool, device=device)
    if P == 0:
        return present
    view, proj, _, _ = _settings_tensors(raster_settings, device)
    with torch.cuda.device(device):
        stream = ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)
        _native.check(_native.lib().gsr_mark_visible(P, _ptr(pos), _ptr(view), _ptr(proj), _ptr(present), stream),
                      "gsr_mark_visible")
    return present


# --------------------------------------------------------------------------------------------
# autograd bridge + nn.Module, as the reference's dependency exposes them
# --------------------------------------------------------------------------------------------
def _cpu_deep_copy_tuple(args):
    return tuple(a.detach().cpu().clone() if isinstance(a, torch.Tensor) else a for a in args)


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                raster_settings):
        rs = raster_settings
        P = means3D.shape[0] if means3D.dim() == 2 else -1
        if P == 0:
            # the reference's glue allocates a zero image and skips the core call when P == 0
            ctx.empty = True
            ctx.shapes = [t.shape for t in (means3D, means2D, sh, colors_precomp, opacities, scales,
                                            rotations, cov3Ds_precomp)]
            color = torch.zeros((3, int(rs.image_height), int(rs.image_width)), dtype=torch.float32,
                                device=means3D.device)
            radii = torch.zeros(0, dtype=torch.int32, device=means3D.device)
            ctx.mark_non_differentiable(radii)
            return color, radii
        ctx.empty = False
        args = (means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, rs)
        if rs.debug:
            snapshot = _cpu_deep_copy_tuple(args)
            try:
                color, radii, rctx, inputs = rasterize_forward(*args)
            except Exception:
                torch.save(snapshot, "snapshot_fw.dump")
                print("\nAn error occured in forward. Please forward snapshot_fw.dump for debugging.")
                raise
        else:
            color, radii, rctx, inputs = rasterize_forward(*args)
        ctx.raster_settings = rs
        ctx.rctx = rctx
        ctx.save_for_backward(*inputs)   # in-place edits before backward are detected; freed with the graph
        ctx.mark_non_differentiable(radii)
        return color, radii

    @staticmethod
    def backward(ctx, grad_out_color, _grad_radii):
        if ctx.empty:
            return (None,) * 9
        rs = ctx.raster_settings
        inputs = ctx.saved_tensors   # raises autograd's own error after the graph was freed
        # The workspaces (ctx.rctx) live exactly as long as the graph: a second backward under
        # retain_graph=True recomputes from them, as the reference's saved byte tensors allow.
        if rs.debug:
            try:
                grads = rasterize_backward(ctx.rctx, inputs, rs, grad_out_color)
            except Exception:
                torch.save(_cpu_deep_copy_tuple(tuple(inputs) + (grad_out_color,)), "snapshot_bw.dump")
                print("\nAn error occured in backward. Writing snapshot_bw.dump for debugging.\n")
                raise
        else:
            grads = rasterize_backward(ctx.rctx, inputs, rs, grad_out_color)
        return grads + (None,)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                     cov3Ds_precomp, raster_settings)


class GaussianRasterizer(nn.Module):
    """Drop-in for the module constructed at gaussian_renderer/__init__.py:57 and called at :257-265."""

    def __init__(self, raster_settings: GaussianRasterizationSettings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions: torch.Tensor) -> torch.Tensor:
        with torch.no_grad():
            return mark_visible(positions, self.raster_settings)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None,
                rotations=None, cov3D_precomp=None):
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception("Please provide excatly one of either SHs or precomputed colors!")
        if ((scales is None or rotations is None) and cov3D_precomp is None) or (
                (scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception("Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!")
        empty = lambda: torch.empty(0, dtype=torch.float32, device=means3D.device)
        shs = empty() if shs is None else shs
        colors_precomp = empty() if colors_precomp is None else colors_precomp
        scales = empty() if scales is None else scales
        rotations = empty() if rotations is None else rotations
        cov3D_precomp = empty() if cov3D_precomp is None else cov3D_precomp
        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations,
                                   cov3D_precomp, self.raster_settings)

    def forward_raw(self, means3D, means2D, features_dc, features_rest, opacity_raw, scaling_raw, rotation_raw):
        """Same result as `forward(means3D, means2D, sigmoid(opacity_raw), shs=cat(features_dc,
        features_rest), scales=exp(scaling_raw), rotations=normalize(rotation_raw))` — what render() builds
        from the model's getters (gaussian_renderer/__init__.py:59-86) — with the activations and their
        backward fused into the rasterizer's kernels; differentiable w.r.t. the raw tensors."""
        if means3D.shape[0] == 0:
            raise RuntimeError("forward_raw needs at least one Gaussian")
        return _RasterizeGaussiansRaw.apply(means3D, means2D, features_dc, features_rest, opacity_raw, scaling_raw,
                                            rotation_raw, self.raster_settings)
