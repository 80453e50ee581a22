"""Gaussian-sharded multi-view step: N ranks, rank r owns parameter slice r and renders view r.

Why.  Replicated-parameter data parallelism (multiview.py) must all-reduce 59 floats = 236 B per
Gaussian and step; at 3 M Gaussians that is 708 MB against a 2.3 ms step (measured: 1.3 ms at 2
GPUs, 63 % scaling).  Nothing downstream of `preprocess` reads raw parameters until
`preprocess_backward`, and both are per-Gaussian.  So the Gaussians are sharded for those two
stages and the views for everything between them:

    forward   rank r: preprocess(slice r, camera v) for v = 0..N-1   -> geometry rows of slice r per view
              all-to-all of geometry rows (depths, means2D, conic_opacity, rgb, tiles_touched, radii:
              52 B per Gaussian and view)                             -> full-P geometry of view r
              depth order, binning, blend of view r                   (unchanged kernels)
    backward  blend backward of view r                                -> grad_accum[P,12] of view r
              all-to-all of grad_accum slices (48 B per Gaussian)     -> accumulators of slice r, all views
              preprocess_backward(slice r, camera v), summed over v   -> parameter gradients of slice r

Per rank and step 100 B x P x (N-1)/N cross NVLink instead of ~2 x 236 B x P x (N-1)/N, no rank
holds more than its slice of parameters, gradients and optimizer state, and the per-(view, Gaussian)
arithmetic is exactly that of the single-GPU path (same kernels, row offsets only), so images,
radii, key lists and per-view gradients are unchanged; the sum over views runs in view order.

Two transports.  exchange="peer" (default): every rank's geometry workspace lives in symmetric
memory (torch.distributed._symmetric_memory: peer-mapped over NVLink); the preprocess kernel of rank
r writes view v's rows STRAIGHT INTO rank v's workspace (remote stores, no staging, no copy kernel)
and preprocess_backward reads view v's accumulators straight from rank v's workspace (remote loads),
so each exchange is fused into the kernel that produces / consumes the data; two device-side
barriers per step order the phases (the workspace is double-buffered across steps, which removes the
third: "peers are done reading the previous step").  exchange="nccl": staged rows + one batched group of NCCL
point-to-point operations per exchange (measured 0.43 + 0.27 ms at 2 GPUs against 0.17 ms of peer
copies, hence not the default).

`ShardPlan` and `all_to_all_rows` are device-agnostic (exercised on CPU with gloo in
tests/test_sharded_gloo.py); `ShardedViewRasterizer` drives the C ABI and needs CUDA.
"""
from __future__ import annotations

import ctypes
from typing import Dict, List, Optional, Sequence

import torch
import torch.distributed as dist

from . import _native
from .rasterizer import (GaussianRasterizationSettings, RasterContext, _dense, _pinned_info, _ptr, _render_stage,
                         _settings_tensors)

__all__ = ["ShardPlan", "all_to_all_rows", "ShardedViewRasterizer", "GEOMETRY_ARRAYS"]

SLICE_ALIGN = 128  # rows; keeps every per-Gaussian array of a slice 16-byte aligned (needs a multiple of 4)

# per-Gaussian arrays of the geometry workspace that travel in the forward exchange
GEOMETRY_ARRAYS = ("depths", "means2D", "conic_opacity", "rgb", "tiles_touched")


class ShardPlan:
    """Contiguous slices of P Gaussians over `world` ranks; every slice but the last is a multiple of
    SLICE_ALIGN rows, so global index order is slice order (the depth sort's tie-break stays by index)."""

    def __init__(self, P: int, world: int):
        if P < 0 or world < 1:
            raise ValueError("P >= 0 and world >= 1 required")
        per = -(-P // world)
        per = -(-per // SLICE_ALIGN) * SLICE_ALIGN
        self.P, self.world = P, world
        self.starts = [min(P, r * per) for r in range(world)]
        self.sizes = [min(P, (r + 1) * per) - self.starts[r] for r in range(world)]

    def rows(self, rank: int) -> slice:
        return slice(self.starts[rank], self.starts[rank] + self.sizes[rank])

    def slot_stride(self, rank: int) -> int:
        """Row stride between the per-view slots of rank's send workspace."""
        return max(SLICE_ALIGN, -(-self.sizes[rank] // SLICE_ALIGN) * SLICE_ALIGN)


def all_to_all_rows(pairs: Sequence, group=None) -> None:
    """pairs = [(send_views, recv_views), ...]; for each pair send_views[j] goes to rank j and
    recv_views[j] is filled from rank j (contiguous tensors, sizes agreed by construction).  All
    arrays travel in ONE batched group of point-to-point operations (one NCCL launch); the local
    chunk is a device copy.  Empty chunks are skipped on both sides."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    ops = []
    for send, recv in pairs:
        if len(send) != world or len(recv) != world:
            raise ValueError("one send and one receive view per rank required")
        recv[rank].copy_(send[rank])
        for j in range(world):
            if j == rank:
                continue
            peer = dist.get_global_rank(group, j) if group is not None else j
            if send[j].numel():
                ops.append(dist.P2POp(dist.isend, send[j], peer, group))
            if recv[j].numel():
                ops.append(dist.P2POp(dist.irecv, recv[j], peer, group))
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()


def _ws_arrays(ws: torch.Tensor, L: _native.Layout, P: int) -> Dict[str, torch.Tensor]:
    """Typed [P, k] views of the per-Gaussian arrays of a geometry workspace."""
    def v(off, dt, k):
        n = P * k * 4
        return ws[off:off + n].view(dt).view(P, k)
    return dict(depths=v(L.g_depths, torch.float32, 1), means2D=v(L.g_means2D, torch.float32, 2),
                conic_opacity=v(L.g_conic_opacity, torch.float32, 4), rgb=v(L.g_rgb, torch.float32, 4),
                tiles_touched=v(L.g_tiles_touched, torch.int32, 1), grad_accum=v(L.g_grad_accum, torch.float32, 12))




class ShardedViewRasterizer:
    """One instance per rank.  `forward` takes THIS rank's parameter slice and the N cameras of the
    step and returns the image and radii of view `rank`; `backward` takes dL/dpixel of that image and
    returns the parameter gradients of the slice summed over all N views, plus the per-view
    dL/dmeans2D rows of the slice ([N, Ps, 3], for the densification statistics of
    scene/gaussian_model.py:775-777).  SH + scale/rotation inputs (the training configuration)."""

    def __init__(self, P: int, M: int, width: int, height: int, device, group=None, exchange: str = "peer"):
        if not (dist.is_available() and dist.is_initialized()):
            raise RuntimeError("torch.distributed must be initialised (one process per GPU)")
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("the B200 rasterizer has no CPU path: tensors must be on a CUDA device")
        if exchange not in ("peer", "nccl"):
            raise ValueError("exchange must be 'peer' or 'nccl'")
        self.group, self.exchange = group, exchange
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.P, self.M, self.W, self.H = P, M, width, height
        self.plan = ShardPlan(P, self.world)
        self.Ps = self.plan.sizes[self.rank]
        self.start = self.plan.starts[self.rank]
        self.lib = _native.lib()
        dev = self.device
        self.L_full = _native.workspace_layout(P, width, height, 0)
        self.geom_bytes, _, self.image_bytes = _native.workspace_bytes(P, width, height, 0)
        if exchange == "peer":
            import torch.distributed._symmetric_memory as symm_mem
            # TWO workspaces, used on alternate steps: the rows a peer stores into my workspace for step k+1 must
            # not land while any rank still reads step k's.  With one buffer that takes a barrier at the top of
            # every step; with two, the buffer written in step k+1 was last read in step k-1, and every rank has
            # passed step k's barriers — which follow its step k-1 reads in stream order — before anyone stores.
            self.radii_off = (self.geom_bytes + 255) // 256 * 256
            self.half = (self.radii_off + 4 * max(P, 1) + 255) // 256 * 256
            self.symm = symm_mem.empty(2 * self.half, dtype=torch.uint8, device=dev)
            self.hdl = symm_mem.rendezvous(self.symm, group if group is not None else dist.group.WORLD)
            self.peer_base = [int(x) for x in self.hdl.buffer_ptrs]
            self._bufs = []
            for b in (0, self.half):
                ws = self.symm[b:b + self.geom_bytes]
                self._bufs.append((ws, self.symm[b + self.radii_off:b + self.radii_off + 4 * P].view(torch.int32),
                                   _ws_arrays(ws, self.L_full, P)))
            self.parity = 1
            self.geom_ws, self.radii_ws, self.full = self._bufs[self.parity]
        else:
            self.stride = self.plan.slot_stride(self.rank)
            self.P_send = self.world * self.stride
            gb_send, _, _ = _native.workspace_bytes(self.P_send, width, height, 0)
            self.send_ws = torch.empty(gb_send, dtype=torch.uint8, device=dev)
            self.send_radii = torch.zeros(self.P_send, dtype=torch.int32, device=dev)
            self.send = _ws_arrays(self.send_ws, _native.workspace_layout(self.P_send, width, height, 0), self.P_send)
            self.geom_ws = torch.empty(self.geom_bytes, dtype=torch.uint8, device=dev)
            self.radii_ws = torch.empty(P, dtype=torch.int32, device=dev)
            self.full = _ws_arrays(self.geom_ws, self.L_full, P)
        # parameter gradients of the slice (summed over views) and per-view dL/dmeans2D
        f = lambda *s: torch.empty(s, dtype=torch.float32, device=dev)
        Ps = max(self.Ps, 1)
        self.grads = dict(means3D=f(Ps, 3), shs=f(Ps, M, 3), opacities=f(Ps, 1), scales=f(Ps, 3), rotations=f(Ps, 4))
        self.grad_means2D = f(self.world, Ps, 3)
        self._state = None

    # ------------------------------------------------------------------------------------------
    def _slot(self, v: int) -> slice:
        return slice(v * self.stride, v * self.stride + self.Ps)

    def _view_order(self):
        """Views starting at the rank's own, so that at any moment the ranks address different peers."""
        return [(self.rank + k) % self.world for k in range(self.world)]

    def _view_slots(self, chunk, cams, cameras, scal):
        """gsr_view_slot array for the views of `chunk` (+ the workspace row count and first row, which
        are the same for every view of a transport)."""
        arr = (_native.ViewSlot * len(chunk))()
        P_ws = row0 = 0
        for n, v in enumerate(chunk):
            view, proj, campos, _ = cams[v]
            ws_ptr, P_ws, row0, radii_ptr = self._target(v)
            tfx, tfy = (float(cameras[v].tanfovx), float(cameras[v].tanfovy)) if cameras is not None else scal[v][1:]
            arr[n] = _native.ViewSlot(view.data_ptr(), proj.data_ptr(), campos.data_ptr(), tfx, tfy, ws_ptr.value,
                                      radii_ptr.value, self.grad_means2D[v].data_ptr())
        return arr, P_ws, row0

    def _target(self, v: int):
        """(geom_ws pointer, rows in it, first row, radii pointer) a slot call for view v addresses."""
        if self.exchange == "peer":
            base = self.peer_base[v] + self.parity * self.half
            return (ctypes.c_void_p(base), self.P, self.start,
                    ctypes.c_void_p(base + self.radii_off + 4 * self.start))
        return (_ptr(self.send_ws), self.P_send, v * self.stride,
                ctypes.c_void_p(self.send_radii.data_ptr() + 4 * v * self.stride))

    def forward(self, params: Dict[str, torch.Tensor], cameras: Sequence[GaussianRasterizationSettings]):
        """params: means3D [Ps,3], shs [Ps,M,3], opacities [Ps,1], scales [Ps,3], rotations [Ps,4] of
        slice `rank` (ShardPlan.rows).  cameras[v]: settings of view v (same image size, sh_degree,
        scale_modifier on every rank).  Returns (color [3,H,W], radii [P] int32 of view `rank`)."""
        N, r, dev, L = self.world, self.rank, self.device, self.lib
        if len(cameras) != N:
            raise ValueError("one camera per rank required")
        P, Ps, W, H, M = self.P, self.Ps, self.W, self.H, self.M
        rs = cameras[r]
        if int(rs.image_width) != W or int(rs.image_height) != H:
            raise ValueError("camera image size does not match this rasterizer")
        p = {k: _dense(params[k], dev) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
        if Ps and (p["means3D"] is None or p["means3D"].shape[0] != Ps or p["shs"].shape[1] != M):
            raise ValueError("parameter slice does not match ShardPlan.rows(rank)")
        debug = 1 if rs.debug else 0
        D = int(rs.sh_degree)
        cams = [_settings_tensors(c, dev) for c in cameras]   # (view, proj, campos, bg) per view
        dev_idx = dev.index if dev.index is not None else torch.cuda.current_device()
        ctx = RasterContext()
        ctx.P, ctx.D, ctx.M, ctx.W, ctx.H, ctx.reruns = P, D, M, W, H, 0
        peer = self.exchange == "peer"
        with torch.cuda.device(dev):
            stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
            ctx.geom_ws = self.geom_ws
            ctx.image_ws = torch.empty(self.image_bytes, dtype=torch.uint8, device=dev)
            color = torch.empty((3, H, W), dtype=torch.float32, device=dev)
            if peer:   # no barrier here: this step's rows go to the workspace last read two steps ago (see __init__)
                self.parity ^= 1
                self.geom_ws, self.radii_ws, self.full = self._bufs[self.parity]
                ctx.geom_ws = self.geom_ws
            order = self._view_order()
            step = _native.MAX_VIEWS if peer else 1   # staged (nccl) slots sit at different rows: one view per call
            for c0 in range(0, N, step):
                chunk = order[c0:c0 + step]
                slots, P_ws, row0 = self._view_slots(chunk, cams, cameras, None)
                _native.check(L.gsr_preprocess_views(
                    P_ws, row0, Ps, D, M, W, H, _ptr(p["means3D"]), _ptr(p["shs"]), None, _ptr(p["opacities"]),
                    _ptr(p["scales"]), float(rs.scale_modifier), _ptr(p["rotations"]), None, len(chunk), slots,
                    stream, debug), "gsr_preprocess_views")
            if peer:
                self.hdl.barrier(channel=1)   # all slices of this rank's view have landed
            else:
                pairs = [([self.send[name][self._slot(v)] for v in range(N)],
                          [self.full[name][self.plan.rows(j)] for j in range(N)]) for name in GEOMETRY_ARRAYS]
                pairs.append(([self.send_radii[self._slot(v)] for v in range(N)],
                              [self.radii_ws[self.plan.rows(j)] for j in range(N)]))
                all_to_all_rows(pairs, self.group)
            info = _pinned_info(dev_idx)
            _native.check(L.gsr_depth_order(P, W, H, _ptr(ctx.geom_ws), ctypes.c_void_p(info.data_ptr()), stream,
                                            debug), "gsr_depth_order")
            _render_stage(L, ctx, P, W, H, cams[r][3], self.radii_ws, color, info, None, dev, dev_idx, stream, debug)
            radii = self.radii_ws.clone()     # the workspace copy is overwritten by the next step
        self._state = (ctx, p, cams, [(float(c.scale_modifier), float(c.tanfovx), float(c.tanfovy)) for c in cameras],
                       D, debug)
        return color, radii

    def backward(self, dL_dpix: torch.Tensor):
        """Returns (grads, grad_means2D): grads[name] = dL/d(parameter rows of this slice) summed over
        the N views (buffers owned by this object, overwritten by the next call); grad_means2D [N,Ps,3]."""
        if self._state is None:
            raise RuntimeError("backward() needs a preceding forward()")
        ctx, p, cams, scal, D, debug = self._state
        N, r, dev, L = self.world, self.rank, self.device, self.lib
        P, Ps, W, H, M = self.P, self.Ps, self.W, self.H, self.M
        g = _dense(dL_dpix, dev)
        if g is None or tuple(g.shape) != (3, H, W):
            raise RuntimeError("dL_dpix must be [3,H,W]")
        with torch.cuda.device(dev):
            stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
            _native.check(L.gsr_blend_backward(P, W, H, _ptr(cams[r][3]), _ptr(ctx.geom_ws), _ptr(ctx.binning_ws),
                                               ctx.capacity, _ptr(ctx.image_ws), _ptr(g), stream, debug),
                          "gsr_blend_backward")
            if self.exchange == "peer":
                self.hdl.barrier(channel=2)   # every view's accumulators are complete
            else:
                all_to_all_rows([([self.full["grad_accum"][self.plan.rows(j)] for j in range(N)],
                                  [self.send["grad_accum"][self._slot(v)] for v in range(N)])], self.group)
            G = self.grads
            order = self._view_order()
            step = _native.MAX_VIEWS if self.exchange == "peer" else 1
            for c0 in range(0, N, step):
                chunk = order[c0:c0 + step]
                slots, P_ws, row0 = self._view_slots(chunk, cams, None, scal)
                _native.check(L.gsr_preprocess_backward_views(
                    P_ws, row0, Ps, D, M, W, H, _ptr(p["means3D"]), _ptr(p["shs"]), None, _ptr(p["scales"]),
                    scal[r][0], _ptr(p["rotations"]), None, len(chunk), slots, None, _ptr(G["opacities"]),
                    _ptr(G["means3D"]), None, _ptr(G["shs"]), _ptr(G["scales"]), _ptr(G["rotations"]),
                    1 if c0 > 0 else 0, stream, debug), "gsr_preprocess_backward_views")
        return G, self.grad_means2D
