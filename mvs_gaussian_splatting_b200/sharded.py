"""Gaussian-sharded multi-view step: N ranks, a batch of V views per step (V a multiple of N); rank r owns
parameter slice r and renders views r, r+N, ... (V = N: one view per rank).

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

__all__ = ["ShardPlan", "ViewPlan", "all_to_all_rows", "ShardedViewRasterizer", "GEOMETRY_ARRAYS"]

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


class ViewPlan:
    """Which rank renders which view of a V-view batch ("rank r takes views r, r+n, ..."):
    view v is rendered by rank v % N as its local view v // N."""

    def __init__(self, n_views: int, world: int):
        if world < 1 or n_views < 1 or n_views % world:
            raise ValueError("the number of views per step must be a positive multiple of the world size")
        self.V, self.world, self.L = n_views, world, n_views // world

    def owner(self, v: int) -> int:
        return v % self.world

    def local_index(self, v: int) -> int:
        return v // self.world

    def local_views(self, rank: int) -> List[int]:
        return [rank + l * self.world for l in range(self.L)]

    def order(self, rank: int) -> List[int]:
        """All V views starting at the rank's own, so that at any moment the ranks address different peers."""
        return [(rank + k) % self.V for k in range(self.V)]


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
    """One instance per rank.  A step renders a BATCH of V views (V a multiple of the world size N; BASELINE.json
    configs[3]: 8 views over 2 / 4 / 8 GPUs): rank r owns parameter slice r and renders views r, r+N, r+2N, ...
    ("rank r takes views r, r+n, ...").  `forward` takes THIS rank's parameter slice and the V cameras
    and returns the images and radii of the rank's L = V/N views; `backward` takes dL/dpixel of those images and
    returns the parameter gradients of the slice summed over all V views, plus the per-view dL/dmeans2D rows of
    the slice ([V, Ps, 3]) that `densification_stats` (train.py:130-131, scene/gaussian_model.py:775-777) consumes
    together with the per-view radii of the slice.  SH + scale/rotation inputs (the training configuration).

    With one process and no process group (world 1) the same algorithm runs on one GPU with plain local
    workspaces (exchange "local"): the multi-view kernels read the parameters once for the whole batch — the
    single-GPU point of the scaling curve."""

    def __init__(self, P: int, M: int, width: int, height: int, device, group=None, exchange: str = "peer",
                 n_views: Optional[int] = None, peer_accumulator_ring: Optional[bool] = None,
                 standalone: bool = False):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("the B200 rasterizer has no CPU path: tensors must be on a CUDA device")
        # standalone: this process renders the whole batch by itself (world 1, "local" transport) even though a
        # process group exists — bench.py measures the single-GPU point of the scaling curve inside an N-rank run
        if dist.is_available() and dist.is_initialized() and not standalone:
            self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        else:
            self.world, self.rank = 1, 0
        if self.world == 1:
            exchange = "local"
        if exchange not in ("peer", "nccl", "local"):
            raise ValueError("exchange must be 'peer', 'nccl' or 'local'")
        self.group, self.exchange = group, exchange
        V = self.world if n_views is None else int(n_views)
        self.views = ViewPlan(V, self.world)
        self.L = self.views.L
        self.V = V
        self.P, self.M, self.W, self.H = P, M, width, height
        self.plan = ShardPlan(P, self.world)
        self.Ps = self.plan.sizes[self.rank]
        self.start = self.plan.starts[self.rank]
        self.lib = _native.lib()
        dev, L = self.device, self.L
        self.L_full = _native.workspace_layout(P, width, height, 0)
        self.geom_bytes, _, self.image_bytes = _native.workspace_bytes(P, width, height, 0)
        self.radii_off = (self.geom_bytes + 255) // 256 * 256
        self.half = (self.radii_off + 4 * max(P, 1) + 255) // 256 * 256      # one view: geometry workspace + radii
        if exchange in ("peer", "local"):
            # 2 x L view workspaces, the two sets used on alternate steps: the rows a peer stores into my workspace
            # for step k+1 must not land while any rank still reads step k's.  With one set that takes a barrier at
            # the top of every step; with two, the set written in step k+1 was last read in step k-1, and every rank
            # has passed step k's barriers — which follow its step k-1 reads in stream order — before anyone stores.
            nbuf = 2 if exchange == "peer" else 1
            if exchange == "peer":
                import torch.distributed._symmetric_memory as symm_mem
                self.symm = symm_mem.empty(nbuf * L * self.half, dtype=torch.uint8, device=dev)
                self.hdl = symm_mem.rendezvous(self.symm, group if group is not None else dist.group.WORLD)
                self.peer_base = [int(x) for x in self.hdl.buffer_ptrs]
            else:
                self.symm = torch.empty(nbuf * L * self.half, dtype=torch.uint8, device=dev)
                self.hdl, self.peer_base = None, [self.symm.data_ptr()]
            self._bufs = []
            for b in range(nbuf):
                views = []
                for l in range(L):
                    o = (b * L + l) * self.half
                    ws = self.symm[o:o + self.geom_bytes]
                    views.append((ws, self.symm[o + self.radii_off:o + self.radii_off + 4 * P].view(torch.int32),
                                  _ws_arrays(ws, self.L_full, P)))
                self._bufs.append(views)
            self.nbuf, self.parity = nbuf, nbuf - 1
        else:
            self.stride = self.plan.slot_stride(self.rank)
            self.P_send = V * self.stride
            gb_send, _, _ = _native.workspace_bytes(self.P_send, width, height, 0)
            self.send_ws = torch.empty(gb_send, dtype=torch.uint8, device=dev)
            self.send_radii = torch.zeros(self.P_send, dtype=torch.int32, device=dev)
            self.send = _ws_arrays(self.send_ws, _native.workspace_layout(self.P_send, width, height, 0), self.P_send)
            views = []
            for l in range(L):
                ws = torch.empty(self.geom_bytes, dtype=torch.uint8, device=dev)
                views.append((ws, torch.empty(P, dtype=torch.int32, device=dev), _ws_arrays(ws, self.L_full, P)))
            self._bufs, self.nbuf, self.parity = [views], 1, 0
        # parameter gradients of the slice (summed over views), per-view dL/dmeans2D, and the slice owner's local
        # copies of the forward rows its backward needs again (gsr_view_slot.radii_local / conic_local)
        f = lambda *s: torch.empty(s, dtype=torch.float32, device=dev)
        Ps = max(self.Ps, 1)
        self.grads = dict(means3D=f(Ps, 3), shs=f(Ps, M, 3), opacities=f(Ps, 1), scales=f(Ps, 3), rotations=f(Ps, 4))
        self.grad_means2D = f(V, Ps, 3)
        self.radii_local = torch.zeros((V, Ps), dtype=torch.int32, device=dev)
        self.conic_local = f(V, Ps, 4)
        self._state = None
        self._slot_cache: Dict = {}   # (ids of the settings objects, workspace set) -> camera tensors + gsr_view_slot arrays
        # GSR_BWD_PEER_ACCUMULATORS: accumulator rows fetched by bulk copies one view ahead (2 GPUs: 1.55 -> 1.39 ms).
        # `peer_accumulator_ring` (constructor argument, None = on exactly when the accumulators sit on a peer) lets a
        # test run the ring kernel on local memory; no environment variable selects a kernel.
        ring = (exchange == "peer") if peer_accumulator_ring is None else bool(peer_accumulator_ring)
        self._acc_flag = 2 if ring else 0

    # ------------------------------------------------------------------------------------------
    def local_views(self) -> List[int]:
        """Global indices of the views this rank renders, in the order `forward` returns them."""
        return self.views.local_views(self.rank)

    def _slot(self, v: int) -> slice:
        return slice(v * self.stride, v * self.stride + self.Ps)

    def _view_order(self):
        return self.views.order(self.rank)

    def _view_slots(self, chunk, cams, tanfov):
        """gsr_view_slot array for the views of `chunk` (+ the workspace row count and first row, which
        are the same for every view of a transport)."""
        arr = (_native.ViewSlot * len(chunk))()
        P_ws = row0 = 0
        for n, v in enumerate(chunk):
            view, proj, campos, _ = cams[v]
            ws_ptr, P_ws, row0, radii_ptr = self._target(v)
            arr[n] = _native.ViewSlot(view.data_ptr(), proj.data_ptr(), campos.data_ptr(), tanfov[v][0], tanfov[v][1],
                                      ws_ptr, radii_ptr, self.grad_means2D[v].data_ptr(),
                                      self.radii_local[v].data_ptr(), self.conic_local[v].data_ptr())
        return arr, P_ws, row0

    def _target(self, v: int):
        """(geom_ws address, rows in it, first row, radii address) a slot call for view v addresses: the workspace
        of the rank that renders v (rank v % N, its local view v // N), or this rank's staging slot."""
        if self.exchange in ("peer", "local"):
            base = self.peer_base[self.views.owner(v)] + (self.parity * self.L + self.views.local_index(v)) * self.half
            return base, self.P, self.start, base + self.radii_off + 4 * self.start
        return (self.send_ws.data_ptr(), self.P_send, v * self.stride,
                self.send_radii.data_ptr() + 4 * v * self.stride)

    def _camera_tensors(self, cameras):
        """(view, proj, campos, bg) device tensors per camera."""
        return [_settings_tensors(c, self.device) for c in cameras]   # cached per settings object (rasterizer.py)

    def _step_plan(self, cameras):
        """Camera tensors, tan(fov) pairs and the gsr_view_slot arrays of every launch chunk for this camera LIST
        and the current workspace set.  A training loop hands the same list object (or one of a few) every step
        and every address in a slot is fixed for the life of this object, so the ~10 pointer look-ups per view
        and call are made once per (list, workspace set) instead of twice per step.  The entry keeps the settings
        objects alive and a hit requires the very same objects in the same order, so a recycled id or a list edited
        in place is a miss, never a stale plan."""
        key = (tuple(id(c) for c in cameras), self.parity)   # a fresh LIST of the same settings objects still hits
        hit = self._slot_cache.get(key)
        if hit is not None and len(hit[0]) == len(cameras) and all(a is b for a, b in zip(hit[0], cameras)):
            return hit[1:]     # the same settings objects in the same order (the entry keeps them alive: no recycled ids)
        cams = self._camera_tensors(cameras)
        # like rasterizer._settings_cache: a plan is kept only when its camera tensors ARE the caller's tensors.  Where
        # a camera had to be copied (CPU, strided or non-float32 blocks) the slots point at the copies, and a caller
        # that refreshes its own tensors in place would be ignored by every later hit.
        own = all(t is src for c, ts in zip(cameras, cams)
                  for t, src in zip(ts, (c.viewmatrix, c.projmatrix, c.campos, c.bg)))
        tanfov = [(float(c.tanfovx), float(c.tanfovy)) for c in cameras]
        order = self._view_order()
        step = 1 if self.exchange == "nccl" else _native.MAX_VIEWS   # staged slots sit at different rows: one view per call
        chunks = []
        for c0 in range(0, self.V, step):
            chunk = order[c0:c0 + step]
            chunks.append((len(chunk),) + self._view_slots(chunk, cams, tanfov))
        if own:
            if len(self._slot_cache) >= 16:
                self._slot_cache.pop(next(iter(self._slot_cache)))
            self._slot_cache[key] = (tuple(cameras), cams, tanfov, chunks)
        return cams, tanfov, chunks

    def forward(self, params: Dict[str, torch.Tensor], cameras: Sequence[GaussianRasterizationSettings]):
        """params: means3D [Ps,3], shs [Ps,M,3], opacities [Ps,1], scales [Ps,3], rotations [Ps,4] of
        slice `rank` (ShardPlan.rows).  cameras[v]: settings of view v, v = 0..V-1 (same image size, sh_degree,
        scale_modifier on every rank).  Returns (color [L,3,H,W], radii [L,P] int32) of views `local_views()`."""
        N, r, dev, lib = self.world, self.rank, self.device, self.lib
        V, L = self.V, self.L
        if len(cameras) != V:
            raise ValueError(f"{V} cameras per step required ({L} per rank)")
        P, Ps, W, H, M = self.P, self.Ps, self.W, self.H, self.M
        rs = cameras[r]
        if int(rs.image_width) != W or int(rs.image_height) != H:
            raise ValueError("camera image size does not match this rasterizer")
        p = {k: _dense(params[k], dev) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
        if Ps and (p["means3D"] is None or p["means3D"].shape[0] != Ps or p["shs"].shape[1] != M):
            raise ValueError("parameter slice does not match ShardPlan.rows(rank)")
        debug = 1 if rs.debug else 0
        D = int(rs.sh_degree)
        dev_idx = dev.index if dev.index is not None else torch.cuda.current_device()
        staged = self.exchange == "nccl"
        with torch.cuda.device(dev):
            stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
            # no barrier here: this step's rows go to the workspace set last read two steps ago (see __init__)
            self.parity = (self.parity + 1) % self.nbuf
            bufs = self._bufs[self.parity]
            cams, tanfov, chunks = self._step_plan(cameras)   # (view, proj, campos, bg) per view + the slot arrays
            for n_chunk, slots, P_ws, row0 in chunks:
                _native.check(lib.gsr_preprocess_views(
                    P_ws, row0, Ps, D, M, W, H, _ptr(p["means3D"]), _ptr(p["shs"]), None, _ptr(p["opacities"]),
                    _ptr(p["scales"]), float(rs.scale_modifier), _ptr(p["rotations"]), None, n_chunk, slots,
                    stream, debug), "gsr_preprocess_views")
            if self.exchange == "peer":
                self.hdl.barrier(channel=1)   # all slices of this rank's views have landed
            elif staged:
                pairs = []
                for l in range(L):
                    full, radii_ws = bufs[l][2], bufs[l][1]
                    for name in GEOMETRY_ARRAYS:
                        pairs.append(([self.send[name][self._slot(j + l * N)] for j in range(N)],
                                      [full[name][self.plan.rows(j)] for j in range(N)]))
                    pairs.append(([self.send_radii[self._slot(j + l * N)] for j in range(N)],
                                  [radii_ws[self.plan.rows(j)] for j in range(N)]))
                all_to_all_rows(pairs, self.group)
            info = _pinned_info(dev_idx)
            ctxs, colors = [], torch.empty((L, 3, H, W), dtype=torch.float32, device=dev)
            for l in range(L):
                v = r + l * N
                ctx = RasterContext()
                ctx.P, ctx.D, ctx.M, ctx.W, ctx.H, ctx.reruns = P, D, M, W, H, 0
                ctx.geom_ws = bufs[l][0]
                ctx.image_ws = torch.empty(self.image_bytes, dtype=torch.uint8, device=dev)
                _native.check(lib.gsr_depth_order(P, W, H, _ptr(ctx.geom_ws), ctypes.c_void_p(info.data_ptr()), stream,
                                                  debug), "gsr_depth_order")
                _render_stage(lib, ctx, P, W, H, cams[v][3], bufs[l][1], colors[l], info, None, dev, dev_idx, stream,
                              debug)
                ctxs.append(ctx)
            # the workspace copies are overwritten two steps later: hand out stable radii
            radii = torch.stack([bufs[l][1] for l in range(L)]) if L > 1 else bufs[0][1].clone().unsqueeze(0)
        self._state = (ctxs, p, cams, chunks, float(rs.scale_modifier), D, debug)
        return colors, radii

    def backward(self, dL_dpix: torch.Tensor):
        """dL_dpix: [L,3,H,W] (or [3,H,W] when L == 1).  Returns (grads, grad_means2D): grads[name] =
        dL/d(parameter rows of this slice) summed over the V views (buffers owned by this object, overwritten by
        the next call); grad_means2D [V,Ps,3], one block per view."""
        if self._state is None:
            raise RuntimeError("backward() needs a preceding forward()")
        ctxs, p, cams, chunks, scale_modifier, D, debug = self._state
        N, r, dev, lib = self.world, self.rank, self.device, self.lib
        P, Ps, W, H, M, V, L = self.P, self.Ps, self.W, self.H, self.M, self.V, self.L
        g = _dense(dL_dpix, dev)
        if g is not None and g.dim() == 3 and L == 1:
            g = g.unsqueeze(0)
        if g is None or tuple(g.shape) != (L, 3, H, W):
            raise RuntimeError(f"dL_dpix must be [{L},3,{H},{W}]")
        staged = self.exchange == "nccl"
        with torch.cuda.device(dev):
            stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
            for l, ctx in enumerate(ctxs):
                _native.check(lib.gsr_blend_backward(P, W, H, _ptr(cams[r + l * N][3]), _ptr(ctx.geom_ws),
                                                     _ptr(ctx.binning_ws), ctx.capacity, _ptr(ctx.image_ws), _ptr(g[l]),
                                                     stream, debug), "gsr_blend_backward")
            if self.exchange == "peer":
                self.hdl.barrier(channel=2)   # every view's accumulators are complete
            elif staged:
                bufs = self._bufs[self.parity]
                all_to_all_rows([([bufs[l][2]["grad_accum"][self.plan.rows(j)] for j in range(N)],
                                  [self.send["grad_accum"][self._slot(j + l * N)] for j in range(N)])
                                 for l in range(L)], self.group)
            G = self.grads
            for ci, (n_chunk, slots, P_ws, row0) in enumerate(chunks):
                _native.check(lib.gsr_preprocess_backward_views(
                    P_ws, row0, Ps, D, M, W, H, _ptr(p["means3D"]), _ptr(p["shs"]), None, _ptr(p["scales"]),
                    scale_modifier, _ptr(p["rotations"]), None, n_chunk, slots, None, _ptr(G["opacities"]),
                    _ptr(G["means3D"]), None, _ptr(G["shs"]), _ptr(G["scales"]), _ptr(G["rotations"]),
                    (1 if ci > 0 else 0) | self._acc_flag,   # GSR_BWD_ACCUMULATE | GSR_BWD_PEER_ACCUMULATORS
                    stream, debug), "gsr_preprocess_backward_views")
        return G, self.grad_means2D

    def slice_radii(self) -> torch.Tensor:
        """[V, Ps] int32: the radii of this rank's slice in each view of the last forward (train.py:130 reads
        `radii` of the rendered view; here every view of the batch contributes)."""
        return self.radii_local & _native.LOCAL_RADIUS_MASK

    def densification_stats(self, max_radii2D: torch.Tensor, xyz_gradient_accum: torch.Tensor, denom: torch.Tensor):
        """train.py:130-131 + scene/gaussian_model.py:775-777 for the slice, once per view of the batch in view
        order: max_radii2D = max(., radii_v) and xyz_gradient_accum += |dL/dmeans2D_v[:, :2]|, denom += 1 on the rows
        view v sees.  In place on the slice owner's [Ps] / [Ps,1] / [Ps,1] statistics; call after `backward`."""
        if self.Ps == 0:
            return
        radii = self.slice_radii()
        with torch.cuda.device(self.device):
            stream = ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
            for v in range(self.V):
                _native.check(self.lib.gsr_densification_stats(
                    self.Ps, _ptr(radii[v]), _ptr(self.grad_means2D[v]), _ptr(max_radii2D), _ptr(xyz_gradient_accum),
                    _ptr(denom), stream), "gsr_densification_stats")
