"""Multi-view data parallelism for the rasterizer path (BASELINE.json north_star).

The reference is single-GPU (utils/general_utils.py:133 pins cuda:0; one camera per step,
train.py:81-91).  Here a batch of views is split across ranks — one process per GPU, every rank
holds all Gaussians — each rank runs fwd+bwd for its own cameras, and the parameter gradients
(59 floats per Gaussian: the six Adam groups of scene/gaussian_model.py:245-252) are summed with
ONE all-reduce over NCCL/NVLink.  The backward kernel writes straight into the flat
`GradientArena`, so there is no packing copy before the collective.  Densification statistics the
training loop keeps per Gaussian (train.py:130-131: max radii, visibility counts) reduce alongside.

`render_view` is injected so the same host logic is exercised on CPU (gloo, world_size 2) in
tests/test_multiview_gloo.py; in production it is `cuda_render_view` below.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Sequence

import torch
import torch.distributed as dist

from .rasterizer import GradientArena, rasterize_backward, rasterize_forward

__all__ = ["partition_views", "multiview_step", "cuda_render_view", "ViewStats"]


def partition_views(n_views: int, rank: int, world_size: int) -> List[int]:
    """Round-robin camera subset of `rank`: views rank, rank+world, ... (every view exactly once)."""
    if not (0 <= rank < world_size):
        raise ValueError("rank must be in [0, world_size)")
    return list(range(rank, n_views, world_size))


class ViewStats:
    """Per-Gaussian statistics a training loop derives from each view (train.py:127-131):
    element-wise max of radii over the batch and the number of views each Gaussian was visible in."""

    def __init__(self, P: int, device):
        self.max_radii = torch.zeros(P, dtype=torch.int32, device=device)
        self.visible_count = torch.zeros(P, dtype=torch.int32, device=device)

    def update(self, radii: torch.Tensor) -> None:
        torch.maximum(self.max_radii, radii, out=self.max_radii)
        self.visible_count += (radii > 0).to(torch.int32)


def cuda_render_view(params: dict, settings, dL_dpix: torch.Tensor, arena: GradientArena):
    """fwd+bwd of one view through the C ABI; gradients land in `arena`. Returns (color, radii)."""
    color, radii, ctx, inputs = rasterize_forward(params["means3D"], params["shs"], None, params["opacities"],
                                                  params["scales"], params["rotations"], None, settings)
    rasterize_backward(ctx, inputs, settings, dL_dpix, arena=arena)
    return color, radii


def multiview_step(view_ids: Sequence[int], render_view: Callable[[int, GradientArena], torch.Tensor],
                   arena: GradientArena, accum: Optional[GradientArena] = None, stats: Optional[ViewStats] = None,
                   group=None) -> GradientArena:
    """Runs `render_view(view_id, arena)` (which must OVERWRITE arena with that view's gradients and
    return its radii) for this rank's views, sums them locally, then all-reduces once.

    With a single local view the arena itself is reduced in place (no accumulation pass); with more,
    `accum` (same shape) receives the local sum.  A rank with no views contributes zeros.  Returns
    the arena holding the batch-summed gradients (identical on every rank)."""
    if len(view_ids) > 1 and accum is None:
        raise ValueError("more than one local view needs an accumulation arena")
    out = arena
    if len(view_ids) == 0:
        arena.flat.zero_()
    for n, v in enumerate(view_ids):
        radii = render_view(v, arena)
        if stats is not None and radii is not None:
            stats.update(radii)
        if len(view_ids) > 1:
            if n == 0:
                accum.flat.copy_(arena.flat)
            else:
                accum.flat.add_(arena.flat)
            out = accum
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(out.flat, op=dist.ReduceOp.SUM, group=group)
        if stats is not None:
            dist.all_reduce(stats.max_radii, op=dist.ReduceOp.MAX, group=group)
            dist.all_reduce(stats.visible_count, op=dist.ReduceOp.SUM, group=group)
    return out
