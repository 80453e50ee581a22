"""The step after backward: fused Adam over the Gaussian model's parameter groups and the
per-Gaussian densification statistics, on the library's sm_100a kernels (csrc/gsr_optim.cu).

`FusedAdam` is constructed exactly like the optimizer of the reference
(scene/gaussian_model.py:245-262: `torch.optim.Adam(l, lr=0.0, eps=1e-15)` with one named group per
parameter tensor) and IS a `torch.optim.Optimizer`: `param_groups`, `state[param]["exp_avg" | "exp_avg_sq" |
"step"]`, `zero_grad`, `state_dict` behave as the reference's densification code expects
(scene/gaussian_model.py:360-470 edits those entries in place).  Only `step()` differs: ONE kernel launch
for all groups instead of a chain of elementwise kernels per group.  No CPU path.
"""
from __future__ import annotations

import ctypes
from typing import Optional

import torch

from . import _native

__all__ = ["FusedAdam", "densification_stats"]


def _stream(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8):
        if not 0.0 <= lr or not 0.0 <= eps or not (0.0 <= betas[0] < 1.0) or not (0.0 <= betas[1] < 1.0):
            raise ValueError("invalid Adam hyper-parameter")
        super().__init__(params, dict(lr=lr, betas=tuple(betas), eps=eps))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        L = _native.lib()
        # (device, betas, eps, step) -> list of (param, grad, state, lr); normally ONE bucket = one launch
        buckets = {}
        for group in self.param_groups:
            for p in group["params"]:
                if p.grad is None:
                    continue
                if p.device.type != "cuda":
                    raise RuntimeError("FusedAdam has no CPU path: parameters must be on a CUDA device")
                if p.dtype != torch.float32 or not p.is_contiguous() or p.data_ptr() % 16:
                    raise RuntimeError("FusedAdam needs contiguous, 16-byte aligned float32 parameters")
                if p.grad.is_sparse:
                    raise RuntimeError("FusedAdam does not support sparse gradients")
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = torch.tensor(0.0)
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                st["step"] += 1
                g = p.grad
                if g.dtype != torch.float32 or not g.is_contiguous() or g.data_ptr() % 16:
                    g = g.to(torch.float32).contiguous().clone()
                for k in ("exp_avg", "exp_avg_sq"):   # the reference's densification code may install new tensors
                    if not st[k].is_contiguous() or st[k].data_ptr() % 16:
                        st[k] = st[k].contiguous().clone()
                key = (p.device, tuple(group["betas"]), float(group["eps"]), int(st["step"].item()))
                buckets.setdefault(key, []).append((p, g, st, float(group["lr"])))
        for (device, betas, eps, step), items in buckets.items():
            with torch.cuda.device(device):
                for c0 in range(0, len(items), _native.MAX_ADAM_GROUPS):
                    chunk = items[c0:c0 + _native.MAX_ADAM_GROUPS]
                    arr = (_native.AdamGroup * len(chunk))()
                    for n, (p, g, st, lr) in enumerate(chunk):
                        arr[n] = _native.AdamGroup(p.data_ptr(), g.data_ptr(), st["exp_avg"].data_ptr(),
                                                   st["exp_avg_sq"].data_ptr(), p.numel(), lr)
                    _native.check(L.gsr_adam_step(len(chunk), arr, betas[0], betas[1], eps, step, _stream(device)),
                                  "gsr_adam_step")
        return loss


def densification_stats(radii: torch.Tensor, viewspace_grad: torch.Tensor, max_radii2D: torch.Tensor,
                        xyz_gradient_accum: torch.Tensor, denom: torch.Tensor) -> None:
    """train.py:130-131 + scene/gaussian_model.py:775-777 in one launch, in place: for Gaussians with
    radii > 0, max_radii2D = max(max_radii2D, radii); xyz_gradient_accum += ||viewspace_grad[:, :2]||;
    denom += 1.  viewspace_grad is means2D.grad [P,3]."""
    device = radii.device
    if device.type != "cuda":
        raise RuntimeError("densification_stats has no CPU path")
    P = int(radii.shape[0])
    if P == 0:
        return
    for t, shape in ((max_radii2D, (P,)), (xyz_gradient_accum, (P, 1)), (denom, (P, 1))):
        if t.dtype != torch.float32 or not t.is_contiguous() or t.numel() != P:
            raise RuntimeError("statistics tensors must be contiguous float32 with one entry per Gaussian")
    r = radii.to(torch.int32).contiguous()
    g = viewspace_grad.to(torch.float32).contiguous()
    if tuple(g.shape) != (P, 3):
        raise RuntimeError("viewspace_grad must be [P,3]")
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    with torch.cuda.device(device):
        _native.check(_native.lib().gsr_densification_stats(P, p(r), p(g), p(max_radii2D), p(xyz_gradient_accum),
                                                            p(denom), _stream(device)), "gsr_densification_stats")
