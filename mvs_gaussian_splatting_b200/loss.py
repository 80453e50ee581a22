"""The step between forward and backward: the photometric loss of the training loop on the
library's sm_100a kernels (csrc/gsr_loss.cu) — two launches (one forward, one backward) instead of five
grouped 11x11 convolutions, ~15 elementwise kernels and their autograd.

Mirrors /root/reference/utils/loss_utils.py: `l1_loss(network_output, gt)` (:17-18) and
`ssim(img1, img2, window_size=11, size_average=True)` (:39-63) keep their names, argument meaning and
results; `l1_dssim_loss(image, gt, lambda_dssim)` is the combination train.py:99-101 forms from them,
`(1 - lambda) * l1 + lambda * (1 - ssim)`, evaluated in ONE pass over the two images.  All three are
differentiable w.r.t. the first argument (the rendered image).  No CPU path.
"""
from __future__ import annotations

import ctypes

import torch

from . import _native

__all__ = ["l1_loss", "ssim", "l1_dssim_loss", "photometric_terms"]


def _ptr(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _stream(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _as_target(image: torch.Tensor, gt) -> torch.Tensor:
    """`gt` as the reference's torch expressions accept it: a tensor of the image's shape, a tensor that
    broadcasts against it, or a Python scalar (train.py:105 calls `l1_loss(opacity[mask], 1)`); expanded
    to the image's shape on its device."""
    if not isinstance(gt, torch.Tensor):
        gt = torch.as_tensor(gt, dtype=torch.float32, device=image.device)
    elif gt.device != image.device:
        raise ValueError("image and gt must be on the same device")
    if gt.shape != image.shape:
        gt = gt.to(torch.float32).expand(image.shape)   # raises if not broadcastable
    return gt


def _planes(image: torch.Tensor, gt: torch.Tensor):
    if image.device.type != "cuda":
        raise RuntimeError("the fused photometric loss has no CPU path: tensors must be on a CUDA device")
    if image.shape != gt.shape or image.dim() < 2:
        raise ValueError(f"image {tuple(image.shape)} and gt {tuple(gt.shape)} must have the same [..., H, W] shape")
    if gt.device != image.device:
        raise ValueError("image and gt must be on the same device")
    H, W = int(image.shape[-2]), int(image.shape[-1])
    C = 1
    for d in image.shape[:-2]:
        C *= int(d)
    if C * H * W == 0:
        raise ValueError("empty image")
    return C, H, W


class _PhotometricLoss(torch.autograd.Function):
    """out3 = [a0 + a_l1 * l1 + a_ssim * ssim, l1, ssim]; gradient flows from out3[0] only."""

    @staticmethod
    def forward(ctx, image, gt, a0, a_l1, a_ssim):
        C, H, W = _planes(image, gt)
        x = image.detach().to(torch.float32).contiguous()
        y = gt.detach().to(torch.float32).contiguous()
        need_grad = bool(ctx.needs_input_grad[0])
        save = int(need_grad and float(a_ssim) != 0.0)   # the L1-only path keeps no per-pixel maps
        L = _native.lib()
        nbytes = ctypes.c_size_t()
        _native.check(L.gsr_loss_workspace_bytes(C, H, W, save, ctypes.byref(nbytes)),
                      "gsr_loss_workspace_bytes")
        ws = torch.empty(nbytes.value, dtype=torch.uint8, device=x.device)
        out3 = torch.empty(3, dtype=torch.float32, device=x.device)
        with torch.cuda.device(x.device):
            _native.check(L.gsr_loss_forward(C, H, W, _ptr(x), _ptr(y), float(a0), float(a_l1), float(a_ssim),
                                             _ptr(ws), save, _ptr(out3), _stream(x.device)),
                          "gsr_loss_forward")
        ctx.save_for_backward(x, y, ws)
        ctx.dims, ctx.coef, ctx.in_dtype = (C, H, W), (float(a_l1), float(a_ssim)), image.dtype
        return out3

    @staticmethod
    def backward(ctx, g_out3):
        x, y, ws = ctx.saved_tensors
        C, H, W = ctx.dims
        up = g_out3[0:1].to(torch.float32).contiguous()   # l1 / ssim entries are reported values, not graph outputs
        grad = torch.empty_like(x)
        with torch.cuda.device(x.device):
            _native.check(_native.lib().gsr_loss_backward(C, H, W, _ptr(x), _ptr(y), ctx.coef[0], ctx.coef[1],
                                                          _ptr(ws), _ptr(up), _ptr(grad), _stream(x.device)),
                          "gsr_loss_backward")
        return grad.to(ctx.in_dtype), None, None, None, None


def photometric_terms(image: torch.Tensor, gt: torch.Tensor, a0: float, a_l1: float, a_ssim: float) -> torch.Tensor:
    """Device tensor [3] = (a0 + a_l1 * l1 + a_ssim * ssim, l1, ssim) from one pass; differentiate entry 0."""
    return _PhotometricLoss.apply(image, gt, a0, a_l1, a_ssim)


def l1_dssim_loss(image: torch.Tensor, gt: torch.Tensor, lambda_dssim: float = 0.2) -> torch.Tensor:
    """train.py:100-101: (1 - lambda_dssim) * l1_loss(image, gt) + lambda_dssim * (1 - ssim(image, gt))."""
    lam = float(lambda_dssim)
    return photometric_terms(image, gt, lam, 1.0 - lam, -lam)[0]


def l1_loss(network_output: torch.Tensor, gt: torch.Tensor) -> torch.Tensor:
    """utils/loss_utils.py:17-18: mean |network_output - gt|.  `gt` may be a scalar or any tensor that
    broadcasts against `network_output` (train.py:105); inputs with fewer than two dimensions are treated
    as one row."""
    gt = _as_target(network_output, gt)
    if network_output.numel() == 0:
        return torch.abs(network_output - gt).mean()   # nan, as torch gives for an empty selection
    if network_output.dim() < 3:   # not an image (e.g. the [K,1] opacity rows of train.py:105): one long row
        network_output, gt = network_output.reshape(1, -1), gt.reshape(1, -1)
    return photometric_terms(network_output, gt, 0.0, 1.0, 0.0)[0]


def ssim(img1: torch.Tensor, img2: torch.Tensor, window_size: int = 11, size_average: bool = True) -> torch.Tensor:
    """utils/loss_utils.py:39-63 with the default 11-tap sigma-1.5 window and size_average=True."""
    if window_size != 11:
        raise NotImplementedError("only the reference's default window_size=11 is built")
    if not size_average:
        raise NotImplementedError("size_average=False (per-batch means) is not built; the training loop uses True")
    return photometric_terms(img1, img2, 0.0, 0.0, 1.0)[0]
