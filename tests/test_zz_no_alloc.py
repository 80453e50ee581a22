"""gsr_loss_value_and_grad_static on the GPU, through the C ABI.

Some hosts of this library cannot allocate device memory at the point where they call it: a plugin whose
render / evaluate callback is forbidden to allocate, a service that handed out all of its device memory at
start-up, a loop that shares the GPU with other streams and cannot afford cudaMalloc's device-wide wait.
They cannot give gsr_loss_forward / gsr_loss_backward the workspace those calls need, so the library offers
one call that keeps that workspace in an array built into the .so (up to a 3 x 1080 x 1920 image).

What is checked here: the loss scalars and the image gradient of the new call equal those of the existing
workspace pair on the same inputs, within the tolerance of the other loss tests, on two calls in a row — the
second one smaller than the first and with a different upstream gradient, so it runs on top of the first
call's leftovers in the shared array — plus the L1-only path, the size limit, and that free device memory does
not move across a call.  Nothing here depends on what ran before it in the process.

This file sorts last on purpose: `pytest -x` stops at the first failure, and a failure here must not hide the
parity tests.
"""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
TOL = 1e-4   # the tolerance tests/test_gpu_loss.py uses for floating-point outputs


def _p(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _stream(dev):
    return ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


def _images(shape, seed, dev):
    rng = np.random.default_rng(seed)
    y = rng.random(shape).astype(np.float32)
    x = np.clip(y + 0.1 * rng.normal(size=shape), 0, 1).astype(np.float32)
    return torch.from_numpy(x).to(dev), torch.from_numpy(y).to(dev)


def _workspace_pair(native, dev, x, y, a0, a_l1, a_ssim, upstream):
    """The existing entry points: the caller allocates the workspace (here: torch does)."""
    L = native.lib()
    C, H, W = x.shape
    nbytes = ctypes.c_size_t()
    native.check(L.gsr_loss_workspace_bytes(C, H, W, 1, ctypes.byref(nbytes)), "gsr_loss_workspace_bytes")
    ws = torch.empty(nbytes.value, dtype=torch.uint8, device=dev)
    out3 = torch.empty(3, dtype=torch.float32, device=dev)
    grad = torch.empty_like(x)
    native.check(L.gsr_loss_forward(C, H, W, _p(x), _p(y), a0, a_l1, a_ssim, _p(ws), 1, _p(out3), _stream(dev)),
                 "gsr_loss_forward")
    native.check(L.gsr_loss_backward(C, H, W, _p(x), _p(y), a_l1, a_ssim, _p(ws), _p(upstream), _p(grad),
                                     _stream(dev)), "gsr_loss_backward")
    torch.cuda.synchronize(dev)
    return out3.cpu().numpy(), grad.cpu().numpy()


def _static_call(native, dev, x, y, a0, a_l1, a_ssim, upstream):
    """The new entry point: inputs and outputs only, no workspace argument."""
    C, H, W = x.shape
    out3 = torch.full((3,), float("nan"), dtype=torch.float32, device=dev)
    grad = torch.full_like(x, float("nan"))
    torch.cuda.synchronize(dev)
    free_before = torch.cuda.mem_get_info(dev)[0]
    st = native.lib().gsr_loss_value_and_grad_static(C, H, W, _p(x), _p(y), a0, a_l1, a_ssim, _p(upstream),
                                                     _p(out3), _p(grad), _stream(dev))
    torch.cuda.synchronize(dev)
    free_after = torch.cuda.mem_get_info(dev)[0]
    native.check(st, "gsr_loss_value_and_grad_static")
    return out3.cpu().numpy(), grad.cpu().numpy(), free_before, free_after


def _same(got, want, n_elems):
    from test_loss_cpu import close_grad
    np.testing.assert_allclose(got[0], want[0], rtol=TOL, atol=1e-7)   # loss, l1, ssim
    assert np.isfinite(got[1]).all()                                   # every gradient element was written
    close_grad(got[1], want[1], n_elems, TOL)


def test_two_calls_in_a_row_match_the_workspace_entry_points(cuda, native):
    lam = 0.2
    with torch.cuda.device(cuda):
        # call 1: spans tile borders in both directions; fills a large part of the shared array
        x1, y1 = _images((3, 203, 331), 11, cuda)
        want1 = _workspace_pair(native, cuda, x1, y1, lam, 1.0 - lam, -lam, None)
        got1 = _static_call(native, cuda, x1, y1, lam, 1.0 - lam, -lam, None)
        _same(got1, want1, x1.numel())
        # call 2, straight after: a smaller image and an upstream gradient of 3, on top of call 1's leftovers
        x2, y2 = _images((3, 70, 45), 12, cuda)
        up = torch.tensor([3.0], dtype=torch.float32, device=cuda)
        want2 = _workspace_pair(native, cuda, x2, y2, lam, 1.0 - lam, -lam, up)
        got2 = _static_call(native, cuda, x2, y2, lam, 1.0 - lam, -lam, up)
        _same(got2, want2, x2.numel())
        # by now the array's address is known and its module loaded: a call leaves free device memory as it was
        assert got2[2] == got2[3], "device memory moved across a call that must not allocate"
        # the same inputs again give the same answer (nothing accumulates from call to call)
        got3 = _static_call(native, cuda, x2, y2, lam, 1.0 - lam, -lam, up)
        _same(got3, want2, x2.numel())
        assert got3[2] == got3[3]


def test_l1_only_path_and_the_size_limit(cuda, native):
    L = native.lib()
    with torch.cuda.device(cuda):
        x, y = _images((3, 64, 100), 13, cuda)
        want = _workspace_pair(native, cuda, x, y, 0.0, 1.0, 0.0, None)
        got = _static_call(native, cuda, x, y, 0.0, 1.0, 0.0, None)
        _same(got, want, x.numel())
        # one element row beyond the built-in scratch: refused with its own status before any launch
        C, W = 3, 1920
        H = native.LOSS_STATIC_MAX_ELEMS // (C * W) + 1
        before = L.gsr_launch_count()
        sentinel = torch.full((3,), -7.0, dtype=torch.float32, device=cuda)
        st = L.gsr_loss_value_and_grad_static(C, H, W, _p(x), _p(y), 0.2, 0.8, -0.2, None, _p(sentinel), _p(x),
                                              _stream(cuda))
        torch.cuda.synchronize(cuda)
        assert st == native.GSR_ERR_SCRATCH_LIMIT
        assert b"scratch" in L.gsr_status_string(st)
        assert L.gsr_launch_count() == before and (sentinel == -7.0).all()
        assert L.gsr_loss_value_and_grad_static(3, 8, 8, None, _p(y), 0.2, 0.8, -0.2, None, _p(sentinel), _p(x),
                                                _stream(cuda)) == native.GSR_ERR_INVALID_ARGUMENT


def test_largest_supported_image_uses_the_whole_array(cuda, native):
    """C * H * W == GSR_LOSS_STATIC_MAX_ELEMS exactly: the last map element sits at the array's last bytes."""
    lam = 0.2
    with torch.cuda.device(cuda):
        x, y = _images((3, 1080, 1920), 14, cuda)
        assert x.numel() == native.LOSS_STATIC_MAX_ELEMS
        want = _workspace_pair(native, cuda, x, y, lam, 1.0 - lam, -lam, None)
        got = _static_call(native, cuda, x, y, lam, 1.0 - lam, -lam, None)
        _same(got, want, x.numel())
