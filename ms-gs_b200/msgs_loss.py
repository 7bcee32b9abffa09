"""Fused training loss of MS-GS (include/msgs_loss.h).

`l1_ssim_loss(image, gt, lambda_dssim)` = `(1 - lambda) * l1_loss(image, gt) + lambda * (1 - ssim(image, gt))`
(train.py:210-211; utils/loss_utils.py:17-64), differentiable w.r.t. `image`: one CUDA launch forward (+ one reduction)
and one backward instead of five grouped conv2d and ~15 element-wise launches each way.  No CPU path."""
import ctypes
import os
from math import exp

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libmsgs_raster.so")
if not os.path.exists(_LIB_PATH):
    raise ImportError(f"{_LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'`")
_lib = ctypes.CDLL(_LIB_PATH)
_vp, _ci, _cf = ctypes.c_void_p, ctypes.c_int, ctypes.c_float
_lib.msgs_l1_ssim_forward.restype = _ci
_lib.msgs_l1_ssim_forward.argtypes = [_ci, _ci, _ci, _cf] + [_vp] * 7 + [_vp]
_lib.msgs_l1_ssim_backward.restype = _ci
_lib.msgs_l1_ssim_backward.argtypes = [_ci, _ci, _ci, _cf] + [_vp] * 8 + [_vp]
_taps = {}


def window_taps(device):
    """gaussian(11, 1.5) of utils/loss_utils.py:24-26, built with the same float32 torch ops, cached per device."""
    key = str(device)
    if key not in _taps:
        g = torch.Tensor([exp(-(x - 11 // 2) ** 2 / float(2 * 1.5 ** 2)) for x in range(11)])
        _taps[key] = (g / g.sum()).to(device)
    return _taps[key]


class _L1SSIM(torch.autograd.Function):
    @staticmethod
    def forward(ctx, image, gt, lam):
        for name, t in (("image", image), ("gt", gt)):
            if not (torch.is_tensor(t) and t.is_cuda and t.dtype == torch.float32 and t.dim() == 3):
                raise RuntimeError(f"l1_ssim_loss: {name} must be a (C,H,W) fp32 CUDA tensor (there is no CPU path)")
        if image.shape != gt.shape:
            raise RuntimeError("l1_ssim_loss: image and gt differ in shape")
        img, g = image.contiguous(), gt.contiguous()
        C, H, W = img.shape
        dev = img.device
        w = window_taps(dev)
        maps = torch.empty((4, C, H, W), dtype=torch.float32, device=dev)
        st = _vp(torch.cuda.current_stream(dev).cuda_stream)
        with torch.cuda.device(dev):
            rc = _lib.msgs_l1_ssim_forward(C, H, W, float(lam), _vp(img.data_ptr()), _vp(g.data_ptr()), _vp(w.data_ptr()),
                                           _vp(maps[0].data_ptr()), _vp(maps[1].data_ptr()), _vp(maps[2].data_ptr()),
                                           _vp(maps[3].data_ptr()), st)
        if rc != 0:
            raise RuntimeError(f"msgs_l1_ssim_forward failed ({rc})")
        ctx.save_for_backward(img, g, w, maps)
        ctx.lam = float(lam)
        return maps[0].mean()

    @staticmethod
    def backward(ctx, grad_loss):
        img, g, w, maps = ctx.saved_tensors
        C, H, W = img.shape
        dev = img.device
        up = grad_loss.to(torch.float32).reshape(1).contiguous()
        out = torch.empty_like(img)
        with torch.cuda.device(dev):
            rc = _lib.msgs_l1_ssim_backward(C, H, W, ctx.lam, _vp(img.data_ptr()), _vp(g.data_ptr()), _vp(w.data_ptr()),
                                            _vp(maps[1].data_ptr()), _vp(maps[2].data_ptr()), _vp(maps[3].data_ptr()),
                                            _vp(up.data_ptr()), _vp(out.data_ptr()),
                                            _vp(torch.cuda.current_stream(dev).cuda_stream))
        if rc != 0:
            raise RuntimeError(f"msgs_l1_ssim_backward failed ({rc})")
        return out, None, None


def l1_ssim_loss(image, gt, lambda_dssim=0.2):
    return _L1SSIM.apply(image, gt, lambda_dssim)
