"""Fused training loss of MS-GS (include/msgs_loss.h).

`l1_ssim_loss(image, gt, lambda_dssim)` = `(1 - lambda) * l1_loss(image, gt) + lambda * (1 - ssim(image, gt))`
(train.py:210-211; utils/loss_utils.py:17-64), differentiable w.r.t. `image`: one CUDA launch forward (+ one reduction)
and one backward instead of five grouped conv2d and ~15 element-wise launches each way.
`l1_ssim_loss_batch(images, gts, weights, lambda_dssim)` = the sum over the images of a step (the 7 scales of a view) of
`weight_i * loss_i` (train.py:212-214: 1 at the finest scale, 0.1 at the coarse ones) — ONE forward launch (per-image means
reduced inside it) and ONE backward launch for all of them.  No CPU path."""
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


class _LossImage(ctypes.Structure):          # msgs_loss_image (include/msgs_loss.h)
    _fields_ = [("C", _ci), ("H", _ci), ("W", _ci), ("weight", _cf), ("img", _vp), ("gt", _vp), ("dmu", _vp), ("de11", _vp),
                ("de12", _vp), ("loss_px", _vp), ("dL_dimg", _vp)]


MAX_IMAGES = 8
_lib.msgs_l1_ssim_forward_batch.restype = _ci
_lib.msgs_l1_ssim_forward_batch.argtypes = [_ci, _vp, _cf, _vp, _vp, _vp]
_lib.msgs_l1_ssim_backward_batch.restype = _ci
_lib.msgs_l1_ssim_backward_batch.argtypes = [_ci, _vp, _cf, _vp, _vp, _vp]
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


def _descriptors(imgs, gts, weights, maps, grads):
    arr = (_LossImage * len(imgs))()
    for i, (im, g, w, m) in enumerate(zip(imgs, gts, weights, maps)):
        C, H, W = im.shape
        n = C * H * W
        arr[i] = _LossImage(C, H, W, float(w), im.data_ptr(), g.data_ptr(), m.data_ptr(), m.data_ptr() + 4 * n,
                            m.data_ptr() + 8 * n, None, grads[i].data_ptr() if grads is not None else None)
    return arr


class _L1SSIMBatch(torch.autograd.Function):
    @staticmethod
    def forward(ctx, lam, weights, gts, *images):
        n = len(images)
        if not 0 < n <= MAX_IMAGES or len(gts) != n or len(weights) != n:
            raise RuntimeError(f"l1_ssim_loss_batch: 1..{MAX_IMAGES} images with one gt and one weight each")
        for name, ts in (("image", images), ("gt", gts)):
            for t in ts:
                if not (torch.is_tensor(t) and t.is_cuda and t.dtype == torch.float32 and t.dim() == 3):
                    raise RuntimeError(f"l1_ssim_loss_batch: {name} must be a (C,H,W) fp32 CUDA tensor (there is no CPU path)")
        dev = images[0].device
        imgs = [t.contiguous() for t in images]
        gs = [t.contiguous() for t in gts]
        for a, b in zip(imgs, gs):
            if a.shape != b.shape or a.device != dev or b.device != dev:
                raise RuntimeError("l1_ssim_loss_batch: image and gt differ in shape or device")
        w = window_taps(dev)
        maps = [torch.empty(3 * a.numel(), dtype=torch.float32, device=dev) for a in imgs]
        losses = torch.empty(n, dtype=torch.float32, device=dev)
        desc = _descriptors(imgs, gs, weights, maps, None)
        with torch.cuda.device(dev):
            rc = _lib.msgs_l1_ssim_forward_batch(n, ctypes.addressof(desc), float(lam), _vp(w.data_ptr()), _vp(losses.data_ptr()),
                                                 _vp(torch.cuda.current_stream(dev).cuda_stream))
        if rc != 0:
            raise RuntimeError(f"msgs_l1_ssim_forward_batch failed ({rc})")
        ctx.save_for_backward(w, *imgs, *gs, *maps)
        ctx.lam, ctx.weights, ctx.n = float(lam), [float(x) for x in weights], n
        ctx.mark_non_differentiable(losses)
        return losses.sum(), losses

    @staticmethod
    def backward(ctx, grad_total, _grad_losses):
        n = ctx.n
        w, rest = ctx.saved_tensors[0], ctx.saved_tensors[1:]
        imgs, gs, maps = rest[:n], rest[n:2 * n], rest[2 * n:]
        dev = imgs[0].device
        up = grad_total.to(torch.float32).reshape(1).contiguous()
        grads = [torch.empty_like(a) for a in imgs]
        desc = _descriptors(imgs, gs, ctx.weights, maps, grads)
        with torch.cuda.device(dev):
            rc = _lib.msgs_l1_ssim_backward_batch(n, ctypes.addressof(desc), ctx.lam, _vp(w.data_ptr()), _vp(up.data_ptr()),
                                                  _vp(torch.cuda.current_stream(dev).cuda_stream))
        if rc != 0:
            raise RuntimeError(f"msgs_l1_ssim_backward_batch failed ({rc})")
        return (None, None, None, *grads)


def l1_ssim_loss_batch(images, gts, weights, lambda_dssim=0.2):
    """(total, per_image): total = sum_i weights[i] * ((1 - lambda) L1_i + lambda (1 - SSIM_i)), differentiable w.r.t. every
    image; per_image = the weighted terms (detached; what train.py:396 logs).  Up to 8 images of any sizes per call."""
    return _L1SSIMBatch.apply(lambda_dssim, list(weights), list(gts), *images)


_ones = {}


def l1_ssim_loss_and_grad_batch(images, gts, weights, lambda_dssim=0.2):
    """(total, per_image, grads) without autograd: total as l1_ssim_loss_batch returns it and grads[i] = d total / d images[i]
    (the upstream gradient of a training loss is 1): the forward launch and the backward launch go out back to back.  For a
    caller that drives the rasterizer's backward itself (msgs_train.MultiScaleStep)."""
    n = len(images)
    if not 0 < n <= MAX_IMAGES or len(gts) != n or len(weights) != n:
        raise RuntimeError(f"l1_ssim_loss_and_grad_batch: 1..{MAX_IMAGES} images with one gt and one weight each")
    dev = images[0].device
    imgs, gs = [], []
    for a, b in zip(images, gts):
        if not (a.is_cuda and b.is_cuda and a.dtype == torch.float32 and b.dtype == torch.float32 and a.dim() == 3 and
                a.shape == b.shape):
            raise RuntimeError("l1_ssim_loss_and_grad_batch: (C,H,W) fp32 CUDA image / gt pairs of equal shape (no CPU path)")
        imgs.append(a.detach().contiguous())
        gs.append(b.contiguous())
    w = window_taps(dev)
    key = str(dev)
    if key not in _ones:
        _ones[key] = torch.ones(1, dtype=torch.float32, device=dev)
    total_el = sum(a.numel() for a in imgs)
    maps_all = torch.empty(3 * total_el, dtype=torch.float32, device=dev)
    grads_all = torch.empty(total_el, dtype=torch.float32, device=dev)
    losses = torch.empty(n, dtype=torch.float32, device=dev)
    maps, grads, mo, go = [], [], 0, 0
    for a in imgs:
        maps.append(maps_all[mo:mo + 3 * a.numel()])
        grads.append(grads_all[go:go + a.numel()].view(a.shape))
        mo += 3 * a.numel()
        go += a.numel()
    desc = _descriptors(imgs, gs, weights, maps, grads)
    st = _vp(torch.cuda.current_stream(dev).cuda_stream)
    with torch.cuda.device(dev):
        rc = _lib.msgs_l1_ssim_forward_batch(n, ctypes.addressof(desc), float(lambda_dssim), _vp(w.data_ptr()),
                                             _vp(losses.data_ptr()), st)
        if rc == 0:
            rc = _lib.msgs_l1_ssim_backward_batch(n, ctypes.addressof(desc), float(lambda_dssim), _vp(w.data_ptr()),
                                                  _vp(_ones[key].data_ptr()), st)
    if rc != 0:
        raise RuntimeError(f"msgs_l1_ssim_*_batch failed ({rc})")
    return losses.sum(), losses, grads
