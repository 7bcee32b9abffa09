"""Fused parameter activations for the rasterizer's caller (include/msgs_activations.h).

`activate_gaussians(features_dc, features_rest, scaling, rotation, opacity)` returns what `gaussian_renderer.render()`
obtains from five `GaussianModel` getters (gaussian_renderer/__init__.py:57-89; scene/gaussian_model.py:127-155):
`shs = cat(dc, rest)` (P,1+D,3), `scales = exp(.)`, `rotations = normalize(.)`, `opacities = sigmoid(.)` — ONE kernel
forward and ONE backward (hand-written CUDA in libmsgs_raster.so) instead of ≈ 7 element-wise launches each way.
No CPU path: CPU tensors raise, a missing library fails the import."""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libmsgs_raster.so")
if not os.path.exists(_LIB_PATH):
    raise ImportError(f"{_LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'`")
_lib = ctypes.CDLL(_LIB_PATH)
_vp, _ci = ctypes.c_void_p, ctypes.c_int
_lib.msgs_activate.restype = _ci
_lib.msgs_activate.argtypes = [_ci, _ci] + [_vp] * 9 + [_vp]
_lib.msgs_activate_backward.restype = _ci
_lib.msgs_activate_backward.argtypes = [_ci, _ci] + [_vp] * 12 + [_vp]


def _ptr(t):
    return _vp(t.data_ptr()) if t is not None else _vp(0)


def _check(name, t, shape):
    if not (torch.is_tensor(t) and t.is_cuda and t.dtype == torch.float32):
        raise RuntimeError(f"activate_gaussians: {name} must be an fp32 CUDA tensor (there is no CPU path)")
    if tuple(t.shape) != shape:
        raise RuntimeError(f"activate_gaussians: {name} has shape {tuple(t.shape)}, expected {shape}")
    t = t.contiguous()
    # the kernels use 16-byte loads / stores (msgs_activate* return MSGS_ERR_ARG on misaligned pointers): a contiguous view
    # with a storage offset that is not a multiple of 4 floats (rotation[1:], a field of a flat buffer with odd P) is cloned
    return t if t.data_ptr() % 16 == 0 else t.clone()


class _Activate(torch.autograd.Function):
    @staticmethod
    def forward(ctx, features_dc, features_rest, scaling, rotation, opacity):
        P, D = features_dc.shape[0], features_rest.shape[1]
        dc, rest = _check("features_dc", features_dc, (P, 1, 3)), _check("features_rest", features_rest, (P, D, 3))
        sraw, rraw, oraw = _check("scaling", scaling, (P, 3)), _check("rotation", rotation, (P, 4)), _check("opacity", opacity, (P, 1))
        dev = dc.device
        opt = dict(dtype=torch.float32, device=dev)
        shs, scales = torch.empty((P, 1 + D, 3), **opt), torch.empty((P, 3), **opt)
        rots, opac = torch.empty((P, 4), **opt), torch.empty((P, 1), **opt)
        with torch.cuda.device(dev):
            rc = _lib.msgs_activate(P, D, _ptr(dc), _ptr(rest), _ptr(sraw), _ptr(rraw), _ptr(oraw), _ptr(shs), _ptr(scales),
                                    _ptr(rots), _ptr(opac), _vp(torch.cuda.current_stream(dev).cuda_stream))
        if rc != 0:
            raise RuntimeError(f"msgs_activate failed ({rc})")
        ctx.save_for_backward(scales, rraw, opac)
        ctx.set_materialize_grads(False)        # an unused output's gradient arrives as None (= NULL = zero), not as zeros
        ctx.dims = (P, D)
        return shs, scales, rots, opac

    @staticmethod
    def backward(ctx, g_shs, g_scales, g_rots, g_opac):
        scales, rraw, opac = ctx.saved_tensors
        P, D = ctx.dims
        dev = scales.device
        opt = dict(dtype=torch.float32, device=dev)
        need = ctx.needs_input_grad
        g = [g_shs, g_scales, g_rots, g_opac]
        g = [x.contiguous() if x is not None else None for x in g]
        g = [x if (x is None or x.data_ptr() % 16 == 0) else x.clone() for x in g]
        out = [torch.empty(s, **opt) if n else None
               for n, s in zip(need, ((P, 1, 3), (P, D, 3), (P, 3), (P, 4), (P, 1)))]
        with torch.cuda.device(dev):
            rc = _lib.msgs_activate_backward(P, D, _ptr(scales), _ptr(rraw), _ptr(opac), _ptr(g[0]), _ptr(g[1]), _ptr(g[2]),
                                             _ptr(g[3]), _ptr(out[0]), _ptr(out[1]), _ptr(out[2]), _ptr(out[3]), _ptr(out[4]),
                                             _vp(torch.cuda.current_stream(dev).cuda_stream))
        if rc != 0:
            raise RuntimeError(f"msgs_activate_backward failed ({rc})")
        return tuple(out)


def activate_gaussians(features_dc, features_rest, scaling, rotation, opacity):
    """(shs (P,1+D,3), scales (P,3), rotations (P,4), opacities (P,1)) from the raw GaussianModel parameters."""
    return _Activate.apply(features_dc, features_rest, scaling, rotation, opacity)
