"""Fused per-step statistics update for the MS-GS training loop (include/msgs_stats.h).

`update_stats(...)` does, for one rendered (view, scale) unit and in ONE kernel launch without host synchronisation, what
train.py:241-250 does with `update_base_gaussian_mask`, `update_pixel_sizes`, the max_radii2D line and
`add_densification_stats` (scene/gaussian_model.py:662-705).  All state tensors are updated in place.  No CPU path."""
import ctypes
import os

import torch

PIXEL_SIZES, DENSIFY, BASE_MASK = 1, 2, 4

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libmsgs_raster.so")
if not os.path.exists(_LIB_PATH):
    raise ImportError(f"{_LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'`")
_lib = ctypes.CDLL(_LIB_PATH)
_vp, _ci = ctypes.c_void_p, ctypes.c_int
_lib.msgs_update_stats.restype = _ci
_lib.msgs_update_stats.argtypes = [_ci] * 4 + [_vp] * 10 + [_vp]


def _arg(name, t, dtype, numel, needed):
    if t is None:
        if needed:
            raise RuntimeError(f"update_stats: {name} is required for the requested updates")
        return _vp(0)
    if not (torch.is_tensor(t) and t.is_cuda and t.dtype == dtype and t.is_contiguous() and t.numel() == numel):
        raise RuntimeError(f"update_stats: {name} must be a contiguous {dtype} CUDA tensor of {numel} elements "
                           "(updated in place; there is no CPU path)")
    return _vp(t.data_ptr())


def update_stats(reso_lvl, reso_lvls, radii, pixel_sizes=None, target_reso_lvl=None, dL_dmeans2D=None,
                 max_pixel_sizes=None, min_pixel_sizes=None, max_radii2D=None, xyz_gradient_accum=None, denom=None,
                 base_gaussian_mask=None, what=PIXEL_SIZES | DENSIFY):
    """`what` = PIXEL_SIZES | DENSIFY | BASE_MASK selects the updates (train.py applies them under different conditions)."""
    P, L = radii.numel(), int(reso_lvls)
    f, px, dn, bm = torch.float32, bool(what & PIXEL_SIZES), bool(what & DENSIFY), bool(what & BASE_MASK)
    args = [_arg("radii", radii, torch.int32, P, True), _arg("pixel_sizes", pixel_sizes, f, P, px),
            _arg("target_reso_lvl", target_reso_lvl, torch.int64, P, px), _arg("dL_dmeans2D", dL_dmeans2D, f, 3 * P, dn),
            _arg("max_pixel_sizes", max_pixel_sizes, f, P, px), _arg("min_pixel_sizes", min_pixel_sizes, f, P, px),
            _arg("max_radii2D", max_radii2D, f, P, dn), _arg("xyz_gradient_accum", xyz_gradient_accum, f, P * L, dn),
            _arg("denom", denom, f, P * L, dn), _arg("base_gaussian_mask", base_gaussian_mask, torch.bool, P, bm)]
    dev = radii.device
    with torch.cuda.device(dev):
        rc = _lib.msgs_update_stats(P, L, int(reso_lvl), int(what), *args, _vp(torch.cuda.current_stream(dev).cuda_stream))
    if rc != 0:
        raise RuntimeError(f"msgs_update_stats failed ({rc})")
