"""Fused per-step statistics update for the MS-GS training loop (include/msgs_stats.h).

`update_stats(...)` does, for one rendered (view, scale) unit and in ONE kernel launch without host synchronisation, what
train.py:241-250 does with `update_base_gaussian_mask`, `update_pixel_sizes`, the max_radii2D line and
`add_densification_stats` (scene/gaussian_model.py:662-705).  `update_stats_units(...)` does it for every unit of a step
(the scales of a view) in one launch.  `StatsDelta` is the sharded step's form: the rank's units write observations into a
delta buffer, ONE typed-lane collective reduces it over the ranks (sum / max / min lanes), every rank applies the result.
All state tensors are updated in place.  No CPU path."""
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
_ll = ctypes.c_longlong


class _StatsUnit(ctypes.Structure):         # msgs_stats_unit (include/msgs_stats.h)
    _fields_ = [("reso_lvl", _ci), ("what", _ci), ("radii", _vp), ("pixel_sizes", _vp), ("dL_dmeans2D", _vp)]


MAX_UNITS = 8
_lib.msgs_update_stats_units.restype = _ci
_lib.msgs_update_stats_units.argtypes = [_ci, _ci, _ci, _vp] + [_vp] * 7 + [_vp]
_lib.msgs_stats_delta_words.restype = None
_lib.msgs_stats_delta_words.argtypes = [_ci, _ci, ctypes.POINTER(_ll), ctypes.POINTER(_ll), ctypes.POINTER(_ll)]
_lib.msgs_stats_delta_fast_words.restype = None
_lib.msgs_stats_delta_fast_words.argtypes = [_ci, _ci, ctypes.POINTER(_ll), ctypes.POINTER(_ll), ctypes.POINTER(_ll)]
_lib.msgs_stats_delta_reset.restype = _ci
_lib.msgs_stats_delta_reset.argtypes = [_ci, _ci, _vp, _vp]
_lib.msgs_stats_delta_units.restype = _ci
_lib.msgs_stats_delta_units.argtypes = [_ci, _ci, _ci, _vp, _vp, _vp, _vp]
_lib.msgs_stats_delta_apply.restype = _ci
_lib.msgs_stats_delta_apply.argtypes = [_ci, _ci, _vp, _vp] + [_vp] * 6 + [_vp]


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


def _pack_units(P, units):
    """units: sequence of (reso_lvl, what, radii, pixel_sizes, dL_dmeans2D) -> (ctypes array, tensors kept alive)."""
    if not 0 < len(units) <= MAX_UNITS:
        raise RuntimeError(f"1..{MAX_UNITS} units per call")
    arr = (_StatsUnit * len(units))()
    f = torch.float32
    for k, (lvl, what, radii, px, g2d) in enumerate(units):
        arr[k] = _StatsUnit(int(lvl), int(what), _arg("radii", radii, torch.int32, P, True).value,
                            _arg("pixel_sizes", px, f, P, bool(what & PIXEL_SIZES)).value,
                            _arg("dL_dmeans2D", g2d, f, 3 * P, bool(what & DENSIFY)).value)
    return arr


def update_stats_units(units, reso_lvls, target_reso_lvl, max_pixel_sizes, min_pixel_sizes, max_radii2D, xyz_gradient_accum,
                       denom, base_gaussian_mask=None):
    """Every unit of a step — `units` = [(reso_lvl, what, radii, pixel_sizes, dL_dmeans2D), ...], at most 8 — applied in
    order by ONE launch; bit-identical to calling `update_stats` once per unit."""
    P, L, f = units[0][2].numel(), int(reso_lvls), torch.float32
    arr = _pack_units(P, units)
    anyw = 0
    for u in units:
        anyw |= int(u[1])
    px, dn, bm = bool(anyw & PIXEL_SIZES), bool(anyw & DENSIFY), bool(anyw & BASE_MASK)
    args = [_arg("target_reso_lvl", target_reso_lvl, torch.int64, P, px), _arg("max_pixel_sizes", max_pixel_sizes, f, P, px),
            _arg("min_pixel_sizes", min_pixel_sizes, f, P, px), _arg("max_radii2D", max_radii2D, f, P, dn),
            _arg("xyz_gradient_accum", xyz_gradient_accum, f, P * L, dn), _arg("denom", denom, f, P * L, dn),
            _arg("base_gaussian_mask", base_gaussian_mask, torch.bool, P, bm)]
    dev = units[0][2].device
    with torch.cuda.device(dev):
        rc = _lib.msgs_update_stats_units(P, L, len(units), ctypes.addressof(arr), *args,
                                          _vp(torch.cuda.current_stream(dev).cuda_stream))
    if rc != 0:
        raise RuntimeError(f"msgs_update_stats_units failed ({rc})")


def delta_words(P, L):
    a, b, c = _ll(), _ll(), _ll()
    _lib.msgs_stats_delta_words(int(P), int(L), ctypes.byref(a), ctypes.byref(b), ctypes.byref(c))
    return a.value, b.value, c.value


class StatsDelta:
    """One step's statistics observations of this rank (include/msgs_stats.h, "the sharded step").

        d = StatsDelta(P, L, device, symmetric=True)     # symmetric: the buffer is NVSwitch-multicast mapped
        d.reset(); d.add_units(units, target)            # per step; add_units once per view group
        d.allreduce(rank, world)                         # ONE launch: f32 sum + u32 max + u32 min lanes
        d.apply(target, state...)                        # every rank: identical state afterwards"""

    def __init__(self, P, L, device, symmetric=False, group=None):
        self.P, self.L, self.device = int(P), int(L), torch.device(device)
        self.n_sum, self.n_max, self.n_min = delta_words(P, L)
        n = self.n_sum + self.n_max + self.n_min
        self.handle = None
        if symmetric:
            import torch.distributed as dist
            import torch.distributed._symmetric_memory as symm_mem
            self.buf = symm_mem.empty(max(n, 4), dtype=torch.float32, device=device)
            self.handle = symm_mem.rendezvous(self.buf, group if group is not None else dist.group.WORLD)
            if not self.handle.multicast_ptr:
                raise RuntimeError("symmetric memory has no multicast mapping on this system")
        else:
            self.buf = torch.empty(max(n, 4), dtype=torch.float32, device=device)
        self.reset()

    def _st(self):
        return _vp(torch.cuda.current_stream(self.device).cuda_stream)

    def reset(self):
        with torch.cuda.device(self.device):
            rc = _lib.msgs_stats_delta_reset(self.P, self.L, _vp(self.buf.data_ptr()), self._st())
        if rc != 0:
            raise RuntimeError(f"msgs_stats_delta_reset failed ({rc})")

    def add_units(self, units, target_reso_lvl):
        arr = _pack_units(self.P, units)
        with torch.cuda.device(self.device):
            rc = _lib.msgs_stats_delta_units(self.P, self.L, len(units), ctypes.addressof(arr),
                                             _arg("target_reso_lvl", target_reso_lvl, torch.int64, self.P, True),
                                             _vp(self.buf.data_ptr()), self._st())
        if rc != 0:
            raise RuntimeError(f"msgs_stats_delta_units failed ({rc})")

    def allreduce(self, rank, world, ctas=0):
        """Stream-ordered: barrier -> typed-lane NVLS kernel over the multicast mapping -> barrier."""
        if self.handle is None:
            raise RuntimeError("StatsDelta(symmetric=True) is needed for the collective")
        import msgs_optim
        msgs_optim.stats_allreduce_multimem(self.handle, self.n_sum, self.n_max, self.n_min, rank, world, self.device, ctas)

    def allreduce_fast(self, rank, world, ctas=0, barriers=True):
        """Only the lanes the next render depends on (pixel-size observations, base mask: 3 words per Gaussian); apply() then
        adds this RANK's gradient-norm sums / visibility counts / radii, so the caller must hand it rank-local accumulators
        (msgs_train.MultiScaleStep sums them over the ranks when densification consumes them)."""
        if self.handle is None:
            raise RuntimeError("StatsDelta(symmetric=True) is needed for the collective")
        import msgs_optim
        off, n_max, n_min = _ll(), _ll(), _ll()
        _lib.msgs_stats_delta_fast_words(self.P, self.L, ctypes.byref(off), ctypes.byref(n_max), ctypes.byref(n_min))
        msgs_optim.stats_allreduce_multimem(self.handle, 0, n_max.value, n_min.value, rank, world, self.device, ctas,
                                            offset_words=off.value, barriers=barriers)

    def apply(self, target_reso_lvl, max_pixel_sizes, min_pixel_sizes, max_radii2D, xyz_gradient_accum, denom,
              base_gaussian_mask=None):
        P, L, f = self.P, self.L, torch.float32
        args = [_arg("target_reso_lvl", target_reso_lvl, torch.int64, P, True), _vp(self.buf.data_ptr()),
                _arg("max_pixel_sizes", max_pixel_sizes, f, P, True), _arg("min_pixel_sizes", min_pixel_sizes, f, P, True),
                _arg("max_radii2D", max_radii2D, f, P, True), _arg("xyz_gradient_accum", xyz_gradient_accum, f, P * L, True),
                _arg("denom", denom, f, P * L, True), _arg("base_gaussian_mask", base_gaussian_mask, torch.bool, P, False)]
        with torch.cuda.device(self.device):
            rc = _lib.msgs_stats_delta_apply(P, L, *args, self._st())
        if rc != 0:
            raise RuntimeError(f"msgs_stats_delta_apply failed ({rc})")
