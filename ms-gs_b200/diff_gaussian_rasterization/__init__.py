"""Drop-in `diff_gaussian_rasterization` package backed by the B200-native rasterizer (libmsgs_raster.so).

Public surface kept from the reference package
(/root/reference/submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py):
  * GaussianRasterizationSettings  — 15-field NamedTuple (:197-212)
  * GaussianRasterizer(nn.Module)  — forward(...) (:230-269), markVisible(positions) (:219-228)
  * rasterize_gaussians(...)       — 14 positional arguments (:21-52)
with the same argument order, the same 5 returned tensors (color, acc_pixel_size, depth, radii, pixel_sizes),
the same 14-slot gradient tuple, the same exception messages and the same debug-snapshot behaviour.
`gaussian_renderer.render()` (gaussian_renderer/__init__.py:14,55,94-108) runs unmodified on top of it.
"""
from typing import NamedTuple

import torch
import torch.nn as nn

from . import _C

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians", "rasterize_gaussians_units"]


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool
    filter_small: bool
    filter_large: bool
    fade_size: float


def _snapshot(args):
    """CPU copies of every tensor argument, taken before the call so a crash cannot corrupt them (:17-19)."""
    return tuple(a.detach().cpu().clone() if isinstance(a, torch.Tensor) else a for a in args)


def _guarded(fn, args, debug, dump_name, label):
    """Run one `_C` call; in debug mode dump the inputs to `dump_name` if it raises (:105-113, :165-173)."""
    if not debug:
        return fn(*args)
    saved = _snapshot(args)
    try:
        return fn(*args)
    except Exception:
        torch.save(saved, dump_name)
        print(f"\nAn error occured in {label}. Please forward {dump_name} for debugging.")
        raise


# Gradient sink (extension; sharded training).  A data-parallel trainer that sums the per-Gaussian gradients
# over the ranks wants them in ONE flat buffer (one collective, or the symmetric buffer of the NVLS all-reduce) instead
# of five separately allocated tensors.  `set_gradient_sink({"mean3D", "sh", "opacity", "scale", "rot"} -> tensors)`
# makes the backward of both autograd Functions write / accumulate those five gradients into the given tensors (e.g.
# `msgs_sharding.GradientExchange.shaped()`) and return fresh views of them, so that autograd installs the views as the
# parameters' `.grad` without a copy; `gradient_sink_installed(params)` tells the caller whether that happened (it then
# all-reduces the flat buffer once; otherwise it falls back to one collective per `.grad`).  None switches it off.  The
# sink serves ONE backward per step (a second render of the same step would overwrite it): multi-unit steps go through
# `rasterize_gaussians_units`.
_gradient_sink = None
_SINK_KEYS = ("mean3D", "sh", "opacity", "scale", "rot")
# Slice hook (extension; sharded training): set_backward_slices([0, b1, ..., P], hook) makes the backward of
# `rasterize_gaussians_units` issue the preprocess backward of a step that consists of ONE view group slice by slice of the
# Gaussians and call hook(k, n_slices, stream) right after slice k has been enqueued on `stream` — with a gradient sink
# installed the sink then holds the FINAL gradients of Gaussians [b_k, b_k+1), and a data-parallel trainer can start
# exchanging them while the next slice is being computed.  A step with several view groups is not sliced (its gradients
# are only final after the last group) and the hook is not called.  None switches it off.
_backward_slices = None


def set_backward_slices(bounds, hook):
    global _backward_slices
    _backward_slices = None if bounds is None else (list(bounds), hook)


def set_gradient_sink(views):
    global _gradient_sink
    if views is not None:
        missing = [k for k in _SINK_KEYS if k not in views]
        if missing:
            raise Exception(f'gradient sink lacks {missing}')
        for k in _SINK_KEYS:
            v = views[k]
            if not (torch.is_tensor(v) and v.dtype == torch.float32 and v.is_contiguous()):
                raise Exception(f"gradient sink['{k}'] must be a contiguous fp32 tensor")
        views = {k: views[k] for k in _SINK_KEYS}
    _gradient_sink = views


def gradient_sink_installed(means3D, shs, opacities, scales, rotations):
    """True when every parameter's .grad is (a view of) its sink tensor, i.e. the flat buffer holds the step's gradients."""
    sink = _gradient_sink
    if sink is None:
        return False
    pairs = ((means3D, "mean3D"), (shs, "sh"), (opacities, "opacity"), (scales, "scale"), (rotations, "rot"))
    return all(p.grad is not None and p.grad.data_ptr() == sink[k].data_ptr() and p.grad.numel() == sink[k].numel()
               for p, k in pairs)


def _sink_for(P, M, inputs=()):
    """The sink's tensors viewed in the gradient shapes (fresh views: autograd may adopt them), or None.

    `inputs`: the forward's parameter tensors.  If one of them already HOLDS a .grad that lives in the sink (a second
    backward before .grad was reset, or optimizer.zero_grad(set_to_none=False)), writing the new gradient into the sink
    and returning a view of it would make autograd add the buffer to itself (2x the gradient).  The sink is then not used
    for this backward: the gradients come back in fresh tensors and autograd accumulates them into the existing .grad —
    which IS the sink memory, so the flat buffer still ends up holding the accumulated gradient."""
    sink = _gradient_sink
    if sink is None:
        return None
    ptrs = {v.data_ptr() for v in sink.values()}
    for t in inputs:
        g = getattr(t, "grad", None)
        if g is not None and g.data_ptr() in ptrs:
            return None
    shapes = {"mean3D": (P, 3), "sh": (P, M, 3), "opacity": (P, 1), "scale": (P, 3), "rot": (P, 4)}
    out = {}
    for k, shp in shapes.items():
        n = 1
        for d in shp:
            n *= d
        if sink[k].numel() != n:
            raise Exception(f"gradient sink['{k}'] has {sink[k].numel()} elements, the backward writes {n}")
        out[k] = sink[k].view(shp)
    return out


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, max_pixel_sizes, min_pixel_sizes, opacities,
                occ_multiplier, dc_delta, scales, rotations, cov3Ds_precomp, base_mask, raster_settings):
        rs = raster_settings
        # argument order of _C.rasterize_gaussians (rasterize_points.h:18-47)
        call = (rs.bg, means3D, colors_precomp, max_pixel_sizes, min_pixel_sizes, opacities, occ_multiplier, dc_delta,
                scales, rotations, rs.scale_modifier, cov3Ds_precomp, base_mask, rs.viewmatrix, rs.projmatrix,
                rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, sh, rs.sh_degree, rs.campos, rs.prefiltered,
                rs.filter_small, rs.filter_large, rs.fade_size, rs.debug)
        (num_rendered, color, acc_pixel_size, depth, radii, pixel_sizes, geom_buf, binning_buf, img_buf) = _guarded(
            _C.rasterize_gaussians, call, rs.debug, "snapshot_fw.dump", "forward")

        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, base_mask, radii, sh,
                              geom_buf, binning_buf, img_buf, pixel_sizes, max_pixel_sizes, min_pixel_sizes, opacities)
        ctx.mark_non_differentiable(radii)
        return color, acc_pixel_size, depth, radii, pixel_sizes

    @staticmethod
    def backward(ctx, grad_out_color, grad_acc_pixel_size, grad_out_depth, grad_radii, grad_pixel_sizes):
        # Only grad_out_color is consumed, as in the reference (quirk Q8, :126-162).
        rs = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, base_mask, radii, sh, geom_buf, binning_buf,
         img_buf, pixel_sizes, max_pixel_sizes, min_pixel_sizes, opacities) = ctx.saved_tensors
        # argument order of _C.rasterize_gaussians_backward (rasterize_points.h:49-80)
        call = (rs.bg, means3D, radii, colors_precomp, pixel_sizes, max_pixel_sizes, min_pixel_sizes, scales,
                rotations, rs.scale_modifier, cov3Ds_precomp, base_mask, rs.viewmatrix, rs.projmatrix, rs.tanfovx,
                rs.tanfovy, grad_out_color, sh, rs.sh_degree, rs.campos, geom_buf, ctx.num_rendered, binning_buf,
                img_buf, rs.filter_small, rs.filter_large, rs.fade_size, rs.debug)
        # gradients of inputs that do not require one are neither allocated nor written (the reference fills and returns
        # all of them; autograd drops them): forward input slots 1 means2D, 3 colors_precomp, 7 occ_multiplier, 8 dc_delta,
        # 11 cov3Ds_precomp
        need = ctx.needs_input_grad
        skip = tuple(n for n, i in (("means2D", 1), ("colors", 3), ("occ_multiplier", 7), ("dc_delta", 8), ("cov3D", 11))
                     if not need[i])
        # SH + scale/rotation inputs (the training configuration) with a sink set: the five parameter gradients are
        # written into the caller's flat buffer (set_gradient_sink above)
        sink = None
        if _gradient_sink is not None and sh.numel() != 0:
            sink = _sink_for(means3D.size(0), sh.size(1), (means3D, sh, opacities, scales, rotations))
        (g_means2D, g_colors_precomp, g_opacities, g_occ_multiplier, g_dc_delta, g_means3D, g_cov3Ds_precomp, g_sh,
         g_scales, g_rotations) = _guarded(lambda *a: _C.rasterize_gaussians_backward(*a, skip=skip, out=sink), call,
                                           rs.debug, "snapshot_bw.dump", "backward")
        # one slot per forward input (:178-193); max/min_pixel_sizes, base_mask and the settings get None
        return (g_means3D, g_means2D, g_sh, g_colors_precomp, None, None, g_opacities, g_occ_multiplier, g_dc_delta,
                g_scales, g_rotations, g_cov3Ds_precomp, None, None)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, max_pixel_sizes, min_pixel_sizes, opacities,
                        occ_multiplier, dc_delta, scales, rotations, cov3Ds_precomp, base_mask, raster_settings):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, max_pixel_sizes, min_pixel_sizes,
                                     opacities, occ_multiplier, dc_delta, scales, rotations, cov3Ds_precomp,
                                     base_mask, raster_settings)


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        """Boolean mask of the points whose view-space depth exceeds the near plane (0.2)."""
        rs = self.raster_settings
        with torch.no_grad():
            return _C.mark_visible(positions, rs.viewmatrix, rs.projmatrix)

    def forward(self, means3D, means2D, opacities, occ_multiplier=None, dc_delta=None, shs=None, colors_precomp=None,
                base_mask=None, scales=None, rotations=None, cov3D_precomp=None, max_pixel_sizes=None,
                min_pixel_sizes=None):
        have_sh, have_col = shs is not None, colors_precomp is not None
        if have_sh == have_col:
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        have_sr = scales is not None and rotations is not None
        any_sr = scales is not None or rotations is not None
        have_cov = cov3D_precomp is not None
        if (not have_sr and not have_cov) or (any_sr and have_cov):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')

        empty = torch.Tensor([])     # "not provided" is an empty tensor -> NULL pointer in the core
        shs = empty if shs is None else shs
        colors_precomp = empty if colors_precomp is None else colors_precomp
        scales = empty if scales is None else scales
        rotations = empty if rotations is None else rotations
        cov3D_precomp = empty if cov3D_precomp is None else cov3D_precomp

        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, max_pixel_sizes, min_pixel_sizes, opacities,
                                   occ_multiplier, dc_delta, scales, rotations, cov3D_precomp, base_mask,
                                   self.raster_settings)


# ------------------------------------------------------------------------------------------------------------------
# Extension: all (view, scale) units of one training step in ONE autograd node.
#
# The reference renders one (view, scale) pair per iteration (train.py:152-204); a step that renders several units of
# the same Gaussians through `rasterize_gaussians` pays, per unit, a dense 364 B/Gaussian gradient write plus autograd's
# accumulation of it, and runs the units back to back although they are independent.  `rasterize_gaussians_units`
# renders the units on round-robin CUDA streams (msgs_sharding.UnitStreams) and its backward ADDS every unit's training
# gradients into one set of accumulators (msgs_backward_accumulate: reductions, rendered Gaussians only), so autograd
# sees one gradient per parameter.  Per unit the results are those of `rasterize_gaussians` with that unit's settings.
# ------------------------------------------------------------------------------------------------------------------
_unit_streams = {}


def _streams_for(device):
    try:
        import msgs_sharding                   # ms-gs_b200/msgs_sharding.py (the directory this package lives in)
    except ImportError:
        import os
        import sys
        sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        import msgs_sharding
    key = (device.type, device.index)
    if key not in _unit_streams:
        _unit_streams[key] = msgs_sharding.UnitStreams(device)
    return _unit_streams[key]


def _view_groups(settings, view_ids=None):
    """Consecutive units that share the camera — declared by the caller through `view_ids` (equal id = same viewmatrix,
    projmatrix and campos values), else recognised by the camera tensors being the same storage — and the FoV, flags and
    SH degree, and differ only in the image size / background: the scales of one view.  MSGS_VIEW_GROUPS=0 turns
    grouping off."""
    import os
    groups = []
    on = os.environ.get("MSGS_VIEW_GROUPS", "1") != "0"

    def key(i):
        rs = settings[i]
        cam = (("id", view_ids[i]) if view_ids is not None else
               (rs.viewmatrix.data_ptr(), rs.projmatrix.data_ptr(), rs.campos.data_ptr()))
        return (cam, float(rs.tanfovx),
                float(rs.tanfovy), float(rs.scale_modifier), int(rs.sh_degree), bool(rs.prefiltered), bool(rs.debug),
                bool(rs.filter_small), bool(rs.filter_large))
    for i in range(len(settings)):
        if on and groups and len(groups[-1]) < _C.MAX_GROUP and key(groups[-1][0]) == key(i):
            groups[-1].append(i)
        else:
            groups.append([i])
    return groups


def _units_forward_impl(ctx, means3D, sh, opacities, scales, rotations, max_pixel_sizes, min_pixel_sizes, occ_multiplier,
                        dc_delta, base_mask, settings, aux, view_ids, raw):
        """forward of every unit (shared by the autograd Function and rasterize_gaussians_units_begin); `ctx` receives what
        the backward needs"""
        dev = means3D.device
        us = _streams_for(dev)
        cur = torch.cuda.current_stream(dev)
        empty = torch.Tensor([])
        saved, colors = [], []
        us.begin()
        # first every unit's kernels go out on its stream (msgs_forward_begin: no waiting), then the read-backs
        pending = [None] * len(settings)
        ctx.groups = _view_groups(settings, view_ids)
        ctx.unit_streams = ust = us.plan(ctx.groups, [int(rs.image_height) * int(rs.image_width) for rs in settings])
        for grp in ctx.groups:
            rs = settings[grp[0]]
            if len(grp) == 1 or means3D.size(0) == 0:
                for i in grp:
                    rs = settings[i]
                    with torch.cuda.stream(ust[i]):
                        pending[i] = _C.rasterize_gaussians_begin(
                            rs.bg, means3D, empty, max_pixel_sizes, min_pixel_sizes, opacities, occ_multiplier, dc_delta,
                            scales, rotations, rs.scale_modifier, empty, base_mask, rs.viewmatrix, rs.projmatrix, rs.tanfovx,
                            rs.tanfovy, rs.image_height, rs.image_width, sh, rs.sh_degree, rs.campos, rs.prefiltered,
                            rs.filter_small, rs.filter_large, rs.fade_size, rs.debug, raw=raw)
                continue
            # the scales of one view: one preprocess launch + one depth sort for the whole group
            units = [(settings[i].bg, settings[i].image_height, settings[i].image_width, ust[i]) for i in grp]
            pend = _C.rasterize_gaussians_group_begin(
                units, means3D, empty, max_pixel_sizes, min_pixel_sizes, opacities, occ_multiplier, dc_delta, scales,
                rotations, rs.scale_modifier, empty, base_mask, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, sh,
                rs.sh_degree, rs.campos, rs.prefiltered, rs.filter_small, rs.filter_large, rs.fade_size, rs.debug, raw=raw)
            for i, pf in zip(grp, pend):
                pending[i] = pf
        in_group = {i for grp in ctx.groups if len(grp) > 1 and means3D.size(0) != 0 for i in grp}
        for i, pend in enumerate(pending):
            if i in in_group:
                # a view group's tensors live on the caller's stream (allocated there; us.begin / us.end order the unit
                # streams around the step), and the library remembers each unit's stream: no stream switch, no record_stream
                (R, color, acc_px, depth, radii, pixel_sizes, gb, bb, ib) = pend.finish()
            else:
                with torch.cuda.stream(ust[i]):
                    (R, color, acc_px, depth, radii, pixel_sizes, gb, bb, ib) = pend.finish()
                    for t_ in (color, acc_px, depth, radii, pixel_sizes):
                        t_.record_stream(cur)          # produced on a unit stream, consumed by the caller's stream
            saved.append((R, radii, pixel_sizes, gb, bb, ib))
            colors.append(color)
            aux.append({"acc_pixel_size": acc_px, "depth": depth, "radii": radii, "pixel_sizes": pixel_sizes})
        us.end()
        ctx.settings, ctx.saved_units, ctx.raw = settings, saved, bool(raw)
        ctx.save_for_backward(means3D, sh, scales, rotations, base_mask, max_pixel_sizes, min_pixel_sizes, opacities)
        return tuple(colors)


def _units_backward_impl(ctx, grad_colors):
        """backward of every unit; returns (accumulators dict, per-unit dL_dmeans2D list)"""
        means3D, sh, scales, rotations, base_mask, max_pixel_sizes, min_pixel_sizes, opacities = ctx.saved_tensors
        oraw = opacities if ctx.raw else None       # raw mode: the backward chains sigmoid' and needs the raw opacities
        dev = means3D.device
        P, M = means3D.size(0), sh.size(1)
        us = _streams_for(dev)
        ust = ctx.unit_streams
        cur = torch.cuda.current_stream(dev)
        empty = torch.Tensor([])
        f32 = dict(dtype=torch.float32, device=dev)
        # set_gradient_sink: the units accumulate into the caller's flat buffer
        accum = _sink_for(P, M, (means3D, sh, opacities, scales, rotations))
        if accum is not None:
            for a in accum.values():
                a.zero_()
        else:
            accum = {"mean3D": torch.zeros((P, 3), **f32), "sh": torch.zeros((P, M, 3), **f32),
                     "opacity": torch.zeros((P, 1), **f32), "scale": torch.zeros((P, 3), **f32),
                     "rot": torch.zeros((P, 4), **f32)}
        g2ds = [None] * len(ctx.settings)
        us.begin()
        grouped = set()
        # sliced issue only when the step is ONE view group and its gradients land in the caller's sink
        sl = _backward_slices if (_backward_slices is not None and len(ctx.groups) == 1 and _gradient_sink is not None
                                  and accum[_SINK_KEYS[0]].data_ptr() == _gradient_sink[_SINK_KEYS[0]].data_ptr()) else None
        for grp in ctx.groups:
            # a view group whose units all have a gradient: blend backward per unit (unit streams), then ONE preprocess
            # backward for the group on its first unit's stream (_C.rasterize_gaussians_group_backward_accumulate)
            if len(grp) < 2 or M != 16 or P == 0 or any(grad_colors[i] is None for i in grp):
                continue
            grouped.update(grp)
            rs = ctx.settings[grp[0]]
            # ONE library call for the group: every unit's blend backward on its stream (coarsest first) behind the
            # caller's stream, then the group's preprocess backward on the caller's stream behind all of them
            big = max(grp, key=lambda i: ctx.settings[i].image_height * ctx.settings[i].image_width)
            g2 = _C.rasterize_gaussians_group_backward_full(
                [(ctx.settings[i].bg, ctx.settings[i].image_height, ctx.settings[i].image_width, ctx.saved_units[i][1],
                  ctx.saved_units[i][2], ctx.saved_units[i][3], ctx.saved_units[i][0], ctx.saved_units[i][4],
                  ctx.saved_units[i][5], grad_colors[i], ust[i]) for i in grp],
                means3D, max_pixel_sizes, min_pixel_sizes, scales, rotations, rs.scale_modifier, empty, base_mask,
                rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, sh, rs.sh_degree, rs.campos, rs.filter_small,
                rs.filter_large, rs.debug, accum, want_means2D=True, opacity_raw=oraw,
                stream=ust[big],         # the group's stream (it already waits for the caller's: us.begin())
                slices=sl[0] if sl else None, slice_hook=sl[1] if sl else None)
            for i, t_ in zip(grp, g2):
                g2ds[i] = t_
        for i, (rs, (R, radii, pixel_sizes, gb, bb, ib)) in enumerate(zip(ctx.settings, ctx.saved_units)):
            g = grad_colors[i]
            if g is None or i in grouped:
                continue
            with torch.cuda.stream(ust[i]):
                g2d = _C.rasterize_gaussians_backward_accumulate(
                    rs.bg, means3D, radii, empty, pixel_sizes, max_pixel_sizes, min_pixel_sizes, scales, rotations,
                    rs.scale_modifier, empty, base_mask, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, g, sh,
                    rs.sh_degree, rs.campos, gb, R, bb, ib, rs.filter_small, rs.filter_large, rs.fade_size, rs.debug,
                    accum, want_means2D=True, opacity_raw=oraw)
                g2d.record_stream(cur)
            g2ds[i] = g2d                            # this unit's dL_dmeans2D (per-view densification statistic)
        us.end()
        ctx.saved_units = None                       # release the state buffers
        return accum, g2ds


class _RasterizeUnits(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, sh, opacities, scales, rotations, max_pixel_sizes, min_pixel_sizes, occ_multiplier,
                dc_delta, base_mask, settings, aux, view_ids, raw, *means2D_list):
        return _units_forward_impl(ctx, means3D, sh, opacities, scales, rotations, max_pixel_sizes, min_pixel_sizes,
                                   occ_multiplier, dc_delta, base_mask, settings, aux, view_ids, raw)

    @staticmethod
    def backward(ctx, *grad_colors):
        accum, g2ds = _units_backward_impl(ctx, grad_colors)
        return (accum["mean3D"], accum["sh"], accum["opacity"], accum["scale"], accum["rot"], None, None, None, None,
                None, None, None, None, None, *g2ds)


class _UnitsHandle:
    """what rasterize_gaussians_units_begin hands to rasterize_gaussians_units_backward (the autograd ctx's stand-in)"""
    saved_tensors = ()

    def save_for_backward(self, *tensors):
        self.saved_tensors = tensors


def rasterize_gaussians_units_begin(means3D, shs, opacities, scales, rotations, max_pixel_sizes, min_pixel_sizes,
                                    occ_multiplier, dc_delta, base_mask, settings_list, view_ids=None, raw=False):
    """`rasterize_gaussians_units` without autograd, for a caller that drives the backward itself (msgs_train.MultiScaleStep:
    the loss kernel produces dL_dimage directly, so neither autograd's graph nor its engine is needed on the step's critical
    path).  Returns (units, handle): `units` as rasterize_gaussians_units returns them ("render" does not require grad);
    rasterize_gaussians_units_backward(handle, grads) runs the backward."""
    settings_list = list(settings_list)
    if not settings_list:
        return [], None
    h, aux = _UnitsHandle(), []
    with torch.no_grad():
        colors = _units_forward_impl(h, means3D, shs, opacities, scales, rotations, max_pixel_sizes, min_pixel_sizes,
                                     occ_multiplier, dc_delta, base_mask, settings_list, aux,
                                     list(view_ids) if view_ids is not None else None, bool(raw))
    for a, c in zip(aux, colors):
        a["render"] = c
    return aux, h


def rasterize_gaussians_units_backward(handle, grad_renders):
    """Backward of rasterize_gaussians_units_begin: grad_renders[i] = dL/d(units[i]["render"]) (3,H,W) or None.  Returns
    (grads, dL_dmeans2D): grads = {"mean3D", "sh", "opacity", "scale", "rot"} summed over the units — the gradient sink's
    tensors when one is installed (set_gradient_sink), fresh tensors otherwise; raw-parameter gradients if the forward was
    raw — and every unit's dL_dmeans2D (P,3).  One call per handle."""
    if handle is None:
        return {}, []
    if handle.saved_units is None:
        raise RuntimeError("rasterize_gaussians_units_backward: the handle has already been used")
    with torch.no_grad():
        return _units_backward_impl(handle, tuple(grad_renders))


def rasterize_gaussians_units(means3D, means2D_list, shs, opacities, scales, rotations, max_pixel_sizes,
                              min_pixel_sizes, occ_multiplier, dc_delta, base_mask, settings_list, view_ids=None, raw=False):
    """Render every (view, scale) unit of `settings_list` (GaussianRasterizationSettings each) from the same Gaussians.

    `means2D_list[i]` is unit i's screen-space points tensor ((P,3) zeros with requires_grad, as
    gaussian_renderer/__init__.py:27 creates per render): its .grad receives the unit's dL_dmeans2D, the per-view
    densification statistic of train.py:249-251.  Returns one dict per unit with "render" (3,H,W; differentiable),
    "acc_pixel_size", "depth", "radii" and "pixel_sizes" as `GaussianRasterizer.forward` returns them.  SH +
    scale/rotation inputs only (the training configuration).  Gradients reach means3D, shs, opacities, scales and
    rotations once, summed over the units.

    `view_ids` (optional, one int per unit): consecutive units with the same id are the scales of ONE view (same
    viewmatrix / projmatrix / campos values, FoV and flags; they differ in image size only) and are rendered as a view
    group — one preprocess launch and one depth sort for all of them (_C.rasterize_gaussians_group_begin).  Without it
    units are grouped when their camera tensors are the same storage.  Results are bit-identical either way.

    `raw=True` (msgs_*_raw): `opacities`, `scales`, `rotations` are GaussianModel's RAW parameters
    (_opacity, _scaling, _rotation) and `shs` the (P,16,3) concatenation of _features_dc / _features_rest; sigmoid / exp /
    F.normalize (gaussian_renderer/__init__.py:57-89) run inside the preprocess kernels and the gradients that reach the
    inputs are those of the raw parameters — bit-identical to msgs_activations.activate_gaussians followed by this call."""
    settings_list = list(settings_list)
    means2D_list = list(means2D_list)
    if len(means2D_list) != len(settings_list):
        raise Exception('Please provide one means2D tensor per unit!')
    if not settings_list:
        return []
    aux = []
    colors = _RasterizeUnits.apply(means3D, shs, opacities, scales, rotations, max_pixel_sizes, min_pixel_sizes,
                                   occ_multiplier, dc_delta, base_mask, settings_list, aux,
                                   list(view_ids) if view_ids is not None else None, bool(raw), *means2D_list)
    for a, c in zip(aux, colors):
        a["render"] = c
    return aux
