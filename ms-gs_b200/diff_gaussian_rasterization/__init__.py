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

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians"]


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
                              geom_buf, binning_buf, img_buf, pixel_sizes, max_pixel_sizes, min_pixel_sizes)
        ctx.mark_non_differentiable(radii)
        return color, acc_pixel_size, depth, radii, pixel_sizes

    @staticmethod
    def backward(ctx, grad_out_color, grad_acc_pixel_size, grad_out_depth, grad_radii, grad_pixel_sizes):
        # Only grad_out_color is consumed, as in the reference (quirk Q8, :126-162).
        rs = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, base_mask, radii, sh, geom_buf, binning_buf,
         img_buf, pixel_sizes, max_pixel_sizes, min_pixel_sizes) = ctx.saved_tensors
        # argument order of _C.rasterize_gaussians_backward (rasterize_points.h:49-80)
        call = (rs.bg, means3D, radii, colors_precomp, pixel_sizes, max_pixel_sizes, min_pixel_sizes, scales,
                rotations, rs.scale_modifier, cov3Ds_precomp, base_mask, rs.viewmatrix, rs.projmatrix, rs.tanfovx,
                rs.tanfovy, grad_out_color, sh, rs.sh_degree, rs.campos, geom_buf, ctx.num_rendered, binning_buf,
                img_buf, rs.filter_small, rs.filter_large, rs.fade_size, rs.debug)
        (g_means2D, g_colors_precomp, g_opacities, g_occ_multiplier, g_dc_delta, g_means3D, g_cov3Ds_precomp, g_sh,
         g_scales, g_rotations) = _guarded(_C.rasterize_gaussians_backward, call, rs.debug, "snapshot_bw.dump",
                                           "backward")
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
