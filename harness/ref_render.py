"""`gaussian_renderer.render()` (gaussian_renderer/__init__.py:18-128) and the GaussianModel getters it reads
(scene/gaussian_model.py:127-175), restated for the GPU box — /root/reference does not travel there.  TEST INFRASTRUCTURE.

`render()` below makes the calls the reference's makes — the same 15 GaussianRasterizationSettings keywords (:36-52), the same
13 rasterizer keywords (:94-108), None for what is not provided, the three colour / covariance modes selected by `pipe`
(:68-89) — against whatever `diff_gaussian_rasterization` module it is given, and returns the same dict (:112-119).
tests/test_render_glue.py checks, where /root/reference exists, that these keyword sets ARE the reference's (parsed from its
source), so the restatement cannot drift from what the reference calls."""
import math
from types import SimpleNamespace

import torch


def eval_sh(deg, sh, dirs):
    """utils/sh_utils.py:57-112 (degrees 0..3): sh (..., C, (deg + 1)^2), dirs (..., 3) -> (..., C)"""
    C0, C1 = 0.28209479177387814, 0.4886025119029199
    C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396]
    C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658,
          1.445305721320277, -0.5900435899266435]
    res = C0 * sh[..., 0]
    if deg > 0:
        x, y, z = dirs[..., 0:1], dirs[..., 1:2], dirs[..., 2:3]
        res = res - C1 * y * sh[..., 1] + C1 * z * sh[..., 2] - C1 * x * sh[..., 3]
        if deg > 1:
            xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
            res = (res + C2[0] * xy * sh[..., 4] + C2[1] * yz * sh[..., 5] + C2[2] * (2.0 * zz - xx - yy) * sh[..., 6] +
                   C2[3] * xz * sh[..., 7] + C2[4] * (xx - yy) * sh[..., 8])
            if deg > 2:
                res = (res + C3[0] * y * (3 * xx - yy) * sh[..., 9] + C3[1] * xy * z * sh[..., 10] +
                       C3[2] * y * (4 * zz - xx - yy) * sh[..., 11] + C3[3] * z * (2 * zz - 3 * xx - 3 * yy) * sh[..., 12] +
                       C3[4] * x * (4 * zz - xx - yy) * sh[..., 13] + C3[5] * z * (xx - yy) * sh[..., 14] +
                       C3[6] * x * (xx - 3 * yy) * sh[..., 15])
    return res


class Model:
    """The attributes of GaussianModel that render() reads, over raw parameter tensors (leaf tensors with requires_grad)."""

    def __init__(self, xyz, features_dc, features_rest, scaling, rotation, opacity, occ_multiplier, dc_delta, base_mask,
                 max_pixel_sizes, min_pixel_sizes, sh_degree=3):
        self._xyz, self._features_dc, self._features_rest = xyz, features_dc, features_rest
        self._scaling, self._rotation, self._opacity = scaling, rotation, opacity
        self._occ_multiplier, self._dc_delta = occ_multiplier, dc_delta
        self.base_gaussian_mask, self.max_pixel_sizes, self.min_pixel_sizes = base_mask, max_pixel_sizes, min_pixel_sizes
        self.active_sh_degree = self.max_sh_degree = sh_degree

    get_xyz = property(lambda s: s._xyz)
    get_features = property(lambda s: torch.cat((s._features_dc, s._features_rest), dim=1))          # :139-146
    get_scaling = property(lambda s: torch.exp(s._scaling))                                          # :127-129
    get_rotation = property(lambda s: torch.nn.functional.normalize(s._rotation))                    # :131-133
    get_opacity = property(lambda s: torch.sigmoid(s._opacity))                                      # :148-150
    get_occ_multiplier = property(lambda s: s._occ_multiplier)                                       # :152-157 (multi_occ off)
    get_dc_delta = property(lambda s: s._dc_delta)
    get_base_mask = property(lambda s: s.base_gaussian_mask)
    get_max_pixel_sizes = property(lambda s: s.max_pixel_sizes)
    get_min_pixel_sizes = property(lambda s: s.min_pixel_sizes)

    def get_covariance(self, scaling_modifier=1):
        """build_covariance_from_scaling_rotation (:33-37) with utils/general_utils.py:64-110"""
        s, q = scaling_modifier * self.get_scaling, self._rotation
        q = q / torch.sqrt((q * q).sum(dim=1))[:, None]
        r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
        R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                         2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                         2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], dim=1).view(-1, 3, 3)
        L = R @ torch.diag_embed(s)
        cov = L @ L.transpose(1, 2)
        return torch.stack([cov[:, 0, 0], cov[:, 0, 1], cov[:, 0, 2], cov[:, 1, 1], cov[:, 1, 2], cov[:, 2, 2]], dim=1)


def camera(cam, device):
    """scene/cameras.py attributes render() reads, from a harness.scenes.Camera"""
    tt = lambda a: torch.from_numpy(a).to(device)
    return SimpleNamespace(FoVx=2.0 * math.atan(cam.tanfovx), FoVy=2.0 * math.atan(cam.tanfovy), image_height=cam.H,
                           image_width=cam.W, world_view_transform=tt(cam.viewmatrix), full_proj_transform=tt(cam.projmatrix),
                           camera_center=tt(cam.campos))


def render(dgr, viewpoint_camera, pc, pipe, bg_color, scaling_modifier=1.0, override_color=None, filter_small=False,
           filter_large=False, fade_size=1.0):
    screenspace_points = torch.zeros_like(pc.get_xyz, dtype=pc.get_xyz.dtype, requires_grad=True, device=pc.get_xyz.device) + 0
    try:
        screenspace_points.retain_grad()
    except Exception:
        pass
    raster_settings = dgr.GaussianRasterizationSettings(
        image_height=int(viewpoint_camera.image_height), image_width=int(viewpoint_camera.image_width),
        tanfovx=math.tan(viewpoint_camera.FoVx * 0.5), tanfovy=math.tan(viewpoint_camera.FoVy * 0.5), bg=bg_color,
        scale_modifier=scaling_modifier, viewmatrix=viewpoint_camera.world_view_transform,
        projmatrix=viewpoint_camera.full_proj_transform, sh_degree=pc.active_sh_degree,
        campos=viewpoint_camera.camera_center, prefiltered=False, debug=pipe.debug, filter_small=filter_small,
        filter_large=filter_large, fade_size=fade_size)
    rasterizer = dgr.GaussianRasterizer(raster_settings=raster_settings)
    scales = rotations = cov3D_precomp = None
    if pipe.compute_cov3D_python:
        cov3D_precomp = pc.get_covariance(scaling_modifier)
    else:
        scales, rotations = pc.get_scaling, pc.get_rotation
    shs = colors_precomp = None
    if override_color is None:
        if pipe.convert_SHs_python:
            shs_view = pc.get_features.transpose(1, 2).view(-1, 3, (pc.max_sh_degree + 1) ** 2)
            dir_pp = pc.get_xyz - viewpoint_camera.camera_center.repeat(pc.get_features.shape[0], 1)
            sh2rgb = eval_sh(pc.active_sh_degree, shs_view, dir_pp / dir_pp.norm(dim=1, keepdim=True))
            colors_precomp = torch.clamp_min(sh2rgb + 0.5, 0.0)
        else:
            shs = pc.get_features
    else:
        colors_precomp = override_color
    rendered_image, acc_pixel_size, depth, radii, pixel_sizes = rasterizer(
        means3D=pc.get_xyz, means2D=screenspace_points, shs=shs, colors_precomp=colors_precomp,
        max_pixel_sizes=pc.get_max_pixel_sizes, min_pixel_sizes=pc.get_min_pixel_sizes, opacities=pc.get_opacity,
        occ_multiplier=pc.get_occ_multiplier, dc_delta=pc.get_dc_delta, scales=scales, rotations=rotations,
        cov3D_precomp=cov3D_precomp, base_mask=pc.get_base_mask)
    return {"render": rendered_image, "acc_pixel_size": acc_pixel_size, "depth": depth, "viewspace_points": screenspace_points,
            "visibility_filter": radii > 0, "radii": radii, "pixel_sizes": pixel_sizes}
