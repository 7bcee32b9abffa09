"""Seeded synthetic random-Gaussian scenes of the BASELINE.json shapes.

numpy only (PCG64 streams are identical on every machine), fp32 outputs.  Camera matrices follow the reference's
conventions: world_view_transform = getWorld2View2(R, T).T, full_proj_transform = world_view @ projection.T,
camera_center = inverse(world_view)[3, :3]  (/root/reference/scene/cameras.py:54-57,
/root/reference/utils/graphics_utils.py:38-71), i.e. row-vector matrices that the CUDA side reads column-major.
"""
import math
from dataclasses import dataclass, field

import numpy as np

CONFIGS = {
    # name: (P, W, H)
    "C1": (16384, 200, 200),
    "C2": (1_200_000, 1237, 822),
    "C4": (3_000_000, 1237, 822),
    "C5": (6_000_000, 3840, 2160),
}


def _projection(znear, zfar, tanx, tany):
    # getProjectionMatrix (utils/graphics_utils.py:51-71) restated with tan(fov/2) given directly
    top, right = tany * znear, tanx * znear
    P = np.zeros((4, 4), dtype=np.float32)
    P[0, 0] = 2.0 * znear / (2 * right)
    P[1, 1] = 2.0 * znear / (2 * top)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


@dataclass
class Camera:
    W: int
    H: int
    tanfovx: float
    tanfovy: float
    viewmatrix: np.ndarray      # (4,4) f32, world_view_transform (transposed convention)
    projmatrix: np.ndarray      # (4,4) f32, full_proj_transform
    campos: np.ndarray          # (3,) f32

    def at_scale(self, s):
        """Same view at 1/s resolution: int(W/s) x int(H/s) (utils/camera_utils.py:38-39), FoV unchanged."""
        return Camera(int(self.W / s), int(self.H / s), self.tanfovx, self.tanfovy, self.viewmatrix,
                      self.projmatrix, self.campos)


def make_camera(W, H, yaw_deg=0.0, centre_z=12.75, focal_factor=1.1):
    f = focal_factor * W
    tanx, tany = W / (2.0 * f), H / (2.0 * f)
    a = math.radians(yaw_deg)
    Ry = np.array([[math.cos(a), 0, math.sin(a)], [0, 1, 0], [-math.sin(a), 0, math.cos(a)]], dtype=np.float64)
    centre = np.array([0.0, 0.0, centre_z])
    pos = centre + Ry @ np.array([0.0, 0.0, -centre_z])
    R = Ry                               # camera-to-world rotation, as 3DGS stores it
    T = -(R.T @ pos)
    Rt = np.zeros((4, 4))
    Rt[:3, :3] = R.T
    Rt[:3, 3] = T
    Rt[3, 3] = 1.0
    w2v = np.float32(Rt)                                 # getWorld2View2 with zero translate, unit scale
    view = np.ascontiguousarray(w2v.T)                   # .transpose(0, 1)
    proj = np.ascontiguousarray(_projection(0.01, 100.0, tanx, tany).T)
    full = np.ascontiguousarray((view @ proj).astype(np.float32))
    campos = np.linalg.inv(view.astype(np.float64))[3, :3].astype(np.float32)
    return Camera(W, H, float(tanx), float(tany), view, full, campos)


@dataclass
class Scene:
    P: int
    sh_degree: int
    means3D: np.ndarray         # (P,3)
    scales: np.ndarray          # (P,3)
    rotations: np.ndarray       # (P,4) normalised
    opacities: np.ndarray       # (P,1)
    shs: np.ndarray             # (P,16,3)
    max_pixel_sizes: np.ndarray # (P,)
    min_pixel_sizes: np.ndarray # (P,)
    base_mask: np.ndarray       # (P,) bool
    occ_multiplier: np.ndarray  # (P,4,1) dead input
    dc_delta: np.ndarray        # (P,12,1) dead input
    bg: np.ndarray = field(default_factory=lambda: np.zeros(3, dtype=np.float32))
    scale_modifier: float = 1.0


def make_scene(P, W, H, seed, sh_degree=3, focal_factor=1.1, splat_scale=0.004):
    """Variant A (plain): MS-GS state unset (-1 / False).  W, H only fix the field of view the points fill.
    splat_scale: median world-space sigma per unit depth (0.004 -> sigma ~ 5 px at W = 1237)."""
    rng = np.random.default_rng(seed)
    f = focal_factor * W
    tanx, tany = W / (2.0 * f), H / (2.0 * f)
    z = rng.uniform(0.5, 25.0, P)
    near = rng.random(P) < 0.01                          # ~1 % behind / at the near plane: exercises culling
    z = np.where(near, rng.uniform(-1.0, 0.2, P), z)
    x = rng.uniform(-1.15, 1.15, P) * tanx * np.abs(z)
    y = rng.uniform(-1.15, 1.15, P) * tany * np.abs(z)
    means = np.stack([x, y, z], 1).astype(np.float32)
    zz = np.maximum(np.abs(z), 0.05)
    scales = np.exp(rng.normal(np.log(splat_scale * zz)[:, None], 0.7, (P, 3))).astype(np.float32)
    q = rng.normal(0, 1, (P, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    opac = (1.0 / (1.0 + np.exp(-rng.normal(0, 2.0, (P, 1))))).astype(np.float32)
    shs = np.concatenate([rng.normal(0, 1.0, (P, 1, 3)), rng.normal(0, 0.2, (P, 15, 3))], 1).astype(np.float32)
    return Scene(P, sh_degree, means, scales, q.astype(np.float32), opac, shs,
                 np.full(P, -1.0, np.float32), np.full(P, -1.0, np.float32), np.zeros(P, bool),
                 np.ones((P, 4, 1), np.float32), np.zeros((P, 12, 1), np.float32))


def apply_ms_state(scene, pixel_sizes, seed):
    """Variant B (README configuration): min/max pixel sizes derived from a scale-1 warm-up render's
    pixel_sizes, 5 % base Gaussians.  Used with filter_small = filter_large = True."""
    rng = np.random.default_rng(seed + 7919)
    P = scene.P
    ps = np.asarray(pixel_sizes, dtype=np.float32)
    mn = np.where(ps > 0, ps * rng.uniform(0.5, 1.0, P).astype(np.float32), -1.0).astype(np.float32)
    large = rng.random(P) < 0.15
    mx = np.where(large & (mn > 0), mn * rng.uniform(2.0, 8.0, P).astype(np.float32), -1.0).astype(np.float32)
    scene.min_pixel_sizes = mn
    scene.max_pixel_sizes = mx
    scene.base_mask = rng.random(P) < 0.05
    return scene


MS_SCALES = [1, 2, 4, 8, 16, 32, 64]      # --ms_train_max_scale 6 (train.py:50)
