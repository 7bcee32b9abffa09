"""Plain data types of the training step (no CUDA library behind them): the step driver (msgs_train.py) and the test /
bench harness around the reference both describe a step with these."""
from dataclasses import dataclass
from typing import List

import torch

# reference learning rates (arguments/__init__.py OptimizationParams): position_lr_init * spatial_lr_scale, feature_lr,
# feature_lr / 20, opacity_lr, scaling_lr, rotation_lr
DEFAULT_LRS = {"xyz": 0.00016, "f_dc": 0.0025, "f_rest": 0.0025 / 20.0, "opacity": 0.05, "scaling": 0.005, "rotation": 0.001}


@dataclass
class TrainView:
    """One training camera and its ground-truth images at the scales rendered this step (scene/cameras.py; the per-scale
    camera stacks of scene/__init__.py: image i is the view at 1 / 2**levels[i] resolution, utils/camera_utils.py:38-39)."""
    viewmatrix: torch.Tensor          # (4,4) world_view_transform, device
    projmatrix: torch.Tensor          # (4,4) full_proj_transform, device
    campos: torch.Tensor              # (3,) camera_center, device
    tanfovx: float
    tanfovy: float
    sizes: List[tuple]                # [(H, W)] per scale
    gts: List[torch.Tensor]           # [(3,H,W)] device, ground truth per scale
    levels: List[int]                 # reso_idx per scale (0 = finest)
    view_id: int = 0


@dataclass
class MSState:
    """Per-Gaussian MS-GS bookkeeping (scene/gaussian_model.py:60-75, 662-705), updated in place by the step."""
    target_reso_lvl: torch.Tensor     # (P,) int64
    max_pixel_sizes: torch.Tensor     # (P,) f32
    min_pixel_sizes: torch.Tensor     # (P,) f32
    base_gaussian_mask: torch.Tensor  # (P,) bool
    max_radii2D: torch.Tensor         # (P,) f32
    xyz_gradient_accum: torch.Tensor  # (P,L,1) f32
    denom: torch.Tensor               # (P,L,1) f32
    occ_multiplier: torch.Tensor      # (P,4,1) dead input of the rasterizer (quirk Q9)
    dc_delta: torch.Tensor            # (P,12,1) dead input
    reso_lvls: int = 1

    @staticmethod
    def fresh(P: int, L: int, device, max_pixel_sizes=None, min_pixel_sizes=None, base_mask=None, target=None):
        f = dict(dtype=torch.float32, device=device)
        return MSState(
            target_reso_lvl=(torch.zeros(P, dtype=torch.int64, device=device) if target is None else target),
            max_pixel_sizes=(torch.full((P,), -1.0, **f) if max_pixel_sizes is None else max_pixel_sizes),
            min_pixel_sizes=(torch.full((P,), -1.0, **f) if min_pixel_sizes is None else min_pixel_sizes),
            base_gaussian_mask=(torch.zeros(P, dtype=torch.bool, device=device) if base_mask is None else base_mask),
            max_radii2D=torch.zeros(P, **f), xyz_gradient_accum=torch.zeros((P, L, 1), **f),
            denom=torch.zeros((P, L, 1), **f), occ_multiplier=torch.ones((P, 4, 1), **f),
            dc_delta=torch.zeros((P, 12, 1), **f), reso_lvls=L)
