"""CPU restatement (numpy, fp32) of the parameter activations on the caller side of the rasterizer — SURVEY.md §8 row f1.

TEST INFRASTRUCTURE ONLY: imported by tests/ (and nothing in the product path).  Pinned against outputs of the reference
itself: tests/golden/activations_golden.npz is produced by tests/golden/make_activations_golden.py from the reference's
own `GaussianModel` getters (scene/gaussian_model.py:127-155) and their autograd gradients.

    get_scaling   exp(_scaling)                                    scene/gaussian_model.py:127-129 (activation :37)
    get_rotation  torch.nn.functional.normalize(_rotation)         :131-133 (:47)   x / max(||x||_2, 1e-12)
    get_opacity   sigmoid(_opacity)                                :153-155 (:44)
    get_features  cat((_features_dc, _features_rest), dim=1)       :143-150
"""
import numpy as np

EPS = np.float32(1e-12)


def activate(features_dc, features_rest, scaling_raw, rotation_raw, opacity_raw):
    f32 = np.float32
    shs = np.concatenate([features_dc.astype(f32), features_rest.astype(f32)], axis=1)
    scales = np.exp(scaling_raw.astype(f32))
    n = np.sqrt((rotation_raw.astype(f32) ** 2).sum(axis=1, keepdims=True, dtype=f32))
    rotations = rotation_raw / np.maximum(n, EPS)
    opacities = (f32(1) / (f32(1) + np.exp(-opacity_raw.astype(f32)))).astype(f32)
    return shs, scales.astype(f32), rotations.astype(f32), opacities


def activate_backward(scales, rotation_raw, opacities, g_shs, g_scales, g_rotations, g_opacities):
    """Gradients of (features_dc, features_rest, scaling_raw, rotation_raw, opacity_raw); computed in fp64, returned fp32."""
    g_dc, g_rest = g_shs[:, :1, :].copy(), g_shs[:, 1:, :].copy()
    g_sraw = g_scales.astype(np.float64) * scales
    q, g = rotation_raw.astype(np.float64), g_rotations.astype(np.float64)
    n = np.sqrt((q ** 2).sum(axis=1, keepdims=True))
    big = n > float(EPS)
    nn = np.where(big, n, 1.0)
    y = q / nn
    g_rraw = np.where(big, (g - y * (y * g).sum(axis=1, keepdims=True)) / nn, g / float(EPS))
    o = opacities.astype(np.float64)
    g_oraw = g_opacities * o * (1.0 - o)
    return (g_dc.astype(np.float32), g_rest.astype(np.float32), g_sraw.astype(np.float32), g_rraw.astype(np.float32),
            g_oraw.astype(np.float32))
