"""CPU restatement (numpy, fp32) of the per-step statistics updates of the MS-GS training loop — SURVEY.md §8 row f2.

TEST INFRASTRUCTURE ONLY.  Pinned against outputs of the reference itself: tests/golden/stats_golden.npz is produced by
tests/golden/make_stats_golden.py from the reference's own `GaussianModel.update_pixel_sizes` /
`add_densification_stats` / `update_base_gaussian_mask` (scene/gaussian_model.py:662-705) and train.py:249."""
import numpy as np


def update_stats(L, lvl, what, radii, px, target, g2d, max_px, min_px, max_r, accum, denom, base):
    """In place on the state arrays (max_px, min_px, max_r (P); accum, denom (P,L,1); base (P) bool)."""
    f32 = np.float32
    vis = radii > 0
    if what & 4:                                                   # gaussian_model.py:703-705
        base |= vis
    if what & 1:                                                   # gaussian_model.py:662-686
        m = vis & (target == lvl)
        if lvl > 0:
            max_px[m] = np.maximum(max_px[m] * f32(0.95), px[m])
        if lvl < L - 1:
            cur = np.maximum(min_px[m] * f32(1.05), f32(-1))
            p = px[m]
            min_px[m] = np.where(cur < 0, np.where(p > 0, p, cur), np.where(p > 0, np.minimum(cur, p), cur))
    if what & 2:                                                   # train.py:249; gaussian_model.py:698-701
        max_r[vis] = np.maximum(max_r[vis], radii[vis].astype(f32))
        n = np.sqrt(g2d[vis, 0] * g2d[vis, 0] + g2d[vis, 1] * g2d[vis, 1]).astype(f32)
        accum[vis, lvl, 0] += n
        denom[vis, lvl, 0] += f32(1)
