"""Pins the CPU oracle (oracle/msgs_oracle.cpp) against outputs of the reference CUDA rasterizer itself
(tests/golden/*.npz, produced on a B200 by tests/golden/make_golden.py through oracle/_ref).  The reference ships
no tests or golden vectors of its own, so these fixtures ARE the pin.

CPU libm / no-FMA arithmetic differs from the GPU's in the last bits, so: integer outputs get a small mismatch
budget (a radius or a rectangle edge that flips on a rounding boundary), float outputs a tolerance; the list
order is compared where the depth bits agree."""
import hashlib
import os

import numpy as np
import pytest

from harness import oracle_bind, scenes

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = {
    # name: (P, W, H, seed, variant, scale, precomp)   — must match tests/golden/make_golden.py
    "plain_2k": (2048, 96, 64, 101, "A", 1, False),
    "ms_2k_s2": (2048, 96, 64, 102, "B", 2, False),
    "precomp_1k": (1024, 80, 48, 103, "A", 1, True),
}


def _checksum(sc, cam):
    h = hashlib.sha256()
    for a in (sc.means3D, sc.scales, sc.rotations, sc.opacities, sc.shs, sc.max_pixel_sizes, sc.min_pixel_sizes,
              sc.base_mask.astype(np.uint8), cam.viewmatrix, cam.projmatrix, cam.campos):
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()


def _load(name):
    P, W, H, seed, variant, scale, precomp = CASES[name]
    g = np.load(os.path.join(GOLD, name + ".npz"))
    sc = scenes.make_scene(P, W, H, seed)
    cam1 = scenes.make_camera(W, H)
    if variant == "B":
        scenes.apply_ms_state(sc, g["warm_pixel_sizes"], seed)
    cam = cam1.at_scale(scale)
    return g, sc, cam, precomp


@pytest.mark.parametrize("name", list(CASES))
def test_inputs_regenerate_identically(name):
    g, sc, cam, _ = _load(name)
    assert _checksum(sc, cam) == str(g["input_sha256"]), "scene generator drifted from the one the fixtures were made with"


@pytest.mark.parametrize("name", list(CASES))
def test_oracle_forward_matches_reference_cuda(built, name):
    g, sc, cam, precomp = _load(name)
    P = sc.P
    kw = dict(colors_precomp=g["colors_precomp"], cov3D_precomp=g["cov3D_precomp"]) if precomp else {}
    o = oracle_bind.Oracle().forward(sc, cam, bool(g["filter_small"]), bool(g["filter_large"]), **kw)
    budget = 1 + P // 1000      # measured on these fixtures: 0 mismatches
    assert int((o.radii != g["radii"]).sum()) <= budget
    assert int(((o.radii > 0) != (g["radii"] > 0)).sum()) <= 1, "visibility must agree (one boundary flip allowed)"
    assert int((o.buffer("tiles_touched") != g["tiles_touched"]).sum()) <= budget
    assert abs(o.R - int(g["R"])) <= 4 * budget
    both = (o.radii > 0) & (g["radii"] > 0)
    np.testing.assert_allclose(o.pixel_sizes, g["pixel_sizes"], rtol=2e-5, atol=1e-6)   # Q5: also where radii == 0
    np.testing.assert_allclose(o.buffer("depths")[both], g["depths"][both], rtol=1e-6)
    np.testing.assert_allclose(o.buffer("means2D").reshape(P, 2)[both], g["means2D"][both], rtol=0, atol=1e-4)
    np.testing.assert_allclose(o.buffer("conic_opacity").reshape(P, 4)[both], g["conic_opacity"][both], rtol=1e-4, atol=1e-6)
    if o.R == int(g["R"]):
        same = o.buffer("vals_sorted") == g["point_list"].view(np.uint32)
        assert same.mean() > 0.995, "sorted list order differs beyond depth-bit ties"
        assert (o.buffer("ranges").reshape(-1, 2) == g["ranges"].view(np.uint32)).mean() > 0.99
    assert (o.buffer("n_contrib") != g["n_contrib"].view(np.uint32)).mean() < 0.01
    for mine, ref, tol in ((o.color, g["color"], 1e-5), (o.buffer("final_T"), g["final_T"], 1e-5)):
        assert np.quantile(np.abs(mine.reshape(-1) - ref.reshape(-1)), 0.999) < tol
    for mine, ref in ((o.acc_px, g["acc_px"]), (o.depth, g["depth"])):
        err = np.abs(mine - ref) / (np.abs(ref) + 1.0)
        assert np.quantile(err, 0.999) < 1e-5


@pytest.mark.parametrize("name", list(CASES))
@pytest.mark.parametrize("acc64", [False, True])
def test_oracle_backward_matches_reference_cuda(built, name, acc64):
    g, sc, cam, precomp = _load(name)
    kw = dict(colors_precomp=g["colors_precomp"], cov3D_precomp=g["cov3D_precomp"]) if precomp else {}
    o = oracle_bind.Oracle().forward(sc, cam, bool(g["filter_small"]), bool(g["filter_large"]), **kw)
    grads = o.backward(g["dpix"], acc64=acc64)
    for n in ("mean2D", "color", "opacity", "mean3D", "cov3D", "sh", "scale", "rot"):
        ref = g["grad_" + n].astype(np.float64)
        mine = grads[n].reshape(ref.shape).astype(np.float64)
        if ref.size == 0 or not np.any(ref):
            assert not np.any(mine)
            continue
        # Gaussians whose forward differs by a rounding flip (radius / n_contrib) dominate the max error: compare in
        # the l2 sense over the tensor and on the bulk of the elements
        rel = np.linalg.norm(mine - ref) / (np.linalg.norm(ref) + 1e-30)
        assert rel < 1e-5, f"{n}: rel l2 {rel}"      # measured: <= 9e-7
        err = np.abs(mine - ref) / (np.abs(ref) + np.abs(ref).mean() + 1e-30)
        assert np.quantile(err, 0.99) < 1e-5, f"{n}: 99th percentile error {np.quantile(err, 0.99)}"


def test_oracle_mark_visible():
    sc = scenes.make_scene(5000, 200, 200, 9)
    cam = scenes.make_camera(200, 200, yaw_deg=10.0)
    vis = oracle_bind.mark_visible(sc.means3D, cam.viewmatrix, cam.projmatrix)
    z = sc.means3D.astype(np.float64) @ cam.viewmatrix[:3, 2].astype(np.float64) + cam.viewmatrix[3, 2]
    assert int((vis != (z > 0.2)).sum()) <= 2
    assert 0 < vis.sum() < 5000
