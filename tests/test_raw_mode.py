"""Raw-parameter mode of the rasterizer (msgs_*_raw in include/msgs_raster.h): exp / F.normalize /
sigmoid of GaussianModel's getters (gaussian_renderer/__init__.py:57-89) run inside the preprocess kernels and their
backward inside the preprocess backward.  Checked against the composition it replaces — msgs_activations.activate_gaussians
(pinned to the reference's own getters + autograd by tests/test_zz_activations.py) followed by the activated-input call:
forward outputs bit-identical, raw-parameter gradients equal up to atomic-order noise.  CPU: exports + signatures."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_raw_entry_points_are_exported_with_matching_signatures(built):
    import diff_gaussian_rasterization as dgr
    lib = dgr._C.lib
    src = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "msgs_raster.h")).read(), flags=re.S)
    protos = dict(re.findall(r"\b(msgs_[a-z_0-9]+)\s*\(([^()]*)\)\s*;", src))
    for name in ("msgs_forward_begin_raw", "msgs_forward_group_begin_raw", "msgs_backward_raw", "msgs_backward_group_accumulate_raw",
                 "msgs_speculation_stats"):
        assert name in protos and hasattr(lib, name), name
        assert len(getattr(lib, name).argtypes) == len(protos[name].split(",")), name
    assert "msgs_backward_accumulate_multicast" not in protos and not hasattr(lib, "msgs_backward_accumulate_multicast")
    # raw variants take their namesake's arguments (+ the raw opacities in the backward)
    assert len(protos["msgs_forward_begin_raw"].split(",")) == len(protos["msgs_forward_begin"].split(","))
    assert len(protos["msgs_backward_raw"].split(",")) == len(protos["msgs_backward"].split(",")) + 3
    assert len(protos["msgs_backward_group_accumulate_raw"].split(",")) == len(protos["msgs_backward_group_accumulate"].split(",")) + 1


def _raw_scene(P, W, H, seed):
    from harness import scenes
    sc = scenes.make_scene(P, W, H, seed)
    dev = "cuda"
    tt = lambda a, dt=torch.float32: torch.from_numpy(a).to(device=dev, dtype=dt)
    g = torch.Generator().manual_seed(seed)
    raw = dict(xyz=tt(sc.means3D), shs=tt(sc.shs), scaling=torch.log(tt(sc.scales)),
               rotation=tt(sc.rotations) * (0.5 + 2.0 * torch.rand(P, 1, generator=g).to(dev)),      # unnormalised
               opacity=torch.logit(tt(sc.opacities).clamp(1e-4, 1 - 1e-4)))
    return sc, raw


@pytest.mark.gpu
@pytest.mark.parametrize("P,wh,scales_", [(16384, (200, 200), [1, 2, 8]), (300, (64, 48), [1]), (200_000, (640, 400), [1, 2, 4, 16])])
def test_raw_mode_equals_activate_then_rasterize(built, P, wh, scales_):
    import diff_gaussian_rasterization as dgr
    import msgs_activations
    from harness import scenes
    W, H = wh
    sc, raw = _raw_scene(P, W, H, 5)
    cam0 = scenes.make_camera(W, H)
    dev = "cuda"
    tt = lambda a, dt=torch.float32: torch.from_numpy(a).to(device=dev, dtype=dt)
    view, proj, campos, bg = tt(cam0.viewmatrix), tt(cam0.projmatrix), tt(cam0.campos), tt(sc.bg)
    # MS-GS state so that the filters act (pixel sizes of a first raw render)
    def settings(c, fs):
        return dgr.GaussianRasterizationSettings(c.H, c.W, c.tanfovx, c.tanfovy, bg, 1.0, view, proj, sc.sh_degree, campos,
                                                 False, False, fs, fs, 1.0)
    fixed = dict(max_pixel_sizes=tt(sc.max_pixel_sizes), min_pixel_sizes=tt(sc.min_pixel_sizes), occ=tt(sc.occ_multiplier),
                 dc=tt(sc.dc_delta), base=tt(sc.base_mask, torch.bool))
    with torch.no_grad():
        warm = dgr.rasterize_gaussians_units(raw["xyz"], [torch.zeros_like(raw["xyz"])], raw["shs"], raw["opacity"], raw["scaling"],
                                             raw["rotation"], fixed["max_pixel_sizes"], fixed["min_pixel_sizes"], fixed["occ"],
                                             fixed["dc"], fixed["base"], [settings(cam0, False)], raw=True)
    scenes.apply_ms_state(sc, warm[0]["pixel_sizes"].cpu().numpy(), 5)
    fixed.update(max_pixel_sizes=tt(sc.max_pixel_sizes), min_pixel_sizes=tt(sc.min_pixel_sizes), base=tt(sc.base_mask, torch.bool))
    cams = [cam0.at_scale(s) for s in scales_]
    rs = [settings(c, True) for c in cams]
    gen = torch.Generator().manual_seed(17)
    dpix = [torch.randn((3, c.H, c.W), generator=gen).to(dev) for c in cams]
    vids = [0] * len(cams)

    def run(raw_mode):
        leaves = {k: v.clone().requires_grad_(True) for k, v in raw.items()}
        m2 = [torch.zeros_like(leaves["xyz"], requires_grad=True) for _ in cams]
        if raw_mode:
            shs, scales, rots, opac = leaves["shs"], leaves["scaling"], leaves["rotation"], leaves["opacity"]
        else:
            shs, scales, rots, opac = msgs_activations.activate_gaussians(leaves["shs"][:, :1].contiguous(), leaves["shs"][:, 1:].contiguous(),
                                                                          leaves["scaling"], leaves["rotation"], leaves["opacity"])
        out = dgr.rasterize_gaussians_units(leaves["xyz"], m2, shs, opac, scales, rots, fixed["max_pixel_sizes"],
                                            fixed["min_pixel_sizes"], fixed["occ"], fixed["dc"], fixed["base"], rs, view_ids=vids,
                                            raw=raw_mode)
        sum((u["render"] * d).sum() for u, d in zip(out, dpix)).backward()
        torch.cuda.synchronize()
        return out, leaves, m2

    oa, la, ma = run(True)
    ob, lb, mb = run(False)
    for ua, ub in zip(oa, ob):
        for k in ("render", "acc_pixel_size", "depth", "radii", "pixel_sizes"):
            a, b = ua[k].detach(), ub[k].detach()
            assert torch.equal(a, b), f"{k} differs between the raw-parameter path and activate + rasterize"
    assert int((oa[0]["radii"] > 0).sum()) > 0
    for k in raw:
        a, b = la[k].grad.double(), lb[k].grad.double()
        assert float(b.abs().max()) > 0, k
        rel = float((a - b).norm() / (b.norm() + 1e-30))
        # two separate backward passes: the blend backward's reductions arrive in a different order each run (measured on
        # B200 at P = 200 000: 3.4e-5 for scaling, the field with the most cancellation); the arithmetic is the same
        assert rel < 1e-4, f"gradient of raw {k}: rel l2 {rel}"
    for a, b in zip(ma, mb):
        assert float((a.grad - b.grad).abs().max()) <= 1e-3 * float(b.grad.abs().max()) + 1e-12


@pytest.mark.gpu
def test_gradient_sink_survives_a_kept_grad(built):
    """with a sink installed, a parameter whose .grad already lives in the sink (zero_grad(set_to_none=False), or a
    second backward of the step) must not get its gradient doubled: the second backward accumulates as plain autograd does."""
    import diff_gaussian_rasterization as dgr
    import msgs_sharding
    from harness import scenes
    P, W, H = 4096, 128, 96
    sc = scenes.make_scene(P, W, H, 3)
    cam = scenes.make_camera(W, H)
    dev = "cuda"
    tt = lambda a, dt=torch.float32: torch.from_numpy(a).to(device=dev, dtype=dt)
    leaves = {k: tt(getattr(sc, k)).requires_grad_(True) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
    rs = dgr.GaussianRasterizationSettings(H, W, cam.tanfovx, cam.tanfovy, tt(sc.bg), 1.0, tt(cam.viewmatrix), tt(cam.projmatrix),
                                           sc.sh_degree, tt(cam.campos), False, False, False, False, 1.0)
    dpix = torch.randn((3, H, W), generator=torch.Generator().manual_seed(1)).to(dev)

    def backward_once():
        m2 = torch.zeros_like(leaves["means3D"], requires_grad=True)
        color = dgr.GaussianRasterizer(rs)(means3D=leaves["means3D"], means2D=m2, opacities=leaves["opacities"],
                                           occ_multiplier=tt(sc.occ_multiplier), dc_delta=tt(sc.dc_delta), shs=leaves["shs"],
                                           base_mask=tt(sc.base_mask, torch.bool), scales=leaves["scales"], rotations=leaves["rotations"],
                                           max_pixel_sizes=tt(sc.max_pixel_sizes), min_pixel_sizes=tt(sc.min_pixel_sizes))[0]
        (color * dpix).sum().backward()
        torch.cuda.synchronize()

    backward_once()                                     # no sink: the plain gradients
    plain = {k: v.grad.clone() for k, v in leaves.items()}
    for v in leaves.values():
        v.grad = None
    ex = msgs_sharding.GradientExchange(P, dev)
    dgr.set_gradient_sink(ex.shaped())
    try:
        backward_once()                                 # .grad adopts the sink views
        assert dgr.gradient_sink_installed(*(leaves[k] for k in ("means3D", "shs", "opacities", "scales", "rotations")))
        backward_once()                                 # .grad kept: must ACCUMULATE (2x the plain gradient), not 4x / 2x garbage
        for k, v in leaves.items():
            a, b = v.grad.double(), 2.0 * plain[k].double()
            assert float((a - b).norm() / (b.norm() + 1e-30)) < 1e-4, k
        for v in leaves.values():                       # zero_grad(set_to_none=False)
            v.grad.zero_()
        backward_once()
        for k, v in leaves.items():
            a, b = v.grad.double(), plain[k].double()
            assert float((a - b).norm() / (b.norm() + 1e-30)) < 1e-4, k
    finally:
        dgr.set_gradient_sink(None)
