"""Generates tests/golden/*.npz: outputs of the UNMODIFIED reference CUDA rasterizer (oracle/_ref, compiled in place
from /root/reference for sm_100a) on small seeded scenes.  Run ON THE GPU BOX:  python tests/golden/make_golden.py
The fixtures pin the CPU oracle (tests/test_oracle_golden.py) and, through it, every CPU-side check.
Inputs are regenerated from the seed by harness.scenes (numpy PCG64: machine-independent); a checksum of the
inputs is stored so that a drift of the generator is detected."""
import hashlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from harness import refcuda_bind, scenes  # noqa: E402

CASES = {
    # name: (P, W, H, seed, variant, scale, precomp)
    "plain_2k": (2048, 96, 64, 101, "A", 1, False),
    "ms_2k_s2": (2048, 96, 64, 102, "B", 2, False),
    "precomp_1k": (1024, 80, 48, 103, "A", 1, True),
}


def input_checksum(sc, cam):
    h = hashlib.sha256()
    for a in (sc.means3D, sc.scales, sc.rotations, sc.opacities, sc.shs, sc.max_pixel_sizes, sc.min_pixel_sizes,
              sc.base_mask.astype(np.uint8), cam.viewmatrix, cam.projmatrix, cam.campos):
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()


def build_case(name):
    P, W, H, seed, variant, scale, precomp = CASES[name]
    sc = scenes.make_scene(P, W, H, seed)
    cam1 = scenes.make_camera(W, H)
    return sc, cam1, cam1.at_scale(scale), variant, seed, precomp


def main():
    out_dir = os.path.dirname(os.path.abspath(__file__))
    for name in CASES:
        sc, cam1, cam, variant, seed, precomp = build_case(name)
        fs = fl = False
        extra = {}
        if variant == "B":
            warm = refcuda_bind.RefRasterizer().forward(refcuda_bind.scene_to_torch(sc, cam1), cam1)
            extra["warm_pixel_sizes"] = warm.pixel_sizes.cpu().numpy()
            scenes.apply_ms_state(sc, extra["warm_pixel_sizes"], seed)
            fs = fl = True
        t = refcuda_bind.scene_to_torch(sc, cam)
        colors = cov = None
        if precomp:
            ref0 = refcuda_bind.RefRasterizer().forward(t, cam)
            cov = ref0.state()["cov3D"].clone()
            cov[ref0.radii <= 0] = torch.tensor([1e-4, 0, 0, 1e-4, 0, 1e-4], device="cuda")
            colors = torch.rand((sc.P, 3), generator=torch.Generator().manual_seed(seed)).cuda()
            extra["colors_precomp"] = colors.cpu().numpy()
            extra["cov3D_precomp"] = cov.cpu().numpy()
        ref = refcuda_bind.RefRasterizer().forward(t, cam, fs, fl, colors_precomp=colors, cov3D_precomp=cov)
        s = ref.state()
        dpix = torch.randn((3, cam.H, cam.W), generator=torch.Generator().manual_seed(seed + 1)).cuda()
        g = ref.backward(dpix)
        vis = (ref.radii > 0)
        d = dict(input_sha256=np.array(input_checksum(sc, cam)), filter_small=fs, filter_large=fl, R=ref.R,
                 radii=ref.radii.cpu().numpy(), pixel_sizes=ref.pixel_sizes.cpu().numpy(),
                 tiles_touched=s["tiles_touched"].cpu().numpy(),
                 depths=torch.where(vis, s["depths"], torch.zeros_like(s["depths"])).cpu().numpy(),
                 means2D=torch.where(vis[:, None], s["means2D"], torch.zeros_like(s["means2D"])).cpu().numpy(),
                 conic_opacity=torch.where(vis[:, None], s["conic_opacity"], torch.zeros_like(s["conic_opacity"])).cpu().numpy(),
                 ranges=s["ranges"].cpu().numpy(), n_contrib=s["n_contrib"].cpu().numpy(),
                 final_T=s["final_T"].cpu().numpy(), color=ref.color.cpu().numpy(), acc_px=ref.acc_px.cpu().numpy(),
                 depth=ref.depth.cpu().numpy(), dpix=dpix.cpu().numpy(), **extra)
        if ref.R > 0:
            d["keys_sorted"] = s["keys_sorted"].cpu().numpy().view(np.uint64)
            d["point_list"] = s["point_list"].cpu().numpy()
        for k, v in g.items():
            d["grad_" + k] = v.cpu().numpy()
        path = os.path.join(out_dir, name + ".npz")
        np.savez_compressed(path, **d)
        print(name, "R", ref.R, "visible", int(vis.sum()), os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
