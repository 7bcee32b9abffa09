"""Diagnostic run on a B200: product (libmsgs_raster.so) vs the in-place reference CUDA build, stage by stage.
Prints a JSON report (per-stage mismatch counts and error norms) for looking INTO a parity failure; the pass / fail bars live in
tests/test_gpu_parity.py.  Usage: python tools/parity_report.py [C1|C2|...] [--variant A|B] [--scale s]"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from harness import scenes, refcuda_bind, msgs_state  # noqa: E402


def cmp_exact(name, a, b, rep, mask=None):
    a, b = a.cpu(), b.cpu()
    if mask is not None:
        a, b = a[mask], b[mask]
    if a.dtype.is_floating_point:
        neq = (a.view(torch.int32) != b.view(torch.int32))
        both_nan = torch.isnan(a) & torch.isnan(b)
        neq = neq & ~both_nan
    else:
        neq = a != b
    n = int(neq.sum())
    rep[name] = {"mismatch": n, "of": a.numel()}
    if n and a.dtype.is_floating_point:
        d = (a.double() - b.double()).abs()
        rep[name]["max_abs"] = float(d.max())
        rep[name]["max_rel"] = float((d / (b.double().abs() + 1e-30))[neq].max())
    return n


def cmp_close(name, a, b, rep):
    a, b = a.double().cpu(), b.double().cpu()
    d = (a - b).abs()
    scale = b.abs().mean() + 1e-30
    rep[name] = {"max_abs": float(d.max()), "rel_l2": float(d.norm() / (b.norm() + 1e-30)),
                 "max_rel_scaled": float((d / (b.abs() + scale)).max()), "ref_mean_abs": float(scale)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("config", nargs="?", default="C1")
    ap.add_argument("--variant", default="A")
    ap.add_argument("--scale", type=int, default=1)
    ap.add_argument("--seed", type=int, default=1234)
    ap.add_argument("--P", type=int, default=0)
    ap.add_argument("--timing", type=int, default=0)
    args = ap.parse_args()
    P, W, H = scenes.CONFIGS[args.config]
    if args.P:
        P = args.P
    dgr = msgs_state.import_product()
    C = dgr._C
    sc = scenes.make_scene(P, W, H, args.seed)
    cam1 = scenes.make_camera(W, H)
    rep = {"config": args.config, "P": P, "variant": args.variant, "scale": args.scale}
    fs = fl = False
    if args.variant == "B":
        t1 = refcuda_bind.scene_to_torch(sc, cam1)
        warm = refcuda_bind.RefRasterizer().forward(t1, cam1)
        scenes.apply_ms_state(sc, warm.pixel_sizes.cpu().numpy(), args.seed)
        fs = fl = True
    cam = cam1.at_scale(args.scale)
    W, H = cam.W, cam.H
    t = refcuda_bind.scene_to_torch(sc, cam)
    empty = torch.empty(0)

    ref = refcuda_bind.RefRasterizer().forward(t, cam, fs, fl)
    rs = ref.state()
    out = C.rasterize_gaussians(t["bg"], t["means3D"], empty, t["max_pixel_sizes"], t["min_pixel_sizes"],
                                t["opacities"], t["occ_multiplier"], t["dc_delta"], t["scales"], t["rotations"],
                                sc.scale_modifier, empty, t["base_mask"], t["viewmatrix"], t["projmatrix"],
                                cam.tanfovx, cam.tanfovy, H, W, t["shs"], sc.sh_degree, t["campos"], False, fs, fl,
                                1.0, False)
    torch.cuda.synchronize()
    R, color, acc_px, depth, radii, pixel_sizes, gb, bb, ib = out
    rep["R"] = {"mine": R, "ref": ref.R}
    ms = msgs_state.decode(gb, bb, ib, P, R, W, H)
    vis = (ref.radii > 0).cpu()
    rep["visible"] = int(vis.sum())
    cmp_exact("radii", radii, ref.radii, rep)
    cmp_exact("pixel_sizes", pixel_sizes, ref.pixel_sizes, rep)
    cmp_exact("tiles_touched", ms["tiles_touched"], rs["tiles_touched"], rep)
    cmp_exact("depths", ms["depths"], rs["depths"], rep, vis)
    cmp_exact("means2D", ms["means2D"], rs["means2D"], rep, vis)
    cmp_exact("conic_opacity", ms["conic_opacity"], rs["conic_opacity"], rep, vis)
    cmp_exact("rgb", ms["rgb"], rs["rgb"], rep, vis)
    cmp_exact("cov3D", ms["cov3D"], rs["cov3D"], rep, vis)
    if R == ref.R and R > 0:
        cmp_exact("keys_sorted", ms["keys_sorted"], rs["keys_sorted"], rep)
        cmp_exact("point_list", ms["point_list"], rs["point_list"], rep)
    cmp_exact("ranges", ms["ranges"], rs["ranges"], rep)
    cmp_exact("n_contrib", ms["n_contrib"], rs["n_contrib"], rep)
    cmp_exact("final_T", ms["final_T"], rs["final_T"], rep)
    cmp_exact("color", color, ref.color, rep)
    cmp_exact("acc_px", acc_px, ref.acc_px, rep)
    cmp_exact("depth_img", depth, ref.depth, rep)
    rep["sum_contrib"] = int(rs["n_contrib"].long().sum())

    g = torch.Generator(device="cpu").manual_seed(99)
    dpix = torch.randn((3, H, W), generator=g).cuda()
    gref = {k: v.clone() for k, v in ref.backward(dpix).items()}
    gref2 = {k: v.clone() for k, v in ref.backward(dpix).items()}     # atomic-order noise floor of the reference
    mine = C.rasterize_gaussians_backward(t["bg"], t["means3D"], radii, empty, pixel_sizes, t["max_pixel_sizes"],
                                          t["min_pixel_sizes"], t["scales"], t["rotations"], sc.scale_modifier, empty,
                                          t["base_mask"], t["viewmatrix"], t["projmatrix"], cam.tanfovx, cam.tanfovy,
                                          dpix, t["shs"], sc.sh_degree, t["campos"], gb, R, bb, ib, fs, fl, 1.0, False)
    torch.cuda.synchronize()
    names = ["mean2D", "color", "opacity", "occ", "dc", "mean3D", "cov3D", "sh", "scale", "rot"]
    rep["grads"], rep["grads_ref_vs_ref"] = {}, {}
    for n, m in zip(names, mine):
        cmp_close(n, m.reshape(gref[n].shape), gref[n], rep["grads"])
        cmp_close(n, gref2[n], gref[n], rep["grads_ref_vs_ref"])

    if args.timing:
        def timeit(fn, n):
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(n):
                fn()
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1) / n

        def my_fwd():
            return C.rasterize_gaussians(t["bg"], t["means3D"], empty, t["max_pixel_sizes"], t["min_pixel_sizes"],
                                         t["opacities"], t["occ_multiplier"], t["dc_delta"], t["scales"],
                                         t["rotations"], sc.scale_modifier, empty, t["base_mask"], t["viewmatrix"],
                                         t["projmatrix"], cam.tanfovx, cam.tanfovy, H, W, t["shs"], sc.sh_degree,
                                         t["campos"], False, fs, fl, 1.0, False)

        def my_bwd():
            return C.rasterize_gaussians_backward(t["bg"], t["means3D"], radii, empty, pixel_sizes,
                                                  t["max_pixel_sizes"], t["min_pixel_sizes"], t["scales"],
                                                  t["rotations"], sc.scale_modifier, empty, t["base_mask"],
                                                  t["viewmatrix"], t["projmatrix"], cam.tanfovx, cam.tanfovy, dpix,
                                                  t["shs"], sc.sh_degree, t["campos"], gb, R, bb, ib, fs, fl, 1.0,
                                                  False)
        n = args.timing
        rep["ms"] = {"mine_fwd": timeit(my_fwd, n), "mine_bwd": timeit(my_bwd, n),
                     "ref_fwd": timeit(lambda: ref.forward(t, cam, fs, fl), n),
                     "ref_bwd": timeit(lambda: ref.backward(dpix), n)}
    print(json.dumps(rep, indent=1))
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", f"parity_{args.config}_{args.variant}_s{args.scale}.json"), "w") as f:
        json.dump(rep, f, indent=1)


if __name__ == "__main__":
    main()
