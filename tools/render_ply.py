"""Headless viewer path (viewer.py:37-72 / render.py, BASELINE configs[4]) end to end: MS-GS model file -> msgs_ply loader ->
activations -> forward-only rasterizer -> image file (binary PPM; there is no window system on the box).
python tools/render_ply.py [model.ply] [out.ppm] [W H]     — without a model file a seeded synthetic one is written first."""
import os, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "ms-gs_b200"))
import numpy as np
import torch
import msgs_ply
from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
from harness import scenes, ply_io

path = sys.argv[1] if len(sys.argv) > 1 and sys.argv[1] != "-" else None
out = sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "gpurun_out", "render_ply.ppm")
W, H = (int(sys.argv[3]), int(sys.argv[4])) if len(sys.argv) > 4 else (1237, 822)
dev = torch.device("cuda", 0)
tmp = None
if path is None:
    # a synthetic trained model: the bench scene's Gaussians stored as RAW parameters (log scale, logit opacity)
    P = 200_000
    sc = scenes.make_scene(P, W, H, 11)
    m = ply_io.random_model(P, 4, 4, seed=11)
    m["xyz"] = sc.means3D.astype(np.float32)
    m["scaling"] = np.log(sc.scales).astype(np.float32)
    op = np.clip(sc.opacities.reshape(P, 1), 1e-4, 1 - 1e-4)
    m["opacity"] = np.log(op / (1 - op)).astype(np.float32)
    m["rotation"] = sc.rotations.astype(np.float32)
    m["features_dc"], m["features_rest"] = sc.shs[:, :1].copy(), sc.shs[:, 1:].copy()
    m["occ_multiplier"][:] = 1.0
    m["dc_delta"][:] = 0.0
    m["base_gaussian_mask"][:] = False
    m["max_pixel_sizes"][:] = -1.0
    m["min_pixel_sizes"][:] = -1.0
    tmp = tempfile.TemporaryDirectory()
    path = os.path.join(tmp.name, "point_cloud.ply")
    ply_io.write_ms_ply(path, m)
t0 = time.perf_counter()
t, info = msgs_ply.load_ply(path, device=dev)
torch.cuda.synchronize()
t1 = time.perf_counter()
cam = scenes.make_camera(W, H)
tt = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
settings = GaussianRasterizationSettings(H, W, cam.tanfovx, cam.tanfovy, torch.zeros(3, device=dev), 1.0, tt(cam.viewmatrix),
                                         tt(cam.projmatrix), 3 if info["sh_coeffs"] == 16 else 0, tt(cam.campos), False, False,
                                         True, True, 1.0)
with torch.no_grad():
    color, acc_px, depth, radii, px = GaussianRasterizer(settings)(
        means3D=t["means3D"], means2D=torch.zeros_like(t["means3D"]), shs=t["shs"], opacities=torch.sigmoid(t["opacity"]),
        scales=torch.exp(t["scales"]), rotations=torch.nn.functional.normalize(t["rotations"]),
        occ_multiplier=t["occ_multiplier"], dc_delta=t["dc_delta"], base_mask=t["base_mask"],
        max_pixel_sizes=t["max_pixel_sizes"], min_pixel_sizes=t["min_pixel_sizes"])
    img = (color.clamp(0, 1) * 255).to(torch.uint8).permute(1, 2, 0).contiguous().cpu().numpy()
os.makedirs(os.path.dirname(out), exist_ok=True)
with open(out, "wb") as f:
    f.write(b"P6\n%d %d\n255\n" % (W, H))
    f.write(img.tobytes())
print(f"{info['num_points']} Gaussians loaded in {(t1 - t0) * 1e3:.1f} ms ({info['data_bytes'] / 1e6:.0f} MB file), "
      f"visible {(radii > 0).sum().item()}, mean colour {img.mean():.1f}/255 -> {out}")
