"""Small end-to-end cases for compute-sanitizer (memcheck / racecheck / initcheck / synccheck; tools/sanitize.sh):
C1-sized forward + backward (plain and MS-GS variant), the precomputed colour + covariance path, a forward-only render,
a 4-scale view-group training step (raw parameters, loss, statistics, Adam) twice, u64 sort.  Prints SANITIZE_CASE_OK."""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "ms-gs_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import diff_gaussian_rasterization as dgr
import msgs_optim, msgs_train
from harness import scenes
from test_gpu_train_step import _setup, _clone_state

dev = torch.device("cuda", 0)
P, W, H = int(os.environ.get("SAN_P", "4096")), 200, 200
tt = lambda a, dt=torch.float32: torch.from_numpy(np.ascontiguousarray(a)).to(device=dev, dtype=dt)
for variant in ("plain", "ms"):
    sc = scenes.make_scene(P, W, H, seed=3)
    if variant == "ms":
        rng = np.random.default_rng(1)
        sc.min_pixel_sizes = np.where(rng.random(P) < 0.5, rng.uniform(0.5, 6.0, P), -1.0).astype(np.float32)
        sc.max_pixel_sizes = np.where(rng.random(P) < 0.15, rng.uniform(4.0, 20.0, P), -1.0).astype(np.float32)
        sc.base_mask = rng.random(P) < 0.05
    cam = scenes.make_camera(W, H, yaw_deg=5.0)
    leaf = {k: tt(getattr(sc, k)).requires_grad_(True) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
    m2 = torch.zeros_like(leaf["means3D"], requires_grad=True)
    rs = dgr.GaussianRasterizationSettings(H, W, cam.tanfovx, cam.tanfovy, tt(sc.bg), 1.0, tt(cam.viewmatrix), tt(cam.projmatrix), 3,
                                           tt(cam.campos), False, False, variant == "ms", variant == "ms", 1.0)
    out = dgr.GaussianRasterizer(rs)(means3D=leaf["means3D"], means2D=m2, opacities=leaf["opacities"], occ_multiplier=tt(sc.occ_multiplier),
                                     dc_delta=tt(sc.dc_delta), shs=leaf["shs"], base_mask=tt(sc.base_mask, torch.bool), scales=leaf["scales"],
                                     rotations=leaf["rotations"], max_pixel_sizes=tt(sc.max_pixel_sizes), min_pixel_sizes=tt(sc.min_pixel_sizes))
    (out[0] * torch.rand_like(out[0])).sum().backward()
    torch.cuda.synchronize()
    print(variant, "visible", int((out[3] > 0).sum()))
# forward only + mark_visible
with torch.no_grad():
    out = dgr.GaussianRasterizer(rs)(means3D=leaf["means3D"], means2D=m2, opacities=leaf["opacities"], occ_multiplier=tt(sc.occ_multiplier),
                                     dc_delta=tt(sc.dc_delta), shs=leaf["shs"], base_mask=tt(sc.base_mask, torch.bool), scales=leaf["scales"],
                                     rotations=leaf["rotations"], max_pixel_sizes=tt(sc.max_pixel_sizes), min_pixel_sizes=tt(sc.min_pixel_sizes))
    vis = dgr.GaussianRasterizer(rs).markVisible(leaf["means3D"])
torch.cuda.synchronize()
# training step on a 4-scale view group, twice (speculative buffer: miss then hit)
Pt = int(os.environ.get("SAN_PT", "6000"))
sc, raw, state, views = _setup(Pt, 160, 112, [1, 2, 4, 8], 1, dev, seed=9)
flat = msgs_optim.FlatGaussians(Pt, dev)
flat.load(raw["xyz"], raw["f_dc"], raw["f_rest"], raw["scaling"], raw["rotation"], raw["opacity"])
tr = msgs_train.MultiScaleStep(flat, _clone_state(state), torch.zeros(3, device=dev), sc.sh_degree)
for _ in range(2):
    loss = tr.step(views)
torch.cuda.synchronize()
print("train step loss", float(loss))
# u64 sort
k = torch.randint(0, 2 ** 44, (5000,), device=dev, dtype=torch.int64)
v = torch.arange(5000, device=dev, dtype=torch.int32)
if hasattr(dgr._C, "sort_pairs"):
    ks, vs = dgr._C.sort_pairs(k, v, 44)
    torch.cuda.synchronize()
    assert bool((ks[1:] >= ks[:-1]).all())
print("SANITIZE_CASE_OK")
