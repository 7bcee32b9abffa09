"""Live / reserved device memory around multi-unit steps (diagnostic): python tools/mem_probe.py [C4]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "ms-gs_b200"))
import torch
import bench
import msgs_sharding

cfg = sys.argv[1] if len(sys.argv) > 1 else "C4"
dev = torch.device("cuda", 0)
wl = bench.build_workload(cfg, 0, 1, dev)
r = bench.MineRunner(wl, dev)
ex = msgs_sharding.GradientExchange(wl["P"], dev)
gb = lambda x: round(x / 2**30, 2)


def show(tag):
    torch.cuda.synchronize()
    print(tag, "allocated", gb(torch.cuda.memory_allocated()), "GiB  reserved", gb(torch.cuda.memory_reserved()),
          "GiB  peak allocated", gb(torch.cuda.max_memory_allocated()), flush=True)


show("start")
for i in range(3):
    ex.zero()
    r.step(accum=ex.shaped())
    show(f"value step {i}")
dgr = r.dgr
t, sc = wl["t"], wl["scene"]
params = {k: t[k].clone().requires_grad_(True) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
m2 = [torch.zeros_like(params["means3D"], requires_grad=True) for _ in wl["cams"]]
rs = [dgr.GaussianRasterizationSettings(c.H, c.W, c.tanfovx, c.tanfovy, t["bg"], sc.scale_modifier,
                                        torch.from_numpy(c.viewmatrix).to(dev), torch.from_numpy(c.projmatrix).to(dev),
                                        sc.sh_degree, torch.from_numpy(c.campos).to(dev), False, False, True, True, 0.0)
      for c in wl["cams"]]
show("e2e setup")
for i in range(3):
    for p in params.values():
        p.grad = None
    units = dgr.rasterize_gaussians_units(params["means3D"], m2, params["shs"], params["opacities"], params["scales"],
                                          params["rotations"], t["max_pixel_sizes"], t["min_pixel_sizes"],
                                          t["occ_multiplier"], t["dc_delta"], t["base_mask"], rs)
    show(f"e2e fwd {i}")
    sum(u["render"].sum() for u in units).backward()
    del units
    show(f"e2e bwd {i}")
