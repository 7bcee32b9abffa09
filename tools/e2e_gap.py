"""Where does the end-to-end step lose time against the device-resident one?  python tools/e2e_gap.py [C2]
Variants (CUDA-event timed, 30 steps each): A device-resident _C calls; B autograd API + (color*dpix).sum();
C autograd API + L2 loss against a device target; D = C + pinned-host target H2D on a copy stream (bench.py's e2e leg)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "ms-gs_b200"))
import torch
import bench

cfg = sys.argv[1] if len(sys.argv) > 1 else "C2"
dev = torch.device("cuda", 0)
wl = bench.build_workload(cfg, 0, 1, dev)
r = bench.MineRunner(wl, dev)
dgr = r.dgr
t, sc = wl["t"], wl["scene"]
ct, dpix = r.cam_t[0], r.dpix[0]
cam = ct["cam"]
params = {k: t[k].clone().requires_grad_(True) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
means2D = torch.zeros_like(params["means3D"], requires_grad=True)
rs = dgr.GaussianRasterizationSettings(cam.H, cam.W, cam.tanfovx, cam.tanfovy, t["bg"], sc.scale_modifier, ct["view"],
                                       ct["proj"], sc.sh_degree, ct["campos"], False, False, False, False, 0.0)
tgt_host = torch.rand((3, cam.H, cam.W)).pin_memory()
tgt = tgt_host.to(dev)
copy_stream = torch.cuda.Stream(device=dev)


def render():
    for p in params.values():
        p.grad = None
    means2D.grad = None
    return dgr.GaussianRasterizer(rs)(means3D=params["means3D"], means2D=means2D, opacities=params["opacities"],
                                      occ_multiplier=t["occ_multiplier"], dc_delta=t["dc_delta"], shs=params["shs"],
                                      base_mask=t["base_mask"], scales=params["scales"], rotations=params["rotations"],
                                      max_pixel_sizes=t["max_pixel_sizes"], min_pixel_sizes=t["min_pixel_sizes"])[0]


def A():
    r.step()


def B():
    (render() * dpix).sum().backward()


def C():
    ((render() - tgt).square().sum() * 0.5).backward()


def D():
    main = torch.cuda.current_stream()
    with torch.cuda.stream(copy_stream):
        copy_stream.wait_event(D.consumed) if D.consumed is not None else None
        tgt.copy_(tgt_host, non_blocking=True)
        copied = copy_stream.record_event()
    color = render()
    main.wait_event(copied)
    ((color - tgt).square().sum() * 0.5).backward()
    D.consumed = main.record_event()


D.consumed = None
for name, fn in (("A device-resident", A), ("B autograd + dot loss", B), ("C autograd + L2 loss", C), ("D + target H2D", D)):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    N = 30
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    w0 = time.perf_counter(); e0.record()
    for _ in range(N):
        fn()
    e1.record(); torch.cuda.synchronize(); w1 = time.perf_counter()
    print(f"{cfg} {name:24s} device/step {e0.elapsed_time(e1) / N * 1e3:8.1f} us   wall/step {(w1 - w0) / N * 1e6:8.1f} us")
