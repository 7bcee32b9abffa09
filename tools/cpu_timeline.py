"""Where does the host spend its time in one fwd+bwd step?  python tools/cpu_timeline.py [C2|C3]
Prints mean host microseconds inside the forward call, the backward call and between them, plus the device-timed
step, for the device-resident loop of bench.py (diagnostic; not a benchmark)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "ms-gs_b200"))
import torch
import bench

cfg = sys.argv[1] if len(sys.argv) > 1 else "C2"
dev = torch.device("cuda", 0)
wl = bench.build_workload(cfg, 0, 1, dev)
r = bench.MineRunner(wl, dev)
C, t, e, sc = r.C, wl["t"], r.empty, wl["scene"]
ct, dpix = r.cam_t[0], r.dpix[0]
cam = ct["cam"]

def fwd():
    return C.rasterize_gaussians(t["bg"], t["means3D"], e, t["max_pixel_sizes"], t["min_pixel_sizes"], t["opacities"],
                                 t["occ_multiplier"], t["dc_delta"], t["scales"], t["rotations"], sc.scale_modifier, e,
                                 t["base_mask"], ct["view"], ct["proj"], cam.tanfovx, cam.tanfovy, cam.H, cam.W, t["shs"],
                                 sc.sh_degree, ct["campos"], False, False, False, 0.0, False)

def bwd(f):
    R, color, acc, depth, radii, px, gb, bb, ib = f
    return C.rasterize_gaussians_backward(t["bg"], t["means3D"], radii, e, px, t["max_pixel_sizes"], t["min_pixel_sizes"],
                                          t["scales"], t["rotations"], sc.scale_modifier, e, t["base_mask"], ct["view"],
                                          ct["proj"], cam.tanfovx, cam.tanfovy, dpix, t["shs"], sc.sh_degree, ct["campos"],
                                          gb, R, bb, ib, False, False, 0.0, False)

for mode in ("run-ahead", "sync-each-step", "fwd-only"):
    for _ in range(5):
        f = fwd(); bwd(f)
    torch.cuda.synchronize()
    N = 30
    tf = tb = 0.0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    w0 = time.perf_counter(); e0.record()
    for _ in range(N):
        a = time.perf_counter(); f = fwd(); b = time.perf_counter()
        if mode != "fwd-only":
            g = bwd(f)
        c = time.perf_counter()
        tf += b - a; tb += c - b
        if mode == "sync-each-step":
            torch.cuda.synchronize()
    e1.record(); torch.cuda.synchronize(); w1 = time.perf_counter()
    print(f"{cfg} {mode:15s} host fwd call {tf/N*1e6:8.1f} us  bwd call {tb/N*1e6:8.1f} us  wall/step {(w1-w0)/N*1e6:8.1f} us  device/step {e0.elapsed_time(e1)/N*1e3:8.1f} us")
