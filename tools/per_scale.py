"""Per-(scale) device time of the C3 multi-scale step, stage by stage (diagnostic): python tools/per_scale.py [C3|C4]"""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "ms-gs_b200"))
import torch
import bench
from harness import scenes

cfg = sys.argv[1] if len(sys.argv) > 1 else "C3"
dev = torch.device("cuda", 0)
wl = bench.build_workload(cfg, 0, 1, dev)
probe = bench.MineRunner(dict(wl, cams=[wl["probe_cam"]], ms=False, forward_only=True), dev)
probe.step()
scenes.apply_ms_state(wl["scene"], probe.pixel_sizes().cpu().numpy(), 1236)
sc = wl["scene"]
wl["t"].update(max_pixel_sizes=torch.from_numpy(sc.max_pixel_sizes).to(dev), min_pixel_sizes=torch.from_numpy(sc.min_pixel_sizes).to(dev),
               base_mask=torch.from_numpy(sc.base_mask).to(dev))
lib = probe.lib
nst = lib.msgs_profile_num_stages()
names = [lib.msgs_profile_stage_name(i).decode() for i in range(nst)]
print("scale  WxH        visible        R   " + " ".join(f"{n[:10]:>10s}" for n in names) + "      total")
import msgs_sharding
ex = msgs_sharding.GradientExchange(wl["P"], dev)
acc = ex.shaped() if (len(sys.argv) > 2 and sys.argv[2] == "acc") else None
for k, cam in enumerate(wl["cams"][:len(wl["scales"])]):
    r = bench.MineRunner(dict(wl, cams=[cam]), dev)
    for _ in range(3):
        r.step(accum=acc)
    torch.cuda.synchronize()
    lib.msgs_profile_enable(1)
    N = 5
    for _ in range(N):
        r.step(accum=acc)
    torch.cuda.synchronize()
    ms, cnt = (ctypes.c_double * nst)(), (ctypes.c_longlong * nst)()
    lib.msgs_profile_collect(ms, cnt, nst)
    lib.msgs_profile_enable(0)
    st = r.stats()
    print(f"{wl['scales'][k]:5d}  {cam.W:4d}x{cam.H:<4d} {st['visible']:9d} {st['R']:9d}   " +
          " ".join(f"{ms[i] / N * 1e3:10.1f}" for i in range(nst)) + f" {sum(ms) / N * 1e3:10.1f} us")
