"""One profiled fwd+bwd step of a bench config between cudaProfilerStart/Stop (for `ncu --profile-from-start off`):
python tools/ncu_step.py [C2] [acc]   (acc: accumulating multi-unit step = view groups)"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

cfg = sys.argv[1] if len(sys.argv) > 1 else "C2"
dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
wl = bench.build_workload(cfg, 0, 1, dev)
r = bench.MineRunner(wl, dev)
acc = None
if len(sys.argv) > 2 and sys.argv[2] == "acc":
    sys.path.insert(0, os.path.join(ROOT, "ms-gs_b200"))
    import msgs_sharding
    if cfg in ("C3", "C4"):      # the MS-GS per-Gaussian state of the multi-scale configs (as bench.py sets it up)
        from harness import scenes
        probe = bench.MineRunner(dict(wl, cams=[wl["probe_cam"]], ms=False, forward_only=True), dev)
        probe.step()
        scenes.apply_ms_state(wl["scene"], probe.pixel_sizes().cpu().numpy(), 1236)
        sc = wl["scene"]
        wl["t"].update(max_pixel_sizes=torch.from_numpy(sc.max_pixel_sizes).to(dev),
                       min_pixel_sizes=torch.from_numpy(sc.min_pixel_sizes).to(dev), base_mask=torch.from_numpy(sc.base_mask).to(dev))
    acc = msgs_sharding.GradientExchange(wl["P"], dev).shaped()
for _ in range(3):
    r.step(accum=acc)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
r.step(accum=acc)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print("profiled one step of", cfg, r.stats())
