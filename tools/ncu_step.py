"""One profiled fwd+bwd step of a bench config between cudaProfilerStart/Stop (for `ncu --profile-from-start off`):
python tools/ncu_step.py [C2]"""
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
for _ in range(3):
    r.step()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
r.step()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print("profiled one step of", cfg, r.stats())
