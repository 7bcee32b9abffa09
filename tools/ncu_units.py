"""One C3 training step between cudaProfilerStart / Stop (for `ncu --profile-from-start off -k regex:blend ...`).
python tools/ncu_units.py [C3]"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
cfg = sys.argv[1] if len(sys.argv) > 1 else "C3"
dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
arm = bench.MineTrainArm(cfg, 0, 1, dev)
for _ in range(4):
    arm.step_resident()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
arm.step_resident()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print("done")
