"""Per-step host/device times of the device-resident loop (diagnostic): python tools/step_jitter.py [steps]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "ms-gs_b200"))
import torch
import bench

N = int(sys.argv[1]) if len(sys.argv) > 1 else 60
dev = torch.device("cuda", 0)
wl = bench.build_workload("C2", 0, 1, dev)
r = bench.MineRunner(wl, dev)
for _ in range(5):
    r.step()
torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(N + 1)]
host = []
ev[0].record()
for i in range(N):
    a = time.perf_counter()
    r.step()
    host.append(time.perf_counter() - a)
    ev[i + 1].record()
torch.cuda.synchronize()
devms = [ev[i].elapsed_time(ev[i + 1]) for i in range(N)]
print("device ms/step:", " ".join(f"{x:.2f}" for x in devms))
print("host   ms/step:", " ".join(f"{x*1e3:.2f}" for x in host))
print("total", sum(devms) / N)
