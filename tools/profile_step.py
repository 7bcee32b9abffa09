"""Host / device timeline of the C3 training step (GPU): python cProfile of the host side of 30 steps + torch.profiler
kernel table of 3 steps.  python tools/profile_step.py [C3|C4]"""
import cProfile, io, os, pstats, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
cfg = sys.argv[1] if len(sys.argv) > 1 else "C3"
dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
arm = bench.MineTrainArm(cfg, 0, 1, dev)
for _ in range(10):
    arm.step_resident()
torch.cuda.synchronize()
# (1) host enqueue time vs GPU time
N = 30
t0 = time.perf_counter()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(N):
    arm.step_resident()
t1 = time.perf_counter()
e1.record()
torch.cuda.synchronize()
t2 = time.perf_counter()
print(f"host loop {1e3 * (t1 - t0) / N:.3f} ms/step, until idle {1e3 * (t2 - t0) / N:.3f} ms/step, gpu events {e0.elapsed_time(e1) / N:.3f} ms/step")
# (2) python profile of the host side
pr = cProfile.Profile()
pr.enable()
for _ in range(N):
    arm.step_resident()
pr.disable()
torch.cuda.synchronize()
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(45)
print(s.getvalue()[:9000])
# (3) kernels
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3):
        arm.step_resident()
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=40, max_name_column_width=70))
try:
    prof.export_chrome_trace(os.path.join(ROOT, "gpurun_out", f"trace_{cfg}.json"))
except Exception as ex:
    print("trace export failed", ex)
