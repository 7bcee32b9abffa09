"""Host-side timeline of the training step (no profiler): perf_counter stamps around the phases of
MultiScaleStep.render_and_backward, averaged over N steps, next to the GPU time per step.  python tools/host_timeline.py [C3]"""
import os, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
import msgs_train, msgs_loss
import diff_gaussian_rasterization as dgr
cfg = sys.argv[1] if len(sys.argv) > 1 else "C3"
dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
arm = bench.MineTrainArm(cfg, 0, 1, dev)
stamps = {}
def wrap(obj, name, key):
    f = getattr(obj, name)
    def g(*a, **k):
        t0 = time.perf_counter()
        r = f(*a, **k)
        stamps.setdefault(key, []).append(time.perf_counter() - t0)
        return r
    setattr(obj, name, g)
wrap(dgr, "rasterize_gaussians_units", "units_forward")
wrap(msgs_loss, "l1_ssim_loss_batch", "loss_forward")
wrap(torch.Tensor, "backward", "backward_call")
wrap(arm.trainer, "update_statistics", "statistics")
wrap(arm.trainer.adam, "step", "adam")
wrap(arm.trainer, "render_and_backward", "render_and_backward")
wrap(dgr._C, "rasterize_gaussians_group_begin", "  group_begin")
wrap(dgr._C, "rasterize_gaussians_backward_accumulate", "  unit_blend_bwd(x7)")
wrap(dgr._C, "rasterize_gaussians_group_backward_accumulate", "  group_pre_bwd")
orig_finish = dgr._C.PendingForward.finish
def fin(self):
    t0 = time.perf_counter(); r = orig_finish(self); stamps.setdefault("  finish(x7)", []).append(time.perf_counter() - t0); return r
dgr._C.PendingForward.finish = fin
for _ in range(10):
    arm.step_resident()
torch.cuda.synchronize()
stamps.clear()
N = 40
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter(); e0.record()
for _ in range(N):
    arm.step_resident()
e1.record(); t1 = time.perf_counter()
torch.cuda.synchronize()
print(f"host loop {1e3 * (t1 - t0) / N:.3f} ms/step; gpu {e0.elapsed_time(e1) / N:.3f} ms/step")
for k, v in stamps.items():
    print(f"{k:28s} {1e6 * sum(v) / N:9.1f} us/step  ({len(v) // N} calls)")
