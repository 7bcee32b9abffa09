"""torch.profiler trace of the sharded C3 step on rank 0 (torchrun, N ranks).  Writes gpurun_out/trace_dist.json"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dev = torch.device("cuda", torch.cuda.current_device())
torch.distributed.init_process_group("nccl", device_id=dev)
arm = bench.MineTrainArm("C3", rank, world, dev)
for _ in range(12):
    arm.step_resident()
torch.cuda.synchronize(); torch.distributed.barrier()
from torch.profiler import profile, ProfilerActivity
if rank in (0, 1, world - 1):
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(4):
            arm.step_resident()
        torch.cuda.synchronize()
    prof.export_chrome_trace(os.path.join(ROOT, "gpurun_out", f"trace_dist_r{rank}.json"))
else:
    for _ in range(4):
        arm.step_resident()
    torch.cuda.synchronize()
torch.distributed.barrier()
torch.distributed.destroy_process_group()
