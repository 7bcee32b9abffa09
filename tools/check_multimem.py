"""torchrun check (N >= 2 GPUs): the fused backward + multicast reduction gives the same summed gradients as the plain
backward followed by an NCCL all-reduce.  python -m torch.distributed.run --nproc-per-node 2 tools/check_multimem.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "ms-gs_b200"))
import torch
import torch.distributed as dist
import bench
import msgs_sharding

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
cfg = sys.argv[1] if len(sys.argv) > 1 else "C1"
wl = bench.build_workload(cfg, rank, world, dev)
r = bench.MineRunner(wl, dev)
P = wl["P"]
# reference: plain backward per unit, summed locally, NCCL all-reduce
ex = msgs_sharding.GradientExchange(P, dev)
ex.zero()
for g in r.step():
    ex.add(g)
want = {k: v.clone() for k, v in ex.reduce().items()}
# fused: multimem accumulate
mc = msgs_sharding.SymmetricGradientExchange(P, dev)
worst = 0.0
for it in range(3):                      # several steps: the zero / barrier protocol must hold across steps
    mc.begin_step()
    r.step(accum=mc.multicast_ptrs(), multicast=True)
    got = mc.end_step()
    torch.cuda.synchronize()
    for k in want:
        rel = float((got[k].double() - want[k].double()).norm() / (want[k].double().norm() + 1e-30))
        worst = max(worst, rel)
t = torch.tensor([worst], device=dev)
dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    print(f"multimem vs NCCL on {cfg} x{world}: worst rel l2 over fields / steps / ranks = {float(t):.3e}")
    assert float(t) < 1e-4
dist.destroy_process_group()
