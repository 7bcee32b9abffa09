"""torchrun check + timing (N >= 2 GPUs on one NVSwitch box) of the hand-written two-shot NVLS all-reduce
(msgs_allreduce_multimem, csrc/exchange.cu) against NCCL on the gradient buffer of P Gaussians (59 floats each):
python -m torch.distributed.run --nproc-per-node 8 tools/check_nvls.py [P=1200000] [ctas=0]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "ms-gs_b200"))
import torch
import torch.distributed as dist
import msgs_sharding

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
P = int(sys.argv[1]) if len(sys.argv) > 1 else 1_200_000
ctas = int(sys.argv[2]) if len(sys.argv) > 2 else 0
sym = msgs_sharding.SymmetricGradientExchange(P, dev)
n = sym.flat.numel()
worst = 0.0
for it in range(3):                     # several rounds: the barrier protocol must hold across steps
    g = torch.Generator(device=dev).manual_seed(1000 * it + rank)
    mine = torch.randn(n, generator=g, device=dev)
    want = mine.clone()
    dist.all_reduce(want)
    sym.flat.copy_(mine)
    sym.allreduce(ctas)
    torch.cuda.synchronize()
    worst = max(worst, float((sym.flat.double() - want.double()).abs().max() / want.double().abs().max()))
t = torch.tensor([worst], device=dev)
dist.all_reduce(t, op=dist.ReduceOp.MAX)


def timed(fn, iters=20):
    for _ in range(3):
        fn()
    dist.barrier(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record(); torch.cuda.synchronize()
    ms = torch.tensor([a.elapsed_time(b) / iters], device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms)


buf = torch.randn(n, device=dev)
ms_nccl = timed(lambda: dist.all_reduce(buf))
mb = n * 4 / 1e6
if rank == 0:
    print(f"NVLS two-shot vs NCCL all-reduce, {world} ranks, {mb:.0f} MB fp32: max abs diff / max|x| = {float(t):.2e}; "
          f"NCCL {ms_nccl:.3f} ms ({mb / ms_nccl:.0f} GB/s algbw)")
    assert float(t) < 1e-5
for c in ([ctas] if ctas else [148, 296, 74]):
    ms_nvls = timed(lambda: sym.allreduce(c))
    if rank == 0:
        print(f"  NVLS ctas={c}: {ms_nvls:.3f} ms ({mb / ms_nvls:.0f} GB/s algbw, both barriers included)")
dist.destroy_process_group()
