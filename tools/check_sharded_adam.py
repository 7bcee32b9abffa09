"""2+ ranks under torchrun (one GPU each): the sharded optimiser step — reduce-scatter + Adam + all-gather as ONE NVLS kernel
(msgs_adam_step_multimem) — against the single-GPU fused step (msgs_adam_step) on the gradients summed with NCCL, and the
typed-lane statistics all-reduce (sum / max / min) against torch.distributed.all_reduce.  Prints SHARDED_ADAM_OK (rank 0).
  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/check_sharded_adam.py [P]"""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "ms-gs_b200"))
import msgs_optim  # noqa: E402


def main():
    P = int(sys.argv[1]) if len(sys.argv) > 1 else 300_001
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    dev = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("nccl", device_id=dev)
    g = torch.Generator().manual_seed(11)                       # identical initial parameters on every rank
    init = torch.randn(msgs_optim.FlatLayout(P).total, generator=g).to(dev)
    sharded = msgs_optim.FlatGaussians(P, dev, symmetric=True)
    single = msgs_optim.FlatGaussians(P, dev)
    sharded.params.copy_(init)
    single.params.copy_(init)
    ad_s = msgs_optim.FusedAdam(sharded, rank=rank, world_size=world)
    ad_1 = msgs_optim.FusedAdam(single)
    worst = 0.0
    for step in range(1, 6):
        gr = torch.Generator().manual_seed(100 * step + rank)  # every rank its own gradients (its own views)
        mine = torch.randn(init.numel(), generator=gr).to(dev) * 10.0 ** (step - 3)
        sharded.grads.copy_(mine)
        total = mine.clone()
        dist.all_reduce(total)                                  # the single-GPU step sees the summed gradients
        single.grads.copy_(total)
        ad_s.step()
        ad_1.step()
        torch.cuda.synchronize()
        # fp32 sums of 2 (N) terms: the switch and NCCL may add in a different order for N > 2; N = 2 is exact
        a, b = sharded.params.double(), single.params.double()
        err = float(((a - b).abs() / (b.abs() + 0.1)).max())
        worst = max(worst, err)
        lim = 0.0 if world == 2 else 1e-5
        if err > lim:
            print(f"rank {rank}: step {step}: parameters differ from the single-GPU step: {err}", flush=True)
            sys.exit(1)
    # every replica identical
    chk = sharded.params.double().sum().reshape(1)
    lo, hi = chk.clone(), chk.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    if float(lo) != float(hi):
        print(f"rank {rank}: replicas diverged", flush=True)
        sys.exit(1)

    # typed lanes: sum | max | min over non-negative floats
    import torch.distributed._symmetric_memory as symm_mem
    n_sum, n_max, n_min = 4 * 1001, 4 * 333, 4 * 77
    buf = symm_mem.empty(n_sum + n_max + n_min, dtype=torch.float32, device=dev)
    h = symm_mem.rendezvous(buf, dist.group.WORLD)
    gr = torch.Generator().manual_seed(7 + rank)
    vals = (torch.rand(buf.numel(), generator=gr) * 100).to(dev)
    vals[n_sum + 5] = float("inf")                              # "nothing observed" markers survive min / max
    buf.copy_(vals)
    want = [vals[:n_sum].clone(), vals[n_sum:n_sum + n_max].clone(), vals[n_sum + n_max:].clone()]
    dist.all_reduce(want[0])
    dist.all_reduce(want[1], op=dist.ReduceOp.MAX)
    dist.all_reduce(want[2], op=dist.ReduceOp.MIN)
    torch.cuda.synchronize()
    msgs_optim.stats_allreduce_multimem(h, n_sum, n_max, n_min, rank, world, dev)
    torch.cuda.synchronize()
    ok = (torch.allclose(buf[:n_sum], want[0], rtol=1e-6, atol=0) and torch.equal(buf[n_sum:n_sum + n_max], want[1])
          and torch.equal(buf[n_sum + n_max:], want[2]))
    if not ok:
        print(f"rank {rank}: typed-lane all-reduce differs", flush=True)
        sys.exit(1)
    dist.barrier()
    if rank == 0:
        print(f"SHARDED_ADAM_OK world={world} P={P} worst_rel={worst:.3e}")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
