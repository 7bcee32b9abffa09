"""2+ ranks under torchrun (one GPU each): the sharded multi-scale training step (msgs_train.MultiScaleStep with
world_size = N: every rank renders ITS view, statistics deltas reduced by the typed-lane NVLS kernel, reduce-scatter + Adam
+ all-gather as ONE multimem kernel) against the single-GPU step over ALL the ranks' views, computed redundantly on every
rank.  Prints SHARDED_STEP_OK (rank 0).
  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/check_sharded_step.py [P]"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "ms-gs_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import msgs_optim  # noqa: E402
import msgs_train  # noqa: E402
import msgs_types  # noqa: E402
from test_gpu_train_step import _clone_state, _setup  # noqa: E402


def fail(rank, msg):
    print(f"rank {rank}: {msg}", flush=True)
    sys.exit(1)


def main():
    P = int(sys.argv[1]) if len(sys.argv) > 1 else 40_000
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    dev = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("nccl", device_id=dev)
    W, H, scales = 320, 208, [1, 2, 4, 8]
    sc, raw, state, views = _setup(P, W, H, scales, world, dev)         # identical on every rank (seeded on the host)
    bg = torch.zeros(3, device=dev)
    lrs = dict(msgs_types.DEFAULT_LRS)

    sharded = msgs_optim.FlatGaussians(P, dev, symmetric=True)
    single = msgs_optim.FlatGaussians(P, dev)
    for f in (sharded, single):
        f.load(raw["xyz"], raw["f_dc"], raw["f_rest"], raw["scaling"], raw["rotation"], raw["opacity"])
    tr_s = msgs_train.MultiScaleStep(sharded, _clone_state(state), bg, sc.sh_degree, lrs, rank=rank, world_size=world)
    tr_1 = msgs_train.MultiScaleStep(single, _clone_state(state), bg, sc.sh_degree, lrs)
    for step in range(3):
        ls = tr_s.step([views[rank]])
        l1 = tr_1.step(views)
        tr_s.sync_statistics()          # fold the rank-local densification sums in (the step does it every 100 steps)
        torch.cuda.synchronize()
        tot = ls.clone().reshape(1)
        dist.all_reduce(tot)
        if abs(float(tot) - float(l1)) > 1e-5 * abs(float(l1)):
            fail(rank, f"step {step}: summed loss {float(tot)} vs single-GPU {float(l1)}")
        # Adam normalises by sqrt(v): an element whose gradient is at the summation-order noise floor may flip sign, so the
        # bar is a fraction of elements (tests/test_gpu_train_step.py), not a maximum
        for name in ("sh", "rot", "mean3D", "scale", "opacity"):
            a, b = sharded.p[name].double(), single.p[name].double()
            lr = {"sh": lrs["f_rest"], "rot": lrs["rotation"], "mean3D": lrs["xyz"], "scale": lrs["scaling"], "opacity": lrs["opacity"]}[name]
            frac = float(((a - b).abs() > 0.1 * lr).double().mean())
            if frac > 2e-3:
                fail(rank, f"step {step}: {name}: {frac:.2e} of the elements differ from the single-GPU step")
        s, o = tr_s.state, tr_1.state
        # step 0 starts from identical parameters: the statistics must agree exactly.  From step 1 on the two models have
        # each taken an Adam step and differ in ~1e-4 of their elements by ~lr (see above), which moves the observed pixel
        # size of a few Gaussians in the last bits and a handful across a visibility boundary: counted budget.
        budget = 0 if step == 0 else P // 500
        if int((s.denom != o.denom).sum()) > budget or int((s.max_radii2D != o.max_radii2D).sum()) > budget:
            fail(rank, f"step {step}: denom / max_radii2D differ from the single-GPU step")
        for a, b, n in ((s.max_pixel_sizes, o.max_pixel_sizes, "max_px"), (s.min_pixel_sizes, o.min_pixel_sizes, "min_px")):
            bad = int((a != b).sum()) if step == 0 else int(((a - b).abs() > 1e-3 * b.abs().clamp_min(1e-3)).sum())
            if bad > budget:
                fail(rank, f"step {step}: {n} differs from the single-GPU step on {bad} Gaussians")
        a, b = s.xyz_gradient_accum.double(), o.xyz_gradient_accum.double()
        if float((a - b).norm() / b.norm().clamp_min(1e-30)) > (1e-5 if step == 0 else 2e-3):
            fail(rank, f"step {step}: gradient-norm accumulators differ")
    # every replica identical (parameters and statistics)
    for t in (sharded.params, tr_s.state.min_pixel_sizes, tr_s.state.xyz_gradient_accum):
        chk = t.double().sum().reshape(1)
        lo, hi = chk.clone(), chk.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        if float(lo) != float(hi):
            fail(rank, "replicas diverged")
    dist.barrier()
    if rank == 0:
        print("SHARDED_STEP_OK", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
