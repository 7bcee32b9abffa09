"""Probe of torch symmetric memory / NVSwitch multicast on this box (torchrun, N >= 2): prints multicast support and
times NCCL all-reduce vs symm_mem multimem all-reduce of the C2 gradient volume (1.2 M x 59 floats)."""
import os
import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
n = 1_200_000 * 59
a = torch.ones(n, device=dev)
for _ in range(3):
    dist.all_reduce(a)
torch.cuda.synchronize(); dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    dist.all_reduce(a)
e1.record(); torch.cuda.synchronize()
if rank == 0:
    print(f"NCCL all_reduce {n * 4 / 1e6:.0f} MB x{world}: {e0.elapsed_time(e1) / 10:.3f} ms")
try:
    t = symm_mem.empty(n, dtype=torch.float32, device=dev)
    h = symm_mem.rendezvous(t, dist.group.WORLD)
    if rank == 0:
        print("symm_mem ok: multicast_ptr", hex(h.multicast_ptr), "world", h.world_size, "buffer_ptrs", [hex(p) for p in h.buffer_ptrs][:2],
              "has_multicast_support", getattr(type(h), "has_multicast_support", None))
    t.fill_(1.0)
    torch.cuda.synchronize(); dist.barrier()
    for _ in range(3):
        torch.ops.symm_mem.multimem_all_reduce_(t, "sum", dist.group.WORLD.group_name)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(10):
        torch.ops.symm_mem.multimem_all_reduce_(t, "sum", dist.group.WORLD.group_name)
    e1.record(); torch.cuda.synchronize()
    if rank == 0:
        print(f"symm_mem multimem_all_reduce_: {e0.elapsed_time(e1) / 10:.3f} ms")
    e0.record()
    for _ in range(10):
        h.barrier(channel=0)
    e1.record(); torch.cuda.synchronize()
    if rank == 0:
        print(f"symm_mem barrier: {e0.elapsed_time(e1) / 10 * 1e3:.1f} us")
except Exception as ex:
    if rank == 0:
        print("symm_mem failed:", repr(ex)[:500])
dist.destroy_process_group()
