"""Is the step jitter of the device-resident loop the caching allocator?  (diagnostic)
python tools/alloc_probe.py [steps]   -> per-step host ms in fwd / bwd calls + cudaMalloc/cudaFree counters,
for (A) fresh gradient tensors every step and (B) preallocated gradient tensors (out=)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "ms-gs_b200"))
import torch
import bench
import msgs_sharding

N = int(sys.argv[1]) if len(sys.argv) > 1 else 60
dev = torch.device("cuda", 0)
wl = bench.build_workload("C2", 0, 1, dev)
r = bench.MineRunner(wl, dev)


def counters():
    s = torch.cuda.memory_stats(dev)
    return s.get("num_device_alloc", 0), s.get("num_device_free", 0), s.get("num_alloc_retries", 0), s["reserved_bytes.all.current"] >> 20


for mode in ("fresh", "out", "fresh", "out"):
    flat = msgs_sharding.GradientExchange(wl["P"], dev) if mode == "out" else None
    kw = dict(grad_out=flat.shaped()) if flat is not None else {}
    for _ in range(8):
        r.step(**kw)
    torch.cuda.synchronize()
    c0 = counters()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(N + 1)]
    host, mallocs = [], []
    ev[0].record()
    for i in range(N):
        a = time.perf_counter()
        m0 = counters()[0]
        r.step(**kw)
        host.append(time.perf_counter() - a)
        mallocs.append(counters()[0] - m0)
        ev[i + 1].record()
    torch.cuda.synchronize()
    c1 = counters()
    slow = [(i, round(ev[i].elapsed_time(ev[i + 1]), 2), mallocs[i]) for i in range(N) if ev[i].elapsed_time(ev[i + 1]) > 1.8 or mallocs[i]]
    print("   steps slow or with cudaMalloc (step, ms, mallocs):", slow[:40])
    d = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(N))
    print(f"{mode:6s} device ms/step min {d[0]:.2f} med {d[N // 2]:.2f} p90 {d[int(N * .9)]:.2f} max {d[-1]:.2f} mean {sum(d) / N:.2f}"
          f" | host mean {sum(host) / N * 1e3:.2f} | cudaMalloc {c1[0] - c0[0]} cudaFree {c1[1] - c0[1]} retries {c1[2] - c0[2]} reserved {c1[3]} MiB")
