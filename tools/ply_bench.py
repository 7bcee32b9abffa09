"""Host-side measurement of the model-file loader (f4): python tools/ply_bench.py [P=1200000] [L=4] [D=4]
Times msgs_ply.load_ply (all host threads; file in the page cache) against a plain memcpy of the same number of bytes
(the host-memory roofline of a one-pass conversion: file bytes read + fp32 bytes written).  CPU only.  The comparison
with the restated reference loader lives in tests/test_ply_loader.py (only tests may execute oracle/)."""
import os, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "ms-gs_b200"))
import numpy as np
import msgs_ply
from harness import ply_io

P = int(sys.argv[1]) if len(sys.argv) > 1 else 1_200_000
L = int(sys.argv[2]) if len(sys.argv) > 2 else 4
D = int(sys.argv[3]) if len(sys.argv) > 3 else 4
m = ply_io.random_model(P, L, D, seed=1)
with tempfile.TemporaryDirectory() as d:
    path = os.path.join(d, "point_cloud.ply")
    ply_io.write_ms_ply(path, m)
    size = os.path.getsize(path)

    def best(fn, n=5):
        ts = []
        for _ in range(n):
            t0 = time.perf_counter(); r = fn(); ts.append(time.perf_counter() - t0)
        return min(ts), r

    t_mine, (t, info) = best(lambda: msgs_ply.load_ply(path))
    out_bytes = sum(v.numel() * v.element_size() for v in t.values())
    t_one, _ = best(lambda: msgs_ply.load_ply(path, threads=1), 3)
    src = np.fromfile(path, dtype=np.uint8)
    dst = np.empty(size + out_bytes, dtype=np.uint8)
    t_cp, _ = best(lambda: (np.copyto(dst[:size], src), np.copyto(dst[size:], dst[:out_bytes])))
    moved = size + out_bytes
    print(f"model file: P={P} L={L} D={D}: {size / 1e6:.1f} MB ({info['row_bytes']} B/Gaussian) -> {out_bytes / 1e6:.1f} MB of fp32 tensors; "
          f"host cores {os.cpu_count()}")
    print(f"  msgs_ply.load_ply, all threads : {t_mine * 1e3:8.1f} ms  {moved / t_mine / 1e9:6.2f} GB/s moved "
          f"({size / t_mine / 1e9:.2f} GB/s of file)")
    print(f"  msgs_ply.load_ply, 1 thread    : {t_one * 1e3:8.1f} ms  {moved / t_one / 1e9:6.2f} GB/s moved")
    print(f"  numpy memcpy of the same bytes : {t_cp * 1e3:8.1f} ms  {moved / t_cp / 1e9:6.2f} GB/s (1 thread) -> loader at "
          f"{t_cp / t_mine:.2f}x of it")
