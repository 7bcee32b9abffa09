"""Memory-safety evidence without compute-sanitizer.  `compute-sanitizer` is closed on this GPU pool
(profiles/r02_compute_sanitizer_closed.txt: every tool exits 86 with the operators' notice), so the checks it would make are
approximated with what the pool allows:

  initcheck  ->  POISON: the caching allocator's free blocks are filled with 0x00, then with 0xFF (NaN patterns), before the same
                 forward + backward + training step; every output that read memory it did not write would differ between the
                 two runs.  Forward outputs must be bit-identical, gradients identical up to atomic-order noise.
  memcheck   ->  tools/sanitize_case.py with PYTORCH_NO_CUDA_MEMORY_CACHING=1 (every tensor its own cudaMalloc: an access past a
                 buffer's allocation faults instead of landing in a neighbouring cached block) and CUDA_LAUNCH_BLOCKING=1.
  racecheck  ->  the forward has no order-dependent arithmetic: 8 repetitions must agree bit for bit (a race in the onesweep
                 look-back, the counting / placement passes or the speculative buffer shows up as a differing list or image)."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "ms-gs_b200"))
pytestmark = pytest.mark.gpu


def _poison(dev, byte):
    """fill (most of) the allocator's reusable memory with `byte`, then give it back to the cache"""
    torch.cuda.synchronize()
    blocks = []
    for mb in (512, 256, 128, 64, 32, 16, 8, 4, 2, 1):
        for _ in range(3):
            blocks.append(torch.empty(mb << 20, dtype=torch.uint8, device=dev).fill_(byte))
    torch.cuda.synchronize()
    del blocks


def _case(dev, P=20_000, W=320, H=208):
    import diff_gaussian_rasterization as dgr
    import msgs_optim
    import msgs_train
    from harness import scenes
    from test_gpu_train_step import _clone_state, _setup
    sc, raw, state, views = _setup(P, W, H, [1, 2, 4, 8], 1, dev, seed=21)
    flat = msgs_optim.FlatGaussians(P, dev)
    flat.load(raw["xyz"], raw["f_dc"], raw["f_rest"], raw["scaling"], raw["rotation"], raw["opacity"])
    tr = msgs_train.MultiScaleStep(flat, _clone_state(state), torch.zeros(3, device=dev), sc.sh_degree)
    loss, units, m2 = tr.render_and_backward(views)
    torch.cuda.synchronize()
    fwd = [u[k].clone() for u in units for k in ("render", "acc_pixel_size", "depth", "radii", "pixel_sizes")]
    grads = [flat.grads.clone()] + [m.clone() for m in m2]
    tr.update_statistics(views, units, m2)
    tr.adam.step()
    torch.cuda.synchronize()
    return fwd, grads, [flat.params.clone(), tr.state.min_pixel_sizes.clone(), tr.state.max_pixel_sizes.clone(), tr.state.denom.clone()]


def test_results_do_not_depend_on_what_freed_memory_holds(built):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    dev = torch.device("cuda", 0)
    _case(dev)                                    # warm the allocator so that both runs reuse cached blocks
    _poison(dev, 0x00)
    f0, g0, s0 = _case(dev)
    _poison(dev, 0xFF)
    f1, g1, s1 = _case(dev)
    for a, b in zip(f0, f1):
        assert torch.equal(a, b), "a forward output depends on uninitialised memory"
    for a, b in zip(g0, g1):
        assert bool(torch.isfinite(b).all())
        d = (a.double() - b.double()).norm() / a.double().norm().clamp_min(1e-30)
        assert float(d) < 1e-5, f"a gradient depends on uninitialised memory (rel {float(d):.2e})"
    for a, b in zip(s0[1:], s1[1:]):
        assert torch.equal(a, b)


def test_forward_is_bit_reproducible(built):
    import diff_gaussian_rasterization as dgr
    from harness import msgs_state, scenes
    dev = torch.device("cuda", 0)
    P, W, H = 60_000, 400, 300
    sc = scenes.make_scene(P, W, H, seed=8)
    cam = scenes.make_camera(W, H, yaw_deg=-9.0)
    tt = lambda a, dt=torch.float32: torch.from_numpy(np.ascontiguousarray(a)).to(device=dev, dtype=dt)
    C, e = dgr._C, torch.empty(0)
    args = (tt(sc.bg), tt(sc.means3D), e, tt(sc.max_pixel_sizes), tt(sc.min_pixel_sizes), tt(sc.opacities), tt(sc.occ_multiplier),
            tt(sc.dc_delta), tt(sc.scales), tt(sc.rotations), 1.0, e, tt(sc.base_mask, torch.bool), tt(cam.viewmatrix),
            tt(cam.projmatrix), cam.tanfovx, cam.tanfovy, H, W, tt(sc.shs), 3, tt(cam.campos), False, False, False, 1.0, False)
    first = None
    for rep in range(8):
        R, color, acc, depth, radii, px, gb, bb, ib = C.rasterize_gaussians(*args)
        st = msgs_state.decode(gb, bb, ib, P, R, W, H)
        cur = (R, color.clone(), acc.clone(), depth.clone(), radii.clone(), px.clone(), st["point_list"].clone(),
               st["ranges"].clone(), st["n_contrib"].clone(), st["final_T"].clone())
        if first is None:
            first = cur
            continue
        assert cur[0] == first[0]
        for a, b in zip(cur[1:], first[1:]):
            assert torch.equal(a, b), f"repetition {rep} differs"


def test_cases_run_with_one_cudamalloc_per_tensor(built):
    env = dict(os.environ, PYTORCH_NO_CUDA_MEMORY_CACHING="1", CUDA_LAUNCH_BLOCKING="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "sanitize_case.py")], env=env, capture_output=True, text=True,
                       timeout=900)
    assert r.returncode == 0 and "SANITIZE_CASE_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-3000:]
