"""The multi-scale training step (ms-gs_b200/msgs_train.py; SURVEY.md §8 rows f2 + f3 + e) against the reference's training
iteration: harness/ref_train.py = the reference's torch code around its UNMODIFIED CUDA rasterizer (oracle/_ref), i.e.
activations -> rasterizer -> L1 + SSIM (x0.1 on coarse scales) -> backward -> update_pixel_sizes / max_radii2D /
add_densification_stats -> torch.optim.Adam, per (view, scale) unit as train.py:152-260, 416-418 runs them.

Bars.  Loss: 1e-4 relative (images are within 1e-4).  Statistics: denom / max_radii2D exact up to a counted handful of
Gaussians whose radius sits on a rounding boundary (the raw-parameter forward evaluates exp / sigmoid with ex2, torch with
its own expf: last-ulp differences in scale), pixel-size state 1e-5 relative, gradient-norm accumulator 2e-3 relative l2.
Parameters after Adam: Adam divides by sqrt(v), so the first steps move every rendered parameter by ~lr * sign(g): an
element whose gradient is at the atomic-order noise floor may flip sign.  Bar: <= 2e-3 of the elements of a tensor off
by more than 0.1 * lr after each of 3 steps (measured: ~1e-4), and the rest within that."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "ms-gs_b200"))

from harness import scenes  # noqa: E402

pytestmark = pytest.mark.gpu


def _setup(P, W, H, scales, n_views, dev, seed=77):
    import msgs_types
    sc = scenes.make_scene(P, W, H, seed=seed)
    rng = np.random.default_rng(seed + 1)
    tt = lambda a, dt=torch.float32: torch.from_numpy(np.ascontiguousarray(a)).to(device=dev, dtype=dt)
    logit = lambda x: np.log(np.clip(x, 1e-6, 1 - 1e-6) / (1 - np.clip(x, 1e-6, 1 - 1e-6))).astype(np.float32)
    raw = {"xyz": tt(sc.means3D), "f_dc": tt(sc.shs[:, :1]), "f_rest": tt(sc.shs[:, 1:]), "opacity": tt(logit(sc.opacities)),
           "scaling": tt(np.log(sc.scales)), "rotation": tt(sc.rotations * rng.uniform(0.5, 2.0, (P, 1)).astype(np.float32))}
    L = len(scales)
    # MS-GS state: a third of the Gaussians have initialised pixel sizes, 15 % belong to a coarser level
    mn = np.where(rng.random(P) < 0.35, rng.uniform(0.5, 6.0, P), -1.0).astype(np.float32)
    large = (rng.random(P) < 0.15) & (mn > 0)
    mx = np.where(large, mn * rng.uniform(2, 8, P), -1.0).astype(np.float32)
    target = np.where(large, rng.integers(1, max(2, L), P), 0).astype(np.int64)
    base = rng.random(P) < 0.05
    state = msgs_types.MSState.fresh(P, L, dev, tt(mx), tt(mn), tt(base, torch.bool), tt(target, torch.int64))
    g = torch.Generator().manual_seed(seed)
    views = []
    for v in range(n_views):
        cam = scenes.make_camera(W, H, yaw_deg=-12.0 + 24.0 * v / max(1, n_views - 1) if n_views > 1 else 0.0)
        sizes, gts, levels = [], [], []
        for lvl, s in enumerate(scales):
            cs = cam.at_scale(s)
            sizes.append((cs.H, cs.W))
            gts.append(torch.rand((3, cs.H, cs.W), generator=g).to(dev))
            levels.append(lvl)
        views.append(msgs_types.TrainView(tt(cam.viewmatrix), tt(cam.projmatrix), tt(cam.campos), cam.tanfovx, cam.tanfovy,
                                          sizes, gts, levels, view_id=v))
    return sc, raw, state, views


def _clone_state(st):
    import msgs_types
    return msgs_types.MSState(**{k: (v.clone() if torch.is_tensor(v) else v) for k, v in st.__dict__.items()})


def _compare_params(flat, ref_p, lrs, what):
    pairs = (("xyz", flat.xyz), ("f_dc", flat.features_dc), ("f_rest", flat.features_rest), ("opacity", flat.opacity),
             ("scaling", flat.scaling), ("rotation", flat.rotation))
    for name, mine in pairs:
        a, b = mine.detach().double().cpu(), ref_p[name].detach().double().cpu()
        off = ((a - b).abs() > 0.1 * lrs[name])
        frac = float(off.double().mean())
        assert frac <= 2e-3, f"{what}: {name}: {frac:.2e} of the elements differ by more than 0.1 lr"
        assert float((a - b).abs().max()) <= 2.5 * lrs[name] * 3, f"{what}: {name}: max diff {(a - b).abs().max():.3e}"


def _compare_stats(st, ref, what, budget=4):
    """budget: Gaussians allowed to differ.  Step 0 starts from identical parameters (a handful of radii on a rounding
    boundary); from step 1 on the two models have taken one Adam step each and differ in ~1e-4 of their elements by ~lr
    (module docstring), which moves a few more Gaussians across a visibility / radius boundary at some scale."""
    assert int((st.denom != ref.denom).sum()) <= budget, what
    assert int((st.max_radii2D != ref.max_radii2D).sum()) <= budget, what
    for a, b, n in ((st.max_pixel_sizes, ref.max_px, "max_px"), (st.min_pixel_sizes, ref.min_px, "min_px")):
        bad = ((a - b).abs() > (1e-5 if budget <= 4 else 1e-3) * b.abs().clamp_min(1e-3))
        assert int(bad.sum()) <= budget, f"{what}: {n}: {int(bad.sum())} differ"
    a, b = st.xyz_gradient_accum.double(), ref.accum.double()
    assert float((a - b).norm() / b.norm().clamp_min(1e-30)) < 2e-3, what


@pytest.mark.parametrize("n_views", [1, 2])
def test_training_step_matches_the_reference_iteration(built, n_views):
    from harness import ref_train, refcuda_bind
    if not refcuda_bind.available():
        pytest.skip("oracle/_ref not built")
    import msgs_optim
    import msgs_train
    import msgs_types
    dev = torch.device("cuda", 0)
    P, W, H, scales = 30_000, 320, 208, [1, 2, 4, 8]
    sc, raw, state, views = _setup(P, W, H, scales, n_views, dev)
    bg = torch.zeros(3, device=dev)
    lrs = dict(msgs_types.DEFAULT_LRS)
    flat = msgs_optim.FlatGaussians(P, dev)
    flat.load(raw["xyz"], raw["f_dc"], raw["f_rest"], raw["scaling"], raw["rotation"], raw["opacity"])
    mine = msgs_train.MultiScaleStep(flat, _clone_state(state), bg, sc.sh_degree, lrs)
    ref = ref_train.RefTrainer(raw, _clone_state(state), bg, sc.sh_degree, lrs, stats_what=3, sync=True)
    for step in range(3):
        vs = views if n_views > 1 else [views[0]]
        lm = float(mine.step(vs))
        lr_ = float(ref.step(vs))
        torch.cuda.synchronize()
        assert abs(lm - lr_) <= (1e-4 if step == 0 else 1e-3 * n_views ** 2) * abs(lr_), f"step {step}: loss {lm} vs {lr_}"
        if n_views == 1:
            # one view per step: the statistics are those of per-unit updates in order, exactly the reference's sequence
            _compare_stats(mine.state, ref, f"step {step}", 4 if step == 0 else P // 300)
        if n_views == 1 or step == 0:
            # several views per step: from step 1 on the pixel-size state differs by design (next comment), some Gaussians
            # pass the MS-GS filters on one side only and the two models part ways (measured: 3 % of xyz after step 1)
            _compare_params(flat, ref.p, lrs, f"views={n_views} step {step}")
    if n_views > 1:
        # several views per step go through the delta form (one observation per Gaussian and step: the 0.95 / 1.05 ageing is
        # applied once, include/msgs_stats.h) — denominators and gradient norms still add up exactly as per-unit updates do
        assert float((mine.state.denom != ref.denom).double().mean()) <= 5e-2
        a, b = mine.state.xyz_gradient_accum.double(), ref.accum.double()
        assert float((a - b).norm() / b.norm()) < 0.1


def test_snapshot_restore_replays_the_same_steps(built):
    import msgs_optim
    import msgs_train
    dev = torch.device("cuda", 0)
    P, W, H, scales = 8_000, 160, 112, [1, 2, 4]
    sc, raw, state, views = _setup(P, W, H, scales, 1, dev, seed=5)
    flat = msgs_optim.FlatGaussians(P, dev)
    flat.load(raw["xyz"], raw["f_dc"], raw["f_rest"], raw["scaling"], raw["rotation"], raw["opacity"])
    tr = msgs_train.MultiScaleStep(flat, state, torch.zeros(3, device=dev), sc.sh_degree)
    snap = tr.snapshot()
    l1 = [float(tr.step(views)) for _ in range(3)]
    p1 = flat.params.clone()
    tr.restore(snap)
    l2 = [float(tr.step(views)) for _ in range(3)]
    # atomic-order noise only
    assert np.allclose(l1, l2, rtol=1e-5)
    assert float((flat.params - p1).abs().max()) <= 3 * 0.05 + 1e-6
    assert float(((flat.params - p1).abs() > 1e-5).double().mean()) < 2e-3


def test_sharded_step_two_ranks(built):
    """2 GPUs: the sharded step (each rank renders one view; statistics deltas reduced by the typed-lane NVLS kernel;
    reduce-scatter + Adam + all-gather in one kernel) equals the single-GPU step over both views."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run under gpurun --gpus 2)")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                        "127.0.0.1", "--master-port", "29613", os.path.join(ROOT, "tools", "check_sharded_step.py")],
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and "SHARDED_STEP_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
