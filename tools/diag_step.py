"""diagnostic: where does the training step differ from the reference iteration? (GPU)"""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "ms-gs_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import msgs_loss, msgs_optim, msgs_train, msgs_types
from harness import ref_train
from test_gpu_train_step import _setup, _clone_state
dev = torch.device("cuda", 0)
g = torch.Generator().manual_seed(1)
for (H, W) in [(26, 40), (52, 80), (104, 160), (208, 320), (822, 1237)]:
    a = torch.rand((3, H, W), generator=g).to(dev).requires_grad_(True)
    b = torch.rand((3, H, W), generator=g).to(dev)
    l_ref = ref_train.training_loss(a, b, 0.2); l_ref.backward(); gr = a.grad.clone(); a.grad = None
    l_mine = msgs_loss.l1_ssim_loss(a, b, 0.2); l_mine.backward(); gm = a.grad.clone(); a.grad = None
    tot, per = msgs_loss.l1_ssim_loss_batch([a], [b], [1.0], 0.2); tot.backward(); gb = a.grad.clone(); a.grad = None
    print(f"loss {H}x{W}: ref {float(l_ref):.7f} single {float(l_mine):.7f} batch {float(tot):.7f}  grad maxdiff single {float((gm-gr).abs().max()):.3e} batch {float((gb-gr).abs().max()):.3e} (max |g| {float(gr.abs().max()):.3e})")
P, W, H, scales = 30_000, 320, 208, [1, 2, 4, 8]
sc, raw, state, views = _setup(P, W, H, scales, 1, dev)
bg = torch.zeros(3, device=dev)
flat = msgs_optim.FlatGaussians(P, dev)
flat.load(raw["xyz"], raw["f_dc"], raw["f_rest"], raw["scaling"], raw["rotation"], raw["opacity"])
mine = msgs_train.MultiScaleStep(flat, _clone_state(state), bg, sc.sh_degree)
ref = ref_train.RefTrainer(raw, _clone_state(state), bg, sc.sh_degree, msgs_types.DEFAULT_LRS, sync=True)
loss, units, m2 = mine.render_and_backward(views)
v = views[0]
for i, ((h, w), gt, lvl) in enumerate(zip(v.sizes, v.gts, v.levels)):
    img, radii, px, _ = ref.render(v, h, w)
    u = units[i]
    print(f"unit {i} {w}x{h}: image maxdiff {float((u['render'] - img).abs().max()):.3e} radii mismatches {int((u['radii'] != radii).sum())} "
          f"px maxrel {float(((u['pixel_sizes'] - px).abs() / px.abs().clamp_min(1e-6)).max()):.3e} "
          f"loss mine {float(msgs_loss.l1_ssim_loss(u['render'].detach(), gt, 0.2)):.6f} ref {float(ref_train.training_loss(img.detach(), gt, 0.2)):.6f}")
print("total mine", float(loss))
