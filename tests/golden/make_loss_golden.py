"""Generates tests/golden/loss_golden.npz: the reference's OWN training loss run here on CPU (row f2).

`utils.loss_utils` is imported from /root/reference; `loss = (1 - lambda) * l1_loss(image, gt) + lambda * (1 - ssim(image,
gt))` exactly as train.py:210-211 writes it, the gradient w.r.t. the image from autograd, and the window taps
`gaussian(11, 1.5)` the reference builds.  Run:  python tests/golden/make_loss_golden.py"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, "/root/reference")
from utils.loss_utils import gaussian, l1_loss, ssim  # noqa: E402

out = {"w1d": gaussian(11, 1.5).numpy()}
for tag, (C, H, W, lam, seed) in {"c3_40x52": (3, 40, 52, 0.2, 1), "c3_7x9": (3, 7, 9, 0.2, 2), "c1_1x1": (1, 1, 1, 0.5, 3),
                                  "c3_23x11_l0": (3, 23, 11, 0.0, 4), "c3_16x16_l1": (3, 16, 16, 1.0, 5)}.items():
    g = torch.Generator().manual_seed(seed)
    gt = torch.rand(C, H, W, generator=g)
    img = (gt + 0.2 * torch.randn(C, H, W, generator=g)).clamp(0, 1).requires_grad_(True)
    loss = (1.0 - lam) * l1_loss(img, gt) + lam * (1.0 - ssim(img, gt))
    (loss * 1.7).backward()
    out[f"{tag}/img"], out[f"{tag}/gt"] = img.detach().numpy(), gt.numpy()
    out[f"{tag}/lam"], out[f"{tag}/loss"], out[f"{tag}/grad_x1.7"] = np.float32(lam), loss.detach().numpy(), img.grad.numpy()
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "loss_golden.npz"), **out)
print("wrote", len(out), "arrays")
