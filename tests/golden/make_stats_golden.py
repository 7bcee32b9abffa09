"""Generates tests/golden/stats_golden.npz: the reference's OWN per-step statistics updates run here on CPU (row f2).

`scene.gaussian_model.GaussianModel` is imported from /root/reference with the absent third-party modules stubbed (as in
make_activations_golden.py); its methods `update_pixel_sizes`, `add_densification_stats`, `update_base_gaussian_mask`
(scene/gaussian_model.py:662-705) and the max_radii2D line of train.py:249 (restated literally below) are applied to
seeded state for a sequence of rendered units; the state after every unit is stored.
Run:  python tests/golden/make_stats_golden.py"""
import importlib.abc
import importlib.machinery
import os
import sys
from unittest import mock

import numpy as np
import torch


class _StubMissing(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    def find_spec(self, name, path=None, target=None):
        return importlib.machinery.ModuleSpec(name, self, is_package=True)

    def create_module(self, spec):
        m = mock.MagicMock(name=spec.name)
        m.__path__, m.__spec__, m.__name__, m.__loader__ = [], spec, spec.name, self
        return m

    def exec_module(self, module):
        pass


sys.path.insert(0, "/root/reference")
sys.meta_path.append(_StubMissing())
from scene.gaussian_model import GaussianModel  # noqa: E402

STATE = ("max_pixel_sizes", "min_pixel_sizes", "max_radii2D", "xyz_gradient_accum", "denom", "base_gaussian_mask")
out = {}
for tag, (P, L, seed) in {"p1000_l4": (1000, 4, 11), "p37_l1": (37, 1, 12), "p1_l3": (1, 3, 13)}.items():
    g = torch.Generator().manual_seed(seed)
    pc = GaussianModel(3, reso_lvls=L)
    pc.max_pixel_sizes = torch.where(torch.rand(P, generator=g) < 0.3, -torch.ones(P), torch.rand(P, generator=g) * 20)
    pc.min_pixel_sizes = torch.where(torch.rand(P, generator=g) < 0.3, -torch.ones(P), torch.rand(P, generator=g) * 3)
    pc.max_radii2D = torch.rand(P, generator=g) * 30
    pc.target_reso_lvl = torch.randint(0, L, (P,), generator=g, dtype=torch.long)
    pc.xyz_gradient_accum = torch.rand(P, L, 1, generator=g)
    pc.denom = torch.randint(0, 5, (P, L, 1), generator=g).float()
    pc.base_gaussian_mask = torch.rand(P, generator=g) < 0.2
    out[f"{tag}/target_reso_lvl"] = pc.target_reso_lvl.numpy().copy()
    for k in STATE:
        out[f"{tag}/init/{k}"] = getattr(pc, k).numpy().copy()
    units = [(lvl, what) for lvl in range(L) for what in (7, 3, 1, 2, 4)][: 3 * L + 2]
    out[f"{tag}/units"] = np.array(units, np.int32)
    for u, (lvl, what) in enumerate(units):
        radii = torch.randint(-1, 40, (P,), generator=g, dtype=torch.int32) * (torch.rand(P, generator=g) < 0.7).int()
        px = torch.where(torch.rand(P, generator=g) < 0.15, torch.zeros(P), torch.rand(P, generator=g) * 25)
        vp = torch.zeros(P, 3, requires_grad=True)
        vp.grad = torch.randn(P, 3, generator=g)
        vis = radii > 0
        if what & 4:
            pc.update_base_gaussian_mask(vis)                                                        # train.py:241
        if what & 1:
            pc.update_pixel_sizes(vis, px, lvl, 1000)                                                # train.py:245
        if what & 2:
            pc.max_radii2D[vis] = torch.max(pc.max_radii2D[vis], radii[vis])                         # train.py:249
            pc.add_densification_stats(vp, vis, reso_lvl=lvl)                                        # train.py:250
        out[f"{tag}/u{u}/radii"], out[f"{tag}/u{u}/pixel_sizes"] = radii.numpy().copy(), px.numpy().copy()
        out[f"{tag}/u{u}/grad"] = vp.grad.numpy().copy()
        for k in STATE:
            out[f"{tag}/u{u}/{k}"] = getattr(pc, k).numpy().copy()
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "stats_golden.npz"), **out)
print("wrote", len(out), "arrays")
