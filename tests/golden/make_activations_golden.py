"""Generates tests/golden/activations_golden.npz: the reference's OWN activations (row f1 of SURVEY.md §8) run here on CPU.

The reference is imported from /root/reference (this container only): `scene.gaussian_model.GaussianModel` with the
absent third-party modules (plyfile, simple_knn, open3d.ml, ...) replaced by inert stand-ins — none of them is touched by the
getters.  A GaussianModel is filled with seeded raw parameters and its properties `get_scaling`, `get_rotation`,
`get_opacity`, `get_features` (scene/gaussian_model.py:127-155) are evaluated; the gradients come from autograd of those
same properties under seeded output gradients.  Run:  python tests/golden/make_activations_golden.py"""
import importlib
import importlib.abc
import importlib.machinery
import os
import sys
from unittest import mock

import numpy as np
import torch

REF = "/root/reference"
sys.path.insert(0, REF)


class _StubMissing(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    """Last on sys.meta_path: any module no real finder knows becomes an inert stand-in (package-like MagicMock)."""

    def find_spec(self, name, path=None, target=None):
        return importlib.machinery.ModuleSpec(name, self, is_package=True)

    def create_module(self, spec):
        m = mock.MagicMock(name=spec.name)
        m.__path__, m.__spec__, m.__name__, m.__loader__ = [], spec, spec.name, self
        return m

    def exec_module(self, module):
        pass


sys.meta_path.append(_StubMissing())
from scene.gaussian_model import GaussianModel  # noqa: E402

out = {}
for tag, (P, deg, seed) in {"deg3": (257, 3, 1), "deg2": (101, 2, 2), "deg0": (5, 0, 3), "one": (1, 3, 4)}.items():
    g = torch.Generator().manual_seed(seed)
    D = (deg + 1) ** 2 - 1
    pc = GaussianModel(deg)
    raw = dict(_features_dc=torch.randn(P, 1, 3, generator=g), _features_rest=torch.randn(P, D, 3, generator=g) * 0.2,
               _scaling=torch.randn(P, 3, generator=g) * 2 - 3, _rotation=torch.randn(P, 4, generator=g),
               _opacity=torch.randn(P, 1, generator=g) * 3)
    if tag == "deg3":
        raw["_rotation"][0] = 0.0                       # F.normalize's eps clamp
        raw["_rotation"][1] = torch.tensor([1e-20, 0.0, 0.0, 0.0])
    for k, v in raw.items():
        setattr(pc, k, v.clone().requires_grad_(True))
    pc.cat_feature = None
    outs = dict(shs=pc.get_features, scales=pc.get_scaling, rotations=pc.get_rotation, opacities=pc.get_opacity)
    gouts = {k: torch.randn(v.shape, generator=g) for k, v in outs.items()}
    sum((outs[k] * gouts[k]).sum() for k in outs).backward()
    for k, v in raw.items():
        out[f"{tag}/in/{k}"] = v.numpy()
        out[f"{tag}/grad/{k}"] = getattr(pc, k).grad.numpy()
    for k in outs:
        out[f"{tag}/out/{k}"] = outs[k].detach().numpy()
        out[f"{tag}/gout/{k}"] = gouts[k].numpy()
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "activations_golden.npz"), **out)
print("wrote", len(out), "arrays")
