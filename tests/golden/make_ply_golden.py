"""Generates tests/golden/ply_ms_200.ply + ply_golden.npz: an MS-GS model file written by the reference's OWN
`GaussianModel.save_ply` and read back by its OWN `load_ply` + `get_features` (scene/gaussian_model.py:293-344, 358-417,
127-153) — row f4 of SURVEY.md §8.

Runs in this container only (imports /root/reference).  Third-party modules the reference imports but this image lacks are
stubbed (as in make_activations_golden.py); `plyfile` is replaced by harness/plyfile_standin.py (the PLY container format,
restated).  load_ply creates its tensors with device="cuda": there is no GPU here, so the device keyword is dropped (the values do
not depend on it).  Also asserts that harness/ply_io.py (the writer the other loader tests use) produces the same bytes as
save_ply.   Run:  python tests/golden/make_ply_golden.py"""
import importlib.abc
import importlib.machinery
import os
import sys
from unittest import mock

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = "/root/reference"
sys.path.insert(0, REF)
sys.path.insert(0, ROOT)
from harness import ply_io, plyfile_standin  # noqa: E402

sys.modules["plyfile"] = plyfile_standin


class _StubMissing(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    def find_spec(self, name, path=None, target=None):
        return importlib.machinery.ModuleSpec(name, self, is_package=True)

    def create_module(self, spec):
        m = mock.MagicMock(name=spec.name)
        m.__path__, m.__spec__, m.__name__, m.__loader__ = [], spec, spec.name, self
        return m

    def exec_module(self, module):
        pass


sys.meta_path.append(_StubMissing())
if not hasattr(np, "core"):
    import numpy._core as _core
    np.core = _core
from scene.gaussian_model import GaussianModel  # noqa: E402

# no GPU here: device="cuda" -> CPU (values are unaffected)
_tensor, _to = torch.tensor, torch.Tensor.to
torch.tensor = lambda *a, **k: _tensor(*a, **{kk: vv for kk, vv in k.items() if kk != "device"})
torch.Tensor.to = lambda self, *a, **k: _to(self, *a, **{kk: ("cpu" if kk == "device" else vv) for kk, vv in k.items()})

P, seed = 200, 17
m = ply_io.random_model(P, 4, 4, seed=seed)
pc = GaussianModel(3)
assert pc.n_lvl_occ == 4 and pc.n_lvl_dc == 4, (pc.n_lvl_occ, pc.n_lvl_dc)
t = torch.from_numpy
pc._xyz, pc._features_dc, pc._features_rest = t(m["xyz"]), t(m["features_dc"]), t(m["features_rest"])
pc._opacity, pc._occ_multiplier, pc._dc_delta = t(m["opacity"]), t(m["occ_multiplier"]), t(m["dc_delta"])
pc._scaling, pc._rotation = t(m["scaling"]), t(m["rotation"])
pc.base_gaussian_mask, pc.max_pixel_sizes, pc.min_pixel_sizes = t(m["base_gaussian_mask"]), t(m["max_pixel_sizes"]), t(m["min_pixel_sizes"])

out_dir = os.path.dirname(os.path.abspath(__file__))
path = os.path.join(out_dir, "ply_ms_200.ply")
pc.save_ply(path)                                           # the reference's writer
tmp = path + ".harness"
ply_io.write_ms_ply(tmp, m)
assert open(tmp, "rb").read() == open(path, "rb").read(), "harness/ply_io.py does not write what save_ply writes"
os.remove(tmp)

pc2 = GaussianModel(3)
pc2.load_ply(path)                                          # the reference's reader
pc2.cat_feature = None
npz = dict(xyz=pc2._xyz.detach().numpy(), features=pc2.get_features.detach().numpy(), opacity=pc2._opacity.detach().numpy(),
           scaling=pc2._scaling.detach().numpy(), rotation=pc2._rotation.detach().numpy(),
           occ_multiplier=pc2._occ_multiplier.detach().numpy(), dc_delta=pc2._dc_delta.detach().numpy(),
           base_gaussian_mask=pc2.base_gaussian_mask.numpy(), max_pixel_sizes=pc2.max_pixel_sizes.numpy(),
           min_pixel_sizes=pc2.min_pixel_sizes.numpy(), seed=np.int64(seed))
np.savez_compressed(os.path.join(out_dir, "ply_golden.npz"), **npz)
print("wrote", path, os.path.getsize(path), "bytes;", {k: getattr(v, "shape", None) for k, v in npz.items()})
