"""Flat parameter store + fused Adam for the MS-GS training loop (device half; include/msgs_optim.h).

The reference keeps eight separately allocated parameter tensors and one torch.optim.Adam param group per tensor
(scene/gaussian_model.py:229-252) and steps them with ~9 foreach launches (train.py:416-418).  Here every trainable tensor
lives in ONE flat fp32 buffer, field by field, in the layout of the flat gradient buffer the rasterizer's backward writes
(`msgs_sharding.GRAD_FIELDS`: sh 48, rot 4, mean3D 3, scale 3, opacity 1 floats per Gaussian, every field 16-byte aligned):

    FlatLayout(P)            offsets / sizes of the fields
    FlatGaussians(P, dev)    parameters (raw, as GaussianModel stores them: _xyz, _features_dc/_rest as one (P,16,3) tensor,
                             _scaling (log), _rotation (unnormalised), _opacity (logit)) + the gradient buffer, as views of
                             two flat buffers (plain CUDA memory, or symmetric memory for the sharded step)
    FusedAdam(flat, lrs)     torch.optim.Adam(lr per group, eps=1e-15) semantics in ONE launch per step (msgs_adam_step);
                             sharded: reduce-scatter + Adam on the rank's 1/N shard of the state + all-gather of the updated
                             parameters as ONE NVLS kernel (msgs_adam_step_multimem)

No CPU path: CPU tensors raise, a missing library fails the import."""
import ctypes
import os
from typing import Dict, Optional

import torch

from msgs_sharding import GRAD_FIELDS

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libmsgs_raster.so")
if not os.path.exists(_LIB_PATH):
    raise ImportError(f"{_LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'`")
_lib = ctypes.CDLL(_LIB_PATH)
_vp, _ci, _ll, _d = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_double


class _AdamField(ctypes.Structure):
    """msgs_adam_field (include/msgs_optim.h)"""
    _fields_ = [("begin", _ll), ("end", _ll), ("lr", _d), ("lr_alt", _d), ("period", _ci), ("alt_count", _ci)]


_lib.msgs_adam_step.restype = _ci
_lib.msgs_adam_step.argtypes = [_vp, _vp, _vp, _vp, _ll, ctypes.POINTER(_AdamField), _ci, _d, _d, _d, _ci, _vp]
_lib.msgs_adam_step_multimem.restype = _ci
_lib.msgs_adam_step_multimem.argtypes = [_vp, _vp, _vp, _vp, _vp, _ll, ctypes.POINTER(_AdamField), _ci, _d, _d, _d, _ci, _ci, _ci,
                                         _ci, _vp]
_lib.msgs_adam_step_multimem_range.restype = _ci
_lib.msgs_adam_step_multimem_range.argtypes = [_vp, _vp, _vp, _vp, _vp, _ll, ctypes.POINTER(_AdamField), _ci, _d, _d, _d, _ci, _ll, _ll,
                                               _ci, _ci, _ci, _vp]
_lib.msgs_adam_shard_floats.restype = _ll
_lib.msgs_adam_shard_floats.argtypes = [_ll, _ci]
_lib.msgs_adam_shard_begin.restype = _ll
_lib.msgs_adam_shard_begin.argtypes = [_ll, _ci, _ci]
_lib.msgs_stats_allreduce_multimem.restype = _ci
_lib.msgs_stats_allreduce_multimem.argtypes = [_vp, _ll, _ll, _ll, _ci, _ci, _ci, _vp]

from msgs_types import DEFAULT_LRS      # noqa: E402  (the reference learning rates)
# flat field -> (lr name, lr name of the first `alt_count` floats of every `period`-float row)
_FIELD_LR = {"sh": ("f_rest", "f_dc", 48, 3), "rot": ("rotation", None, 0, 0), "mean3D": ("xyz", None, 0, 0),
             "scale": ("scaling", None, 0, 0), "opacity": ("opacity", None, 0, 0)}


class FlatLayout:
    """Field offsets (in floats) of the flat 59-float-per-Gaussian buffer; every field starts 16-byte aligned."""

    def __init__(self, P: int):
        self.P = int(P)
        self.offset: Dict[str, int] = {}
        self.size: Dict[str, int] = {}
        off = 0
        for name, n in GRAD_FIELDS:
            self.offset[name] = off
            self.size[name] = self.P * n
            off += (self.P * n + 3) // 4 * 4
        self.total = off

    def views(self, flat: torch.Tensor) -> Dict[str, torch.Tensor]:
        P = self.P
        shp = {"sh": (P, 16, 3), "rot": (P, 4), "mean3D": (P, 3), "scale": (P, 3), "opacity": (P, 1)}
        return {k: flat[self.offset[k]:self.offset[k] + self.size[k]].view(shp[k]) for k, _ in GRAD_FIELDS}


class FlatGaussians:
    """Raw Gaussian parameters and their gradients as views of two flat buffers.

    symmetric=True (sharded step): both buffers are torch.distributed._symmetric_memory allocations mapped through the
    NVSwitch multicast object (same layout on every rank); `params_mc` / `grads_mc` are their multicast addresses."""

    def __init__(self, P: int, device, symmetric: bool = False, group=None):
        self.layout = FlatLayout(P)
        self.P, self.device, self.symmetric = int(P), device, bool(symmetric)
        n = self.layout.total
        self.params_mc = self.grads_mc = 0
        if symmetric:
            import torch.distributed as dist
            import torch.distributed._symmetric_memory as symm_mem
            self.group = group if group is not None else dist.group.WORLD
            self.params = symm_mem.empty(n, dtype=torch.float32, device=device)
            self.grads = symm_mem.empty(n, dtype=torch.float32, device=device)
            self.params_handle = symm_mem.rendezvous(self.params, self.group)
            self.grads_handle = symm_mem.rendezvous(self.grads, self.group)
            if not self.params_handle.multicast_ptr or not self.grads_handle.multicast_ptr:
                raise RuntimeError("symmetric memory has no multicast mapping on this system")
            self.params_mc, self.grads_mc = int(self.params_handle.multicast_ptr), int(self.grads_handle.multicast_ptr)
            self.params.zero_()
            self.grads.zero_()
        else:
            self.params = torch.zeros(n, dtype=torch.float32, device=device)
            self.grads = torch.zeros(n, dtype=torch.float32, device=device)
        self.p = self.layout.views(self.params)        # {"sh", "rot", "mean3D", "scale", "opacity"} raw parameters
        self.g = self.layout.views(self.grads)         # the shapes _C.rasterize_gaussians_backward_accumulate adds into

    # GaussianModel's names (scene/gaussian_model.py:127-155)
    @property
    def xyz(self): return self.p["mean3D"]
    @property
    def features(self): return self.p["sh"]
    @property
    def features_dc(self): return self.p["sh"][:, :1, :]
    @property
    def features_rest(self): return self.p["sh"][:, 1:, :]
    @property
    def scaling(self): return self.p["scale"]
    @property
    def rotation(self): return self.p["rot"]
    @property
    def opacity(self): return self.p["opacity"]

    def load(self, xyz, features_dc, features_rest, scaling, rotation, opacity):
        """Copy a model in (raw values, as GaussianModel / the PLY loader hold them)."""
        with torch.no_grad():
            self.p["mean3D"].copy_(xyz)
            self.p["sh"][:, :1, :].copy_(features_dc)
            self.p["sh"][:, 1:, :].copy_(features_rest)
            self.p["scale"].copy_(scaling)
            self.p["rot"].copy_(rotation)
            self.p["opacity"].copy_(opacity.reshape(self.P, 1))

    def zero_grad(self):
        self.grads.zero_()


class FusedAdam:
    """torch.optim.Adam(param groups with the reference's learning rates, betas (0.9, 0.999), eps 1e-15) over a
    FlatGaussians in ONE launch per step.  `lrs`: {"xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation"} -> lr
    (set_lr changes one, e.g. the scheduled position learning rate of update_learning_rate, gaussian_model.py:286-292).

    world_size > 1 (flat.symmetric): step() is reduce-scatter + Adam + all-gather in one NVLS kernel; this rank holds only
    its 1/N shard of exp_avg / exp_avg_sq; the call is bracketed by stream-ordered barriers across the ranks."""

    def __init__(self, flat: FlatGaussians, lrs: Optional[Dict[str, float]] = None, betas=(0.9, 0.999), eps: float = 1e-15,
                 rank: int = 0, world_size: int = 1, slices: int = 1):
        self.flat, self.betas, self.eps = flat, (float(betas[0]), float(betas[1])), float(eps)
        self.lrs = dict(DEFAULT_LRS)
        if lrs:
            self.lrs.update(lrs)
        self.rank, self.world = int(rank), int(world_size)
        self.t = 0
        n = flat.layout.total
        # Sharded: the flat buffer is cut into `slices` RANGES — the SH rows (81 % of the floats) of successive slices of the
        # Gaussians, the last range also holding the four small fields behind them — and every range is sharded over the
        # ranks on its own.  step() issues them back to back; step_range(k) lets the caller issue range k as soon as the
        # backward has produced the gradients of Gaussian slice k (msgs_train: the exchange of slice k runs under the
        # preprocess backward of slice k + 1).  slice_bounds = the Gaussian boundaries (multiples of 128).
        self.slice_bounds = [0, flat.P]
        self.ranges = [(0, n)]
        if self.world > 1:
            if not flat.symmetric:
                raise RuntimeError("the sharded step needs FlatGaussians(symmetric=True)")
            S = max(1, min(int(slices), flat.P // 128 if flat.P >= 128 else 1))
            gb = [min(flat.P, (flat.P * k // S + 127) // 128 * 128) for k in range(S)] + [flat.P]
            self.slice_bounds = gb
            off = flat.layout.offset["sh"]
            if off != 0:
                raise RuntimeError("the sliced step expects the SH field first in the flat layout")
            self.ranges = [(48 * gb[k], 48 * gb[k + 1] if k + 1 < S else n) for k in range(S)]
            self._state_off, tot = [], 0
            for b, e in self.ranges:
                self._state_off.append(tot)
                tot += int(_lib.msgs_adam_shard_floats(e - b, self.world))
            ns = tot
            self.shard_begin = int(_lib.msgs_adam_shard_begin(self.ranges[0][1] - self.ranges[0][0], self.rank, self.world))
        else:
            ns, self.shard_begin = n, 0
        self.exp_avg = torch.zeros(ns, dtype=torch.float32, device=flat.device)
        self.exp_avg_sq = torch.zeros(ns, dtype=torch.float32, device=flat.device)
        self._fields = (_AdamField * len(GRAD_FIELDS))()
        self._issued = 0

    def set_lr(self, name: str, lr: float):
        if name not in self.lrs:
            raise KeyError(name)
        self.lrs[name] = float(lr)

    def _fill_fields(self):
        lay = self.flat.layout
        for i, (name, _) in enumerate(GRAD_FIELDS):
            lr_name, alt_name, period, alt = _FIELD_LR[name]
            f = self._fields[i]
            f.begin, f.end = lay.offset[name], lay.offset[name] + (lay.size[name] + 3) // 4 * 4
            f.lr = self.lrs[lr_name]
            f.lr_alt = self.lrs[alt_name] if alt_name else 0.0
            f.period, f.alt_count = period, alt
        return self._fields

    def begin_step(self):
        """Sharded, sliced issue: call once per step before the first step_range()."""
        self.t += 1
        self._issued = 0
        self._fill_fields()

    def step_range(self, k: int, ctas: int = 0, barrier_before: bool = True):
        """Range k of this step (begin_step() first) on the CURRENT stream, bracketed by stream-ordered barriers across the
        ranks: before — every rank's gradients of the range are complete; after — every rank's parameter stores have landed."""
        flat = self.flat
        b, e = self.ranges[k]
        st = _vp(torch.cuda.current_stream(flat.device).cuda_stream)
        so = self._state_off[k]
        with torch.cuda.device(flat.device):
            # ONE barrier in front of every range: every rank has finished the previous range's kernel (its parameter stores
            # have landed everywhere) and has this range's gradients complete in its copy; one more behind the last range
            if barrier_before:
                flat.grads_handle.barrier(channel=k & 1)
            rc = _lib.msgs_adam_step_multimem_range(
                _vp(flat.params_mc), _vp(flat.grads_mc), _vp(flat.params.data_ptr()), _vp(self.exp_avg.data_ptr() + 4 * so),
                _vp(self.exp_avg_sq.data_ptr() + 4 * so), flat.layout.total, self._fields, len(GRAD_FIELDS), self.betas[0],
                self.betas[1], self.eps, self.t, b, e, self.rank, self.world, int(ctas), st)
            if k == len(self.ranges) - 1:
                flat.grads_handle.barrier(channel=(k + 1) & 1)
        if rc != 0:
            raise RuntimeError(f"msgs_adam_step_multimem_range failed ({rc})")
        self._issued = max(self._issued, k + 1)

    def finish_step(self, ctas: int = 0, first_barrier: bool = True):
        """Issue the ranges of this step that step_range() has not issued yet (first_barrier=False: the caller has just put
        a cross-rank barrier on this stream behind the complete backward)."""
        for i, k in enumerate(range(self._issued, len(self.ranges))):
            self.step_range(k, ctas, barrier_before=(first_barrier or i > 0))

    def step(self, ctas: int = 0):
        """One optimiser step on the current stream, from flat.grads (this rank's gradients; sharded: summed over the
        ranks inside the kernel).  The caller zeroes the gradients afterwards (zero_grad / the next step's writes)."""
        flat = self.flat
        if not flat.params.is_cuda:
            raise RuntimeError("FusedAdam: CUDA tensors only (there is no CPU path)")
        if self.world > 1:
            self.begin_step()
            self.finish_step(ctas)
            return
        self.t += 1
        fields = self._fill_fields()
        st = _vp(torch.cuda.current_stream(flat.device).cuda_stream)
        n = flat.layout.total
        with torch.cuda.device(flat.device):
            rc = _lib.msgs_adam_step(_vp(flat.params.data_ptr()), _vp(flat.grads.data_ptr()), _vp(self.exp_avg.data_ptr()),
                                     _vp(self.exp_avg_sq.data_ptr()), n, fields, len(GRAD_FIELDS), self.betas[0],
                                     self.betas[1], self.eps, self.t, st)
        if rc != 0:
            raise RuntimeError(f"msgs_adam_step failed ({rc})")


def stats_allreduce_multimem(handle, n_sum: int, n_max: int, n_min: int, rank: int, world: int, device, ctas: int = 0,
                             offset_words: int = 0, barriers: bool = True):
    """Typed-lane all-reduce (sum | max | min) of a symmetric statistics buffer (from word `offset_words` on; a multiple of
    4); barriers around it as for the gradients (barriers=False: the caller brackets it with barriers of its own)."""
    if barriers:
        handle.barrier(channel=0)
    rc = _lib.msgs_stats_allreduce_multimem(_vp(int(handle.multicast_ptr) + 4 * int(offset_words)), n_sum, n_max, n_min, rank,
                                            world, int(ctas), _vp(torch.cuda.current_stream(device).cuda_stream))
    if rc != 0:
        raise RuntimeError(f"msgs_stats_allreduce_multimem failed ({rc})")
    if barriers:
        handle.barrier(channel=1)
