"""(view, scale) sharding of the multi-scale training step and the gradient / statistics exchange that follows it.

The reference is single-process, single-GPU: it renders ONE (view, scale) pair per iteration
(train.py:152-204).  Every `rasterize_gaussians` call is a pure function of the (replicated) Gaussian tensors and one
camera at one resolution (rasterizer_impl.cu:200-367 allocates all state per call), so a step made of several
(view, scale) units shards by units with no data-path collective; the only exchange is AFTER the local backward passes:

  * per-Gaussian gradients of the trainable tensors — xyz 3, f_dc 3 + f_rest 45 (one (P,16,3) tensor here), opacity 1,
    scaling 3, rotation 4 = 59 floats = 236 B per Gaussian (scene/gaussian_model.py:235-246) — SUM;
  * densification statistics consumed by scene/gaussian_model.py: sum of |dL_dmeans2D[:, :2]| and visibility counts
    (:698-701) SUM, max_radii2D (train.py:249) MAX, per-level observed pixel sizes (:663-686) MIN / MAX.

One process per GPU (torchrun); `torch.distributed` supplies the communicator: NCCL over NVLink / NVSwitch on the
B200 box, gloo in the CPU tests (tests/test_sharding.py, world_size 2).  Nothing here touches the rasterizer kernels.
"""
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

# floats per Gaussian of the tensors that training all-reduces, in exchange order (largest first so that the big
# transfer is in flight while the small ones are being enqueued)
GRAD_FIELDS: Tuple[Tuple[str, int], ...] = (("sh", 48), ("rot", 4), ("mean3D", 3), ("scale", 3), ("opacity", 1))
GRAD_FLOATS = sum(n for _, n in GRAD_FIELDS)        # 59


@dataclass(frozen=True)
class Unit:
    """One render unit: view `view` at resolution scale `scale` (image W x H)."""
    view: int
    scale: int
    W: int
    H: int


def unit_cost(P: int, u: Unit, a: float = 1.0, c: float = 6.0) -> float:
    """Relative cost a*P + c*pixels: a per-Gaussian floor (preprocess fwd/bwd, depth sort: independent of the image
    size) plus a per-pixel term for binning and blending.  c/a = 6 fits the measured C3 stage times on B200
    (1.2 M Gaussians: ~0.45 ms floor per unit, ~1.0 ms of blend + placement at 1237x822)."""
    return a * P + c * u.W * u.H


def make_units(n_views: int, scales: Sequence[int], W: int, H: int) -> List[Unit]:
    """All (view, scale) pairs of a step; image sizes as utils/camera_utils.py:38-39 (int(W / s), int(H / s))."""
    return [Unit(v, s, int(W / s), int(H / s)) for v in range(n_views) for s in scales]


def assign_units(units: Sequence[Unit], world_size: int, P: int) -> List[List[int]]:
    """Indices of `units` per rank.  Whole views per rank when the views divide evenly (every rank then gets the same
    unit mix and a view's scales stay together); otherwise longest-processing-time-first on `unit_cost`."""
    if world_size < 1:
        raise ValueError("world_size must be >= 1")
    views = sorted({u.view for u in units})
    per_view: Dict[int, List[int]] = {v: [i for i, u in enumerate(units) if u.view == v] for v in views}
    same_mix = len({tuple(sorted((units[i].scale, units[i].W, units[i].H) for i in idx)) for idx in per_view.values()}) == 1
    if views and len(views) % world_size == 0 and same_mix:
        out: List[List[int]] = [[] for _ in range(world_size)]
        for k, v in enumerate(views):
            out[k % world_size].extend(per_view[v])
        return out
    order = sorted(range(len(units)), key=lambda i: (-unit_cost(P, units[i]), i))
    load = [0.0] * world_size
    out = [[] for _ in range(world_size)]
    for i in order:
        r = min(range(world_size), key=lambda k: (load[k], k))
        out[r].append(i)
        load[r] += unit_cost(P, units[i])
    for lst in out:
        lst.sort()
    return out


class UnitStreams:
    """CUDA streams for the (view, scale) units of one rank's step.

    The units of a step are independent rasterize calls; at the coarse scales of a multi-scale step
    their kernels are far too small to fill 148 SMs (a 19x12 image is 2 tiles) and each forward has one host read-back,
    so running the units back to back leaves the GPU mostly idle.  Unit i runs on stream i mod n instead:

        us.begin()                                   # unit streams wait for what the current stream has enqueued
        for i, unit in enumerate(units):
            with torch.cuda.stream(us.stream(i)):
                ...forward + backward of the unit (accumulating backward: reductions, safe to run concurrently)...
        us.end()                                     # the current stream waits for every unit stream

    n = 1 (or MSGS_UNIT_STREAMS=1) degenerates to the current stream."""

    def __init__(self, device, n: Optional[int] = None):
        import os
        if n is None:
            n = int(os.environ.get("MSGS_UNIT_STREAMS", "8"))
        self.device = device
        # Streams 0 and 1 (default priority) carry the full-resolution unit of a view; the others (high priority) carry the
        # coarse scales.  A coarse unit is a few tiles with very long lists: its blend kernels are a handful of CTAs that run
        # for as long as the full-resolution blend (C3, 19x12: 0.46 ms forward, 0.66 ms backward on 2 tiles).  At equal
        # priority the work distributor hands out ALL 4056 CTAs of the full-resolution kernel before the first CTA of a
        # kernel launched after it, so the long coarse kernels started at the big kernel's tail and became the step's
        # critical path (profiles/r02d_profile_step_c3.log); with priority they start at once on the SMs they need and the
        # full-resolution kernel fills the rest.
        prio = os.environ.get("MSGS_UNIT_PRIORITY", "1") != "0"
        self.streams = ([torch.cuda.Stream(device=device, priority=(-1 if (prio and k >= 2) else 0)) for k in range(n)]
                        if n > 1 else [])
        self._rr = 0

    def plan(self, groups: Sequence[Sequence[int]], pixels: Sequence[int]) -> List["torch.cuda.Stream"]:
        """Stream of every unit of a step.  `groups`: unit indices per view group; pixels[i] = H * W of unit i.  The largest
        unit of a group goes to a default-priority stream (two of them, alternating by group), the others round-robin over
        the high-priority ones."""
        cur = torch.cuda.current_stream(self.device)
        out: List = [cur] * len(pixels)
        if not self.streams:
            return out
        n = len(self.streams)
        lo = self.streams[:2] if n > 2 else self.streams[:1]
        hi = self.streams[len(lo):] or lo
        for gi, grp in enumerate(groups):
            big = max(grp, key=lambda i: (pixels[i], -i))
            for i in grp:
                if i == big:
                    out[i] = lo[gi % len(lo)]
                else:
                    out[i] = hi[self._rr % len(hi)]
                    self._rr += 1
        return out

    def begin(self):
        if self.streams:
            ev = torch.cuda.current_stream(self.device).record_event()
            for s in self.streams:
                s.wait_event(ev)

    def stream(self, i: int):
        return self.streams[i % len(self.streams)] if self.streams else torch.cuda.current_stream(self.device)

    def end(self):
        if self.streams:
            cur = torch.cuda.current_stream(self.device)
            for s in self.streams:
                cur.wait_event(s.record_event())


class GradientExchange:
    """Flat per-field gradient accumulators for one rank's units + the all-reduce that ends the step.

    add(grads)   accumulate one unit's gradients (dict with the GRAD_FIELDS names; shapes (P, ...));
    reduce()     SUM over ranks, in place, one collective per field, largest first; returns the dict of views.
    The buffers are allocated once (P x 59 floats) and reused every step; zero() starts the next step."""

    def __init__(self, P: int, device, group: Optional[dist.ProcessGroup] = None):
        self.P, self.group = P, group
        self.flat = torch.zeros(P * GRAD_FLOATS, dtype=torch.float32, device=device)
        self.views: Dict[str, torch.Tensor] = {}
        off = 0
        for name, n in GRAD_FIELDS:
            self.views[name] = self.flat[off:off + P * n]
            off += P * n
        self.units_added = 0

    def zero(self):
        self.flat.zero_()
        self.units_added = 0

    def shaped(self) -> Dict[str, torch.Tensor]:
        """The accumulators in the shapes `_C.rasterize_gaussians_backward_accumulate` adds into (zero() first)."""
        P, v = self.P, self.views
        return {"sh": v["sh"].view(P, 16, 3), "rot": v["rot"].view(P, 4), "mean3D": v["mean3D"].view(P, 3),
                "scale": v["scale"].view(P, 3), "opacity": v["opacity"].view(P, 1)}

    def add(self, grads: Dict[str, torch.Tensor]):
        for name, _ in GRAD_FIELDS:
            g = grads[name].reshape(-1)
            if g.numel() != self.views[name].numel():
                raise ValueError(f"gradient '{name}' has {g.numel()} elements, expected {self.views[name].numel()}")
            if self.units_added == 0:
                self.views[name].copy_(g)
            else:
                self.views[name].add_(g)
        self.units_added += 1

    def reduce(self, async_op: bool = False):
        """All-reduce (SUM) every field.  With async_op the collectives are only enqueued; wait() completes them."""
        self._work = []
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(self.group) > 1:
            for name, _ in GRAD_FIELDS:
                w = dist.all_reduce(self.views[name], op=dist.ReduceOp.SUM, group=self.group, async_op=async_op)
                if async_op:
                    self._work.append(w)
                if name == "sh":
                    # the four small fields travel as one collective (their slices are contiguous behind "sh")
                    rest = self.flat[self.views["sh"].numel():]
                    w = dist.all_reduce(rest, op=dist.ReduceOp.SUM, group=self.group, async_op=async_op)
                    if async_op:
                        self._work.append(w)
                    break
        return self.views

    def wait(self):
        for w in getattr(self, "_work", []):
            w.wait()
        self._work = []
        return self.views

    def bytes_per_step(self) -> int:
        return self.flat.numel() * 4


class SymmetricGradientExchange:
    """Gradient accumulators in a symmetric buffer (same size and layout on every rank, torch.distributed._symmetric_memory,
    mapped through the NVSwitch multicast object) + the hand-written two-shot NVLS all-reduce over it
    (msgs_allreduce_multimem, csrc/exchange.cu).  The units' backward passes write / accumulate this rank's gradients into
    shaped() with plain local stores; allreduce() then sums the ranks' copies in place.  Field offsets are 16-byte aligned.
    Needs multicast support (NVSwitch); use GradientExchange (NCCL all-reduce) otherwise.  `available()` tells.
    (The sharded TRAINING step does not all-reduce at all: msgs_optim.FusedAdam fuses reduce-scatter + Adam + all-gather.)"""

    @staticmethod
    def available(device) -> bool:
        try:
            import torch.distributed._symmetric_memory as symm_mem  # noqa: F401
            return torch.cuda.is_available() and dist.is_initialized() and dist.get_world_size() > 1
        except Exception:
            return False

    def __init__(self, P: int, device, group: Optional[dist.ProcessGroup] = None):
        import torch.distributed._symmetric_memory as symm_mem
        self.P = P
        self.group = group if group is not None else dist.group.WORLD
        offs, off = {}, 0
        for name, n in GRAD_FIELDS:
            offs[name] = off
            off += (P * n + 3) // 4 * 4                      # keep every field 16-byte aligned
        self.flat = symm_mem.empty(off, dtype=torch.float32, device=device)
        self.handle = symm_mem.rendezvous(self.flat, self.group)
        if not self.handle.multicast_ptr:
            raise RuntimeError("symmetric memory has no multicast mapping on this system")
        self.views = {name: self.flat[offs[name]:offs[name] + P * n] for name, n in GRAD_FIELDS}
        self.flat.zero_()
        self._chan = 0

    def shaped(self) -> Dict[str, torch.Tensor]:
        P, v = self.P, self.views
        return {"sh": v["sh"].view(P, 16, 3), "rot": v["rot"].view(P, 4), "mean3D": v["mean3D"].view(P, 3),
                "scale": v["scale"].view(P, 3), "opacity": v["opacity"].view(P, 1)}

    def bytes_per_step(self) -> int:
        return self.P * GRAD_FLOATS * 4

    # ---- two-shot all-reduce over the multicast mapping (msgs_allreduce_multimem, csrc/exchange.cu) ----
    def zero(self):
        self.flat.zero_()

    def allreduce(self, ctas: int = 0):
        """In-place fp32 sum over the ranks of the whole symmetric buffer, on the current stream: the units' backward
        passes have written / accumulated this rank's gradients into shaped() (plain local stores, no multicast); then
        barrier (every rank's copy is complete) -> hand-written NVLS kernel (rank r: multimem.ld_reduce + multimem.st
        over slice r) -> barrier (every rank's stores have landed; nobody re-zeroes a copy that is still being read)."""
        import ctypes
        lib = _raster_lib()
        self.handle.barrier(channel=0)
        rc = lib.msgs_allreduce_multimem(ctypes.c_void_p(int(self.handle.multicast_ptr)), ctypes.c_size_t(self.flat.numel()),
                                         dist.get_rank(self.group), dist.get_world_size(self.group), int(ctas),
                                         ctypes.c_void_p(torch.cuda.current_stream(self.flat.device).cuda_stream))
        if rc < 0:
            raise RuntimeError(f"msgs_allreduce_multimem failed ({rc})")
        self.handle.barrier(channel=1)
        return self.views


def _raster_lib():
    import ctypes
    import os
    lib = getattr(_raster_lib, "lib", None)
    if lib is None:
        path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libmsgs_raster.so")
        if not os.path.exists(path):
            raise ImportError(f"{path} not found: build it first (python -c 'import __graft_entry__ as g; g.build()')")
        lib = ctypes.CDLL(path)
        lib.msgs_allreduce_multimem.restype = ctypes.c_int
        lib.msgs_allreduce_multimem.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                                ctypes.c_void_p]
        _raster_lib.lib = lib
    return lib


class DensificationStats:
    """Per-step statistics of scene/gaussian_model.py that must agree on every rank before densify / prune decisions
    (SURVEY.md §8e).  Local updates follow the reference's formulas; reduce() makes them global."""

    def __init__(self, P: int, levels: int, device):
        f = dict(dtype=torch.float32, device=device)
        self.grad_accum = torch.zeros((P, levels), **f)          # xyz_gradient_accum[:, lvl]   (gaussian_model.py:698-701)
        self.denom = torch.zeros((P, levels), **f)               # denom[:, lvl]
        self.max_radii2D = torch.zeros((P,), **f)                # train.py:249
        self.px_min = torch.full((P, levels), float("inf"), **f)  # observed pixel sizes per level (gaussian_model.py:663-686)
        self.px_max = torch.full((P, levels), float("-inf"), **f)
        self.seen = torch.zeros((P,), dtype=torch.bool, device=device)   # visibility_filter OR (update_base_gaussian_mask :703-705)

    def update(self, level: int, radii: torch.Tensor, pixel_sizes: torch.Tensor, dL_dmeans2D: torch.Tensor):
        vis = radii > 0
        self.grad_accum[:, level] += torch.where(vis, dL_dmeans2D[:, :2].norm(dim=-1), torch.zeros_like(self.max_radii2D))
        self.denom[:, level] += vis.to(torch.float32)
        self.max_radii2D = torch.maximum(self.max_radii2D, torch.where(vis, radii.to(torch.float32), self.max_radii2D))
        have = pixel_sizes > 0                                   # train.py:296 relies on pixel_sizes > 0 (quirk Q5)
        self.px_min[:, level] = torch.where(have, torch.minimum(self.px_min[:, level], pixel_sizes), self.px_min[:, level])
        self.px_max[:, level] = torch.where(have, torch.maximum(self.px_max[:, level], pixel_sizes), self.px_max[:, level])
        self.seen |= vis

    def reduce(self, group: Optional[dist.ProcessGroup] = None):
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
            return self
        dist.all_reduce(self.grad_accum, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(self.denom, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(self.max_radii2D, op=dist.ReduceOp.MAX, group=group)
        dist.all_reduce(self.px_min, op=dist.ReduceOp.MIN, group=group)
        dist.all_reduce(self.px_max, op=dist.ReduceOp.MAX, group=group)
        seen = self.seen.to(torch.uint8)
        dist.all_reduce(seen, op=dist.ReduceOp.MAX, group=group)
        self.seen = seen.bool()
        return self
