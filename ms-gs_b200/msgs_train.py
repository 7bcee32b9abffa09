"""The multi-scale training step of MS-GS as ONE driver call (SURVEY.md §8 rows f2 + f3 + e).

What train.py does per iteration for ONE (view, scale) pair (train.py:152-260, 416-418) —

    render(...)                                  gaussian_renderer/__init__.py:57-108 (activations + rasterizer)
    loss = (1 - l) L1 + l (1 - SSIM); loss *= 1 if reso_idx == 0 else 0.1; loss.backward()         train.py:208-215
    update_base_gaussian_mask / update_pixel_sizes / max_radii2D / add_densification_stats          train.py:239-250
    optimizer.step(); optimizer.zero_grad(set_to_none=True)                                         train.py:416-418

— `MultiScaleStep.step(views)` does for EVERY (view, scale) unit this rank holds in the step (BASELINE configs[2]: "all
scales per step"; configs[3]: views sharded over the ranks), with these launches:

    1  raw-parameter view-group forward per view  (exp / sigmoid / normalize inside the preprocess kernel; one depth sort)
       + per unit: counting, scans, placement, blend forward on the unit's stream
    1  L1 + SSIM forward for the <= 8 images of a view, per-image weight and mean inside        (msgs_loss)
    1  L1 + SSIM backward for the same images
       per unit: blend backward; 1 preprocess backward per view writing RAW-parameter gradients into the flat buffer
    1  statistics update for all units of a view                                                (msgs_stats)
    1  Adam over the flat parameter buffer                                                      (msgs_optim)

Sharded (one process per GPU, Gaussians replicated in symmetric memory): the statistics observations go into a delta
buffer that ONE typed-lane NVLS kernel reduces (sum / max / min lanes), and the optimiser launch is reduce-scatter + Adam
on the rank's 1/N shard of the state + all-gather of the updated parameters as ONE kernel over the NVSwitch multicast
mapping — there is no separate gradient all-reduce.  The step of N ranks equals the single-GPU step over the union of their
units (gradients and losses are SUMS over units, as a single GPU that renders several units before optimizer.step() forms
them); tools/check_sharded_step.py checks exactly that on 2 GPUs.

No CPU path: everything here drives the CUDA library."""
import os
import sys
from typing import Dict, List, Optional, Sequence

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
if _HERE not in sys.path:
    sys.path.insert(0, _HERE)

import diff_gaussian_rasterization as dgr      # noqa: E402
import msgs_loss                                # noqa: E402
import msgs_optim                               # noqa: E402
import msgs_stats                               # noqa: E402


from msgs_types import MSState, TrainView     # noqa: E402,F401  (plain dataclasses, no CUDA library behind them)


class MultiScaleStep:
    """flat: msgs_optim.FlatGaussians (symmetric=True when world_size > 1); state: MSState.

    stats_what: msgs_stats.PIXEL_SIZES | DENSIFY | BASE_MASK applied to every unit (train.py switches them by iteration
    ranges: pixel sizes from iteration 250, densification statistics until densify_until_iter, the base mask after it at the
    coarsest scale only — `base_mask_level` restricts BASE_MASK to that level)."""

    def __init__(self, flat: "msgs_optim.FlatGaussians", state: MSState, bg: torch.Tensor, sh_degree: int = 3,
                 lrs: Optional[Dict[str, float]] = None, lambda_dssim: float = 0.2, filter_small: bool = True,
                 filter_large: bool = True, fade_size: float = 1.0, scale_modifier: float = 1.0, rank: int = 0,
                 world_size: int = 1, stats_what: int = msgs_stats.PIXEL_SIZES | msgs_stats.DENSIFY,
                 base_mask_level: Optional[int] = None, coarse_loss_weight: float = 0.1, stats_sync_interval: int = 100):
        if not flat.params.is_cuda:
            raise RuntimeError("MultiScaleStep: CUDA tensors only (there is no CPU path)")
        self.flat, self.state, self.bg, self.sh_degree = flat, state, bg, int(sh_degree)
        self.lam, self.fs, self.fl, self.fade = float(lambda_dssim), bool(filter_small), bool(filter_large), float(fade_size)
        self.mod = float(scale_modifier)
        self.rank, self.world = int(rank), int(world_size)
        self.what, self.base_level, self.coarse_w = int(stats_what), base_mask_level, float(coarse_loss_weight)
        # sharded: the exchange + optimiser launch can be issued in MSGS_EXCHANGE_SLICES ranges, each as soon as the backward
        # has produced that slice's gradients.  Default 1 (one launch after the backward): measured on 2 and 8 B200
        # (profiles/r02n_*, r02o_*) the sliced issue overlaps as intended but gains nothing — both kernels want the whole
        # register file / thread slots of an SM, so the preprocess backward slows down by what the exchange takes
        # (84 -> 178 us per slice under a 160 us exchange kernel): 3.69 ms vs 3.60 ms per step at N = 8.
        self.slices = max(1, int(os.environ.get("MSGS_EXCHANGE_SLICES", "1"))) if self.world > 1 else 1
        self.adam = msgs_optim.FusedAdam(flat, lrs, rank=self.rank, world_size=self.world, slices=self.slices)
        self.comm = torch.cuda.Stream(device=flat.device, priority=-1) if self.world > 1 else None
        dev = flat.device
        self.dev = dev
        # leaf views of the flat parameter buffer (same storage): what autograd differentiates
        self.leaf = {k: flat.p[k].detach().requires_grad_(True) for k in ("mean3D", "sh", "opacity", "scale", "rot")}
        self._m2: List[torch.Tensor] = []
        # MSGS_STEP_AUTOGRAD=1: go through the public autograd API instead of the direct calls (same kernels, same results)
        self.use_autograd = os.environ.get("MSGS_STEP_AUTOGRAD", "0") == "1"
        self.delta = None                  # statistics delta buffer (several views per step, or several ranks)
        self._local, self._unsynced = None, 0      # sharded: rank-local densification accumulators since the last sync
        self.stats_sync_interval = int(stats_sync_interval)
        self.last = {}
        self.side = torch.cuda.Stream(device=dev)

    # ---- pieces ---------------------------------------------------------------------------------------------------
    def _means2D(self, n):
        P = self.flat.P
        while len(self._m2) < n:        # gaussian_renderer/__init__.py:27 makes one per render; only its .grad matters
            self._m2.append(torch.zeros((P, 3), dtype=torch.float32, device=self.dev, requires_grad=True))
        for m in self._m2[:n]:
            m.grad = None
        return self._m2[:n]

    def _what_for(self, level):
        w = self.what
        if self.base_level is not None and level != self.base_level:
            w &= ~msgs_stats.BASE_MASK
        return w

    def _slice_done(self, k, n, stream):
        """diff_gaussian_rasterization's slice hook: Gaussian slice k of n has its final gradients in flat.grads (enqueued on
        `stream`): the comm stream takes range k (reduce-scatter + Adam + all-gather of that slice's SH rows; the last range
        also carries the small fields)."""
        if n != len(self.adam.ranges):
            return                        # not sliced the way the optimiser is: finish_step() issues everything afterwards
        self.comm.wait_event(stream.record_event())
        with torch.cuda.stream(self.comm):
            self.adam.step_range(k)

    def render_and_backward(self, views: Sequence[TrainView], before_loss=None, overlap_optimizer=False, after_forward=None):
        """forward of every unit, loss, backward: the raw-parameter gradients are in flat.grads afterwards.  Returns
        (total loss (device scalar), per-unit outputs, per-unit means2D tensors).  `before_loss()` (optional) runs after
        the forward has been enqueued and before the loss is: the place to make the stream wait for ground-truth images
        that are still travelling on a copy stream."""
        st, leaf = self.state, self.leaf
        settings, view_ids, gts, weights = [], [], [], []
        for v in views:
            for (H, W), gt, lvl in zip(v.sizes, v.gts, v.levels):
                settings.append(dgr.GaussianRasterizationSettings(
                    int(H), int(W), v.tanfovx, v.tanfovy, self.bg, self.mod, v.viewmatrix, v.projmatrix, self.sh_degree,
                    v.campos, False, False, self.fs, self.fl, self.fade))
                view_ids.append(v.view_id)
                gts.append(gt)
                weights.append(1.0 if lvl == 0 else self.coarse_w)       # train.py:212-214
        dgr.set_gradient_sink(self.flat.g)
        if overlap_optimizer:
            self.adam.begin_step()
            dgr.set_backward_slices(self.adam.slice_bounds, self._slice_done)
        try:
            if self.use_autograd:
                # the public autograd path (rasterize_gaussians_units -> l1_ssim_loss_batch -> backward())
                for p in leaf.values():
                    p.grad = None             # zero_grad(set_to_none=True), train.py:418 (the sink needs it: ADVICE r01)
                m2 = self._means2D(len(settings))
                units = dgr.rasterize_gaussians_units(
                    leaf["mean3D"], m2, leaf["sh"], leaf["opacity"], leaf["scale"], leaf["rot"], st.max_pixel_sizes,
                    st.min_pixel_sizes, st.occ_multiplier, st.dc_delta, st.base_gaussian_mask, settings, view_ids=view_ids,
                    raw=True)
                if after_forward is not None:
                    after_forward(units)
                if before_loss is not None:
                    before_loss()
                total = None
                for b in range(0, len(units), msgs_loss.MAX_IMAGES):          # one launch pair per <= 8 images
                    sl = slice(b, b + msgs_loss.MAX_IMAGES)
                    t, _ = msgs_loss.l1_ssim_loss_batch([u["render"] for u in units[sl]], gts[sl], weights[sl], self.lam)
                    total = t if total is None else total + t
                total.backward()
                if not dgr.gradient_sink_installed(leaf["mean3D"], leaf["sh"], leaf["opacity"], leaf["scale"], leaf["rot"]):
                    # autograd cloned instead of adopting the views: bring the gradients into the flat buffer
                    for k in ("mean3D", "sh", "opacity", "scale", "rot"):
                        self.flat.g[k].copy_(leaf[k].grad)
                    if overlap_optimizer and self.adam._issued:
                        raise RuntimeError("sliced exchange issued but autograd did not adopt the gradient sink")
                m2 = [m.grad for m in m2]
            else:
                # the same calls without autograd between them: the loss launch pair yields d total / d image directly (the
                # upstream gradient of the training loss is 1), so neither the graph nor the autograd engine sits on the
                # step's critical path (host time between the last read-back and the backward launches: 0.65 -> 0.35 ms)
                p = self.flat.p
                units, handle = dgr.rasterize_gaussians_units_begin(
                    p["mean3D"], p["sh"], p["opacity"], p["scale"], p["rot"], st.max_pixel_sizes, st.min_pixel_sizes,
                    st.occ_multiplier, st.dc_delta, st.base_gaussian_mask, settings, view_ids=view_ids, raw=True)
                if after_forward is not None:
                    after_forward(units)
                if before_loss is not None:
                    before_loss()
                total, grads = None, []
                for b in range(0, len(units), msgs_loss.MAX_IMAGES):
                    sl = slice(b, b + msgs_loss.MAX_IMAGES)
                    t, _, g = msgs_loss.l1_ssim_loss_and_grad_batch([u["render"] for u in units[sl]], gts[sl], weights[sl],
                                                                    self.lam)
                    total = t if total is None else total + t
                    grads += g
                _, m2 = dgr.rasterize_gaussians_units_backward(handle, grads)
        finally:
            dgr.set_gradient_sink(None)
            dgr.set_backward_slices(None, None)
        return total.detach(), units, m2          # m2[i] = unit i's dL_dmeans2D (P,3)

    def _per_view_units(self, views, units, m2):
        per_view, k = [], 0
        for v in views:
            us = []
            for lvl in v.levels:
                u = units[k]
                us.append((lvl, self._what_for(lvl), u["radii"], u["pixel_sizes"], m2[k]))
                k += 1
            per_view.append(us)
        return per_view

    def _stats_observe(self, views, units, m2, which=7, reset=True):
        """this rank's observations of the step -> delta buffer (one launch per view group).  `which` masks the per-unit
        updates: PIXEL_SIZES | BASE_MASK need only the forward's radii / pixel sizes, DENSIFY also dL_dmeans2D (m2)."""
        st, L = self.state, self.state.reso_lvls
        if self.delta is None:
            self.delta = msgs_stats.StatsDelta(self.flat.P, L, self.dev, symmetric=self.world > 1)
        d = self.delta
        if reset:
            d.reset()
        m2 = m2 if m2 is not None else [None] * len(units)
        for us in self._per_view_units(views, units, m2):
            us = [(lvl, w & which, r, px, g) for (lvl, w, r, px, g) in us]
            for b in range(0, len(us), msgs_stats.MAX_UNITS):
                d.add_units(us[b:b + msgs_stats.MAX_UNITS], st.target_reso_lvl)

    def _stats_apply(self):
        """(reduced) delta -> state.  Sharded: only what the NEXT render reads (pixel-size observations -> max /
        min_pixel_sizes, base mask: 3 words per Gaussian) has been reduced; the gradient-norm sums, visibility counts and
        radii (2 L + 1 words per Gaussian, 30 % of the gradient volume at L = 7) are pure sums / maxima that only
        densification reads: they go to rank-local accumulators and are summed over the ranks by sync_statistics() —
        automatically every `stats_sync_interval` steps (densification_interval, train.py:253) and before anything reads
        state.xyz_gradient_accum / denom / max_radii2D."""
        st, d = self.state, self.delta
        base = st.base_gaussian_mask if (self.what & msgs_stats.BASE_MASK) else None
        if self.world == 1:
            d.apply(st.target_reso_lvl, st.max_pixel_sizes, st.min_pixel_sizes, st.max_radii2D, st.xyz_gradient_accum,
                    st.denom, base)
            return
        if self._local is None:
            self._local = {"max_radii2D": torch.zeros_like(st.max_radii2D), "accum": torch.zeros_like(st.xyz_gradient_accum),
                           "denom": torch.zeros_like(st.denom)}
        lo = self._local
        d.apply(st.target_reso_lvl, st.max_pixel_sizes, st.min_pixel_sizes, lo["max_radii2D"], lo["accum"], lo["denom"], base)
        self._unsynced += 1
        if self.stats_sync_interval and self._unsynced >= self.stats_sync_interval:
            self.sync_statistics()

    def update_statistics(self, views: Sequence[TrainView], units, m2):
        """train.py:239-250 for every unit of the step, on the current stream (sharded: including the exchange)."""
        st, L = self.state, self.state.reso_lvls
        if self.world == 1 and len(views) == 1:
            base = st.base_gaussian_mask if (self.what & msgs_stats.BASE_MASK) else None
            for us in self._per_view_units(views, units, m2):
                for b in range(0, len(us), msgs_stats.MAX_UNITS):
                    msgs_stats.update_stats_units(us[b:b + msgs_stats.MAX_UNITS], L, st.target_reso_lvl, st.max_pixel_sizes,
                                                  st.min_pixel_sizes, st.max_radii2D, st.xyz_gradient_accum, st.denom, base)
            return
        # several views and / or several ranks: observations -> delta -> (typed-lane all-reduce) -> apply
        self._stats_observe(views, units, m2)
        if self.world > 1:
            self.delta.allreduce_fast(self.rank, self.world)
        self._stats_apply()

    def sync_statistics(self):
        """Sharded step: fold every rank's local densification accumulators into the replicated state (sum / sum / max over
        the ranks).  Collective: every rank must call it at the same step.  No-op on one rank or when nothing is pending."""
        if self.world == 1 or self._local is None or self._unsynced == 0:
            return
        import torch.distributed as dist
        lo, st = self._local, self.state
        dist.all_reduce(lo["accum"], op=dist.ReduceOp.SUM)
        dist.all_reduce(lo["denom"], op=dist.ReduceOp.SUM)
        dist.all_reduce(lo["max_radii2D"], op=dist.ReduceOp.MAX)
        st.xyz_gradient_accum += lo["accum"]
        st.denom += lo["denom"]
        torch.maximum(st.max_radii2D, lo["max_radii2D"], out=st.max_radii2D)
        for t in lo.values():
            t.zero_()
        self._unsynced = 0

    # ---- the step ---------------------------------------------------------------------------------------------------
    def step(self, views: Sequence[TrainView], optimizer: bool = True, statistics: bool = True, before_loss=None):
        """One training step over this rank's views.  Returns the rank's summed loss as a device scalar (no host sync)."""
        sharded = optimizer and self.world > 1
        do_stats = bool(statistics and self.what)
        main = torch.cuda.current_stream(self.dev)
        early = sharded and do_stats and self.slices == 1
        FAST = msgs_stats.PIXEL_SIZES | msgs_stats.BASE_MASK

        def after_forward(units):
            # Sharded: what the NEXT render needs from this step's statistics — the pixel-size observations and the base
            # mask — depends on the forward's radii / pixel sizes only, so those lanes are observed AND reduced over the ranks
            # (typed-lane NVLS kernel, 3 words per Gaussian, its own barriers) on the side stream WHILE the loss and the
            # backward run.  Behind the backward only the exchange + optimiser kernel is left on the critical path
            # (profiles/r02t_trace_n8_before.txt: the statistics chain used to be the step's tail: its 80 us observation
            # kernel took 250-680 us beside the NVLS kernel and its barriers ended 0.17 ms after the optimiser's).
            self.side.wait_event(main.record_event())
            with torch.cuda.stream(self.side):
                self._stats_observe(views, units, None, which=FAST, reset=True)
                self.delta.allreduce_fast(self.rank, self.world)

        loss, units, m2 = self.render_and_backward(views, before_loss, overlap_optimizer=sharded,
                                                   after_forward=after_forward if early else None)
        done = None
        if early:
            ev = main.record_event()
            self.side.wait_event(ev)
            with torch.cuda.stream(self.side):        # rank-local densification sums + apply, under the optimiser kernel
                self._stats_observe(views, units, m2, which=msgs_stats.DENSIFY, reset=False)
                self._stats_apply()
                done = self.side.record_event()
            self.comm.wait_event(ev)
            with torch.cuda.stream(self.comm):
                self.adam.finish_step()
            main.wait_event(self.comm.record_event())
        else:
            # the statistics update reads what the backward left (radii, pixel sizes, dL_dmeans2D) and touches none of the
            # optimiser's buffers: it runs on a side stream under the Adam launch
            if do_stats:
                self.side.wait_event(main.record_event())
                with torch.cuda.stream(self.side):
                    self.update_statistics(views, units, m2)
                    done = self.side.record_event()
            if sharded:
                # (sliced issue: the ranges the backward has not issued follow the whole backward)
                self.comm.wait_event(main.record_event())
                with torch.cuda.stream(self.comm):
                    self.adam.finish_step()
                main.wait_event(self.comm.record_event())
            elif optimizer:
                self.adam.step()
        if done is not None:
            main.wait_event(done)
        self.last = {"units": units, "loss": loss}
        return loss

    # ---- snapshots (checkpoint / resume of the step's state; bench.py restarts every timed region from the same state) ----
    def snapshot(self):
        st = self.state
        return {"params": self.flat.params.clone(), "exp_avg": self.adam.exp_avg.clone(), "exp_avg_sq": self.adam.exp_avg_sq.clone(),
                "t": self.adam.t, "state": {k: getattr(st, k).clone() for k in (
                    "target_reso_lvl", "max_pixel_sizes", "min_pixel_sizes", "base_gaussian_mask", "max_radii2D",
                    "xyz_gradient_accum", "denom")}}

    def restore(self, snap):
        """In place (the symmetric buffers keep their addresses); every rank restores its own replica and shard."""
        self.flat.params.copy_(snap["params"])
        self.adam.exp_avg.copy_(snap["exp_avg"])
        self.adam.exp_avg_sq.copy_(snap["exp_avg_sq"])
        self.adam.t = int(snap["t"])
        for k, v in snap["state"].items():
            getattr(self.state, k).copy_(v)
        if self._local is not None:
            for t in self._local.values():
                t.zero_()
        self._unsynced = 0
