#!/usr/bin/env python
"""bench.py — MS-GS multi-scale training step (and rasterizer fwd+bwd) throughput on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl mine|reference] [--config C3|C4|C2|C5|C1]

Default workload C3 = BASELINE.json configs[2], the configuration north_star's target is quoted on: 1.2 M Gaussians,
--ms_train_max_scale 6 (7 scales 1237x822 .. 19x12), filter_small + filter_large, ALL scales per step.  One "step" =
one iteration of the train.py loop for every (view, scale) unit of the step: render forward -> L1 + SSIM loss (x0.1 on the
coarse scales) -> backward -> per-level statistics (update_pixel_sizes, max_radii2D, add_densification_stats) -> Adam
(train.py:152-260, 416-418).  Every step renders another view (a pool of 8 yawed cameras per rank), so num_rendered varies
as in training.  N > 1 (torchrun): every rank renders its own view of the replicated Gaussians (weak scaling, the
(view, scale) sharding of SURVEY.md §8e); the statistics deltas are reduced with one typed-lane NVLS kernel and the
optimiser launch is reduce-scatter + Adam on the rank's shard + all-gather in one kernel; the line additionally carries
the C4 strong-scaling figure (3 M Gaussians, 8 views x 6 scales per step) under `extras`.
--config C2 / C5 / C1: the rasterizer alone (one forward + backward with a dense random grad_out_color; C5 forward only).

Prints ONE JSON line (rank 0).  `value` = whole-job it/s with every input resident in HBM; `e2e` = the same step through
the public API with the step's host inputs (camera, ground-truth images) copied from pinned host memory and the loss
read back inside the timed region; `roofline` = the dominant kernel of the scale-1 unit (the C2-shaped unit: per-stage work
is only well defined per unit), timed live with CUDA events on the launching stream, plus the whole-unit fraction;
`cpu_baseline` = the CPU oracle (port of the reference path) on the host cores; `extras` (N = 1) = the C2 and C5 figures.
--impl reference runs the UNMODIFIED reference CUDA rasterizer (oracle/_ref, built in place for sm_100a) inside the
reference's own training iteration restated with torch ops (harness/ref_train.py) — the speed bar BASELINE.json names.
"""
import argparse
import contextlib
import ctypes
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "ms-gs_b200"))
from harness import scenes  # noqa: E402

METRIC = "fwd_bwd_iterations_per_s"
UNIT = "it/s"
POOL_VIEWS = 8          # training views per rank the steps rotate over (train.py:193-204 pops a random camera per iteration)


# ------------------------------------------------------------------------------------------------------------
def pin_rank_to_cores(local_rank, local_world):
    """One process per GPU, each on its own physical cores (what numactl / torchrun launch scripts do): every rank's host
    thread waits for read-backs by spinning, and the sharded step is coupled by barriers, so a rank whose thread is migrated or
    shares a core with another rank's makes ALL ranks wait.  Whole cores (all hyper-thread siblings) per rank; MSGS_PIN_CORES=0
    switches it off.  Returns the CPU set, or None."""
    if os.environ.get("MSGS_PIN_CORES", "1") == "0" or not hasattr(os, "sched_setaffinity"):
        return None
    try:
        avail = sorted(os.sched_getaffinity(0))
        cores, seen = [], set()
        for c in avail:
            if c in seen:
                continue
            try:
                sib = open(f"/sys/devices/system/cpu/cpu{c}/topology/thread_siblings_list").read().strip()
                ids = set()
                for part in sib.split(","):
                    a, _, b = part.partition("-")
                    ids.update(range(int(a), int(b or a) + 1))
            except Exception:
                ids = {c}
            ids &= set(avail)
            seen |= ids
            cores.append(sorted(ids))
        per = len(cores) // max(1, local_world)
        if per < 1:
            return None
        mine = sorted(x for core in cores[local_rank * per:(local_rank + 1) * per] for x in core)
        os.sched_setaffinity(0, mine)
        return mine
    except Exception:
        return None


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)", float(d.get("sm_max_mhz", 1965.0))
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)", 1965.0


def ncu_traffic(stage):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `stage` from the committed ncu --set full capture of the
    scale-1 unit (profiles/ncu_traffic_c2.json, written from the .ncu-rep each round), or None."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic_c2.json")
    if not os.path.exists(p):
        return None
    try:
        return json.load(open(p))["stages"].get(stage)
    except Exception:
        return None


class ClockSampler:
    """SM clock / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe).  In-process NVML
    (pynvml) in a thread, initialised before the warm-up so that no driver initialisation lands inside the timed
    region; falls back to an `nvidia-smi -lms` child process (started early, first sample awaited)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx, self.rows, self.proc, self.nvml, self.thread = gpu_index, [], None, None, None
        self.stop_flag = threading.Event()
        self.period = float(os.environ.get("MSGS_CLOCK_PERIOD_S", "0.01"))
        self.wake = threading.Event()
        self.active = threading.Event()

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[self.idx]) if vis and all(v.strip().isdigit() for v in vis.split(",")) else self.idx
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.nvml = pynvml
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.thread = threading.Thread(target=self._poll_nvml, daemon=True)
            self.thread.start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.idx}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read_smi, daemon=True)
            self.thread.start()
            t0 = time.time()
            while not self.rows and time.time() - t0 < 5.0:      # driver/NVML start-up happens here, not in the timed region
                time.sleep(0.05)
        except Exception:
            self.proc = None

    def mark(self, on):
        """only samples taken while marked (the timed region) are reported"""
        (self.active.set if on else self.active.clear)()
        if on:
            self.wake.set()          # take a sample right at the start of the timed region

    def _poll_nvml(self):
        n = self.nvml
        bits = {"hw_slowdown": n.nvmlClocksEventReasonHwSlowdown if hasattr(n, "nvmlClocksEventReasonHwSlowdown") else 0x8,
                "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}
        while not self.stop_flag.is_set():
            try:
                mhz = float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM))
                try:
                    r = int(n.nvmlDeviceGetCurrentClocksEventReasons(self.handle))
                except Exception:
                    r = int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle))
                if self.active.is_set():
                    self.rows.append((mhz, self.max_mhz, [k for k, b in bits.items() if r & b]))
            except Exception:
                pass
            self.wake.wait(timeout=self.period)
            self.wake.clear()

    def _read_smi(self):
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.proc.stdout:
            c = [x.strip() for x in line.split(",")]
            try:
                row = (float(c[1]), float(c[2]), [n for n, v in zip(names, c[5:9]) if v.lower().startswith("active")])
            except Exception:
                continue
            if self.active.is_set() or not self.rows:
                self.rows.append(row)

    def stop(self):
        if self.nvml is None and self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml / nvidia-smi unavailable"]}
        self.stop_flag.set()
        if self.proc is not None:
            time.sleep(0.15)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
        elif self.thread is not None:
            self.thread.join(timeout=1)
        sm = [r[0] for r in self.rows]
        mx = [r[1] for r in self.rows]
        reasons = sorted({x for r in self.rows for x in r[2]})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm), "source": "nvml" if self.nvml else "nvidia-smi"}


# name -> scene shape (harness/scenes.py), resolution scales per view, views per step ("world" = one per GPU: weak
# scaling; a number = fixed total: strong scaling), MS-GS state + filters (variant B of SURVEY.md §8d), kind:
# "train" = the whole training iteration per unit, "raster" = rasterizer forward (+ backward) only
CFG = {
    "C1": dict(scene="C1", seed=1235, scales=[1], views="world", ms=False, forward_only=False, kind="raster"),
    "C2": dict(scene="C2", seed=1236, scales=[1], views="world", ms=False, forward_only=False, kind="raster"),
    "C3": dict(scene="C2", seed=1236, scales=scenes.MS_SCALES, views="world", ms=True, forward_only=False, kind="train"),
    "C4": dict(scene="C4", seed=1238, scales=[1, 2, 4, 8, 16, 32], views=8, ms=True, forward_only=False, kind="train"),
    # C5: the same on-screen footprint statistics as C2 (sigma ~ 5 px) at 3840 px width
    "C5": dict(scene="C5", seed=1239, scales=[1], views="world", ms=True, forward_only=True, kind="raster",
               splat_scale=0.004 * 1237 / 3840),
}
WORKLOADS = {"C1": "C1: 16384 Gaussians SH3, 200x200, rasterizer fwd+bwd",
             "C2": "C2 (BASELINE configs[1]): 1.2M Gaussians SH3, 1237x822 camera, single-scale rasterizer fwd+bwd",
             "C3": "C3 (BASELINE configs[2]): 1.2M Gaussians, --ms_train --ms_train_max_scale 6 --filter_small (+ filter_large), "
                   "all 7 scales 1237x822..19x12 of one view per step: render fwd -> L1+SSIM loss -> bwd -> per-level "
                   "statistics -> Adam; a different view every step",
             "C4": "C4 (BASELINE configs[3]): 3M Gaussians, 8 views x 6 scales (1237x822..38x25) per step sharded as (view, scale) "
                   "units by whole views, filter_small+filter_large, full training iteration, gradients + statistics reduced",
             "C5": "C5 (BASELINE configs[4]): 6M Gaussians, 3840x2160, filter_small+filter_large, forward only (viewer / render.py path)"}


def view_camera(W, H, v):
    """camera of global view index v: view 0 looks down +z, the others are yawed by -20..+20 degrees around the scene centre"""
    return scenes.make_camera(W, H, yaw_deg=0.0 if v == 0 else (-20.0 + 40.0 * (v % 8) / 7.0))


def pool_camera(W, H, v):
    """Training-view pool of the weak-scaling configs: 8 cameras yawed by -20..+20 degrees around the scene centre (their cost
    differs by ~10 %: num_rendered changes with the yaw)."""
    return scenes.make_camera(W, H, yaw_deg=-20.0 + 40.0 * (v % 8) / 7.0)


def build_workload(cfg, rank, world, dev):
    """Scene (replicated on every rank) + this rank's (view, scale) units as CUDA tensors.  Returns dict."""
    import msgs_sharding
    c = CFG[cfg]
    P, W, H = scenes.CONFIGS[c["scene"]]
    sc = scenes.make_scene(P, W, H, seed=c["seed"], splat_scale=c.get("splat_scale", 0.004))
    n_views = world if c["views"] == "world" else int(c["views"])
    units = msgs_sharding.make_units(n_views, c["scales"], W, H)
    mine = msgs_sharding.assign_units(units, world, P)[rank]
    cams = [view_camera(W, H, units[i].view).at_scale(units[i].scale) for i in mine]
    tt = lambda a, dt=torch.float32: torch.from_numpy(np.ascontiguousarray(a)).to(device=dev, dtype=dt)
    t = dict(means3D=tt(sc.means3D), shs=tt(sc.shs), opacities=tt(sc.opacities), scales=tt(sc.scales),
             rotations=tt(sc.rotations), max_pixel_sizes=tt(sc.max_pixel_sizes), min_pixel_sizes=tt(sc.min_pixel_sizes),
             base_mask=tt(sc.base_mask, torch.bool), occ_multiplier=tt(sc.occ_multiplier), dc_delta=tt(sc.dc_delta),
             bg=tt(sc.bg), sh_degree=sc.sh_degree, scale_modifier=sc.scale_modifier)
    return dict(scene=sc, cams=cams, t=t, P=P, W=W, H=H, scales=c["scales"], ms=c["ms"], forward_only=c["forward_only"],
                n_views=n_views, units_total=len(units), unit_ids=mine, probe_cam=view_camera(W, H, 0),
                view_of_unit=[units[i].view for i in mine],
                scaling="weak" if c["views"] == "world" else "strong")


def apply_ms_variant(wl, Runner, dev):
    """README configuration (variant B of SURVEY.md §8d): MS-GS min/max pixel sizes from a scale-1 warm-up render of view 0
    (the same on every rank, so the replicated scene state stays identical)."""
    probe = Runner(dict(wl, cams=[wl["probe_cam"]], ms=False, forward_only=True), dev)
    probe.step()
    ps = probe.pixel_sizes().cpu().numpy()
    scenes.apply_ms_state(wl["scene"], ps, 1236)
    sc = wl["scene"]
    wl["t"].update(max_pixel_sizes=torch.from_numpy(sc.max_pixel_sizes).to(dev), min_pixel_sizes=torch.from_numpy(sc.min_pixel_sizes).to(dev),
                   base_mask=torch.from_numpy(sc.base_mask).to(dev))
    del probe


# ------------------------------------------------------------------------------------------------------------
class MineRunner:
    name = "mine"

    def __init__(self, wl, dev):
        import diff_gaussian_rasterization as dgr
        self.dgr, self.C, self.wl, self.dev = dgr, dgr._C, wl, dev
        self.lib = dgr._C.lib
        self.lib.msgs_launch_count.restype = ctypes.c_longlong
        self.lib.msgs_profile_stage_name.restype = ctypes.c_char_p
        self.empty = torch.empty(0)
        t = wl["t"]
        self.cam_t = []
        for cam in wl["cams"]:
            self.cam_t.append(dict(view=torch.from_numpy(cam.viewmatrix).to(dev), proj=torch.from_numpy(cam.projmatrix).to(dev),
                                   campos=torch.from_numpy(cam.campos).to(dev), cam=cam))
        g = torch.Generator(device="cpu").manual_seed(99)
        self.dpix = [torch.randn((3, c["cam"].H, c["cam"].W), generator=g).to(dev) for c in self.cam_t]
        self.fs = self.fl = bool(wl["ms"])          # multi-scale configs run the README flags (variant B)
        self.fwd_only = bool(wl.get("forward_only"))
        self.last = None
        import msgs_sharding
        self.unit_streams = msgs_sharding.UnitStreams(dev)
        # view groups: consecutive units that share the camera and differ only in the image size (the scales of one
        # view) go through msgs_forward_group_begin — one preprocess launch + one depth sort for the group
        self.groups = []
        use_groups = os.environ.get("MSGS_VIEW_GROUPS", "1") != "0"
        for i, c in enumerate(self.cam_t):
            cam = c["cam"]
            if use_groups and self.groups:
                j = self.groups[-1][0]
                ref = self.cam_t[j]["cam"]
                if (len(self.groups[-1]) < self.C.MAX_GROUP and np.array_equal(ref.viewmatrix, cam.viewmatrix)
                        and np.array_equal(ref.projmatrix, cam.projmatrix) and np.array_equal(ref.campos, cam.campos)
                        and ref.tanfovx == cam.tanfovx and ref.tanfovy == cam.tanfovy):
                    self.groups[-1].append(i)
                    continue
            self.groups.append([i])

    def launches(self):
        return int(self.lib.msgs_launch_count())

    def step(self, accum=None, grad_out=None):
        """device-resident fwd(+bwd) for every (view, scale) unit of this rank.  Returns each unit's gradient dict, or —
        with `accum` (msgs_sharding.GradientExchange.shaped()) — adds the units' training gradients into it."""
        t, C, e = self.wl["t"], self.C, self.empty
        sc = self.wl["scene"]
        outs = []
        # several units per step: unit i runs on stream i mod n (msgs_sharding.UnitStreams) — needs the accumulating
        # backward (reductions into shared accumulators) or no backward at all
        us = self.unit_streams if (len(self.cam_t) > 1 and (accum is not None or self.fwd_only)) else None
        if us is not None:
            us.begin()
        try:
            self._units(us, accum, grad_out, outs)
        finally:
            if us is not None:
                us.end()
        return outs

    def _units(self, us, accum, grad_out, outs):
        t, C, e = self.wl["t"], self.C, self.empty
        sc = self.wl["scene"]
        ctx = (lambda i: torch.cuda.stream(us.stream(i))) if us is not None else (lambda i: contextlib.nullcontext())
        units = list(zip(self.cam_t, self.dpix))

        def begin(ct):
            cam = ct["cam"]
            return C.rasterize_gaussians_begin(t["bg"], t["means3D"], e, t["max_pixel_sizes"], t["min_pixel_sizes"],
                                               t["opacities"], t["occ_multiplier"], t["dc_delta"], t["scales"], t["rotations"],
                                               sc.scale_modifier, e, t["base_mask"], ct["view"], ct["proj"], cam.tanfovx,
                                               cam.tanfovy, cam.H, cam.W, t["shs"], sc.sh_degree, ct["campos"], False,
                                               self.fs, self.fl, 0.0, False)

        # with unit streams: every unit's forward kernels go out first (msgs_forward_begin), then the read-backs and
        # the backward passes; without: unit after unit
        pending = []
        if us is not None:
            pending = self._begin_all(us)
            if (accum is not None and not self.fwd_only and any(len(g) > 1 for g in self.groups)
                    and os.environ.get("MSGS_GROUP_BWD", "1") != "0"):
                return self._units_grouped(us, accum, pending)
        for i, (ct, dpix) in enumerate(units):
            with ctx(i):
                cam = ct["cam"]
                f = (pending[i] if us is not None else begin(ct)).finish()
                R, color, acc, depth, radii, px, gb, bb, ib = f
                self.last = dict(R=R, n_contrib_buf=(ib, cam), radii=radii, px=px, color=color)
                if self.fwd_only:
                    continue
                if accum is not None:
                    C.rasterize_gaussians_backward_accumulate(
                        t["bg"], t["means3D"], radii, e, px, t["max_pixel_sizes"], t["min_pixel_sizes"], t["scales"],
                        t["rotations"], sc.scale_modifier, e, t["base_mask"], ct["view"], ct["proj"], cam.tanfovx, cam.tanfovy,
                        dpix, t["shs"], sc.sh_degree, ct["campos"], gb, R, bb, ib, self.fs, self.fl, 0.0, False, accum,
                        want_means2D=True)
                    continue
                out = C.rasterize_gaussians_backward(t["bg"], t["means3D"], radii, e, px, t["max_pixel_sizes"],
                                                     t["min_pixel_sizes"], t["scales"], t["rotations"], sc.scale_modifier,
                                                     e, t["base_mask"], ct["view"], ct["proj"], cam.tanfovx, cam.tanfovy,
                                                     dpix, t["shs"], sc.sh_degree, ct["campos"], gb, R, bb, ib, self.fs,
                                                     self.fl, 0.0, False, out=grad_out)
                # grads that training all-reduces (names of msgs_sharding.GRAD_FIELDS)
                outs.append({"mean3D": out[5], "sh": out[7], "opacity": out[2], "scale": out[8], "rot": out[9]})

    def _units_grouped(self, us, accum, pending):
        """view groups, accumulating backward: per unit read-back + blend backward on the unit's stream, then ONE
        preprocess backward per group (msgs_backward_group_accumulate) on the group's first stream"""
        t, C, e = self.wl["t"], self.C, self.empty
        sc = self.wl["scene"]
        dev = self.dev

        def bwd(i, f, part):
            ct, dpix = self.cam_t[i], self.dpix[i]
            cam = ct["cam"]
            R, color, acc, depth, radii, px, gb, bb, ib = f
            return C.rasterize_gaussians_backward_accumulate(
                t["bg"], t["means3D"], radii, e, px, t["max_pixel_sizes"], t["min_pixel_sizes"], t["scales"], t["rotations"],
                sc.scale_modifier, e, t["base_mask"], ct["view"], ct["proj"], cam.tanfovx, cam.tanfovy, dpix, t["shs"],
                sc.sh_degree, ct["campos"], gb, R, bb, ib, self.fs, self.fl, 0.0, False, accum, want_means2D=(part == 3),
                part=part)

        for grp in self.groups:
            fin, evs = [], []
            for i in grp:
                with torch.cuda.stream(us.stream(i)):
                    f = pending[i].finish()
                    fin.append(f)
                    self.last = dict(R=f[0], n_contrib_buf=(f[8], self.cam_t[i]["cam"]), radii=f[4], px=f[5], color=f[1])
                    bwd(i, f, 1 if len(grp) > 1 else 3)
                    evs.append(torch.cuda.current_stream(dev).record_event())
            if len(grp) == 1:
                continue
            ct0 = self.cam_t[grp[0]]
            cam0 = ct0["cam"]
            with torch.cuda.stream(us.stream(grp[0])):
                st0 = torch.cuda.current_stream(dev)
                for ev in evs[1:]:
                    st0.wait_event(ev)
                C.rasterize_gaussians_group_backward_accumulate(
                    [(self.cam_t[i]["cam"].H, self.cam_t[i]["cam"].W, f[4], f[5], f[6]) for i, f in zip(grp, fin)],
                    t["means3D"], t["max_pixel_sizes"], t["min_pixel_sizes"], t["scales"], t["rotations"], sc.scale_modifier, e,
                    t["base_mask"], ct0["view"], ct0["proj"], cam0.tanfovx, cam0.tanfovy, t["shs"], sc.sh_degree, ct0["campos"],
                    self.fs, self.fl, False, accum, want_means2D=True)

    def _begin_all(self, us):
        """every unit's forward kernels enqueued (no read-back yet): view groups through the group entry point, single
        units through rasterize_gaussians_begin, each unit on its stream"""
        t, C, e = self.wl["t"], self.C, self.empty
        sc = self.wl["scene"]
        pending = [None] * len(self.cam_t)
        for grp in self.groups:
            ct0 = self.cam_t[grp[0]]
            cam0 = ct0["cam"]
            if len(grp) == 1:
                i = grp[0]
                with torch.cuda.stream(us.stream(i)):
                    pending[i] = C.rasterize_gaussians_begin(
                        t["bg"], t["means3D"], e, t["max_pixel_sizes"], t["min_pixel_sizes"], t["opacities"],
                        t["occ_multiplier"], t["dc_delta"], t["scales"], t["rotations"], sc.scale_modifier, e, t["base_mask"],
                        ct0["view"], ct0["proj"], cam0.tanfovx, cam0.tanfovy, cam0.H, cam0.W, t["shs"], sc.sh_degree,
                        ct0["campos"], False, self.fs, self.fl, 0.0, False)
                continue
            units = [(t["bg"], self.cam_t[i]["cam"].H, self.cam_t[i]["cam"].W, us.stream(i)) for i in grp]
            pend = C.rasterize_gaussians_group_begin(
                units, t["means3D"], e, t["max_pixel_sizes"], t["min_pixel_sizes"], t["opacities"], t["occ_multiplier"],
                t["dc_delta"], t["scales"], t["rotations"], sc.scale_modifier, e, t["base_mask"], ct0["view"], ct0["proj"],
                cam0.tanfovx, cam0.tanfovy, t["shs"], sc.sh_degree, ct0["campos"], False, self.fs, self.fl, 0.0, False)
            for i, pf in zip(grp, pend):
                pending[i] = pf
        return pending

    def pixel_sizes(self):
        return self.last["px"]

    def stats(self):
        """R, visible count and sum of n_contrib of the last unit (implementation-independent work counters)."""
        ib, cam = self.last["n_contrib_buf"]
        o4 = (ctypes.c_uint64 * 4)()
        self.lib.msgs_image_layout(ib.data_ptr(), cam.W, cam.H, self.wl["P"], o4)
        off = o4[1] - ib.data_ptr()
        ncon = ib[off:off + 4 * cam.W * cam.H].view(torch.int32)
        return dict(R=int(self.last["R"]), visible=int((self.last["radii"] > 0).sum()), sum_n_contrib=int(ncon.long().sum()))


class RefRunner:
    """The unmodified reference CUDA rasterizer (oracle/_ref/libref_rasterizer.so) on the same inputs."""
    name = "reference"

    def __init__(self, wl, dev):
        from harness import refcuda_bind
        if not refcuda_bind.available():
            raise FileNotFoundError(refcuda_bind.SO)
        self.rb, self.wl, self.dev = refcuda_bind, wl, dev
        t = wl["t"]
        self.units = []
        g = torch.Generator(device="cpu").manual_seed(99)
        shared = refcuda_bind.RefRasterizer(sync=False)
        for cam in wl["cams"]:
            tt = dict(t)
            tt.update(viewmatrix=torch.from_numpy(cam.viewmatrix).to(dev), projmatrix=torch.from_numpy(cam.projmatrix).to(dev),
                      campos=torch.from_numpy(cam.campos).to(dev))
            # ONE rasterizer object for all units (they run one after the other): each object keeps its state buffers
            # (24 B per instance + sort temp), 48 of them do not fit 180 GB at C4
            self.units.append((tt, cam, torch.randn((3, cam.H, cam.W), generator=g).to(dev), shared))
        self.fs = self.fl = bool(wl["ms"])
        self.fwd_only = bool(wl.get("forward_only"))
        self.last = None

    def launches(self):
        return 0

    def step(self):
        outs = []
        for tt, cam, dpix, r in self.units:
            r.forward(tt, cam, self.fs, self.fl)
            self.last = r
            if self.fwd_only:
                continue
            g = r.backward(dpix)
            outs.append({k: g[k] for k in ("mean3D", "sh", "opacity", "scale", "rot")})
        return outs

    def pixel_sizes(self):
        return self.last.pixel_sizes

    def stats(self):
        s = self.last.state()
        return dict(R=int(self.last.R), visible=int((self.last.radii > 0).sum()), sum_n_contrib=int(s["n_contrib"].long().sum()))


# ------------------------------------------------------------------------------------------------------------
MIN_WARM_S = float(os.environ.get("MSGS_MIN_WARM_S", "0.4"))


def timed_loop(step_fn, steps, warmup, dist, sampler=None, on_start=None, on_end=None):
    # W warm-up steps, and keep stepping until the GPU has been under load for MIN_WARM_S: the first steps after an
    # idle period (scene generation on the host takes seconds) run at ramping clocks / power state on some boxes
    # (every rank must run the same number of steps — the step holds collectives — so multi-GPU runs use a count)
    t0, n = time.time(), 0
    while n < warmup or (n < 25 if dist else (time.time() - t0 < MIN_WARM_S and n < 100 * max(1, warmup))):
        step_fn()
        n += 1
        if n >= warmup:
            torch.cuda.synchronize()
    torch.cuda.synchronize()
    if dist:
        torch.distributed.barrier()
    torch.cuda.synchronize()
    if on_start is not None:
        on_start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if sampler is not None:
        sampler.mark(True)
    e0.record()
    for _ in range(steps):
        step_fn()
    if on_end is not None:
        on_end()            # e.g. make the timed stream wait for copies still in flight on a side stream
    e1.record()
    torch.cuda.synchronize()
    if sampler is not None:
        sampler.mark(False)
    ms = e0.elapsed_time(e1)
    if dist:
        tm = torch.tensor([ms], device="cuda")
        torch.distributed.all_reduce(tm, op=torch.distributed.ReduceOp.MAX)
        torch.distributed.barrier()
        ms = float(tm.item())
    return ms


# ------------------------------------------------------------------------------------------------------------
# training configs (C3, C4): the whole train.py iteration per (view, scale) unit
def _logit(x):
    x = np.clip(x.astype(np.float64), 1e-6, 1 - 1e-6)
    return np.log(x / (1 - x)).astype(np.float32)


class TrainSetup:
    """Everything both arms share: the raw (pre-activation) Gaussian parameters as GaussianModel stores them, the MS-GS
    per-Gaussian state, this rank's training views with their ground-truth images at every scale (device resident for
    `value`, pinned host copies for `e2e`), and which views a step renders."""

    def __init__(self, cfg, rank, world, dev, render_fn_factory, Runner):
        c = CFG[cfg]
        self.cfg, self.rank, self.world, self.dev = cfg, rank, world, dev
        wl = build_workload(cfg, rank, world, dev)
        apply_ms_variant(wl, Runner, dev)
        self.wl = wl
        sc = wl["scene"]
        P, W, H = wl["P"], wl["W"], wl["H"]
        self.P, self.W, self.H, self.scales = P, W, H, list(c["scales"])
        self.L = len(self.scales)
        rng = np.random.default_rng(c["seed"] + 31)
        tt = lambda a, dt=torch.float32: torch.from_numpy(np.ascontiguousarray(a)).to(device=dev, dtype=dt)
        # raw parameters (scene/gaussian_model.py:127-155 inverts to the scene's activated values; quaternions get a
        # random norm so that F.normalize has work to do)
        qn = rng.uniform(0.5, 2.0, (P, 1)).astype(np.float32)
        self.raw = {"xyz": tt(sc.means3D), "f_dc": tt(sc.shs[:, :1, :]), "f_rest": tt(sc.shs[:, 1:, :]),
                    "opacity": tt(_logit(sc.opacities)), "scaling": tt(np.log(sc.scales)), "rotation": tt(sc.rotations * qn)}
        # Gaussians with a max pixel size are the "large" ones of a coarser level (variant B: 15 %)
        target = np.where(sc.max_pixel_sizes > 0, rng.integers(1, max(2, self.L), P), 0).astype(np.int64)
        self.target = tt(target, torch.int64)
        self.bg = wl["t"]["bg"]
        self.sh_degree = sc.sh_degree
        # views: weak scaling -> a pool of POOL_VIEWS cameras, one per step (rank r renders view (k + r) mod POOL at step k);
        # strong scaling -> this rank's share of the step's fixed views, all of them every step
        if wl["scaling"] == "weak":
            self.view_ids = list(range(POOL_VIEWS))
            self.views_per_step = 1
        else:
            self.view_ids = sorted(set(wl["view_of_unit"]))
            self.views_per_step = len(self.view_ids)
        render = render_fn_factory(self)
        g = torch.Generator(device="cpu").manual_seed(4321 + c["seed"])
        self.views, self.host = {}, {}
        import msgs_types
        weak = wl["scaling"] == "weak"
        for v in self.view_ids:
            cam = pool_camera(W, H, v) if weak else view_camera(W, H, v)
            sizes, gts, levels, pinned = [], [], [], []
            for lvl, s in enumerate(self.scales):
                cs = cam.at_scale(s)
                # ground truth = the initial model's own render of this view + noise: the loss and its gradients are those
                # of a model near convergence (most of training), with no systematic drift of the synthetic scene
                img = (render(cs) + 0.05 * torch.randn((3, cs.H, cs.W), generator=g).to(dev)).clamp_(0, 1)
                sizes.append((cs.H, cs.W))
                gts.append(img)
                levels.append(lvl)
                pinned.append(img.cpu().pin_memory())
            self.views[v] = msgs_types.TrainView(tt(cam.viewmatrix), tt(cam.projmatrix), tt(cam.campos), cam.tanfovx,
                                                 cam.tanfovy, sizes, gts, levels, view_id=v)
            self.host[v] = dict(cam=torch.from_numpy(np.concatenate([cam.viewmatrix.reshape(-1), cam.projmatrix.reshape(-1),
                                                                     cam.campos.reshape(-1)])).pin_memory(), gts=pinned)
        self.units_per_step = self.views_per_step * self.L
        self.px_per_step = sum(h * w for (h, w) in self.views[self.view_ids[0]].sizes) * self.views_per_step

    def step_view_ids(self, k):
        if self.wl["scaling"] == "weak":
            # Weak scaling = the per-GPU work of a step does not depend on N: at step k EVERY rank renders pool view k mod 8
            # (its own forward, backward, statistics and gradient contribution; the exchange volume is what it is for any
            # views), so a rank's work at N = 8 is exactly the N = 1 work and the ratio of the two lines isolates the sharded
            # step's overhead.  MSGS_VIEW_POOL=spread: rank r renders view (k + r) mod 8 — every step then also waits for the
            # most expensive of N different views (a property of the views, not of the step: + 0.1 ms at N = 8).
            if os.environ.get("MSGS_VIEW_POOL", "same") == "spread":
                return [self.view_ids[(k + self.rank) % len(self.view_ids)]]
            return [self.view_ids[k % len(self.view_ids)]]
        return self.view_ids

    def fresh_state(self):
        import msgs_types
        sc, dev = self.wl["scene"], self.dev
        tt = lambda a, dt=torch.float32: torch.from_numpy(np.ascontiguousarray(a)).to(device=dev, dtype=dt)
        return msgs_types.MSState.fresh(self.P, self.L, dev, tt(sc.max_pixel_sizes), tt(sc.min_pixel_sizes),
                                        tt(sc.base_mask, torch.bool), self.target.clone())


def activated(raw):
    """scene/gaussian_model.py:127-155"""
    return dict(means3D=raw["xyz"], shs=torch.cat((raw["f_dc"], raw["f_rest"]), dim=1).contiguous(),
                opacities=torch.sigmoid(raw["opacity"]), scales=torch.exp(raw["scaling"]),
                rotations=torch.nn.functional.normalize(raw["rotation"]))


class MineTrainArm:
    name = "mine"

    def __init__(self, cfg, rank, world, dev):
        import diff_gaussian_rasterization as dgr
        import msgs_optim
        import msgs_train
        self.dgr, self.dev, self.rank, self.world = dgr, dev, rank, world
        self.lib = dgr._C.lib
        self.lib.msgs_launch_count.restype = ctypes.c_longlong

        def render_factory(setup):
            act = activated(setup.raw)
            t = setup.wl["t"]

            def render(cam):
                rs = dgr.GaussianRasterizationSettings(cam.H, cam.W, cam.tanfovx, cam.tanfovy, setup.bg, 1.0,
                                                       torch.from_numpy(cam.viewmatrix).to(dev), torch.from_numpy(cam.projmatrix).to(dev),
                                                       setup.sh_degree, torch.from_numpy(cam.campos).to(dev), False, False, True, True, 1.0)
                with torch.no_grad():
                    color = dgr.GaussianRasterizer(rs)(
                        means3D=act["means3D"], means2D=torch.zeros_like(act["means3D"]), opacities=act["opacities"],
                        occ_multiplier=t["occ_multiplier"], dc_delta=t["dc_delta"], shs=act["shs"], base_mask=t["base_mask"],
                        scales=act["scales"], rotations=act["rotations"], max_pixel_sizes=t["max_pixel_sizes"],
                        min_pixel_sizes=t["min_pixel_sizes"])[0]
                return color
            return render
        self.setup = s = TrainSetup(cfg, rank, world, dev, render_factory, MineRunner)
        self.flat = msgs_optim.FlatGaussians(s.P, dev, symmetric=world > 1)
        r = s.raw
        self.flat.load(r["xyz"], r["f_dc"], r["f_rest"], r["scaling"], r["rotation"], r["opacity"])
        self.trainer = msgs_train.MultiScaleStep(self.flat, s.fresh_state(), s.bg, s.sh_degree, rank=rank, world_size=world)
        self.snap = self.trainer.snapshot()
        self.k = 0
        self.loss = None
        # ---- e2e plumbing: double-buffered device copies of the step's cameras and ground-truth images ----
        self.copy_stream = torch.cuda.Stream(device=dev)
        n = s.views_per_step
        self.cam_dev = [[torch.empty(35, device=dev) for _ in range(n)] for _ in range(2)]
        v0 = s.views[s.view_ids[0]]
        self.gt_dev = [[[torch.empty_like(g) for g in v0.gts] for _ in range(n)] for _ in range(2)]
        self.consumed = [None, None]
        self.loss_host = [torch.zeros(1).pin_memory() for _ in range(2)]
        self.loss_ready = [None, None]
        self.last_loss = None
        self.h2d = n * (35 * 4 + sum(g.numel() * 4 for g in v0.gts))

    def launches(self):
        return int(self.lib.msgs_launch_count())

    def reset(self):
        self.trainer.restore(self.snap)
        self.k = 0

    def step_resident(self):
        s = self.setup
        self.loss = self.trainer.step([s.views[v] for v in s.step_view_ids(self.k)])
        self.k += 1

    def step_e2e(self):
        """the public step call with HOST inputs: per step the camera (35 floats) and the ground-truth images of the step's
        views are copied from pinned host memory (copy stream; the images are only needed at loss time, so their copy runs
        under the forward), and the loss scalar goes back to pinned host memory (read by the host one step later)."""
        import msgs_train
        s, kb = self.setup, self.k & 1
        main = torch.cuda.current_stream()
        vids = s.step_view_ids(self.k)
        views = []
        with torch.cuda.stream(self.copy_stream):
            for j, v in enumerate(vids):
                self.cam_dev[kb][j].copy_(s.host[v]["cam"], non_blocking=True)
            cams_copied = self.copy_stream.record_event()
            if self.consumed[kb] is not None:
                self.copy_stream.wait_event(self.consumed[kb])       # the loss of two steps ago has read these buffers
            for j, v in enumerate(vids):
                for d, h in zip(self.gt_dev[kb][j], s.host[v]["gts"]):
                    d.copy_(h, non_blocking=True)
            gts_copied = self.copy_stream.record_event()
        main.wait_event(cams_copied)
        for j, v in enumerate(vids):
            ref, d = s.views[v], self.cam_dev[kb][j]
            views.append(msgs_train.TrainView(d[0:16].view(4, 4), d[16:32].view(4, 4), d[32:35], ref.tanfovx, ref.tanfovy,
                                              ref.sizes, self.gt_dev[kb][j], ref.levels, view_id=v))
        loss = self.trainer.step(views, before_loss=lambda: main.wait_event(gts_copied))
        self.consumed[kb] = main.record_event()
        self.loss_host[kb].copy_(loss.reshape(1), non_blocking=True)
        self.loss_ready[kb] = main.record_event()
        if self.loss_ready[kb ^ 1] is not None:
            self.loss_ready[kb ^ 1].synchronize()
            self.last_loss = float(self.loss_host[kb ^ 1][0])
        self.k += 1

    def phase_times(self, steps=5):
        """CUDA-event times of the step's phases on the main stream (the unit streams are joined at the phase boundaries)."""
        s, tr = self.setup, self.trainer
        acc = [0.0, 0.0, 0.0]
        for _ in range(steps):
            views = [s.views[v] for v in s.step_view_ids(self.k)]
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
            ev[0].record()
            loss, units, m2 = tr.render_and_backward(views)
            ev[1].record()
            tr.update_statistics(views, units, m2)
            ev[2].record()
            tr.adam.step()
            ev[3].record()
            torch.cuda.synchronize()
            for i in range(3):
                acc[i] += ev[i].elapsed_time(ev[i + 1]) / steps
            self.k += 1
        return {"render_loss_backward_ms": acc[0], "statistics_ms": acc[1], "optimizer_ms": acc[2]}


class RefTrainArm:
    """The reference's training iteration (harness/ref_train.py: torch activations, loss, statistics, torch.optim.Adam)
    around its UNMODIFIED CUDA rasterizer (oracle/_ref).  Loads nothing of ms-gs_b200/."""
    name = "reference"

    def __init__(self, cfg, rank, world, dev):
        from harness import ref_train, refcuda_bind
        self.dev = dev

        def render_factory(setup):
            act = activated(setup.raw)
            t = setup.wl["t"]

            def render(cam):
                tt = dict(t)
                tt.update(act)
                tt.update(viewmatrix=torch.from_numpy(cam.viewmatrix).to(dev), projmatrix=torch.from_numpy(cam.projmatrix).to(dev),
                          campos=torch.from_numpy(cam.campos).to(dev))
                return refcuda_bind.RefRasterizer(sync=False).forward(tt, cam, True, True).color.clone()
            return render
        self.setup = s = TrainSetup(cfg, rank, world, dev, render_factory, RefRunner)
        import msgs_types
        self.make = lambda: ref_train.RefTrainer(s.raw, s.fresh_state(), s.bg, s.sh_degree, msgs_types.DEFAULT_LRS,
                                                 stats_what=3, sync=False)
        self.trainer = self.make()
        self.k = 0

    def launches(self):
        return 0

    def reset(self):
        self.trainer = self.make()
        self.k = 0

    def step_resident(self):
        s = self.setup
        self.loss = self.trainer.step([s.views[v] for v in s.step_view_ids(self.k)])
        self.k += 1


# ------------------------------------------------------------------------------------------------------------
def cpu_baseline(cfg, budget_s=20.0):
    """The CPU oracle (port of the reference path, OpenMP over all host cores) on a bounded sample of the workload:
    the full-size scene, view 0 at scale 1 (the dominant unit of a multi-scale step), one fwd(+bwd), repeated while
    the budget allows (at most 3 times)."""
    from harness import oracle_bind
    c = CFG[cfg]
    P, W, H = scenes.CONFIGS[c["scene"]]
    sc = scenes.make_scene(P, W, H, seed=c["seed"], splat_scale=c.get("splat_scale", 0.004))
    cam = scenes.make_camera(W, H)
    dpix = np.random.default_rng(99).normal(0, 1, (3, H, W)).astype(np.float32)
    o = oracle_bind.Oracle()
    times = []
    t_all = time.time()
    while True:
        t0 = time.time()
        o.forward(sc, cam)
        if not c["forward_only"]:
            o.backward(dpix)
        times.append(time.time() - t0)
        if time.time() - t_all + times[-1] > budget_s or len(times) >= 3:
            break
    best = min(times)
    units = len(c["scales"]) * (1 if c["views"] == "world" else int(c["views"]))
    what = "forward" if c["forward_only"] else "fwd+bwd"
    note = "" if units == 1 else (f"; the step of this config has {units} (view, scale) units of which this is the most "
                                  f"expensive one, plus loss / statistics / Adam, so the CPU time of a full step is larger")
    return {"value": 1.0 / best, "unit": UNIT, "cores": os.cpu_count(), "kind": "port",
            "sample": f"{len(times)} full rasterizer {what} iteration(s) of the {c['scene']} scene, view 0 at scale 1 (best of {len(times)}: "
                      f"{best:.2f} s){note}; oracle/msgs_oracle.cpp, g++ -O2 -fopenmp, {os.cpu_count()} threads"}


def reference_arm_cpu(args):
    """Fallback reference arm: the CPU port (oracle/) of the reference path on all host cores, a bounded sample."""
    cb = cpu_baseline(args.config, budget_s=60.0)
    return {"metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1000.0 / cb["value"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "impl": "reference",
            "config": {"workload": WORKLOADS[args.config]}, "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}


REF_CPU_NOTE = ("the reference has NO CPU implementation of this path: this arm is its own CUDA rasterizer "
                "(cuda_rasterizer/*.cu, unmodified, compiled in place for sm_100a -> oracle/_ref) on the same B200, same inputs")


# ------------------------------------------------------------------------------------------------------------
def stage_profile(runner, step_fn, prof_steps, dev):
    """per-stage CUDA-event times of the rasterizer (msgs_profile_*), units back to back on one stream"""
    import msgs_sharding
    lib = runner.lib
    concurrent = runner.unit_streams
    runner.unit_streams = msgs_sharding.UnitStreams(dev, 1)
    lib.msgs_profile_enable(1)
    nst = lib.msgs_profile_num_stages()
    ms_acc, cnt_acc = (ctypes.c_double * nst)(), (ctypes.c_longlong * nst)()
    for _ in range(prof_steps):
        step_fn()
    torch.cuda.synchronize()
    lib.msgs_profile_collect(ms_acc, cnt_acc, nst)
    lib.msgs_profile_enable(0)
    runner.unit_streams = concurrent
    return [(lib.msgs_profile_stage_name(i).decode(), cnt_acc[i], ms_acc[i]) for i in range(nst) if cnt_acc[i]]


def roofline_of_unit(runner, wl, st, prof, prof_steps, unit_ms, clocks, hbm_peak, peak_src, sm_max):
    """roofline of the dominant kernel of ONE (view, scale) unit + the whole-unit fraction
    sum_stage max(bytes / HBM peak, flops / FP32 peak) / measured fwd+bwd time of the unit."""
    R, P, vis, E = st["R"], wl["P"], st["visible"], st["sum_n_contrib"]
    clk = (clocks or {}).get("sm_mhz") or sm_max
    fp32_peak = 148 * 128 * 2 * clk * 1e6 / 1e12       # TFLOP/s at the clock seen during the run
    fwd_only = bool(wl.get("forward_only"))
    alg = {   # algorithmic bytes (hbm) or flops (fp32) per launch — DESIGN.md §5
        "preprocess_fwd": ("hbm", 53.0 * P + 192.0 * vis + 28.0 * P + 73.0 * vis),
        "depth_sort": ("hbm", (4.0 + 12.0 + 2 * 16.0 + 12.0) * P),    # histogram | pass 0 (no value read) | 2 full passes | last pass (no key write)
        "chunk_count": ("hbm", 20.0 * vis), "chunk_scan": ("hbm", 0.0), "tile_scan": ("hbm", 0.0), "list_count": ("hbm", 0.0),
        "place": ("hbm", 12.0 * vis + 4.0 * R),
        "preprocess_bwd": ("hbm", 53.0 * P + (192.0 + 24 + 48 + 1) * vis + 364.0 * P),
        "blend_fwd": ("fp32", 36.0 * E), "blend_bwd": ("fp32", 62.0 * E)}
    stages, ideal_ms = [], 0.0
    for name, n, ms in prof:
        per = ms / n
        kind, work = alg.get(name, ("hbm", 0.0))
        ach = None if not work else (work / (per * 1e-3) / (1e9 if kind == "hbm" else 1e12))
        peak = hbm_peak if kind == "hbm" else fp32_peak
        if work:
            ideal_ms += (n // prof_steps) * work / (peak * (1e9 if kind == "hbm" else 1e12)) * 1e3
        stages.append({"stage": name, "launches_per_step": int(n) // prof_steps, "ms_per_launch": per, "ms_per_step": ms / prof_steps,
                       "bound": kind, "achieved": ach, "unit": "GB/s" if kind == "hbm" else "TFLOP/s",
                       "frac": None if ach is None else ach / peak})
    stages.sort(key=lambda s: -s["ms_per_step"])
    top = stages[0]
    roof = {"kernel": top["stage"], "bound": "hbm" if top["bound"] == "hbm" else "fp32", "achieved": top["achieved"],
            "peak": hbm_peak if top["bound"] == "hbm" else fp32_peak, "unit": top["unit"], "frac": top["frac"],
            "traffic": ncu_traffic(top["stage"]),
            "peak_source": peak_src if top["bound"] == "hbm" else f"148 SMs x 128 FP32 lanes x 2 x {clk:.0f} MHz (clock sampled during the run)",
            "ms_per_launch": top["ms_per_launch"], "share_of_unit": top["ms_per_step"] / unit_ms,
            "unit_measured": f"{wl['W']}x{wl['H']} unit of view 0 ({'forward' if fwd_only else 'forward + backward'}): {unit_ms:.3f} ms",
            "whole_unit": {"ideal_ms": ideal_ms, "measured_ms": unit_ms, "frac": ideal_ms / unit_ms,
                           "how": "sum over stages of max(algorithmic bytes / HBM peak, flops / FP32 peak) / measured time"},
            "note": ("bound 'fp32': the dominant kernel is limited by the FP32 pipe / issue slots (per-pixel 2x2 quadratic form + ex2 + "
                     "blend recurrence; no dense contraction, so no tensor-core work on this path); its DRAM traffic "
                     "(roofline.traffic) is < 1 % of what HBM could move in the kernel's time; the HBM-bound stages are "
                     "listed with their own fractions under 'stages'") if top["bound"] != "hbm" else None}
    return roof, stages


def raster_bench(cfg, impl, steps, warmup, dev, sampler=None, repeats=2, want_e2e=True):
    """single-GPU rasterizer-only figures of a raster config (C2: fwd+bwd, C5: forward): dict(value, ms_per_step, e2e, ...)"""
    wl = build_workload(cfg, 0, 1, dev)
    Runner = MineRunner if impl == "mine" else RefRunner
    if wl["ms"]:
        apply_ms_variant(wl, Runner, dev)
    runner = Runner(wl, dev)
    l0 = runner.launches()
    regions = [timed_loop(runner.step, steps, warmup, False, sampler) / steps for _ in range(repeats)]
    ms = min(regions)
    n_l = (runner.launches() - l0)
    st = runner.stats()
    out = {"workload": WORKLOADS[cfg], "value": 1000.0 / ms, "unit": UNIT if not wl["forward_only"] else "frames/s", "ms_per_step": ms,
           "timed_regions_ms_per_step": [round(x, 4) for x in regions], "num_rendered_R": st["R"], "visible": st["visible"],
           "sum_n_contrib": st["sum_n_contrib"],
           "gaussian_pixel_evals_per_s": (1 if wl["forward_only"] else 2) * st["sum_n_contrib"] * 1000.0 / ms}
    e2e = None
    if impl == "mine" and want_e2e:
        e2e = raster_e2e(runner, wl, steps, warmup, dev, repeats)
        out["e2e"] = e2e
    return out, runner, wl, st


def raster_e2e(runner, wl, steps, warmup, dev, repeats):
    """the public autograd API (GaussianRasterizer) with host inputs: camera + target image H2D, loss scalar D2H per step;
    forward-only configs copy every rendered frame to pinned host memory instead (viewer path)."""
    dgr = runner.dgr
    t, sc = wl["t"], wl["scene"]
    fo = wl["forward_only"]
    cam = wl["cams"][0]
    params = {k: t[k].clone().requires_grad_(not fo) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
    means2D = torch.zeros_like(params["means3D"], requires_grad=not fo)
    cam_host = torch.from_numpy(np.concatenate([cam.viewmatrix.reshape(-1), cam.projmatrix.reshape(-1), cam.campos.reshape(-1)])).pin_memory()
    cam_dev = [torch.empty(35, device=dev) for _ in range(2)]
    target = torch.rand((3, cam.H, cam.W)).pin_memory()
    dev_target = torch.empty((3, cam.H, cam.W), device=dev)
    copy_stream, up_stream = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
    loss_host = [torch.zeros(1).pin_memory() for _ in range(2)]
    loss_ready, frame_done = [None, None], [None, None]
    frame_host = [torch.empty((3, cam.H, cam.W)).pin_memory() for _ in range(2)] if fo else None
    k = [0]
    consumed = [None]

    def settings(d):
        return dgr.GaussianRasterizationSettings(cam.H, cam.W, cam.tanfovx, cam.tanfovy, t["bg"], sc.scale_modifier, d[0:16].view(4, 4),
                                                 d[16:32].view(4, 4), sc.sh_degree, d[32:35], False, False, runner.fs, runner.fl, 1.0)

    def render(rs):
        return dgr.GaussianRasterizer(rs)(
            means3D=params["means3D"], means2D=means2D, opacities=params["opacities"], occ_multiplier=t["occ_multiplier"],
            dc_delta=t["dc_delta"], shs=params["shs"], base_mask=t["base_mask"], scales=params["scales"],
            rotations=params["rotations"], max_pixel_sizes=t["max_pixel_sizes"], min_pixel_sizes=t["min_pixel_sizes"])

    def step_train():
        for p in params.values():
            p.grad = None
        means2D.grad = None
        kb, main = k[0] & 1, torch.cuda.current_stream()
        with torch.cuda.stream(copy_stream):
            cam_dev[kb].copy_(cam_host, non_blocking=True)
            cam_copied = copy_stream.record_event()
            if consumed[0] is not None:
                copy_stream.wait_event(consumed[0])
            dev_target.copy_(target, non_blocking=True)
            copied = copy_stream.record_event()
        main.wait_event(cam_copied)
        color = render(settings(cam_dev[kb]))[0]
        main.wait_event(copied)
        loss = (color - dev_target).square().sum() * 0.5
        loss.backward()
        consumed[0] = main.record_event()
        loss_host[kb].copy_(loss.detach().reshape(1), non_blocking=True)
        loss_ready[kb] = main.record_event()
        if loss_ready[kb ^ 1] is not None:
            loss_ready[kb ^ 1].synchronize()
        k[0] += 1

    def step_view():
        kb, main = k[0] & 1, torch.cuda.current_stream()
        with torch.no_grad():
            if frame_done[kb] is not None:
                frame_done[kb].synchronize()         # the copy that last used this pinned buffer (two steps ago)
            with torch.cuda.stream(up_stream):       # not the copy stream: that one is busy with the previous frame
                cam_dev[kb].copy_(cam_host, non_blocking=True)
                cam_copied = up_stream.record_event()
            main.wait_event(cam_copied)
            color = render(settings(cam_dev[kb]))[0]
            rendered = main.record_event()
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(rendered)
                frame_host[kb].copy_(color, non_blocking=True)
                color.record_stream(copy_stream)
                frame_done[kb] = copy_stream.record_event()
        k[0] += 1

    fn = step_view if fo else step_train
    regions = [timed_loop(fn, steps, warmup, False, on_end=lambda: torch.cuda.current_stream().wait_stream(copy_stream)) / steps
               for _ in range(repeats)]
    ms = min(regions)
    return {"value": 1000.0 / ms, "unit": UNIT if not fo else "frames/s",
            "h2d_bytes_per_step": 35 * 4 + (0 if fo else target.numel() * 4), "d2h_bytes_per_step": 4 if not fo else target.numel() * 4,
            "ms_per_step": ms, "timed_regions_ms_per_step": [round(x, 4) for x in regions],
            "what": ("GaussianRasterizer autograd API; per step: pinned target image (copy stream, under the forward) + camera H2D, L2 "
                     "loss, backward, loss scalar D2H (read one step later)") if not fo else
                    ("GaussianRasterizer API under no_grad; per step: camera H2D, forward, rendered frame D2H into one of two pinned "
                     "buffers on the copy stream (frame k travels while frame k + 1 is rendered)")}


# ------------------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="mine", choices=["mine", "reference"])
    ap.add_argument("--config", default="C3", choices=sorted(CFG))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    dist = world > 1
    if args.impl == "reference" and world > 1:
        # reference arm under torchrun: rank 0 alone runs the single-GPU reference; the others exit without work
        if rank != 0:
            return
        dist, world = False, 1
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the rasterizer has no CPU path")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if dist:
        pin_rank_to_cores(local_rank, int(os.environ.get("LOCAL_WORLD_SIZE", world)))
    if dist:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.distributed.init_process_group("nccl", device_id=dev)

    if args.impl == "reference":
        from harness import refcuda_bind
        if not refcuda_bind.available():
            # oracle/_ref (the reference CUDA build) did not travel: the CPU port of the reference path is the arm
            if rank == 0:
                print(json.dumps(reference_arm_cpu(args)))
            return

    hbm_peak, peak_src, sm_max = load_peaks()
    repeats = max(1, int(os.environ.get("MSGS_REPEATS", "3")))
    sampler = ClockSampler(local_rank)
    if rank == 0 and not os.environ.get("MSGS_NO_CLOCK_SAMPLER"):
        sampler.start()
    kind = CFG[args.config]["kind"]
    extras = {}

    if kind == "raster":
        if dist:
            raise SystemExit("rasterizer-only configs (C1, C2, C5) are single-GPU lines; the sharded step is --config C3 / C4")
        out, runner, wl, st = raster_bench(args.config, args.impl, args.steps, args.warmup, dev, sampler, repeats)
        clocks = sampler.stop()
        line = {"metric": METRIC, "value": out["value"], "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": out["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic", "impl": args.impl,
                "config": {"workload": WORKLOADS[args.config], "P": wl["P"], "image": [wl["W"], wl["H"]], "num_rendered_R": st["R"],
                           "visible": st["visible"], "sum_n_contrib": st["sum_n_contrib"],
                           "l2_policy": "per-step inputs exceed the 126 MB L2; no explicit flush"},
                "gaussian_pixel_evals_per_s": out["gaussian_pixel_evals_per_s"], "clocks": clocks,
                "gpu_launches": (runner.launches() if args.impl == "mine" else 0),
                "timed_regions_ms_per_step": out["timed_regions_ms_per_step"],
                "e2e": out.get("e2e") or {"value": out["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        if args.impl == "mine":
            prof_steps = 5
            prof = stage_profile(runner, runner.step, prof_steps, dev)
            line["roofline"], line["stages"] = roofline_of_unit(runner, wl, st, prof, prof_steps, out["ms_per_step"], clocks, hbm_peak,
                                                                peak_src, sm_max)
            if not args.no_cpu_baseline:
                line["cpu_baseline"] = cpu_baseline(args.config)
        else:
            line["cpu_baseline"] = {"value": out["value"], "unit": UNIT, "cores": 0, "kind": "reference", "sample": REF_CPU_NOTE}
        print(json.dumps(line))
        return

    # ---------------- training configs ----------------
    Arm = MineTrainArm if args.impl == "mine" else RefTrainArm
    arm = Arm(args.config, rank, world, dev)
    s = arm.setup
    spec0 = None
    if args.impl == "mine":
        spec0 = arm.dgr._C.speculation_stats(reset=True)

    regions, total_ms, l0, l1 = [], None, 0, 0
    for rep in range(repeats):
        arm.reset()
        lc = []
        ms = timed_loop(arm.step_resident, args.steps, args.warmup, dist, sampler if rank == 0 else None,
                        on_start=lambda: lc.append(arm.launches()))
        regions.append(ms / args.steps)
        if total_ms is None or ms < total_ms:
            total_ms, l0, l1 = ms, lc[0], arm.launches()
    clocks = sampler.stop() if rank == 0 else None
    ms_per_step = total_ms / args.steps
    weak = s.wl["scaling"] == "weak"
    value = (world if weak else 1) * 1000.0 / ms_per_step
    spec = arm.dgr._C.speculation_stats() if args.impl == "mine" else None

    e2e = None
    if args.impl == "mine":
        e2e_regions = []
        for rep in range(repeats):
            arm.reset()
            e2e_regions.append(timed_loop(arm.step_e2e, args.steps, args.warmup, dist,
                                          on_end=lambda: torch.cuda.current_stream().wait_stream(arm.copy_stream)) / args.steps)
        e2e_ms = min(e2e_regions)
        e2e = {"value": (world if weak else 1) * 1000.0 / e2e_ms, "unit": UNIT, "h2d_bytes_per_step": int(arm.h2d), "d2h_bytes_per_step": 4,
               "ms_per_step": e2e_ms, "timed_regions_ms_per_step": [round(x, 4) for x in e2e_regions],
               "what": "msgs_train.MultiScaleStep.step (rasterize_gaussians_units raw-parameter autograd node -> l1_ssim_loss_batch -> "
                       "backward -> statistics -> fused Adam); per step: camera + the view's ground-truth images of all scales H2D "
                       "from pinned memory (copy stream; the images travel under the forward), loss scalar D2H into pinned memory "
                       "(read one step later); Gaussian parameters and optimiser state stay resident (they are the model)"}
        arm.reset()
        phases = arm.phase_times()
    loss_val = float(arm.loss) if getattr(arm, "loss", None) is not None else None

    per_rank_ms = None
    if dist:
        # load balance diagnostic: every rank's own step time without optimiser / exchange
        arm.reset()
        torch.cuda.synchronize()
        torch.distributed.barrier()
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c0.record()
        for _ in range(5):
            vs = [s.views[v] for v in s.step_view_ids(arm.k)]
            arm.trainer.step(vs, optimizer=False, statistics=False)
            arm.k += 1
        c1.record()
        torch.cuda.synchronize()
        mine_ms = torch.tensor([c0.elapsed_time(c1) / 5], device=dev)
        allms = [torch.zeros(1, device=dev) for _ in range(world)]
        torch.distributed.all_gather(allms, mine_ms)
        per_rank_ms = [round(float(x.item()), 3) for x in allms]
        torch.distributed.barrier()

    # ---- extras: at N > 1 the C4 strong-scaling figure of north_star's multi-GPU workload (every rank takes part) ----
    if dist and args.config == "C3" and args.impl == "mine" and not args.no_extras:
        del arm
        torch.cuda.empty_cache()
        a4 = MineTrainArm("C4", rank, world, dev)
        r4 = []
        for rep in range(2):
            a4.reset()
            r4.append(timed_loop(a4.step_resident, max(3, args.steps // 4), 3, dist) / max(3, args.steps // 4))
        extras["C4"] = {"workload": WORKLOADS["C4"], "value": 1000.0 / min(r4), "unit": UNIT, "ms_per_step": min(r4), "scaling": "strong",
                        "timed_regions_ms_per_step": [round(x, 4) for x in r4], "views_per_rank": a4.setup.views_per_step,
                        "note": "whole 8-view step per second; device-resident inputs; the same step driver as the headline"}
        del a4
        arm = None
    if rank != 0:
        if dist:
            torch.distributed.destroy_process_group()
        return

    P = s.P
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": s.wl["scaling"], "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "impl": args.impl,
        "config": {"workload": WORKLOADS[args.config], "P": P, "image": [s.W, s.H], "scales": s.scales,
                   "units_per_step_per_gpu": s.units_per_step, "views_per_step_per_gpu": s.views_per_step,
                   "pixels_per_step_per_gpu": s.px_per_step, "view_pool": len(s.view_ids),
                   "step": "render fwd -> L1+SSIM (lambda 0.2, x0.1 on coarse scales) -> bwd -> update_pixel_sizes + max_radii2D + "
                           "add_densification_stats -> Adam (reference learning rates, eps 1e-15); a different view every step (pool of 8, "
                           "yaw -20..+20 degrees; at a given step every rank renders the same pool view, so per-GPU work is the N = 1 "
                           "work: weak scaling; MSGS_VIEW_POOL=spread gives every rank its own view); "
                           "ground truth = initial render + noise; every timed region starts from the same model state",
                   "last_loss": loss_val,
                   "l2_policy": f"per-step inputs (SH rows {P * 192 / 1e6:.0f} MB + splat records {P * 64 / 1e6:.0f} MB + parameters, "
                                f"gradients and Adam state {P * 236 * 4 / 1e6:.0f} MB) exceed the 126 MB L2; no explicit flush",
                   "parallelism": (f"{world} ranks, (view, scale) units sharded by whole views, Gaussians replicated in symmetric memory; "
                                   "statistics deltas: one typed-lane NVLS all-reduce; gradients: reduce-scatter + Adam on the rank's "
                                   "1/N shard + all-gather of the parameters fused in ONE multimem kernel (csrc/optim.cu)")
                   if world > 1 else "single GPU"},
        "clocks": clocks, "gpu_launches": (l1 - l0) if args.impl == "mine" else 0,
        "timed_regions_ms_per_step": [round(x, 4) for x in regions],
    }
    if per_rank_ms is not None:
        line["config"]["per_rank_step_ms_without_optimizer_and_exchange"] = per_rank_ms
    if args.impl == "mine":
        line["e2e"] = e2e
        line["step_phases"] = phases
        line["speculation"] = {"forwards": spec[0], "buffer_large_enough": spec[1], "relaunched": spec[2],
                               "hit_rate": (spec[1] / spec[0]) if spec[0] else None,
                               "what": "num_rendered changes every step (rotating views); a miss re-runs placement + forward blend"}
    else:
        line["e2e"] = {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
        line["cpu_baseline"] = {"value": value, "unit": UNIT, "cores": 0, "kind": "reference",
                                "sample": REF_CPU_NOTE + f"; {args.steps} full training iterations per timed region (torch activations, "
                                          "loss, statistics and torch.optim.Adam around it as train.py has them); 0 host cores do arithmetic"}

    # ---- N = 1: roofline of the scale-1 unit (the C2-shaped unit of the step) and the C2 / C5 figures ----
    if world == 1:
        del arm
        torch.cuda.empty_cache()
        if args.impl == "mine":
            # the scale-1 unit of view 0 with the step's MS-GS state and filters, rasterizer only
            wl = build_workload(args.config, 0, 1, dev)
            apply_ms_variant(wl, MineRunner, dev)
            wl1 = dict(wl, cams=[wl["probe_cam"]], scales=[1])
            r1 = MineRunner(wl1, dev)
            unit_ms = min(timed_loop(r1.step, 10, 3, False) / 10 for _ in range(2))
            st1 = r1.stats()
            prof_steps = 5
            prof = stage_profile(r1, r1.step, prof_steps, dev)
            line["roofline"], line["stages"] = roofline_of_unit(r1, wl1, st1, prof, prof_steps, unit_ms, clocks, hbm_peak, peak_src, sm_max)
            line["config"].update(num_rendered_R=st1["R"], visible=st1["visible"], sum_n_contrib=st1["sum_n_contrib"])
            line["gaussian_pixel_evals_per_s"] = 2 * st1["sum_n_contrib"] * 1000.0 / unit_ms
            del r1, wl, wl1
            torch.cuda.empty_cache()
        if not args.no_extras:
            for xc in ("C2", "C5"):
                try:
                    out, _r, _wl, _st = raster_bench(xc, args.impl, 10, 3, dev, None, 2)
                    extras[xc] = out
                    del _r, _wl
                    torch.cuda.empty_cache()
                except Exception as ex:
                    extras[xc] = {"failed": repr(ex)}
        if args.impl == "mine" and not args.no_cpu_baseline:
            try:
                line["cpu_baseline"] = cpu_baseline(args.config)
            except Exception as ex:   # the oracle is test infrastructure; its absence must not hide the GPU number
                line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "port", "sample": f"failed: {ex}"}
    if extras:
        line["extras"] = extras
    print(json.dumps(line))
    if dist:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
