"""Fused per-step statistics update (include/msgs_stats.h) against golden vectors produced by the
reference's own GaussianModel methods (tests/golden/make_stats_golden.py).  CPU: numpy oracle, the kernel's thread body run
on the host, exports / signature.  GPU (-m gpu): the CUDA kernel through `msgs_stats.update_stats`.  Integer-like state
(denom, max_radii2D, base mask) and the pixel-size state bit-exact; the gradient-norm accumulator within 1e-6 relative."""
import ctypes
import os
import subprocess

import numpy as np
import pytest
import torch

from oracle import stats_oracle as so

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
Z = np.load(os.path.join(ROOT, "tests", "golden", "stats_golden.npz"))
TAGS = ("p1000_l4", "p37_l1", "p1_l3")
STATE = ("max_pixel_sizes", "min_pixel_sizes", "max_radii2D", "xyz_gradient_accum", "denom", "base_gaussian_mask")


def _check(tag, u, st):
    for k in STATE:
        want = Z[f"{tag}/u{u}/{k}"]
        got = np.asarray(st[k]).reshape(want.shape)
        if k == "xyz_gradient_accum":
            assert np.allclose(got, want, rtol=1e-6, atol=0), f"{tag} unit {u}: {k}"
        else:
            assert np.array_equal(got, want), f"{tag} unit {u}: {k} ({int((got != want).sum())} differ)"


def _run(tag, step):
    """step(L, lvl, what, radii, px, target, grad, state) updates state in place; checked after every unit."""
    st = {k: Z[f"{tag}/init/{k}"].copy() for k in STATE}
    target = Z[f"{tag}/target_reso_lvl"]
    L = st["denom"].shape[1]
    for u, (lvl, what) in enumerate(Z[f"{tag}/units"]):
        step(L, int(lvl), int(what), Z[f"{tag}/u{u}/radii"], Z[f"{tag}/u{u}/pixel_sizes"], target, Z[f"{tag}/u{u}/grad"], st)
        _check(tag, u, st)


@pytest.mark.parametrize("tag", TAGS)
def test_oracle_matches_reference_golden(tag):
    _run(tag, lambda L, lvl, what, r, px, t, g, st: so.update_stats(L, lvl, what, r, px, t, g, *(st[k] for k in STATE)))


@pytest.fixture(scope="module")
def host_lib(tmp_path_factory):
    so_path = tmp_path_factory.mktemp("stats") / "libstats_host.so"
    subprocess.check_call(["nvcc", "-O1", "-std=c++17", "-shared", "-Xcompiler", "-fPIC,-ffp-contract=off", "-o", str(so_path),
                           os.path.join(ROOT, "tests", "stats_host_harness.cu")])
    lib = ctypes.CDLL(str(so_path))
    lib.host_update_stats.argtypes = [ctypes.c_int] * 4 + [ctypes.c_void_p] * 10
    lib.host_update_stats_units.argtypes = [ctypes.c_int] * 3 + [ctypes.c_void_p] * 12
    lib.host_stats_delta_units.argtypes = [ctypes.c_int] * 3 + [ctypes.c_void_p] * 12
    lib.host_stats_delta_apply.argtypes = [ctypes.c_int] * 2 + [ctypes.c_void_p] * 13
    for f in (lib.host_update_stats, lib.host_update_stats_units, lib.host_stats_delta_units, lib.host_stats_delta_apply):
        f.restype = None
    return lib


def _ptr(a):
    return ctypes.c_void_p(a.ctypes.data)


def _unit_args(tag, us):
    """ctypes arguments (lvl[], what[], radii*[], px*[], g2d*[]) for units `us` of golden case `tag` + the arrays kept alive."""
    units = Z[f"{tag}/units"]
    lvl = np.ascontiguousarray(units[us, 0].astype(np.int32))
    what = np.ascontiguousarray(units[us, 1].astype(np.int32))
    keep = [[np.ascontiguousarray(Z[f"{tag}/u{u}/{k}"]) for u in us] for k in ("radii", "pixel_sizes", "grad")]
    ptrs = [np.array([a.ctypes.data for a in arrs], np.uint64) for arrs in keep]
    return [_ptr(lvl), _ptr(what)] + [_ptr(p) for p in ptrs], (lvl, what, keep, ptrs)


@pytest.mark.parametrize("tag", TAGS)
def test_units_body_on_host_equals_the_sequence(host_lib, tag):
    """All units in one pass == the reference's methods applied unit after unit (chunks of <= 8 units)."""
    st = {k: Z[f"{tag}/init/{k}"].copy() for k in STATE}
    target = np.ascontiguousarray(Z[f"{tag}/target_reso_lvl"])
    n, L = len(Z[f"{tag}/units"]), st["denom"].shape[1]
    for u0 in range(0, n, 8):
        us = list(range(u0, min(n, u0 + 8)))
        args, keep = _unit_args(tag, us)
        host_lib.host_update_stats_units(len(target), L, len(us), *args, _ptr(target), *(_ptr(st[k]) for k in STATE[:5]),
                                         _ptr(st["base_gaussian_mask"].view(np.uint8)))
        _check(tag, us[-1], st)


@pytest.mark.parametrize("tag", TAGS)
def test_delta_then_apply_on_host_equals_the_direct_update(host_lib, tag):
    """One unit per step through the delta lanes (reset -> delta -> apply): the sharded step's path at world size 1 is
    bit-identical to the reference's direct updates, unit after unit."""
    st = {k: Z[f"{tag}/init/{k}"].copy() for k in STATE}
    target = np.ascontiguousarray(Z[f"{tag}/target_reso_lvl"])
    P, L = len(target), st["denom"].shape[1]
    for u in range(len(Z[f"{tag}/units"])):
        lanes = [np.zeros(P * L, np.float32), np.zeros(P * L, np.float32), np.zeros(P, np.float32), np.zeros(P, np.float32),
                 np.full(P, np.inf, np.float32), np.zeros(P, np.float32)]
        args, keep = _unit_args(tag, [u])
        host_lib.host_stats_delta_units(P, L, 1, *args, _ptr(target), *(_ptr(a) for a in lanes))
        host_lib.host_stats_delta_apply(P, L, _ptr(target), *(_ptr(a) for a in lanes), *(_ptr(st[k]) for k in STATE[:5]),
                                        _ptr(st["base_gaussian_mask"].view(np.uint8)))
        _check(tag, u, st)


def test_delta_lanes_reduce_over_ranks_on_host(host_lib):
    """Two ranks' deltas reduced lane by lane (f32 add | u32 max | u32 min on the bit patterns, as the switch does) and
    applied == numpy restatement of the documented semantics: one ageing per step with the largest / smallest observation."""
    tag = "p1000_l4"
    target = np.ascontiguousarray(Z[f"{tag}/target_reso_lvl"])
    P, L = len(target), 4
    units = Z[f"{tag}/units"]
    ranks = [[u for u in range(len(units)) if units[u, 1] == 7][:2], [u for u in range(len(units)) if units[u, 1] == 3][:2]]
    per_rank = []
    for us in ranks:
        lanes = [np.zeros(P * L, np.float32), np.zeros(P * L, np.float32), np.zeros(P, np.float32), np.zeros(P, np.float32),
                 np.full(P, np.inf, np.float32), np.zeros(P, np.float32)]
        args, keep = _unit_args(tag, us)
        host_lib.host_stats_delta_units(P, L, len(us), *args, _ptr(target), *(_ptr(a) for a in lanes))
        per_rank.append(lanes)
    a, b = per_rank
    red = [a[0] + b[0], a[1] + b[1]] + [np.maximum(a[k].view(np.uint32), b[k].view(np.uint32)).view(np.float32) for k in (2, 3)]
    red += [np.minimum(a[4].view(np.uint32), b[4].view(np.uint32)).view(np.float32),
            np.maximum(a[5].view(np.uint32), b[5].view(np.uint32)).view(np.float32)]
    assert np.array_equal(red[2], np.maximum(a[2], b[2])) and np.array_equal(red[4], np.minimum(a[4], b[4]))   # order-preserving bits
    st = {k: Z[f"{tag}/init/{k}"].copy() for k in STATE}
    init = {k: v.copy() for k, v in st.items()}
    host_lib.host_stats_delta_apply(P, L, _ptr(target), *(_ptr(np.ascontiguousarray(x)) for x in red),
                                    *(_ptr(st[k]) for k in STATE[:5]), _ptr(st["base_gaussian_mask"].view(np.uint8)))
    # restatement
    obs = {}
    vis_any, base_any = np.zeros(P, bool), np.zeros(P, bool)
    acc, den, rad = np.zeros((P, L), np.float32), np.zeros((P, L), np.float32), np.zeros(P, np.float32)
    for us in ranks:
        for u in us:
            lvl, what = int(units[u, 0]), int(units[u, 1])
            r, px, g = Z[f"{tag}/u{u}/radii"], Z[f"{tag}/u{u}/pixel_sizes"], Z[f"{tag}/u{u}/grad"]
            vis = r > 0
            for i in np.nonzero(vis & (target == lvl))[0]:
                if what & 1:
                    obs.setdefault(i, []).append(float(px[i]))
            if what & 2:
                acc[vis, lvl] += np.sqrt(g[vis, 0] ** 2 + g[vis, 1] ** 2).astype(np.float32)
                den[vis, lvl] += 1
                rad[vis] = np.maximum(rad[vis], r[vis].astype(np.float32))
                vis_any |= vis
            if what & 4:
                base_any |= vis
    want_max, want_min = init["max_pixel_sizes"].copy(), init["min_pixel_sizes"].copy()
    for i, ps in obs.items():
        hi = np.float32(max(max(ps), 0.0))
        pos = [p for p in ps if p > 0]
        if target[i] > 0:
            want_max[i] = max(np.float32(want_max[i] * np.float32(0.95)), hi)
        if target[i] < L - 1:
            m = max(np.float32(want_min[i] * np.float32(1.05)), np.float32(-1))
            want_min[i] = (np.float32(min(pos)) if pos else m) if m < 0 else (min(m, np.float32(min(pos))) if pos else m)
    assert np.array_equal(st["max_pixel_sizes"], want_max)
    assert np.array_equal(st["min_pixel_sizes"], want_min)
    assert np.array_equal(st["max_radii2D"], np.maximum(init["max_radii2D"], rad))
    assert np.allclose(st["xyz_gradient_accum"].reshape(P, L), init["xyz_gradient_accum"].reshape(P, L) + acc, rtol=1e-6)
    assert np.array_equal(st["denom"].reshape(P, L), init["denom"].reshape(P, L) + den)
    assert np.array_equal(st["base_gaussian_mask"], init["base_gaussian_mask"] | base_any)


@pytest.mark.parametrize("tag", TAGS)
def test_thread_body_on_host(host_lib, tag):
    lib = host_lib

    def step(L, lvl, what, r, px, t, g, st):
        ins = [np.ascontiguousarray(a) for a in (r, px, t, g)]
        base = st["base_gaussian_mask"].view(np.uint8)
        arrs = ins + [st[k] for k in STATE[:5]] + [base]
        lib.host_update_stats(len(r), L, lvl, what, *(ctypes.c_void_p(a.ctypes.data) for a in arrs))
    _run(tag, step)


def test_library_exports_and_signature(built):
    import re
    import msgs_stats
    src = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "msgs_stats.h")).read(), flags=re.S)
    protos = dict(re.findall(r"\b(msgs_[a-z_]+)\s*\(([^()]*)\)\s*;", src))
    assert set(protos) == {"msgs_update_stats", "msgs_update_stats_units", "msgs_stats_delta_words", "msgs_stats_delta_fast_words", "msgs_stats_delta_reset",
                           "msgs_stats_delta_units", "msgs_stats_delta_apply"}
    lib = ctypes.CDLL(os.path.join(ROOT, "ms-gs_b200", "libmsgs_raster.so"))
    for name, args in protos.items():
        assert hasattr(lib, name)
        kinds = ["ptr" if "*" in a else "int" for a in args.split(",")]
        got = ["int" if t is ctypes.c_int else "ptr" for t in getattr(msgs_stats._lib, name).argtypes]
        assert got == kinds, name
    assert ctypes.sizeof(msgs_stats._StatsUnit) == 8 + 3 * 8
    assert msgs_stats.delta_words(5, 3) == (32, 8 + 12, 8) and msgs_stats.delta_words(0, 1) == (0, 0, 0)
    with pytest.raises(RuntimeError):
        msgs_stats.update_stats(0, 1, torch.ones(4, dtype=torch.int32), what=msgs_stats.BASE_MASK,
                                base_gaussian_mask=torch.zeros(4, dtype=torch.bool))        # CPU tensors: loud failure


@pytest.mark.gpu
@pytest.mark.parametrize("tag", TAGS)
def test_gpu_update_stats_matches_golden(built, tag):
    import msgs_stats

    def step(L, lvl, what, r, px, t, g, st):
        d = {k: torch.from_numpy(st[k]).cuda() for k in STATE}
        msgs_stats.update_stats(lvl, L, torch.from_numpy(r).cuda(), torch.from_numpy(px).cuda(), torch.from_numpy(t).cuda(),
                                torch.from_numpy(g).cuda(), d["max_pixel_sizes"], d["min_pixel_sizes"], d["max_radii2D"],
                                d["xyz_gradient_accum"], d["denom"], d["base_gaussian_mask"], what=what)
        torch.cuda.synchronize()
        for k in STATE:
            st[k][...] = d[k].cpu().numpy()
    _run(tag, step)


@pytest.mark.gpu
@pytest.mark.parametrize("tag", TAGS)
def test_gpu_units_and_delta_paths_match_golden(built, tag):
    """GPU: all units in one launch (chunks of 8) and the delta path (one unit per step) both reproduce the reference."""
    import msgs_stats
    target = torch.from_numpy(Z[f"{tag}/target_reso_lvl"]).cuda()
    units = Z[f"{tag}/units"]
    P, L = target.numel(), Z[f"{tag}/init/denom"].shape[1]

    def unit(u):
        return (int(units[u, 0]), int(units[u, 1]), torch.from_numpy(Z[f"{tag}/u{u}/radii"]).cuda(),
                torch.from_numpy(Z[f"{tag}/u{u}/pixel_sizes"]).cuda(), torch.from_numpy(Z[f"{tag}/u{u}/grad"]).cuda())

    d = {k: torch.from_numpy(Z[f"{tag}/init/{k}"].copy()).cuda() for k in STATE}
    for u0 in range(0, len(units), 8):
        us = list(range(u0, min(len(units), u0 + 8)))
        msgs_stats.update_stats_units([unit(u) for u in us], L, target, *(d[k] for k in STATE))
        _check(tag, us[-1], {k: d[k].cpu().numpy() for k in STATE})
    d = {k: torch.from_numpy(Z[f"{tag}/init/{k}"].copy()).cuda() for k in STATE}
    delta = msgs_stats.StatsDelta(P, L, "cuda")
    for u in range(len(units)):
        delta.reset()
        delta.add_units([unit(u)], target)
        delta.apply(target, *(d[k] for k in STATE))
        _check(tag, u, {k: d[k].cpu().numpy() for k in STATE})
