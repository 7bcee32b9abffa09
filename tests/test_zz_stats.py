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


@pytest.mark.parametrize("tag", TAGS)
def test_thread_body_on_host(tmp_path_factory, tag):
    so_path = tmp_path_factory.mktemp("stats") / "libstats_host.so"
    subprocess.check_call(["nvcc", "-O1", "-std=c++17", "-shared", "-Xcompiler", "-fPIC,-ffp-contract=off", "-o", str(so_path),
                           os.path.join(ROOT, "tests", "stats_host_harness.cu")])
    lib = ctypes.CDLL(str(so_path))
    lib.host_update_stats.argtypes = [ctypes.c_int] * 4 + [ctypes.c_void_p] * 10
    lib.host_update_stats.restype = None

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
    (name, args), = re.findall(r"\b(msgs_[a-z_]+)\s*\(([^()]*)\)\s*;", src)
    assert name == "msgs_update_stats"
    assert hasattr(ctypes.CDLL(os.path.join(ROOT, "ms-gs_b200", "libmsgs_raster.so")), name)
    kinds = ["ptr" if "*" in a else "int" for a in args.split(",")]
    got = ["ptr" if t is ctypes.c_void_p else "int" for t in msgs_stats._lib.msgs_update_stats.argtypes]
    assert got == kinds
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
