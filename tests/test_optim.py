"""Fused Adam over the flat parameter buffer (device half; include/msgs_optim.h) against
torch.optim.Adam configured as the reference configures it (scene/gaussian_model.py:229-252: one param group per tensor,
lr per group, eps=1e-15, default betas; train.py:416-418 steps it once per iteration).

CPU: the kernel's per-element body (csrc/optim_body.cuh) compiled for the host and run over every chunk of a flat buffer
against torch.optim.Adam on CPU tensors; layout / exports / signatures; argument errors.  GPU (-m gpu): the CUDA kernel
against torch.optim.Adam on CUDA tensors (<= 2 ulp per element after every step: both sides are fp32 and differ at most in
FMA contraction), full C2 size; the sharded NVLS kernel against the single-GPU step on the summed gradients (2 ranks, runs
where 2 GPUs are visible: tools/check_sharded_adam.py under torchrun)."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LRS = {"xyz": 0.00016 * 3.7, "f_dc": 0.0025, "f_rest": 0.0025 / 20.0, "opacity": 0.05, "scaling": 0.005, "rotation": 0.001}


def _ref_model(P, seed, device):
    """Six leaf tensors as GaussianModel holds them + the reference's optimiser."""
    g = torch.Generator().manual_seed(seed)
    mk = lambda *s: torch.randn(*s, generator=g).to(device).requires_grad_(True)
    t = {"xyz": mk(P, 3), "f_dc": mk(P, 1, 3), "f_rest": mk(P, 15, 3), "opacity": mk(P, 1), "scaling": mk(P, 3), "rotation": mk(P, 4)}
    opt = torch.optim.Adam([{"params": [t[k]], "lr": LRS[k], "name": k} for k in ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")],
                           lr=0.0, eps=1e-15)
    return t, opt


def _grads(P, step, seed, device, sparse=True):
    g = torch.Generator().manual_seed(seed * 1000 + step)
    vis = (torch.rand(P, generator=g) < 0.7) if sparse else torch.ones(P, dtype=torch.bool)     # invisible Gaussians: zero gradient
    mk = lambda *s: (torch.randn(*s, generator=g) * vis.view(P, *([1] * (len(s) - 1))) * 10 ** torch.randn(1, generator=g).item()).to(device)
    return {"xyz": mk(P, 3), "f_dc": mk(P, 1, 3), "f_rest": mk(P, 15, 3), "opacity": mk(P, 1), "scaling": mk(P, 3), "rotation": mk(P, 4)}


def _ulp_diff(a, b):
    a, b = a.detach().cpu().contiguous().view(torch.int32).long(), b.detach().cpu().contiguous().view(torch.int32).long()
    # map the sign-magnitude float order onto a monotone integer line
    a = torch.where(a < 0, -(a & 0x7FFFFFFF), a)
    b = torch.where(b < 0, -(b & 0x7FFFFFFF), b)
    return int((a - b).abs().max())


def _flat_fill(lay, flat_views, t):
    with torch.no_grad():
        flat_views["mean3D"].copy_(t["xyz"])
        flat_views["sh"][:, :1].copy_(t["f_dc"])
        flat_views["sh"][:, 1:].copy_(t["f_rest"])
        flat_views["scale"].copy_(t["scaling"])
        flat_views["rot"].copy_(t["rotation"])
        flat_views["opacity"].copy_(t["opacity"])


def _compare(flat_views, t, tol_ulp, what, lrs=None, envelope=None):
    """Element-wise: within tol_ulp ulp, where the ulp is that of max(|p|, lr) — an update that nearly cancels the
    parameter (p' ~ 0) keeps the absolute rounding error of its two operands, not the relative one of the difference.
    `envelope` (dict, updated in place): running max of |p| over the steps compared so far — a rounding difference made
    while |p| was larger stays in p when it shrinks into the next binade (seen on B200: 2 ulp at step 3 = 4 ulp at step 4)."""
    lrs = lrs or LRS
    pairs = (("mean3D", t["xyz"], "xyz"), ("scale", t["scaling"], "scaling"), ("rot", t["rotation"], "rotation"),
             ("opacity", t["opacity"], "opacity"), (None, t["f_dc"], "f_dc"), (None, t["f_rest"], "f_rest"))
    for k, ref, lr_name in pairs:
        mine = flat_views[k] if k else (flat_views["sh"][:, :1] if lr_name == "f_dc" else flat_views["sh"][:, 1:])
        a, b = mine.detach().cpu().double(), ref.detach().cpu().double()
        mag = b.abs()
        if envelope is not None:
            mag = envelope[lr_name] = torch.maximum(mag, envelope.get(lr_name, mag))
        tol = tol_ulp * 2.0 ** -23 * torch.maximum(mag, torch.full_like(b, 2 * lrs[lr_name]))
        bad = int(((a - b).abs() > tol).sum())
        assert bad == 0, f"{what}: {lr_name}: {bad} elements differ by more than {tol_ulp} ulp (max {float((a - b).abs().max()):.3e})"


def test_flat_layout_and_exports(built):
    import msgs_optim
    from msgs_sharding import GRAD_FIELDS, GRAD_FLOATS
    for P in (1, 3, 1000, 1_200_000):
        lay = msgs_optim.FlatLayout(P)
        off = 0
        for name, n in GRAD_FIELDS:
            assert lay.offset[name] == off and lay.offset[name] % 4 == 0 and lay.size[name] == P * n
            off += (P * n + 3) // 4 * 4
        assert lay.total == off and lay.total >= P * GRAD_FLOATS and lay.total % 4 == 0
    # every prototype of the header is exported; ctypes argument counts match
    src = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "msgs_optim.h")).read(), flags=re.S)
    protos = re.findall(r"\b(msgs_[a-z_]+)\s*\(([^()]*)\)\s*;", src)
    assert {n for n, _ in protos} == {"msgs_adam_step", "msgs_adam_step_multimem", "msgs_adam_step_multimem_range", "msgs_adam_shard_floats", "msgs_adam_shard_begin",
                                      "msgs_stats_allreduce_multimem"}
    lib = ctypes.CDLL(os.path.join(ROOT, "ms-gs_b200", "libmsgs_raster.so"))
    for name, args in protos:
        assert hasattr(lib, name)
        assert len(getattr(msgs_optim._lib, name).argtypes) == len(args.split(",")), name
    # shards tile the flat index space, 16-byte aligned
    lib.msgs_adam_shard_floats.restype = lib.msgs_adam_shard_begin.restype = ctypes.c_longlong
    lib.msgs_adam_shard_floats.argtypes = [ctypes.c_longlong, ctypes.c_int]
    lib.msgs_adam_shard_begin.argtypes = [ctypes.c_longlong, ctypes.c_int, ctypes.c_int]
    for n, w in ((70_800_000, 8), (236, 3), (4, 8), (0, 2)):
        per = lib.msgs_adam_shard_floats(n, w)
        assert per % 4 == 0 and per * w >= n
        begins = [lib.msgs_adam_shard_begin(n, r, w) for r in range(w)]
        assert begins[0] == 0 and all(b % 4 == 0 and b <= n for b in begins) and begins == sorted(begins)
    # bad arguments fail loudly (no device needed)
    f = (msgs_optim._AdamField * 1)()
    f[0].begin, f[0].end, f[0].lr = 0, 8, 0.1
    assert msgs_optim._lib.msgs_adam_step(None, None, None, None, 8, f, 1, 0.9, 0.999, 1e-15, 0, None) < 0      # step 0
    assert msgs_optim._lib.msgs_adam_step(None, None, None, None, 6, f, 1, 0.9, 0.999, 1e-15, 1, None) < 0      # n % 4
    assert msgs_optim._lib.msgs_adam_step(None, None, None, None, 8, f, 1, 0.9, 0.999, 1e-15, 1, None) < 0      # NULL buffers
    f[0].end = 12
    assert msgs_optim._lib.msgs_adam_step(None, None, None, None, 8, f, 1, 0.9, 0.999, 1e-15, 1, None) < 0      # field past the end


@pytest.mark.parametrize("P", [1, 7, 1000])
def test_adam_body_on_host_matches_torch_adam(built, tmp_path_factory, P):
    import msgs_optim
    so_path = tmp_path_factory.mktemp("optim") / "liboptim_host.so"
    subprocess.check_call(["nvcc", "-O1", "-std=c++17", "-shared", "-Xcompiler", "-fPIC", "-o", str(so_path),
                           os.path.join(ROOT, "tests", "optim_host_harness.cu")], stderr=subprocess.DEVNULL)
    host = ctypes.CDLL(str(so_path))
    host.host_adam_step.restype = ctypes.c_int
    host.host_adam_step.argtypes = [ctypes.c_void_p] * 4 + [ctypes.c_longlong, ctypes.POINTER(msgs_optim._AdamField), ctypes.c_int,
                                                            ctypes.c_double, ctypes.c_double, ctypes.c_double, ctypes.c_int]
    t, opt = _ref_model(P, 5, "cpu")
    lay = msgs_optim.FlatLayout(P)
    bufs = {k: torch.zeros(lay.total) for k in ("p", "g", "m", "v")}
    pv, gv = lay.views(bufs["p"]), lay.views(bufs["g"])
    _flat_fill(lay, pv, t)

    class _Flat:                      # the part of FlatGaussians FusedAdam._fill_fields needs
        layout = lay
        device = "cpu"
        symmetric = False
    ad = msgs_optim.FusedAdam.__new__(msgs_optim.FusedAdam)
    ad.flat, ad.lrs = _Flat, dict(LRS)
    ad._fields = (msgs_optim._AdamField * 5)()
    for step in range(1, 8):
        g = _grads(P, step, 5, "cpu")
        for k in t:
            t[k].grad = g[k].clone()
        opt.step()
        _flat_fill(lay, gv, g)
        rc = host.host_adam_step(bufs["p"].data_ptr(), bufs["g"].data_ptr(), bufs["m"].data_ptr(), bufs["v"].data_ptr(), lay.total,
                                 ad._fill_fields(), 5, 0.9, 0.999, 1e-15, step)
        assert rc == 0
        _compare(pv, t, 4, f"step {step}")
    # padding between the fields is never touched
    for name in ("mean3D", "scale", "opacity"):
        pad = bufs["p"][lay.offset[name] + lay.size[name]:(lay.offset[name] + lay.size[name] + 3) // 4 * 4]
        assert float(pad.abs().sum()) == 0.0


def test_fused_adam_rejects_cpu_tensors(built):
    import msgs_optim
    flat = msgs_optim.FlatGaussians(16, "cpu")
    with pytest.raises(RuntimeError):
        msgs_optim.FusedAdam(flat).step()


# ------------------------------------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("P", [1, 1000, 1_200_000])
def test_fused_adam_matches_torch_adam_on_gpu(built, P):
    import msgs_optim
    dev = torch.device("cuda", 0)
    t, opt = _ref_model(P, 9, dev)
    flat = msgs_optim.FlatGaussians(P, dev)
    flat.load(t["xyz"], t["f_dc"], t["f_rest"], t["scaling"], t["rotation"], t["opacity"])
    ad = msgs_optim.FusedAdam(flat, LRS)
    steps = 3 if P > 100_000 else 12
    env = {}
    for step in range(1, steps + 1):
        if step == 3:                                            # update_learning_rate (gaussian_model.py:286-292)
            ad.set_lr("xyz", LRS["xyz"] * 0.5)
            opt.param_groups[0]["lr"] = LRS["xyz"] * 0.5
        g = _grads(P, step, 9, dev)
        for k in t:
            t[k].grad = g[k].clone()
        opt.step()
        _flat_fill(flat.layout, flat.g, g)
        ad.step()
        torch.cuda.synchronize()
        _compare(flat.p, t, 2, f"P={P} step {step}", envelope=env)
    # optimiser state too
    st = opt.state[t["scaling"]]
    m = flat.layout.views(ad.exp_avg)["scale"]
    v = flat.layout.views(ad.exp_avg_sq)["scale"]
    for mine, ref in ((m, st["exp_avg"]), (v, st["exp_avg_sq"])):
        a, b = mine.double(), ref.double()
        assert bool(((a - b).abs() <= 4 * 2.0 ** -23 * torch.maximum(b.abs(), b.abs().mean().expand_as(b))).all())


@pytest.mark.gpu
def test_sharded_adam_two_ranks(built):
    """2 GPUs: parameters after one sharded step (reduce-scatter + Adam + all-gather as one NVLS kernel) equal the
    single-GPU fused step on the summed gradients, on both ranks; typed-lane statistics all-reduce (sum / max / min)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run under gpurun --gpus 2)")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                        "127.0.0.1", "--master-port", "29611", os.path.join(ROOT, "tools", "check_sharded_adam.py")],
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and "SHARDED_ADAM_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
