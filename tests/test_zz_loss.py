"""Fused L1 + SSIM training loss (include/msgs_loss.h) against golden vectors produced by the
reference's own utils/loss_utils.py and autograd (tests/golden/make_loss_golden.py).  CPU: the kernels' thread bodies run
on the host, exports / signatures, window taps.  GPU (-m gpu): the CUDA kernels through `msgs_loss.l1_ssim_loss`.
Tolerances (fp32 summation order of the 121-tap windows differs from conv2d's): loss 2e-5 relative; image gradient
|diff| <= 2e-4 * max|gradient|."""
import ctypes
import os
import subprocess

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
Z = np.load(os.path.join(ROOT, "tests", "golden", "loss_golden.npz"))
TAGS = ("c3_40x52", "c3_7x9", "c1_1x1", "c3_23x11_l0", "c3_16x16_l1")


def _check(tag, loss, grad):
    want_l, want_g = float(Z[f"{tag}/loss"]), Z[f"{tag}/grad_x1.7"]
    assert abs(loss - want_l) <= 2e-5 * abs(want_l) + 1e-7, f"{tag}: loss {loss} vs {want_l}"
    assert grad.shape == want_g.shape
    assert np.abs(grad - want_g).max() <= 2e-4 * np.abs(want_g).max() + 1e-9, f"{tag}: gradient"


@pytest.fixture(scope="module")
def host_lib(tmp_path_factory):
    so = tmp_path_factory.mktemp("loss") / "libloss_host.so"
    subprocess.check_call(["nvcc", "-O1", "-std=c++17", "-shared", "-Xcompiler", "-fPIC,-ffp-contract=off", "-o", str(so),
                           os.path.join(ROOT, "tests", "loss_host_harness.cu")])
    lib = ctypes.CDLL(str(so))
    lib.host_l1_ssim_forward.argtypes = [ctypes.c_int] * 3 + [ctypes.c_float] + [ctypes.c_void_p] * 7
    lib.host_l1_ssim_backward.argtypes = [ctypes.c_int] * 3 + [ctypes.c_float] + [ctypes.c_void_p] * 8
    lib.host_l1_ssim_forward.restype = lib.host_l1_ssim_backward.restype = None
    return lib


@pytest.mark.parametrize("tag", TAGS)
def test_thread_bodies_on_host(host_lib, tag):
    img, gt, lam, w = Z[f"{tag}/img"].copy(), Z[f"{tag}/gt"].copy(), float(Z[f"{tag}/lam"]), Z["w1d"].copy()
    C, H, W = img.shape
    maps = np.full((4, C, H, W), np.nan, np.float32)
    p = lambda a: ctypes.c_void_p(a.ctypes.data)   # noqa: E731
    host_lib.host_l1_ssim_forward(C, H, W, lam, p(img), p(gt), p(w), p(maps[0]), p(maps[1]), p(maps[2]), p(maps[3]))
    assert not np.isnan(maps).any()
    up, grad = np.array([1.7], np.float32), np.full((C, H, W), np.nan, np.float32)
    host_lib.host_l1_ssim_backward(C, H, W, lam, p(img), p(gt), p(w), p(maps[1]), p(maps[2]), p(maps[3]), p(up), p(grad))
    assert not np.isnan(grad).any()
    _check(tag, float(maps[0].astype(np.float64).mean()), grad)


def test_exports_signatures_and_taps(built):
    import re
    import msgs_loss
    src = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "msgs_loss.h")).read(), flags=re.S)
    protos = dict(re.findall(r"\b(msgs_[a-z0-9_]+)\s*\(([^()]*)\)\s*;", src))
    assert set(protos) == {"msgs_l1_ssim_forward", "msgs_l1_ssim_backward"}
    lib = ctypes.CDLL(os.path.join(ROOT, "ms-gs_b200", "libmsgs_raster.so"))
    kind = {ctypes.c_void_p: "ptr", ctypes.c_int: "int", ctypes.c_float: "float"}
    for name, args in protos.items():
        assert hasattr(lib, name)
        kinds = ["ptr" if "*" in a else ("float" if a.strip().startswith("float") else "int") for a in args.split(",")]
        assert [kind[t] for t in getattr(msgs_loss._lib, name).argtypes] == kinds, name
    assert np.array_equal(msgs_loss.window_taps("cpu").numpy(), Z["w1d"])
    with pytest.raises(RuntimeError):
        msgs_loss.l1_ssim_loss(torch.zeros(3, 4, 4), torch.zeros(3, 4, 4))


@pytest.mark.gpu
@pytest.mark.parametrize("tag", TAGS)
def test_gpu_loss_matches_reference_golden(built, tag):
    import msgs_loss
    img = torch.from_numpy(Z[f"{tag}/img"]).cuda().requires_grad_(True)
    gt = torch.from_numpy(Z[f"{tag}/gt"]).cuda()
    loss = msgs_loss.l1_ssim_loss(img, gt, float(Z[f"{tag}/lam"]))
    (loss * 1.7).backward()
    torch.cuda.synchronize()
    _check(tag, float(loss), img.grad.cpu().numpy())
