"""Fused L1 + SSIM training loss (include/msgs_loss.h) against golden vectors produced by the
reference's own utils/loss_utils.py and autograd (tests/golden/make_loss_golden.py).  CPU: the kernels' thread bodies run
on the host, exports / signatures, window taps.  GPU (-m gpu): the CUDA kernels through `msgs_loss.l1_ssim_loss`.
Tolerances (fp32 summation order of the separable 11 + 11-tap windows differs from conv2d's): loss 2e-5 relative; image gradient
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
    lib.host_l1_ssim_backward.argtypes = ([ctypes.c_int] * 3 + [ctypes.c_float] + [ctypes.c_void_p] * 6 + [ctypes.c_float] * 2
                                          + [ctypes.c_void_p])
    lib.host_l1_ssim_forward.restype, lib.host_l1_ssim_backward.restype = ctypes.c_double, None
    return lib


@pytest.mark.parametrize("tag", TAGS)
def test_thread_bodies_on_host(host_lib, tag):
    img, gt, lam, w = Z[f"{tag}/img"].copy(), Z[f"{tag}/gt"].copy(), float(Z[f"{tag}/lam"]), Z["w1d"].copy()
    C, H, W = img.shape
    maps = np.full((4, C, H, W), np.nan, np.float32)
    p = lambda a: ctypes.c_void_p(a.ctypes.data)   # noqa: E731
    mean = host_lib.host_l1_ssim_forward(C, H, W, lam, p(img), p(gt), p(w), p(maps[0]), p(maps[1]), p(maps[2]), p(maps[3]))
    assert not np.isnan(maps).any()
    grad = np.full((C, H, W), np.nan, np.float32)
    host_lib.host_l1_ssim_backward(C, H, W, lam, p(img), p(gt), p(w), p(maps[1]), p(maps[2]), p(maps[3]), 1.7, 1.0, p(grad))
    assert not np.isnan(grad).any()
    _check(tag, float(maps[0].astype(np.float64).mean()), grad)
    assert abs(mean - float(maps[0].astype(np.float64).mean())) < 1e-6          # the in-kernel reduction's operand
    # the per-image weight (train.py:212-214's multiplier) scales the gradient and nothing else
    g01 = np.empty_like(grad)
    host_lib.host_l1_ssim_backward(C, H, W, lam, p(img), p(gt), p(w), p(maps[1]), p(maps[2]), p(maps[3]), 17.0, 0.1, p(g01))
    assert np.abs(g01 - grad).max() <= 1e-6 * np.abs(grad).max()


def test_exports_signatures_and_taps(built):
    import re
    import msgs_loss
    src = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "msgs_loss.h")).read(), flags=re.S)
    protos = dict(re.findall(r"\b(msgs_[a-z0-9_]+)\s*\(([^()]*)\)\s*;", src))
    assert set(protos) == {"msgs_l1_ssim_forward", "msgs_l1_ssim_backward", "msgs_l1_ssim_forward_batch",
                           "msgs_l1_ssim_backward_batch"}
    lib = ctypes.CDLL(os.path.join(ROOT, "ms-gs_b200", "libmsgs_raster.so"))
    kind = {ctypes.c_void_p: "ptr", ctypes.c_int: "int", ctypes.c_float: "float"}
    for name, args in protos.items():
        assert hasattr(lib, name)
        kinds = ["ptr" if "*" in a else ("float" if a.strip().startswith("float") else "int") for a in args.split(",")]
        assert [kind[t] for t in getattr(msgs_loss._lib, name).argtypes] == kinds, name
    # the descriptor struct mirrors msgs_loss_image field by field
    body = re.search(r"typedef struct \{(.*?)\} msgs_loss_image;", src, re.S).group(1)
    names = []
    for decl in body.split(";"):
        decl = decl.replace("const", "").strip()
        if decl:
            names += [n.strip(" *") for n in decl.split(None, 1)[1].split(",")]
    assert names == [f[0] for f in msgs_loss._LossImage._fields_], names
    assert ctypes.sizeof(msgs_loss._LossImage) == 16 + 7 * 8
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


@pytest.mark.gpu
def test_gpu_batched_loss_matches_reference_golden(built):
    """All golden images in ONE forward and ONE backward launch, with train.py:212-214's multipliers."""
    import msgs_loss
    lam = float(Z[f"{TAGS[0]}/lam"])
    tags = [t for t in TAGS if float(Z[f"{t}/lam"]) == lam]
    assert len(tags) >= 2
    imgs = [torch.from_numpy(Z[f"{t}/img"]).cuda().requires_grad_(True) for t in tags]
    gts = [torch.from_numpy(Z[f"{t}/gt"]).cuda() for t in tags]
    weights = [1.0] + [0.1] * (len(tags) - 1)
    total, per = msgs_loss.l1_ssim_loss_batch(imgs, gts, weights, lam)
    (total * 1.7).backward()
    torch.cuda.synchronize()
    want_total = 0.0
    for t, im, w, l in zip(tags, imgs, weights, per.tolist()):
        want_g = Z[f"{t}/grad_x1.7"] * w
        want_l = float(Z[f"{t}/loss"]) * w
        want_total += want_l
        assert abs(l - want_l) <= 2e-5 * abs(want_l) + 1e-7, t
        assert np.abs(im.grad.cpu().numpy() - want_g).max() <= 2e-4 * np.abs(want_g).max() + 1e-9, t
    assert abs(float(total) - want_total) <= 2e-5 * want_total


@pytest.mark.gpu
def test_gpu_batched_loss_large_image_against_torch_conv(built):
    """A full-size image (tile borders, ragged right/bottom tiles) against the conv2d formulation of utils/loss_utils.py:39-64
    restated with torch ops on the GPU (the restatement itself is pinned by the golden vectors above in
    tests/golden/make_loss_golden.py's terms: same window, same padding, same constants)."""
    import torch.nn.functional as F
    import msgs_loss
    g = torch.Generator(device="cuda").manual_seed(3)
    C, H, W = 3, 545, 980
    gt = torch.rand((C, H, W), device="cuda", generator=g)
    img = (gt + 0.1 * torch.randn((C, H, W), device="cuda", generator=g)).clamp(0, 1).requires_grad_(True)
    w1 = msgs_loss.window_taps("cuda")
    win = (w1[:, None] @ w1[None, :]).expand(C, 1, 11, 11).contiguous()

    def ref(a, b, lam=0.2):
        conv = lambda t: F.conv2d(t[None], win, padding=5, groups=C)[0]      # noqa: E731
        mu1, mu2 = conv(a), conv(b)
        s1, s2, s12 = conv(a * a) - mu1 * mu1, conv(b * b) - mu2 * mu2, conv(a * b) - mu1 * mu2
        ssim = ((2 * mu1 * mu2 + 0.01 ** 2) * (2 * s12 + 0.03 ** 2)) / ((mu1 * mu1 + mu2 * mu2 + 0.01 ** 2) * (s1 + s2 + 0.03 ** 2))
        return (1 - lam) * (a - b).abs().mean() + lam * (1 - ssim.mean())

    prev = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        want = ref(img, gt)
        want_g, = torch.autograd.grad(want, img)
    finally:
        torch.backends.cudnn.allow_tf32 = prev
    total, per = msgs_loss.l1_ssim_loss_batch([img], [gt], [1.0], 0.2)
    total.backward()
    assert abs(float(total) - float(want)) <= 2e-5 * float(want)
    assert float((img.grad - want_g).abs().max()) <= 2e-4 * float(want_g.abs().max())
