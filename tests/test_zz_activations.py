"""Fused parameter activations (include/msgs_activations.h).

CPU: the numpy oracle against golden vectors produced by the reference's own GaussianModel getters and their autograd
(tests/golden/make_activations_golden.py), the exported symbols and the ctypes signatures against the header.
GPU (-m gpu): the CUDA kernels through `msgs_activations.activate_gaussians` against the oracle and the golden vectors —
SH rows and their gradients bit-exact (pure data movement), exp / sigmoid / normalize within 4e-6 relative (device expf /
division differ from the host libm by a few ulp), gradients within 1e-5 relative."""
import ctypes
import os

import numpy as np
import pytest
import torch

from oracle import activations_oracle as ao

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden", "activations_golden.npz")
TAGS = ("deg3", "deg2", "deg0", "one")
RAW = ("_features_dc", "_features_rest", "_scaling", "_rotation", "_opacity")
OUTS = ("shs", "scales", "rotations", "opacities")


def _case(tag):
    z = np.load(GOLD)
    return ({k: z[f"{tag}/in/{k}"] for k in RAW}, {k: z[f"{tag}/out/{k}"] for k in OUTS},
            {k: z[f"{tag}/gout/{k}"] for k in OUTS}, {k: z[f"{tag}/grad/{k}"] for k in RAW})


def _close(a, b, rtol, what, row_scale=False):
    """|a - b| <= rtol * |b|; with row_scale (normalize backward: g - y (y.g) cancels) |b| is replaced by the largest
    magnitude in b's row, the scale the fp32 rounding of the cancelling terms has."""
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    assert a.shape == b.shape, f"{what}: shape {a.shape} != {b.shape}"
    ref = np.abs(b).max(axis=-1, keepdims=True) if row_scale and b.size else np.abs(b)
    err = np.abs(a - b) - rtol * ref
    assert (err <= 1e-30).all(), f"{what}: max excess {err.max():.3e} (rtol {rtol})"


@pytest.mark.parametrize("tag", TAGS)
def test_oracle_matches_reference_golden(tag):
    raw, outs, gouts, grads = _case(tag)
    got = dict(zip(OUTS, ao.activate(*(raw[k] for k in RAW))))
    assert np.array_equal(got["shs"], outs["shs"])
    for k in ("scales", "rotations", "opacities"):
        _close(got[k], outs[k], 1e-6, f"{tag} {k}")
    g = ao.activate_backward(outs["scales"], raw["_rotation"], outs["opacities"], gouts["shs"], gouts["scales"],
                             gouts["rotations"], gouts["opacities"])
    assert np.array_equal(g[0], grads["_features_dc"]) and np.array_equal(g[1], grads["_features_rest"])
    for got_g, k in zip(g[2:], RAW[2:]):
        _close(got_g, grads[k], 2e-5, f"{tag} grad {k}", row_scale=(k == "_rotation"))


def test_library_exports_and_signatures(built):
    import re
    import msgs_activations
    src = open(os.path.join(ROOT, "include", "msgs_activations.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    protos = dict(re.findall(r"\b(msgs_[a-z_]+)\s*\(([^()]*)\)\s*;", src))
    assert set(protos) == {"msgs_activate", "msgs_activate_backward"}
    lib = ctypes.CDLL(os.path.join(ROOT, "ms-gs_b200", "libmsgs_raster.so"))
    for name, args in protos.items():
        assert hasattr(lib, name), f"libmsgs_raster.so does not export {name}"
        kinds = ["ptr" if "*" in a else "int" for a in args.split(",")]
        got = ["ptr" if t is ctypes.c_void_p else "int" for t in getattr(msgs_activations._lib, name).argtypes]
        assert got == kinds, f"{name}: header {kinds} != binding {got}"


def test_cpu_tensors_are_rejected(built):
    import msgs_activations
    raw, *_ = _case("one")
    with pytest.raises(RuntimeError):
        msgs_activations.activate_gaussians(*(torch.from_numpy(raw[k]) for k in RAW))


@pytest.mark.gpu
@pytest.mark.parametrize("tag", TAGS)
def test_gpu_activations_match_oracle_and_golden(built, tag):
    import msgs_activations
    raw, outs, gouts, grads = _case(tag)
    t = {k: torch.from_numpy(raw[k]).cuda().requires_grad_(True) for k in RAW}
    got = dict(zip(OUTS, msgs_activations.activate_gaussians(*(t[k] for k in RAW))))
    sum((got[k] * torch.from_numpy(gouts[k]).cuda()).sum() for k in OUTS).backward()
    torch.cuda.synchronize()
    ora = dict(zip(OUTS, ao.activate(*(raw[k] for k in RAW))))
    assert np.array_equal(got["shs"].detach().cpu().numpy(), outs["shs"])
    for k in ("scales", "rotations", "opacities"):
        _close(got[k].detach().cpu().numpy(), outs[k], 4e-6, f"{tag} {k} vs golden")
        _close(got[k].detach().cpu().numpy(), ora[k], 4e-6, f"{tag} {k} vs oracle")
    assert np.array_equal(t["_features_dc"].grad.cpu().numpy(), grads["_features_dc"])
    assert np.array_equal(t["_features_rest"].grad.cpu().numpy(), grads["_features_rest"])
    og = ao.activate_backward(got["scales"].detach().cpu().numpy(), raw["_rotation"], got["opacities"].detach().cpu().numpy(),
                              gouts["shs"], gouts["scales"], gouts["rotations"], gouts["opacities"])
    for k, o in zip(RAW[2:], og[2:]):
        _close(t[k].grad.cpu().numpy(), grads[k], 2e-5, f"{tag} grad {k} vs golden", row_scale=(k == "_rotation"))
        _close(t[k].grad.cpu().numpy(), o, 2e-5, f"{tag} grad {k} vs oracle", row_scale=(k == "_rotation"))


@pytest.mark.gpu
def test_gpu_activations_full_size_and_partial_grads(built):
    """C2 size (1.2 M Gaussians, D = 15) against torch's own ops on the device; only some inputs require a gradient;
    P = 0 is a no-op."""
    import msgs_activations
    g = torch.Generator(device="cuda").manual_seed(3)
    P, D = 1_200_000, 15
    dc = torch.randn(P, 1, 3, device="cuda", generator=g)
    rest = torch.randn(P, D, 3, device="cuda", generator=g)
    s = (torch.randn(P, 3, device="cuda", generator=g) * 2 - 3).requires_grad_(True)
    r = torch.randn(P, 4, device="cuda", generator=g).requires_grad_(True)
    o = torch.randn(P, 1, device="cuda", generator=g) * 3
    shs, sc, ro, op = msgs_activations.activate_gaussians(dc, rest, s, r, o)
    assert torch.equal(shs, torch.cat((dc, rest), dim=1))
    s2, r2 = s.detach().clone().requires_grad_(True), r.detach().clone().requires_grad_(True)
    sc2, ro2, op2 = torch.exp(s2), torch.nn.functional.normalize(r2), torch.sigmoid(o)
    for a, b in ((sc, sc2), (ro, ro2), (op, op2)):
        assert float(((a - b).abs() - 4e-6 * b.abs()).max()) <= 1e-30
    w1, w2 = torch.randn_like(sc), torch.randn_like(ro)
    ((sc * w1).sum() + (ro * w2).sum()).backward()
    ((sc2 * w1).sum() + (ro2 * w2).sum()).backward()
    for a, b in ((s.grad, s2.grad), (r.grad, r2.grad)):
        rel = float((a.double() - b.double()).norm() / (b.double().norm() + 1e-30))
        assert rel < 1e-5, rel
    e = torch.empty
    out = msgs_activations.activate_gaussians(e(0, 1, 3, device="cuda"), e(0, D, 3, device="cuda"), e(0, 3, device="cuda"),
                                              e(0, 4, device="cuda"), e(0, 1, device="cuda"))
    assert [tuple(x.shape) for x in out] == [(0, 16, 3), (0, 3), (0, 4), (0, 1)]


@pytest.fixture(scope="module")
def host_bodies(tmp_path_factory):
    """The product kernels' per-thread bodies (csrc/activations_body.cuh) compiled as HOST code."""
    import subprocess
    so = tmp_path_factory.mktemp("act") / "libact_host.so"
    subprocess.check_call(["nvcc", "-O1", "-std=c++17", "-shared", "-Xcompiler", "-fPIC,-ffp-contract=off", "-o", str(so),
                           os.path.join(ROOT, "tests", "activations_host_harness.cu")])
    lib = ctypes.CDLL(str(so))
    lib.host_activate.argtypes = [ctypes.c_int] * 2 + [ctypes.c_void_p] * 9
    lib.host_activate_backward.argtypes = [ctypes.c_int] * 2 + [ctypes.c_void_p] * 12
    lib.host_activate.restype = lib.host_activate_backward.restype = None
    return lib


def _aligned(shape, fill=np.nan):
    """fp32 array whose data pointer is 16-byte aligned (the bodies use float4 accesses), NaN-filled plus a guard tail."""
    n = int(np.prod(shape))
    raw = np.full(n + 8 + 4, fill, np.float32)
    off = (-raw.ctypes.data // 4) % 4
    return raw[off:off + n].reshape(shape), raw[off + n:off + n + 8]


def _in(a):
    out, _ = _aligned(a.shape)
    out[...] = a
    return out


@pytest.mark.parametrize("tag", TAGS)
def test_thread_bodies_on_host(host_bodies, tag):
    """Same source as the CUDA kernels, executed thread by thread on the host: SH rows and their gradients bit-exact
    against the reference-generated golden vectors, every output element written exactly within bounds (NaN canaries and
    guard tails), activations / gradients within the GPU test's tolerances; NULL gradient pointers = zero / skipped."""
    raw, outs, gouts, grads = _case(tag)
    P, D = raw["_features_dc"].shape[0], raw["_features_rest"].shape[1]
    ins = [_in(raw[k]) for k in RAW]
    shapes = {"shs": (P, 1 + D, 3), "scales": (P, 3), "rotations": (P, 4), "opacities": (P, 1)}
    res = {k: _aligned(s) for k, s in shapes.items()}
    p = lambda a: ctypes.c_void_p(a.ctypes.data) if a is not None else ctypes.c_void_p(0)   # noqa: E731
    host_bodies.host_activate(P, D, *(p(a) for a in ins), *(p(res[k][0]) for k in OUTS))
    for k in OUTS:
        assert not np.isnan(res[k][0]).any(), f"{k}: element not written"
        assert np.isnan(res[k][1]).all(), f"{k}: write past the end"
    assert np.array_equal(res["shs"][0], outs["shs"])
    for k in ("scales", "rotations", "opacities"):
        _close(res[k][0], outs[k], 4e-6, f"{tag} {k}")
    gin = [_in(gouts[k]) for k in OUTS]
    gres = [_aligned(raw[k].shape) for k in RAW]
    host_bodies.host_activate_backward(P, D, p(res["scales"][0]), p(ins[3]), p(res["opacities"][0]), *(p(a) for a in gin),
                                       *(p(a[0]) for a in gres))
    for k, (a, guard) in zip(RAW, gres):
        assert not np.isnan(a).any() and np.isnan(guard).all(), f"grad {k}: not written / overrun"
    assert np.array_equal(gres[0][0], grads["_features_dc"]) and np.array_equal(gres[1][0], grads["_features_rest"])
    for k, (a, _) in list(zip(RAW, gres))[2:]:
        _close(a, grads[k], 2e-5, f"{tag} grad {k}", row_scale=(k == "_rotation"))
    # NULL output gradients read as zero; NULL input-gradient pointers are skipped
    g2 = [_aligned(raw[k].shape) for k in RAW]
    host_bodies.host_activate_backward(P, D, p(res["scales"][0]), p(ins[3]), p(res["opacities"][0]), p(None), p(gin[1]),
                                       p(None), p(None), p(g2[0][0]), p(None), p(g2[2][0]), p(g2[3][0]), p(g2[4][0]))
    assert not g2[0][0].any() and not g2[3][0].any() and not g2[4][0].any()
    assert np.isnan(g2[1][0]).all()
    assert np.array_equal(g2[2][0], gres[2][0])
