"""ctypes binding of oracle/_ref/libref_rasterizer.so — the UNMODIFIED reference CUDA rasterizer
(/root/reference/submodules/diff-gaussian-rasterization/cuda_rasterizer/*.cu) compiled in place for sm_100a by
oracle/Makefile, plus the thin C shim oracle/ref_shim.cu.  TEST INFRASTRUCTURE (parity bar + speed bar): only
tests/, smoke() and bench.py's baseline legs may import this.  Needs a GPU."""
import ctypes
import os

import torch

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(_ROOT, "oracle", "_ref", "libref_rasterizer.so")

_ALLOC = ctypes.CFUNCTYPE(ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t)
_lib = None


def available():
    return os.path.exists(SO)


def lib():
    global _lib
    if _lib is None:
        l = ctypes.CDLL(SO)
        vp, i, f = ctypes.c_void_p, ctypes.c_int, ctypes.c_float
        l.ref_forward.restype = i
        l.ref_forward.argtypes = ([_ALLOC, vp, _ALLOC, vp, _ALLOC, vp, i, i, i, vp, i, i] + [vp] * 8
                                  + [vp, f, vp, vp, vp, vp, vp, vp, f, f, i, i, i, f] + [vp] * 5 + [i])
        l.ref_backward.restype = i
        l.ref_backward.argtypes = ([i, i, i, i, vp, i, i] + [vp] * 6 + [vp, f, vp, vp, vp, vp, vp, vp, f, f, vp]
                                   + [vp, vp, vp, i, i, f, vp] + [vp] * 11 + [i])
        l.ref_mark_visible.argtypes = [i, vp, vp, vp, vp]
        l.ref_geom_layout.argtypes = [vp, ctypes.c_size_t, ctypes.POINTER(ctypes.c_uint64)]
        l.ref_image_layout.argtypes = [vp, ctypes.c_size_t, ctypes.POINTER(ctypes.c_uint64)]
        l.ref_binning_layout.argtypes = [vp, ctypes.c_size_t, ctypes.POINTER(ctypes.c_uint64)]
        _lib = l
    return _lib


class _Buf:
    def __init__(self, dev):
        self.t = torch.empty(0, dtype=torch.uint8, device=dev)

        def cb(_ctx, n):
            self.t = torch.empty(int(n) + 256, dtype=torch.uint8, device=dev)
            return self.t.data_ptr()

        self.cb = _ALLOC(cb)


def _p(t):
    return None if t is None or t.numel() == 0 else t.data_ptr()


def _view(ptr, count, dtype, owner):
    """Tensor view of `count` elements at raw device address `ptr` inside byte tensor `owner`."""
    off = ptr - owner.data_ptr()
    nbytes = count * torch.empty(0, dtype=dtype).element_size()
    return owner[off:off + nbytes].view(dtype)


class RefRasterizer:
    """forward / backward of the reference CUDA core on torch CUDA tensors (legacy default stream, as the
    reference launches everything there — torch's default stream IS the legacy stream, so ordering against torch
    ops holds without explicit synchronisation).  sync=True (tests) brackets every call with
    torch.cuda.synchronize(); the bench passes sync=False so the reference is not charged for syncs it never makes
    (its own single blocking cudaMemcpy at rasterizer_impl.cu:305 remains)."""

    def __init__(self, sync=True):
        self.sync = sync

    def forward(self, t, cam, filter_small=False, filter_large=False, colors_precomp=None, cov3D_precomp=None):
        """t: dict of CUDA tensors (means3D, shs, opacities, scales, rotations, max_pixel_sizes, min_pixel_sizes,
        base_mask, occ_multiplier, dc_delta, bg, viewmatrix, projmatrix, campos) + sh_degree, scale_modifier."""
        dev = t["means3D"].device
        P, W, H = t["means3D"].shape[0], cam.W, cam.H
        self.t, self.cam, self.P, self.W, self.H = t, cam, P, W, H
        self.fs, self.fl = int(filter_small), int(filter_large)
        self.colors_precomp, self.cov3D_precomp = colors_precomp, cov3D_precomp
        self.shs = t["shs"] if colors_precomp is None else None
        self.scales = t["scales"] if cov3D_precomp is None else None
        self.rots = t["rotations"] if cov3D_precomp is None else None
        self.M = self.shs.shape[1] if self.shs is not None else 0
        f32 = dict(dtype=torch.float32, device=dev)
        self.color = torch.zeros((3, H, W), **f32)
        self.acc_px = torch.zeros((H, W), **f32)
        self.depth = torch.zeros((H, W), **f32)
        self.radii = torch.zeros((P,), dtype=torch.int32, device=dev)
        self.pixel_sizes = torch.zeros((P,), **f32)
        self.geom, self.binning, self.img = _Buf(dev), _Buf(dev), _Buf(dev)
        if self.sync:
            torch.cuda.synchronize()
        self.R = lib().ref_forward(
            self.geom.cb, None, self.binning.cb, None, self.img.cb, None, P, int(t["sh_degree"]), self.M,
            _p(t["bg"]), W, H, _p(t["means3D"]), _p(self.shs), _p(colors_precomp), _p(t["max_pixel_sizes"]),
            _p(t["min_pixel_sizes"]), _p(t["opacities"]), _p(t["occ_multiplier"]), _p(t["dc_delta"]),
            _p(self.scales), float(t["scale_modifier"]), _p(self.rots), _p(cov3D_precomp), _p(t["base_mask"]),
            _p(t["viewmatrix"]), _p(t["projmatrix"]), _p(t["campos"]), float(cam.tanfovx), float(cam.tanfovy),
            0, self.fs, self.fl, 1.0, _p(self.color), _p(self.acc_px), _p(self.depth), _p(self.radii),
            _p(self.pixel_sizes), 0)
        if self.sync:
            torch.cuda.synchronize()
        if self.R < 0:
            raise RuntimeError("reference forward failed")
        return self

    def backward(self, dL_dpix):
        t, cam, P, M = self.t, self.cam, self.P, self.M
        dev = dL_dpix.device
        z = lambda *s: torch.zeros(s, dtype=torch.float32, device=dev)   # rasterize_points.cu:184-194
        g = dict(mean2D=z(P, 3), conic=z(P, 2, 2), opacity=z(P, 1), occ=z(P, 4, 1), dc=z(P, 12, 1), color=z(P, 3),
                 mean3D=z(P, 3), cov3D=z(P, 6), sh=z(P, M, 3), scale=z(P, 3), rot=z(P, 4))
        if self.sync:
            torch.cuda.synchronize()
        rc = lib().ref_backward(
            P, int(t["sh_degree"]), M, self.R, _p(t["bg"]), self.W, self.H, _p(t["means3D"]), _p(self.shs),
            _p(self.colors_precomp), _p(self.pixel_sizes), _p(t["max_pixel_sizes"]), _p(t["min_pixel_sizes"]),
            _p(self.scales), float(t["scale_modifier"]), _p(self.rots), _p(self.cov3D_precomp), _p(t["base_mask"]),
            _p(t["viewmatrix"]), _p(t["projmatrix"]), _p(t["campos"]), float(cam.tanfovx), float(cam.tanfovy),
            _p(self.radii), _p(self.geom.t), _p(self.binning.t), _p(self.img.t), self.fs, self.fl, 1.0,
            _p(dL_dpix.contiguous()), _p(g["mean2D"]), _p(g["conic"]), _p(g["opacity"]), _p(g["occ"]), _p(g["dc"]),
            _p(g["color"]), _p(g["mean3D"]), _p(g["cov3D"]), _p(g["sh"]), _p(g["scale"]), _p(g["rot"]), 0)
        if self.sync:
            torch.cuda.synchronize()
        if rc != 0:
            raise RuntimeError("reference backward failed")
        self.grads = g
        return g

    # ---- private state of the reference, for bit-exact comparisons ----
    def state(self):
        P, R, N = self.P, self.R, self.W * self.H
        out = (ctypes.c_uint64 * 12)()
        lib().ref_geom_layout(self.geom.t.data_ptr(), P, out)
        g = self.geom.t
        s = dict(depths=_view(out[0], P, torch.float32, g), clamped=_view(out[1], 3 * P, torch.uint8, g),
                 means2D=_view(out[3], 2 * P, torch.float32, g).view(P, 2),
                 cov3D=_view(out[4], 6 * P, torch.float32, g).view(P, 6),
                 conic_opacity=_view(out[5], 4 * P, torch.float32, g).view(P, 4),
                 rgb=_view(out[8], 3 * P, torch.float32, g).view(P, 3),
                 tiles_touched=_view(out[9], P, torch.int32, g), point_offsets=_view(out[11], P, torch.int32, g))
        o3 = (ctypes.c_uint64 * 3)()
        lib().ref_image_layout(self.img.t.data_ptr(), N, o3)
        i = self.img.t
        T = ((self.W + 15) // 16) * ((self.H + 15) // 16)
        s.update(final_T=_view(o3[0], N, torch.float32, i), n_contrib=_view(o3[1], N, torch.int32, i),
                 ranges=_view(o3[2], 2 * T, torch.int32, i).view(T, 2))
        if R > 0:
            o4 = (ctypes.c_uint64 * 4)()
            lib().ref_binning_layout(self.binning.t.data_ptr(), R, o4)
            b = self.binning.t
            s.update(point_list=_view(o4[0], R, torch.int32, b), keys_sorted=_view(o4[2], R, torch.int64, b))
        return s


def scene_to_torch(scene, cam, device="cuda"):
    """numpy Scene/Camera (harness.scenes) -> dict of CUDA tensors in the layout both rasterizers take."""
    tt = lambda a, dt=torch.float32: torch.from_numpy(a).to(device=device, dtype=dt).contiguous()
    return dict(means3D=tt(scene.means3D), shs=tt(scene.shs), opacities=tt(scene.opacities), scales=tt(scene.scales),
                rotations=tt(scene.rotations), max_pixel_sizes=tt(scene.max_pixel_sizes),
                min_pixel_sizes=tt(scene.min_pixel_sizes), base_mask=tt(scene.base_mask, torch.bool),
                occ_multiplier=tt(scene.occ_multiplier), dc_delta=tt(scene.dc_delta), bg=tt(scene.bg),
                viewmatrix=tt(cam.viewmatrix), projmatrix=tt(cam.projmatrix), campos=tt(cam.campos),
                sh_degree=scene.sh_degree, scale_modifier=scene.scale_modifier)
