"""ctypes binding of the CPU oracle (oracle/libmsgs_oracle.so).  TEST INFRASTRUCTURE: only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this."""
import ctypes
import os
import subprocess

import numpy as np

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_SO = os.path.join(_ROOT, "oracle", "libmsgs_oracle.so")


def build():
    subprocess.check_call(["make", "-s", "-C", os.path.join(_ROOT, "oracle"), "oracle"])


def _load():
    if not os.path.exists(_SO):
        build()
    lib = ctypes.CDLL(_SO)
    vp, i, f = ctypes.c_void_p, ctypes.c_int, ctypes.c_float
    lib.msgs_oracle_new.restype = vp
    lib.msgs_oracle_free.argtypes = [vp]
    lib.msgs_oracle_forward.restype = i
    lib.msgs_oracle_forward.argtypes = [vp, i, i, i, vp, i, i] + [vp] * 6 + [vp, f, vp, vp, vp, vp, vp, vp, f, f, i, i] + [vp] * 5
    lib.msgs_oracle_backward.restype = None
    lib.msgs_oracle_backward.argtypes = ([vp, vp] + [vp] * 6 + [vp, f, vp, vp, vp, vp, vp, vp, f, f, vp, i, i, vp]
                                         + [vp] * 9 + [i])
    lib.msgs_oracle_mark_visible.argtypes = [i, vp, vp, vp, vp]
    lib.msgs_oracle_buffer.restype = vp
    lib.msgs_oracle_buffer.argtypes = [vp, ctypes.c_char_p, ctypes.POINTER(ctypes.c_uint64)]
    lib.msgs_oracle_sum_contrib.restype = ctypes.c_uint64
    lib.msgs_oracle_sum_contrib.argtypes = [vp]
    return lib


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = _load()
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def _c(a, dt=np.float32):
    return None if a is None else np.ascontiguousarray(a, dtype=dt)


_BUF_DTYPES = {"depths": np.float32, "means2D": np.float32, "cov3D": np.float32, "conic_opacity": np.float32,
               "rgb": np.float32, "clamped": np.uint8, "tiles_touched": np.uint32, "point_offsets": np.uint32,
               "keys_unsorted": np.uint64, "vals_unsorted": np.uint32, "keys_sorted": np.uint64,
               "vals_sorted": np.uint32, "ranges": np.uint32, "final_T": np.float32, "n_contrib": np.uint32}


class Oracle:
    """One forward (+ optional backward) of the CPU restatement on numpy inputs."""

    def __init__(self):
        self.h = lib().msgs_oracle_new()

    def __del__(self):
        try:
            lib().msgs_oracle_free(self.h)
        except Exception:
            pass

    def forward(self, scene, cam, filter_small=False, filter_large=False, colors_precomp=None, cov3D_precomp=None):
        P, W, H = scene.P, cam.W, cam.H
        self.scene, self.cam = scene, cam
        self.fs, self.fl = int(filter_small), int(filter_large)
        self.colors_precomp = _c(colors_precomp)
        self.cov3D_precomp = _c(cov3D_precomp)
        use_sh = colors_precomp is None
        use_sr = cov3D_precomp is None
        self.M = scene.shs.shape[1] if use_sh else 0
        self.a = dict(bg=_c(scene.bg), means=_c(scene.means3D), shs=_c(scene.shs) if use_sh else None,
                      maxp=_c(scene.max_pixel_sizes), minp=_c(scene.min_pixel_sizes), op=_c(scene.opacities),
                      scales=_c(scene.scales) if use_sr else None, rot=_c(scene.rotations) if use_sr else None,
                      mask=_c(scene.base_mask, np.uint8), view=_c(cam.viewmatrix), proj=_c(cam.projmatrix),
                      campos=_c(cam.campos))
        a = self.a
        self.color = np.zeros((3, H, W), np.float32)
        self.acc_px = np.zeros((H, W), np.float32)
        self.depth = np.zeros((H, W), np.float32)
        self.radii = np.zeros(P, np.int32)
        self.pixel_sizes = np.zeros(P, np.float32)
        self.R = lib().msgs_oracle_forward(
            self.h, P, scene.sh_degree, self.M, _p(a["bg"]), W, H, _p(a["means"]), _p(a["shs"]),
            _p(self.colors_precomp), _p(a["maxp"]), _p(a["minp"]), _p(a["op"]), _p(a["scales"]),
            ctypes.c_float(scene.scale_modifier), _p(a["rot"]), _p(self.cov3D_precomp), _p(a["mask"]),
            _p(a["view"]), _p(a["proj"]), _p(a["campos"]), ctypes.c_float(cam.tanfovx), ctypes.c_float(cam.tanfovy),
            self.fs, self.fl, _p(self.color), _p(self.acc_px), _p(self.depth), _p(self.radii), _p(self.pixel_sizes))
        return self

    def backward(self, dL_dpix, acc64=False):
        s, cam, a, P, M = self.scene, self.cam, self.a, self.scene.P, self.M
        d = _c(dL_dpix)
        g = dict(mean2D=np.zeros((P, 3), np.float32), conic=np.zeros((P, 4), np.float32),
                 opacity=np.zeros((P, 1), np.float32), color=np.zeros((P, 3), np.float32),
                 mean3D=np.zeros((P, 3), np.float32), cov3D=np.zeros((P, 6), np.float32),
                 sh=np.zeros((P, max(M, 1), 3), np.float32), scale=np.zeros((P, 3), np.float32),
                 rot=np.zeros((P, 4), np.float32))
        lib().msgs_oracle_backward(
            self.h, _p(a["bg"]), _p(a["means"]), _p(a["shs"]), _p(self.colors_precomp), _p(self.pixel_sizes),
            _p(a["maxp"]), _p(a["minp"]), _p(a["scales"]), ctypes.c_float(s.scale_modifier), _p(a["rot"]),
            _p(self.cov3D_precomp), _p(a["mask"]), _p(a["view"]), _p(a["proj"]), _p(a["campos"]),
            ctypes.c_float(cam.tanfovx), ctypes.c_float(cam.tanfovy), _p(self.radii), self.fs, self.fl, _p(d),
            _p(g["mean2D"]), _p(g["conic"]), _p(g["opacity"]), _p(g["color"]), _p(g["mean3D"]), _p(g["cov3D"]),
            _p(g["sh"]), _p(g["scale"]), _p(g["rot"]), int(acc64))
        if M == 0:
            g["sh"] = np.zeros((P, 0, 3), np.float32)
        self.grads = g
        return g

    def buffer(self, name):
        n = ctypes.c_uint64(0)
        ptr = lib().msgs_oracle_buffer(self.h, name.encode(), ctypes.byref(n))
        dt = np.dtype(_BUF_DTYPES[name])
        if not ptr or n.value == 0:
            return np.zeros(0, dt)
        buf = (ctypes.c_char * (n.value * dt.itemsize)).from_address(ptr)
        return np.frombuffer(buf, dtype=dt).copy()

    @property
    def sum_contrib(self):
        return int(lib().msgs_oracle_sum_contrib(self.h))


def mark_visible(means3D, viewmatrix, projmatrix):
    m = _c(means3D)
    out = np.zeros(m.shape[0], np.uint8)
    lib().msgs_oracle_mark_visible(m.shape[0], _p(m), _p(_c(viewmatrix)), _p(_c(projmatrix)), _p(out))
    return out.astype(bool)
