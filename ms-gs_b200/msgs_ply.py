"""ctypes binding of libmsgs_ply.so (include/msgs_ply.h): MS-GS model files straight into the rasterizer's input layout.

`load_ply(path)` replaces GaussianModel.load_ply (/root/reference/scene/gaussian_model.py:358-417) + get_features'
torch.cat (:127-153) for the forward-only viewer / render.py path (viewer.py:37-72, BASELINE configs[4]): the file is
memory-mapped and converted by host threads in one pass into pinned fp32 buffers, which travel to the device with one
asynchronous copy each.  Values are RAW parameters (pre-activation), as the reference's loader leaves them; the caller
applies exp / sigmoid / normalize as GaussianModel's properties do.  No fallback: raises if the library is missing."""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libmsgs_ply.so")
if not os.path.exists(_LIB_PATH):
    raise ImportError(f"{_LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'`")
_lib = ctypes.CDLL(_LIB_PATH)


class PlyInfo(ctypes.Structure):
    """msgs_ply_info"""
    _fields_ = [("num_points", ctypes.c_int64), ("sh_coeffs", ctypes.c_int), ("n_lvl_occ", ctypes.c_int),
                ("n_lvl_dc", ctypes.c_int), ("has_ms_attributes", ctypes.c_int), ("row_bytes", ctypes.c_int64),
                ("data_bytes", ctypes.c_int64)]


_vp = ctypes.c_void_p
_lib.msgs_ply_open.restype = _vp
_lib.msgs_ply_open.argtypes = [ctypes.c_char_p, ctypes.POINTER(PlyInfo)]
_lib.msgs_ply_read.restype = ctypes.c_int
_lib.msgs_ply_read.argtypes = [_vp] * 11 + [ctypes.c_int]
_lib.msgs_ply_close.argtypes = [_vp]
_lib.msgs_ply_last_error.restype = ctypes.c_char_p
lib = _lib


def _err(where):
    msg = _lib.msgs_ply_last_error()
    raise RuntimeError(f"{where}: {msg.decode() if msg else 'failed'}")


def load_ply(path, device=None, pin_memory=None, threads=0):
    """Returns (tensors, info): tensors = {"means3D" (P,3), "shs" (P,M,3), "opacity" (P,1), "scales" (P,3),
    "rotations" (P,4), "occ_multiplier" (P,L,1), "dc_delta" (P,3D,1), "base_mask" (P,) bool, "max_pixel_sizes" (P,),
    "min_pixel_sizes" (P,)} — fp32 host tensors (pinned when a device is given or pin_memory=True), or device tensors
    when `device` is given (copied with non_blocking=True on the current stream).  A plain 3DGS file (no MS-GS
    attributes) gets occ_multiplier = ones (P,1,1), dc_delta = zeros (P,3,1), base_mask = False, max / min = -1."""
    info = PlyInfo()
    h = _lib.msgs_ply_open(os.fsencode(path), ctypes.byref(info))
    if not h:
        _err("load_ply")
    try:
        P, M, L, D = int(info.num_points), int(info.sh_coeffs), int(info.n_lvl_occ), int(info.n_lvl_dc)
        pin = (device is not None and torch.cuda.is_available()) if pin_memory is None else bool(pin_memory)
        mk = lambda shape, dt=torch.float32: torch.empty(shape, dtype=dt, pin_memory=pin)
        t = {"means3D": mk((P, 3)), "shs": mk((P, M, 3)), "opacity": mk((P, 1)), "scales": mk((P, 3)),
             "rotations": mk((P, 4)), "occ_multiplier": mk((P, max(L, 1), 1)), "dc_delta": mk((P, 3 * max(D, 1), 1)),
             "base_mask": mk((P,), torch.uint8), "max_pixel_sizes": mk((P,)), "min_pixel_sizes": mk((P,))}
        if L == 0:
            t["occ_multiplier"].fill_(1.0)
        if D == 0:
            t["dc_delta"].zero_()
        ptr = lambda k, want=True: ctypes.c_void_p(t[k].data_ptr()) if (want and P > 0) else None
        rc = _lib.msgs_ply_read(h, ptr("means3D"), ptr("shs"), ptr("opacity"), ptr("scales"), ptr("rotations"),
                                ptr("occ_multiplier", L > 0), ptr("dc_delta", D > 0), ptr("base_mask"),
                                ptr("max_pixel_sizes"), ptr("min_pixel_sizes"), int(threads))
        if rc < 0:
            _err("load_ply")
    finally:
        _lib.msgs_ply_close(h)
    t["base_mask"] = t["base_mask"].view(torch.bool) if P > 0 else torch.zeros((0,), dtype=torch.bool)
    if device is not None:
        t = {k: v.to(device, non_blocking=True) for k, v in t.items()}
    return t, {"num_points": P, "sh_coeffs": M, "n_lvl_occ": L, "n_lvl_dc": D, "has_ms_attributes": bool(info.has_ms_attributes),
               "row_bytes": int(info.row_bytes), "data_bytes": int(info.data_bytes)}
