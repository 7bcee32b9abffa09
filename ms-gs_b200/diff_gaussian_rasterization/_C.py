"""ctypes binding of libmsgs_raster.so — the `_C` module of the drop-in `diff_gaussian_rasterization` package.

Mirrors the three functions the reference registers through pybind11
(/root/reference/submodules/diff-gaussian-rasterization/ext.cpp:15-19, rasterize_points.h:18-85) with the
same argument order, the same returned tuples and the same error behaviour (RuntimeError when means3D is not
(N,3), rasterize_points.cu:66-68).  PyTorch is used for device memory and streams only; every computation runs
in the hand-written sm_100a kernels of ms-gs_b200/csrc through the C ABI of include/msgs_raster.h.

There is NO fallback: if the shared library is missing or does not load, importing this module raises.
"""
import ctypes
import os
import threading

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(os.path.dirname(_HERE), "libmsgs_raster.so")

if not os.path.exists(_LIB_PATH):
    raise ImportError(
        f"{_LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
        f"(or `make -C ms-gs_b200/csrc`). The B200 rasterizer has no CPU / PyTorch fallback."
    )
_lib = ctypes.CDLL(_LIB_PATH)

_ALLOC = ctypes.CFUNCTYPE(ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t)
_vp, _i, _f, _sz = ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_size_t

_lib.msgs_abi_version.restype = _i
_lib.msgs_last_error.restype = ctypes.c_char_p
_lib.msgs_mark_visible.restype = _i
_lib.msgs_mark_visible.argtypes = [_i, _vp, _vp, _vp, _vp, _vp]
_lib.msgs_forward.restype = _i
_lib.msgs_forward.argtypes = (
    [_ALLOC, _vp, _ALLOC, _vp, _ALLOC, _vp, _i, _i, _i, _vp, _i, _i]      # allocators, P D M, bg, W, H
    + [_vp] * 8 + [_vp, _f, _vp, _vp, _vp]                                   # means3D..dc_delta, scales, mod, rot, cov3D, mask
    + [_vp, _vp, _vp, _f, _f, _i, _i, _i, _f]                                # view, proj, campos, tanfov, flags, fade
    + [_vp] * 5 + [_i, _vp]                                                  # outputs, debug, stream
)
_lib.msgs_forward_begin.restype = _i
_lib.msgs_forward_begin.argtypes = _lib.msgs_forward.argtypes
_lib.msgs_forward_finish.restype = _i
_lib.msgs_forward_finish.argtypes = [_i]


class _GroupUnit(ctypes.Structure):
    """msgs_group_unit (include/msgs_raster.h)"""
    _fields_ = [("width", _i), ("height", _i), ("background", _vp), ("out_color", _vp), ("out_acc_pixel_size", _vp),
                ("out_depth", _vp), ("radii", _vp), ("pixel_sizes", _vp), ("geom_alloc", _ALLOC), ("geom_ctx", _vp),
                ("binning_alloc", _ALLOC), ("binning_ctx", _vp), ("img_alloc", _ALLOC), ("img_ctx", _vp), ("stream", _vp)]


_lib.msgs_forward_group_begin.restype = _i
_lib.msgs_forward_group_begin.argtypes = (
    [_i, ctypes.POINTER(_GroupUnit), ctypes.POINTER(_i), _i, _i, _i]           # n_units, units, handles, P D M
    + [_vp] * 8 + [_vp, _f, _vp, _vp, _vp]                                   # means3D..dc_delta, scales, mod, rot, cov3D, mask
    + [_vp, _vp, _vp, _f, _f, _i, _i, _i, _f, _i]                            # view, proj, campos, tanfov, flags, fade, debug
)


class _GroupBwdUnit(ctypes.Structure):
    """msgs_group_bwd_unit (include/msgs_raster.h)"""
    _fields_ = [("width", _i), ("height", _i), ("radii", _vp), ("pixel_sizes", _vp), ("geom_buffer", _vp),
                ("dL_dmean2D", _vp)]


_lib.msgs_backward_group_accumulate.restype = _i
_lib.msgs_backward_group_accumulate.argtypes = (
    [_i, ctypes.POINTER(_GroupBwdUnit), _i, _i, _i]                          # n_units, units, P D M
    + [_vp] * 4 + [_vp, _f, _vp, _vp, _vp]                                   # means3D, sh, max/min px, scales, mod, rot, cov3D, mask
    + [_vp, _vp, _vp, _f, _f, _i, _i]                                        # view, proj, campos, tanfov, flags
    + [_vp] * 5 + [_i, _vp]                                                  # accumulators, debug, stream
)


class _GroupBwdBlend(ctypes.Structure):
    """msgs_group_bwd_blend (include/msgs_raster.h)"""
    _fields_ = [("background", _vp), ("binning_buffer", _vp), ("image_buffer", _vp), ("R", _i), ("dL_dpix", _vp), ("stream", _vp)]


_lib.msgs_backward_group_full.restype = _i
_lib.msgs_backward_group_full.argtypes = ([_vp, ctypes.POINTER(_GroupBwdBlend)] + list(_lib.msgs_backward_group_accumulate.argtypes))
_lib.msgs_backward_group_blend.restype = _i
_lib.msgs_backward_group_blend.argtypes = _lib.msgs_backward_group_full.argtypes
_lib.msgs_backward_group_accumulate_part.restype = _i
_lib.msgs_backward_group_accumulate_part.argtypes = ([_vp, _i, _i] + list(_lib.msgs_backward_group_accumulate.argtypes))
_lib.msgs_backward.restype = _i
_lib.msgs_backward.argtypes = (
    [_i, _i, _i, _i, _vp, _i, _i]                                            # P D M R bg W H
    + [_vp] * 6 + [_vp, _f, _vp, _vp, _vp]                                   # means3D..min_px, scales, mod, rot, cov3D, mask
    + [_vp, _vp, _vp, _f, _f, _vp]                                           # view, proj, campos, tanfov, radii
    + [_vp, _vp, _vp, _i, _i, _f, _vp]                                       # buffers, flags, fade, dL_dpix
    + [_vp] * 11 + [_i, _vp]                                                 # grads, debug, stream
)
_lib.msgs_backward_accumulate.restype = _i
_lib.msgs_backward_accumulate.argtypes = _lib.msgs_backward.argtypes
_lib.msgs_backward_accumulate_part.restype = _i
_lib.msgs_backward_accumulate_part.argtypes = _lib.msgs_backward.argtypes + [_i, _i, _i]
# raw-parameter entry points: same arguments (+ the raw opacities in the backward)
_lib.msgs_forward_begin_raw.restype = _i
_lib.msgs_forward_begin_raw.argtypes = _lib.msgs_forward.argtypes
_lib.msgs_forward_group_begin_raw.restype = _i
_lib.msgs_forward_group_begin_raw.argtypes = _lib.msgs_forward_group_begin.argtypes
_lib.msgs_backward_raw.restype = _i
_lib.msgs_backward_raw.argtypes = [_vp, _i, _i] + _lib.msgs_backward.argtypes
_lib.msgs_backward_group_accumulate_raw.restype = _i
_lib.msgs_backward_group_accumulate_raw.argtypes = [_vp] + _lib.msgs_backward_group_accumulate.argtypes
_lib.msgs_speculation_stats.restype = None
_lib.msgs_speculation_stats.argtypes = [ctypes.POINTER(ctypes.c_longlong), _i]
for _n in ("msgs_geom_bytes", "msgs_binning_bytes", "msgs_sort_temp_bytes"):
    getattr(_lib, _n).restype = _sz
    getattr(_lib, _n).argtypes = [_i]
_lib.msgs_image_bytes.restype = _sz
_lib.msgs_image_bytes.argtypes = [_i, _i, _i]
_lib.msgs_geom_layout.argtypes = [_vp, _i, ctypes.POINTER(ctypes.c_uint64)]
_lib.msgs_image_layout.argtypes = [_vp, _i, _i, _i, ctypes.POINTER(ctypes.c_uint64)]
_lib.msgs_binning_layout.argtypes = [_vp, _i, ctypes.POINTER(ctypes.c_uint64)]
_lib.msgs_sort_pairs.restype = _i
_lib.msgs_sort_pairs.argtypes = [_vp] * 6 + [_i, _i, _vp, _vp]
# the same sort with its scratch in a static device array inside the library (nothing allocated; n <= 2^21)
_lib.msgs_sort_pairs_static.restype = _i
_lib.msgs_sort_pairs_static.argtypes = [_vp] * 6 + [_i, _i, _vp]

if _lib.msgs_abi_version() != 1:
    raise ImportError("libmsgs_raster.so ABI version mismatch")

lib = _lib          # exposed for tests / bench
LIB_PATH = _LIB_PATH


def _err(code, where):
    msg = _lib.msgs_last_error()
    raise RuntimeError(f"{where} failed ({code}): {msg.decode() if msg else ''}")


def _ptr(t):
    """Device pointer of a tensor, or NULL for an empty tensor (the reference's `optional` convention)."""
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()


def _cuda_f32(t):
    return t.contiguous()


class _FixedBuffer(ctypes.Structure):
    """msgs_fixed_buffer (include/msgs_raster.h)"""
    _fields_ = [("ptr", ctypes.c_void_p), ("bytes", ctypes.c_size_t), ("fallback", _ALLOC), ("fallback_ctx", ctypes.c_void_p),
                ("requested", ctypes.c_size_t), ("served_fixed", _i)]


_lib.msgs_fixed_alloc.restype = ctypes.c_void_p
_lib.msgs_fixed_alloc.argtypes = [ctypes.c_void_p, ctypes.c_size_t]
_FIXED_ALLOC = ctypes.cast(_lib.msgs_fixed_alloc, _ALLOC)
_lib.msgs_binning_hint_bytes.restype = ctypes.c_size_t
_lib.msgs_binning_hint_bytes.argtypes = [_i, _i, _i]
_size_cache = {}


def _geom_bytes(P):
    k = ("g", P)
    if k not in _size_cache:
        _size_cache[k] = int(_lib.msgs_geom_bytes(P))
    return _size_cache[k]


def _image_bytes(W, H, P):
    k = ("i", W, H, P)
    if k not in _size_cache:
        _size_cache[k] = int(_lib.msgs_image_bytes(W, H, P))
    return _size_cache[k]


class _Buffer:
    """One resizable byte tensor handed to the core through an allocator callback (the analogue of resizeFunctional,
    rasterize_points.cu:27-33).  `nbytes` > 0: the tensor is allocated up front and the core is given the library's own
    msgs_fixed_alloc with it — no call back into Python when the request fits (geometry / image buffers: always; binning
    buffer: when the speculative size was guessed from msgs_binning_hint_bytes); any other request reaches `_alloc`."""

    def __init__(self, device, nbytes=0):
        self.device = device
        self.tensor = torch.empty(int(nbytes), dtype=torch.uint8, device=device)

        def _alloc(_ctx, n):
            self.tensor = torch.empty(int(n), dtype=torch.uint8, device=self.device)
            return self.tensor.data_ptr()

        self.cb_py = _ALLOC(_alloc)
        self.fixed = _FixedBuffer(self.tensor.data_ptr() if nbytes else None, int(nbytes), self.cb_py, None, 0, 0)
        self.cb, self.ctx = _FIXED_ALLOC, ctypes.addressof(self.fixed)

    def release(self):
        """The tensor, with the callback dropped: the callback's closure refers back to this object, and a buffer left
        in that reference cycle keeps its (hundreds of MB of) device memory alive until the cyclic GC runs."""
        t, self.tensor, self.cb_py, self.fixed = self.tensor, None, None, None
        return t


def _stream_ptr(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


class PendingForward:
    """A forward call whose kernels are enqueued but whose num_rendered has not been read back yet
    (msgs_forward_begin / msgs_forward_finish).  finish() returns what `rasterize_gaussians` returns."""

    def __init__(self, handle, device, outs, bufs, keep):
        self.handle, self.device, self.outs, self.bufs, self.keep = handle, device, outs, bufs, keep
        self.thread = threading.get_ident()        # handles index a per-host-thread table of the library

    def __del__(self):
        if getattr(self, "handle", None) is not None:      # never finished: give the slot back
            try:
                # only the creating thread owns the slot (a GC / autograd thread would index ITS table: skip, no CUDA call)
                if threading.get_ident() == self.thread:
                    _lib.msgs_forward_finish(self.handle)
                for b in self.bufs:
                    b.release()
            except Exception:
                pass

    def finish(self):
        if self.handle is None:
            raise RuntimeError("PendingForward.finish() called twice")
        if threading.get_ident() != self.thread:
            raise RuntimeError("PendingForward.finish() must be called from the host thread that began the forward")
        h, self.handle = self.handle, None
        with torch.cuda.device(self.device):
            rendered = _lib.msgs_forward_finish(h)
        if rendered < 0:
            _err(rendered, "rasterize_gaussians")
        state = tuple(b.release() for b in self.bufs)
        outs, self.outs, self.bufs, self.keep = self.outs, None, None, None
        return (rendered, *outs, *state)


def rasterize_gaussians_begin(background, means3D, colors, max_pixel_sizes, min_pixel_sizes, opacity, occ_multiplier,
                              dc_delta, scales, rotations, scale_modifier, cov3D_precomp, base_mask, viewmatrix,
                              projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos, prefiltered,
                              filter_small, filter_large, fade_size, debug, raw=False):
    """Extension: first half of `rasterize_gaussians` (same arguments).  raw=True (msgs_forward_begin_raw): `scales`,
    `rotations`, `opacity` are GaussianModel's raw parameters; exp / normalize / sigmoid run inside the preprocess kernel.  Everything is enqueued on the current stream —
    placement + blend speculatively — and a PendingForward is returned without waiting for the device, so the forwards
    of several independent units (on their own streams) can be in flight before the first read-back.  finish() must be
    called from the same host thread, under the same current stream."""
    if means3D.dim() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    if not means3D.is_cuda:
        raise RuntimeError("diff_gaussian_rasterization (B200): tensors must be CUDA tensors; there is no CPU path")
    dev = means3D.device
    P, H, W = means3D.size(0), int(image_height), int(image_width)
    fopt = dict(dtype=torch.float32, device=dev)
    out_color = torch.empty((3, H, W), **fopt)
    out_acc = torch.empty((H, W), **fopt)
    out_depth = torch.empty((H, W), **fopt)
    radii = torch.empty((P,), dtype=torch.int32, device=dev)
    pixel_sizes = torch.empty((P,), **fopt)
    with torch.cuda.device(dev):
        hint = int(_lib.msgs_binning_hint_bytes(P, W, H)) if (P > 0 and not debug) else 0
    geom = _Buffer(dev, _geom_bytes(P) if P > 0 else 0)
    binning = _Buffer(dev, hint)
    img = _Buffer(dev, _image_bytes(W, H, P) if P > 0 else 0)
    M = sh.size(1) if sh.numel() != 0 else 0

    keep = [_cuda_f32(t) for t in (background, means3D, sh, colors, max_pixel_sizes, min_pixel_sizes, opacity,
                                   occ_multiplier, dc_delta, scales, rotations, cov3D_precomp, base_mask,
                                   viewmatrix, projmatrix, campos)]
    (bg_c, means_c, sh_c, col_c, maxp_c, minp_c, op_c, occm_c, dcd_c, sc_c, rot_c, cov_c, mask_c, view_c, proj_c,
     cam_c) = keep
    with torch.cuda.device(dev):
        handle = (_lib.msgs_forward_begin_raw if raw else _lib.msgs_forward_begin)(
            geom.cb, geom.ctx, binning.cb, binning.ctx, img.cb, img.ctx, P, int(degree), int(M),
            _ptr(bg_c), W, H,
            _ptr(means_c), _ptr(sh_c), _ptr(col_c), _ptr(maxp_c), _ptr(minp_c), _ptr(op_c), _ptr(occm_c), _ptr(dcd_c),
            _ptr(sc_c), float(scale_modifier), _ptr(rot_c), _ptr(cov_c), _ptr(mask_c),
            _ptr(view_c), _ptr(proj_c), _ptr(cam_c), float(tan_fovx), float(tan_fovy),
            int(bool(prefiltered)), int(bool(filter_small)), int(bool(filter_large)), float(fade_size),
            out_color.data_ptr(), out_acc.data_ptr(), out_depth.data_ptr(),
            _ptr(radii), _ptr(pixel_sizes), int(bool(debug)), _stream_ptr(dev))
    if handle < 0:
        for b in (geom, binning, img):
            b.release()
        _err(handle, "rasterize_gaussians")
    return PendingForward(handle, dev, (out_color, out_acc, out_depth, radii, pixel_sizes), (geom, binning, img), keep)


MAX_GROUP = 8


def rasterize_gaussians_group_begin(units, means3D, colors, max_pixel_sizes, min_pixel_sizes, opacity, occ_multiplier,
                                    dc_delta, scales, rotations, scale_modifier, cov3D_precomp, base_mask, viewmatrix,
                                    projmatrix, tan_fovx, tan_fovy, sh, degree, campos, prefiltered, filter_small,
                                    filter_large, fade_size, debug, raw=False):
    """Extension (msgs_forward_group_begin[_raw], include/msgs_raster.h): `rasterize_gaussians_begin` for a VIEW GROUP — up to 8
    units that share the camera, the Gaussians and the flags and differ only in the image size (the scales of one view of
    a multi-scale step).  `units` = [(background, image_height, image_width, stream)], stream a torch.cuda.Stream or None
    (= the current stream).  One preprocess launch and ONE depth sort serve all units (on the first unit's stream);
    counting, placement and blend run per unit on the unit's stream.  Returns one PendingForward per unit; each finish()
    returns what `rasterize_gaussians` returns for that unit, bit-identical to a separate call."""
    if means3D.dim() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    if not means3D.is_cuda:
        raise RuntimeError("diff_gaussian_rasterization (B200): tensors must be CUDA tensors; there is no CPU path")
    n = len(units)
    if not 1 <= n <= MAX_GROUP:
        raise RuntimeError(f"a view group holds 1..{MAX_GROUP} units")
    dev = means3D.device
    P = means3D.size(0)
    if P == 0:
        raise RuntimeError("rasterize_gaussians_group_begin needs at least one Gaussian")
    fopt = dict(dtype=torch.float32, device=dev)
    M = sh.size(1) if sh.numel() != 0 else 0
    keep = [_cuda_f32(t) for t in (means3D, sh, colors, max_pixel_sizes, min_pixel_sizes, opacity, occ_multiplier,
                                   dc_delta, scales, rotations, cov3D_precomp, base_mask, viewmatrix, projmatrix, campos)]
    (means_c, sh_c, col_c, maxp_c, minp_c, op_c, occm_c, dcd_c, sc_c, rot_c, cov_c, mask_c, view_c, proj_c, cam_c) = keep
    cur = torch.cuda.current_stream(dev)
    arr = (_GroupUnit * n)()
    handles = (_i * n)()
    per_unit = []
    for k, (bg, H, W, stream) in enumerate(units):
        H, W = int(H), int(W)
        st = stream if stream is not None else cur
        bg_c = _cuda_f32(bg)
        # everything a unit owns is allocated on the CALLER's stream: the caller orders the unit streams around the step
        # (msgs_sharding.UnitStreams.begin / end), so blocks freed after end() are reused stream-ordered; the state buffers'
        # sizes are known here, so the core gets them without calling back into Python (msgs_fixed_alloc)
        outs = (torch.empty((3, H, W), **fopt), torch.empty((H, W), **fopt), torch.empty((H, W), **fopt),
                torch.empty((P,), dtype=torch.int32, device=dev), torch.empty((P,), **fopt))
        hint = int(_lib.msgs_binning_hint_bytes(P, W, H)) if not debug else 0
        bufs = (_Buffer(dev, _geom_bytes(P)), _Buffer(dev, hint), _Buffer(dev, _image_bytes(W, H, P)))
        u = arr[k]
        u.width, u.height, u.background = W, H, bg_c.data_ptr()
        u.out_color, u.out_acc_pixel_size, u.out_depth = outs[0].data_ptr(), outs[1].data_ptr(), outs[2].data_ptr()
        u.radii, u.pixel_sizes = outs[3].data_ptr(), outs[4].data_ptr()
        u.geom_alloc, u.binning_alloc, u.img_alloc = bufs[0].cb, bufs[1].cb, bufs[2].cb
        u.geom_ctx, u.binning_ctx, u.img_ctx = bufs[0].ctx, bufs[1].ctx, bufs[2].ctx
        u.stream = st.cuda_stream
        per_unit.append((outs, bufs, st, bg_c))
    with torch.cuda.device(dev), torch.cuda.stream(per_unit[0][2]):
        rc = (_lib.msgs_forward_group_begin_raw if raw else _lib.msgs_forward_group_begin)(
            n, arr, handles, P, int(degree), int(M),
            _ptr(means_c), _ptr(sh_c), _ptr(col_c), _ptr(maxp_c), _ptr(minp_c), _ptr(op_c), _ptr(occm_c), _ptr(dcd_c),
            _ptr(sc_c), float(scale_modifier), _ptr(rot_c), _ptr(cov_c), _ptr(mask_c),
            _ptr(view_c), _ptr(proj_c), _ptr(cam_c), float(tan_fovx), float(tan_fovy),
            int(bool(prefiltered)), int(bool(filter_small)), int(bool(filter_large)), float(fade_size), int(bool(debug)))
    if rc < 0:
        for _, bufs, _, _ in per_unit:
            for b in bufs:
                b.release()
        _err(rc, "rasterize_gaussians (view group)")
    group_geom = per_unit[0][1][0]
    pend = []
    for k, (outs, bufs, st, bg_c) in enumerate(per_unit):
        # every unit keeps the group's depth order (unit 0's geometry buffer) alive until it is finished
        pend.append(PendingForward(int(handles[k]), dev, outs, bufs, keep + [bg_c, arr, group_geom.tensor]))
    return pend


def rasterize_gaussians(background, means3D, colors, max_pixel_sizes, min_pixel_sizes, opacity, occ_multiplier,
                        dc_delta, scales, rotations, scale_modifier, cov3D_precomp, base_mask, viewmatrix,
                        projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos, prefiltered,
                        filter_small, filter_large, fade_size, debug):
    """RasterizeGaussiansCUDA (rasterize_points.cu:35-139): returns
    (rendered, out_color, out_acc_pixel_size, out_depth, radii, pixel_sizes, geomBuffer, binningBuffer, imgBuffer)."""
    return rasterize_gaussians_begin(background, means3D, colors, max_pixel_sizes, min_pixel_sizes, opacity,
                                     occ_multiplier, dc_delta, scales, rotations, scale_modifier, cov3D_precomp,
                                     base_mask, viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width,
                                     sh, degree, campos, prefiltered, filter_small, filter_large, fade_size, debug).finish()


def rasterize_gaussians_backward(background, means3D, radii, colors, pixel_sizes, max_pixel_sizes, min_pixel_sizes,
                                 scales, rotations, scale_modifier, cov3D_precomp, base_mask, viewmatrix, projmatrix,
                                 tan_fovx, tan_fovy, dL_dout_color, sh, degree, campos, geomBuffer, R, binningBuffer,
                                 imageBuffer, filter_small, filter_large, fade_size, debug, out=None, skip=()):
    """RasterizeGaussiansBackwardCUDA (rasterize_points.cu:141-241): returns (dL_dmeans2D, dL_dcolors, dL_dopacity,
    dL_docc_multiplier, dL_ddc_delta, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations).

    `out` (extension, optional): {"mean3D", "sh", "opacity", "scale", "rot"} -> preallocated contiguous fp32 CUDA tensors
    to write those gradients into (e.g. views of one flat buffer that is all-reduced as a whole).
    `skip` (extension, optional): names among {"means2D", "colors", "occ_multiplier", "dc_delta", "cov3D"} whose gradient
    the caller does not need (autograd's needs_input_grad): not allocated, not written, returned as None.  dL_dconic,
    which the reference computes but never returns (rasterize_points.cu:240), is not materialised at all."""
    dev = means3D.device
    P = means3D.size(0)
    H, W = dL_dout_color.size(1), dL_dout_color.size(2)
    M = sh.size(1) if sh.numel() != 0 else 0
    fopt = dict(dtype=torch.float32, device=dev)
    # the core writes every element (zeros included): no torch.zeros fills (364 B per Gaussian in the reference)
    mk = torch.empty if P > 0 else torch.zeros
    dL_dmeans3D = mk((P, 3), **fopt)
    dL_dmeans2D = mk((P, 3), **fopt) if "means2D" not in skip else None
    dL_dcolors = mk((P, 3), **fopt) if "colors" not in skip else None
    dL_dconic = None
    dL_dopacity = mk((P, 1), **fopt)
    dL_docc = mk((P, 4, 1), **fopt) if "occ_multiplier" not in skip else None
    dL_ddc = mk((P, 12, 1), **fopt) if "dc_delta" not in skip else None
    dL_dcov3D = mk((P, 6), **fopt) if "cov3D" not in skip else None
    dL_dsh = mk((P, M, 3), **fopt)
    dL_dscales = mk((P, 3), **fopt)
    dL_drot = mk((P, 4), **fopt)
    if out:
        def pick(key, default):
            o = out.get(key)
            if o is None:
                return default
            if not (o.is_cuda and o.dtype == torch.float32 and o.is_contiguous() and o.numel() == default.numel()):
                raise RuntimeError(f"out['{key}'] must be a contiguous fp32 CUDA tensor of {default.numel()} elements")
            return o.view(default.shape)
        dL_dmeans3D, dL_dsh, dL_dopacity = pick("mean3D", dL_dmeans3D), pick("sh", dL_dsh), pick("opacity", dL_dopacity)
        dL_dscales, dL_drot = pick("scale", dL_dscales), pick("rot", dL_drot)
    if P != 0:
        keep = [t.contiguous() for t in (background, means3D, sh, colors, pixel_sizes, max_pixel_sizes,
                                         min_pixel_sizes, scales, rotations, cov3D_precomp, base_mask, viewmatrix,
                                         projmatrix, campos, radii, dL_dout_color)]
        (bg_c, means_c, sh_c, col_c, px_c, maxp_c, minp_c, sc_c, rot_c, cov_c, mask_c, view_c, proj_c, cam_c,
         radii_c, dpix_c) = keep
        with torch.cuda.device(dev):
            rc = _lib.msgs_backward(
                P, int(degree), int(M), int(R), _ptr(bg_c), W, H,
                _ptr(means_c), _ptr(sh_c), _ptr(col_c), _ptr(px_c), _ptr(maxp_c), _ptr(minp_c),
                _ptr(sc_c), float(scale_modifier), _ptr(rot_c), _ptr(cov_c), _ptr(mask_c),
                _ptr(view_c), _ptr(proj_c), _ptr(cam_c), float(tan_fovx), float(tan_fovy), _ptr(radii_c),
                _ptr(geomBuffer), _ptr(binningBuffer), _ptr(imageBuffer),
                int(bool(filter_small)), int(bool(filter_large)), float(fade_size), _ptr(dpix_c),
                _ptr(dL_dmeans2D), _ptr(dL_dconic), _ptr(dL_dopacity), _ptr(dL_docc), _ptr(dL_ddc), _ptr(dL_dcolors),
                _ptr(dL_dmeans3D), _ptr(dL_dcov3D), _ptr(dL_dsh), _ptr(dL_dscales), _ptr(dL_drot),
                int(bool(debug)), _stream_ptr(dev))
        if rc < 0:
            _err(rc, "rasterize_gaussians_backward")
    return (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_docc, dL_ddc, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales,
            dL_drot)


def rasterize_gaussians_backward_accumulate(background, means3D, radii, colors, pixel_sizes, max_pixel_sizes,
                                            min_pixel_sizes, scales, rotations, scale_modifier, cov3D_precomp, base_mask,
                                            viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color, sh, degree, campos,
                                            geomBuffer, R, binningBuffer, imageBuffer, filter_small, filter_large,
                                            fade_size, debug, accum, want_means2D=True, part=3,
                                            first=0, count=-1, means2D_out=None, opacity_raw=None):
    """Extension (msgs_backward_accumulate, include/msgs_raster.h): same inputs as rasterize_gaussians_backward, but
    the training gradients of this (view, scale) unit are ADDED into the caller's tensors
    accum = {"mean3D": (P,3), "sh": (P,M,3), "opacity": (P,1), "scale": (P,3), "rot": (P,4)} (fp32, contiguous, zeroed
    by the caller once per step), touching only the Gaussians the unit rendered.  Returns dL_dmeans2D (P,3) of this
    unit (the per-view densification statistic), or None when want_means2D is False.

    opacity_raw (msgs_backward_raw): `scales` / `rotations` are the RAW parameters and opacity_raw the raw opacities (P,1);
    the gradients added are those of the raw parameters (the forward must have been a raw forward too)."""
    dev = means3D.device
    P = means3D.size(0)
    H, W = dL_dout_color.size(1), dL_dout_color.size(2)
    M = sh.size(1) if sh.numel() != 0 else 0
    if P == 0:
        return torch.zeros((0, 3), dtype=torch.float32, device=dev) if want_means2D else None
    for k in ("mean3D", "sh", "opacity", "scale", "rot"):
        a = accum[k]
        if not (a.is_cuda and a.dtype == torch.float32 and a.is_contiguous() and a.size(0) == P):
            raise RuntimeError(f"accum['{k}'] must be a contiguous fp32 CUDA tensor with {P} rows")
    aptr = {k: accum[k].data_ptr() for k in ("mean3D", "sh", "opacity", "scale", "rot")}
    fn = _lib.msgs_backward_accumulate
    partial = (part != 3 or first != 0 or count >= 0)
    raw_args = ()
    if opacity_raw is not None:
        if first != 0 or count >= 0:
            raise RuntimeError("the raw-parameter backward has no Gaussian slices")
        oraw_c = opacity_raw.contiguous()
        fn, partial, raw_args = _lib.msgs_backward_raw, False, (oraw_c.data_ptr(), 1, int(part))
    elif partial:
        # msgs_backward_accumulate_part: part bit 0 = blend backward, bit 1 = preprocess backward over the Gaussians
        # [first, first + count); the caller passes ONE means2D_out tensor to all slices of a unit
        fn = _lib.msgs_backward_accumulate_part
    if means2D_out is not None:
        dL_dmeans2D = means2D_out
    else:
        dL_dmeans2D = torch.empty((P, 3), dtype=torch.float32, device=dev) if (want_means2D and (part & 2)) else None
    keep = [t.contiguous() for t in (background, means3D, sh, colors, pixel_sizes, max_pixel_sizes, min_pixel_sizes, scales,
                                     rotations, cov3D_precomp, base_mask, viewmatrix, projmatrix, campos, radii,
                                     dL_dout_color)]
    (bg_c, means_c, sh_c, col_c, px_c, maxp_c, minp_c, sc_c, rot_c, cov_c, mask_c, view_c, proj_c, cam_c, radii_c,
     dpix_c) = keep
    with torch.cuda.device(dev):
        rc = fn(
            *raw_args, P, int(degree), int(M), int(R), _ptr(bg_c), W, H,
            _ptr(means_c), _ptr(sh_c), _ptr(col_c), _ptr(px_c), _ptr(maxp_c), _ptr(minp_c),
            _ptr(sc_c), float(scale_modifier), _ptr(rot_c), _ptr(cov_c), _ptr(mask_c),
            _ptr(view_c), _ptr(proj_c), _ptr(cam_c), float(tan_fovx), float(tan_fovy), _ptr(radii_c),
            _ptr(geomBuffer), _ptr(binningBuffer), _ptr(imageBuffer),
            int(bool(filter_small)), int(bool(filter_large)), float(fade_size), _ptr(dpix_c),
            _ptr(dL_dmeans2D), None, aptr["opacity"], None, None, None,
            aptr["mean3D"], None, aptr["sh"], aptr["scale"], aptr["rot"],
            int(bool(debug)), _stream_ptr(dev), *((int(part), int(first), int(count)) if partial else ()))
    if rc < 0:
        _err(rc, "rasterize_gaussians_backward_accumulate")
    return dL_dmeans2D


def rasterize_gaussians_group_backward_accumulate(units, means3D, max_pixel_sizes, min_pixel_sizes, scales, rotations,
                                                  scale_modifier, cov3D_precomp, base_mask, viewmatrix, projmatrix, tan_fovx,
                                                  tan_fovy, sh, degree, campos, filter_small, filter_large, debug, accum,
                                                  want_means2D=True, opacity_raw=None):
    """Extension (msgs_backward_group_accumulate): the preprocess backward of a view group in ONE pass over the Gaussians,
    on the current stream.  `units` = [(image_height, image_width, radii, pixel_sizes, geomBuffer)] of the group's units,
    whose blend backward (rasterize_gaussians_backward_accumulate(..., part=1)) has already been enqueued and which the
    current stream has been made to wait for.  Adds the group's training gradients into `accum` (as
    rasterize_gaussians_backward_accumulate does per unit) and returns each unit's dL_dmeans2D (P,3) (None entries when
    want_means2D is False)."""
    dev = means3D.device
    P = means3D.size(0)
    M = sh.size(1) if sh.numel() != 0 else 0
    n = len(units)
    if not 1 <= n <= MAX_GROUP:
        raise RuntimeError(f"a view group holds 1..{MAX_GROUP} units")
    for k in ("mean3D", "sh", "opacity", "scale", "rot"):
        a = accum[k]
        if not (a.is_cuda and a.dtype == torch.float32 and a.is_contiguous() and a.size(0) == P):
            raise RuntimeError(f"accum['{k}'] must be a contiguous fp32 CUDA tensor with {P} rows")
    if P == 0:
        return [torch.zeros((0, 3), dtype=torch.float32, device=dev) if want_means2D else None for _ in units]
    keep = [t.contiguous() for t in (means3D, sh, max_pixel_sizes, min_pixel_sizes, scales, rotations, cov3D_precomp,
                                     base_mask, viewmatrix, projmatrix, campos)]
    (means_c, sh_c, maxp_c, minp_c, sc_c, rot_c, cov_c, mask_c, view_c, proj_c, cam_c) = keep
    arr = (_GroupBwdUnit * n)()
    g2ds = []
    for k, (H, W, radii, px, gb) in enumerate(units):
        radii_c, px_c = radii.contiguous(), px.contiguous()
        keep += [radii_c, px_c, gb]
        g2d = torch.empty((P, 3), dtype=torch.float32, device=dev) if want_means2D else None
        g2ds.append(g2d)
        u = arr[k]
        u.width, u.height, u.radii, u.pixel_sizes = int(W), int(H), radii_c.data_ptr(), px_c.data_ptr()
        u.geom_buffer, u.dL_dmean2D = gb.data_ptr(), _ptr(g2d)
    raw_args, fn = (), _lib.msgs_backward_group_accumulate
    if opacity_raw is not None:          # raw-parameter mode: gradients of _scaling / _rotation / _opacity (msgs_*_raw)
        oraw_c = opacity_raw.contiguous()
        keep.append(oraw_c)
        raw_args, fn = (oraw_c.data_ptr(),), _lib.msgs_backward_group_accumulate_raw
    with torch.cuda.device(dev):
        rc = fn(
            *raw_args, n, arr, P, int(degree), int(M), _ptr(means_c), _ptr(sh_c), _ptr(maxp_c), _ptr(minp_c),
            _ptr(sc_c), float(scale_modifier), _ptr(rot_c), _ptr(cov_c), _ptr(mask_c),
            _ptr(view_c), _ptr(proj_c), _ptr(cam_c), float(tan_fovx), float(tan_fovy),
            int(bool(filter_small)), int(bool(filter_large)),
            accum["opacity"].data_ptr(), accum["mean3D"].data_ptr(), accum["sh"].data_ptr(), accum["scale"].data_ptr(),
            accum["rot"].data_ptr(), int(bool(debug)), _stream_ptr(dev))
    if rc < 0:
        _err(rc, "rasterize_gaussians_group_backward_accumulate")
    return g2ds


def rasterize_gaussians_group_backward_full(units, means3D, max_pixel_sizes, min_pixel_sizes, scales, rotations,
                                            scale_modifier, cov3D_precomp, base_mask, viewmatrix, projmatrix, tan_fovx, tan_fovy,
                                            sh, degree, campos, filter_small, filter_large, debug, accum, want_means2D=True,
                                            opacity_raw=None, stream=None, slices=None, slice_hook=None):
    """Extension (msgs_backward_group_full): the whole backward of a view group in ONE library call.  `units` =
    [(background, image_height, image_width, radii, pixel_sizes, geomBuffer, R, binningBuffer, imageBuffer, dL_dout_color,
    stream)]; every unit's blend backward goes out on its stream (coarsest first) behind what the CURRENT stream holds, then
    the group's preprocess backward runs on the current stream behind all of them (`stream`: use that torch stream as the
    group's instead of the current one; tensors stay on the caller's stream, which must order the streams around the call as
    msgs_sharding.UnitStreams.begin / end do).  Adds the group's training gradients into `accum`; returns each unit's
    dL_dmeans2D (P,3).

    `slices` = Gaussian boundaries [0, b1, ..., P] (multiples of 128): the preprocess backward is issued slice by slice
    (msgs_backward_group_blend, then msgs_backward_group_accumulate_part per slice) and `slice_hook(k, n_slices, group_stream)`
    is called right after slice k has been enqueued — the place to start exchanging that slice's gradients."""
    dev = means3D.device
    P = means3D.size(0)
    M = sh.size(1) if sh.numel() != 0 else 0
    n = len(units)
    if not 1 <= n <= MAX_GROUP:
        raise RuntimeError(f"a view group holds 1..{MAX_GROUP} units")
    for k in ("mean3D", "sh", "opacity", "scale", "rot"):
        a = accum[k]
        if not (a.is_cuda and a.dtype == torch.float32 and a.is_contiguous() and a.size(0) == P):
            raise RuntimeError(f"accum['{k}'] must be a contiguous fp32 CUDA tensor with {P} rows")
    if P == 0:
        return [torch.zeros((0, 3), dtype=torch.float32, device=dev) if want_means2D else None for _ in units]
    keep = [t.contiguous() for t in (means3D, sh, max_pixel_sizes, min_pixel_sizes, scales, rotations, cov3D_precomp,
                                     base_mask, viewmatrix, projmatrix, campos)]
    (means_c, sh_c, maxp_c, minp_c, sc_c, rot_c, cov_c, mask_c, view_c, proj_c, cam_c) = keep
    arr, bl = (_GroupBwdUnit * n)(), (_GroupBwdBlend * n)()
    g2ds = []
    g2d_all = torch.empty((n, P, 3), dtype=torch.float32, device=dev) if want_means2D else None     # one allocation
    for k, (bg, H, W, radii, px, gb, R, bb, ib, dpix, unit_stream) in enumerate(units):
        radii_c, px_c, bg_c, dpix_c = radii.contiguous(), px.contiguous(), bg.contiguous(), dpix.contiguous()
        if dpix_c.size(1) != H or dpix_c.size(2) != W:
            raise RuntimeError("dL_dout_color does not have the unit's image size")
        keep += [radii_c, px_c, bg_c, dpix_c, gb, bb, ib]
        g2d = g2d_all[k] if want_means2D else None
        g2ds.append(g2d)
        u = arr[k]
        u.width, u.height, u.radii, u.pixel_sizes = int(W), int(H), radii_c.data_ptr(), px_c.data_ptr()
        u.geom_buffer, u.dL_dmean2D = gb.data_ptr(), _ptr(g2d)
        b = bl[k]
        b.background, b.binning_buffer, b.image_buffer, b.R = bg_c.data_ptr(), _ptr(bb), ib.data_ptr(), int(R)
        b.dL_dpix, b.stream = dpix_c.data_ptr(), unit_stream.cuda_stream
    oraw_c = opacity_raw.contiguous() if opacity_raw is not None else None
    gstream = ctypes.c_void_p(stream.cuda_stream) if stream is not None else _stream_ptr(dev)
    tail = (n, arr, P, int(degree), int(M), _ptr(means_c), _ptr(sh_c), _ptr(maxp_c), _ptr(minp_c),
            _ptr(sc_c), float(scale_modifier), _ptr(rot_c), _ptr(cov_c), _ptr(mask_c),
            _ptr(view_c), _ptr(proj_c), _ptr(cam_c), float(tan_fovx), float(tan_fovy),
            int(bool(filter_small)), int(bool(filter_large)),
            accum["opacity"].data_ptr(), accum["mean3D"].data_ptr(), accum["sh"].data_ptr(), accum["scale"].data_ptr(),
            accum["rot"].data_ptr(), int(bool(debug)), gstream)
    with torch.cuda.device(dev):
        if slices is None or len(slices) <= 2:
            rc = _lib.msgs_backward_group_full(_ptr(oraw_c), bl, *tail)
            if rc >= 0 and slice_hook is not None:
                slice_hook(0, 1, stream if stream is not None else torch.cuda.current_stream(dev))
        else:
            rc = _lib.msgs_backward_group_blend(_ptr(oraw_c), bl, *tail)
            ns = len(slices) - 1
            for k in range(ns):
                if rc < 0:
                    break
                last = k == ns - 1
                rc = _lib.msgs_backward_group_accumulate_part(_ptr(oraw_c), int(slices[k]), -1 if last else int(slices[k + 1] - slices[k]),
                                                              *tail)
                if rc >= 0 and slice_hook is not None:
                    slice_hook(k, ns, stream if stream is not None else torch.cuda.current_stream(dev))
    if rc < 0:
        _err(rc, "rasterize_gaussians_group_backward_full")
    return g2ds


def mark_visible(means3D, viewmatrix, projmatrix):
    """markVisible (rasterize_points.cu:243-262)."""
    if not means3D.is_cuda:
        raise RuntimeError("diff_gaussian_rasterization (B200): tensors must be CUDA tensors; there is no CPU path")
    P = means3D.size(0)
    present = torch.zeros((P,), dtype=torch.bool, device=means3D.device)
    if P != 0:
        m, v, p = means3D.contiguous(), viewmatrix.contiguous(), projmatrix.contiguous()
        with torch.cuda.device(means3D.device):
            rc = _lib.msgs_mark_visible(P, m.data_ptr(), v.data_ptr(), p.data_ptr(), present.data_ptr(),
                                        _stream_ptr(means3D.device))
        if rc < 0:
            _err(rc, "mark_visible")
    return present


def speculation_stats(reset=False):
    """(forwards finished, speculative binning buffer large enough, placement + blend launched again) of this host thread."""
    out = (ctypes.c_longlong * 3)()
    _lib.msgs_speculation_stats(out, int(bool(reset)))
    return int(out[0]), int(out[1]), int(out[2])
