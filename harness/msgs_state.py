"""Decode the private state buffers of libmsgs_raster.so (geometry / image / binning byte tensors returned by
`_C.rasterize_gaussians`) into named torch views, through the library's own layout introspection entry points
(include/msgs_raster.h: msgs_geom_layout / msgs_image_layout / msgs_binning_layout).  Parity tests only."""
import ctypes
import os
import sys

import torch

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG_DIR = os.path.join(_ROOT, "ms-gs_b200")


def import_product():
    """Import the drop-in package from ms-gs_b200/ (directory name is not a Python identifier)."""
    if PKG_DIR not in sys.path:
        sys.path.insert(0, PKG_DIR)
    import diff_gaussian_rasterization as dgr
    return dgr


def _view(ptr, count, dtype, owner):
    off = ptr - owner.data_ptr()
    nbytes = count * torch.empty(0, dtype=dtype).element_size()
    return owner[off:off + nbytes].view(dtype)


def decode(geom_buf, binning_buf, img_buf, P, R, W, H):
    dgr = import_product()
    lib = dgr._C.lib
    T = ((W + 15) // 16) * ((H + 15) // 16)
    o8 = (ctypes.c_uint64 * 8)()
    lib.msgs_geom_layout(geom_buf.data_ptr(), P, o8)
    rec = _view(o8[0], 16 * P, torch.float32, geom_buf).view(P, 4, 4)
    meta = _view(o8[7], 16, torch.int32, geom_buf)
    s = dict(rec=rec, means2D=rec[:, 0, 0:2], depths=rec[:, 0, 2], blend_px=rec[:, 0, 3],
             conic_opacity=rec[:, 1, :], rgb=rec[:, 2, 0:3],
             cov3D=_view(o8[1], 6 * P, torch.float32, geom_buf).view(P, 6),
             clamped_bits=_view(o8[2], P, torch.uint8, geom_buf),
             tiles_touched=_view(o8[3], P, torch.int32, geom_buf),
             rect=_view(o8[4], 4 * P, torch.int16, geom_buf).view(P, 4),
             gacc=_view(o8[5], 12 * P, torch.float32, geom_buf).view(P, 12), meta=meta)
    o4 = (ctypes.c_uint64 * 4)()
    lib.msgs_image_layout(img_buf.data_ptr(), W, H, P, o4)
    s.update(final_T=_view(o4[0], W * H, torch.float32, img_buf), n_contrib=_view(o4[1], W * H, torch.int32, img_buf),
             ranges=_view(o4[2], 2 * T, torch.int32, img_buf).view(T, 2),
             tile_max_contrib=_view(o4[3], T, torch.int32, img_buf))
    if R > 0:
        lib.msgs_binning_layout(binning_buf.data_ptr(), R, o4)
        point_list = _view(o4[0], R, torch.int32, binning_buf)
        # The library never materialises the reference's 64-bit keys (tile << 32 | float_bits(depth)); the sorted
        # key sequence is reconstructed from its own outputs: the tile of list position i comes from `ranges`,
        # the depth bits from the splat record of point_list[i].
        r = s["ranges"].long()
        counts = (r[:, 1] - r[:, 0])
        tile_of = torch.repeat_interleave(torch.arange(T, device=r.device), counts)
        starts_ok = bool((counts.sum() == R).item())
        dbits = s["depths"].contiguous().view(torch.int32)[point_list.long()].long() & 0xFFFFFFFF
        keys = (tile_of << 32) | dbits if starts_ok else torch.zeros(R, dtype=torch.int64, device=r.device)
        s.update(point_list=point_list, keys_sorted=keys, depth_order=_view(o8[6], P, torch.int32, geom_buf))
    return s
