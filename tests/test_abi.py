"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports every symbol include/msgs_raster.h
declares, and the Python package mirrors the reference's public surface.  No compute calls (no GPU here)."""
import ctypes
import inspect
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "msgs_raster.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(msgs_[a-z_]+)\s*\(", src)) - {"msgs_alloc_fn"})


def test_header_declares_expected_entry_points():
    syms = _declared_symbols()
    for s in ("msgs_forward", "msgs_backward", "msgs_mark_visible", "msgs_sort_pairs", "msgs_geom_bytes"):
        assert s in syms


def test_library_exports_every_declared_symbol(built):
    lib = ctypes.CDLL(os.path.join(ROOT, "ms-gs_b200", "libmsgs_raster.so"))
    for s in _declared_symbols():
        assert hasattr(lib, s), f"libmsgs_raster.so does not export {s}"
    lib.msgs_abi_version.restype = ctypes.c_int
    assert lib.msgs_abi_version() == 1


def test_buffer_size_queries(built):
    lib = ctypes.CDLL(os.path.join(ROOT, "ms-gs_b200", "libmsgs_raster.so"))
    for n in ("msgs_geom_bytes", "msgs_binning_bytes"):
        getattr(lib, n).restype = ctypes.c_size_t
        getattr(lib, n).argtypes = [ctypes.c_int]
    lib.msgs_image_bytes.restype = ctypes.c_size_t
    lib.msgs_image_bytes.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int]
    assert lib.msgs_geom_bytes(0) > 0
    # per Gaussian: 64 B record + 24 B cov3D + 1 B clamp + 4 B tiles + 2 x 8 B rect + 16 B depth-sort ping-pong
    # + 48 B accumulators (+ look-back words of the depth sort)
    per = (lib.msgs_geom_bytes(1_000_000) - lib.msgs_geom_bytes(0)) / 1e6
    assert 172 <= per <= 176
    # per instance: the Gaussian index only (the reference keeps 24 B per instance + sort temp)
    assert 4 * 1_000_000 <= lib.msgs_binning_bytes(1_000_000) <= 4 * 1_000_000 + 4096
    # per pixel final_T + n_contrib, plus the (chunks x tiles) count matrix of the binning stage
    assert lib.msgs_image_bytes(1237, 822, 1_200_000) >= 8 * 1237 * 822 + 4 * 200 * 4056


def test_python_surface_matches_reference(built):
    import diff_gaussian_rasterization as dgr
    assert dgr.GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
        "sh_degree", "campos", "prefiltered", "debug", "filter_small", "filter_large", "fade_size")
    assert list(inspect.signature(dgr.rasterize_gaussians).parameters) == [
        "means3D", "means2D", "sh", "colors_precomp", "max_pixel_sizes", "min_pixel_sizes", "opacities",
        "occ_multiplier", "dc_delta", "scales", "rotations", "cov3Ds_precomp", "base_mask", "raster_settings"]
    assert list(inspect.signature(dgr.GaussianRasterizer.forward).parameters) == [
        "self", "means3D", "means2D", "opacities", "occ_multiplier", "dc_delta", "shs", "colors_precomp", "base_mask",
        "scales", "rotations", "cov3D_precomp", "max_pixel_sizes", "min_pixel_sizes"]
    assert hasattr(dgr.GaussianRasterizer, "markVisible")
    for fn in ("rasterize_gaussians", "rasterize_gaussians_backward", "mark_visible"):
        assert callable(getattr(dgr._C, fn))


def test_argument_validation_messages(built):
    import torch
    import diff_gaussian_rasterization as dgr
    z = torch.zeros(4, 3)
    r = dgr.GaussianRasterizer(None)
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(z, z, z[:, :1])
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(z, z, z[:, :1], shs=z, colors_precomp=z)
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(z, z, z[:, :1], shs=z)
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(z, z, z[:, :1], shs=z, scales=z, rotations=z, cov3D_precomp=z)


def test_cpu_tensors_are_rejected_loudly(built):
    """No CPU fallback: a CPU tensor must raise, not silently compute somewhere else."""
    import torch
    import diff_gaussian_rasterization as dgr
    with pytest.raises(RuntimeError, match="no CPU path"):
        dgr._C.mark_visible(torch.zeros(4, 3), torch.eye(4), torch.eye(4))
    with pytest.raises(RuntimeError, match="must have dimensions"):
        dgr._C.rasterize_gaussians(*([torch.zeros(3)] + [torch.zeros(5)] + [None] * 25))
