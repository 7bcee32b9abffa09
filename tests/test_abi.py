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
    # per instance: the Gaussian index (4 B) + room for one 8-byte level-1 entry of the two-level placement (only
    # ~ R / 13 of them are ever written at C2; the reference keeps 24 B per instance + sort temp)
    assert 12 * 1_000_000 <= lib.msgs_binning_bytes(1_000_000) <= 12 * 1_000_000 + 4096
    # per pixel final_T + n_contrib, plus the (chunk groups x tiles) count matrix and the (super-block x list chunk)
    # offsets of the binning stage
    assert lib.msgs_image_bytes(1237, 822, 1_200_000) >= 8 * 1237 * 822 + 4 * 16 * 4056 + 4 * 70 * 1172
    # the sort's scratch: histograms + tickets + one look-back row per 8192-pair tile + one per group of 16 tiles; what
    # msgs_sort_pairs_static can sort (MSGS_SORT_STATIC_MAX_N pairs) has to fit the 540 KiB static device array it owns
    lib.msgs_sort_temp_bytes.restype = ctypes.c_size_t
    lib.msgs_sort_temp_bytes.argtypes = [ctypes.c_int]
    hdr = open(os.path.join(ROOT, "include", "msgs_raster.h")).read()
    max_n = eval(re.search(r"#define MSGS_SORT_STATIC_MAX_N\s+\(([^)]*)\)", hdr).group(1))
    assert max_n == 1 << 21
    assert 0 < lib.msgs_sort_temp_bytes(1) <= lib.msgs_sort_temp_bytes(max_n) <= 540 * 1024
    assert lib.msgs_sort_temp_bytes(1 << 27) < (1 << 27) // 8192 * 1024 * 1.1 + (1 << 16)


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


def test_group_structs_match_the_header(tmp_path):
    """The ctypes mirrors of msgs_group_unit / msgs_group_bwd_unit (_C.py) have the size and field offsets the C compiler
    gives the structs declared in include/msgs_raster.h (checked with gcc: the header must also compile as plain C)."""
    import subprocess
    import sys
    sys.path.insert(0, os.path.join(ROOT, "ms-gs_b200"))
    from diff_gaussian_rasterization import _C
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "msgs_raster.h"\n'
                   'int main(void) { printf("%zu %zu %zu %zu %zu %zu\\n", sizeof(msgs_group_unit), offsetof(msgs_group_unit, background),'
                   ' offsetof(msgs_group_unit, stream), sizeof(msgs_group_bwd_unit), offsetof(msgs_group_bwd_unit, geom_buffer),'
                   ' offsetof(msgs_group_bwd_unit, dL_dmean2D)); return 0; }\n')
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-std=c99", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    got = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    U, B = _C._GroupUnit, _C._GroupBwdUnit
    want = [ctypes.sizeof(U), U.background.offset, U.stream.offset, ctypes.sizeof(B), B.geom_buffer.offset, B.dL_dmean2D.offset]
    assert got == want


def test_view_groups_host_logic():
    """_view_groups (units API): consecutive units that share the camera (declared through view_ids, or the same tensor
    storage) and the flags form a group of at most 8; anything else starts a new group."""
    import sys
    import torch
    sys.path.insert(0, os.path.join(ROOT, "ms-gs_b200"))
    import diff_gaussian_rasterization as dgr
    v, p, c = torch.eye(4), torch.eye(4), torch.zeros(3)
    v2 = torch.eye(4)

    def rs(H, W, view=v, fs=True, deg=3):
        return dgr.GaussianRasterizationSettings(H, W, 0.5, 0.4, torch.zeros(3), 1.0, view, p, deg, c, False, False, fs, True, 1.0)

    s = [rs(100, 200), rs(50, 100), rs(25, 50), rs(100, 200, view=v2), rs(50, 100, view=v2), rs(25, 50, view=v2, fs=False)]
    assert dgr._view_groups(s) == [[0, 1, 2], [3, 4], [5]]
    assert dgr._view_groups(s, view_ids=[0, 0, 1, 1, 1, 1]) == [[0, 1], [2, 3, 4], [5]]
    assert dgr._view_groups([rs(10, 10)] * 11) == [list(range(8)), [8, 9, 10]]
    assert dgr._view_groups([rs(10, 10), rs(10, 10, deg=2)]) == [[0], [1]]
    os.environ["MSGS_VIEW_GROUPS"] = "0"
    try:
        assert dgr._view_groups(s) == [[i] for i in range(6)]
    finally:
        del os.environ["MSGS_VIEW_GROUPS"]


def _prototypes(header):
    """{name: [kind per parameter]} of the `msgs_*` prototypes in an include/ header; kind = ptr | int | float | size."""
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = re.sub(r"//[^\n]*", "", src)
    src = re.sub(r"typedef\s+struct\s+\w+\s*\{.*?\}\s*\w+\s*;", "", src, flags=re.S)
    src = re.sub(r"typedef[^;]*;", "", src)
    protos = {}
    for name, args in re.findall(r"\b(msgs_[a-z_0-9]+)\s*\(([^()]*)\)\s*;", src):
        kinds = []
        for a in [x.strip() for x in args.split(",")]:
            if a in ("void", ""):
                continue
            if "*" in a or "[" in a or a.startswith("msgs_alloc_fn"):
                kinds.append("ptr")
            elif re.match(r"(const\s+)?float\b", a):
                kinds.append("float")
            elif re.match(r"(const\s+)?size_t\b", a):
                kinds.append("size")
            elif re.match(r"(const\s+)?(int|int32_t)\b", a):
                kinds.append("int")
            else:
                raise AssertionError(f"{header}: {name}: parameter '{a}' not understood")
        protos[name] = kinds
    return protos


def _ctypes_kind(t):
    if t in (ctypes.c_void_p, ctypes.c_char_p) or hasattr(t, "contents") or hasattr(t, "_flags_") and hasattr(t, "_argtypes_"):
        return "ptr"
    if isinstance(t, type) and (issubclass(t, ctypes._Pointer) or issubclass(t, ctypes._CFuncPtr)):
        return "ptr"
    return {ctypes.c_int: "int", ctypes.c_float: "float", ctypes.c_size_t: "size", ctypes.c_int64: "i64",
            ctypes.c_uint64: "size"}.get(t, str(t))


def test_ctypes_signatures_match_the_header(built):
    """Every binding the Python side declares (argtypes) has the parameter list of the header prototype: same count,
    pointer / int / float / size_t in the same positions — a drifted prototype would otherwise corrupt arguments silently."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "ms-gs_b200"))
    try:
        import diff_gaussian_rasterization._C as C
        import msgs_ply
        import msgs_sharding
    finally:
        sys.path.pop(0)
    checked = 0
    for header, lib in (("msgs_raster.h", C._lib), ("msgs_ply.h", msgs_ply._lib)):
        for name, kinds in _prototypes(header).items():
            fn = getattr(lib, name)
            if fn.argtypes is None:
                # bound elsewhere: the exchange kernel lazily by msgs_sharding (below), the profiling hooks by bench.py
                assert not kinds or name == "msgs_allreduce_multimem" or name.startswith("msgs_profile_"), \
                    f"{name}: takes {len(kinds)} arguments but the binding declares no argtypes"
                continue
            got = [_ctypes_kind(t) for t in fn.argtypes]
            assert got == kinds, f"{name}: header {kinds} != binding {got}"
            checked += 1
    assert checked >= 15
    # the exchange kernel is bound lazily by msgs_sharding (it needs a symmetric buffer to be useful)
    lib = msgs_sharding._raster_lib()
    got = [_ctypes_kind(t) for t in lib.msgs_allreduce_multimem.argtypes]
    assert got == _prototypes("msgs_raster.h")["msgs_allreduce_multimem"]
