"""CPU restatement (numpy) of the reference's model-file reader — TEST INFRASTRUCTURE ONLY (imported by tests/ and
tools/ply_bench.py's baseline leg, never by the product).

Follows GaussianModel.load_ply (/root/reference/scene/gaussian_model.py:358-417) attribute by attribute, with the same
name-sorted column gathering, float64 temporaries, reshapes and transposes, and get_features (:127-153: cat of
features_dc and features_rest along dim 1).  The reference reads the file with the third-party `plyfile` package
(environment.yml; absent from this image and from /root/reference): its binary reader is `numpy.fromfile` of a structured
dtype built from the header's property list (plyfile.py, PlyElement._read_bin) — restated in `_read_vertex_table`.
The reference ships no fixture or test for this path: parity unpinned; the pin used here is the writer
`harness/ply_io.write_ms_ply`, which restates GaussianModel.save_ply (:314-344), and a round trip through both."""
import numpy as np

_TYPES = {"float": "<f4", "float32": "<f4", "double": "<f8", "float64": "<f8", "uchar": "u1", "uint8": "u1",
          "char": "i1", "int8": "i1", "ushort": "<u2", "uint16": "<u2", "short": "<i2", "int16": "<i2",
          "uint": "<u4", "uint32": "<u4", "int": "<i4", "int32": "<i4"}


def _read_vertex_table(path):
    """header -> numpy structured dtype -> one fromfile of the vertex rows (what plyfile does for binary files)"""
    with open(path, "rb") as f:
        if f.readline().strip() != b"ply":
            raise ValueError("not a PLY file")
        fields, count, in_vertex, fmt = [], None, False, None
        while True:
            line = f.readline()
            if not line:
                raise ValueError("no end_header")
            tok = line.decode("ascii").split()
            if not tok:
                continue
            if tok[0] == "end_header":
                break
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                in_vertex = tok[1] == "vertex"
                if in_vertex:
                    count = int(tok[2])
            elif tok[0] == "property" and in_vertex:
                fields.append((tok[2], _TYPES[tok[1]]))
        if fmt != "binary_little_endian":
            raise ValueError("only binary_little_endian")
        return np.fromfile(f, dtype=np.dtype(fields), count=count)


def load_ply(path, n_lvl_occ, n_lvl_dc, max_sh_degree=3):
    """gaussian_model.py:358-417 — returns the tensors load_ply assigns, as fp32 numpy arrays in the reference's shapes."""
    v = _read_vertex_table(path)
    names = v.dtype.names
    P = v.shape[0]
    xyz = np.stack((np.asarray(v["x"]), np.asarray(v["y"]), np.asarray(v["z"])), axis=1)             # :361-363
    opacities = np.asarray(v["opacity"])[..., np.newaxis]                                            # :364
    occ = np.zeros((P, n_lvl_occ, 1))
    for i in range(n_lvl_occ):
        occ[:, i, 0] = np.asarray(v[f"occ_multiplier_{i}"])                                          # :365-367
    dcd = np.zeros((P, n_lvl_dc * 3, 1))
    for i in range(n_lvl_dc):
        for j in range(3):
            dcd[:, i * 3 + j, 0] = np.asarray(v[f"dc_delta_{i}_{j}"])                                # :369-372
    fdc = np.zeros((P, 3, 1))
    for c in range(3):
        fdc[:, c, 0] = np.asarray(v[f"f_dc_{c}"])                                                    # :374-377
    extra = sorted([n for n in names if n.startswith("f_rest_")], key=lambda x: int(x.split("_")[-1]))
    assert len(extra) == 3 * (max_sh_degree + 1) ** 2 - 3                                             # :381
    fx = np.zeros((P, len(extra)))
    for idx, n in enumerate(extra):
        fx[:, idx] = np.asarray(v[n])
    fx = fx.reshape((P, 3, (max_sh_degree + 1) ** 2 - 1))                                             # :386
    sn = sorted([n for n in names if n.startswith("scale_")], key=lambda x: int(x.split("_")[-1]))
    scales = np.stack([np.asarray(v[n]) for n in sn], axis=1).astype(np.float64)
    rn = sorted([n for n in names if n.startswith("rot")], key=lambda x: int(x.split("_")[-1]))
    rots = np.stack([np.asarray(v[n]) for n in rn], axis=1).astype(np.float64)
    f32 = lambda a: np.ascontiguousarray(a, dtype=np.float32)
    out = dict(xyz=f32(xyz),
               features_dc=f32(fdc.transpose(0, 2, 1)),          # .transpose(1, 2).contiguous()  :404
               features_rest=f32(fx.transpose(0, 2, 1)),         # :405
               opacity=f32(opacities), occ_multiplier=f32(occ), dc_delta=f32(dcd), scaling=f32(scales), rotation=f32(rots),
               base_gaussian_mask=np.asarray(v["base_gaussian_mask"]).astype(bool),                   # :399
               max_pixel_sizes=f32(np.asarray(v["max_pixel_sizes"])), min_pixel_sizes=f32(np.asarray(v["min_pixel_sizes"])))
    out["features"] = np.concatenate((out["features_dc"], out["features_rest"]), axis=1)             # get_features :141-144
    return out
