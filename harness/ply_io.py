"""Writer of MS-GS model files for the tests / the loader benchmark: restates GaussianModel.save_ply
(/root/reference/scene/gaussian_model.py:293-344) — attribute order of construct_list_of_attributes, float32 columns,
`base_gaussian_mask` as a 1-byte column (numpy bool -> PLY uchar in plyfile), binary_little_endian, one vertex element —
and a seeded random model in the reference's parameter shapes."""
import numpy as np


def attribute_list(n_lvl_occ, n_lvl_dc, n_rest=45, ms=True):
    l = ["x", "y", "z", "nx", "ny", "nz"] + [f"f_dc_{i}" for i in range(3)] + [f"f_rest_{i}" for i in range(n_rest)] + ["opacity"]
    if ms:
        l += [f"occ_multiplier_{i}" for i in range(n_lvl_occ)]
        l += [f"dc_delta_{i}_{j}" for i in range(n_lvl_dc) for j in range(3)]
    l += [f"scale_{i}" for i in range(3)] + [f"rot_{i}" for i in range(4)]
    if ms:
        l += ["base_gaussian_mask", "max_pixel_sizes", "min_pixel_sizes"]
    return l


def random_model(P, n_lvl_occ=4, n_lvl_dc=4, seed=0, M=16):
    """parameter tensors in GaussianModel's own shapes (:52-66): _features_dc (P,1,3), _features_rest (P,M-1,3), ..."""
    g = np.random.default_rng(seed)
    f = lambda *s: g.standard_normal(s).astype(np.float32)
    return dict(xyz=f(P, 3), features_dc=f(P, 1, 3), features_rest=f(P, M - 1, 3), opacity=f(P, 1),
                occ_multiplier=f(P, n_lvl_occ, 1), dc_delta=f(P, n_lvl_dc * 3, 1), scaling=f(P, 3), rotation=f(P, 4),
                base_gaussian_mask=g.random(P) < 0.1,
                max_pixel_sizes=np.where(g.random(P) < 0.2, g.random(P) * 8, -1).astype(np.float32),
                min_pixel_sizes=(g.random(P) * 3).astype(np.float32))


def write_ms_ply(path, m, ms=True):
    """save_ply (:314-344): f_dc / f_rest are the (P,K,3) parameters transposed to (P,3,K) and flattened."""
    P = m["xyz"].shape[0]
    L, D = m["occ_multiplier"].shape[1], m["dc_delta"].shape[1] // 3
    n_rest = m["features_rest"].shape[1] * 3
    names = attribute_list(L, D, n_rest, ms)
    dtype = np.dtype([(n, "u1" if n == "base_gaussian_mask" else "<f4") for n in names])
    cols = [m["xyz"], np.zeros_like(m["xyz"]), m["features_dc"].transpose(0, 2, 1).reshape(P, 3),
            m["features_rest"].transpose(0, 2, 1).reshape(P, n_rest), m["opacity"]]
    if ms:
        cols += [m["occ_multiplier"].reshape(P, L), m["dc_delta"].reshape(P, 3 * D)]
    cols += [m["scaling"], m["rotation"]]
    flat = np.concatenate(cols, axis=1).astype(np.float32)
    rows = np.empty(P, dtype=dtype)
    fnames = [n for n in names if n not in ("base_gaussian_mask", "max_pixel_sizes", "min_pixel_sizes")]
    assert flat.shape[1] == len(fnames)
    for k, n in enumerate(fnames):
        rows[n] = flat[:, k]
    if ms:
        rows["base_gaussian_mask"] = m["base_gaussian_mask"].astype(np.uint8)
        rows["max_pixel_sizes"] = m["max_pixel_sizes"]
        rows["min_pixel_sizes"] = m["min_pixel_sizes"]
    header = "ply\nformat binary_little_endian 1.0\nelement vertex %d\n" % P
    header += "".join("property %s %s\n" % ("uchar" if n == "base_gaussian_mask" else "float", n) for n in names)
    header += "end_header\n"
    with open(path, "wb") as f:
        f.write(header.encode("ascii"))
        rows.tofile(f)
    return names
