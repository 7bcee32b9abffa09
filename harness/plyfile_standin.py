"""Minimal stand-in for the third-party `plyfile` package (absent in this image, not vendored by the reference): exactly the
calls GaussianModel.save_ply / load_ply make (scene/gaussian_model.py:340-343, 358-417) —

    PlyElement.describe(structured_array, 'vertex');  PlyData([el]).write(path)
    PlyData.read(path).elements[0][name];  .elements[0].properties[i].name

— for ONE element in `binary_little_endian 1.0`, which is what plyfile writes by default on a little-endian host
(plyfile 0.7-1.0, PlyData(text=False, byte_order='=')).  Type names follow plyfile's table: numpy 'f4' -> float, 'b1' / 'u1' ->
uchar (plyfile lists ('uchar', 'b1') before ('uchar', 'u1'), so a bool column is written as uchar and read back as u1), 'i4' -> int,
'f8' -> double.  TEST INFRASTRUCTURE: tests/golden/make_ply_golden.py injects it as `plyfile` so that the reference's own
save_ply / load_ply can run here; nothing under ms-gs_b200/ imports it."""
import numpy as np

_TO_PLY = {"f4": "float", "f8": "double", "b1": "uchar", "u1": "uchar", "i1": "char", "i2": "short", "u2": "ushort", "i4": "int",
           "u4": "uint"}
_FROM_PLY = {"float": "<f4", "float32": "<f4", "double": "<f8", "float64": "<f8", "uchar": "u1", "uint8": "u1", "char": "i1",
             "int8": "i1", "short": "<i2", "int16": "<i2", "ushort": "<u2", "uint16": "<u2", "int": "<i4", "int32": "<i4",
             "uint": "<u4", "uint32": "<u4"}


class PlyProperty:
    def __init__(self, name, ply_type):
        self.name, self.ply_type = name, ply_type


class PlyElement:
    def __init__(self, name, data):
        self.name, self.data = name, data
        self.properties = tuple(PlyProperty(n, _TO_PLY[data.dtype[n].str.lstrip("<>=|")]) for n in data.dtype.names)

    @staticmethod
    def describe(data, name):
        if data.dtype.names is None or data.ndim != 1:
            raise ValueError("only one-dimensional structured arrays are supported")
        return PlyElement(name, data)

    def __getitem__(self, key):
        return self.data[key]

    @property
    def count(self):
        return len(self.data)


class PlyData:
    def __init__(self, elements=(), text=False, byte_order="="):
        self.elements = list(elements)

    def write(self, path):
        (el,) = self.elements
        head = ["ply", "format binary_little_endian 1.0", f"element {el.name} {el.count}"]
        head += [f"property {p.ply_type} {p.name}" for p in el.properties]
        head.append("end_header")
        packed = np.dtype([(n, el.data.dtype[n].newbyteorder("<") if el.data.dtype[n].itemsize > 1 else el.data.dtype[n])
                           for n in el.data.dtype.names])
        rows = np.empty(el.count, dtype=packed)               # packed (no alignment padding), little endian
        for n in packed.names:
            rows[n] = el.data[n]
        with open(path, "wb") as f:
            f.write(("\n".join(head) + "\n").encode("ascii"))
            rows.tofile(f)

    @staticmethod
    def read(path):
        with open(path, "rb") as f:
            raw = f.read()
        end = raw.index(b"end_header\n") + len(b"end_header\n")
        lines = raw[:end].decode("ascii").split("\n")
        if lines[0] != "ply" or lines[1] != "format binary_little_endian 1.0":
            raise ValueError("stand-in reads binary_little_endian files only")
        name, count, fields = None, 0, []
        for ln in lines[2:]:
            w = ln.split()
            if not w or w[0] in ("comment", "obj_info", "end_header"):
                continue
            if w[0] == "element":
                if name is not None:
                    raise ValueError("stand-in reads one element")
                name, count = w[1], int(w[2])
            elif w[0] == "property":
                fields.append((w[2], _FROM_PLY[w[1]]))
        data = np.frombuffer(raw, dtype=np.dtype(fields), count=count, offset=end).copy()
        return PlyData([PlyElement(name, data)])
