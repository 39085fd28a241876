"""Minimal stand-in for the `plyfile` package, enough for what the reference does with it: write one `vertex`
element made from a numpy structured array (scene/dataset_readers.py:118-133, scene/gaussian_model.py:190-208) and
read such a file back by property name (dataset_readers.py:110-116, gaussian_model.py:215-249). Binary
little-endian PLY with scalar properties only; not a general PLY library. It is only picked up when the real
package is absent (event_3dgs_b200/dropin sits last on the path for this name: see INTEGRATION.md)."""
import numpy as np

_PLY_TYPES = {"f4": "float", "f8": "double", "u1": "uchar", "i1": "char", "u2": "ushort", "i2": "short", "u4": "uint", "i4": "int"}
_NP_TYPES = {v: k for k, v in _PLY_TYPES.items()}
_NP_TYPES.update({"float32": "f4", "float64": "f8", "uint8": "u1", "int8": "i1", "uint16": "u2", "int16": "i2", "uint32": "u4", "int32": "i4"})


class PlyProperty:
    def __init__(self, name, val_dtype):
        self.name, self.val_dtype = name, val_dtype


class PlyElement:
    def __init__(self, name, data):
        self.name, self.data = name, data
        self.properties = tuple(PlyProperty(n, data.dtype[n].str.lstrip("<|=")) for n in data.dtype.names)

    @staticmethod
    def describe(data, name):
        if data.dtype.names is None:
            raise ValueError("PlyElement.describe wants a numpy structured array")
        return PlyElement(name, np.asarray(data))

    @property
    def count(self):
        return len(self.data)

    def __getitem__(self, key):
        return self.data[key]

    def __len__(self):
        return len(self.data)


class PlyData:
    def __init__(self, elements=(), text=False, byte_order="<"):
        self.elements = list(elements)

    def __getitem__(self, name):
        for e in self.elements:
            if e.name == name:
                return e
        raise KeyError(name)

    def write(self, path):
        with open(path, "wb") as f:
            head = ["ply", "format binary_little_endian 1.0"]
            for e in self.elements:
                head.append("element %s %d" % (e.name, e.count))
                for p in e.properties:
                    head.append("property %s %s" % (_PLY_TYPES[p.val_dtype], p.name))
            head.append("end_header")
            f.write(("\n".join(head) + "\n").encode("ascii"))
            for e in self.elements:
                le = e.data.astype(e.data.dtype.newbyteorder("<"), copy=False)
                f.write(np.ascontiguousarray(le).tobytes())

    @staticmethod
    def read(path):
        with open(path, "rb") as f:
            if f.readline().strip() != b"ply":
                raise ValueError("%s is not a PLY file" % path)
            fmt, elements, cur = None, [], None
            while True:
                line = f.readline()
                if not line:
                    raise ValueError("unterminated PLY header")
                tok = line.decode("ascii").split()
                if not tok or tok[0] == "comment":
                    continue
                if tok[0] == "format":
                    fmt = tok[1]
                elif tok[0] == "element":
                    cur = [tok[1], int(tok[2]), []]
                    elements.append(cur)
                elif tok[0] == "property":
                    if tok[1] == "list":
                        raise ValueError("list properties are not supported by this stand-in")
                    cur[2].append((tok[2], _NP_TYPES[tok[1]]))
                elif tok[0] == "end_header":
                    break
            out = []
            for name, count, props in elements:
                if fmt == "ascii":
                    rows = [tuple(f.readline().split()) for _ in range(count)]
                    data = np.array(rows, dtype=[(n, t) for n, t in props])
                else:
                    order = "<" if fmt == "binary_little_endian" else ">"
                    dt = np.dtype([(n, order + t if t[1] != "1" else t) for n, t in props])
                    data = np.frombuffer(f.read(dt.itemsize * count), dtype=dt, count=count).astype([(n, t) for n, t in props])
                out.append(PlyElement(name, data))
        return PlyData(out)
