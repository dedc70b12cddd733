"""Minimal stand-in for the `plyfile` package (absent from this image): the subset the reference's callers use --
`PlyElement.describe(structured_array, 'vertex')`, `PlyData([el]).write(path)`, `PlyData.read(path)`,
`plydata['vertex']` / `plydata.elements[0]` with `el['name']` column access and `el.properties[i].name`
(scene/dataset_readers.py:120-143, scene/gaussian_model.py:240-314).  binary_little_endian 1.0, one element,
scalar properties.  Test infrastructure only."""
import numpy as np

_PLY2NP = {"float": "f4", "float32": "f4", "double": "f8", "float64": "f8", "uchar": "u1", "uint8": "u1", "char": "i1",
           "int8": "i1", "short": "i2", "int16": "i2", "ushort": "u2", "uint16": "u2", "int": "i4", "int32": "i4",
           "uint": "u4", "uint32": "u4"}
_NP2PLY = {"f4": "float", "f8": "double", "u1": "uchar", "i1": "char", "i2": "short", "u2": "ushort", "i4": "int", "u4": "uint"}


class PlyProperty:
    def __init__(self, name, np_type):
        self.name, self.np_type = name, np_type


class PlyElement:
    def __init__(self, name, data):
        self.name, self.data = name, data
        self.properties = [PlyProperty(n, data.dtype[n].str.lstrip("<|=")) for n in data.dtype.names]

    @staticmethod
    def describe(data, name):
        return PlyElement(name, np.asarray(data))

    def __getitem__(self, key):
        return self.data[key]

    def __len__(self):
        return len(self.data)

    count = property(lambda self: len(self.data))


class PlyData:
    def __init__(self, elements):
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
                head.append("element %s %d" % (e.name, len(e.data)))
                for p in e.properties:
                    head.append("property %s %s" % (_NP2PLY[p.np_type], p.name))
            head.append("end_header")
            f.write(("\n".join(head) + "\n").encode("ascii"))
            for e in self.elements:
                le = e.data.astype(e.data.dtype.newbyteorder("<"), copy=False)
                f.write(le.tobytes())

    @staticmethod
    def read(path):
        with open(path, "rb") as f:
            raw = f.read()
        end = raw.index(b"end_header\n") + len(b"end_header\n")
        lines = raw[:end].decode("ascii").split("\n")
        assert lines[0].strip() == "ply" and "binary_little_endian" in lines[1], "stub reads binary_little_endian only"
        elems, cur = [], None
        for ln in lines[2:]:
            t = ln.split()
            if not t:
                continue
            if t[0] == "element":
                cur = [t[1], int(t[2]), []]
                elems.append(cur)
            elif t[0] == "property":
                assert t[1] != "list", "stub: scalar properties only"
                cur[2].append((t[2], "<" + _PLY2NP[t[1]]))
        out, off = [], end
        for name, count, props in elems:
            dt = np.dtype(props)
            out.append(PlyElement(name, np.frombuffer(raw, dtype=dt, count=count, offset=off).copy()))
            off += dt.itemsize * count
        return PlyData(out)
