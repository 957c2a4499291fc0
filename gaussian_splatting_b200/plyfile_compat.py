"""The subset of the third-party `plyfile` API that the reference uses, over plyio's reader / writer.

The reference imports `from plyfile import PlyData, PlyElement` at module level (scene/gaussian_model.py:19,
scene/dataset_readers.py:22); `plyfile` is an unpinned pip dependency (environment.yml:8) that this image does not
have.  Calls the reference makes, all supported here:
    PlyData.read(path)                         gaussian_model.py:330, dataset_readers.py:121
    plydata.elements[0]["x"], plydata['vertex'], vertices['x']          :343-372, dataset_readers.py:122-125
    plydata.elements[0].properties -> objects with .name                :351,360,366
    PlyElement.describe(structured_array, 'vertex'); PlyData([el]).write(path)     :319-321, dataset_readers.py:141-143
`install()` registers this module as `plyfile` only when the real package is not importable."""
import importlib.util
import sys

import numpy as np

from . import plyio


class PlyProperty:
    def __init__(self, name, val_dtype):
        self.name = name
        self.val_dtype = val_dtype

    def __repr__(self):
        return f"PlyProperty({self.name!r}, {self.val_dtype!r})"


class PlyElement:
    def __init__(self, name, data):
        self.name = name
        self.data = data

    @staticmethod
    def describe(data, name, **_):
        if not isinstance(data, np.ndarray) or data.dtype.names is None or data.ndim != 1:
            raise TypeError("only one-dimensional structured arrays are supported")
        return PlyElement(name, data)

    @property
    def properties(self):
        return tuple(PlyProperty(n, self.data.dtype[n].str[1:]) for n in self.data.dtype.names)

    @property
    def count(self):
        return int(self.data.shape[0])

    def __getitem__(self, key):
        return self.data[key]

    def __len__(self):
        return self.count


class PlyData:
    def __init__(self, elements=(), text=False, byte_order="<", **_):
        if text or byte_order not in ("<", "="):
            raise NotImplementedError("only binary_little_endian output is supported")
        self.elements = list(elements)

    @staticmethod
    def read(stream):
        if not isinstance(stream, (str, bytes)) and not hasattr(stream, "__fspath__"):
            raise NotImplementedError("PlyData.read needs a path")
        return PlyData([PlyElement("vertex", plyio.read_ply_table(stream))])

    def __getitem__(self, name):
        for el in self.elements:
            if el.name == name:
                return el
        raise KeyError(name)

    def write(self, stream):
        if len(self.elements) != 1:
            raise NotImplementedError("exactly one element is supported")
        data = self.elements[0].data
        names = list(data.dtype.names)
        types = [data.dtype[n].str[1:] for n in names]
        little = data.astype(np.dtype([(n, "<" + t) for n, t in zip(names, types)]), copy=False)
        plyio.write_ply_table(stream, names, types, little, element=self.elements[0].name)


def install():
    """Make `import plyfile` resolve to this module unless the real package exists."""
    if "plyfile" in sys.modules:
        return sys.modules["plyfile"]
    if importlib.util.find_spec("plyfile") is not None:
        import plyfile
        return plyfile
    sys.modules["plyfile"] = sys.modules[__name__]
    return sys.modules[__name__]
