"""On-disk formats of the Gaussian model.

Host-side mirror of the reference's I/O, same names and file bytes:
    save_ply(gm, path) / load_ply(gm, path, use_train_test_exp=False)   scene/gaussian_model.py:303-321, 329-380
    construct_list_of_attributes(gm)                                    scene/gaussian_model.py:285-301
    storePly(path, xyz, rgb) / fetchPly(path)                           scene/dataset_readers.py:120-143
    save_checkpoint(gm, iteration, path) / load_checkpoint(path)        train.py:62,256 + GaussianModel.capture (:94-108)

The reference writes through the third-party `plyfile` package (environment.yml:8, unpinned, absent from
/root/reference and from this image).  What it produces is the PLY 1.0 format: an ASCII header
(`ply / format binary_little_endian 1.0 / element vertex N / property float <name> ... / end_header`) followed by
N packed little-endian rows.  This module writes exactly that and reads any PLY whose first element has scalar
properties (binary little/big endian or ascii, any property order and numeric type), by property NAME as the
reference does.

B200 path: the row matrix [P][17 + 3K] is built (save) or scattered back into the six model tensors (load) by ONE
kernel each (libgsrast.so: gsr_ply_pack / gsr_ply_unpack, csrc/densify.cu) and crosses PCIe as one contiguous copy
through pinned memory, instead of the reference's seven `.cpu()` copies plus a per-row Python `tuple` loop (:318)
and 62 per-property column gathers (:349-372).  No CPU path for the model tensors: they must be CUDA tensors.
"""
import json
import os

import numpy as np
import torch
from torch import nn

from . import _lib

# PLY scalar type names (both spellings the format allows) -> numpy codes
_PLY_TYPES = {"char": "i1", "uchar": "u1", "short": "i2", "ushort": "u2", "int": "i4", "uint": "u4", "float": "f4",
              "double": "f8", "int8": "i1", "uint8": "u1", "int16": "i2", "uint16": "u2", "int32": "i4", "uint32": "u4",
              "float32": "f4", "float64": "f8"}
_PLY_NAMES = {"i1": "char", "u1": "uchar", "i2": "short", "u2": "ushort", "i4": "int", "u4": "uint", "f4": "float",
              "f8": "double"}


def _require_cuda(t, what):
    if not t.is_cuda:
        raise RuntimeError(f"{what} must be a CUDA tensor (the Gaussian file path has no CPU implementation)")


_device_ctx = torch.cuda.device


def _stream(device):
    return torch.cuda.current_stream(device).cuda_stream


def _to_host(rows):
    host = torch.empty(rows.shape, dtype=rows.dtype).pin_memory()
    host.copy_(rows, non_blocking=True)
    torch.cuda.current_stream(rows.device).synchronize()
    return host


def _to_device(host, device):
    return host.pin_memory().to(device, non_blocking=True)


# ------------------------------------------------------------------ generic PLY table reader / writer (host)
def write_ply_table(path, names, types, table, element="vertex"):
    """`table`: structured or plain 2-D array with one column per property, written as one element."""
    count = int(table.shape[0])
    lines = ["ply", "format binary_little_endian 1.0", f"element {element} {count}"]
    lines += [f"property {_PLY_NAMES[t]} {n}" for n, t in zip(names, types)]
    lines.append("end_header")
    with open(path, "wb") as f:
        f.write(("\n".join(lines) + "\n").encode("ascii"))
        if table.dtype.names is None:
            assert table.dtype == np.dtype("<" + types[0]) and all(t == types[0] for t in types)
            np.ascontiguousarray(table).tofile(f)
        else:
            table.tofile(f)


def read_ply_table(path):
    """First element of a PLY file as a numpy structured array (native byte order), like
    `PlyData.read(path).elements[0]`: fields are addressed by property name."""
    with open(path, "rb") as f:
        if f.readline().strip() != b"ply":
            raise ValueError(f"{path}: not a PLY file")
        fmt, elements = None, []
        while True:
            line = f.readline()
            if not line:
                raise ValueError(f"{path}: unexpected end of header")
            tok = line.decode("ascii", "replace").split()
            if not tok or tok[0] in ("comment", "obj_info"):
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                elements.append((tok[1], int(tok[2]), []))
            elif tok[0] == "property":
                if tok[1] == "list":
                    if len(elements) == 1:
                        raise ValueError(f"{path}: list properties in the first element are not supported")
                    elements[-1][2].append((tok[-1], None))
                else:
                    if tok[1] not in _PLY_TYPES:
                        raise ValueError(f"{path}: unknown property type {tok[1]}")
                    elements[-1][2].append((tok[2], _PLY_TYPES[tok[1]]))
            elif tok[0] == "end_header":
                break
        if not elements or fmt not in ("binary_little_endian", "binary_big_endian", "ascii"):
            raise ValueError(f"{path}: unsupported PLY header")
        _, count, props = elements[0]
        if fmt == "ascii":
            cols = len(props)
            vals = []
            while len(vals) < count * cols:
                line = f.readline()
                if not line:
                    raise ValueError(f"{path}: truncated data")
                vals += line.split()
            out = np.empty(count, dtype=[(n, "=" + t) for n, t in props])
            for c, (n, t) in enumerate(props):
                out[n] = np.array(vals[c:count * cols:cols], dtype="f8" if t[0] == "f" else "i8").astype(t)
            return out
        order = "<" if fmt == "binary_little_endian" else ">"
        dt = np.dtype([(n, order + t) for n, t in props])
        data = np.fromfile(f, dtype=dt, count=count)
        if data.shape[0] != count:
            raise ValueError(f"{path}: truncated data")
        return data.astype(dt.newbyteorder("="), copy=False)


# ------------------------------------------------------------------ Gaussian model files
def construct_list_of_attributes(gm):
    """gaussian_model.py:285-301: x y z nx ny nz f_dc_* f_rest_* opacity scale_* rot_*."""
    names = ["x", "y", "z", "nx", "ny", "nz"]
    names += [f"f_dc_{i}" for i in range(gm._features_dc.shape[1] * gm._features_dc.shape[2])]
    names += [f"f_rest_{i}" for i in range(gm._features_rest.shape[1] * gm._features_rest.shape[2])]
    names.append("opacity")
    names += [f"scale_{i}" for i in range(gm._scaling.shape[1])]
    names += [f"rot_{i}" for i in range(gm._rotation.shape[1])]
    return names


def _model_tensors(gm):
    ts = []
    for attr in ("_xyz", "_features_dc", "_features_rest", "_opacity", "_scaling", "_rotation"):
        t = getattr(gm, attr).detach()
        _require_cuda(t, f"GaussianModel.{attr}")
        if t.dtype != torch.float32:
            raise RuntimeError(f"GaussianModel.{attr} must be float32")
        ts.append(t if t.is_contiguous() else t.contiguous())
    xyz, f_dc, f_rest, opacity, scaling, rotation = ts
    P = int(xyz.shape[0])
    if tuple(xyz.shape) != (P, 3) or tuple(f_dc.shape) != (P, 1, 3) or f_rest.dim() != 3 or f_rest.shape[0] != P or \
            f_rest.shape[2] != 3 or tuple(opacity.shape) != (P, 1) or tuple(scaling.shape) != (P, 3) or \
            tuple(rotation.shape) != (P, 4):
        raise RuntimeError("unexpected GaussianModel tensor shapes (expected xyz (P,3), f_dc (P,1,3), f_rest (P,K,3), "
                           "opacity (P,1), scaling (P,3), rotation (P,4))")
    return ts, P, int(f_rest.shape[1])


@torch.no_grad()
def pack_rows(gm):
    """The PLY row matrix [P][17 + 3K] of the model, on the model's device."""
    lib = _lib.load()
    (xyz, f_dc, f_rest, opacity, scaling, rotation), P, K = _model_tensors(gm)
    rows = torch.empty((P, 17 + 3 * K), dtype=torch.float32, device=xyz.device)
    if P:
        with _device_ctx(xyz.device):
            _lib.check(lib.gsr_ply_pack(P, K, xyz.data_ptr(), f_dc.data_ptr(), f_rest.data_ptr() if K else None,
                                        opacity.data_ptr(), scaling.data_ptr(), rotation.data_ptr(), rows.data_ptr(),
                                        _stream(xyz.device)))
    return rows


@torch.no_grad()
def save_ply(gm, path):
    """GaussianModel.save_ply (:303-321): same header, same row bytes."""
    d = os.path.dirname(path)
    if d:
        os.makedirs(d, exist_ok=True)                 # mkdir_p (utils/system_utils.py)
    names = construct_list_of_attributes(gm)
    rows = pack_rows(gm)
    host = _to_host(rows).numpy() if rows.numel() else np.zeros(tuple(rows.shape), np.float32)
    write_ply_table(path, names, ["f4"] * len(names), host)


def _sorted_by_index(names, prefix):
    return sorted((n for n in names if n.startswith(prefix)), key=lambda x: int(x.split("_")[-1]))


@torch.no_grad()
def load_ply(gm, path, use_train_test_exp=False, device="cuda"):
    """GaussianModel.load_ply (:329-380): properties are looked up by name (f_rest_*, scale_*, rot* sorted by their
    numeric suffix), the six tensors become fresh nn.Parameters on `device`, active_sh_degree = max_sh_degree."""
    lib = _lib.load()
    el = read_ply_table(path)
    if use_train_test_exp:                            # :331-341
        exposure_file = os.path.join(os.path.dirname(path), os.pardir, os.pardir, "exposure.json")
        if os.path.exists(exposure_file):
            with open(exposure_file, "r") as f:
                exposures = json.load(f)
            gm.pretrained_exposures = {k: torch.FloatTensor(v).requires_grad_(False).to(device)
                                       for k, v in exposures.items()}
            print("Pretrained exposures loaded.")
        else:
            print(f"No exposure to be loaded at {exposure_file}")
            gm.pretrained_exposures = None
    names = list(el.dtype.names)
    rest = _sorted_by_index(names, "f_rest_")
    assert len(rest) == 3 * (gm.max_sh_degree + 1) ** 2 - 3          # :353
    scale_names = _sorted_by_index(names, "scale_")
    rot_names = _sorted_by_index(names, "rot")
    if len(scale_names) != 3 or len(rot_names) != 4:
        raise RuntimeError(f"{path}: expected 3 scale_* and 4 rot_* properties")
    K = len(rest) // 3
    canon = ["x", "y", "z", "nx", "ny", "nz", "f_dc_0", "f_dc_1", "f_dc_2"] + rest + ["opacity"] + scale_names + rot_names
    P = int(el.shape[0])
    if names == canon and all(el.dtype[n] == np.dtype("f4") for n in names):
        rows = np.ascontiguousarray(el).view(np.float32).reshape(P, len(canon))   # the file's rows ARE the canonical
                                                                                  # matrix: no host gather
    else:
        rows = np.empty((P, len(canon)), np.float32)
        for c, n in enumerate(canon):
            rows[:, c] = el[n] if n in names else 0.0   # the reference never reads nx ny nz
    dev = torch.device(device)
    new = {"_xyz": (P, 3), "_features_dc": (P, 1, 3), "_features_rest": (P, K, 3), "_opacity": (P, 1), "_scaling": (P, 3),
           "_rotation": (P, 4)}
    t = {k: torch.empty(s, dtype=torch.float32, device=dev) for k, s in new.items()}
    if P:
        _require_cuda(t["_xyz"], "load_ply target device")
        with _device_ctx(dev):
            drows = _to_device(torch.from_numpy(np.ascontiguousarray(rows)), dev)
            _lib.check(lib.gsr_ply_unpack(P, K, drows.data_ptr(), t["_xyz"].data_ptr(), t["_features_dc"].data_ptr(),
                                          t["_features_rest"].data_ptr() if K else None, t["_opacity"].data_ptr(),
                                          t["_scaling"].data_ptr(), t["_rotation"].data_ptr(), _stream(dev)))
    for k, v in t.items():
        setattr(gm, k, nn.Parameter(v.requires_grad_(True)))
    gm.active_sh_degree = gm.max_sh_degree
    return gm


# ------------------------------------------------------------------ input point clouds (dataset_readers.py:120-143)
def storePly(path, xyz, rgb):
    xyz = np.asarray(xyz)
    rgb = np.asarray(rgb)
    dtype = [("x", "<f4"), ("y", "<f4"), ("z", "<f4"), ("nx", "<f4"), ("ny", "<f4"), ("nz", "<f4"),
             ("red", "u1"), ("green", "u1"), ("blue", "u1")]
    el = np.zeros(xyz.shape[0], dtype=dtype)          # normals = zeros (:134)
    for c, n in enumerate(("x", "y", "z")):
        el[n] = xyz[:, c]
    for c, n in enumerate(("red", "green", "blue")):
        el[n] = rgb[:, c]
    write_ply_table(path, [d[0] for d in dtype], ["f4"] * 6 + ["u1"] * 3, el)


def fetchPly(path):
    """(positions (N,3), colors (N,3) in [0,1], normals (N,3)) — the fields of BasicPointCloud (:120-126)."""
    v = read_ply_table(path)
    positions = np.vstack([v["x"], v["y"], v["z"]]).T
    colors = np.vstack([v["red"], v["green"], v["blue"]]).T / 255.0
    normals = np.vstack([v["nx"], v["ny"], v["nz"]]).T
    return positions, colors, normals


# ------------------------------------------------------------------ checkpoints (train.py:62,256)
def capture(gm):
    """GaussianModel.capture (:94-108): the 12-tuple train.py pickles."""
    return (gm.active_sh_degree, gm._xyz, gm._features_dc, gm._features_rest, gm._scaling, gm._rotation, gm._opacity,
            gm.max_radii2D, gm.xyz_gradient_accum, gm.denom, gm.optimizer.state_dict(), gm.spatial_lr_scale)


def save_checkpoint(gm, iteration, path):
    """train.py:256: torch.save((gaussians.capture(), iteration), model_path + "/chkpnt<iteration>.pth")."""
    torch.save((capture(gm), iteration), path)


def load_checkpoint(path, map_location=None):
    """train.py:62: (model_params, first_iter) = torch.load(checkpoint); model_params feeds GaussianModel.restore."""
    return torch.load(path, map_location=map_location, weights_only=False)
