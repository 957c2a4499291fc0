"""CPU oracle for SURVEY 8(f) rank 4: the Gaussian model's PLY file.  TEST INFRASTRUCTURE ONLY (imported by tests/).

Restates GaussianModel.save_ply / load_ply (scene/gaussian_model.py:285-321, 329-380) value by value with
`struct` and Python loops (small cases only), independently of the product's numpy/CUDA row path.

Pin: the reference writes through the third-party `plyfile` package (environment.yml:8, unpinned), which is absent from
/root/reference and from this image, and the tree holds no .ply file, so there is no reference FIXTURE.  The row layout
(attribute order, the f_rest transposition) is pinned by the reference's own code instead: tests/test_plyio_cpu.py runs
its unmodified GaussianModel.save_ply / load_ply on gaussian_splatting_b200.plyfile_compat and requires its file to
equal save_ply_bytes() byte for byte.  What stays unpinned is only plyfile's header spelling, restated here from its
published behaviour — PLY 1.0 for a structured array of '<f4' fields:
    ply\\nformat binary_little_endian 1.0\\nelement vertex <N>\\nproperty float <name>\\n...end_header\\n
followed by N rows of little-endian float32 in property order."""
import struct

import numpy as np


def attribute_names(n_dc, n_rest):                      # construct_list_of_attributes (:285-301)
    l = ["x", "y", "z", "nx", "ny", "nz"]
    l += ["f_dc_{}".format(i) for i in range(n_dc)]
    l += ["f_rest_{}".format(i) for i in range(n_rest)]
    l.append("opacity")
    l += ["scale_{}".format(i) for i in range(3)]
    l += ["rot_{}".format(i) for i in range(4)]
    return l


def save_ply_bytes(xyz, f_dc, f_rest, opacity, scaling, rotation):
    """Bytes of the file save_ply writes.  f_dc (P,1,3), f_rest (P,K,3): `.transpose(1, 2).flatten(start_dim=1)` (:308-309)
    makes the columns channel-major: f_rest_{c*K + k} = f_rest[p, k, c]."""
    P, K = xyz.shape[0], f_rest.shape[1]
    names = attribute_names(3, 3 * K)
    head = "ply\nformat binary_little_endian 1.0\nelement vertex %d\n" % P
    head += "".join("property float %s\n" % n for n in names) + "end_header\n"
    out = [head.encode("ascii")]
    for p in range(P):
        row = [xyz[p, 0], xyz[p, 1], xyz[p, 2], 0.0, 0.0, 0.0]
        row += [f_dc[p, 0, c] for c in range(3)]
        row += [f_rest[p, k, c] for c in range(3) for k in range(K)]
        row += [opacity[p, 0]] + [scaling[p, j] for j in range(3)] + [rotation[p, j] for j in range(4)]
        out.append(struct.pack("<%df" % len(row), *[float(v) for v in row]))
    return b"".join(out)


def load_ply_bytes(data, max_sh_degree):
    """load_ply (:329-380) on the bytes of a binary_little_endian float-only file: dict of the six model arrays."""
    end = data.index(b"end_header\n") + len(b"end_header\n")
    lines = data[:end].decode("ascii").split("\n")
    assert lines[0] == "ply" and lines[1] == "format binary_little_endian 1.0"
    P = int(lines[2].split()[2])
    names = [l.split()[2] for l in lines if l.startswith("property ")]
    assert all(l.split()[1] == "float" for l in lines if l.startswith("property "))
    vals = struct.unpack("<%df" % (P * len(names)), data[end:end + 4 * P * len(names)])
    col = {n: [vals[p * len(names) + c] for p in range(P)] for c, n in enumerate(names)}
    by_index = lambda prefix: sorted([n for n in names if n.startswith(prefix)], key=lambda x: int(x.split("_")[-1]))
    extra = by_index("f_rest_")
    assert len(extra) == 3 * (max_sh_degree + 1) ** 2 - 3
    K = len(extra) // 3
    f = lambda rows: np.array(rows, np.float64).astype(np.float32)
    xyz = f([[col["x"][p], col["y"][p], col["z"][p]] for p in range(P)]).reshape(P, 3)
    opac = f([[col["opacity"][p]] for p in range(P)]).reshape(P, 1)
    dc = f([[[col["f_dc_%d" % c][p] for c in range(3)]] for p in range(P)]).reshape(P, 1, 3)
    # features_extra.reshape(P, 3, K) then .transpose(1, 2) (:358,375): [p, k, c] = column c*K + k
    rest = f([[[col[extra[c * K + k]][p] for c in range(3)] for k in range(K)] for p in range(P)]).reshape(P, K, 3)
    scale = f([[col[n][p] for n in by_index("scale_")] for p in range(P)]).reshape(P, 3)
    rot = f([[col[n][p] for n in by_index("rot")] for p in range(P)]).reshape(P, 4)
    return {"xyz": xyz, "f_dc": dc, "f_rest": rest, "opacity": opac, "scaling": scale, "rotation": rot}
