"""On-disk formats either side of the rasterizer -- host-side, numpy only.

* Gaussian PLY, exactly the file `Ubody_Gaussian.save_gaussian_ply` writes
  (models/UbodyAvatar/ubody_gaussian.py:350-420 via plyfile: binary little-endian, one `vertex` element,
  17 float32 properties x y z nx ny nz f_dc_0..2 opacity scale_0..2 rot_0..3 -- the standard 3DGS layout
  with zero higher-order SH).  Stored values are the *pre-activation* ones: colour / C0 ("RGB to SHS",
  :357), inverse_sigmoid(opacity) (:358), log(scale) (:360), raw wxyz rotation (:362).
* Avatar state: the tensor dictionary `debug_assets.py:97` / `debug_frame.py:73` load with
  `load_state_dict(torch.load(path))` -- the `_smplx_*` / `_uv_*` attributes of `Ubody_Gaussian`
  (ubody_gaussian.py:245-289) -- mapped to the keyword arguments of `deform_gaussians`.
  (`create_avatar.py:70` pickles the whole nn.Module; unpickling that needs GUAVA's class definitions and
  is out of scope -- save `module.state_dict()` instead.)
plyfile is not a dependency: the PLY header/body codec below is self-contained and also reads ASCII and
big-endian files with other scalar property types.
"""
from __future__ import annotations

import os
from typing import Dict, Optional

import numpy as np
import torch

C0 = 0.28209479177387814          # SH band-0 constant (ubody_gaussian.py:357)
PLY_ATTRIBUTES = ("x", "y", "z", "nx", "ny", "nz", "f_dc_0", "f_dc_1", "f_dc_2", "opacity",
                  "scale_0", "scale_1", "scale_2", "rot_0", "rot_1", "rot_2", "rot_3")   # construct_list_of_attributes :408-420

_PLY_TYPES = {"char": "i1", "int8": "i1", "uchar": "u1", "uint8": "u1", "short": "i2", "int16": "i2",
              "ushort": "u2", "uint16": "u2", "int": "i4", "int32": "i4", "uint": "u4", "uint32": "u4",
              "float": "f4", "float32": "f4", "double": "f8", "float64": "f8"}


def _np(t):
    return t.detach().cpu().numpy() if torch.is_tensor(t) else np.asarray(t)


def inverse_sigmoid(x):
    """utils/general_utils.py inverse_sigmoid: log(x / (1 - x))."""
    return np.log(x / (1 - x))


def write_gaussian_ply(path: str, xyz, features_color, opacity, scaling, rotation) -> None:
    """Write one cloud ([P,3], [P,>=3], [P,1], [P,3], [P,4] wxyz; activated values as the rasterizer takes
    them) in the reference's GS_canonical.ply format."""
    xyz = _np(xyz).astype(np.float32).reshape(-1, 3)
    P = xyz.shape[0]
    fc = _np(features_color).astype(np.float32)
    rgb = fc.reshape(P, fc.shape[-1])[:, :3]
    cols = np.concatenate([
        xyz, np.zeros_like(xyz), rgb / np.float32(C0),
        inverse_sigmoid(_np(opacity).astype(np.float32).reshape(P, 1)),
        np.log(_np(scaling).astype(np.float32).reshape(P, 3)),
        _np(rotation).astype(np.float32).reshape(P, 4)], axis=1).astype("<f4")
    header = "ply\nformat binary_little_endian 1.0\nelement vertex %d\n" % P
    header += "".join("property float %s\n" % a for a in PLY_ATTRIBUTES) + "end_header\n"
    tmp = path + ".tmp"
    with open(tmp, "wb") as f:
        f.write(header.encode("ascii"))
        f.write(np.ascontiguousarray(cols).tobytes())
    os.replace(tmp, path)


def read_ply_vertices(path: str) -> Dict[str, np.ndarray]:
    """The scalar properties of the `vertex` element of any PLY (ascii / binary LE / binary BE)."""
    with open(path, "rb") as f:
        if f.readline().strip() != b"ply":
            raise ValueError(f"{path}: not a PLY file")
        fmt, elements, cur = None, [], None
        while True:
            line = f.readline()
            if not line:
                raise ValueError(f"{path}: truncated PLY header")
            tok = line.decode("ascii", "replace").split()
            if not tok or tok[0] in ("comment", "obj_info"):
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                cur = dict(name=tok[1], count=int(tok[2]), props=[])
                elements.append(cur)
            elif tok[0] == "property":
                if tok[1] == "list":
                    cur["props"].append(("list", tok[2], tok[3], tok[4]))
                else:
                    if tok[1] not in _PLY_TYPES:
                        raise ValueError(f"{path}: unknown property type {tok[1]}")
                    cur["props"].append((tok[2], _PLY_TYPES[tok[1]]))
            elif tok[0] == "end_header":
                break
        if fmt not in ("ascii", "binary_little_endian", "binary_big_endian"):
            raise ValueError(f"{path}: unsupported PLY format {fmt}")
        for el in elements:
            scalar = all(p[0] != "list" for p in el["props"])
            if el["name"] != "vertex":
                if not scalar or fmt == "ascii":
                    if fmt == "ascii":
                        for _ in range(el["count"]):
                            f.readline()
                        continue
                    raise ValueError(f"{path}: a list element precedes the vertex element")
                f.seek(el["count"] * sum(np.dtype(t).itemsize for _, t in el["props"]), os.SEEK_CUR)
                continue
            if not scalar:
                raise ValueError(f"{path}: list properties in the vertex element")
            n = el["count"]
            if fmt == "ascii":
                rows = np.array([f.readline().split() for _ in range(n)], dtype=np.float64).reshape(n, len(el["props"]))
                return {name: rows[:, i].astype(t) for i, (name, t) in enumerate(el["props"])}
            order = "<" if fmt == "binary_little_endian" else ">"
            dt = np.dtype([(name, order + t) for name, t in el["props"]])
            raw = f.read(n * dt.itemsize)
            if len(raw) != n * dt.itemsize:
                raise ValueError(f"{path}: truncated PLY body ({len(raw)} of {n * dt.itemsize} bytes)")
            arr = np.frombuffer(raw, dtype=dt, count=n)
            return {name: np.ascontiguousarray(arr[name]).astype(t) for name, t in el["props"]}
    raise ValueError(f"{path}: no vertex element")


def read_gaussian_ply(path: str, num_channels: int = 32, normalize_rotation: bool = True,
                      device: Optional[str] = None) -> Dict[str, torch.Tensor]:
    """Inverse of `write_gaussian_ply`: activated tensors with a leading batch axis of 1, keyed like the
    reference's `deformed_assets` (ready for `render_views`).  Channels 0-2 of `features_color` are the RGB
    the file stores (f_dc * C0); the remaining channels, which the PLY format has no room for, are zero."""
    v = read_ply_vertices(path)
    missing = [a for a in PLY_ATTRIBUTES if a not in v and not a.startswith("n")]
    if missing:
        raise ValueError(f"{path}: missing Gaussian properties {missing}")
    col = lambda *names: np.stack([v[n].astype(np.float32) for n in names], 1)
    P = len(v["x"])
    feat = np.zeros((P, num_channels), np.float32)
    feat[:, :3] = col("f_dc_0", "f_dc_1", "f_dc_2") * np.float32(C0)
    rot = col("rot_0", "rot_1", "rot_2", "rot_3")
    if normalize_rotation:   # the rasterizer uses the quaternion as-is (forward.cu:123)
        rot = rot / np.maximum(np.linalg.norm(rot, axis=1, keepdims=True), 1e-12).astype(np.float32)
    out = dict(xyz=col("x", "y", "z"), features_color=feat,
               opacity=(1.0 / (1.0 + np.exp(-col("opacity")))).astype(np.float32),
               scaling=np.exp(col("scale_0", "scale_1", "scale_2")), rotation=rot)
    return {k: torch.from_numpy(np.ascontiguousarray(a))[None].to(device or "cpu") for k, a in out.items()}


_AVATAR_KEYS = {   # Ubody_Gaussian attribute -> deform_gaussians keyword
    "_smplx_rotation": "smplx_rotation", "_smplx_scaling": "smplx_scaling", "_smplx_opacity": "smplx_opacity",
    "_smplx_features_color": "smplx_features_color", "_uv_local_xyz": "uv_local_xyz", "_uv_rotation": "uv_rotation",
    "_uv_scaling": "uv_scaling", "_uv_opacity": "uv_opacity", "_uv_features_color": "uv_features_color",
    "_uv_binding_face": "uv_binding_face", "_uv_face_bary": "uv_face_bary",
}


def save_avatar_state(path: str, **tensors) -> None:
    """Store the per-avatar tensors under the reference's attribute names (a plain tensor dictionary)."""
    inv = {v: k for k, v in _AVATAR_KEYS.items()}
    unknown = [k for k in tensors if k not in inv]
    if unknown:
        raise KeyError(f"unknown avatar tensors {unknown}")
    torch.save({inv[k]: (t.detach().cpu() if torch.is_tensor(t) else torch.as_tensor(t)) for k, t in tensors.items()}, path)


def load_avatar_state(path: str, device: Optional[str] = None) -> Dict[str, torch.Tensor]:
    """Read a tensor dictionary saved from `Ubody_Gaussian` (state_dict or attribute dump) and return the
    per-avatar keyword arguments of `deform_gaussians`; keys of other sub-modules (ehm.*, smplx.*) are ignored."""
    state = torch.load(path, map_location="cpu", weights_only=True)
    if not isinstance(state, dict):
        raise ValueError(f"{path}: expected a tensor dictionary")
    out = {}
    for k, t in state.items():
        name = _AVATAR_KEYS.get(k.split(".")[-1])
        if name is not None and torch.is_tensor(t):
            out[name] = t.to(device or "cpu")
    missing = sorted(set(_AVATAR_KEYS.values()) - set(out))
    if missing:
        raise KeyError(f"{path}: avatar tensors missing: {missing}")
    return out
