"""The on-disk format either side of the render path: Gaussian-model .ply files as the
reference writes and reads them (/root/reference/scene/gaussian_model.py:279-309 save_ply, :316-358 load_ply;
called from scene/__init__.py:77-87), without `plyfile` and without per-property Python loops.

The body of such a file is a row-major [P, n_props] little-endian binary32 matrix.  `load_ply` reads it with
ONE read straight into pinned host memory, copies it to the device once, and de-interleaves it into the six
parameter tensors with one kernel (csrc/gsr_ply.cu: `gsr_ply_unpack`); `save_ply` is the reverse
(`gsr_ply_pack`, one device-to-host copy, one write).  Only the header is parsed in Python: property names
are resolved to columns the way load_ply does (by name; f_rest_* / scale_* / rot* ordered by numeric suffix).
The tensors never take a CPU detour: there is no CPU path for the (de)interleave.
"""
from __future__ import annotations

import ctypes
import os
from typing import Dict, List, Tuple

import numpy as np
import torch

from . import _native

__all__ = ["load_ply", "save_ply", "read_header", "resolve_columns", "attribute_names", "header_bytes"]

_FLOAT_TYPES = ("float", "float32")


def attribute_names(n_rest: int = 45) -> List[str]:
    """scene/gaussian_model.py:279-291 construct_list_of_attributes."""
    return (["x", "y", "z", "nx", "ny", "nz"] + [f"f_dc_{i}" for i in range(3)]
            + [f"f_rest_{i}" for i in range(n_rest)] + ["opacity"] + [f"scale_{i}" for i in range(3)]
            + [f"rot_{i}" for i in range(4)])


def header_bytes(n_vertex: int, names: List[str]) -> bytes:
    """The header `plyfile` emits for PlyData([PlyElement.describe(elements, 'vertex')]) with 'f4' fields on a
    little-endian host (scene/gaussian_model.py:303-309)."""
    lines = ["ply", "format binary_little_endian 1.0", f"element vertex {n_vertex}"]
    lines += [f"property float {n}" for n in names]
    lines.append("end_header")
    return ("\n".join(lines) + "\n").encode("ascii")


def read_header(f) -> Tuple[int, List[str], int]:
    """Parses a PLY header from a binary file object.  Returns (vertex count, property names, body offset).
    Only what the reference writes is accepted: binary little-endian, a `vertex` element of binary32
    properties (further elements must be empty)."""
    if f.readline().strip() != b"ply":
        raise ValueError("not a PLY file")
    fmt, n_vertex, names, current = None, None, [], None
    while True:
        line = f.readline()
        if not line:
            raise ValueError("PLY header is not terminated by end_header")
        tok = line.decode("ascii", "replace").split()
        if not tok or tok[0] in ("comment", "obj_info"):
            continue
        if tok[0] == "format":
            fmt = tok[1]
        elif tok[0] == "element":
            current = tok[1]
            if current == "vertex":
                n_vertex = int(tok[2])
            elif int(tok[2]) != 0:
                raise NotImplementedError(f"PLY element '{current}' with {tok[2]} entries: only 'vertex' is supported")
        elif tok[0] == "property" and current == "vertex":
            if tok[1] not in _FLOAT_TYPES:
                raise NotImplementedError(f"PLY property '{tok[-1]}' has type {tok[1]}: only float32 rows are supported")
            names.append(tok[2])
        elif tok[0] == "end_header":
            break
    if fmt != "binary_little_endian":
        raise NotImplementedError(f"PLY format {fmt}: only binary_little_endian is supported")
    if n_vertex is None:
        raise ValueError("PLY file has no vertex element")
    return n_vertex, names, f.tell()


def resolve_columns(names: List[str], max_sh_degree: int | None = 3):
    """Property names -> `gsr_ply_columns`, with load_ply's rules (scene/gaussian_model.py:316-347).  With
    max_sh_degree=None the degree is inferred from the f_rest_* count."""
    if len(names) > _native.PLY_MAX_PROPS:
        raise NotImplementedError(f"{len(names)} properties per vertex: at most {_native.PLY_MAX_PROPS}")
    col = {}
    for i, n in enumerate(names):
        if n in col:
            raise ValueError(f"duplicate PLY property '{n}'")
        col[n] = i
    by_suffix = lambda prefix: sorted((n for n in names if n.startswith(prefix)), key=lambda n: int(n.split("_")[-1]))
    extra, scales, rots = by_suffix("f_rest_"), by_suffix("scale_"), by_suffix("rot")
    if max_sh_degree is None:
        M = {0: 1, 9: 4, 24: 9, 45: 16}.get(len(extra))
        if M is None:
            raise ValueError(f"{len(extra)} f_rest_* properties do not correspond to an SH degree 0..3")
    else:
        M = (max_sh_degree + 1) ** 2
        assert len(extra) == 3 * M - 3, "f_rest_* count does not match max_sh_degree"   # gaussian_model.py:331
    if len(scales) != 3 or len(rots) != 4:
        raise ValueError("expected scale_0..2 and rot_0..3")
    try:
        c = _native.PlyColumns()
        c.n_props, c.n_rest = len(names), 3 * (M - 1)
        for j, k in enumerate(("x", "y", "z")):
            c.xyz[j] = col[k]
        for ch in range(3):
            c.f_dc[ch] = col[f"f_dc_{ch}"]
        for k in range(M - 1):                      # features_rest[p, k, ch] <- f_rest_{ch*(M-1)+k}
            for ch in range(3):
                c.f_rest[k * 3 + ch] = col[extra[ch * (M - 1) + k]]
        c.opacity = col["opacity"]
        for j in range(3):
            c.scale[j] = col[scales[j]]
        for j in range(4):
            c.rot[j] = col[rots[j]]
    except KeyError as e:
        raise ValueError(f"PLY file lacks property {e}") from None
    return c, M


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None and t.numel() else None


def _stream(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def load_ply(path: str, device="cuda", max_sh_degree: int | None = 3) -> Dict[str, torch.Tensor]:
    """Returns the six raw parameter tensors load_ply installs on the model (float32, on `device`):
    xyz [P,3], features_dc [P,1,3], features_rest [P,M-1,3], opacity [P,1], scaling [P,3], rotation [P,4]."""
    device = torch.device(device)
    if device.type != "cuda":
        raise RuntimeError("load_ply de-interleaves on the GPU: device must be a CUDA device")
    with open(path, "rb") as f:
        P, names, offset = read_header(f)
        cols, M = resolve_columns(names, max_sh_degree)
        n = P * len(names)
        host = torch.empty(max(n, 1), dtype=torch.float32, pin_memory=True)
        if n:
            got = f.readinto(memoryview(host.numpy()).cast("B")[: 4 * n])
            if got != 4 * n:
                raise ValueError(f"PLY body truncated: {got} of {4 * n} bytes")
    rows = host[:n].to(device, non_blocking=True)
    out = dict(xyz=torch.empty(P, 3, device=device), features_dc=torch.empty(P, 1, 3, device=device),
               features_rest=torch.empty(P, M - 1, 3, device=device), opacity=torch.empty(P, 1, device=device),
               scaling=torch.empty(P, 3, device=device), rotation=torch.empty(P, 4, device=device))
    if P:
        with torch.cuda.device(device):
            _native.check(_native.lib().gsr_ply_unpack(P, ctypes.byref(cols), _ptr(rows), _ptr(out["xyz"]),
                                                       _ptr(out["features_dc"]), _ptr(out["features_rest"]),
                                                       _ptr(out["opacity"]), _ptr(out["scaling"]), _ptr(out["rotation"]),
                                                       _stream(device)), "gsr_ply_unpack")
    return out


def save_ply(path: str, xyz: torch.Tensor, features_dc: torch.Tensor, features_rest: torch.Tensor,
             opacity: torch.Tensor, scaling: torch.Tensor, rotation: torch.Tensor) -> None:
    """Writes the file save_ply writes (same header, same attribute order, normals zero)."""
    device = xyz.device
    if device.type != "cuda":
        raise RuntimeError("save_ply interleaves on the GPU: tensors must be on a CUDA device")
    P = int(xyz.shape[0])
    n_rest = int(features_rest.shape[1]) * 3 if features_rest.numel() or features_rest.dim() == 3 else 0
    names = attribute_names(n_rest)
    cols, _ = resolve_columns(names, None)
    t = [x.detach().to(torch.float32).contiguous() for x in (xyz, features_dc, features_rest, opacity, scaling, rotation)]
    widths = (3, 3, n_rest, 1, 3, 4)
    for x, w, nm in zip(t, widths, ("xyz", "features_dc", "features_rest", "opacity", "scaling", "rotation")):
        if x.numel() != P * w:
            raise ValueError(f"{nm} has {x.numel()} values, expected {P} x {w}")
    rows = torch.empty(max(P * len(names), 1), dtype=torch.float32, device=device)
    if P:
        with torch.cuda.device(device):
            _native.check(_native.lib().gsr_ply_pack(P, ctypes.byref(cols), *[_ptr(x) for x in t], _ptr(rows),
                                                     _stream(device)), "gsr_ply_pack")
    host = torch.empty(max(P * len(names), 1), dtype=torch.float32, pin_memory=True)
    host.copy_(rows, non_blocking=True)
    torch.cuda.current_stream(device).synchronize()
    d = os.path.dirname(path)
    if d:
        os.makedirs(d, exist_ok=True)                                   # mkdir_p, gaussian_model.py:294
    with open(path, "wb") as f:
        f.write(header_bytes(P, names))
        if P:
            f.write(memoryview(host.numpy()).cast("B")[: 4 * P * len(names)])
