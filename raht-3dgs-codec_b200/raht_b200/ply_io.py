"""Scene I/O in the reference's PLY dialect, built on the device.

The reference exchanges voxelised scenes between its stages as binary PLY files:
  * `save_ply` (python/quality_eval.py:18-117) writes the header, then loops over the Gaussians in Python with six
    small writes each (minutes for a 6 M scene);
  * `read_compressed_3dgs_ply` (python/data_util.py:272-382) parses the 62-float vertex records through a numpy
    structured dtype and a dozen np.stack calls, and takes `voxel_size` / `vmin` from header comments.
Same file format, same signatures and return values here -- but the file body is ONE buffer: a CUDA kernel interleaves
the attribute arrays into vertex records (raht_ply_pack_vertices) or splits them (raht_ply_unpack_vertices), the host
side is a single write / read through pinned memory.  Files written by either implementation are read by the other.
"""
import os
import warnings
from pathlib import Path
from typing import Optional, Tuple

import numpy as np
import torch

from . import _lib as L
from ._lib import lib, check, ptr, stream


def _header(n, color_dim, voxel_size, vmin) -> bytes:
    """The header of quality_eval.py:58-92, byte for byte (every colour property is named f_dc_<i>, as there)."""
    h = [b"ply\n", b"format binary_little_endian 1.0\n"]
    if voxel_size is not None:
        h.append(f"comment voxel_size {voxel_size}\n".encode())
    if vmin is not None:
        v = vmin.detach().cpu().float().numpy() if isinstance(vmin, torch.Tensor) else np.asarray(vmin, dtype=np.float32)
        h.append(f"comment vmin {v[0]} {v[1]} {v[2]}\n".encode())
    h.append(f"element vertex {n}\n".encode())
    for name in ("x", "y", "z", "nx", "ny", "nz"):
        h.append(f"property float {name}\n".encode())
    for i in range(color_dim):
        h.append(f"property float f_dc_{i}\n".encode())
    h.append(b"property float opacity\n")
    for i in range(3):
        h.append(f"property float scale_{i}\n".encode())
    for i in range(4):
        h.append(f"property float rot_{i}\n".encode())
    h.append(b"end_header\n")
    return b"".join(h)


@torch.no_grad()
def save_ply(filepath: str, means: torch.Tensor, quats: torch.Tensor, scales: torch.Tensor, opacities: torch.Tensor,
             colors: torch.Tensor, voxel_size: float = None, vmin: torch.Tensor = None) -> int:
    """quality_eval.save_ply: [N,3] means, [N,4] quats, [N,3] scales, [N] opacities, [N,C] colours -> binary PLY.
    Returns the file size in bytes.  CUDA tensors (this package has no CPU path)."""
    L.require_cuda(means, "means")
    n, cd = means.shape[0], colors.shape[1]
    f32 = lambda t: t.detach().contiguous().float()
    means, quats, scales, opacities, colors = f32(means), f32(quats), f32(scales), f32(opacities).reshape(-1), f32(colors)
    rec = torch.empty(n, 14 + cd, dtype=torch.float32, device=means.device)
    check(lib.raht_ply_pack_vertices(ptr(means), ptr(quats), ptr(scales), ptr(opacities), ptr(colors), cd, n, ptr(rec), stream()),
          "raht_ply_pack_vertices")
    host = torch.empty(rec.shape, dtype=torch.float32).pin_memory()
    host.copy_(rec, non_blocking=True)
    head = _header(n, cd, voxel_size, vmin)
    Path(filepath).parent.mkdir(parents=True, exist_ok=True)
    torch.cuda.current_stream(means.device).synchronize()
    with open(filepath, "wb") as f:
        f.write(head)
        f.write(host.numpy().tobytes() if n else b"")
    return len(head) + n * (14 + cd) * 4


def _parse_header(f):
    """-> (num_vertices, is_binary, voxel_size or None, vmin or None, floats per vertex), data_util.py:296-330."""
    lines = []
    while True:
        raw = f.readline()
        if not raw:
            raise ValueError("PLY header has no end_header")
        line = raw.decode("ascii").strip()
        lines.append(line)
        if line == "end_header":
            break
    nv, is_binary, voxel_size, vmin, props = 0, False, None, None, 0
    for line in lines:
        if line.startswith("format"):
            is_binary = "binary" in line
        elif line.startswith("element vertex"):
            nv = int(line.split()[-1])
        elif line.startswith("comment voxel_size"):
            voxel_size = float(line.split()[-1])
        elif line.startswith("comment vmin"):
            p = line.split()
            vmin = torch.tensor([float(p[2]), float(p[3]), float(p[4])], dtype=torch.float32)
        elif line.startswith("property float"):
            props += 1
    return nv, is_binary, voxel_size, vmin, props


@torch.no_grad()
def read_compressed_3dgs_ply(filename: str, device=None) -> Optional[Tuple[torch.Tensor, torch.Tensor, float, torch.Tensor]]:
    """data_util.read_compressed_3dgs_ply: -> (V_int [N,3] int64, attributes [N,56] fp32 = quats 4 | scales 3 | opacity 1 |
    colours 48, voxel_size, vmin [3]) as CUDA tensors on `device`; None (with a warning) when the file cannot be read,
    like the reference (data_util.py:375-382).  Missing header comments fall back to voxel_size 1.0 / vmin 0 with a
    warning (:319-325)."""
    try:
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        with open(filename, "rb") as f:
            nv, is_binary, voxel_size, vmin, props = _parse_header(f)
            if nv == 0:
                raise ValueError("Could not find vertex count in PLY header")
            if not is_binary:
                raise ValueError("ASCII format not supported for compressed 3DGS PLY. Use binary format.")
            if props < 14:
                raise ValueError(f"{props} float properties per vertex: not a 3DGS vertex record")
            if voxel_size is None:
                warnings.warn("Could not find voxel_size in PLY header comments")
                voxel_size = 1.0
            if vmin is None:
                warnings.warn("Could not find vmin in PLY header comments")
                vmin = torch.zeros(3, dtype=torch.float32)
            cd = props - 14
            host = torch.empty(nv, props, dtype=torch.float32).pin_memory()
            got = f.readinto(memoryview(host.numpy()).cast("B"))
            if got != nv * props * 4:
                raise ValueError(f"truncated PLY body: {got} of {nv * props * 4} bytes")
        rec = host.to(dev, non_blocking=True)
        V_int = torch.empty(nv, 3, dtype=torch.int64, device=dev)
        attrs = torch.empty(nv, 8 + cd, dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            check(lib.raht_ply_unpack_vertices(ptr(rec), cd, nv, ptr(V_int), None, ptr(attrs), stream()), "raht_ply_unpack_vertices")
            torch.cuda.current_stream(dev).synchronize()          # (the pinned staging buffer dies with this call)
        return V_int, attrs, voxel_size, vmin.to(dev)
    except FileNotFoundError:
        warnings.warn(f"File not found: {filename}")
        return None
    except Exception as e:                                         # the reference returns None on any failure
        warnings.warn(f"Error reading compressed 3DGS PLY {filename}: {e}")
        return None


def file_size(n, color_dim=48, voxel_size=None, vmin=None) -> int:
    return len(_header(n, color_dim, voxel_size, vmin)) + n * (14 + color_dim) * 4
