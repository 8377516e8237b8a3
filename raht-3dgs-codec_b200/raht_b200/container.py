"""Compressed-scene container.

The reference never serialises what it codes: encode_3dgs.py:247 takes `len(buf)` of every RLGR stream and drops the
bytes, and the voxel positions travel uncompressed in a PLY (test_voxelize_3dgs.py:284-288).  This module defines the
byte layout a real encoder / decoder pair needs and builds it ON THE DEVICE: one contiguous uint8 tensor holding a small
header, the geometry (Morton codes of the occupied voxels as RLGR-coded Morton deltas, include/raht_b200.h
raht_geometry_symbols), the D attribute streams (the reference-format RLGR bitstreams, unchanged) and, optionally, the
sync indices that let the decoder run chunk-parallel.

Layout (little-endian; every section starts on a multiple of 8 bytes):

    header   "R3GC" | u16 version = 1 | u16 flags | u32 header_bytes | u32 J | u32 D | u32 n_geo = 3 | u32 n_steps
             | u64 n_in | u64 nvox | f32 vmin[3] | f32 0 | f64 voxel_size | f64 width
             | f64 steps[n_steps] (1 = one step for all channels, else D) | u64 geo_bytes[n_geo] | u64 attr_bytes[D]
             | u64 geo_index_entries | u64 attr_index_entries (per channel; 0 = no index) | zero padding to 8
    body     geometry streams, attribute streams (each padded with zero bytes to a multiple of 8),
             geometry sync index (n_geo x entries x 8 B), attribute sync index (D x entries x 8 B)
    flags    bit 0: sync indices present; bit 1: transform ran in fp32 (decoder must do the same)
"""
import struct

import numpy as np
import torch

from . import ops
from . import _lib as L
from ._lib import lib, check, ptr, stream, workspace
from .codec import EncodedScene

MAGIC = b"R3GC"
VERSION = 1
FLAG_INDEX, FLAG_F32 = 1, 2
N_GEO = 3
_FIXED = struct.Struct("<4sHHIIIIIQQ4fdd")


def _pad8(x):
    return (int(x) + 7) // 8 * 8


def geometry_symbols(keys: torch.Tensor, status=None) -> torch.Tensor:
    """(3, n) int32 Morton-delta digits of sorted unique keys (raht_geometry_symbols)."""
    L.require_cuda(keys, "keys")
    n = keys.shape[0]
    ld = max((n + 3) // 4 * 4, 1)
    sym = torch.empty(N_GEO, ld, dtype=torch.int32, device=keys.device)
    own = status is None
    status = ops._status_word(keys.device) if own else status
    check(lib.raht_geometry_symbols(ptr(keys.contiguous()), n, ld, ptr(sym), ptr(status), stream()), "raht_geometry_symbols")
    if own:
        ops.raise_for_status(int(L.read_host(status)[0][0]))
    return sym[:, :n]


def geometry_keys(sym: torch.Tensor) -> torch.Tensor:
    """Inverse of geometry_symbols: (3, n) int32 (row stride >= n) -> (n,) int64 Morton codes."""
    L.require_cuda(sym, "sym")
    n = sym.shape[1]
    if sym.dtype != torch.int32 or sym.stride(1) != 1:
        sym = sym.to(torch.int32).contiguous()
    ld = sym.stride(0) if n > 0 else 1
    keys = torch.empty(n, dtype=torch.int64, device=sym.device)
    ws_bytes = lib.raht_geometry_keys_workspace_bytes(n)
    ws = workspace(ws_bytes, sym.device)
    check(lib.raht_geometry_keys(ptr(sym), n, ld, ptr(keys), ptr(ws), ws_bytes, stream()), "raht_geometry_keys")
    return keys


def _steps_list(step, D):
    if isinstance(step, torch.Tensor):
        step = step.detach().cpu().tolist()
    if isinstance(step, (list, tuple, np.ndarray)):
        vals = [float(s) for s in step]
        if len(vals) not in (1, D):
            raise ValueError(f"{len(vals)} quantisation steps for {D} channels")
        return vals
    return [float(step)]


def _header(J, D, n_in, nvox, vmin, voxel_size, width, steps, geo_sizes, attr_sizes, geo_entries, attr_entries, flags):
    var = struct.pack(f"<{len(steps)}d", *steps) + struct.pack(f"<{N_GEO}Q", *geo_sizes) + struct.pack(f"<{D}Q", *attr_sizes) + \
        struct.pack("<QQ", geo_entries, attr_entries)
    hb = _pad8(_FIXED.size + len(var))
    fixed = _FIXED.pack(MAGIC, VERSION, flags, hb, J, D, N_GEO, len(steps), n_in, nvox, vmin[0], vmin[1], vmin[2], 0.0,
                        voxel_size, width)
    return (fixed + var).ljust(hb, b"\0")


@torch.no_grad()
def pack(enc: EncodedScene, with_index=True) -> torch.Tensor:
    """EncodedScene -> one uint8 CUDA tensor (the container).  Two synchronisations: the sizes of the attribute streams
    (already on the host after the encoder) and of the geometry streams."""
    dev = enc.keys.device
    status = ops._status_word(dev)
    sym = geometry_symbols(enc.keys, status)
    with_index = with_index and enc.streams.sync_index is not None
    geo = ops.rlgr_encode(sym, 0, with_index=with_index, status=status)          # reads the status word (one sync)
    geo_sizes, attr_sizes = geo.sizes(), enc.streams.sizes()
    steps = _steps_list(enc.step, enc.D)
    geo_entries = int(lib.raht_rlgr_sync_entries(enc.nvox)) if with_index else 0
    flags = (FLAG_INDEX if with_index else 0) | (FLAG_F32 if enc.transform_dtype == torch.float32 else 0)
    vmin = [float(x) for x in enc.vmin.detach().cpu().tolist()]
    head = _header(enc.J, enc.D, enc.n_in, enc.nvox, vmin, float(enc.voxel_size), float(enc.width), steps, geo_sizes, attr_sizes,
                   geo_entries, geo_entries, flags)
    geo_body, attr_body = sum(_pad8(s) for s in geo_sizes), sum(_pad8(s) for s in attr_sizes)
    index_bytes = (N_GEO + enc.D) * geo_entries * 8
    out = torch.empty(len(head) + geo_body + attr_body + index_bytes, dtype=torch.uint8, device=dev)
    out[:len(head)].copy_(torch.frombuffer(bytearray(head), dtype=torch.uint8), non_blocking=True)
    o = len(head)
    geo.packed(out[o:o + geo_body], status)
    o += geo_body
    enc.streams.packed(out[o:o + attr_body], status)
    o += attr_body
    if with_index:
        for s in (geo, enc.streams):
            b = s.sync_index.reshape(-1).view(torch.uint8)
            out[o:o + b.numel()].copy_(b)
            o += b.numel()
    bits = int(L.read_host(status)[0][0])                   # raht_streams_pack reports a destination that is too small here
    ops.raise_for_status(bits)
    if bits & L.ST_STREAM_OVERFLOW:
        raise RuntimeError("container.pack: stream sizes changed between the size read and the pack (internal error)")
    return out


def parse_header(head: bytes) -> dict:
    """Header fields from the first bytes of a container (needs at least header_bytes of them; 76 are enough to learn
    header_bytes)."""
    if len(head) < _FIXED.size:
        raise ValueError("container: truncated header")
    magic, version, flags, hb, J, D, n_geo, n_steps, n_in, nvox, vx, vy, vz, _, voxel_size, width = _FIXED.unpack_from(head, 0)
    if magic != MAGIC:
        raise ValueError("container: bad magic")
    if version != VERSION or n_geo != N_GEO or n_steps not in (1, D):
        raise ValueError("container: unsupported version or malformed header")
    h = dict(flags=flags, header_bytes=hb, J=J, D=D, n_in=n_in, nvox=nvox, vmin=(vx, vy, vz), voxel_size=voxel_size, width=width)
    if len(head) >= hb:
        o = _FIXED.size
        h["steps"] = list(struct.unpack_from(f"<{n_steps}d", head, o)); o += 8 * n_steps
        h["geo_bytes"] = list(struct.unpack_from(f"<{N_GEO}Q", head, o)); o += 8 * N_GEO
        h["attr_bytes"] = list(struct.unpack_from(f"<{D}Q", head, o)); o += 8 * D
        h["geo_entries"], h["attr_entries"] = struct.unpack_from("<QQ", head, o)
    return h


def _unpack_streams(body: torch.Tensor, sizes, n, index):
    """body: 8-aligned slice with the padded streams end to end -> RlgrStreams (slot layout the decoder reads)."""
    dev = body.device
    nch = len(sizes)
    offs = np.zeros(nch + 1, dtype=np.int64)
    offs[1:] = np.cumsum([_pad8(s) for s in sizes])
    slot = max(_pad8(max(sizes) if sizes else 0), 8)
    buf = torch.empty(nch * slot, dtype=torch.uint8, device=dev)
    offsets = torch.from_numpy(offs).to(dev)
    check(lib.raht_streams_unpack(ptr(body), ptr(offsets), nch, ptr(buf), slot, stream()), "raht_streams_unpack")
    nbytes = torch.tensor([int(s) for s in sizes], dtype=torch.int64).to(dev)
    return ops.RlgrStreams(buf, slot, nbytes, n, nch, index)


@torch.no_grad()
def unpack(container, device=None) -> EncodedScene:
    """Container (uint8 CUDA tensor, or host bytes / numpy / CPU tensor, copied to `device`) -> EncodedScene with the
    voxel keys decoded on the device.  The attribute streams stay coded: codec.decode_scene() decodes them."""
    if isinstance(container, (bytes, bytearray, memoryview)):
        container = torch.frombuffer(bytearray(container), dtype=torch.uint8)
    elif isinstance(container, np.ndarray):
        container = torch.from_numpy(np.ascontiguousarray(container).view(np.uint8).reshape(-1))
    if not container.is_cuda:
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        head_src = container
        container = container.to(dev)
    else:
        dev = container.device
        head_src = None
    total = container.numel()
    peek = (head_src if head_src is not None else container)[:min(total, _FIXED.size)].cpu().numpy().tobytes()
    hb = parse_header(peek)["header_bytes"]
    if hb > total:
        raise ValueError("container: truncated header")
    h = parse_header((head_src if head_src is not None else container)[:hb].cpu().numpy().tobytes())
    D, nvox = h["D"], h["nvox"]
    # the header is untrusted input: every size the kernels will index with is checked against the buffer first
    if not (1 <= h["J"] <= 21) or not (1 <= D <= 4096) or nvox < 0 or nvox > (1 << 40):
        raise ValueError("container: J, D or nvox out of range")
    if len(h["attr_bytes"]) != D or len(h["geo_bytes"]) != N_GEO:
        raise ValueError("container: malformed stream table")
    if any(s < 0 or s > total for s in h["geo_bytes"] + h["attr_bytes"]) or any(not (st > 0.0) for st in h["steps"]):
        raise ValueError("container: stream size or quantisation step out of range")
    if h["flags"] & FLAG_INDEX:
        want = int(lib.raht_rlgr_sync_entries(nvox))
        if h["geo_entries"] != want or h["attr_entries"] != want:
            raise ValueError("container: sync index does not match the number of voxels")
    elif h["geo_entries"] or h["attr_entries"]:
        raise ValueError("container: index entries without the index flag")
    geo_body, attr_body = sum(_pad8(s) for s in h["geo_bytes"]), sum(_pad8(s) for s in h["attr_bytes"])
    index_bytes = (N_GEO * h["geo_entries"] + D * h["attr_entries"]) * 8
    if hb + geo_body + attr_body + index_bytes > total:
        raise ValueError("container: truncated body")
    o = hb + geo_body + attr_body
    geo_index = attr_index = None
    if h["flags"] & FLAG_INDEX:
        nb = N_GEO * h["geo_entries"] * 8
        geo_index = container[o:o + nb].view(torch.int32).reshape(-1, 2)
        attr_index = container[o + nb:o + nb + D * h["attr_entries"] * 8].view(torch.int32).reshape(-1, 2)
    geo = _unpack_streams(container[hb:hb + geo_body], h["geo_bytes"], nvox, geo_index)
    keys = geometry_keys(ops.rlgr_decode(geo, 0))
    streams = _unpack_streams(container[hb + geo_body:hb + geo_body + attr_body], h["attr_bytes"], nvox, attr_index)
    steps = h["steps"]
    return EncodedScene(J=h["J"], n_in=h["n_in"], nvox=nvox, D=D, step=steps[0] if len(steps) == 1 else steps,
                        vmin=torch.tensor(h["vmin"], dtype=torch.float32, device=dev), voxel_size=h["voxel_size"],
                        width=h["width"], keys=keys, streams=streams,
                        transform_dtype=torch.float32 if h["flags"] & FLAG_F32 else torch.float64)


def write_scene(path, enc: EncodedScene, with_index=True) -> int:
    """Container of `enc` to a file; returns its size in bytes."""
    data = pack(enc, with_index).cpu().numpy()
    with open(path, "wb") as f:
        f.write(data.tobytes())
    return int(data.size)


def read_scene(path, device=None) -> EncodedScene:
    with open(path, "rb") as f:
        return unpack(f.read(), device)
